// omg_ops.cuh -- level operations as device functions over one box, written for an arbitrary
// thread partition (tid of nthr).  The batched kernels run one CTA per box; the tail interpreter
// runs them for every box of a small level inside a single CTA.  These are the general path
// (any box size, any face kind); omg_fast.cuh holds the tuned sm_100a kernels for big levels.
#pragma once
#include "omg_device.cuh"

// smooth_boxes inner loop, one colour half-sweep: src/m_multigrid.f90:412-418 -> box_gs_*
template <int OP>
__device__ void do_smooth(const LevelDev& L, const ParamsDev& P, int b, int colour, int tid, int nthr) {
  double* phi = L.var[0];
  for (int idx = tid; idx < L.cs; idx += nthr) {
    int i, j, k;
    if (!decode_colour(L, colour, idx, i, j, k)) continue;
    phi[(size_t)b * L.bs + colour * L.cs + idx] = gs_at<OP>(L, P, b, i, j, k);
  }
}

// Odd box sizes (only possible on the coarsest level, e.g. 1^3 boxes): the local red-black colouring no
// longer matches across boxes, so a half-sweep reads neighbour cells of the colour being updated.  The
// reference reads them from ghost cells filled BEFORE the half-sweep (Jacobi between boxes), so the new
// values are staged in the res variable (scratch at this point of the coarse solve) and committed after
// a barrier.  phase 0: compute into res; phase 1: commit.
template <int OP>
__device__ void do_smooth_staged(const LevelDev& L, const ParamsDev& P, int b, int colour, int phase, int tid, int nthr) {
  double* phi = L.var[0];
  double* tmp = L.var[3];
  for (int idx = tid; idx < L.cs; idx += nthr) {
    int i, j, k;
    if (!decode_colour(L, colour, idx, i, j, k)) continue;
    const size_t g = (size_t)b * L.bs + colour * L.cs + idx;
    if (phase == 0) tmp[g] = gs_at<OP>(L, P, b, i, j, k); else phi[g] = tmp[g];
  }
}

// ---- mg_smoother_gs: lexicographic Gauss-Seidel inside each box, Jacobi between boxes -----------------
// (box_gs_* with redblack = .false., e.g. src/m_laplacian.f90:99-111: i0 = 1, stride 1, loop order k, j, i.)
// The reference sweeps every box against ghost cells filled BEFORE the sweep, so the ghost faces are first
// snapshotted (do_gs_snapshot = the ghost fill), then each box is swept in place.  A lexicographic sweep only
// depends on already-updated lower neighbours (i-1, j-1, k-1) and not-yet-updated upper ones, so updating the
// hyperplanes i + j + k = w for w = 0, 1, ... in turn performs exactly the same arithmetic on the same values.
__device__ void do_gs_snapshot(const LevelDev& L, const ParamsDev& P, int b, int tid, int nthr) {
  const int nc = L.nc;
  double* out = L.faces[0] + (size_t)b * 6 * nc * nc;
  for (int idx = tid; idx < 6 * nc * nc; idx += nthr) {
    const int f = idx / (nc * nc), r = idx - f * nc * nc, t1 = r % nc, t2 = r / nc;
    const int g = (f & 1) ? nc : -1;
    int i, j, k;
    if (f < 2) { i = g; j = t1; k = t2; } else if (f < 4) { i = t1; j = g; k = t2; } else { i = t1; j = t2; k = g; }
    out[idx] = cell_at<GM_PHI>(L, P, L.var[0], nullptr, b, i, j, k);
  }
}
// all threads of the CTA must call this (it synchronises between hyperplanes)
template <int OP>
__device__ void do_gs_sweep(const LevelDev& L, const ParamsDev& P, int b, int tid, int nthr) {
  const int nc = L.nc;
  double* phi = L.var[0] + (size_t)b * L.bs;
  for (int w = 0; w <= 3 * (nc - 1); ++w) {
    for (int idx = tid; idx < nc * nc; idx += nthr) {
      const int j = idx % nc, k = idx / nc, i = w - j - k;
      if (i >= 0 && i < nc) phi[cidx(L, i, j, k)] = gs_at<OP, GM_STORED>(L, P, b, i, j, k);
    }
    __syncthreads();
  }
}

// residual_box src/m_multigrid.f90:426-436; returns the thread-local max |res|
template <int OP>
__device__ double do_resid(const LevelDev& L, const ParamsDev& P, int b, int tid, int nthr) {
  double m = 0.0;
  for (int idx = tid; idx < L.bs; idx += nthr) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (!decode_colour(L, colour, idx - colour * L.cs, i, j, k)) continue;
    const double r = L.var[1][(size_t)b * L.bs + idx] - op_at<OP>(L, P, b, i, j, k);
    L.var[3][(size_t)b * L.bs + idx] = r;
    m = fmax(m, fabs(r));
  }
  return m;
}

// restrict_onto src/m_restrict.f90:165-214 seen from the child: child box b of Lf writes its
// parent's octant.  0.125 * sum(2x2x2) in Fortran array-element order (i fastest).
__device__ void do_restrict(const LevelDev& Lf, const LevelDev& Lc, int iv, int b, int tid, int nthr) {
  const int p = Lf.parent[b];
  if (p < 0) return;   // parent on another rank: handled by the restrict exchange
  const int off = Lf.coff[b];
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  const int h = Lf.nc >> 1;
  const double* src = Lf.var[iv] + (size_t)b * Lf.bs;
  double* dst = Lc.var[iv] + (size_t)p * Lc.bs;
  for (int idx = tid; idx < h * h * h; idx += nthr) {
    const int I = idx % h, J = (idx / h) % h, K = idx / (h * h);
    double s = 0.0;
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
#pragma unroll
        for (int ii = 0; ii < 2; ++ii) s = s + src[cidx(Lf, 2 * I + ii, 2 * J + jj, 2 * K + kk)];
    dst[cidx(Lc, ox + I, oy + J, oz + K)] = 0.125 * s;
  }
}

// tail of update_coarse src/m_multigrid.f90:369-383 on one parent box: rhs = L(phi) + res; old = phi
template <int OP>
__device__ void do_crhs(const LevelDev& L, const ParamsDev& P, int b, int tid, int nthr) {
  for (int idx = tid; idx < L.bs; idx += nthr) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (!decode_colour(L, colour, idx - colour * L.cs, i, j, k)) continue;
    const size_t g = (size_t)b * L.bs + idx;
    L.var[1][g] = op_at<OP>(L, P, b, i, j, k) + L.var[3][g];
    L.var[2][g] = L.var[0][g];
  }
}

// "old ghost" snapshots: the ghost values that `old(:) = phi(:)` copies along
// (src/m_multigrid.f90:381-382, :121-126) where they cannot be re-derived from old's interior.
__device__ void do_oldpl(const LevelDev& L, const ParamsDev& P, int e, int tid, int nthr) {
  const int b = L.oldpl_list[3 * e], f = L.oldpl_list[3 * e + 1], pl = L.oldpl_list[3 * e + 2];
  const int nc = L.nc;
  for (int idx = tid; idx < nc * nc; idx += nthr) {
    const int t1 = idx % nc, t2 = idx / nc;
    int i, j, k;
    const int g = (f & 1) ? nc : -1;
    if (f < 2) { i = g; j = t1; k = t2; } else if (f < 4) { i = t1; j = g; k = t2; } else { i = t1; j = t2; k = g; }
    L.gpl_old[(size_t)pl * nc * nc + idx] = cell_at<GM_PHI>(L, P, L.var[0], nullptr, b, i, j, k);
  }
}

// correct_children src/m_multigrid.f90:387-402 fused with mg_prolong / prolong_onto / mg_prolong_sparse
// (src/m_prolong.f90:124-156, :159-240) seen from the child box b of Lf:
//   phi_f += P(phi_c - old_c), and the parent's octant of res_c = phi_c - old_c is stored.
__device__ __forceinline__ double corr_at(const LevelDev& Lc, const ParamsDev& P, int p, int i, int j, int k) {
  return cell_at<GM_PHI>(Lc, P, Lc.var[0], nullptr, p, i, j, k) - cell_at<GM_OLD>(Lc, P, Lc.var[2], nullptr, p, i, j, k);
}
__device__ void do_corr_prolong(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int b, int tid, int nthr) {
  const int p = Lf.parent[b];
  if (p == -1) return;
  const int off = Lf.coff[b];
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  const int h = Lf.nc >> 1, pw = h + 2;
  // parent on another rank: its rank sent the (h+2)^3 patch of phi_c - old_c around my octant
  const double* patch = p <= -2 ? Lf.ppatch + (size_t)(-(p + 2)) * pw * pw * pw : nullptr;
  auto corr = [&](int I, int J, int K) -> double {   // octant coordinates, -1 .. h
    if (patch) return patch[((K + 1) * pw + (J + 1)) * pw + (I + 1)];
    return corr_at(Lc, P, p, ox + I, oy + J, oz + K);
  };
  double* phi = Lf.var[0] + (size_t)b * Lf.bs;
  for (int idx = tid; idx < h * h * h; idx += nthr) {
    const int I = idx % h, J = (idx / h) % h, K = idx / (h * h);
    const double c = corr(I, J, K);
    if (!patch) Lc.var[3][(size_t)p * Lc.bs + cidx(Lc, ox + I, oy + J, oz + K)] = c;
    const double f0 = 0.25 * c;
    const double flx = 0.25 * corr(I - 1, J, K), fhx = 0.25 * corr(I + 1, J, K);
    const double fly = 0.25 * corr(I, J - 1, K), fhy = 0.25 * corr(I, J + 1, K);
    const double flz = 0.25 * corr(I, J, K - 1), fhz = 0.25 * corr(I, J, K + 1);
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
#pragma unroll
        for (int ii = 0; ii < 2; ++ii) {
          const double fine = f0 + (ii ? fhx : flx) + (jj ? fhy : fly) + (kk ? fhz : flz);
          const int q = cidx(Lf, 2 * I + ii, 2 * J + jj, 2 * K + kk);
          phi[q] = phi[q] + fine;
        }
  }
}

// prolong_set_buffer (src/m_prolong.f90:91-121) seen from the parent's rank: instead of the prolonged
// fine block (nc^3 values) the (nc/2+2)^3 patch of the correction phi_c - old_c around the child's octant
// is sent; the child's rank applies mg_prolong_sparse itself (same values, 4x fewer bytes).  Also stores
// res_c = phi_c - old_c on the octant (correct_children, src/m_multigrid.f90:396-398).
__device__ void do_pack_prolong(const LevelDev& Lc, const ParamsDev& P, int p, int off, double* __restrict__ out, int nc_f,
                                int tid, int nthr) {
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  const int h = nc_f >> 1, pw = h + 2;
  for (int r = tid; r < pw * pw * pw; r += nthr) {
    const int a = r % pw, bb = (r / pw) % pw, c = r / (pw * pw);
    const int nedge = (a == 0 || a == pw - 1) + (bb == 0 || bb == pw - 1) + (c == 0 || c == pw - 1);
    double v = 0.0;
    if (nedge <= 1) {
      const int ic = ox + a - 1, jc = oy + bb - 1, kc = oz + c - 1;
      v = corr_at(Lc, P, p, ic, jc, kc);
      if (nedge == 0) Lc.var[3][(size_t)p * Lc.bs + cidx(Lc, ic, jc, kc)] = v;
    }
    out[r] = v;
  }
}

// the same for the stand-alone mg_prolong of any variable (src/m_prolong.f90:51-85): the patch holds the coarse VALUES
// (ghost cells through cell_at with the variable's ghost mode); edges and corners are never read
template <int GMV>
__device__ void do_pack_prolong_var(const LevelDev& Lc, const ParamsDev& P, int iv, int p, int off, double* __restrict__ out, int nc_f,
                                    int tid, int nthr) {
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  const int h = nc_f >> 1, pw = h + 2;
  const double* cv = Lc.var[iv]; const double* cf = Lc.faces[iv];
  for (int r = tid; r < pw * pw * pw; r += nthr) {
    const int a = r % pw, bb = (r / pw) % pw, c = r / (pw * pw);
    const int nedge = (a == 0 || a == pw - 1) + (bb == 0 || bb == pw - 1) + (c == 0 || c == pw - 1);
    out[r] = nedge <= 1 ? cell_at<GMV>(Lc, P, cv, cf, p, ox + a - 1, oy + bb - 1, oz + c - 1) : 0.0;
  }
}

// restrict_set_buffer (src/m_restrict.f90:116-163): restricted block of child b for a parent on another rank
__device__ void do_pack_restrict(const LevelDev& Lf, int iv, int b, double* __restrict__ out, int tid, int nthr) {
  const int h = Lf.nc >> 1;
  const double* src = Lf.var[iv] + (size_t)b * Lf.bs;
  for (int idx = tid; idx < h * h * h; idx += nthr) {
    const int I = idx % h, J = (idx / h) % h, K = idx / (h * h);
    double s = 0.0;
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
#pragma unroll
        for (int ii = 0; ii < 2; ++ii) s = s + src[cidx(Lf, 2 * I + ii, 2 * J + jj, 2 * K + kk)];
    out[idx] = 0.125 * s;
  }
}
// restrict_onto for a child on another rank (src/m_restrict.f90:195-212): received block -> parent's octant
__device__ void do_unpack_restrict(const LevelDev& Lc, int iv, int p, int off, const double* __restrict__ in, int h,
                                   int tid, int nthr) {
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  double* dst = Lc.var[iv] + (size_t)p * Lc.bs;
  for (int idx = tid; idx < h * h * h; idx += nthr) {
    const int I = idx % h, J = (idx / h) % h, K = idx / (h * h);
    dst[cidx(Lc, ox + I, oy + J, oz + K)] = in[idx];
  }
}

// buffer_for_nb (src/m_ghost_cells.f90:348-383): the interior layer of box sb that a same-level
// neighbour on another rank needs as its ghost face f (f = face of the RECEIVER)
__device__ void do_pack_face(const LevelDev& L, int sb, int f, double* __restrict__ out, int tid, int nthr, int iv = 0) {
  const int nc = L.nc;
  const int layer = (f & 1) ? 0 : nc - 1;
  const double* v = L.var[iv] + (size_t)sb * L.bs;
  for (int idx = tid; idx < nc * nc; idx += nthr) {
    const int t1 = idx % nc, t2 = idx / nc;
    int i, j, k;
    if (f < 2) { i = layer; j = t1; k = t2; } else if (f < 4) { i = t1; j = layer; k = t2; } else { i = t1; j = t2; k = layer; }
    out[idx] = v[cidx(L, i, j, k)];
  }
}

// generic mg_prolong (src/m_prolong.f90:51-85) of variable iv (ghost mode GMV) into iv_to
template <int GMV>
__device__ void do_prolong_var(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int iv, int iv_to, int add,
                               int b, int tid, int nthr) {
  const int p = Lf.parent[b];
  if (p == -1) return;
  const int off = Lf.coff[b];
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  const int h = Lf.nc >> 1, pw = h + 2;
  const double* cv = Lc.var[iv]; const double* cf = Lc.faces[iv];
  // parent on another rank: the (nc/2+2)^3 patch of coarse values around the octant arrived with the prolong exchange
  const double* patch = p <= -2 ? Lf.ppatch + (size_t)(-(p + 2)) * pw * pw * pw : nullptr;
  double* dst = Lf.var[iv_to] + (size_t)b * Lf.bs;
  auto C = [&](int a, int bb, int c) -> double {
    if (patch) return patch[((size_t)(c + 1) * pw + (bb + 1)) * pw + (a + 1)];
    return cell_at<GMV>(Lc, P, cv, cf, p, ox + a, oy + bb, oz + c);
  };
  for (int idx = tid; idx < h * h * h; idx += nthr) {
    const int I = idx % h, J = (idx / h) % h, K = idx / (h * h);
    const double f0 = 0.25 * C(I, J, K);
    const double flx = 0.25 * C(I - 1, J, K), fhx = 0.25 * C(I + 1, J, K);
    const double fly = 0.25 * C(I, J - 1, K), fhy = 0.25 * C(I, J + 1, K);
    const double flz = 0.25 * C(I, J, K - 1), fhz = 0.25 * C(I, J, K + 1);
    for (int kk = 0; kk < 2; ++kk) for (int jj = 0; jj < 2; ++jj) for (int ii = 0; ii < 2; ++ii) {
      const double fine = f0 + (ii ? fhx : flx) + (jj ? fhy : fly) + (kk ? fhz : flz);
      const int q = cidx(Lf, 2 * I + ii, 2 * J + jj, 2 * K + kk);
      dst[q] = add ? dst[q] + fine : fine;
    }
  }
}

// fill_refinement_bnd + box_gc_for_fine_neighbor + sides_rb (src/m_ghost_cells.f90:287-328, :500-577,
// :769-861) for refinement-boundary entry e of the fine level: materialises the ghost plane.
// GMC = GM_PHI normally, GM_PHI_GZ while the coarse ghosts are the zeros written by mg_fas_fmg.
// GMV selects the variable class (phi: GM_PHI/GM_PHI_GZ with plane pool; stored: faces pool).
// box_gc_for_fine_neighbor (src/m_ghost_cells.f90:500-577): interpolated coarse value for fine ghost cell
// (A, B) (0-based tangential indices) of face f of a fine box with child offset di, from coarse box cb
template <int GMC>
__device__ __forceinline__ double rb_gc_value(const LevelDev& Lc, const ParamsDev& P, int iv, int cb, const int* di, int f,
                                              int A, int B) {
  const int nc = Lc.nc;
  const int d = f >> 1, ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;
  const int cfix = (f & 1) ? 0 : nc - 1;          // coarse interior layer adjacent to the fine box
  const double* cv = Lc.var[iv]; const double* cf = Lc.faces[iv];
  const int ia = (A >> 1), jb = (B >> 1);
  int c[3];
  auto T = [&](int da, int db) {
    c[d] = cfix; c[ta] = di[ta] + ia + da; c[tb] = di[tb] + jb + db;
    return (iv == 0) ? cell_at<GMC>(Lc, P, cv, nullptr, cb, c[0], c[1], c[2])
                     : cell_at<GM_STORED>(Lc, P, cv, cf, cb, c[0], c[1], c[2]);
  };
  const double t0 = T(0, 0);
  const double g1 = 0.125 * (T(1, 0) - T(-1, 0));
  const double g2 = 0.125 * (T(0, 1) - T(0, -1));
  double gc = (A & 1) ? t0 + g1 : t0 - g1;
  gc = (B & 1) ? gc + g2 : gc - g2;
  return gc;
}

// buffer_for_fine_nb (src/m_ghost_cells.f90:385-422): the coarse owner interpolates the face for a fine
// box on another rank (fine child offset `off`, receiver face f) from its coarse box cb
template <int GMC>
__device__ void do_pack_rb_face(const LevelDev& Lc, const ParamsDev& P, int cb, int off, int f, double* __restrict__ out,
                                int tid, int nthr, int iv = 0) {
  const int nc = Lc.nc;
  const int di[3] = {off & 1023, (off >> 10) & 1023, (off >> 20) & 1023};
  for (int idx = tid; idx < nc * nc; idx += nthr) out[idx] = rb_gc_value<GMC>(Lc, P, iv, cb, di, f, idx % nc, idx / nc);
}

// fill_refinement_bnd + box_gc_for_fine_neighbor + sides_rb (src/m_ghost_cells.f90:287-328, :500-577,
// :769-861) for refinement-boundary entry e of the fine level: materialises the ghost plane.
// GMC = GM_PHI normally, GM_PHI_GZ while the coarse ghosts are the zeros written by mg_fas_fmg.
template <int GMC>
__device__ void do_rb_fill(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int iv, int e, int tid, int nthr,
                           const double* rpl = nullptr) {
  const int b = Lf.rb_list[4 * e], f = Lf.rb_list[4 * e + 1], cb = Lf.rb_list[4 * e + 2], pl = Lf.rb_list[4 * e + 3];
  if (cb == -1) return;
  const int nc = Lf.nc;
  const int off = Lf.coff[b];
  const int di[3] = {off & 1023, (off >> 10) & 1023, (off >> 20) & 1023};
  const int d = f >> 1, ta = (d == 0) ? 1 : 0, tb = (d == 2) ? 1 : 2;
  const double* fv = Lf.var[iv] + (size_t)b * Lf.bs;
  const int ix = (f & 1) ? nc - 1 : 0, dix = (f & 1) ? -1 : 1;
  // coarse neighbour on another rank: the interpolated coarse values arrived with the halo exchange
  // (extra variables: with their own exchange, in the receive buffer rpl)
  const double* gcpl = cb <= -2 ? (rpl ? rpl : gpl_cur(Lf)) + (size_t)(-(cb + 2)) * nc * nc : nullptr;
  double* out = (iv == 0) ? gpl_cur(Lf) + (size_t)pl * nc * nc : Lf.faces[iv] + ((size_t)b * 6 + f) * nc * nc;
  for (int idx = tid; idx < nc * nc; idx += nthr) {
    const int A = idx % nc, B = idx / nc;          // 0-based fine tangential indices
    const double gc = gcpl ? gcpl[idx] : rb_gc_value<GMC>(Lc, P, iv, cb, di, f, A, B);
    int q[3], r[3];
    q[d] = ix; r[d] = ix + dix; q[ta] = r[ta] = A; q[tb] = r[tb] = B;
    out[idx] = 0.5 * gc + 0.75 * fv[cidx(Lf, q[0], q[1], q[2])] - 0.25 * fv[cidx(Lf, r[0], r[1], r[2])];
  }
}

// box_op into an arbitrary variable (mg_apply_op, src/m_multigrid.f90:439-456)
template <int OP>
__device__ void do_apply_op(const LevelDev& L, const ParamsDev& P, int i_out, int b, int tid, int nthr) {
  for (int idx = tid; idx < L.bs; idx += nthr) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (!decode_colour(L, colour, idx - colour * L.cs, i, j, k)) continue;
    L.var[i_out][(size_t)b * L.bs + idx] = op_at<OP>(L, P, b, i, j, k);
  }
}

// materialise the ghost faces of a stored-face variable (mg_fill_ghost_cells_lvl for iv >= 5):
// same-level copy and physical boundary (constant bc of that variable); refinement boundaries are
// handled by do_rb_fill.  bc_type / bc_value are the variable's constants for face f.
// rpl / n_recv: planes of this variable received from other ranks (same-level faces are planes 0 .. n_recv-1 of the
// level's ghost plan, in plan order: the face table's FK_PLANE argument indexes them)
__device__ void do_fill_stored(const LevelDev& L, const ParamsDev& P, int iv, const int* bct, const double* bcv,
                               int b, int tid, int nthr, const double* rpl = nullptr, int n_recv = 0) {
  const int nc = L.nc;
  const double* v = L.var[iv];
  for (int idx = tid; idx < 6 * nc * nc; idx += nthr) {
    const int f = idx / (nc * nc), r = idx - f * nc * nc, t1 = r % nc, t2 = r / nc;
    const int kind = L.fkind[b * 6 + f], arg = L.farg[b * 6 + f];
    double* out = L.faces[iv] + ((size_t)b * 6 + f) * nc * nc + r;
    int i1, j1, k1, i2, j2, k2;
    const int n1 = (f & 1) ? nc - 1 : 0, n2 = (f & 1) ? nc - 2 : 1;
    if (f < 2) { i1 = n1; i2 = n2; j1 = j2 = t1; k1 = k2 = t2; }
    else if (f < 4) { j1 = n1; j2 = n2; i1 = i2 = t1; k1 = k2 = t2; }
    else { k1 = n1; k2 = n2; i1 = i2 = t1; j1 = j2 = t2; }
    if (kind == FK_SAME) {
      // neighbour's interior layer facing me
      int ii = i1, jj = j1, kk = k1;
      const int m = (f & 1) ? 0 : nc - 1;
      if (f < 2) ii = m; else if (f < 4) jj = m; else kk = m;
      *out = v[(size_t)arg * L.bs + cidx(L, ii, jj, kk)];
    } else if (kind == FK_PHYS) {
      const int type = bct[f];
      double c0, c1, c2;
      if (type == -10) { c0 = 2; c1 = -1; c2 = 0; }
      else if (type == -11) { c0 = L.dr[f >> 1] * ((f & 1) ? 1.0 : -1.0); c1 = 1; c2 = 0; }
      else { c0 = 0; c1 = 2; c2 = -1; }
      const double x1 = v[(size_t)b * L.bs + cidx(L, i1, j1, k1)];
      const double x2 = nc > 1 ? v[(size_t)b * L.bs + cidx(L, i2, j2, k2)] : 0.0;
      *out = c0 * bcv[f] + c1 * x1 + c2 * x2;
    } else if (kind == FK_PLANE && rpl && arg < n_recv) {
      *out = rpl[(size_t)arg * nc * nc + r];
    }
  }
}

// ---- "simple" levels of the tail: power-of-two boxes whose faces are either physical boundaries (constant values or
// stored planes) or local same-level neighbours (LevelDev::simple; the coarse end of every uniform grid).  The general path
// classifies each of the six neighbours of each cell through cell_at (~900 instructions per cell, ncu r01r); here the
// six in-box neighbours are loaded with clamped indices and the out-of-box one is replaced by the neighbour box's face
// cell or by the ghost value c0*b + c1*phi(1) + c2*phi(2) of cell_at / bc_to_gc (m_ghost_cells.f90:665-766), where
// phi(1) is the cell itself and phi(2) its in-box neighbour on the other side.  All five operators (eps ghost faces from the stored faces pool), GSRB only.
// tests/device_host compiles these functions on the host and checks them bit for bit against the general ones.
struct SimpleBc { double c0[6], c1[6], c2[6], bv[6]; };
__device__ __forceinline__ void simple_bc(const LevelDev& L, const ParamsDev& P, int b, SimpleBc& B) {
#pragma unroll
  for (int f = 0; f < 6; ++f) {
    int type = L.ftype[b * 6 + f];
    if (type == 0) type = P.bc_type[f];
    double c0, c1, c2;
    if (type == -10) { c0 = 2; c1 = -1; c2 = 0; }
    else if (type == -11) { c0 = L.dr[f >> 1] * ((f & 1) ? 1.0 : -1.0); c1 = 1; c2 = 0; }
    else { c0 = 0; c1 = 2; c2 = -1; }
    B.c0[f] = c0; B.c1[f] = c1; B.c2[f] = c2; B.bv[f] = P.bc_value[f];
  }
}
// u[0..5] of cell (colour, idx) of box b of variable `var`; u0 = the cell's own value; kind / arg = the variable's
// face tables (fkind / farg for phi, fkind_old / farg_old for old)
__device__ __forceinline__ void simple_gather(const LevelDev& L, const SimpleBc& B, const int* __restrict__ kind,
                                              const int* __restrict__ arg, const double* __restrict__ var, int b, int colour,
                                              int idx, double u0, double* u) {
  const int nc = L.nc, hx = L.hx;
  const int m = idx & (hx - 1), r = idx >> L.sh_hx, j = r & (nc - 1), k = r >> L.sh_nc;
  const int s = colour ^ ((j + k) & 1);
  const int ocol = (colour ^ 1) * L.cs;
  const double* oth = var + (size_t)b * L.bs + ocol;
  const int ml = m - 1 + s, mh = m + s;
  const int row = (k * nc + j) * hx;
  const double xl = oth[row + (ml < 0 ? 0 : ml)], xh = oth[row + (mh >= hx ? hx - 1 : mh)];
  const double yl = oth[(k * nc + (j > 0 ? j - 1 : 0)) * hx + m], yh = oth[(k * nc + (j < nc - 1 ? j + 1 : j)) * hx + m];
  const double zl = oth[((k > 0 ? k - 1 : 0) * nc + j) * hx + m], zh = oth[((k < nc - 1 ? k + 1 : k) * nc + j) * hx + m];
  // out-of-box neighbour through face f: the cell at `pos` of the neighbour box's other-colour block, or the ghost value
  // with the boundary value b from the constant or from the stored plane (callback / mg_phi_bc_store values) at (t1, t2)
  auto outside = [&](int f, double x2, int pos, int t1, int t2) -> double {
    const int q = b * 6 + f;
    const int a = arg[q];
    if (kind[q] == FK_SAME) return var[(size_t)a * L.bs + ocol + pos];
    const double bv = a >= 0 ? L.bcpl[(size_t)a * nc * nc + t2 * nc + t1] : B.bv[f];
    return B.c0[f] * bv + B.c1[f] * u0 + B.c2[f] * x2;
  };
  const int i = 2 * m + s;
  u[0] = ml < 0 ? outside(0, xh, row + hx - 1, j, k) : xl;
  u[1] = mh >= hx ? outside(1, xl, row, j, k) : xh;
  u[2] = j == 0 ? outside(2, yh, (k * nc + nc - 1) * hx + m, i, k) : yl;
  u[3] = j == nc - 1 ? outside(3, yl, (k * nc) * hx + m, i, k) : yh;
  u[4] = k == 0 ? outside(4, zh, ((nc - 1) * nc + j) * hx + m, i, j) : zl;
  u[5] = k == nc - 1 ? outside(5, zl, j * hx + m, i, j) : zh;
}
// the six neighbours of a variable whose ghost faces are stored (eps: GM_STORED, gather_eps): in-box cells of the other
// colour, out-of-box values from the faces pool [box][face][t2][t1]
__device__ __forceinline__ void simple_gather_stored(const LevelDev& L, const double* __restrict__ var, const double* __restrict__ faces,
                                                     int b, int colour, int idx, double* a) {
  const int nc = L.nc, hx = L.hx;
  const int m = idx & (hx - 1), r = idx >> L.sh_hx, j = r & (nc - 1), k = r >> L.sh_nc;
  const int s = colour ^ ((j + k) & 1), i = 2 * m + s;
  const double* oth = var + (size_t)b * L.bs + (colour ^ 1) * L.cs;
  const double* fc = faces + (size_t)b * 6 * nc * nc;
  const int ml = m - 1 + s, mh = m + s;
  const int row = (k * nc + j) * hx;
  a[0] = ml < 0 ? fc[k * nc + j] : oth[row + ml];
  a[1] = mh >= hx ? fc[nc * nc + k * nc + j] : oth[row + mh];
  a[2] = j == 0 ? fc[2 * nc * nc + k * nc + i] : oth[(k * nc + j - 1) * hx + m];
  a[3] = j == nc - 1 ? fc[3 * nc * nc + k * nc + i] : oth[(k * nc + j + 1) * hx + m];
  a[4] = k == 0 ? fc[4 * nc * nc + j * nc + i] : oth[((k - 1) * nc + j) * hx + m];
  a[5] = k == nc - 1 ? fc[5 * nc * nc + j * nc + i] : oth[((k + 1) * nc + j) * hx + m];
}
// eps coefficients of one cell in face order (gather_eps): a0 = the cell's own, a = the neighbour's
template <int OP>
__device__ __forceinline__ void simple_eps(const LevelDev& L, int b, int colour, int idx, double* a0, double* a) {
  const size_t g = (size_t)b * L.bs + colour * L.cs + idx;
  if (OP == OP_AHELMH) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      double t[6];
      simple_gather_stored(L, L.var[4 + d], L.faces[4 + d], b, colour, idx, t);
      const double c = L.var[4 + d][g];
      a0[2 * d] = c; a0[2 * d + 1] = c; a[2 * d] = t[2 * d]; a[2 * d + 1] = t[2 * d + 1];
    }
  } else {
    const double c = L.var[4][g];
#pragma unroll
    for (int f = 0; f < 6; ++f) a0[f] = c;
    simple_gather_stored(L, L.var[4], L.faces[4], b, colour, idx, a);
  }
}
// one GSRB colour half-sweep over all boxes (do_smooth)
template <int OP>
__device__ void simple_smooth(const LevelDev& L, const ParamsDev& P, int colour, int tid, int nthr) {
  double* phi = L.var[0];
  const double* __restrict__ rhs = L.var[1];
  for (int b = 0; b < L.nb; ++b) {
    SimpleBc B; simple_bc(L, P, b, B);
    for (int idx = tid; idx < L.cs; idx += nthr) {
      const size_t g = (size_t)b * L.bs + colour * L.cs + idx;
      double u[6];
      simple_gather(L, B, L.fkind, L.farg, phi, b, colour, idx, phi[g], u);
      if (OP == OP_LPL || OP == OP_HELMH) phi[g] = gs_lpl_value(L.idr2, L.fac, u, rhs[g]);
      else {
        double a0[6], a[6];
        simple_eps<OP>(L, b, colour, idx, a0, a);
        phi[g] = gs_vlpl_value(L.idr2, u, a0, a, rhs[g], P.lambda, OP != OP_VLPL);
      }
    }
  }
}
template <int OP>
__device__ __forceinline__ double simple_op(const LevelDev& L, const ParamsDev& P, const SimpleBc& B, int b, int q) {
  const int colour = q >= L.cs ? 1 : 0;
  const double u0 = L.var[0][(size_t)b * L.bs + q];
  double u[6];
  simple_gather(L, B, L.fkind, L.farg, L.var[0], b, colour, q - colour * L.cs, u0, u);
  if (OP == OP_LPL) return lpl_value(L.idr2, u0, u);
  if (OP == OP_HELMH) return lpl_value(L.idr2, u0, u) - P.lambda * u0;
  double a0[6], a[6];
  simple_eps<OP>(L, b, colour, q - colour * L.cs, a0, a);
  double v = vlpl_value(L.idr2, u0, u, a0, a);
  if (OP != OP_VLPL) v = v - P.lambda * u0;
  return v;
}
// residual_box on all boxes (do_resid); returns the thread-local max |res|
template <int OP>
__device__ double simple_resid(const LevelDev& L, const ParamsDev& P, int tid, int nthr) {
  double m = 0.0;
  for (int b = 0; b < L.nb; ++b) {
    SimpleBc B; simple_bc(L, P, b, B);
    for (int q = tid; q < L.bs; q += nthr) {
      const size_t g = (size_t)b * L.bs + q;
      const double r = L.var[1][g] - simple_op<OP>(L, P, B, b, q);
      L.var[3][g] = r;
      m = fmax(m, fabs(r));
    }
  }
  return m;
}
// tail of update_coarse on one parent box (do_crhs)
template <int OP>
__device__ void simple_crhs(const LevelDev& L, const ParamsDev& P, int b, int tid, int nthr) {
  SimpleBc B; simple_bc(L, P, b, B);
  for (int q = tid; q < L.bs; q += nthr) {
    const size_t g = (size_t)b * L.bs + q;
    L.var[1][g] = simple_op<OP>(L, P, B, b, q) + L.var[3][g];
    L.var[2][g] = L.var[0][g];
  }
}
// correct_children + mg_prolong_sparse onto the fine box b from its (local) parent on a simple coarse level
// (do_corr_prolong): the correction phi_c - old_c at a coarse cell and at its six neighbours, out-of-box values of phi
// and old through their own face tables
__device__ void simple_corr_prolong(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int b, int tid, int nthr) {
  const int p = Lf.parent[b];
  if (p < 0) return;
  const int off = Lf.coff[b];
  const int ox = off & 1023, oy = (off >> 10) & 1023, oz = (off >> 20) & 1023;
  SimpleBc B; simple_bc(Lc, P, p, B);
  const double* __restrict__ pc = Lc.var[0];
  const double* __restrict__ oc = Lc.var[2];
  double* phi = Lf.var[0] + (size_t)b * Lf.bs;
  const int h = Lf.nc >> 1, sh = Lf.sh_nc - 1;
  for (int idx = tid; idx < h * h * h; idx += nthr) {
    const int I = idx & (h - 1), J = (idx >> sh) & (h - 1), K = idx >> (2 * sh);
    const int ci = ox + I, cj = oy + J, ck = oz + K;
    const int colour = (ci + cj + ck) & 1, idxc = (ck * Lc.nc + cj) * Lc.hx + (ci >> 1);
    const size_t g = (size_t)p * Lc.bs + colour * Lc.cs + idxc;
    const double u0 = pc[g], o0 = oc[g];
    double up[6], uo[6];
    simple_gather(Lc, B, Lc.fkind, Lc.farg, pc, p, colour, idxc, u0, up);
    simple_gather(Lc, B, Lc.fkind_old, Lc.farg_old, oc, p, colour, idxc, o0, uo);
    const double c = u0 - o0;
    Lc.var[3][g] = c;
    const double f0 = 0.25 * c;
    const double flx = 0.25 * (up[0] - uo[0]), fhx = 0.25 * (up[1] - uo[1]);
    const double fly = 0.25 * (up[2] - uo[2]), fhy = 0.25 * (up[3] - uo[3]);
    const double flz = 0.25 * (up[4] - uo[4]), fhz = 0.25 * (up[5] - uo[5]);
#pragma unroll
    for (int kk = 0; kk < 2; ++kk)
#pragma unroll
      for (int jj = 0; jj < 2; ++jj)
#pragma unroll
        for (int ii = 0; ii < 2; ++ii) {
          const double fine = f0 + (ii ? fhx : flx) + (jj ? fhy : fly) + (kk ? fhz : flz);
          const int q = cidx(Lf, 2 * I + ii, 2 * J + jj, 2 * K + kk);
          phi[q] = phi[q] + fine;
        }
  }
}

// block-wide max of non-negative doubles; result valid in thread 0 (and broadcast through smem)
__device__ double block_max(double v, double* smem /* >= 32 doubles */) {
  for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  const int w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) smem[w] = v;
  __syncthreads();
  if (w == 0) {
    v = (int)threadIdx.x < nw ? smem[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (threadIdx.x == 0) smem[0] = v;
  }
  __syncthreads();
  v = smem[0];
  __syncthreads();
  return v;
}
