// Host build of the product's per-cell device code (octree-mg_b200/csrc/omg_device.cuh, omg_ops.cuh) for CPU tests.
//
// Test infrastructure: g++ compiles the SAME headers the kernels are built from (CUDA qualifiers reduced to nothing,
// the few intrinsics stubbed below) with -ffp-contract=off, and tests/test_device_host.py checks on random levels that
// the tail's fast functions (simple_*) produce the bits of the general ones (do_*), i.e. of the path the GPU parity
// tests pin to the oracle.  One "thread" (tid = 0, nthr = 1) walks every loop; a GSRB colour half-sweep only reads the
// other colour and a cell's own value, so the sequential walk performs the same arithmetic as the parallel one.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include <cuda_runtime.h>   // type definitions only (g++ sees __device__ etc. as attributes it ignores)

// ---- stand-ins for device intrinsics -----------------------------------------------------------------------
#ifndef OMG_EMU_THREADS   // fast_emu.cpp (the CTA emulator) brings thread-aware versions of these
struct FakeDim { unsigned x = 0, y = 0, z = 0; };
static FakeDim threadIdx, blockIdx;
static FakeDim blockDim = {1, 1, 1};
static inline void __syncthreads() {}
static inline double __shfl_xor_sync(unsigned, double v, int) { return v; }
static inline unsigned long long atomicMax(unsigned long long* a, unsigned long long v) {
  const unsigned long long o = *a; if (v > o) *a = v; return o;
}
#endif
static inline int __double2hiint(double d) { int64_t b; std::memcpy(&b, &d, 8); return (int)(b >> 32); }
static inline double __hiloint2double(int hi, int lo) {
  const int64_t b = ((int64_t)hi << 32) | (uint32_t)lo; double d; std::memcpy(&d, &b, 8); return d;
}
static inline float __int_as_float(int i) { float f; std::memcpy(&f, &i, 4); return f; }
static inline long long __double_as_longlong(double d) { long long b; std::memcpy(&b, &d, 8); return b; }
static inline double __ldcg(const double* p) { return *p; }

#include "../../octree-mg_b200/csrc/omg_ops.cuh"

namespace {

// a level of nbx x nby x nbz boxes of nc^3 cells in the colour-split layout, boxes in x-fastest order, physical
// boundaries with constant values around the brick, FK_SAME between its boxes
struct HostLevel {
  LevelDev L;
  std::vector<double> var[4];
  std::vector<double> eps[3], epsf[3];   // variables 5..7 (eps) and their stored ghost faces [box][6][nc*nc]
  std::vector<int> fkind, farg, ftype, fkind_old, farg_old, parent, coff;
  std::vector<double> bcpl, gpl;
};
int g_planes = 0;   // 2: physical faces become FK_PLANE ghost planes (as behind refinement / partition boundaries); 1: physical faces carry stored planes of boundary values (callbacks / mg_phi_bc_store) instead of constants

void build_level(HostLevel& H, int nc, int nbx, int nby, int nbz, const int* bc_types, double dr) {
  LevelDev& L = H.L;
  std::memset(&L, 0, sizeof(L));
  const int nb = nbx * nby * nbz;
  L.nc = nc; L.hx = (nc + 1) / 2; L.nb = nb; L.cs = nc * nc * L.hx; L.bs = 2 * L.cs; L.lvl = 0;
  L.sh_nc = L.sh_hx = -1;
  if (nc >= 2 && (nc & (nc - 1)) == 0) { L.sh_nc = 0; while ((1 << L.sh_nc) < nc) ++L.sh_nc; L.sh_hx = L.sh_nc - 1; }
  for (int d = 0; d < 3; ++d) { L.dr[d] = dr * (1.0 + 0.25 * d); L.idr2[d] = 1.0 / (L.dr[d] * L.dr[d]); }
  for (int v = 0; v < 4; ++v) { H.var[v].assign((size_t)nb * L.bs, 0.0); L.var[v] = H.var[v].data(); }
  for (int d = 0; d < 3; ++d) {
    H.eps[d].assign((size_t)nb * L.bs, 1.0); H.epsf[d].assign((size_t)nb * 6 * nc * nc, 1.0);
    L.var[4 + d] = H.eps[d].data(); L.faces[4 + d] = H.epsf[d].data();
  }
  H.fkind.assign(nb * 6, FK_PHYS); H.farg.assign(nb * 6, -1); H.ftype.assign(nb * 6, 0);
  H.parent.assign(nb, -1); H.coff.assign(nb, 0);
  for (int bz = 0; bz < nbz; ++bz) for (int by = 0; by < nby; ++by) for (int bx = 0; bx < nbx; ++bx) {
    const int b = (bz * nby + by) * nbx + bx;
    const int pos[3] = {bx, by, bz}, ext[3] = {nbx, nby, nbz}, stride[3] = {1, nbx, nbx * nby};
    for (int f = 0; f < 6; ++f) {
      const int d = f >> 1, q = pos[d] + ((f & 1) ? 1 : -1);
      if (q >= 0 && q < ext[d]) { H.fkind[b * 6 + f] = FK_SAME; H.farg[b * 6 + f] = b + ((f & 1) ? stride[d] : -stride[d]); }
      else {
        H.ftype[b * 6 + f] = bc_types[f];
        if (g_planes == 2) {
          H.fkind[b * 6 + f] = FK_PLANE;
          H.farg[b * 6 + f] = (int)(H.gpl.size() / ((size_t)nc * nc));
          for (int t = 0; t < nc * nc; ++t) H.gpl.push_back(0.02 * (double)(((b * 6 + f) * 53 + t * 11) % 89) - 0.7);
        } else if (g_planes) {
          H.farg[b * 6 + f] = (int)(H.bcpl.size() / ((size_t)nc * nc));
          for (int t = 0; t < nc * nc; ++t) H.bcpl.push_back(0.01 * (double)((b * 6 + f) * 131 + t * 7 % 97) - 0.3);
        }
      }
    }
  }
  L.bcpl = H.bcpl.data();
  L.gpl = L.gpl_old = H.gpl.data();
  H.fkind_old = H.fkind; H.farg_old = H.farg;
  L.fkind = H.fkind.data(); L.farg = H.farg.data(); L.ftype = H.ftype.data();
  L.fkind_old = H.fkind_old.data(); L.farg_old = H.farg_old.data();
  L.parent = H.parent.data(); L.coff = H.coff.data();
  L.simple = 1;
}

struct Rng {   // xorshift: reproducible values in (-1, 1)
  uint64_t s;
  double next() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (double)(int64_t)(s >> 11) / (double)(1ll << 52) - 1.0; }
};
void randomise(HostLevel& H, uint64_t seed) {
  Rng r{seed * 2654435761u + 88172645463325252ull};
  for (int v = 0; v < 4; ++v) for (double& x : H.var[v]) x = r.next();
  for (int d = 0; d < 3; ++d) {        // positive coefficients
    for (double& x : H.eps[d]) x = 1.5 + r.next();
    for (double& x : H.epsf[d]) x = 1.5 + r.next();
  }
}

ParamsDev params(int op, double lambda, const int* bc_types, const double* bc_values) {
  ParamsDev P;
  std::memset(&P, 0, sizeof(P));
  P.op = op; P.smoother = 2; P.lambda = lambda;
  for (int f = 0; f < 6; ++f) { P.bc_type[f] = bc_types[f]; P.bc_value[f] = bc_values[f]; }
  return P;
}

int differing(const HostLevel& A, const HostLevel& B) {
  int n = 0;
  for (int v = 0; v < 4; ++v) n += std::memcmp(A.var[v].data(), B.var[v].data(), A.var[v].size() * sizeof(double)) != 0;
  return n;
}

template <int OP>
int run_case(int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda, uint64_t seed) {
  HostLevel G, S;                    // general / simple
  build_level(G, nc, nbx, nby, nbz, bc_types, 0.01);
  build_level(S, nc, nbx, nby, nbz, bc_types, 0.01);
  ParamsDev P = params(OP, lambda, bc_types, bc_values);
  const double sum = G.L.idr2[0] + G.L.idr2[1] + G.L.idr2[2];
  G.L.fac = S.L.fac = (OP == OP_HELMH) ? 1.0 / (2 * sum + lambda) : 0.5 / sum;
  randomise(G, seed);
  for (int v = 0; v < 4; ++v) S.var[v] = G.var[v];
  for (int v = 0; v < 4; ++v) S.L.var[v] = S.var[v].data();
  for (int d = 0; d < 3; ++d) { S.eps[d] = G.eps[d]; S.epsf[d] = G.epsf[d]; S.L.var[4 + d] = S.eps[d].data(); S.L.faces[4 + d] = S.epsf[d].data(); }
  int bad = 0;
  // two full GSRB sweeps
  for (int n = 0; n < 4; ++n) {
    for (int b = 0; b < G.L.nb; ++b) do_smooth<OP>(G.L, P, b, n & 1, 0, 1);
    simple_smooth<OP>(S.L, P, n & 1, 0, 1);
    bad += differing(G, S) ? 1 : 0;
  }
  // residual (and its max)
  double mg = 0.0, ms;
  for (int b = 0; b < G.L.nb; ++b) mg = fmax(mg, do_resid<OP>(G.L, P, b, 0, 1));
  ms = simple_resid<OP>(S.L, P, 0, 1);
  bad += (std::memcmp(&mg, &ms, 8) != 0) + (differing(G, S) ? 1 : 0);
  // coarse right-hand side on every box
  for (int b = 0; b < G.L.nb; ++b) do_crhs<OP>(G.L, P, b, 0, 1);
  for (int b = 0; b < S.L.nb; ++b) simple_crhs<OP>(S.L, P, b, 0, 1);
  bad += differing(G, S) ? 1 : 0;
  return bad;
}

// correction + prolongation from a coarse level onto a fine level whose boxes are the octants (same box size) or the
// whole refined box (coarse box of half the size, the single-box coarsening levels)
int run_prolong(int ncf, int nbx, int nby, int nbz, bool half_size_coarse, const int* bc_types, const double* bc_values, uint64_t seed) {
  int bad = 0;
  for (int pass = 0; pass < 2; ++pass) {
    HostLevel F, C;
    if (half_size_coarse) { build_level(F, ncf, 1, 1, 1, bc_types, 0.01); build_level(C, ncf / 2, 1, 1, 1, bc_types, 0.02); }
    else { build_level(F, ncf, 2 * nbx, 2 * nby, 2 * nbz, bc_types, 0.01); build_level(C, ncf, nbx, nby, nbz, bc_types, 0.02); }
    for (int b = 0; b < F.L.nb; ++b) {
      if (half_size_coarse) { F.parent[b] = 0; F.coff[b] = 0; continue; }
      const int fx = b % (2 * nbx), fy = (b / (2 * nbx)) % (2 * nby), fz = b / (4 * nbx * nby);
      F.parent[b] = ((fz / 2) * nby + fy / 2) * nbx + fx / 2;
      const int h = ncf / 2;
      F.coff[b] = ((fx & 1) * h) | (((fy & 1) * h) << 10) | (((fz & 1) * h) << 20);
    }
    ParamsDev P = params(OP_LPL, 0.0, bc_types, bc_values);
    randomise(F, seed); randomise(C, seed + 17);
    static std::vector<double> keep[2][8];
    if (pass == 0) for (int b = 0; b < F.L.nb; ++b) do_corr_prolong(F.L, C.L, P, b, 0, 1);
    else for (int b = 0; b < F.L.nb; ++b) simple_corr_prolong(F.L, C.L, P, b, 0, 1);
    for (int v = 0; v < 4; ++v) {
      if (pass == 0) { keep[0][v] = F.var[v]; keep[1][v] = C.var[v]; }
      else bad += (keep[0][v] != F.var[v]) + (keep[1][v] != C.var[v]);
    }
  }
  return bad;
}

}  // namespace

extern "C" {
void dh_set_planes(int on) { g_planes = on; }
// number of mismatching comparisons (0 = the fast path reproduces the general path bit for bit)
int dh_compare_simple(int op, int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda,
                      unsigned long long seed) {
  switch (op) {
    case OP_LPL: return run_case<OP_LPL>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_VLPL: return run_case<OP_VLPL>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_HELMH: return run_case<OP_HELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    case OP_VHELMH: return run_case<OP_VHELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
    default: return run_case<OP_AHELMH>(nc, nbx, nby, nbz, bc_types, bc_values, lambda, seed);
  }
}
int dh_compare_prolong(int ncf, int nbx, int nby, int nbz, int half_size_coarse, const int* bc_types, const double* bc_values,
                       unsigned long long seed) {
  return run_prolong(ncf, nbx, nby, nbz, half_size_coarse != 0, bc_types, bc_values, seed);
}
// One operation of the GENERAL device path (do_smooth / do_resid, what the batched kernels and the tail run) on a brick
// of nbx x nby x nbz boxes between constant physical boundaries, for the comparison with the oracle on the CPU.
// Arrays: interior cells only, [box (x fastest)][k][j][i].  what = 2: one sweep of mg_smoother_gs; what = 0: one GSRB sweep (colour 0 = the cells the
// reference updates first, then colour 1) of phi; what = 1: res = rhs - L(phi), returns max |res| through *out_max.
int dh_level_op(int what, int op, int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda,
                const double* dr, double* phi, const double* rhs, double* res, double* out_max) {
  HostLevel H;
  build_level(H, nc, nbx, nby, nbz, bc_types, 1.0);
  LevelDev& L = H.L;
  for (int d = 0; d < 3; ++d) { L.dr[d] = dr[d]; L.idr2[d] = 1.0 / (dr[d] * dr[d]); }
  const double sum = L.idr2[0] + L.idr2[1] + L.idr2[2];
  L.fac = (op == OP_HELMH) ? 1.0 / (2 * sum + lambda) : 0.5 / sum;
  ParamsDev P = params(op, lambda, bc_types, bc_values);
  const size_t n3 = (size_t)nc * nc * nc;
  for (int b = 0; b < L.nb; ++b) for (int k = 0; k < nc; ++k) for (int j = 0; j < nc; ++j) for (int i = 0; i < nc; ++i) {
    const size_t src = b * n3 + ((size_t)k * nc + j) * nc + i, dst = (size_t)b * L.bs + cidx(L, i, j, k);
    H.var[0][dst] = phi[src]; H.var[1][dst] = rhs[src];
  }
  double m = 0.0;
  std::vector<double> snap;
  if (what == 0) {
    for (int colour = 0; colour < 2; ++colour)
      for (int b = 0; b < L.nb; ++b) {
        if (op == OP_LPL) do_smooth<OP_LPL>(L, P, b, colour, 0, 1); else do_smooth<OP_HELMH>(L, P, b, colour, 0, 1);
      }
  } else if (what == 2) {
    // mg_smoother_gs: ghost snapshot of every box (= the ghost fill before the sweep), then the in-box sweep over
    // hyperplanes i + j + k = w, which must equal the reference's lexicographic k, j, i loop
    snap.assign((size_t)L.nb * 6 * nc * nc, 0.0);
    L.faces[0] = snap.data();
    P.smoother = 1;
    for (int b = 0; b < L.nb; ++b) do_gs_snapshot(L, P, b, 0, 1);
    for (int b = 0; b < L.nb; ++b) {
      if (op == OP_LPL) do_gs_sweep<OP_LPL>(L, P, b, 0, 1); else do_gs_sweep<OP_HELMH>(L, P, b, 0, 1);
    }
  } else {
    for (int b = 0; b < L.nb; ++b)
      m = fmax(m, op == OP_LPL ? do_resid<OP_LPL>(L, P, b, 0, 1) : do_resid<OP_HELMH>(L, P, b, 0, 1));
  }
  for (int b = 0; b < L.nb; ++b) for (int k = 0; k < nc; ++k) for (int j = 0; j < nc; ++j) for (int i = 0; i < nc; ++i) {
    const size_t dst = b * n3 + ((size_t)k * nc + j) * nc + i, src = (size_t)b * L.bs + cidx(L, i, j, k);
    phi[dst] = H.var[0][src]; res[dst] = H.var[3][src];
  }
  if (out_max) *out_max = m;
  return 0;
}
// The same for the variable-coefficient operators (mg_vlaplacian, mg_vhelmholtz, mg_ahelmholtz): eps_cc holds the n_eps
// coefficient variables as the reference stores them, [variable][box][(nc+2)^3] with their ghost shell (host data); the
// device layout keeps their interior cells colour-split and their ghost cells as stored faces [box][6][nc*nc].
// what = 0: one GSRB sweep, 1: residual + its maximum.
}  // extern "C"
template <int OP>
double level_op_vc(int what, LevelDev& L, const ParamsDev& P) {
  double m = 0.0;
  if (what == 0) { for (int colour = 0; colour < 2; ++colour) for (int b = 0; b < L.nb; ++b) do_smooth<OP>(L, P, b, colour, 0, 1); }
  else for (int b = 0; b < L.nb; ++b) m = fmax(m, do_resid<OP>(L, P, b, 0, 1));
  return m;
}
extern "C" {
int dh_level_op_vc(int what, int op, int nc, int nbx, int nby, int nbz, const int* bc_types, const double* bc_values, double lambda,
                   const double* dr, double* phi, const double* rhs, double* res, const double* eps_cc, int n_eps, double* out_max) {
  HostLevel H;
  build_level(H, nc, nbx, nby, nbz, bc_types, 1.0);
  LevelDev& L = H.L;
  for (int d = 0; d < 3; ++d) { L.dr[d] = dr[d]; L.idr2[d] = 1.0 / (dr[d] * dr[d]); }
  ParamsDev P = params(op, lambda, bc_types, bc_values);
  const size_t n3 = (size_t)nc * nc * nc, n2 = nc + 2, g3 = n2 * n2 * n2;
  for (int b = 0; b < L.nb; ++b) for (int k = 0; k < nc; ++k) for (int j = 0; j < nc; ++j) for (int i = 0; i < nc; ++i) {
    const size_t src = b * n3 + ((size_t)k * nc + j) * nc + i, dst = (size_t)b * L.bs + cidx(L, i, j, k);
    H.var[0][dst] = phi[src]; H.var[1][dst] = rhs[src];
  }
  for (int v = 0; v < n_eps; ++v) for (int b = 0; b < L.nb; ++b) {
    const double* cc = eps_cc + ((size_t)v * L.nb + b) * g3;
    auto at = [&](int i, int j, int k) { return cc[((size_t)(k + 1) * n2 + (j + 1)) * n2 + (i + 1)]; };   // ghosted indices -1 .. nc
    for (int k = 0; k < nc; ++k) for (int j = 0; j < nc; ++j) for (int i = 0; i < nc; ++i) H.eps[v][(size_t)b * L.bs + cidx(L, i, j, k)] = at(i, j, k);
    for (int f = 0; f < 6; ++f) for (int t2 = 0; t2 < nc; ++t2) for (int t1 = 0; t1 < nc; ++t1) {
      const int g = (f & 1) ? nc : -1;
      const double val = f < 2 ? at(g, t1, t2) : f < 4 ? at(t1, g, t2) : at(t1, t2, g);
      H.epsf[v][((size_t)b * 6 + f) * nc * nc + (size_t)t2 * nc + t1] = val;
    }
  }
  double m = 0.0;
  switch (op) {
    case OP_VLPL: m = level_op_vc<OP_VLPL>(what, L, P); break;
    case OP_VHELMH: m = level_op_vc<OP_VHELMH>(what, L, P); break;
    default: m = level_op_vc<OP_AHELMH>(what, L, P); break;
  }
  for (int b = 0; b < L.nb; ++b) for (int k = 0; k < nc; ++k) for (int j = 0; j < nc; ++j) for (int i = 0; i < nc; ++i) {
    const size_t dst = b * n3 + ((size_t)k * nc + j) * nc + i, src = (size_t)b * L.bs + cidx(L, i, j, k);
    phi[dst] = H.var[0][src]; res[dst] = H.var[3][src];
  }
  if (out_max) *out_max = m;
  return 0;
}
// Two levels: a coarse level of ONE box of nc^3 and the fine level of its 2 x 2 x 2 children (same box size), constant
// physical boundaries.  Arrays [box][k][j][i], interior cells; fine boxes in x-fastest octant order.
//   what = 0: update_coarse (src/m_multigrid.f90:347-384) = residual on the fine level, restriction of phi and res,
//             rhs_c = L(phi_c) + res_c and old_c = phi_c on the parent        (do_resid, do_restrict, do_crhs)
//   what = 1: correct_children (:387-402) = res_c = phi_c - old_c, phi_f += prolongation of it (do_corr_prolong)
// simple != 0 runs the tail's fast functions where they exist.  f* / c* = fine / coarse phi, rhs, old, res (in and out).
int dh_two_level(int what, int simple, int op, int nc, const int* bc_types, const double* bc_values, double lambda, const double* dr_f,
                 double* f_phi, double* f_rhs, double* f_old, double* f_res, double* c_phi, double* c_rhs, double* c_old, double* c_res) {
  HostLevel F, C;
  build_level(F, nc, 2, 2, 2, bc_types, 1.0);
  build_level(C, nc, 1, 1, 1, bc_types, 1.0);
  for (int d = 0; d < 3; ++d) {
    F.L.dr[d] = dr_f[d]; F.L.idr2[d] = 1.0 / (dr_f[d] * dr_f[d]);
    C.L.dr[d] = 2 * dr_f[d]; C.L.idr2[d] = 1.0 / (C.L.dr[d] * C.L.dr[d]);
  }
  for (HostLevel* H : {&F, &C}) {
    const double sum = H->L.idr2[0] + H->L.idr2[1] + H->L.idr2[2];
    H->L.fac = (op == OP_HELMH) ? 1.0 / (2 * sum + lambda) : 0.5 / sum;
  }
  const int h = nc / 2;
  for (int b = 0; b < 8; ++b) { F.parent[b] = 0; F.coff[b] = ((b & 1) * h) | ((((b >> 1) & 1) * h) << 10) | ((((b >> 2) & 1) * h) << 20); }
  std::vector<int> parents_list{0};
  C.L.parents_list = parents_list.data(); C.L.n_parents = 1;
  ParamsDev P = params(op, lambda, bc_types, bc_values);
  const size_t n3 = (size_t)nc * nc * nc;
  double* fa[4] = {f_phi, f_rhs, f_old, f_res};
  double* ca[4] = {c_phi, c_rhs, c_old, c_res};
  auto move = [&](HostLevel& H, double** a, bool in) {
    for (int v = 0; v < 4; ++v) for (int b = 0; b < H.L.nb; ++b)
      for (int k = 0; k < nc; ++k) for (int j = 0; j < nc; ++j) for (int i = 0; i < nc; ++i) {
        double& dev = H.var[v][(size_t)b * H.L.bs + cidx(H.L, i, j, k)];
        double& nat = a[v][b * n3 + ((size_t)k * nc + j) * nc + i];
        if (in) dev = nat; else nat = dev;
      }
  };
  move(F, fa, true); move(C, ca, true);
  if (what == 0) {
    if (simple) { if (op == OP_LPL) simple_resid<OP_LPL>(F.L, P, 0, 1); else simple_resid<OP_HELMH>(F.L, P, 0, 1); }
    else for (int b = 0; b < 8; ++b) { if (op == OP_LPL) do_resid<OP_LPL>(F.L, P, b, 0, 1); else do_resid<OP_HELMH>(F.L, P, b, 0, 1); }
    for (int b = 0; b < 8; ++b) { do_restrict(F.L, C.L, 0, b, 0, 1); do_restrict(F.L, C.L, 3, b, 0, 1); }
    if (simple) { if (op == OP_LPL) simple_crhs<OP_LPL>(C.L, P, 0, 0, 1); else simple_crhs<OP_HELMH>(C.L, P, 0, 0, 1); }
    else { if (op == OP_LPL) do_crhs<OP_LPL>(C.L, P, 0, 0, 1); else do_crhs<OP_HELMH>(C.L, P, 0, 0, 1); }
  } else {
    for (int b = 0; b < 8; ++b) { if (simple) simple_corr_prolong(F.L, C.L, P, b, 0, 1); else do_corr_prolong(F.L, C.L, P, b, 0, 1); }
  }
  move(F, fa, false); move(C, ca, false);
  return 0;
}
// ---- prototype (NOT product code): temporal blocking of S GSRB half-sweeps in one pass over phi and rhs ------------
// DESIGN.md "next round" item: the finest-level smoother is at its per-launch HBM roofline, the way past it is to apply
// several half-sweeps while a tile streams through on-chip memory.  This host prototype fixes the dependency logic the
// CUDA kernel will need and checks it bit for bit against S separate half-sweeps of the general device path:
//   * xy tile of tx x ty cells plus a halo of S cells (clipped at the domain), streamed in +z;
//   * at step k half-sweep s (s = 1..S, colour (s-1)&1) is applied IN PLACE to plane k-s+1 on the region
//     tile +- (S-s), stages in increasing order within a step: plane k-s+2 then holds state s-1 and plane k-s has not yet
//     received stage s+1, which is exactly what stage s must read (a Gauss-Seidel update reads the other colour only);
//   * cells outside the region of their stage keep stale values that no later stage of this tile reads;
//   * boundary ghosts c0*b + c1*phi(1) + c2*phi(2) use the cell's own CURRENT value, so the in-place window gives
//     them for free; only the tile interior (stage S region) is written back.
// Window: planes k-S+1 .. k+1 (S + 1 planes; the prototype keeps S + 2) of (tx + 2S) x (ty + 2S) cells -- 5 planes of 56 x 56 doubles = 125 KB
// for S = 4 and 48 x 48 tiles.  Traffic per two full sweeps: phi in + out and rhs once, x (56/48)^2 = 27 B per cell
// against 48 B for four separate half-sweeps.
int dh_fused_proto(int nc, int nbx, int nby, int nbz, int tx, int ty, int S, const int* bc_types, const double* bc_values, unsigned long long seed) {
  HostLevel G;
  build_level(G, nc, nbx, nby, nbz, bc_types, 0.01);
  ParamsDev P = params(OP_LPL, 0.0, bc_types, bc_values);
  const double sum = G.L.idr2[0] + G.L.idr2[1] + G.L.idr2[2];
  G.L.fac = 0.5 / sum;
  randomise(G, seed);
  const int NX = nc * nbx, NY = nc * nby, NZ = nc * nbz;
  auto at = [&](std::vector<double>& v, int x, int y, int z) -> double& {
    const int b = ((z / nc) * nby + y / nc) * nbx + x / nc;
    return v[(size_t)b * G.L.bs + cidx(G.L, x % nc, y % nc, z % nc)];
  };
  // global copies of phi and rhs, then the reference result: S half-sweeps of the general path
  std::vector<double> phi((size_t)NX * NY * NZ), rhs(phi.size()), out(phi.size());
  auto gi = [&](int x, int y, int z) { return ((size_t)z * NY + y) * NX + x; };
  for (int z = 0; z < NZ; ++z) for (int y = 0; y < NY; ++y) for (int x = 0; x < NX; ++x) { phi[gi(x, y, z)] = at(G.var[0], x, y, z); rhs[gi(x, y, z)] = at(G.var[1], x, y, z); }
  for (int s = 1; s <= S; ++s) for (int b = 0; b < G.L.nb; ++b) do_smooth<OP_LPL>(G.L, P, b, (s - 1) & 1, 0, 1);
  // boundary coefficients per face (constant boundaries), as simple_bc
  double c0[6], c1[6], c2[6];
  for (int f = 0; f < 6; ++f) {
    const int type = bc_types[f];
    if (type == -10) { c0[f] = 2; c1[f] = -1; c2[f] = 0; }
    else if (type == -11) { c0[f] = G.L.dr[f >> 1] * ((f & 1) ? 1.0 : -1.0); c1[f] = 1; c2[f] = 0; }
    else { c0[f] = 0; c1[f] = 2; c2[f] = -1; }
  }
  const int H = S, N[3] = {NX, NY, NZ};        // stage 1 updates tile +- (S-1) and reads original values one cell further out
  for (int y0 = 0; y0 < NY; y0 += ty) for (int x0 = 0; x0 < NX; x0 += tx) {
    const int xa = std::max(0, x0 - H), xb = std::min(NX, x0 + tx + H), ya = std::max(0, y0 - H), yb = std::min(NY, y0 + ty + H);
    const int wx = xb - xa, wy = yb - ya;
    std::vector<double> win((size_t)(S + 2) * wx * wy);             // ring of S + 2 planes
    auto W = [&](int x, int y, int z) -> double& { return win[((size_t)((z + S + 2) % (S + 2)) * wy + (y - ya)) * wx + (x - xa)]; };
    for (int y = ya; y < yb; ++y) for (int x = xa; x < xb; ++x) W(x, y, 0) = phi[gi(x, y, 0)];
    for (int k = 0; k < NZ + S; ++k) {                                // step k: plane k + 1 enters (stage 1 at plane k reads it), plane k - S + 1 leaves
      if (k + 1 < NZ) for (int y = ya; y < yb; ++y) for (int x = xa; x < xb; ++x) W(x, y, k + 1) = phi[gi(x, y, k + 1)];
      for (int s = 1; s <= S; ++s) {
        const int z = k - s + 1, colour = (s - 1) & 1, h = S - s;
        if (z < 0 || z >= NZ) continue;
        for (int y = std::max(ya, y0 - h); y < std::min(yb, y0 + ty + h); ++y) for (int x = std::max(xa, x0 - h); x < std::min(xb, x0 + tx + h); ++x) {
          // colour 0 = cells with (x + y + z) even in 0-based box-local indices (boxes have even size: global parity is the same)
          if (((x + y + z) & 1) != colour) continue;
          const int p[3] = {x, y, z};
          const double u0 = W(x, y, z);
          double u[6];
          for (int f = 0; f < 6; ++f) {
            const int d = f >> 1, dir = (f & 1) ? 1 : -1;
            int q[3] = {x, y, z}; q[d] += dir;
            if (q[d] < 0 || q[d] >= N[d]) {                           // physical boundary: ghost from own value and the opposite neighbour
              int r[3] = {x, y, z}; r[d] -= dir;
              u[f] = c0[f] * bc_values[f] + c1[f] * u0 + c2[f] * W(r[0], r[1], r[2]);
            } else u[f] = W(q[0], q[1], q[2]);
          }
          (void)p;
          W(x, y, z) = gs_lpl_value(G.L.idr2, G.L.fac, u, rhs[gi(x, y, z)]);
        }
      }
      const int zo = k - S + 1;                                       // plane that has received all S stages
      if (zo >= 0 && zo < NZ) for (int y = y0; y < std::min(NY, y0 + ty); ++y) for (int x = x0; x < std::min(NX, x0 + tx); ++x) out[gi(x, y, zo)] = W(x, y, zo);
    }
  }
  int bad = 0;
  for (int z = 0; z < NZ; ++z) for (int y = 0; y < NY; ++y) for (int x = 0; x < NX; ++x) {
    const double a = out[gi(x, y, z)], b = at(G.var[0], x, y, z);
    bad += std::memcmp(&a, &b, 8) != 0;
  }
  return bad;
}
// the batched division against the compiler's (host build: exact-reciprocal seed instead of MUFU.RCP64H)
// tolerance mode of the coefficient kernels: the six reciprocals by batch inversion (omg_device.cuh recip6) and the two
// per-cell functions built on it, against the exact ones; returns the number of groups that fell back to exact divisions
int dh_recip6(const double* d, double* out, int n) {
  int fallbacks = 0;
  for (int i = 0; i + 6 <= n; i += 6) if (!recip6(d + i, out + i)) { ++fallbacks; for (int f = 0; f < 6; ++f) out[i + f] = 1.0 / d[i + f]; }
  return fallbacks;
}
void dh_coeff_values(const double* idr2, const double* u, const double* a0, const double* a, double u0, double rhs, double lambda, double* out) {
  out[0] = vlpl_value(idr2, u0, u, a0, a);          out[1] = vlpl_value_tol(idr2, u0, u, a0, a);
  out[2] = gs_vlpl_value(idr2, u, a0, a, rhs, lambda, true); out[3] = gs_vlpl_value_tol(idr2, u, a0, a, rhs, lambda, true);
}
int dh_ddiv(const double* a, const double* b, double* out, int n) {
  for (int i = 0; i + 6 <= n; i += 6) ddiv_batch<6>(a + i, b + i, out + i);
  return 0;
}
}
