// omg_tail.cuh -- the single-CTA "tail" interpreter: the coarse end of the cycle (every level of at most tail_cells
// cells, incl. the coarse solve) as one launch of 1024 threads that walks an op list.  Device code only -- the launcher is
// omgk::tail in omg_kernels.cu -- so that tests/device_host/fast_emu.cpp can run the same source on the host.
#pragma once
#include "omg_kernels.h"
#include "omg_ops.cuh"

#define OMG_SWITCH_OP(op, CALL)                      \
  switch (op) {                                      \
    case OP_LPL: { constexpr int OP = OP_LPL; CALL; } break;       \
    case OP_VLPL: { constexpr int OP = OP_VLPL; CALL; } break;     \
    case OP_HELMH: { constexpr int OP = OP_HELMH; CALL; } break;   \
    case OP_VHELMH: { constexpr int OP = OP_VHELMH; CALL; } break; \
    default: { constexpr int OP = OP_AHELMH; CALL; } break;        \
  }

#ifndef OMG_TAIL_DYN_SMEM
#define OMG_TAIL_DYN_SMEM(name) extern __shared__ __align__(16) double name[]
#endif

namespace {

// Who runs the interpreter: one CTA (k_tail), or a thread-block cluster of CTAs that share the levels through L2 and meet
// at the hardware cluster barrier (k_tail_cluster: the levels of up to tail_cluster_cells cells, i.e. also the 8- and
// 64-box levels of a box-16 hierarchy, whose operations are launch-latency bound as separate kernels).
struct CtaTeam {
  static constexpr bool warp_solve = true;
  __device__ static int tid() { return threadIdx.x; }
  __device__ static int nthr() { return blockDim.x; }
  __device__ static void sync() { __syncthreads(); }
  __device__ static double max(double v, double* red, double*) { return block_max(v, red); }
};
#ifndef OMG_HOST_EMU
struct ClusterTeam {
  static constexpr bool warp_solve = false;
  __device__ static unsigned rank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
  __device__ static unsigned size() { unsigned r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
  __device__ static int tid() { return (int)(rank() * blockDim.x + threadIdx.x); }
  __device__ static int nthr() { return (int)(size() * blockDim.x); }
  // release / acquire at cluster scope: global-memory writes of every CTA before the barrier are visible after it
  __device__ static void sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  }
  // cluster-wide max through a scratch array in global memory (one slot per CTA); the same value on every thread
  __device__ static double max(double v, double* red, double* scratch) {
    const double m = block_max(v, red);
    if (threadIdx.x == 0) scratch[rank()] = m;
    sync();
    double r = 0.0;
    const unsigned n = size();
    for (unsigned q = 0; q < n; ++q) r = fmax(r, __ldcg(scratch + q));
    sync();
    return r;
  }
};
#endif

// ---- coarse solve of a coarsest level of up to FOUR 2^3 boxes (one box: every cubic hierarchy; two or four: domains with
// extents 2 : 1 : 1 or 2 : 2 : 1) by ONE warp: the coarse loop of mg_fas_vcycle (src/m_multigrid.f90:202-208) runs 5-6 times
// per cycle for the Poisson problem and 40-90 times for a Helmholtz / Neumann problem whose constant mode converges slowly;
// as interpreter operations each colour half-sweep and each residual of these few cells costs a 1024-thread barrier and a
// pass through shared memory.  Here lane l keeps cell l & 7 = colour * 4 + k * 2 + j (hx = 1: i = colour ^ ((j + k) & 1)) of
// box l >> 3 in a register; the in-box neighbours in x / y / z are lanes l ^ 4 / l ^ 5 / l ^ 6, the out-of-box one is the
// face cell of the neighbouring box (another lane) or the ghost value of simple_gather's `outside` (same expression, same
// operand order); lanes beyond the last box compute replicas of box 0 that are never stored.  No barrier, no memory
// traffic inside the loop; the results (phi, res, iteration count) are those of the interpreter's loop bit for bit.
__device__ __forceinline__ bool warp_coarse_ok(const LevelDev& L) {
  if (L.nc != 2 || L.nb < 1 || L.nb > 4) return false;
  for (int q = 0; q < 6 * L.nb; ++q) {
    const int kd = L.fkind[q];
    if (kd == FK_SAME) { if (L.farg[q] < 0 || L.farg[q] >= L.nb) return false; }
    else if (kd != FK_PHYS) return false;
  }
  return true;
}
template <int OP>
__device__ void warp_coarse_solve(const LevelDev& L, const ParamsDev& P, int n_sub, int max_it, double* scal) {
  constexpr bool VC = (OP != OP_LPL && OP != OP_HELMH);
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const bool act = lane < 8 * L.nb;
  const int bx = act ? lane >> 3 : 0;
  const int colour = (lane >> 2) & 1, idx = lane & 3, j = idx & 1, k = idx >> 1, i = colour ^ ((j + k) & 1);
  const int g = bx * 8 + colour * 4 + idx;
  SimpleBc B; simple_bc(L, P, bx, B);
  // the face through which this cell looks out of its box, per direction: a neighbouring box (the lane that holds the cell
  // across the face: same j, k for x; row nc - 1 - j for y; plane nc - 1 - k for z; always of the other colour) or a
  // physical boundary with its boundary value (constant or stored plane)
  const int fx = i, fy = 2 + j, fz = 4 + k;
  double c0[3], c1[3], c2[3], bv[3];
  int src[3];
  bool same[3];
#pragma unroll
  for (int f = 0; f < 6; ++f) {
    const int d = f >> 1;
    if (f == (d == 0 ? fx : d == 1 ? fy : fz)) {
      const int q = bx * 6 + f, a = L.farg[q];
      const int t1 = d == 0 ? j : i, t2 = d == 2 ? j : k;
      same[d] = L.fkind[q] == FK_SAME;
      const int pos = d == 0 ? idx : d == 1 ? (k * 2 + (1 - j)) : ((1 - k) * 2 + j);
      src[d] = same[d] ? a * 8 + (colour ^ 1) * 4 + pos : lane;
      c0[d] = B.c0[f]; c1[d] = B.c1[f]; c2[d] = B.c2[f];
      bv[d] = (!same[d] && a >= 0) ? L.bcpl[(size_t)a * 4 + t2 * 2 + t1] : B.bv[f];
    }
  }
  double a0[6], a[6];
  if (VC) simple_eps<OP>(L, bx, colour, idx, a0, a);
  double phi = L.var[0][g];
  const double rhs = L.var[1][g];
  auto gather = [&](double* u) {
    const double x = __shfl_xor_sync(full, phi, 4), y = __shfl_xor_sync(full, phi, 5), z = __shfl_xor_sync(full, phi, 6);
    const double ox = __shfl_sync(full, phi, src[0]), oy = __shfl_sync(full, phi, src[1]), oz = __shfl_sync(full, phi, src[2]);
    const double gx = same[0] ? ox : c0[0] * bv[0] + c1[0] * phi + c2[0] * x;
    const double gy = same[1] ? oy : c0[1] * bv[1] + c1[1] * phi + c2[1] * y;
    const double gz = same[2] ? oz : c0[2] * bv[2] + c1[2] * phi + c2[2] * z;
    u[0] = i ? x : gx; u[1] = i ? gx : x;
    u[2] = j ? y : gy; u[3] = j ? gy : y;
    u[4] = k ? z : gz; u[5] = k ? gz : z;
  };
  double r = 0.0;
  auto resid = [&]() -> double {
    double u[6];
    gather(u);
    double v;
    if (OP == OP_LPL) v = lpl_value(L.idr2, phi, u);
    else if (OP == OP_HELMH) v = lpl_value(L.idr2, phi, u) - P.lambda * phi;
    else { v = vlpl_value(L.idr2, phi, u, a0, a); if (OP != OP_VLPL) v = v - P.lambda * phi; }
    r = rhs - v;
    double m = act ? fabs(r) : 0.0;
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(full, m, o));
    return m;
  };
  const double init_res = resid();
  int it = 0;
  for (int n_it = 1; n_it <= max_it; ++n_it) {
    for (int n = 1; n <= n_sub; ++n) {
      double u[6];
      gather(u);
      double v;
      if (!VC) v = gs_lpl_value(L.idr2, L.fac, u, rhs);
      else v = gs_vlpl_value(L.idr2, u, a0, a, rhs, P.lambda, OP != OP_VLPL);
      if (colour == ((n - 1) & 1)) phi = v;
    }
    const double res = resid();
    it = n_it;
    if (res < scal[SCAL_COARSE_REL] * init_res || res < scal[SCAL_COARSE_ABS]) break;
  }
  if (act) { L.var[0][g] = phi; L.var[3][g] = r; }
  if (lane == 0) scal[SCAL_COARSE_ITERS] = (double)it;
}

// ------------------------------------------------------------------ tail interpreter ---------
template <int OP, class TEAM = CtaTeam>
__device__ void tail_smooth(const LevelDev& L, const ParamsDev& P, int colour, int tid, int nthr) {
  if (L.simple && P.smoother == 2) {
    simple_smooth<OP>(L, P, colour, tid, nthr);
  } else if (P.smoother == 1) {   // mg_smoother_gs
    for (int b = 0; b < L.nb; ++b) do_gs_snapshot(L, P, b, tid, nthr);
    TEAM::sync();
    for (int b = 0; b < L.nb; ++b) do_gs_sweep<OP>(L, P, b, tid, nthr);     // (single CTA only: synchronises inside)
  } else if (L.nc & 1) {   // odd box size: Jacobi between boxes, see do_smooth_staged
    for (int b = 0; b < L.nb; ++b) do_smooth_staged<OP>(L, P, b, colour, 0, tid, nthr);
    TEAM::sync();
    for (int b = 0; b < L.nb; ++b) do_smooth_staged<OP>(L, P, b, colour, 1, tid, nthr);
  } else {
    for (int b = 0; b < L.nb; ++b) do_smooth<OP>(L, P, b, colour, tid, nthr);
  }
}

template <int OP, class TEAM = CtaTeam>
__device__ void tail_run(const LevelDev* levels, const ParamsDev& P, const TailOp* __restrict__ ops, int n_ops,
                         double* scal, double* red) {
  const int tid = TEAM::tid(), nthr = TEAM::nthr();
  double* scratch = scal + SCAL_TEAM;
  for (int o = 0; o < n_ops; ++o) {
    const TailOp op = ops[o];
    const LevelDev& L = levels[op.l];
    // boxes between physical boundaries and local neighbours: neighbour gathers without
    // per-cell face classification (omg_ops.cuh)
    const bool fast = L.simple && P.smoother == 2;
    switch (op.type) {
      case TOP_SMOOTH:
        tail_smooth<OP, TEAM>(L, P, op.a, tid, nthr);
        break;
      case TOP_RESID: {
        double m = 0.0;
        if (fast) m = simple_resid<OP>(L, P, tid, nthr);
        else for (int b = 0; b < L.nb; ++b) m = fmax(m, do_resid<OP>(L, P, b, tid, nthr));
        if (op.a >= 0) { m = block_max(m, red); if (threadIdx.x == 0) atomic_max_nonneg(&scal[op.a], m); }   // one atomic per CTA
      } break;
      case TOP_RESTRICT: {
        const LevelDev& Lc = levels[op.l - 1];
        for (int b = 0; b < L.nb; ++b) { do_restrict(L, Lc, op.a, b, tid, nthr); if (op.b >= 0) do_restrict(L, Lc, op.b, b, tid, nthr); }
      } break;
      case TOP_CRHS:
        if (fast) for (int q = 0; q < L.n_parents; ++q) simple_crhs<OP>(L, P, L.parents_list[q], tid, nthr);
        else for (int q = 0; q < L.n_parents; ++q) do_crhs<OP>(L, P, L.parents_list[q], tid, nthr);
        break;
      case TOP_OLDPL:
        for (int e = 0; e < L.n_oldpl; ++e) do_oldpl(L, P, e, tid, nthr);
        break;
      case TOP_CORR_PROLONG: {
        const LevelDev& Lc = levels[op.l - 1];
        if (fast && Lc.simple) for (int b = 0; b < L.nb; ++b) simple_corr_prolong(L, Lc, P, b, tid, nthr);
        else for (int b = 0; b < L.nb; ++b) do_corr_prolong(L, Lc, P, b, tid, nthr);
      } break;
      case TOP_RB_FILL: {
        const LevelDev& Lc = levels[op.l - 1];
        for (int e = 0; e < L.n_rb; ++e) {
          if (op.a) do_rb_fill<GM_PHI_GZ>(L, Lc, P, 0, e, tid, nthr); else do_rb_fill<GM_PHI>(L, Lc, P, 0, e, tid, nthr);
        }
      } break;
      case TOP_COPY_OLD:
        for (size_t i = tid; i < (size_t)L.nb * L.bs; i += nthr) L.var[2][i] = L.var[0][i];
        break;
      case TOP_COARSE_SOLVE: {
        // coarse loop of mg_fas_vcycle, src/m_multigrid.f90:202-208
        if (TEAM::warp_solve && fast && P.warp_coarse && warp_coarse_ok(L)) {      // up to four 2^3 boxes: one warp, cells in registers
          if (threadIdx.x < 32) warp_coarse_solve<OP>(L, P, op.a, op.b, scal);
          break;
        }
        double m = 0.0;
        if (fast) m = simple_resid<OP>(L, P, tid, nthr);
        else for (int b = 0; b < L.nb; ++b) m = fmax(m, do_resid<OP>(L, P, b, tid, nthr));
        const double init_res = TEAM::max(m, red, scratch);
        int it = 0;
        for (int i = 1; i <= op.b; ++i) {
          for (int n = 1; n <= op.a; ++n) {
            tail_smooth<OP, TEAM>(L, P, (n - 1) & 1, tid, nthr);
            TEAM::sync();
          }
          m = 0.0;
          if (fast) m = simple_resid<OP>(L, P, tid, nthr);
          else for (int b = 0; b < L.nb; ++b) m = fmax(m, do_resid<OP>(L, P, b, tid, nthr));
          const double res = TEAM::max(m, red, scratch);
          it = i;
          if (res < scal[SCAL_COARSE_REL] * init_res || res < scal[SCAL_COARSE_ABS]) break;
        }
        if (tid == 0) scal[SCAL_COARSE_ITERS] = (double)it;
      } break;
      default: break;
    }
    TEAM::sync();
  }
}

// The tail levels are tiny (<= tail_cells cells each), so when phi, rhs, old and res of ALL of them fit in shared
// memory (n_stage > 0) they are staged there for the whole launch: the interpreter then runs on copies of the level
// descriptors whose var[0..3] point into shared memory (generic addressing), and every op of the coarse end of the
// cycle costs shared-memory instead of L2 latency.  Everything else (face tables, planes, eps) stays in global memory.
constexpr int kMaxStageLevels = 12;
constexpr int kMaxStageOps = 192;
__global__ void __launch_bounds__(1024) k_tail(const LevelDev* __restrict__ levels, const __grid_constant__ ParamsDev P,
                                               const TailOp* __restrict__ ops, int n_ops, double* scal, int n_stage) {
  __shared__ double red[32];
  __shared__ LevelDev slev[kMaxStageLevels];
  OMG_TAIL_DYN_SMEM(stage);
  const int tid = threadIdx.x, nthr = blockDim.x;
  if (n_stage > 0) {
    if (tid == 0) {
      size_t off = 0;
      for (int l = 0; l < n_stage; ++l) {
        slev[l] = levels[l];
        const size_t n = (size_t)levels[l].nb * levels[l].bs;
        for (int v = 0; v < 4; ++v) { slev[l].var[v] = stage + off; off += n; }
      }
      // the per-face and per-box tables follow the variables: every boundary cell's neighbour gather starts with a
      // dependent fkind -> farg -> ftype chain, which costs an L2 round trip each when the tables stay in global memory
      int* it = reinterpret_cast<int*>(stage + off);
      for (int l = 0; l < n_stage; ++l) {
        const int nf = levels[l].nb * 6, nb = levels[l].nb;
        slev[l].fkind = it; it += nf; slev[l].farg = it; it += nf; slev[l].ftype = it; it += nf;
        slev[l].fkind_old = it; it += nf; slev[l].farg_old = it; it += nf;
        slev[l].parent = it; it += nb; slev[l].coff = it; it += nb;
      }
    }
    __syncthreads();
    for (int l = 0; l < n_stage; ++l) {
      const int n = levels[l].nb * levels[l].bs;
      for (int v = 0; v < 4; ++v) {
        const double* __restrict__ src = levels[l].var[v];
        double* dst = slev[l].var[v];
        for (int i = tid; i < n; i += nthr) dst[i] = src[i];
      }
      const int nf = levels[l].nb * 6, nb = levels[l].nb;
      for (int i = tid; i < nf; i += nthr) {
        const_cast<int*>(slev[l].fkind)[i] = levels[l].fkind[i]; const_cast<int*>(slev[l].farg)[i] = levels[l].farg[i];
        const_cast<int*>(slev[l].ftype)[i] = levels[l].ftype[i];
        const_cast<int*>(slev[l].fkind_old)[i] = levels[l].fkind_old[i]; const_cast<int*>(slev[l].farg_old)[i] = levels[l].farg_old[i];
      }
      for (int i = tid; i < nb; i += nthr) {
        const_cast<int*>(slev[l].parent)[i] = levels[l].parent[i]; const_cast<int*>(slev[l].coff)[i] = levels[l].coff[i];
      }
    }
    __syncthreads();
  }
  const LevelDev* lv = n_stage > 0 ? slev : levels;
  // the op list too: every op starts with a broadcast load of its descriptor, an L2 round trip from global memory
  __shared__ TailOp sops[kMaxStageOps];
  if (n_ops <= kMaxStageOps) {
    for (int i = tid; i < n_ops; i += nthr) sops[i] = ops[i];
    __syncthreads();
    ops = sops;
  }
  OMG_SWITCH_OP(P.op, tail_run<OP>(lv, P, ops, n_ops, scal, red));
  if (n_stage > 0) {
    __syncthreads();
    for (int l = 0; l < n_stage; ++l) {
      const int n = levels[l].nb * levels[l].bs;
      for (int v = 0; v < 4; ++v) {
        double* __restrict__ dst = levels[l].var[v];
        const double* src = slev[l].var[v];
        for (int i = tid; i < n; i += nthr) dst[i] = src[i];
      }
    }
  }
}

#ifndef OMG_HOST_EMU
// The same interpreter run by a thread-block cluster (up to 8 CTAs x 1024 threads): no staging -- the levels stay in global
// memory, i.e. in L2 --, one hardware cluster barrier (~0.2 us) between operations.  GSRB and even box sizes only (the
// lexicographic and the odd-size smoothers synchronise inside a CTA); the host falls back to k_tail otherwise.
__global__ void __launch_bounds__(1024) k_tail_cluster(const LevelDev* __restrict__ levels, const __grid_constant__ ParamsDev P,
                                                       const TailOp* __restrict__ ops, int n_ops, double* scal) {
  __shared__ double red[32];
  __shared__ TailOp sops[kMaxStageOps];
  if (n_ops <= kMaxStageOps) {
    for (int i = threadIdx.x; i < n_ops; i += blockDim.x) sops[i] = ops[i];
    __syncthreads();
    ops = sops;
  }
  OMG_SWITCH_OP(P.op, (tail_run<OP, ClusterTeam>(levels, P, ops, n_ops, scal, red)));
}
#endif

}  // namespace
