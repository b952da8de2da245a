// omg_kernels.cu -- general-path kernels (one CTA per box, any box size / face kind), the
// single-CTA "tail" interpreter that runs the coarse end of the cycle in one launch, and the
// host<->device layout conversion kernels.  Tuned sm_100a kernels live in omg_fast.cu.
#include "omg_kernels.h"
#include "omg_ops.cuh"

#include "omg_tail.cuh"

namespace {
constexpr int kThreads = 256;

// ------------------------------------------------------------------ batched kernels ----------
template <int OP>
__global__ void __launch_bounds__(kThreads) k_smooth(const __grid_constant__ LevelDev L,
                                                     const __grid_constant__ ParamsDev P, int colour,
                                                     const int* __restrict__ boxlist) {
  do_smooth<OP>(L, P, boxlist ? boxlist[blockIdx.x] : blockIdx.x, colour, threadIdx.x, blockDim.x);
}

// mg_smoother_gs: ghost snapshot (= the ghost fill preceding the sweep) and the in-box lexicographic sweep
__global__ void __launch_bounds__(kThreads) k_gs_snapshot(const __grid_constant__ LevelDev L,
                                                          const __grid_constant__ ParamsDev P) {
  do_gs_snapshot(L, P, blockIdx.x, threadIdx.x, blockDim.x);
}
template <int OP>
__global__ void __launch_bounds__(kThreads) k_smooth_gs(const __grid_constant__ LevelDev L,
                                                        const __grid_constant__ ParamsDev P) {
  do_gs_sweep<OP>(L, P, blockIdx.x, threadIdx.x, blockDim.x);
}

template <int OP>
__global__ void __launch_bounds__(kThreads) k_resid(const __grid_constant__ LevelDev L,
                                                    const __grid_constant__ ParamsDev P, double* maxslot) {
  __shared__ double red[32];
  double m = do_resid<OP>(L, P, blockIdx.x, threadIdx.x, blockDim.x);
  if (maxslot) {
    m = block_max(m, red);
    if (threadIdx.x == 0) atomic_max_nonneg(maxslot, m);
  }
}

__global__ void __launch_bounds__(kThreads) k_restrict(const __grid_constant__ LevelDev Lf,
                                                       const __grid_constant__ LevelDev Lc, int iv0, int iv1) {
  do_restrict(Lf, Lc, iv0, blockIdx.x, threadIdx.x, blockDim.x);
  if (iv1 >= 0) do_restrict(Lf, Lc, iv1, blockIdx.x, threadIdx.x, blockDim.x);
}

template <int OP>
__global__ void __launch_bounds__(kThreads) k_crhs(const __grid_constant__ LevelDev L,
                                                   const __grid_constant__ ParamsDev P) {
  do_crhs<OP>(L, P, L.parents_list[blockIdx.x], threadIdx.x, blockDim.x);
}

__global__ void __launch_bounds__(kThreads) k_oldpl(const __grid_constant__ LevelDev L,
                                                    const __grid_constant__ ParamsDev P) {
  do_oldpl(L, P, blockIdx.x, threadIdx.x, blockDim.x);
}

__global__ void __launch_bounds__(kThreads) k_corr_prolong(const __grid_constant__ LevelDev Lf,
                                                           const __grid_constant__ LevelDev Lc,
                                                           const __grid_constant__ ParamsDev P) {
  do_corr_prolong(Lf, Lc, P, blockIdx.x, threadIdx.x, blockDim.x);
}

__global__ void __launch_bounds__(kThreads) k_prolong_var(const __grid_constant__ LevelDev Lf,
                                                          const __grid_constant__ LevelDev Lc,
                                                          const __grid_constant__ ParamsDev P, int iv, int iv_to,
                                                          int add) {
  if (iv == 0) do_prolong_var<GM_PHI>(Lf, Lc, P, iv, iv_to, add, blockIdx.x, threadIdx.x, blockDim.x);
  else do_prolong_var<GM_STORED>(Lf, Lc, P, iv, iv_to, add, blockIdx.x, threadIdx.x, blockDim.x);
}

__global__ void __launch_bounds__(kThreads) k_rb_fill(const __grid_constant__ LevelDev Lf,
                                                      const __grid_constant__ LevelDev Lc,
                                                      const __grid_constant__ ParamsDev P, int iv, int gz, const double* rpl) {
  if (gz) do_rb_fill<GM_PHI_GZ>(Lf, Lc, P, iv, blockIdx.x, threadIdx.x, blockDim.x, rpl);
  else do_rb_fill<GM_PHI>(Lf, Lc, P, iv, blockIdx.x, threadIdx.x, blockDim.x, rpl);
}

template <int OP>
__global__ void __launch_bounds__(kThreads) k_apply_op(const __grid_constant__ LevelDev L,
                                                       const __grid_constant__ ParamsDev P, int i_out) {
  do_apply_op<OP>(L, P, i_out, blockIdx.x, threadIdx.x, blockDim.x);
}

struct BcVar { int type[6]; double value[6]; };
__global__ void __launch_bounds__(kThreads) k_fill_stored(const __grid_constant__ LevelDev L,
                                                          const __grid_constant__ ParamsDev P, int iv,
                                                          const __grid_constant__ BcVar bc, const double* rpl, int n_recv) {
  do_fill_stored(L, P, iv, bc.type, bc.value, blockIdx.x, threadIdx.x, blockDim.x, rpl, n_recv);
}

// ---- exchange pack / unpack (multi-rank; one CTA per message item) ---------------------------------
// ghost phase: desc[4*e] = {kind (0 same-level, 1 coarse->fine), sender's local box, receiver's face, fine child offset}
__global__ void __launch_bounds__(kThreads) k_pack_ghost(const __grid_constant__ LevelDev L, const __grid_constant__ LevelDev Lc,
                                                         const __grid_constant__ ParamsDev P, const int* __restrict__ desc,
                                                         double* __restrict__ out, int gz, int iv) {
  const int e = blockIdx.x;
  const int kind = desc[4 * e], sb = desc[4 * e + 1], f = desc[4 * e + 2], off = desc[4 * e + 3];
  double* o = out + (size_t)e * L.nc * L.nc;
  if (kind == 0) do_pack_face(L, sb, f, o, threadIdx.x, blockDim.x, iv);
  else if (gz) do_pack_rb_face<GM_PHI_GZ>(Lc, P, sb, off, f, o, threadIdx.x, blockDim.x, iv);
  else do_pack_rb_face<GM_PHI>(Lc, P, sb, off, f, o, threadIdx.x, blockDim.x, iv);
}
// the same items as peer-mapped puts: item e goes straight into plane dst[2e+1] of the plane pool of rank dst[2e]
// (peer_base[rank], mapped over NVLink), into the copy the receiver is NOT reading: (epoch + 1) & 1
__global__ void __launch_bounds__(kThreads) k_put_ghost(const __grid_constant__ LevelDev L, const __grid_constant__ LevelDev Lc,
                                                        const __grid_constant__ ParamsDev P, const int* __restrict__ desc,
                                                        const int* __restrict__ dst, double* const* __restrict__ peer_base,
                                                        const long long* __restrict__ peer_stride, const unsigned long long* __restrict__ epoch, int gz) {
  const int e = blockIdx.x;
  const int kind = desc[4 * e], sb = desc[4 * e + 1], f = desc[4 * e + 2], off = desc[4 * e + 3];
  double* o = peer_base[dst[2 * e]] + (size_t)((*epoch + 1ull) & 1ull) * (size_t)peer_stride[dst[2 * e]] + (size_t)dst[2 * e + 1] * L.nc * L.nc;
  if (kind == 0) do_pack_face(L, sb, f, o, threadIdx.x, blockDim.x);
  else if (gz) do_pack_rb_face<GM_PHI_GZ>(Lc, P, sb, off, f, o, threadIdx.x, blockDim.x);
  else do_pack_rb_face<GM_PHI>(Lc, P, sb, off, f, o, threadIdx.x, blockDim.x);
  __threadfence_system();
}
// restrict phase: desc[e] = child's local box; item = nvar blocks of (nc/2)^3
__global__ void __launch_bounds__(kThreads) k_pack_restrict(const __grid_constant__ LevelDev Lf, const int* __restrict__ desc,
                                                            double* __restrict__ out, int iv0, int iv1) {
  const int h = Lf.nc >> 1, blk = h * h * h, nv = iv1 >= 0 ? 2 : 1;
  double* o = out + (size_t)blockIdx.x * blk * nv;
  do_pack_restrict(Lf, iv0, desc[blockIdx.x], o, threadIdx.x, blockDim.x);
  if (iv1 >= 0) do_pack_restrict(Lf, iv1, desc[blockIdx.x], o + blk, threadIdx.x, blockDim.x);
}
// desc[2*e] = {parent's local box, octant offset}
__global__ void __launch_bounds__(kThreads) k_unpack_restrict(const __grid_constant__ LevelDev Lc, const int* __restrict__ desc,
                                                              const double* __restrict__ in, int h, int iv0, int iv1) {
  const int blk = h * h * h, nv = iv1 >= 0 ? 2 : 1;
  const double* q = in + (size_t)blockIdx.x * blk * nv;
  do_unpack_restrict(Lc, iv0, desc[2 * blockIdx.x], desc[2 * blockIdx.x + 1], q, h, threadIdx.x, blockDim.x);
  if (iv1 >= 0) do_unpack_restrict(Lc, iv1, desc[2 * blockIdx.x], desc[2 * blockIdx.x + 1], q + blk, h, threadIdx.x, blockDim.x);
}
// prolong phase: desc[2*e] = {parent's local box, child's octant offset}; item = (nc/2+2)^3 correction patch
__global__ void __launch_bounds__(kThreads) k_pack_prolong(const __grid_constant__ LevelDev Lc, const __grid_constant__ ParamsDev P,
                                                           const int* __restrict__ desc, double* __restrict__ out, int nc_f) {
  const int pw = (nc_f >> 1) + 2;
  do_pack_prolong(Lc, P, desc[2 * blockIdx.x], desc[2 * blockIdx.x + 1], out + (size_t)blockIdx.x * pw * pw * pw, nc_f,
                  threadIdx.x, blockDim.x);
}

// stand-alone mg_prolong of any variable: patch of coarse values (see do_pack_prolong_var)
__global__ void __launch_bounds__(kThreads) k_pack_prolong_var(const __grid_constant__ LevelDev Lc, const __grid_constant__ ParamsDev P,
                                                               const int* __restrict__ desc, double* __restrict__ out, int nc_f, int iv) {
  const int pw = (nc_f >> 1) + 2;
  double* o = out + (size_t)blockIdx.x * pw * pw * pw;
  if (iv == 0) do_pack_prolong_var<GM_PHI>(Lc, P, iv, desc[2 * blockIdx.x], desc[2 * blockIdx.x + 1], o, nc_f, threadIdx.x, blockDim.x);
  else do_pack_prolong_var<GM_STORED>(Lc, P, iv, desc[2 * blockIdx.x], desc[2 * blockIdx.x + 1], o, nc_f, threadIdx.x, blockDim.x);
}

// regrid: whole boxes from one array to another; pairs[2e] = source box, pairs[2e+1] = destination box, strides in doubles
__global__ void __launch_bounds__(kThreads) k_copy_boxes(double* __restrict__ dst, size_t dstride, const double* __restrict__ src, size_t sstride,
                                                         const int* __restrict__ pairs, size_t n) {
  const double* sp = src + (size_t)pairs[2 * blockIdx.x] * sstride;
  double* dp = dst + (size_t)pairs[2 * blockIdx.x + 1] * dstride;
  for (size_t i = threadIdx.x; i < n; i += blockDim.x) dp[i] = sp[i];
}

// elementwise helpers over whole-level arrays
__global__ void k_fill(double* p, size_t n, double v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void k_copy(double* __restrict__ dst, const double* __restrict__ src, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
// x[i] -= *mean  (subtract_mean src/m_multigrid.f90:261-275)
__global__ void k_sub_scalar(double* p, size_t n, const double* mean) {
  const double m = *mean;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = p[i] - m;
}
__global__ void k_sub_scalar_planes(double* p, size_t n, const double* mean) {
  const double m = *mean;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = p[i] - m;
}

// set_rhs of m_diffusion (src/m_diffusion.f90:159-173): rhs <- f1 * phi + f2 * rhs on the leaves of a level, interior
// cells (the only cells stored).  One CTA per leaf box, 16-B accesses (the box stride 2*cs is even); 24 B per cell.
__global__ void __launch_bounds__(kThreads) k_axpby_leaves(const __grid_constant__ LevelDev L, int iv_x, int iv_y, double f1, double f2) {
  const int b = L.leaves_list[blockIdx.x];
  const double2* x = reinterpret_cast<const double2*>(L.var[iv_x] + (size_t)b * L.bs);
  double2* y = reinterpret_cast<double2*>(L.var[iv_y] + (size_t)b * L.bs);
  for (int idx = threadIdx.x; idx < L.bs / 2; idx += blockDim.x) {
    const double2 xv = x[idx];
    double2 yv = y[idx];
    const double a0 = f1 * xv.x, b0 = f2 * yv.x, a1 = f1 * xv.y, b1 = f2 * yv.y;
    yv.x = a0 + b0; yv.y = a1 + b1;
    y[idx] = yv;
  }
}

// get_sum (src/m_multigrid.f90:278-294): deterministic two-stage sum.  Stage 1: one partial per leaf
// box (fixed in-block tree order); stage 2: one thread adds the partials in box order.
__global__ void __launch_bounds__(kThreads) k_box_sums(const __grid_constant__ LevelDev L, int iv, double w, double* partial) {
  __shared__ double sm[kThreads];
  const int b = L.leaves_list[blockIdx.x];
  const double* v = L.var[iv] + (size_t)b * L.bs;
  double s = 0.0;
  for (int idx = threadIdx.x; idx < L.bs; idx += blockDim.x) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (decode_colour(L, colour, idx - colour * L.cs, i, j, k)) s += v[idx];
  }
  sm[threadIdx.x] = s;
  __syncthreads();
  for (int o = kThreads / 2; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = w * sm[0];
}
__global__ void k_final_sum(const double* partial, int n, double* acc) {
  __shared__ double sm[1024];
  // fixed-shape reduction: thread t sums elements t, t+1024, ...; then tree
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) s += partial[i];
  sm[threadIdx.x] = s;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) sm[threadIdx.x] += sm[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *acc = *acc + sm[0];
}
__global__ void k_scale_scalar(double* acc, double inv_volume) { *acc = *acc / inv_volume; }

// ------------------------------------------------------------------ host <-> device layout ----
// staging slab per box: (nc+2*shell)^3 doubles, reference order cc(i, j, k) i fastest; shell = 1: ghost shell included
// (the reference's cc arrays), shell = 0: interior cells only (mg_copy_from_tree, coupling_amrvac/mod_multigrid_coupling.t:206-234).
__global__ void __launch_bounds__(kThreads) k_unpack(const __grid_constant__ LevelDev L, int iv, const int* __restrict__ boxes,
                                                     const double* __restrict__ stage, int with_faces, int shell) {
  const int b = boxes[blockIdx.x];
  const int nc = L.nc, n2 = nc + 2 * shell;
  const double* s = stage + (size_t)blockIdx.x * n2 * n2 * n2;
  double* v = L.var[iv] + (size_t)b * L.bs;
  for (int idx = threadIdx.x; idx < L.bs; idx += blockDim.x) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (!decode_colour(L, colour, idx - colour * L.cs, i, j, k)) continue;
    v[idx] = s[((size_t)(k + shell) * n2 + (j + shell)) * n2 + (i + shell)];
  }
  if (with_faces && shell && L.faces[iv]) {
    double* fc = L.faces[iv] + (size_t)b * 6 * nc * nc;
    for (int idx = threadIdx.x; idx < 6 * nc * nc; idx += blockDim.x) {
      const int f = idx / (nc * nc), r = idx - f * nc * nc, t1 = r % nc, t2 = r / nc;
      const int g = (f & 1) ? nc + 1 : 0;
      int i, j, k;
      if (f < 2) { i = g; j = t1 + 1; k = t2 + 1; } else if (f < 4) { i = t1 + 1; j = g; k = t2 + 1; } else { i = t1 + 1; j = t2 + 1; k = g; }
      fc[idx] = s[((size_t)k * n2 + j) * n2 + i];
    }
  }
}

// ghost_mode: -1 none, else GhostMode for the six faces
__global__ void __launch_bounds__(kThreads) k_pack(const __grid_constant__ LevelDev L, const __grid_constant__ ParamsDev P, int iv,
                                                   const int* __restrict__ boxes, double* __restrict__ stage, int ghost_mode, int shell) {
  const int b = boxes[blockIdx.x];
  const int nc = L.nc, n2 = nc + 2 * shell;
  double* s = stage + (size_t)blockIdx.x * n2 * n2 * n2;
  const double* v = L.var[iv] + (size_t)b * L.bs;
  if (shell) {
    for (int idx = threadIdx.x; idx < n2 * n2 * n2; idx += blockDim.x) s[idx] = 0.0;
    __syncthreads();
  }
  for (int idx = threadIdx.x; idx < L.bs; idx += blockDim.x) {
    const int colour = idx / L.cs;
    int i, j, k;
    if (!decode_colour(L, colour, idx - colour * L.cs, i, j, k)) continue;
    s[((size_t)(k + shell) * n2 + (j + shell)) * n2 + (i + shell)] = v[idx];
  }
  if (ghost_mode >= 0 && shell) {
    for (int idx = threadIdx.x; idx < 6 * nc * nc; idx += blockDim.x) {
      const int f = idx / (nc * nc), r = idx - f * nc * nc, t1 = r % nc, t2 = r / nc;
      const int g = (f & 1) ? nc : -1;
      int i, j, k;
      if (f < 2) { i = g; j = t1; k = t2; } else if (f < 4) { i = t1; j = g; k = t2; } else { i = t1; j = t2; k = g; }
      double val;
      if (ghost_mode == GM_PHI) val = cell_at<GM_PHI>(L, P, L.var[iv], nullptr, b, i, j, k);
      else if (ghost_mode == GM_OLD) val = cell_at<GM_OLD>(L, P, L.var[iv], nullptr, b, i, j, k);
      else val = cell_at<GM_STORED>(L, P, L.var[iv], L.faces[iv], b, i, j, k);
      s[((size_t)(k + 1) * n2 + (j + 1)) * n2 + (i + 1)] = val;
    }
  }
}

inline int grid_for(size_t n) { size_t g = (n + 255) / 256; return (int)(g > 148 * 16 ? 148 * 16 : (g == 0 ? 1 : g)); }
}  // namespace

// ------------------------------------------------------------------ launch wrappers -----------
namespace omgk {

void smooth(const LevelDev& L, const ParamsDev& P, int colour, const int* list, int n, cudaStream_t s) {
  const int nb = list ? n : L.nb;
  if (nb == 0) return;
  OMG_SWITCH_OP(P.op, (k_smooth<OP><<<nb, kThreads, 0, s>>>(L, P, colour, list)));
}
void gs_snapshot(const LevelDev& L, const ParamsDev& P, cudaStream_t s) {
  if (L.nb == 0) return;
  k_gs_snapshot<<<L.nb, kThreads, 0, s>>>(L, P);
}
void gs_sweep(const LevelDev& L, const ParamsDev& P, cudaStream_t s) {
  if (L.nb == 0) return;
  OMG_SWITCH_OP(P.op, (k_smooth_gs<OP><<<L.nb, kThreads, 0, s>>>(L, P)));
}
void smooth_gs(const LevelDev& L, const ParamsDev& P, cudaStream_t s) { gs_snapshot(L, P, s); gs_sweep(L, P, s); }
void resid(const LevelDev& L, const ParamsDev& P, double* maxslot, cudaStream_t s) {
  if (L.nb == 0) return;
  OMG_SWITCH_OP(P.op, (k_resid<OP><<<L.nb, kThreads, 0, s>>>(L, P, maxslot)));
}
void restrict2(const LevelDev& Lf, const LevelDev& Lc, int iv0, int iv1, cudaStream_t s) {
  if (Lf.nb == 0) return;
  k_restrict<<<Lf.nb, kThreads, 0, s>>>(Lf, Lc, iv0, iv1);
}
void crhs(const LevelDev& L, const ParamsDev& P, cudaStream_t s) {
  if (L.n_parents == 0) return;
  OMG_SWITCH_OP(P.op, (k_crhs<OP><<<L.n_parents, kThreads, 0, s>>>(L, P)));
}
void oldpl(const LevelDev& L, const ParamsDev& P, cudaStream_t s) {
  if (L.n_oldpl == 0) return;
  k_oldpl<<<L.n_oldpl, kThreads, 0, s>>>(L, P);
}
void corr_prolong(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, cudaStream_t s) {
  if (Lf.nb == 0) return;
  k_corr_prolong<<<Lf.nb, kThreads, 0, s>>>(Lf, Lc, P);
}
void prolong_var(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int iv, int iv_to, int add, cudaStream_t s) {
  if (Lf.nb == 0) return;
  k_prolong_var<<<Lf.nb, kThreads, 0, s>>>(Lf, Lc, P, iv, iv_to, add);
}
void rb_fill(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int iv, int gz, cudaStream_t s, const double* rpl) {
  if (Lf.n_rb == 0) return;
  k_rb_fill<<<Lf.n_rb, kThreads, 0, s>>>(Lf, Lc, P, iv, gz, rpl);
}
void apply_op(const LevelDev& L, const ParamsDev& P, int i_out, cudaStream_t s) {
  if (L.nb == 0) return;
  OMG_SWITCH_OP(P.op, (k_apply_op<OP><<<L.nb, kThreads, 0, s>>>(L, P, i_out)));
}
void fill_stored(const LevelDev& L, const ParamsDev& P, int iv, const int* bct, const double* bcv, cudaStream_t s, const double* rpl, int n_recv) {
  if (L.nb == 0) return;
  BcVar bc;
  for (int f = 0; f < 6; ++f) { bc.type[f] = bct[f]; bc.value[f] = bcv[f]; }
  k_fill_stored<<<L.nb, kThreads, 0, s>>>(L, P, iv, bc, rpl, n_recv);
}
void pack_ghost(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, const int* desc, int n, double* out, int gz, cudaStream_t s, int iv) {
  if (n) k_pack_ghost<<<n, kThreads, 0, s>>>(L, Lc, P, desc, out, gz, iv);
}
void pack_prolong_var(const LevelDev& Lc, const ParamsDev& P, const int* desc, int n, double* out, int nc_f, int iv, cudaStream_t s) {
  if (n) k_pack_prolong_var<<<n, kThreads, 0, s>>>(Lc, P, desc, out, nc_f, iv);
}
void put_ghost(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, const int* desc, const int* dst, int n, double* const* peer_base,
               const long long* peer_stride, const unsigned long long* epoch, int gz, cudaStream_t s) {
  if (n) k_put_ghost<<<n, kThreads, 0, s>>>(L, Lc, P, desc, dst, peer_base, peer_stride, epoch, gz);
}
void pack_restrict(const LevelDev& Lf, const int* desc, int n, double* out, int iv0, int iv1, cudaStream_t s) {
  if (n) k_pack_restrict<<<n, kThreads, 0, s>>>(Lf, desc, out, iv0, iv1);
}
void unpack_restrict(const LevelDev& Lc, const int* desc, int n, const double* in, int h, int iv0, int iv1, cudaStream_t s) {
  if (n) k_unpack_restrict<<<n, kThreads, 0, s>>>(Lc, desc, in, h, iv0, iv1);
}
void pack_prolong(const LevelDev& Lc, const ParamsDev& P, const int* desc, int n, double* out, int nc_f, cudaStream_t s) {
  if (n) k_pack_prolong<<<n, kThreads, 0, s>>>(Lc, P, desc, out, nc_f);
}
void copy_boxes(double* dst, const double* src, const int* pairs, int n_pairs, size_t box_doubles, cudaStream_t s) {
  if (n_pairs) k_copy_boxes<<<n_pairs, kThreads, 0, s>>>(dst, box_doubles, src, box_doubles, pairs, box_doubles);
}
void copy_boxes_strided(double* dst, size_t dstride, const double* src, size_t sstride, const int* pairs, int n_pairs, size_t n, cudaStream_t s) {
  if (n_pairs) k_copy_boxes<<<n_pairs, kThreads, 0, s>>>(dst, dstride, src, sstride, pairs, n);
}
void fill(double* p, size_t n, double v, cudaStream_t s) { if (n) k_fill<<<grid_for(n), 256, 0, s>>>(p, n, v); }
void copy(double* dst, const double* src, size_t n, cudaStream_t s) { if (n) k_copy<<<grid_for(n), 256, 0, s>>>(dst, src, n); }
void sub_scalar(double* p, size_t n, const double* mean, cudaStream_t s) { if (n) k_sub_scalar<<<grid_for(n), 256, 0, s>>>(p, n, mean); }
void box_sums(const LevelDev& L, int iv, double w, double* partial, cudaStream_t s) {
  if (L.n_leaves) k_box_sums<<<L.n_leaves, kThreads, 0, s>>>(L, iv, w, partial);
}
void axpby_leaves(const LevelDev& L, int iv_x, int iv_y, double f1, double f2, cudaStream_t s) {
  if (L.n_leaves) k_axpby_leaves<<<L.n_leaves, kThreads, 0, s>>>(L, iv_x, iv_y, f1, f2);
}
void final_sum(const double* partial, int n, double* acc, cudaStream_t s) { if (n) k_final_sum<<<1, 1024, 0, s>>>(partial, n, acc); }
void div_scalar(double* acc, double denom, cudaStream_t s) { k_scale_scalar<<<1, 1, 0, s>>>(acc, denom); }
void tail(const LevelDev* levels, const ParamsDev& P, const TailOp* ops, int n_ops, double* scal, int n_stage, size_t stage_bytes,
          cudaStream_t s) {
  if (!n_ops) return;
  static size_t configured = 0;
  if (n_stage > kMaxStageLevels || stage_bytes > (size_t)200 * 1024) { n_stage = 0; stage_bytes = 0; }
  if (stage_bytes > configured) {
    if (cudaFuncSetAttribute(k_tail, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)stage_bytes) != cudaSuccess) { n_stage = 0; stage_bytes = 0; }
    else configured = stage_bytes;
  }
  k_tail<<<1, 1024, stage_bytes, s>>>(levels, P, ops, n_ops, scal, n_stage);
}
bool tail_cluster(const LevelDev* levels, const ParamsDev& P, const TailOp* ops, int n_ops, double* scal, int cluster, cudaStream_t s) {
  if (!n_ops) return true;
  if (cluster < 2 || cluster > 8) return false;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)cluster); cfg.blockDim = dim3(1024); cfg.dynamicSmemBytes = 0; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = (unsigned)cluster; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, k_tail_cluster, levels, P, ops, n_ops, scal) == cudaSuccess;
}
void unpack(const LevelDev& L, int iv, const int* boxes, int n, const double* stage, int with_faces, int shell, cudaStream_t s) {
  if (n) k_unpack<<<n, kThreads, 0, s>>>(L, iv, boxes, stage, with_faces, shell);
}
void pack(const LevelDev& L, const ParamsDev& P, int iv, const int* boxes, int n, double* stage, int ghost_mode, int shell, cudaStream_t s) {
  if (n) k_pack<<<n, kThreads, 0, s>>>(L, P, iv, boxes, stage, ghost_mode, shell);
}

}  // namespace omgk

// ---- self-test of ddiv_batch (tests/test_gpu_steps.py) -------------------------------------------------------
namespace {
__global__ void k_debug_ddiv(const double* a, const double* b, double* out, long long n, int batched) {
  const long long i0 = 6 * (blockIdx.x * (long long)blockDim.x + threadIdx.x);
  if (i0 >= n) return;
  double x[6], y[6], q[6];
  for (int f = 0; f < 6; ++f) { const long long i = i0 + f < n ? i0 + f : n - 1; x[f] = a[i]; y[f] = b[i]; }
  if (batched) ddiv_batch<6>(x, y, q);
  else for (int f = 0; f < 6; ++f) q[f] = x[f] / y[f];
  for (int f = 0; f < 6; ++f) if (i0 + f < n) out[i0 + f] = q[f];
}
}  // namespace
extern "C" int omg_debug_ddiv(int device, const double* a, const double* b, double* out, int64_t n, int batched) {
  if (n <= 0) return 0;
  double *da = nullptr, *db = nullptr, *dc = nullptr;
  int rc = 1;
  if (cudaSetDevice(device) == cudaSuccess && cudaMalloc(&da, n * 8) == cudaSuccess && cudaMalloc(&db, n * 8) == cudaSuccess &&
      cudaMalloc(&dc, n * 8) == cudaSuccess && cudaMemcpy(da, a, n * 8, cudaMemcpyHostToDevice) == cudaSuccess &&
      cudaMemcpy(db, b, n * 8, cudaMemcpyHostToDevice) == cudaSuccess) {
    const long long thr = (n + 5) / 6;
    k_debug_ddiv<<<(unsigned)((thr + 255) / 256), 256>>>(da, db, dc, n, batched);
    if (cudaMemcpy(out, dc, n * 8, cudaMemcpyDeviceToHost) == cudaSuccess) rc = 0;
  }
  cudaFree(da); cudaFree(db); cudaFree(dc);
  return rc;
}

// ---- omg_debug_ddiv_static: omg_debug_ddiv without any allocation -----------------------------------------------
// No counterpart in the reference; added for callers that may not call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The device staging of a, b and the
// quotients -- three cudaMalloc'ed arrays in omg_debug_ddiv -- lives in one file-scope array instead:
// 3 x OMG_DDIV_STATIC_MAX doubles = 6 MiB, held by every GPU that loads the library until the process exits.
// g_ddiv_scratch is a host-side stand-in in host code, not a device address: the address comes from
// cudaGetSymbolAddress, once per device with that device current.  Nothing in the array has to start at zero (every
// element the kernel reads or the host copies back was written earlier in the same call; the rest keeps the last
// call's contents).  All calls share the array: calls that overlap on one GPU race, the caller serialises them.
#include <cstdio>
#include <mutex>
#ifndef OMG_DDIV_STATIC_MAX
#define OMG_DDIV_STATIC_MAX 262144   /* = include/omg_gpu.h */
#endif
namespace {
static __device__ __align__(256) double g_ddiv_scratch[3 * OMG_DDIV_STATIC_MAX];
constexpr int kMaxDevices = 64;
double* g_ddiv_ptr[kMaxDevices] = {};
std::mutex g_ddiv_mutex;
void put_error(char* err, int cap, const char* fmt, long long a = 0, long long b = 0) {
  if (err && cap > 0) std::snprintf(err, (size_t)cap, fmt, a, b);
}
}  // namespace
extern "C" int omg_debug_ddiv_static(int device, const double* a, const double* b, double* out, int64_t n, int batched,
                                     char* err, int err_cap) {
  if (err && err_cap > 0) err[0] = 0;
  if (n <= 0) return 0;
  if (n > OMG_DDIV_STATIC_MAX) {
    put_error(err, err_cap, "omg_debug_ddiv_static: n = %lld exceeds the static scratch (at most %lld elements per call)", (long long)n, (long long)OMG_DDIV_STATIC_MAX);
    return 2;
  }
  if (!a || !b || !out) { put_error(err, err_cap, "omg_debug_ddiv_static: null array"); return 3; }
  if (device < 0 || device >= kMaxDevices) { put_error(err, err_cap, "omg_debug_ddiv_static: device %lld out of range", device); return 3; }
  std::lock_guard<std::mutex> lock(g_ddiv_mutex);      // one call at a time from this process (other processes: see above)
  cudaError_t e = cudaSetDevice(device);
  if (e != cudaSuccess) { put_error(err, err_cap, "omg_debug_ddiv_static: cudaSetDevice failed (%lld)", (long long)e); return 1; }
  if (!g_ddiv_ptr[device]) {                           // set-up, once per device, with that device current
    double* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_ddiv_scratch);
    if (e != cudaSuccess || !p) { put_error(err, err_cap, "omg_debug_ddiv_static: cudaGetSymbolAddress failed (%lld)", (long long)e); return 1; }
    g_ddiv_ptr[device] = p;
  }
  double* const da = g_ddiv_ptr[device];
  double* const db = da + OMG_DDIV_STATIC_MAX;
  double* const dc = db + OMG_DDIV_STATIC_MAX;
  if ((e = cudaMemcpy(da, a, n * 8, cudaMemcpyHostToDevice)) != cudaSuccess ||
      (e = cudaMemcpy(db, b, n * 8, cudaMemcpyHostToDevice)) != cudaSuccess) {
    put_error(err, err_cap, "omg_debug_ddiv_static: copy to the device failed (%lld)", (long long)e); return 1;
  }
  const long long thr = (n + 5) / 6;
  k_debug_ddiv<<<(unsigned)((thr + 255) / 256), 256>>>(da, db, dc, n, batched);
  if ((e = cudaGetLastError()) != cudaSuccess) { put_error(err, err_cap, "omg_debug_ddiv_static: launch failed (%lld)", (long long)e); return 1; }
  if ((e = cudaMemcpy(out, dc, n * 8, cudaMemcpyDeviceToHost)) != cudaSuccess) {
    put_error(err, err_cap, "omg_debug_ddiv_static: copy from the device failed (%lld)", (long long)e); return 1;
  }
  return 0;
}
