// omg_kernels.h -- host-callable launchers of the device kernels.
#pragma once
#include <cstddef>
#include <cuda_runtime.h>
#include "omg_device.cuh"

enum TailOpType : int {
  TOP_SMOOTH = 0, TOP_RESID, TOP_RESTRICT, TOP_CRHS, TOP_OLDPL, TOP_CORR_PROLONG, TOP_RB_FILL,
  TOP_COPY_OLD, TOP_COARSE_SOLVE
};
struct TailOp { int type, l, a, b; };   // l: index into the device level array (lvl - lowest_lvl)

// slots of the device scalar array
enum ScalSlot : int {
  SCAL_COARSE_REL = 0, SCAL_COARSE_ABS = 1, SCAL_COARSE_ITERS = 2, SCAL_MEAN = 3, SCAL_MAXRES = 4,
  SCAL_RESHIST = 8,            // SCAL_RESHIST + n: max_res of cycle n inside omg_fas_vcycles
  SCAL_TEAM = 8 + 256,         // SCAL_TEAM + r: scratch of CTA r of the cluster tail (cluster-wide max)
  SCAL_COUNT = 8 + 256 + 16
};

namespace omgk {
void smooth(const LevelDev& L, const ParamsDev& P, int colour, const int* list, int n, cudaStream_t s);   // list: subset of boxes or null
// mg_smoother_gs: ghost snapshot into L.faces[0] + in-box lexicographic sweep (two launches)
void smooth_gs(const LevelDev& L, const ParamsDev& P, cudaStream_t s);
void gs_snapshot(const LevelDev& L, const ParamsDev& P, cudaStream_t s);   // the two halves of smooth_gs
void gs_sweep(const LevelDev& L, const ParamsDev& P, cudaStream_t s);
void resid(const LevelDev& L, const ParamsDev& P, double* maxslot, cudaStream_t s);
void restrict2(const LevelDev& Lf, const LevelDev& Lc, int iv0, int iv1, cudaStream_t s);
void crhs(const LevelDev& L, const ParamsDev& P, cudaStream_t s);
void oldpl(const LevelDev& L, const ParamsDev& P, cudaStream_t s);
void corr_prolong(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, cudaStream_t s);
void prolong_var(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int iv, int iv_to, int add, cudaStream_t s);
void rb_fill(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, int iv, int gz, cudaStream_t s, const double* rpl = nullptr);
void apply_op(const LevelDev& L, const ParamsDev& P, int i_out, cudaStream_t s);
void fill_stored(const LevelDev& L, const ParamsDev& P, int iv, const int* bct, const double* bcv, cudaStream_t s, const double* rpl = nullptr, int n_recv = 0);
void pack_ghost(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, const int* desc, int n, double* out, int gz, cudaStream_t s, int iv = 0);
// ghost items stored directly into the receivers' plane pools (peer-mapped), copy (epoch + 1) & 1
void put_ghost(const LevelDev& L, const LevelDev& Lc, const ParamsDev& P, const int* desc, const int* dst, int n, double* const* peer_base,
               const long long* peer_stride, const unsigned long long* epoch, int gz, cudaStream_t s);
void pack_restrict(const LevelDev& Lf, const int* desc, int n, double* out, int iv0, int iv1, cudaStream_t s);
void unpack_restrict(const LevelDev& Lc, const int* desc, int n, const double* in, int h, int iv0, int iv1, cudaStream_t s);
void pack_prolong(const LevelDev& Lc, const ParamsDev& P, const int* desc, int n, double* out, int nc_f, cudaStream_t s);
// regrid (omg_update_tree): whole boxes between arrays, pairs = {source box, destination box}*
void copy_boxes(double* dst, const double* src, const int* pairs, int n_pairs, size_t box_doubles, cudaStream_t s);
void copy_boxes_strided(double* dst, size_t dstride, const double* src, size_t sstride, const int* pairs, int n_pairs, size_t n, cudaStream_t s);
void pack_prolong_var(const LevelDev& Lc, const ParamsDev& P, const int* desc, int n, double* out, int nc_f, int iv, cudaStream_t s);
void fill(double* p, size_t n, double v, cudaStream_t s);
void copy(double* dst, const double* src, size_t n, cudaStream_t s);
void sub_scalar(double* p, size_t n, const double* mean, cudaStream_t s);
void box_sums(const LevelDev& L, int iv, double w, double* partial, cudaStream_t s);
void axpby_leaves(const LevelDev& L, int iv_x, int iv_y, double f1, double f2, cudaStream_t s);   // y <- f1 x + f2 y on leaves
void final_sum(const double* partial, int n, double* acc, cudaStream_t s);
void div_scalar(double* acc, double denom, cudaStream_t s);
// n_stage > 0: the first n_stage levels are staged in shared memory for the launch (stage_bytes = 4 * 8 * their cells)
void tail(const LevelDev* levels, const ParamsDev& P, const TailOp* ops, int n_ops, double* scal, int n_stage, size_t stage_bytes,
          cudaStream_t s);
// the same op list run by a thread-block cluster of `cluster` CTAs (k_tail_cluster); false: launch failed (caller falls back)
bool tail_cluster(const LevelDev* levels, const ParamsDev& P, const TailOp* ops, int n_ops, double* scal, int cluster, cudaStream_t s);
// shell = 1: slabs of (nc+2)^3 with the ghost shell (the reference's cc arrays); 0: interior cells only, nc^3
void unpack(const LevelDev& L, int iv, const int* boxes, int n, const double* stage, int with_faces, int shell, cudaStream_t s);
void pack(const LevelDev& L, const ParamsDev& P, int iv, const int* boxes, int n, double* stage, int ghost_mode, int shell, cudaStream_t s);
}  // namespace omgk
