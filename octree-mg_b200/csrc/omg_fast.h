// omg_fast.h -- tuned sm_100a kernels for levels of box size 16 / 32.  Every face kind is served: FK_SAME and
// FK_PHYS ghost layers are staged / evaluated in shared memory, FK_PLANE faces (refinement boundaries, faces owned
// by another rank) are read from the level's plane pool -- so AMR and distributed levels take this path too.
// Each launcher returns false when it does not apply (other box sizes, ahelmholtz in some modes, all_fast == false);
// the caller then uses the general path of omg_kernels.cu and counts the launch as "general_launches"
// (omg_stat).  Results are bitwise identical between the two paths.
#pragma once
#include <cuda_runtime.h>
#include "omg_device.cuh"

namespace omgf {
// option "pdl": launch the kernels below with programmatic stream serialization (they order themselves with
// griddepcontrol.wait); single-rank cycles only, off by default
void set_pdl(bool on);
// option "ring": persistent ring smoother (k_smooth_ring): 1 = on levels with at least four tiles per SM (default), 2 = on
// every level it serves, 0 = never (one tile per CTA, k_smooth_lpl_fast)
void set_ring(int mode);
// option "store_fine_residual" (default 1): 0 = the residual kernels do not write the level's own residual (mg_ires) --
// for device-resident hosts that never read it; the cycle itself consumes the residual in registers
void set_store_res(int on);
// list != null: only the n boxes of the list (partition-face / interior split of a distributed level)
// fused: the list is the level in launch order, partition-face boxes first; they exchange their faces in-kernel (PutDev)
bool smooth(const LevelDev& L, const ParamsDev& P, int colour, bool all_fast, const int* list, int n, int fused, cudaStream_t s);
bool smooth_applies(const LevelDev& L, const ParamsDev& P);   // would smooth() take this level / operator?
// would smooth() on n boxes of this level run the persistent ring kernel?  With the in-kernel exchange its launch list must
// hold the partition-face boxes LAST: a persistent CTA that meets one waits for the peer's previous half-sweep to end, and
// everything queued behind that tile waits with it (2 GPUs, r02n: 0.385 ms per half-sweep with the natural order)
bool smooth_uses_ring(const LevelDev& L, const ParamsDev& P, int n);
// mg_smoother_gs: the in-box lexicographic sweep after omgk::gs_snapshot (box 16 / 8, Laplacian / Helmholtz)
bool gs_sweep(const LevelDev& L, const ParamsDev& P, cudaStream_t s);
bool resid_restrict(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, bool all_fast, cudaStream_t s);
bool resid_max(const LevelDev& L, const ParamsDev& P, bool all_fast, double* maxslot, cudaStream_t s);
bool crhs(const LevelDev& L, const ParamsDev& P, bool all_fast, cudaStream_t s);
bool corr_prolong(const LevelDev& Lf, const LevelDev& Lc, const ParamsDev& P, bool coarse_all_fast, int fused, cudaStream_t s);   // fused = 1: in-kernel halo put of the corrected faces
}  // namespace omgf
