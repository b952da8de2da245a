// omg_fast.cu -- tuned kernels (placeholder: general path is used until these land)
#include "omg_fast.h"
namespace omgf {
bool smooth(const LevelDev&, const ParamsDev&, int, bool, cudaStream_t) { return false; }
bool resid_restrict(const LevelDev&, const LevelDev&, const ParamsDev&, bool, cudaStream_t) { return false; }
bool corr_prolong(const LevelDev&, const LevelDev&, const ParamsDev&, bool, cudaStream_t) { return false; }
}  // namespace omgf
