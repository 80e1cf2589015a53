// G1 accumulate kernel (msm_acc.cuh) in its own translation unit.
#include "msm_acc.cuh"
namespace b200 {
template void launch_accumulate<G1T>(b200_ctx*, cudaStream_t, const MsmPlan&, const uint32_t*, const uint32_t*, const uint64_t*, const void*,
                                     void*, void*, void*);
}
