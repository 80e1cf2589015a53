// G2 batched-affine pre-reduction rounds (msm_aff.cuh) in their own translation unit.
#include "msm_aff.cuh"
namespace b200 {
template void launch_affine_rounds<G2T>(b200_ctx*, cudaStream_t, const MsmPlan&, int, const uint32_t*, const uint32_t*, const void*,
                                        const uint64_t*, const uint32_t**, const void**, const uint64_t**);
}
