// G2 finalize and bucket-reduction kernels (msm_red.cuh) in their own translation unit.
#include "msm_red.cuh"
namespace b200 {
template void launch_finalize<G2T>(b200_ctx*, cudaStream_t, const MsmPlan&, const uint32_t*, const uint64_t*, void*, const void*, const void*, uint32_t*, size_t);
template void launch_reduce<G2T>(b200_ctx*, cudaStream_t, const MsmPlan&, const void*, void*, void*, void*);
}
