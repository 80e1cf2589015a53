// G2 MSM orchestration (msm_impl.cuh): digits, sort, offsets and the stage launchers.
#include "msm_impl.cuh"
namespace b200 {
B200_INSTANTIATE_GROUP_RUN(G2T, g2)
}
