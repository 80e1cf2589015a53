// G1 group helpers (msm_misc.cuh): window tables, fixed-base generation, additions, final affine.
#include "msm_misc.cuh"
namespace b200 {
B200_INSTANTIATE_GROUP_MISC(G1T, g1)
}
