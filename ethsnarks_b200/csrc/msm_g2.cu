// G2 instantiation of the MSM pipeline and group helpers (see msm_impl.cuh).
#include "msm_impl.cuh"
namespace b200 {
B200_INSTANTIATE_GROUP(G2T, g2)
}
