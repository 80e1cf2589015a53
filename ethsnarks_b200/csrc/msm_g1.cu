// G1 instantiation of the MSM pipeline and group helpers (see msm_impl.cuh).
#include "msm_impl.cuh"
namespace b200 {
B200_INSTANTIATE_GROUP(G1T, g1)
int msm_table_window(size_t n) { return choose_window(n, true); }
}
