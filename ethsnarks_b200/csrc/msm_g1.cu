// G1 MSM orchestration (msm_impl.cuh): digits, sort, offsets and the stage launchers.
#include "msm_impl.cuh"
namespace b200 {
B200_INSTANTIATE_GROUP_RUN(G1T, g1)
int msm_table_window(size_t n) { return choose_window(n, true); }
}
