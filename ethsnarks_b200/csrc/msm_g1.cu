// G1 MSM orchestration (msm_impl.cuh): digits, sort, offsets and the stage launchers.
#include "msm_impl.cuh"
namespace b200 {
B200_INSTANTIATE_GROUP_RUN(G1T, g1)
int msm_table_window(size_t n) { return choose_window(n, true); }
void msm_shared_prepare(b200_ctx* ctx, const void* d_scalars_mont, size_t n, int c, MsmShared* out) {
    if (n == 0 || n >= (1ull << B200_SHARED_VAR_BITS) || (255 + c - 1) / c > 32) throw std::runtime_error("msm_shared_prepare: size out of range");
    const MsmPlan p = make_plan(ctx, n, c, false);
    const MsmSorted so = msm_sort_stage(ctx, ctx->cur(), p, d_scalars_mont, nullptr, n, true, 0, true);
    out->c = c; out->n = (uint32_t)n; out->keys = so.keys; out->vals = so.vals; out->offsets = so.offsets;
}
}
