// G2 group helpers (msm_misc.cuh): window tables, fixed-base generation, additions, final affine.
#include "msm_misc.cuh"
namespace b200 {
B200_INSTANTIATE_GROUP_MISC(G2T, g2)
void g_combine3(b200_ctx* ctx, const void* d_parts, int world, const void* d_alpha_g1, const void* d_beta_g2, void* d_fin) {
    g_combine3_kernel<G2T><<<3, 32, 0, ctx->cur()>>>((const unsigned char*)d_parts, world, (const Affine<Fq>*)d_alpha_g1, (const Affine<Fq2>*)d_beta_g2,
                                                     (unsigned char*)d_fin);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}
}
