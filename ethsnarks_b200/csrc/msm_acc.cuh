#pragma once
// The hot kernel of the MSM (stage 4 of msm_common.cuh's list) and its launcher.
#include "msm_common.cuh"
#include "msm_acc_pair.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------- 4. accumulate
template <typename G>
__global__ void __launch_bounds__(128, G::kAccMinBlocks) msm_accumulate(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                      const uint64_t* __restrict__ offsets, uint32_t nb, uint32_t L,
                                                      const typename G::Aff* __restrict__ bases, typename G::Pt* __restrict__ buckets,
                                                      typename G::Pt* __restrict__ pfirst, typename G::Pt* __restrict__ plast) {
    typedef typename G::Pt Pt;
    typedef typename G::Aff Aff;
    const uint64_t M = offsets[nb];
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t start = t * L;
    if (start >= M) return;
    const uint64_t end = min(start + (uint64_t)L, M);
    uint32_t cur = keys[start];
    bool first_run = true;
    Pt acc = Pt::inf();
    for (uint64_t i = start; i < end; ++i) {
        const uint32_t k = keys[i];
        if (k != cur) {
            if (first_run) pfirst[t] = acc; else buckets[cur] = acc;
            first_run = false;
            acc = Pt::inf();
            cur = k;
        }
        const uint32_t v = vals ? vals[i] : (uint32_t)i;     // no value list: the points are already in list order
        Aff p = bases[v & 0x7fffffffu];
        if (v >> 31) p.y = p.y.neg();
        acc.add_mixed(p);
    }
    if (first_run) pfirst[t] = acc; else plast[t] = acc;
}


template <typename G>
void launch_accumulate(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const uint32_t* skeys, const uint32_t* svals,
                       const uint64_t* offsets, const void* d_bases, void* buckets, void* pfirst, void* plast) {
    typedef typename G::Pt Pt;
    if constexpr (sizeof(Pt) > 128) {
        if (ctx->g2_pair) {     // one point per lane pair (msm_acc_pair.cuh)
            const unsigned grid = (unsigned)((2 * p.nchunks + B200_G2_PAIR_BLOCK - 1) / B200_G2_PAIR_BLOCK);
#define B200_PAIR_LAUNCH(MODE)                                                                                              \
    msm_accumulate_g2_pair<G, MODE><<<grid, B200_G2_PAIR_BLOCK, 0, st>>>(skeys, svals, offsets, p.nb, p.L, (const G2Affine*)d_bases, (G2XYZZ*)buckets, \
                                                          (G2XYZZ*)pfirst, (G2XYZZ*)plast)
            if (ctx->g2_pair == 1) B200_PAIR_LAUNCH(1);          // strict operand preparation, kept as cross-check
            else if (ctx->g2_pair == 3) B200_PAIR_LAUNCH(3);     // lazy operands + out-of-line products
            else B200_PAIR_LAUNCH(2);                            // lazy operands
#undef B200_PAIR_LAUNCH
            ctx->launches++;
            return;
        }
    }
    size_t smem = 0;
    if (sizeof(Pt) > 128 && ctx->g2_acc_smem > 0) {
        smem = (size_t)ctx->g2_acc_smem;
        B200_CUDA_CHECK(cudaFuncSetAttribute(msm_accumulate<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    msm_accumulate<G><<<(unsigned)((p.nchunks + 127) / 128), 128, smem, st>>>(skeys, svals, offsets, p.nb, p.L, (const typename G::Aff*)d_bases,
                                                                            (Pt*)buckets, (Pt*)pfirst, (Pt*)plast);
    ctx->launches++;
}

}  // namespace b200
