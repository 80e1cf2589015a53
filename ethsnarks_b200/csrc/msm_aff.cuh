#pragma once
// Batched-affine pre-reduction rounds of the MSM (between msm_offsets and msm_accumulate).
//
// After the sort, bucket b owns the run [off[b], off[b+1]) of the pair list.  One round adds the points of every run in
// pairs (element 2j + element 2j+1 -> slot j of the bucket's new run, an odd last element is copied), halving every run:
// ceil(n/2).  The additions of a round are independent, so they are done in AFFINE coordinates with one shared inversion
// per thread block (Montgomery's trick): 6 field products per addition - 1 for the running product of the denominators,
// 2 to peel the inverse off on the way back, 3 for the addition itself - plus ~1 for the block-wide product scan,
// against 10 for the XYZZ mixed addition of msm_accumulate.  The single inversion per block is a binary extended Euclid
// (Fp::inv_binary, integer adds/shifts only) run by one thread while the other resident blocks keep the IMAD pipe busy.
// After R rounds the runs are 2^R times shorter and msm_accumulate (XYZZ, any run length) finishes the job on the
// compacted list; the group elements are the same, so the normalised result is bit-identical.
// The reference's bucket loop (multiexp.tcc:469-521) adds Jacobian points one by one; this is a different schedule of the
// same sums.
// STATUS (round 1, measured on B200 at 2^20 points): correct (parity tests, in-the-exponent checks at every round count)
// but NOT faster - one round costs 1.56 ms (G1) / 5.5 ms (G2) for the 6.8 M additions msm_accumulate does in 1.07 / 4.1 ms;
// removing the inversion only brings it to 1.31 / 4.9 ms, so the loss is in the body: two dependent gathers per addition
// (points are re-read on the way back), the prefix array in local memory and the block-wide scan.  It is therefore OFF by
// default (b200_set_option("affine_rounds", R) enables R rounds, -1 picks R from the run length).
#include "msm_common.cuh"
#include <cub/device/device_scan.cuh>

namespace b200 {

// cnt[b] = ceil(run length / 2) for b < nb, cnt[nb] = 0; an exclusive sum of cnt gives the next round's offsets
static __global__ void msm_round_counts(const uint64_t* __restrict__ off, uint32_t nb, uint64_t* __restrict__ cnt) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > nb) return;
    cnt[b] = b < nb ? (off[b + 1] - off[b] + 1) >> 1 : 0;
}

template <typename G, bool FIRST>
B200_DEV typename G::Aff aff_load(const uint32_t* __restrict__ vals, const typename G::Aff* __restrict__ pts, uint64_t i) {
    if (FIRST) {
        const uint32_t v = vals[i];
        typename G::Aff p = pts[v & 0x7fffffffu];
        if (v >> 31) p.y = p.y.neg();
        return p;
    }
    return pts[i];
}

// denominator of the affine addition P + Q and its kind: 0 = chord (den = Qx - Px), 1 = tangent (den = 2 Py),
// 2 = result is P, 3 = result is Q, 4 = result is infinity; kinds >= 2 need no inverse (den = 1)
template <typename F>
B200_DEV int aff_classify(const Affine<F>& P, const Affine<F>& Q, F& den) {
    if (P.is_inf()) { den = F::one(); return 3; }
    if (Q.is_inf()) { den = F::one(); return 2; }
    den = Q.x - P.x;
    if (!den.is_zero()) return 0;
    if (Q.y == P.y) { den = P.y.dbl(); return 1; }    // no point of order 2 in the prime-order groups: den != 0
    den = F::one();
    return 4;
}

#define B200_AFF_C 32          // input positions per thread (<= 32: the leader mask is one word); <= C/2 additions

template <typename G, bool FIRST>
__global__ void __launch_bounds__(128) msm_affine_round(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                        const typename G::Aff* __restrict__ pts, const uint64_t* __restrict__ off,
                                                        const uint64_t* __restrict__ newoff, uint32_t nb,
                                                        uint32_t* __restrict__ keys_out, typename G::Aff* __restrict__ pts_out) {
    typedef typename G::F F;
    typedef typename G::Aff Aff;
    constexpr int C = B200_AFF_C;
    __shared__ F sh_pre[128];
    __shared__ F sh_suf[128];
    __shared__ F sh_inv;
    const uint64_t M = off[nb];
    const uint32_t tid = threadIdx.x;
    const uint64_t start = ((uint64_t)blockIdx.x * 128 + tid) * C;
    const uint64_t end = start < M ? min(start + (uint64_t)C, M) : start;

    // ---- phase 1: copies, and the running product of this thread's denominators
    F prefix[C / 2];
    uint32_t adders = 0;           // bit p set: position start + p is the first element of a pair to add
    int n_add = 0;
    F acc = F::one();
    {
        uint32_t cur = 0xffffffffu;
        uint64_t o0 = 0, o1 = 0, no = 0;
        for (uint64_t i = start; i < end; ++i) {
            const uint32_t k = keys[i];
            if (k != cur) { cur = k; o0 = off[k]; o1 = off[k + 1]; no = newoff[k]; }
            const uint64_t rel = i - o0;
            if (rel & 1) continue;                       // second element of a pair
            if (i + 1 >= o1) {                           // odd last element of its run: copied
                const uint64_t slot = no + (rel >> 1);
                pts_out[slot] = aff_load<G, FIRST>(vals, pts, i);
                keys_out[slot] = k;
                continue;
            }
            const Aff P = aff_load<G, FIRST>(vals, pts, i), Q = aff_load<G, FIRST>(vals, pts, i + 1);
            F den;
            aff_classify(P, Q, den);
            acc = acc * den;
            prefix[n_add++] = acc;
            adders |= 1u << (uint32_t)(i - start);
        }
    }
    // ---- block-wide: inverse of this thread's product = (prod of all)^-1 * (prod of the threads before) * (prod after)
    sh_pre[tid] = acc;
    sh_suf[tid] = acc;
    __syncthreads();
    for (int d = 1; d < 128; d <<= 1) {
        F a, b;
        const bool lo = (int)tid >= d, hi = (int)tid + d < 128;
        if (lo) a = sh_pre[tid - d];
        if (hi) b = sh_suf[tid + d];
        __syncthreads();
        if (lo) sh_pre[tid] = sh_pre[tid] * a;
        if (hi) sh_suf[tid] = sh_suf[tid] * b;
        __syncthreads();
    }
    if (tid == 127) sh_inv = sh_pre[127].inv_binary();
    __syncthreads();
    F run = sh_inv;
    if (tid > 0) run = run * sh_pre[tid - 1];
    if (tid < 127) run = run * sh_suf[tid + 1];

    // ---- phase 2: walk the additions backwards, peeling one inverse off the running inverse each time
    {
        uint32_t cur = 0xffffffffu;
        uint64_t o0 = 0, no = 0;
        while (adders) {
            const uint32_t p = 31u - (uint32_t)__clz(adders);
            adders &= ~(1u << p);
            const uint64_t i = start + p;
            --n_add;
            const uint32_t k = keys[i];
            if (k != cur) { cur = k; o0 = off[k]; no = newoff[k]; }
            const Aff P = aff_load<G, FIRST>(vals, pts, i), Q = aff_load<G, FIRST>(vals, pts, i + 1);
            F den;
            const int kind = aff_classify(P, Q, den);
            F inv = n_add > 0 ? run * prefix[n_add - 1] : run;
            run = run * den;
            Aff R;
            if (kind <= 1) {
                F lambda;
                if (kind == 0) lambda = (Q.y - P.y) * inv;
                else { F xx = P.x.sqr(); lambda = (xx.dbl() + xx) * inv; }
                R.x = lambda.sqr() - P.x - Q.x;
                R.y = lambda * (P.x - R.x) - P.y;
            } else if (kind == 2) R = P;
            else if (kind == 3) R = Q;
            else R = Aff::inf();
            const uint64_t slot = no + ((i - o0) >> 1);
            pts_out[slot] = R;
            keys_out[slot] = k;
        }
    }
}

// Runs n_rounds rounds; on return *keys_out / *pts_out / *off_out describe the compacted list (device pointers).
template <typename G>
void launch_affine_rounds(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, int n_rounds, const uint32_t* skeys, const uint32_t* svals,
                          const void* d_bases, const uint64_t* offsets, const uint32_t** keys_out, const void** pts_out,
                          const uint64_t** off_out) {
    typedef typename G::Aff Aff;
    // sum over buckets of ceil(n/2) <= (pairs + nb) / 2
    const uint64_t bound1 = (p.pairs + p.nb) / 2 + 1;
    Aff* pbuf[2] = {(Aff*)ctx->scratch("msm_aff_pts0", bound1 * sizeof(Aff)), nullptr};
    uint32_t* kbuf[2] = {(uint32_t*)ctx->scratch("msm_aff_keys0", bound1 * 4), nullptr};
    if (n_rounds > 1) {
        const uint64_t bound2 = (bound1 + p.nb) / 2 + 1;
        pbuf[1] = (Aff*)ctx->scratch("msm_aff_pts1", bound2 * sizeof(Aff));
        kbuf[1] = (uint32_t*)ctx->scratch("msm_aff_keys1", bound2 * 4);
    }
    uint64_t* obuf[2] = {(uint64_t*)ctx->scratch("msm_aff_off0", ((size_t)p.nb + 2) * 8), (uint64_t*)ctx->scratch("msm_aff_off1", ((size_t)p.nb + 2) * 8)};
    uint64_t* cnt = (uint64_t*)ctx->scratch("msm_aff_cnt", ((size_t)p.nb + 2) * 8);
    size_t tmp_bytes = 0;
    B200_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(nullptr, tmp_bytes, cnt, obuf[0], (int)(p.nb + 1), st));
    void* tmp = ctx->scratch("msm_aff_scan_tmp", tmp_bytes);

    const uint32_t* keys = skeys;
    const Aff* pts = (const Aff*)d_bases;
    const uint64_t* off = offsets;
    uint64_t bound = p.pairs;                         // upper bound on the live length of the current list
    for (int r = 0; r < n_rounds; ++r) {
        uint64_t* noff = obuf[r & 1];
        msm_round_counts<<<(p.nb + 1 + 255) / 256, 256, 0, st>>>(off, p.nb, cnt);
        B200_CUDA_CHECK(cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, cnt, noff, (int)(p.nb + 1), st));
        const unsigned blocks = (unsigned)((bound + 128ull * B200_AFF_C - 1) / (128ull * B200_AFF_C));
        if (r == 0) msm_affine_round<G, true><<<blocks, 128, 0, st>>>(keys, svals, pts, off, noff, p.nb, kbuf[0], pbuf[0]);
        else msm_affine_round<G, false><<<blocks, 128, 0, st>>>(keys, nullptr, pts, off, noff, p.nb, kbuf[r & 1], pbuf[r & 1]);
        ctx->launches += 3;
        keys = kbuf[r & 1]; pts = pbuf[r & 1]; off = noff;
        bound = (bound + p.nb) / 2 + 1;
    }
    *keys_out = keys; *pts_out = pts; *off_out = off;
}

}  // namespace b200
