#pragma once
// Multi-scalar multiplication on BN254 G1 / G2 for sm_100a.
//
// Replaces the reference's CPU Pippenger (FF = /root/reference/depends/libsnark/depends/libff/libff):
//   multi_exp / multi_exp_with_mixed_addition        FF/algebra/scalar_multiplication/multiexp.tcc:595-714
//   multi_exp_with_density (window loop, combine)     multiexp.tcc:523-593
//   multi_exp_inner_bellman_with_density + get_id     multiexp.tcc:469-521, 167-181
//   kc_multi_exp_with_mixed_addition (sparse gather)  LS/knowledge_commitment/kc_multiexp.tcc:21-136
// Same mathematical result (unique once normalised to affine), different schedule:
//   1. msm_digits      scalar -> standard form -> signed c-bit digits; one (bucket key, point|sign) pair per window.
//                      Zero digits get a sentinel key, so the reference's "scalar == 0" prefilter falls out for free and
//                      "scalar == 1" is just bucket 1 of window 0.
//   2. radix sort of the pairs by bucket key (CUB, toolkit library).
//   3. msm_offsets     bucket start offsets by binary search in the sorted keys.
//   4. msm_accumulate  the hot kernel: the sorted pair list is cut into fixed-length chunks, one thread per chunk, so
//                      the work per thread is the same whatever the scalar distribution (uniform 254-bit or 0/1-heavy).
//                      Each thread walks its chunk with one XYZZ accumulator and mixed additions of affine bases
//                      (8M + 2S); runs that start or end at a chunk edge are emitted as partial sums.
//   5. msm_finalize    per bucket: stitch the partial sums of the chunks it spans (buckets spanning many chunks — the
//                      giant "scalar == 1" bucket of real witnesses — go to a block-per-bucket tree reduction).
//   6. msm_bucket_reduce / msm_window_reduce / msm_horner: sum_b b*B_b per window, then sum_k 2^(ck) W_k.
#include "ec.cuh"
#include "internal.hpp"
#include <stdexcept>
#include <algorithm>
#include <cmath>

namespace b200 {

#ifndef B200_G1_ACC_MINB
#define B200_G1_ACC_MINB 4      // resident 128-thread blocks per SM the accumulate kernel is compiled for (register cap)
#endif
#ifndef B200_G2_ACC_MINB
#define B200_G2_ACC_MINB 2
#endif
struct G1T { typedef Fq F; typedef Affine<Fq> Aff; typedef XYZZ<Fq> Pt; static constexpr int kAccMinBlocks = B200_G1_ACC_MINB; };
struct G2T { typedef Fq2 F; typedef Affine<Fq2> Aff; typedef XYZZ<Fq2> Pt; static constexpr int kAccMinBlocks = B200_G2_ACC_MINB; };

template <typename F> struct Gen;
template <> struct Gen<Fq> {
    static B200_DEV Affine<Fq> get() { Fq one = Fq::one(); return Affine<Fq>{one, one + one}; }   // (1, 2), mcl_bn128_init.cpp:109-111
};
template <> struct Gen<Fq2> {
    static B200_DEV Affine<Fq2> get() {   // mcl_bn128_init.cpp:170-174, Montgomery limbs
        const uint32_t g[4][8] = {
            {0x02bc2026u, 0x8e83b5d1u, 0x497b0172u, 0xdceb1935u, 0x97811adfu, 0xfbb82647u, 0xaf96503bu, 0x19573841u},
            {0xa84c6140u, 0xafb4737du, 0x5802d8c4u, 0x6043dd5au, 0x52a02f86u, 0x09e950fcu, 0x3aea7b6bu, 0x14fef083u},
            {0x886be9f6u, 0x619dfa9du, 0xf59e9b78u, 0xfe7fd297u, 0x231b7dfeu, 0xff9e1a62u, 0xae9e4206u, 0x28fd7eebu},
            {0xc71856eeu, 0x64095b56u, 0x327d3cbbu, 0xdc57f922u, 0x33351076u, 0x55f935beu, 0x93fd6482u, 0x0da4a0e6u}};
        Affine<Fq2> p;
#pragma unroll
        for (int i = 0; i < 8; ++i) { p.x.a.v[i] = g[0][i]; p.x.b.v[i] = g[1][i]; p.y.a.v[i] = g[2][i]; p.y.b.v[i] = g[3][i]; }
        return p;
    }
};

#define B200_BIG_SPAN 8     // a bucket spanning more chunks than this goes to the block-per-bucket reduction (one thread chains the additions below it: 64 cost 0.5 ms per G2 bucket on small real circuits)
struct MsmPlan {
    int c, W, keybits, table;
    uint32_t half, nb, L, S, nseg;
    uint64_t pairs, nchunks;
};


// Each group's kernels are spread over several translation units (msm_<group>_acc.cu, _red.cu, .cu, _misc.cu) so that
// `make -j` compiles them in parallel; the orchestrator reaches the kernels of the other units through these launchers,
// which are explicitly instantiated where the kernels live.
template <typename G>
void launch_accumulate(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const uint32_t* skeys, const uint32_t* svals,
                       const uint64_t* offsets, const void* d_bases, void* buckets, void* pfirst, void* plast);
template <typename G>
void launch_finalize(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const uint32_t* keys, const uint64_t* offsets, void* buckets,
                     const void* pfirst, const void* plast, uint32_t* big, size_t big_cap);
template <typename G>
void launch_reduce(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const void* buckets, void* seg, void* win, void* d_out);

}  // namespace b200
