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
#include <cub/device/device_radix_sort.cuh>
#include <stdexcept>
#include <algorithm>
#include <cmath>

namespace b200 {

struct G1T { typedef Fq F; typedef Affine<Fq> Aff; typedef XYZZ<Fq> Pt; };
struct G2T { typedef Fq2 F; typedef Affine<Fq2> Aff; typedef XYZZ<Fq2> Pt; };

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

// ------------------------------------------------------------------------------------------------- 1. digits
template <bool MONT>
static __global__ void __launch_bounds__(256) msm_digits(const Fr* __restrict__ scalars, const uint32_t* __restrict__ gather, uint32_t n,
                                                  int c, int W, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
                                                  uint32_t stride, int table) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr s = scalars[gather ? gather[i] : i];
    if (MONT) s = s.from_mont();
    uint32_t limb[9];
#pragma unroll
    for (int k = 0; k < 8; ++k) limb[k] = s.v[k];
    limb[8] = 0;
    // table mode: window w of point i uses the precomputed base 2^(c w) P_i at index w * stride + i, and every window
    // feeds the same bucket set, so the key is the digit alone.
    const uint32_t half = 1u << (c - 1), full = 1u << c, mask = full - 1, sentinel = table ? half : (uint32_t)W * half;
    uint32_t carry = 0;
    for (int w = 0; w < W; ++w) {
        uint32_t bit = (uint32_t)w * c, raw = 0;
        if (bit < 256) {
            uint32_t l = bit >> 5, off = bit & 31;
            uint64_t two = ((uint64_t)limb[l + 1] << 32) | limb[l];
            raw = (uint32_t)(two >> off) & mask;
        }
        uint32_t u = raw + carry, d, neg;
        if (u > half) { d = full - u; neg = 1; carry = 1; } else { d = u; neg = 0; carry = 0; }
        keys[(size_t)w * n + i] = d ? (table ? 0u : (uint32_t)w * half) + d - 1 : sentinel;
        vals[(size_t)w * n + i] = (table ? (uint32_t)w * stride + i : i) | (neg << 31);
    }
}

// ------------------------------------------------------------------------------------------------- 3. offsets
// offsets[b] = first position in the sorted key array holding a key >= b, for b in [0, nb]; offsets[nb] = number
// of live (non-sentinel) pairs.
static __global__ void msm_offsets(const uint32_t* __restrict__ keys, uint64_t pairs, uint32_t nb, uint64_t* __restrict__ offsets) {
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > nb) return;
    uint64_t lo = 0, hi = pairs;
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (keys[mid] < b) lo = mid + 1; else hi = mid;
    }
    offsets[b] = lo;
}

// ------------------------------------------------------------------------------------------------- 4. accumulate
template <typename G>
__global__ void __launch_bounds__(128) msm_accumulate(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
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
        const uint32_t v = vals[i];
        Aff p = bases[v & 0x7fffffffu];
        if (v >> 31) p.y = p.y.neg();
        acc.add_mixed(p);
    }
    if (first_run) pfirst[t] = acc; else plast[t] = acc;
}

// ------------------------------------------------------------------------------------------------- 5. finalize
#define B200_BIG_SPAN 64
template <typename G>
__global__ void __launch_bounds__(128) msm_finalize(const uint64_t* __restrict__ offsets, uint32_t nb, uint32_t L,
                                                    typename G::Pt* __restrict__ buckets, const typename G::Pt* __restrict__ pfirst,
                                                    const typename G::Pt* __restrict__ plast, uint32_t* __restrict__ big_count,
                                                    uint32_t* __restrict__ big_list) {
    typedef typename G::Pt Pt;
    uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= nb) return;
    const uint64_t M = offsets[nb];
    const uint64_t s = offsets[b], e = offsets[b + 1];
    if (s == e) { buckets[b] = Pt::inf(); return; }
    const uint64_t t0 = s / L, t1 = (e - 1) / L;
    if (t0 == t1) {
        const bool first = (s == t0 * L);
        const bool last = (e == min((t0 + 1) * (uint64_t)L, M));
        if (first) buckets[b] = pfirst[t0];
        else if (last) buckets[b] = plast[t0];
        return;   // otherwise msm_accumulate already wrote the complete run
    }
    if (t1 - t0 > B200_BIG_SPAN) { big_list[atomicAdd(big_count, 1u)] = b; return; }
    Pt acc = (s == t0 * L) ? pfirst[t0] : plast[t0];
    for (uint64_t t = t0 + 1; t <= t1; ++t) acc.add(pfirst[t]);
    buckets[b] = acc;
}

template <typename Pt>
B200_DEV void block_tree_reduce(Pt& acc, Pt* sh) {   // blockDim.x == 128; result in thread 0's acc
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int stride = 64; stride > 0; stride >>= 1) {
        if ((int)threadIdx.x < stride) {
            Pt a = sh[threadIdx.x];
            a.add(sh[threadIdx.x + stride]);
            sh[threadIdx.x] = a;
        }
        __syncthreads();
    }
    acc = sh[0];
    __syncthreads();
}

template <typename G>
__global__ void __launch_bounds__(128) msm_finalize_big(const uint64_t* __restrict__ offsets, uint32_t L,
                                                        typename G::Pt* __restrict__ buckets, const typename G::Pt* __restrict__ pfirst,
                                                        const typename G::Pt* __restrict__ plast, const uint32_t* __restrict__ big_count,
                                                        const uint32_t* __restrict__ big_list) {
    typedef typename G::Pt Pt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Pt* sh = reinterpret_cast<Pt*>(smem_raw);
    const uint32_t nbig = *big_count;
    for (uint32_t j = blockIdx.x; j < nbig; j += gridDim.x) {
        const uint32_t b = big_list[j];
        const uint64_t s = offsets[b], e = offsets[b + 1];
        const uint64_t t0 = s / L, t1 = (e - 1) / L;
        Pt acc = Pt::inf();
        if (threadIdx.x == 0) acc = (s == t0 * L) ? pfirst[t0] : plast[t0];
        for (uint64_t t = t0 + 1 + threadIdx.x; t <= t1; t += blockDim.x) acc.add(pfirst[t]);
        block_tree_reduce(acc, sh);
        if (threadIdx.x == 0) buckets[b] = acc;
    }
}

// ------------------------------------------------------------------------------------------------- 6. reductions
// one thread per segment of S consecutive buckets of one window: sum_{b in seg} (b+1) * B_b
template <typename G>
__global__ void __launch_bounds__(128) msm_bucket_reduce(const typename G::Pt* __restrict__ buckets, uint32_t half, uint32_t S,
                                                         uint32_t nseg, int W, typename G::Pt* __restrict__ seg) {
    typedef typename G::Pt Pt;
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= nseg * (uint32_t)W) return;
    const uint32_t w = id / nseg, sidx = id % nseg;
    const uint32_t lo = sidx * S, hi = min(lo + S, half);
    const Pt* B = buckets + (size_t)w * half;
    Pt acc = Pt::inf(), sum = Pt::inf();
    for (uint32_t lb = hi; lb-- > lo;) {
        acc.add(B[lb]);
        sum.add(acc);
    }
    if (lo && !acc.is_inf()) sum.add(acc.mul_u32(lo));
    seg[id] = sum;
}
// one block per window: sum of its segment results
template <typename G>
__global__ void __launch_bounds__(128) msm_window_reduce(const typename G::Pt* __restrict__ seg, uint32_t nseg, typename G::Pt* __restrict__ win) {
    typedef typename G::Pt Pt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Pt* sh = reinterpret_cast<Pt*>(smem_raw);
    const uint32_t w = blockIdx.x;
    Pt acc = Pt::inf();
    for (uint32_t s = threadIdx.x; s < nseg; s += blockDim.x) acc.add(seg[(size_t)w * nseg + s]);
    block_tree_reduce(acc, sh);
    if (threadIdx.x == 0) win[w] = acc;
}
// final = sum_k 2^(c k) win[k]   (the reference's combine, multiexp.tcc:583-592)
template <typename G>
__global__ void msm_horner(const typename G::Pt* __restrict__ win, int W, int c, typename G::Pt* __restrict__ out) {
    typedef typename G::Pt Pt;
    if (threadIdx.x || blockIdx.x) return;
    Pt r = Pt::inf();
    for (int w = W - 1; w >= 0; --w) {
        for (int i = 0; i < c; ++i) r.dbl();
        r.add(win[w]);
    }
    *out = r;
}

// generic sum: out[b] = sum of in[b*R*128 .. (b+1)*R*128), one block of 128 threads per output point
template <typename G>
__global__ void __launch_bounds__(128) msm_tree_reduce(const typename G::Pt* __restrict__ in, uint32_t n, uint32_t R,
                                                       typename G::Pt* __restrict__ out) {
    typedef typename G::Pt Pt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Pt* sh = reinterpret_cast<Pt*>(smem_raw);
    const uint64_t base = (uint64_t)blockIdx.x * R * 128;
    Pt acc = Pt::inf();
    for (uint32_t r = 0; r < R; ++r) {
        const uint64_t i = base + (uint64_t)r * 128 + threadIdx.x;
        if (i < n) acc.add(in[i]);
    }
    block_tree_reduce(acc, sh);
    if (threadIdx.x == 0) out[blockIdx.x] = acc;
}

// Fixed-base window tables for a resident key: table[w * n + i] = 2^(c w) * P_i, affine.  The proving key never
// changes between proofs, so (like the reference generator's get_window_table, multiexp.tcc:771-807, but on the prover
// side and held in HBM) every window of every base is materialised once at key load: all windows then share ONE
// bucket set, the per-window bucket reductions and the 254-doubling window combine (multiexp.tcc:583-592) disappear.
template <typename G>
__global__ void __launch_bounds__(128) msm_precompute(const typename G::Aff* __restrict__ bases, size_t n, int c, int W,
                                                      typename G::Aff* __restrict__ table) {
    typedef typename G::Pt Pt;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const typename G::Aff b = bases[i];
    table[i] = b;
    Pt p = Pt::from_affine(b);
    for (int w = 1; w < W; ++w) {
        for (int k = 0; k < c; ++k) p.dbl();
        table[(size_t)w * n + i] = p.to_affine();
    }
}

// ------------------------------------------------------------------------------------------------- host side
struct MsmPlan {
    int c, W, keybits, table;
    uint32_t half, nb, L, S, nseg;
    uint64_t pairs, nchunks;
};

// window size by cost model: N*W mixed adds (+ sort share) against the bucket reduction, which is paid once per window
// without tables and once in total with them
static int choose_window(size_t n, bool table) {
    int best_c = 8;
    double best = 1e300;
    const double add_cost = 10.0, full_cost = 14.0;
    for (int c = 6; c <= 22; ++c) {
        int W = (255 + c - 1) / c;
        double cost = (double)n * W * (add_cost + 1.5 /* sort + digits share */) +
                      (table ? 1.0 : (double)W) * std::ldexp(1.0, c - 1) * (2 * full_cost + 6);
        if (cost < best) { best = cost; best_c = c; }
    }
    return best_c;
}

static MsmPlan make_plan(b200_ctx* ctx, size_t n, int table_c) {
    MsmPlan p;
    int best_c = table_c ? table_c : choose_window(n, false);
    if (ctx->window_override && !table_c) best_c = std::max(4, std::min(24, ctx->window_override));
    p.table = table_c != 0;
    p.c = best_c;
    p.W = (255 + p.c - 1) / p.c;
    p.half = 1u << (p.c - 1);
    p.nb = p.table ? p.half : (uint32_t)p.W * p.half;
    p.keybits = 1;
    while ((1ull << p.keybits) <= p.nb) ++p.keybits;   // keys range over [0, nb] (nb = sentinel)
    p.pairs = (uint64_t)n * p.W;
    uint64_t L = p.pairs / ((uint64_t)ctx->sm_count * 2048);
    L = std::max<uint64_t>(16, std::min<uint64_t>(256, L));
    if (ctx->chunk_override) L = ctx->chunk_override;
    p.L = (uint32_t)L;
    p.nchunks = (p.pairs + L - 1) / L;
    // segments for the bucket reduction: aim at >= 8192 threads in total
    uint32_t S = std::max<uint32_t>(1, p.nb / 16384);
    S = std::min<uint32_t>(S, p.half);
    p.S = S;
    p.nseg = (p.half + S - 1) / S;
    return p;
}

template <typename G>
static void msm_run_t(b200_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_gather, size_t n, bool scalars_mont,
                      void* d_out, int table_c, size_t table_stride) {
    typedef typename G::Pt Pt;
    typedef typename G::Aff Aff;
    cudaStream_t st = ctx->cur();
    if (n == 0) {
        // empty sum = infinity
        Pt* out = (Pt*)d_out;
        msm_horner<G><<<1, 1, 0, st>>>(nullptr, 0, 1, out);
        ctx->launches++;
        return;
    }
    if (n >= (1ull << 31)) throw std::runtime_error("msm: n must be < 2^31");
    if (table_c && (uint64_t)table_stride * ((255 + table_c - 1) / table_c) >= (1ull << 31)) throw std::runtime_error("msm: table too large for 31-bit point indices");
    MsmPlan p = make_plan(ctx, n, table_c);
    uint32_t* keys0 = (uint32_t*)ctx->scratch("msm_keys0", p.pairs * 4);
    uint32_t* keys1 = (uint32_t*)ctx->scratch("msm_keys1", p.pairs * 4);
    uint32_t* vals0 = (uint32_t*)ctx->scratch("msm_vals0", p.pairs * 4);
    uint32_t* vals1 = (uint32_t*)ctx->scratch("msm_vals1", p.pairs * 4);
    uint64_t* offsets = (uint64_t*)ctx->scratch("msm_offsets", ((size_t)p.nb + 2) * 8);
    Pt* buckets = (Pt*)ctx->scratch("msm_buckets", (size_t)p.nb * sizeof(Pt));
    Pt* pfirst = (Pt*)ctx->scratch("msm_pfirst", p.nchunks * sizeof(Pt));
    Pt* plast = (Pt*)ctx->scratch("msm_plast", p.nchunks * sizeof(Pt));
    const size_t big_cap = p.nchunks / B200_BIG_SPAN + 2;
    uint32_t* big = (uint32_t*)ctx->scratch("msm_big", (big_cap + 1) * 4);
    Pt* seg = (Pt*)ctx->scratch("msm_seg", ((size_t)p.nseg * p.W + p.nseg / 64 + 64) * sizeof(Pt));
    Pt* win = (Pt*)ctx->scratch("msm_win", (size_t)p.W * sizeof(Pt));

    ctx->t_begin("msm.digits");
    {
        dim3 grid((unsigned)((n + 255) / 256));
        if (scalars_mont) msm_digits<true><<<grid, 256, 0, st>>>((const Fr*)d_scalars, d_gather, (uint32_t)n, p.c, p.W, keys0, vals0, (uint32_t)table_stride, p.table);
        else msm_digits<false><<<grid, 256, 0, st>>>((const Fr*)d_scalars, d_gather, (uint32_t)n, p.c, p.W, keys0, vals0, (uint32_t)table_stride, p.table);
        ctx->launches++;
    }
    ctx->t_end();
    ctx->t_begin("msm.sort");
    cub::DoubleBuffer<uint32_t> kb(keys0, keys1), vb(vals0, vals1);
    {
        size_t tmp_bytes = 0;
        B200_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, kb, vb, p.pairs, 0, p.keybits, st));
        void* tmp = ctx->scratch("msm_sort_tmp", tmp_bytes);
        B200_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, kb, vb, p.pairs, 0, p.keybits, st));
    }
    const uint32_t* skeys = kb.Current();
    const uint32_t* svals = vb.Current();
    ctx->t_end();
    ctx->t_begin("msm.offsets");
    msm_offsets<<<(p.nb + 1 + 255) / 256, 256, 0, st>>>(skeys, p.pairs, p.nb, offsets);
    ctx->launches++;
    ctx->t_end();
    ctx->t_begin("msm.accumulate");
    msm_accumulate<G><<<(unsigned)((p.nchunks + 127) / 128), 128, 0, st>>>(skeys, svals, offsets, p.nb, p.L, (const Aff*)d_bases, buckets,
                                                                            pfirst, plast);
    ctx->launches++;
    ctx->t_end();
    ctx->t_begin("msm.finalize");
    B200_CUDA_CHECK(cudaMemsetAsync(big, 0, 4, st));
    msm_finalize<G><<<(p.nb + 127) / 128, 128, 0, st>>>(offsets, p.nb, p.L, buckets, pfirst, plast, big, big + 1);
    const size_t smem = 128 * sizeof(Pt);
    static bool attr_set[2] = {false, false};
    const int gi = sizeof(Pt) > 128 ? 1 : 0;
    if (!attr_set[gi]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(msm_finalize_big<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B200_CUDA_CHECK(cudaFuncSetAttribute(msm_window_reduce<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        B200_CUDA_CHECK(cudaFuncSetAttribute(msm_tree_reduce<G>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set[gi] = true;
    }
    msm_finalize_big<G><<<(unsigned)std::min<size_t>(big_cap, 1024), 128, smem, st>>>(offsets, p.L, buckets, pfirst, plast, big, big + 1);
    ctx->launches += 2;
    ctx->t_end();
    ctx->t_begin("msm.reduce");
    if (p.table) {
        // one bucket set: segment sums, then a tree of block reductions straight into the result
        msm_bucket_reduce<G><<<(p.nseg + 127) / 128, 128, 0, st>>>(buckets, p.half, p.S, p.nseg, 1, seg);
        ctx->launches++;
        Pt* cur = seg;
        uint32_t cnt = p.nseg;
        Pt* spare = seg + p.nseg;
        while (true) {
            const uint32_t R = cnt > 128 * 8 ? 8 : (cnt + 127) / 128;
            const uint32_t blocks = (cnt + R * 128 - 1) / (R * 128);
            Pt* dst = blocks == 1 ? (Pt*)d_out : spare;
            msm_tree_reduce<G><<<blocks, 128, smem, st>>>(cur, cnt, R, dst);
            ctx->launches++;
            if (blocks == 1) break;
            cur = dst; cnt = blocks; spare = dst + blocks;
        }
    } else {
        msm_bucket_reduce<G><<<(p.nseg * p.W + 127) / 128, 128, 0, st>>>(buckets, p.half, p.S, p.nseg, p.W, seg);
        msm_window_reduce<G><<<p.W, 128, smem, st>>>(seg, p.nseg, win);
        msm_horner<G><<<1, 32, 0, st>>>(win, p.W, p.c, (Pt*)d_out);
        ctx->launches += 3;
    }
    ctx->t_end();
    B200_CUDA_CHECK(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------- small group helpers
// out = sum(xyzz terms) + sum(affine terms), normalised to affine; layout of out: Aff followed by a u32 infinity flag.
template <typename G>
__global__ void g_finish_kernel(const typename G::Pt* xt, int nx, const typename G::Aff* at, int na, typename G::Aff* out, bool to_std) {
    typedef typename G::Pt Pt;
    if (threadIdx.x || blockIdx.x) return;
    Pt r = Pt::inf();
    for (int i = 0; i < nx; ++i) r.add(xt[i]);
    for (int i = 0; i < na; ++i) r.add_mixed(at[i]);
    typename G::Aff a = r.to_affine();
    uint32_t inf = r.is_inf() ? 1u : 0u;
    if (to_std) {
        // limb-wise from_mont of every Fq coordinate
        Fq* f = reinterpret_cast<Fq*>(&a);
        for (int i = 0; i < (int)(sizeof(a) / sizeof(Fq)); ++i) f[i] = f[i].from_mont();
    }
    *out = a;
    *reinterpret_cast<uint32_t*>(out + 1) = inf;
}
template <typename G>
static void g_finish_t(b200_ctx* ctx, const void* xt, int nx, const void* at, int na, void* out, bool to_std) {
    g_finish_kernel<G><<<1, 32, 0, ctx->stream>>>((const typename G::Pt*)xt, nx, (const typename G::Aff*)at, na, (typename G::Aff*)out, to_std);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

// out[i] = k_i * generator (affine, (0,0) for infinity); replaces batch_exp (multiexp.tcc:838-905) for synthetic keys
template <typename G>
__global__ void __launch_bounds__(128) g_fixed_base_kernel(const Fr* __restrict__ scalars, size_t n, typename G::Aff* __restrict__ out) {
    typedef typename G::Pt Pt;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr k = scalars[i].from_mont();
    const typename G::Aff g = Gen<typename G::F>::get();
    Pt r = Pt::inf();
    for (int l = 7; l >= 0; --l) {
        const uint32_t w = k.v[l];
        for (int b = 31; b >= 0; --b) {
            r.dbl();
            if ((w >> b) & 1) r.add_mixed(g);
        }
    }
    out[i] = r.to_affine();
}
template <typename G>
static void g_fixed_base_t(b200_ctx* ctx, const void* d_scalars, size_t n, void* d_out) {
    if (n == 0) return;
    unsigned grid = (unsigned)((n + 127) / 128);
    g_fixed_base_kernel<G><<<grid, 128, 0, ctx->stream>>>((const Fr*)d_scalars, n, (typename G::Aff*)d_out);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

template <typename G>
__global__ void g_add_affine_kernel(const typename G::Aff* a, const typename G::Aff* b, typename G::Aff* out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    typename G::Pt r = G::Pt::from_affine(a[i]);
    r.add_mixed(b[i]);
    out[i] = r.to_affine();
}
template <typename G>
static void g_add_affine_t(b200_ctx* ctx, const void* a, const void* b, void* out, size_t n) {
    if (n == 0) return;
    unsigned grid = (unsigned)((n + 63) / 64);
    g_add_affine_kernel<G><<<grid, 64, 0, ctx->stream>>>((const typename G::Aff*)a, (const typename G::Aff*)b, (typename G::Aff*)out, n);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

template <typename G>
__global__ void g_scalar_mul_kernel(const typename G::Aff* base, const Fr* scalar, typename G::Pt* out) {
    typedef typename G::Pt Pt;
    if (threadIdx.x || blockIdx.x) return;
    Fr k = scalar->from_mont();
    const typename G::Aff g = *base;
    Pt r = Pt::inf();
    for (int l = 7; l >= 0; --l) {
        const uint32_t w = k.v[l];
        for (int b = 31; b >= 0; --b) {
            r.dbl();
            if ((w >> b) & 1) r.add_mixed(g);
        }
    }
    *out = r;
}
template <typename G>
static void g_scalar_mul_t(b200_ctx* ctx, const void* base, const void* scalar, void* out) {
    g_scalar_mul_kernel<G><<<1, 32, 0, ctx->stream>>>((const typename G::Aff*)base, (const Fr*)scalar, (typename G::Pt*)out);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}


// builds the window table for n resident bases; returns c via the plan the caller chose
template <typename G>
static void g_precompute_t(b200_ctx* ctx, const void* d_bases, size_t n, int c, void* d_table) {
    if (n == 0) return;
    const int W = (255 + c - 1) / c;
    msm_precompute<G><<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>((const typename G::Aff*)d_bases, n, c, W, (typename G::Aff*)d_table);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

// Each group is instantiated in its own translation unit (msm_g1.cu / msm_g2.cu) so the two compile in parallel.
#define B200_INSTANTIATE_GROUP(G, SUFFIX)                                                                                              \
    void msm_run_##SUFFIX(b200_ctx* ctx, const void* b, const void* s, const uint32_t* g, size_t n, bool mont, void* out, int tc,     \
                          size_t ts) {                                                                                                 \
        msm_run_t<G>(ctx, b, s, g, n, mont, out, tc, ts);                                                                              \
    }                                                                                                                                  \
    void g_precompute_##SUFFIX(b200_ctx* ctx, const void* b, size_t n, int c, void* t) { g_precompute_t<G>(ctx, b, n, c, t); }         \
    void g_finish_##SUFFIX(b200_ctx* ctx, const void* xt, int nx, const void* at, int na, void* out, bool to_std) {                   \
        g_finish_t<G>(ctx, xt, nx, at, na, out, to_std);                                                                               \
    }                                                                                                                                  \
    void g_fixed_base_##SUFFIX(b200_ctx* ctx, const void* s, size_t n, void* out) { g_fixed_base_t<G>(ctx, s, n, out); }               \
    void g_add_affine_##SUFFIX(b200_ctx* ctx, const void* a, const void* b, void* out, size_t n) { g_add_affine_t<G>(ctx, a, b, out, n); } \
    void g_scalar_mul_##SUFFIX(b200_ctx* ctx, const void* base, const void* k, void* out) { g_scalar_mul_t<G>(ctx, base, k, out); }

}  // namespace b200
