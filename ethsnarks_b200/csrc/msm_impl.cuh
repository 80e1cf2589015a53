#pragma once
// MSM orchestration for one group: digits, sort, offsets, then the launchers of msm_acc.cuh / msm_red.cuh (see the
// stage list in msm_common.cuh).
#include "msm_common.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace b200 {

// ------------------------------------------------------------------------------------------------- 1. digits
template <bool MONT>
static __global__ void __launch_bounds__(256) msm_digits(const Fr* __restrict__ scalars, const uint32_t* __restrict__ gather, uint32_t n,
                                                  int c, int W, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals,
                                                  uint32_t stride, int table, int packed) {
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
        vals[(size_t)w * n + i] = (packed ? ((uint32_t)w << B200_SHARED_VAR_BITS) | i : table ? (uint32_t)w * stride + i : i) | (neg << 31);
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

// ------------------------------------------------------------------------------------------------- host side
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

static MsmPlan make_plan(b200_ctx* ctx, size_t n, int table_c, bool pair_lanes = false) {
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
    // chunk length: ~4 waves of 512 threads per SM, a multiple of 16 pairs (chunks then start on 64-byte boundaries of the
    // key / value arrays: 2^20 points, 45 -> 48: accumulate 2.28 -> 2.21 ms), and not so short that most runs cross
    // several chunk edges (measured on B200, G1: 2^18 1.68 -> 1.45 ms at 32, 2^19 2.48 -> 2.25 ms at 48)
    uint64_t L = p.pairs / ((uint64_t)ctx->sm_count * 2048);
    L = (L + 15) / 16 * 16;
    L = std::max<uint64_t>(p.pairs >= (6ull << 20) ? 48 : (p.pairs >= (3ull << 19) ? 32 : 16), std::min<uint64_t>(256, L));
    // the lane-pair G2 kernel runs two threads per chunk, so it fills the SMs with 1.5x longer chunks and pays fewer chunk
    // edges in the finalize (measured: 2^19 6.09 -> 5.79 ms at 64, 2^20 9.79 -> 9.72 ms at 80)
    if (pair_lanes && L >= 32) L = std::min<uint64_t>(256, (L * 3 / 2 + 15) / 16 * 16);
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

template <typename Pt>
static __global__ void msm_set_inf(Pt* out) { *out = Pt::inf(); }

// stages 1-3: digits, sort, offsets.  `tag` selects the scratch set (the shared list must survive the consumers' own scratch).
struct MsmSorted { const uint32_t* keys; const uint32_t* vals; uint64_t* offsets; };
static MsmSorted msm_sort_stage(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const void* d_scalars, const uint32_t* d_gather, size_t n,
                                bool scalars_mont, size_t table_stride, bool packed) {
    const char* nm[6] = {"msm_keys0", "msm_keys1", "msm_vals0", "msm_vals1", "msm_offsets", "msm_sort_tmp"};
    const char* nms[6] = {"msm_sh_keys0", "msm_sh_keys1", "msm_sh_vals0", "msm_sh_vals1", "msm_sh_offsets", "msm_sh_sort_tmp"};
    const char** names = packed ? nms : nm;
    uint32_t* keys0 = (uint32_t*)ctx->scratch(names[0], p.pairs * 4);
    uint32_t* keys1 = (uint32_t*)ctx->scratch(names[1], p.pairs * 4);
    uint32_t* vals0 = (uint32_t*)ctx->scratch(names[2], p.pairs * 4);
    uint32_t* vals1 = (uint32_t*)ctx->scratch(names[3], p.pairs * 4);
    uint64_t* offsets = (uint64_t*)ctx->scratch(names[4], ((size_t)p.nb + 2) * 8);
    ctx->t_begin("msm.digits");
    {
        dim3 grid((unsigned)((n + 255) / 256));
        if (scalars_mont) msm_digits<true><<<grid, 256, 0, st>>>((const Fr*)d_scalars, d_gather, (uint32_t)n, p.c, p.W, keys0, vals0, (uint32_t)table_stride, p.table, packed);
        else msm_digits<false><<<grid, 256, 0, st>>>((const Fr*)d_scalars, d_gather, (uint32_t)n, p.c, p.W, keys0, vals0, (uint32_t)table_stride, p.table, packed);
        ctx->launches++;
    }
    ctx->t_end();
    ctx->t_begin("msm.sort");
    cub::DoubleBuffer<uint32_t> kb(keys0, keys1), vb(vals0, vals1);
    {
        size_t tmp_bytes = 0;
        B200_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, kb, vb, p.pairs, 0, p.keybits, st));
        void* tmp = ctx->scratch(names[5], tmp_bytes);
        B200_CUDA_CHECK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, kb, vb, p.pairs, 0, p.keybits, st));
    }
    ctx->t_end();
    ctx->t_begin("msm.offsets");
    msm_offsets<<<(p.nb + 1 + 255) / 256, 256, 0, st>>>(kb.Current(), p.pairs, p.nb, offsets);
    ctx->launches++;
    ctx->t_end();
    return MsmSorted{kb.Current(), vb.Current(), offsets};
}

// shared list -> one consumer's values (internal.hpp: MsmShared)
static __global__ void __launch_bounds__(256) msm_translate(const uint32_t* __restrict__ vals, uint64_t pairs, MsmPos pos, uint32_t zero_index,
                                                            uint32_t* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= pairs) return;
    const uint32_t v = vals[i], idx = v & 0x7fffffffu, var = idx & ((1u << B200_SHARED_VAR_BITS) - 1);
    const uint32_t ps = pos.map ? pos.map[var] : var - pos.off;
    out[i] = ps < pos.count ? (v & 0x80000000u) | ((idx >> B200_SHARED_VAR_BITS) * pos.stride + ps) : zero_index;
}

// stages 4-6: accumulate, finalize, reduce over a sorted list
template <typename G>
static void msm_tail(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const uint32_t* akeys, const uint32_t* avals, const uint64_t* aoff,
                     const void* abases, void* d_out) {
    typedef typename G::Pt Pt;
    Pt* buckets = (Pt*)ctx->scratch("msm_buckets", (size_t)p.nb * sizeof(Pt));
    Pt* pfirst = (Pt*)ctx->scratch("msm_pfirst", p.nchunks * sizeof(Pt));
    Pt* plast = (Pt*)ctx->scratch("msm_plast", p.nchunks * sizeof(Pt));
    const size_t big_cap = p.nchunks / B200_BIG_SPAN + 2;
    uint32_t* big = (uint32_t*)ctx->scratch("msm_big", (big_cap + 1) * 4);
    Pt* seg = (Pt*)ctx->scratch("msm_seg", ((size_t)p.nseg * p.W + p.nseg / 64 + 64) * sizeof(Pt));
    Pt* win = (Pt*)ctx->scratch("msm_win", (size_t)p.W * sizeof(Pt));
    ctx->t_begin("msm.accumulate");
    launch_accumulate<G>(ctx, st, p, akeys, avals, aoff, abases, buckets, pfirst, plast);
    ctx->t_end();
    ctx->t_begin("msm.finalize");
    launch_finalize<G>(ctx, st, p, akeys, aoff, buckets, pfirst, plast, big, big_cap);
    ctx->t_end();
    ctx->t_begin("msm.reduce");
    launch_reduce<G>(ctx, st, p, buckets, seg, win, d_out);
    ctx->t_end();
    B200_CUDA_CHECK(cudaGetLastError());
}

template <typename G>
static void msm_run_t(b200_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_gather, size_t n, bool scalars_mont,
                      void* d_out, int table_c, size_t table_stride) {
    typedef typename G::Pt Pt;
    cudaStream_t st = ctx->cur();
    if (n == 0) {
        // empty sum = infinity
        Pt* out = (Pt*)d_out;
        msm_set_inf<Pt><<<1, 1, 0, st>>>(out);
        ctx->launches++;
        return;
    }
    if (n >= (1ull << 31)) throw std::runtime_error("msm: n must be < 2^31");
    if (table_c && (uint64_t)table_stride * ((255 + table_c - 1) / table_c) >= (1ull << 31)) throw std::runtime_error("msm: table too large for 31-bit point indices");
    MsmPlan p = make_plan(ctx, n, table_c, sizeof(Pt) > 128 && ctx->g2_pair != 0);
    const MsmSorted so = msm_sort_stage(ctx, st, p, d_scalars, d_gather, n, scalars_mont, table_stride, false);
    msm_tail<G>(ctx, st, p, so.keys, so.vals, so.offsets, d_bases, d_out);
}

// One MSM of a shared list (internal.hpp: MsmShared) against this group's window table (W * stride entries + the zero entry).
template <typename G>
static void msm_shared_consume_t(b200_ctx* ctx, const MsmShared& sh, const void* d_table, MsmPos pos, void* d_out) {
    typedef typename G::Pt Pt;
    cudaStream_t st = ctx->cur();
    MsmPlan p = make_plan(ctx, sh.n, sh.c, sizeof(Pt) > 128 && ctx->g2_pair != 0);
    if ((uint64_t)p.W * pos.stride + 1 >= (1ull << 31)) throw std::runtime_error("msm: table too large for 31-bit point indices");
    uint32_t* tvals = (uint32_t*)ctx->scratch("msm_sh_tvals", p.pairs * 4);
    ctx->t_begin("msm.translate");
    msm_translate<<<(unsigned)((p.pairs + 255) / 256), 256, 0, st>>>(sh.vals, p.pairs, pos, (uint32_t)p.W * pos.stride, tvals);
    ctx->launches++;
    ctx->t_end();
    msm_tail<G>(ctx, st, p, sh.keys, tvals, sh.offsets, d_table, d_out);
}

#define B200_INSTANTIATE_GROUP_RUN(G, SUFFIX)                                                                                          \
    void msm_run_##SUFFIX(b200_ctx* ctx, const void* b, const void* s, const uint32_t* g, size_t n, bool mont, void* out, int tc,     \
                          size_t ts) {                                                                                                 \
        msm_run_t<G>(ctx, b, s, g, n, mont, out, tc, ts);                                                                              \
    }                                                                                                                                  \
    void msm_shared_consume_##SUFFIX(b200_ctx* ctx, const MsmShared& sh, const void* table, MsmPos pos, void* out) {                   \
        msm_shared_consume_t<G>(ctx, sh, table, pos, out);                                                                             \
    }

}  // namespace b200
