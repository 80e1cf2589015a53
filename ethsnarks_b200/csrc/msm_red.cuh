#pragma once
// Stages 5 and 6 of the MSM (msm_common.cuh): stitching chunk partial sums into buckets, and the bucket reductions.
#include "msm_common.cuh"
#include "msm_acc_pair.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------- 5. finalize
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

// The same stitching with one thread per CHUNK for the additions: thread t owns the run that crosses from chunk t-1 into
// chunk t, if chunk t is the first one that run continues into.  With runs shorter than a chunk almost every chunk edge
// is such a crossing, so nearly all lanes of a warp execute the one addition together, where the per-bucket kernel above
// has ~half of its lanes idle while the others add.  The memory-only cases (empty bucket, run inside one chunk) stay per
// bucket: thread id handles bucket id and chunk id.
template <typename G>
__global__ void __launch_bounds__(128) msm_finalize_chunks(const uint32_t* __restrict__ keys, const uint64_t* __restrict__ offsets, uint32_t nb,
                                                           uint32_t L, typename G::Pt* __restrict__ buckets,
                                                           const typename G::Pt* __restrict__ pfirst, const typename G::Pt* __restrict__ plast,
                                                           uint32_t* __restrict__ big_count, uint32_t* __restrict__ big_list) {
    typedef typename G::Pt Pt;
    const uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t M = offsets[nb];
    if (id < nb) {
        const uint64_t s = offsets[id], e = offsets[id + 1];
        if (s == e) buckets[id] = Pt::inf();
        else {
            const uint64_t t0 = s / L, t1 = (e - 1) / L;
            if (t0 == t1) {
                if (s == t0 * L) buckets[id] = pfirst[t0];
                else if (e == min((t0 + 1) * (uint64_t)L, M)) buckets[id] = plast[t0];
            }
        }
    }
    const uint64_t pos = id * L;
    if (id == 0 || pos >= M) return;
    const uint32_t b = keys[pos];
    const uint64_t s = offsets[b];
    if (s >= pos || s / L != id - 1) return;          // no crossing here, or an earlier chunk already owns this run
    const uint64_t t0 = id - 1, t1 = (offsets[b + 1] - 1) / L;
    if (t1 - t0 > B200_BIG_SPAN) { big_list[atomicAdd(big_count, 1u)] = b; return; }
    Pt acc = (s == t0 * L) ? pfirst[t0] : plast[t0];
    for (uint64_t t = id; t <= t1; ++t) acc.add(pfirst[t]);
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

// Buckets spanning more than B200_BIG_SPAN chunks - with a real witness (mostly 0/1 values) that is the ONE bucket
// all the "1" scalars fall into, ~10^4 chunk partial sums at 2^20 - are summed in two levels: B200_BIG_SEGS blocks per bucket
// each sum a slice of its chunk partials (msm_finalize_big_seg), then one block per bucket adds the slice sums and the head of
// the run.  One block alone chained ~80 additions per thread (measured, 2^20 0/1 witness: 0.70 ms G1, 1.84 ms G2).  The first
// B200_BIG_FAST entries of the big list take this path; a longer list (many medium-sized runs) falls back to one block each.
#define B200_BIG_SEGS 64
#define B200_BIG_FAST 256
template <typename G>
__global__ void __launch_bounds__(128) msm_finalize_big_seg(const uint64_t* __restrict__ offsets, uint32_t L, const typename G::Pt* __restrict__ pfirst,
                                                            const uint32_t* __restrict__ big_count, const uint32_t* __restrict__ big_list,
                                                            typename G::Pt* __restrict__ segp) {
    typedef typename G::Pt Pt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Pt* sh = reinterpret_cast<Pt*>(smem_raw);
    const uint32_t nbig = min(*big_count, (uint32_t)B200_BIG_FAST);
    for (uint32_t j = blockIdx.y; j < nbig; j += gridDim.y) {
        const uint32_t b = big_list[j];
        const uint64_t t0 = offsets[b] / L, t1 = (offsets[b + 1] - 1) / L;
        const uint64_t per = (t1 - t0 + B200_BIG_SEGS - 1) / B200_BIG_SEGS;          // continuation chunks t0 + 1 .. t1
        const uint64_t lo = t0 + 1 + (uint64_t)blockIdx.x * per, hi = min(lo + per, t1 + 1);
        Pt acc = Pt::inf();
        for (uint64_t t = lo + threadIdx.x; t < hi; t += blockDim.x) acc.add(pfirst[t]);
        block_tree_reduce(acc, sh);
        if (threadIdx.x == 0) segp[(size_t)j * B200_BIG_SEGS + blockIdx.x] = acc;
    }
}
template <typename G>
__global__ void __launch_bounds__(128) msm_finalize_big(const uint64_t* __restrict__ offsets, uint32_t L,
                                                        typename G::Pt* __restrict__ buckets, const typename G::Pt* __restrict__ pfirst,
                                                        const typename G::Pt* __restrict__ plast, const uint32_t* __restrict__ big_count,
                                                        const uint32_t* __restrict__ big_list, const typename G::Pt* __restrict__ segp) {
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
        if (segp && j < B200_BIG_FAST) {
            for (uint32_t x = threadIdx.x; x < B200_BIG_SEGS; x += blockDim.x) acc.add(segp[(size_t)j * B200_BIG_SEGS + x]);
        } else {
            for (uint64_t t = t0 + 1 + threadIdx.x; t <= t1; t += blockDim.x) acc.add(pfirst[t]);
        }
        block_tree_reduce(acc, sh);
        if (threadIdx.x == 0) buckets[b] = acc;
    }
}

// The giant-bucket sums for G2 with one point per LANE PAIR: 64 pairs per 128-thread block, tree through shared memory.  A one-
// thread G2 addition is ~20 us of dependent work, the lane-pair form about half; on small real circuits this finalize is the
// longest link of the B(G2) lane (EdDSA: 0.66 of 1.42 ms).  Same inputs and outputs as msm_finalize_big_seg / msm_finalize_big.
template <int MODE>
B200_DEV void block_tree_reduce_pair(const PairLane<MODE>& ln, HalfXYZZ& acc, G2XYZZ* sh) {     // blockDim.x == 128: 64 pairs; result in pair 0
    const uint32_t pair = threadIdx.x >> 1;
    pair_store(&sh[pair], ln.imag, acc);
    __syncthreads();
    for (uint32_t stride = 32; stride > 0; stride >>= 1) {
        if (pair < stride) {
            pair_add_full(ln, acc, pair_load(&sh[pair + stride], ln.imag));
            pair_store(&sh[pair], ln.imag, acc);
        }
        __syncthreads();
    }
}
template <typename G>
__global__ void __launch_bounds__(128) msm_finalize_big_seg_pair(const uint64_t* __restrict__ offsets, uint32_t L, const G2XYZZ* __restrict__ pfirst,
                                                                 const uint32_t* __restrict__ big_count, const uint32_t* __restrict__ big_list,
                                                                 G2XYZZ* __restrict__ segp) {
    __shared__ G2XYZZ sh[64];
    PairLane<2> ln;
    ln.imag = threadIdx.x & 1;
    ln.mask = 3u << (threadIdx.x & 30);
    const uint32_t pair = threadIdx.x >> 1;
    const uint32_t nbig = min(*big_count, (uint32_t)B200_BIG_FAST);
    for (uint32_t j = blockIdx.y; j < nbig; j += gridDim.y) {
        const uint32_t b = big_list[j];
        const uint64_t t0 = offsets[b] / L, t1 = (offsets[b + 1] - 1) / L;
        const uint64_t per = (t1 - t0 + B200_BIG_SEGS - 1) / B200_BIG_SEGS;
        const uint64_t lo = t0 + 1 + (uint64_t)blockIdx.x * per, hi = min(lo + per, t1 + 1);
        HalfXYZZ acc = ln.inf();
        for (uint64_t t = lo + pair; t < hi; t += 64) pair_add_full(ln, acc, pair_load(&pfirst[t], ln.imag));
        block_tree_reduce_pair(ln, acc, sh);
        if (pair == 0) pair_store(&segp[(size_t)j * B200_BIG_SEGS + blockIdx.x], ln.imag, acc);
        __syncthreads();
    }
}
template <typename G>
__global__ void __launch_bounds__(128) msm_finalize_big_pair(const uint64_t* __restrict__ offsets, uint32_t L, G2XYZZ* __restrict__ buckets,
                                                             const G2XYZZ* __restrict__ pfirst, const G2XYZZ* __restrict__ plast,
                                                             const uint32_t* __restrict__ big_count, const uint32_t* __restrict__ big_list,
                                                             const G2XYZZ* __restrict__ segp) {
    __shared__ G2XYZZ sh[64];
    PairLane<2> ln;
    ln.imag = threadIdx.x & 1;
    ln.mask = 3u << (threadIdx.x & 30);
    const uint32_t pair = threadIdx.x >> 1;
    const uint32_t nbig = *big_count;
    for (uint32_t j = blockIdx.x; j < nbig; j += gridDim.x) {
        const uint32_t b = big_list[j];
        const uint64_t s = offsets[b], e = offsets[b + 1];
        const uint64_t t0 = s / L, t1 = (e - 1) / L;
        HalfXYZZ acc = ln.inf();
        if (pair == 0) acc = pair_load((s == t0 * L) ? &pfirst[t0] : &plast[t0], ln.imag);
        if (segp && j < B200_BIG_FAST) {
            for (uint32_t x = pair; x < B200_BIG_SEGS; x += 64) pair_add_full(ln, acc, pair_load(&segp[(size_t)j * B200_BIG_SEGS + x], ln.imag));
        } else {
            for (uint64_t t = t0 + 1 + pair; t <= t1; t += 64) pair_add_full(ln, acc, pair_load(&pfirst[t], ln.imag));
        }
        block_tree_reduce_pair(ln, acc, sh);
        if (pair == 0) pair_store(&buckets[b], ln.imag, acc);
        __syncthreads();
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
// one block per output: out[w] = sum of seg[w*nseg .. min((w+1)*nseg, total))
template <typename G>
__global__ void __launch_bounds__(128) msm_window_reduce(const typename G::Pt* __restrict__ seg, uint32_t nseg, uint32_t total,
                                                         typename G::Pt* __restrict__ win) {
    typedef typename G::Pt Pt;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    Pt* sh = reinterpret_cast<Pt*>(smem_raw);
    const uint32_t w = blockIdx.x;
    const uint32_t lo = w * nseg, hi = min(lo + nseg, total);
    Pt acc = Pt::inf();
    for (uint32_t s = lo + threadIdx.x; s < hi; s += blockDim.x) acc.add(seg[s]);
    block_tree_reduce(acc, sh);
    if (threadIdx.x == 0) win[w] = acc;
}
// ---- two-dimensional bucket reduction --------------------------------------------------------------------------
// The running-sum reduction above is a chain of 2*S dependent additions per thread; at 2^19 buckets that chain, not the
// arithmetic, sets the time.  Here the buckets of a window are a matrix [Rn rows][S = 2^s columns], b = hi*S + lo, and
//     sum_b (b+1) B_b  =  sum_lo (lo+1) C_lo  +  S * sum_hi hi R_hi ,   C = column sums, R = row sums,
// so the bulk of the work (2 additions per bucket) is plain sums with no dependence between threads (stages A, B).
// The two short weighted sums are taken bit by bit: U_i = sum of the C_lo with bit i of (lo+1) set plus the R_hi with
// bit (i-s) of hi set, then sum_i 2^i U_i with the 2^i applied by i doublings in parallel (stages C, D, E).
// Stage A: K-term partial column sums pc[g][lo] and partial row sums pr[hi][lg].
template <typename G>
__global__ void __launch_bounds__(128) msm_rc_partial(const typename G::Pt* __restrict__ buckets, uint32_t half, uint32_t s, uint32_t K,
                                                      typename G::Pt* __restrict__ pc, typename G::Pt* __restrict__ pr) {
    typedef typename G::Pt Pt;
    const uint32_t S = 1u << s, Rn = half >> s;
    const uint32_t Kc = min(K, Rn), Kr = min(K, S);
    const uint32_t ncol = (Rn / Kc) * S, nrow = Rn * (S / Kr);
    const Pt* B = buckets + (size_t)blockIdx.y * half;
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id < ncol) {
        const uint32_t g = id >> s, lo = id & (S - 1);
        const Pt* src = B + (size_t)g * Kc * S + lo;
        Pt acc = src[0];
        for (uint32_t k = 1; k < Kc; ++k) acc.add(src[(size_t)k * S]);
        pc[(size_t)blockIdx.y * ncol + id] = acc;
        return;
    }
    id -= ncol;
    if (id >= nrow) return;
    const Pt* src = B + (size_t)id * Kr;     // row hi = id / (S/Kr), group lg = id % (S/Kr): consecutive buckets
    Pt acc = src[0];
    for (uint32_t k = 1; k < Kr; ++k) acc.add(src[k]);
    pr[(size_t)blockIdx.y * nrow + id] = acc;
}
#ifndef B200_G2_PAIR_RED_MINB
#define B200_G2_PAIR_RED_MINB 3     // resident blocks per SM the lane-pair full-addition kernels are compiled for
#endif
// Stage A for G2 with one output per LANE PAIR (msm_acc_pair.cuh: pair_add_full); same outputs as msm_rc_partial<G2T>.
template <typename G>
__global__ void __launch_bounds__(128, B200_G2_PAIR_RED_MINB) msm_rc_partial_pair(const G2XYZZ* __restrict__ buckets, uint32_t half, uint32_t s, uint32_t K,
                                                                          G2XYZZ* __restrict__ pc, G2XYZZ* __restrict__ pr) {
    const uint32_t S = 1u << s, Rn = half >> s;
    const uint32_t Kc = min(K, Rn), Kr = min(K, S);
    const uint32_t ncol = (Rn / Kc) * S, nrow = Rn * (S / Kr);
    const G2XYZZ* B = buckets + (size_t)blockIdx.y * half;
    const uint32_t gt = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t id = gt >> 1;
    PairLane<2> ln;
    ln.imag = gt & 1;
    ln.mask = 3u << (threadIdx.x & 30);
    if (id < ncol) {
        const uint32_t g = id >> s, lo = id & (S - 1);
        const G2XYZZ* src = B + (size_t)g * Kc * S + lo;
        HalfXYZZ acc = pair_load(src, ln.imag);
        for (uint32_t k = 1; k < Kc; ++k) pair_add_full(ln, acc, pair_load(src + (size_t)k * S, ln.imag));
        pair_store(&pc[(size_t)blockIdx.y * ncol + id], ln.imag, acc);
        return;
    }
    id -= ncol;
    if (id >= nrow) return;
    const G2XYZZ* src = B + (size_t)id * Kr;
    HalfXYZZ acc = pair_load(src, ln.imag);
    for (uint32_t k = 1; k < Kr; ++k) pair_add_full(ln, acc, pair_load(src + k, ln.imag));
    pair_store(&pr[(size_t)blockIdx.y * nrow + id], ln.imag, acc);
}
// msm_finalize_chunks for G2 with one chunk edge per lane pair
template <typename G>
__global__ void __launch_bounds__(128, B200_G2_PAIR_RED_MINB) msm_finalize_chunks_pair(const uint32_t* __restrict__ keys, const uint64_t* __restrict__ offsets,
                                                                               uint32_t nb, uint32_t L, G2XYZZ* __restrict__ buckets,
                                                                               const G2XYZZ* __restrict__ pfirst, const G2XYZZ* __restrict__ plast,
                                                                               uint32_t* __restrict__ big_count, uint32_t* __restrict__ big_list) {
    const uint64_t gt = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t id = gt >> 1;
    PairLane<2> ln;
    ln.imag = gt & 1;
    ln.mask = 3u << (threadIdx.x & 30);
    const uint64_t M = offsets[nb];
    if (id < nb) {
        const uint64_t s = offsets[id], e = offsets[id + 1];
        if (s == e) pair_store(&buckets[id], ln.imag, ln.inf());
        else {
            const uint64_t t0 = s / L, t1 = (e - 1) / L;
            if (t0 == t1) {
                if (s == t0 * L) pair_store(&buckets[id], ln.imag, pair_load(&pfirst[t0], ln.imag));
                else if (e == min((t0 + 1) * (uint64_t)L, M)) pair_store(&buckets[id], ln.imag, pair_load(&plast[t0], ln.imag));
            }
        }
    }
    const uint64_t pos = id * L;
    if (id == 0 || pos >= M) return;
    const uint32_t b = keys[pos];
    const uint64_t s = offsets[b];
    if (s >= pos || s / L != id - 1) return;
    const uint64_t t0 = id - 1, t1 = (offsets[b + 1] - 1) / L;
    if (t1 - t0 > B200_BIG_SPAN) { if (!ln.imag) big_list[atomicAdd(big_count, 1u)] = b; return; }
    HalfXYZZ acc = pair_load((s == t0 * L) ? &pfirst[t0] : &plast[t0], ln.imag);
    for (uint64_t t = id; t <= t1; ++t) pair_add_full(ln, acc, pair_load(&pfirst[t], ln.imag));
    pair_store(&buckets[b], ln.imag, acc);
}

// Warp-level tree sum of one point per lane; the result is in lane 0.  The stages below run one warp per output so that
// no resident warp sits idle at a block barrier holding 255 registers per thread.
template <typename Pt>
B200_DEV void warp_tree_reduce(Pt& acc) {
    constexpr int NW = sizeof(Pt) / 4;
    const uint32_t lane = threadIdx.x & 31;
#pragma unroll 1
    for (int off = 16; off > 0; off >>= 1) {
        Pt other;
        uint32_t* o = reinterpret_cast<uint32_t*>(&other);
        const uint32_t* a = reinterpret_cast<const uint32_t*>(&acc);
#pragma unroll
        for (int w = 0; w < NW; ++w) o[w] = __shfl_down_sync(0xffffffffu, a[w], off);
        if (lane < (uint32_t)off) acc.add(other);
    }
}
// Stage B: one warp per column (blockIdx.x < S) or row: V[lo] = C_lo, V[S + hi] = R_hi.
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_sum(const typename G::Pt* __restrict__ pc, const typename G::Pt* __restrict__ pr,
                                                 uint32_t half, uint32_t s, uint32_t K, typename G::Pt* __restrict__ V) {
    typedef typename G::Pt Pt;
    const uint32_t S = 1u << s, Rn = half >> s;
    const uint32_t Kc = min(K, Rn), Kr = min(K, S);
    const uint32_t ncol = (Rn / Kc) * S, nrow = Rn * (S / Kr);
    const uint32_t o = blockIdx.x;
    const Pt* src;
    uint32_t cnt, stride;
    if (o < S) { src = pc + (size_t)blockIdx.y * ncol + o; cnt = Rn / Kc; stride = S; }
    else { src = pr + (size_t)blockIdx.y * nrow + (size_t)(o - S) * (S / Kr); cnt = S / Kr; stride = 1; }
    Pt acc = Pt::inf();
    for (uint32_t t = threadIdx.x; t < cnt; t += 32) acc.add(src[(size_t)t * stride]);
    warp_tree_reduce(acc);
    if (threadIdx.x == 0) V[(size_t)blockIdx.y * (S + Rn) + o] = acc;
}
// Stage B as halving passes, one THREAD per addition: pass j adds the upper half of the surviving partial column sums onto
// the lower half (pc is [group][lo], so that is pc[i] += pc[i + n/2]) and the upper half of every row's surviving partial
// sums onto its lower half.  Every pass is one addition deep with as many threads as additions, where the warp-per-output
// kernel above chains cnt/32 + 5 additions in a few hundred warps (G2, 2^19 buckets: 1.14 ms against 7 x ~35 us).
// The last pass of each kind writes V.  colcur = surviving column partials per bucket set (> S), rwl = log2 of the row
// width AFTER this pass; either side may already be finished (colcur == S / rows_active == 0).
template <typename G>
__global__ void __launch_bounds__(128) msm_rc_fold(typename G::Pt* __restrict__ pc, uint32_t ncol, uint32_t colcur,
                                                   typename G::Pt* __restrict__ pr, uint32_t nrow, uint32_t rw0, uint32_t rwl, uint32_t rows_active,
                                                   uint32_t S, uint32_t Rn, typename G::Pt* __restrict__ V) {
    typedef typename G::Pt Pt;
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t nc2 = colcur > S ? colcur >> 1 : 0;
    if (id < nc2) {
        Pt* src = pc + (size_t)blockIdx.y * ncol;
        Pt a = src[id];
        a.add(src[id + nc2]);
        if (nc2 == S) V[(size_t)blockIdx.y * (S + Rn) + id] = a; else src[id] = a;
        return;
    }
    id -= nc2;
    if (!rows_active || id >= (Rn << rwl)) return;
    const uint32_t hi = id >> rwl, lg = id & ((1u << rwl) - 1);
    Pt* src = pr + (size_t)blockIdx.y * nrow + (size_t)hi * rw0;
    Pt a = src[lg];
    a.add(src[lg + (1u << rwl)]);
    if (rwl == 0) V[(size_t)blockIdx.y * (S + Rn) + S + hi] = a; else src[lg] = a;
}
// msm_rc_fold for G2 with one addition per lane pair (half the latency of the one-thread addition per pass)
template <typename G>
__global__ void __launch_bounds__(128, B200_G2_PAIR_RED_MINB) msm_rc_fold_pair(G2XYZZ* __restrict__ pc, uint32_t ncol, uint32_t colcur, G2XYZZ* __restrict__ pr,
                                                                       uint32_t nrow, uint32_t rw0, uint32_t rwl, uint32_t rows_active, uint32_t S,
                                                                       uint32_t Rn, G2XYZZ* __restrict__ V) {
    const uint32_t gt = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t id = gt >> 1;
    PairLane<2> ln;
    ln.imag = gt & 1;
    ln.mask = 3u << (threadIdx.x & 30);
    const uint32_t nc2 = colcur > S ? colcur >> 1 : 0;
    if (id < nc2) {
        G2XYZZ* src = pc + (size_t)blockIdx.y * ncol;
        HalfXYZZ a = pair_load(&src[id], ln.imag);
        pair_add_full(ln, a, pair_load(&src[id + nc2], ln.imag));
        pair_store(nc2 == S ? &V[(size_t)blockIdx.y * (S + Rn) + id] : &src[id], ln.imag, a);
        return;
    }
    id -= nc2;
    if (!rows_active || id >= (Rn << rwl)) return;
    const uint32_t hi = id >> rwl, lg = id & ((1u << rwl) - 1);
    G2XYZZ* src = pr + (size_t)blockIdx.y * nrow + (size_t)hi * rw0;
    HalfXYZZ a = pair_load(&src[lg], ln.imag);
    pair_add_full(ln, a, pair_load(&src[lg + (1u << rwl)], ln.imag));
    pair_store(rwl == 0 ? &V[(size_t)blockIdx.y * (S + Rn) + S + hi] : &src[lg], ln.imag, a);
}
// Stage C: one warp per (weight bit i, slice of B200_RC_SLICE entries of V): sum of the selected C_lo and R_hi.
#define B200_RC_SLICE 128
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_bits(const typename G::Pt* __restrict__ V, uint32_t half, uint32_t s,
                                                  uint32_t nbits, uint32_t nslice, typename G::Pt* __restrict__ Us) {
    typedef typename G::Pt Pt;
    const uint32_t S = 1u << s, Rn = half >> s;
    const uint32_t i = blockIdx.x / nslice, sl = blockIdx.x % nslice;
    const Pt* v = V + (size_t)blockIdx.z * (S + Rn);
    Pt acc = Pt::inf();
    const uint32_t end = min((sl + 1) * B200_RC_SLICE, S + Rn);
    for (uint32_t t = sl * B200_RC_SLICE + threadIdx.x; t < end; t += 32) {
        const uint32_t weight = t < S ? t + 1 : (t - S) << s;
        if ((weight >> i) & 1u) acc.add(v[t]);
    }
    warp_tree_reduce(acc);
    if (threadIdx.x == 0) Us[((size_t)blockIdx.z * nbits + i) * nslice + sl] = acc;
}
// Stage D: one warp per weight bit: U_i = 2^i * (sum of its slices).
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_shift(const typename G::Pt* __restrict__ Us, uint32_t nbits, uint32_t nslice,
                                                   typename G::Pt* __restrict__ U) {
    typedef typename G::Pt Pt;
    const uint32_t i = blockIdx.x;
    const Pt* src = Us + ((size_t)blockIdx.z * nbits + i) * nslice;
    Pt acc = Pt::inf();
    for (uint32_t t = threadIdx.x; t < nslice; t += 32) acc.add(src[t]);
    warp_tree_reduce(acc);
    if (threadIdx.x == 0) {
        for (uint32_t k = 0; k < i; ++k) acc.dbl();
        U[(size_t)blockIdx.z * nbits + i] = acc;
    }
}
// Stage E: one warp per window: sum of its U_i (nbits <= 32).
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_final(const typename G::Pt* __restrict__ U, uint32_t nbits, typename G::Pt* __restrict__ out) {
    typedef typename G::Pt Pt;
    Pt acc = Pt::inf();
    if (threadIdx.x < nbits) acc = U[(size_t)blockIdx.x * nbits + threadIdx.x];
    warp_tree_reduce(acc);
    if (threadIdx.x == 0) out[blockIdx.x] = acc;
}

// ---- stages B - E for G2 with one point per LANE PAIR (16 points per warp).  These stages are chains of dependent group
// operations in a few warps; a lone warp is bound by its own IMAD issue rate, so halving the instructions per operation
// (each lane of a pair computes one Fq2 component, msm_acc_pair.cuh) halves the latency of every link of the chain.
template <int MODE>
B200_DEV void warp_tree_reduce_pair(const PairLane<MODE>& ln, HalfXYZZ& acc) {      // result in pair 0 (lanes 0, 1)
    const uint32_t pair = (threadIdx.x & 31) >> 1;
#pragma unroll 1
    for (int off = 8; off > 0; off >>= 1) {
        HalfXYZZ other;
        uint32_t* o = reinterpret_cast<uint32_t*>(&other);
        const uint32_t* a = reinterpret_cast<const uint32_t*>(&acc);
#pragma unroll
        for (int w = 0; w < (int)(sizeof(HalfXYZZ) / 4); ++w) o[w] = __shfl_down_sync(0xffffffffu, a[w], 2 * off);
        if (pair < (uint32_t)off) pair_add_full(ln, acc, other);
    }
}
B200_DEV PairLane<2> warp_pair_lane() {
    PairLane<2> ln;
    ln.imag = threadIdx.x & 1;
    ln.mask = 3u << (threadIdx.x & 30);
    return ln;
}
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_sum_pair(const G2XYZZ* __restrict__ pc, const G2XYZZ* __restrict__ pr, uint32_t half, uint32_t s, uint32_t K,
                                                      G2XYZZ* __restrict__ V) {
    const uint32_t S = 1u << s, Rn = half >> s;
    const uint32_t Kc = min(K, Rn), Kr = min(K, S);
    const uint32_t ncol = (Rn / Kc) * S, nrow = Rn * (S / Kr);
    const uint32_t o = blockIdx.x;
    const G2XYZZ* src;
    uint32_t cnt, stride;
    if (o < S) { src = pc + (size_t)blockIdx.y * ncol + o; cnt = Rn / Kc; stride = S; }
    else { src = pr + (size_t)blockIdx.y * nrow + (size_t)(o - S) * (S / Kr); cnt = S / Kr; stride = 1; }
    const PairLane<2> ln = warp_pair_lane();
    HalfXYZZ acc = ln.inf();
    for (uint32_t t = threadIdx.x >> 1; t < cnt; t += 16) pair_add_full(ln, acc, pair_load(src + (size_t)t * stride, ln.imag));
    warp_tree_reduce_pair(ln, acc);
    if (threadIdx.x < 2) pair_store(&V[(size_t)blockIdx.y * (S + Rn) + o], ln.imag, acc);
}
#define B200_RC_SLICE_PAIR 64
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_bits_pair(const G2XYZZ* __restrict__ V, uint32_t half, uint32_t s, uint32_t nbits, uint32_t nslice,
                                                       G2XYZZ* __restrict__ Us) {
    const uint32_t S = 1u << s, Rn = half >> s;
    const uint32_t i = blockIdx.x / nslice, sl = blockIdx.x % nslice;
    const G2XYZZ* v = V + (size_t)blockIdx.z * (S + Rn);
    const PairLane<2> ln = warp_pair_lane();
    HalfXYZZ acc = ln.inf();
    const uint32_t end = min((sl + 1) * B200_RC_SLICE_PAIR, S + Rn);
    for (uint32_t t = sl * B200_RC_SLICE_PAIR + (threadIdx.x >> 1); t < end; t += 16) {
        const uint32_t weight = t < S ? t + 1 : (t - S) << s;
        if ((weight >> i) & 1u) pair_add_full(ln, acc, pair_load(&v[t], ln.imag));
    }
    warp_tree_reduce_pair(ln, acc);
    if (threadIdx.x < 2) pair_store(&Us[((size_t)blockIdx.z * nbits + i) * nslice + sl], ln.imag, acc);
}
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_shift_pair(const G2XYZZ* __restrict__ Us, uint32_t nbits, uint32_t nslice, G2XYZZ* __restrict__ U) {
    const uint32_t i = blockIdx.x;
    const G2XYZZ* src = Us + ((size_t)blockIdx.z * nbits + i) * nslice;
    const PairLane<2> ln = warp_pair_lane();
    HalfXYZZ acc = ln.inf();
    for (uint32_t t = threadIdx.x >> 1; t < nslice; t += 16) pair_add_full(ln, acc, pair_load(&src[t], ln.imag));
    warp_tree_reduce_pair(ln, acc);
    if (threadIdx.x < 2) {
        for (uint32_t k = 0; k < i; ++k) acc = pair_dbl(ln, acc);
        pair_store(&U[(size_t)blockIdx.z * nbits + i], ln.imag, acc);
    }
}
template <typename G>
__global__ void __launch_bounds__(32) msm_rc_final_pair(const G2XYZZ* __restrict__ U, uint32_t nbits, G2XYZZ* __restrict__ out) {
    const PairLane<2> ln = warp_pair_lane();
    HalfXYZZ acc = ln.inf();
    for (uint32_t t = threadIdx.x >> 1; t < nbits; t += 16) pair_add_full(ln, acc, pair_load(&U[(size_t)blockIdx.x * nbits + t], ln.imag));
    warp_tree_reduce_pair(ln, acc);
    if (threadIdx.x < 2) pair_store(&out[blockIdx.x], ln.imag, acc);
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

// ------------------------------------------------------------------------------------------------- launchers
template <typename G>
void launch_finalize(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const uint32_t* keys, const uint64_t* offsets, void* buckets_,
                     const void* pfirst_, const void* plast_, uint32_t* big, size_t big_cap) {
    typedef typename G::Pt Pt;
    Pt* buckets = (Pt*)buckets_;
    const Pt* pfirst = (const Pt*)pfirst_;
    const Pt* plast = (const Pt*)plast_;
    const size_t smem = 128 * sizeof(Pt);     // <= 32 KB, inside the default dynamic shared memory limit
    B200_CUDA_CHECK(cudaMemsetAsync(big, 0, 4, st));
    if (ctx->finalize_chunks) {
        const uint64_t threads = std::max<uint64_t>(p.nb, p.nchunks);
        if constexpr (sizeof(Pt) > 128) {
            if (ctx->pair_reduce)
                msm_finalize_chunks_pair<G><<<(unsigned)((2 * threads + 127) / 128), 128, 0, st>>>(
                    keys, offsets, p.nb, p.L, (G2XYZZ*)buckets, (const G2XYZZ*)pfirst, (const G2XYZZ*)plast, big, big + 1);
            else
                msm_finalize_chunks<G><<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(keys, offsets, p.nb, p.L, buckets, pfirst, plast, big, big + 1);
        } else {
            msm_finalize_chunks<G><<<(unsigned)((threads + 127) / 128), 128, 0, st>>>(keys, offsets, p.nb, p.L, buckets, pfirst, plast, big, big + 1);
        }
    } else {
        msm_finalize<G><<<(p.nb + 127) / 128, 128, 0, st>>>(offsets, p.nb, p.L, buckets, pfirst, plast, big, big + 1);
    }
    Pt* segp = nullptr;
    bool pair_big = false;
    if constexpr (sizeof(Pt) > 128) pair_big = ctx->pair_reduce != 0 && ctx->pair_tail != 0;
    if (ctx->finalize_big_seg) {
        segp = (Pt*)ctx->scratch("msm_big_seg", (size_t)B200_BIG_FAST * B200_BIG_SEGS * sizeof(Pt));
        if constexpr (sizeof(Pt) > 128) {
            if (pair_big) msm_finalize_big_seg_pair<G><<<dim3(B200_BIG_SEGS, 8), 128, 0, st>>>(offsets, p.L, (const G2XYZZ*)pfirst, big, big + 1, (G2XYZZ*)segp);
        }
        if (!pair_big) msm_finalize_big_seg<G><<<dim3(B200_BIG_SEGS, 8), 128, smem, st>>>(offsets, p.L, pfirst, big, big + 1, segp);
        ctx->launches++;
    }
    if constexpr (sizeof(Pt) > 128) {
        if (pair_big) msm_finalize_big_pair<G><<<(unsigned)std::min<size_t>(big_cap, 1024), 128, 0, st>>>(
                          offsets, p.L, (G2XYZZ*)buckets, (const G2XYZZ*)pfirst, (const G2XYZZ*)plast, big, big + 1, (const G2XYZZ*)segp);
    }
    if (!pair_big) msm_finalize_big<G><<<(unsigned)std::min<size_t>(big_cap, 1024), 128, smem, st>>>(offsets, p.L, buckets, pfirst, plast, big, big + 1, segp);
    ctx->launches += 2;
}

template <typename G>
void launch_reduce(b200_ctx* ctx, cudaStream_t st, const MsmPlan& p, const void* buckets_, void* seg_, void* win_, void* d_out) {
    typedef typename G::Pt Pt;
    const Pt* B = (const Pt*)buckets_;
    Pt* seg = (Pt*)seg_;
    Pt* win = (Pt*)win_;
    const size_t smem = 128 * sizeof(Pt);
    const int Wr = p.table ? 1 : p.W;                 // bucket sets to reduce
    if (p.half >= 1024 && !ctx->reduce_1d) {
        // two-dimensional reduction (see msm_rc_partial): four launches, no long dependent chain
        const uint32_t s = (uint32_t)(p.c - 1 + 1) / 2, S = 1u << s, Rn = p.half >> s, K = 8;
        const uint32_t Kc = std::min(K, Rn), Kr = std::min(K, S);
        const uint32_t ncol = (Rn / Kc) * S, nrow = Rn * (S / Kr);
        const uint32_t nbits = std::max<uint32_t>(s + 1, (uint32_t)p.c - 1);
        Pt* pc = (Pt*)ctx->scratch("msm_rc_pc", (size_t)Wr * ncol * sizeof(Pt));
        Pt* pr = (Pt*)ctx->scratch("msm_rc_pr", (size_t)Wr * nrow * sizeof(Pt));
        Pt* V = (Pt*)ctx->scratch("msm_rc_v", (size_t)Wr * (S + Rn) * sizeof(Pt));
        Pt* U = (Pt*)ctx->scratch("msm_rc_u", (size_t)Wr * nbits * sizeof(Pt));
        bool pair_a = false, pair_t = false;
        if constexpr (sizeof(Pt) > 128) { pair_a = ctx->pair_reduce != 0; pair_t = ctx->pair_reduce != 0 && ctx->pair_tail != 0; }
        const uint32_t slice = pair_t ? B200_RC_SLICE_PAIR : B200_RC_SLICE;
        const uint32_t nslice = (S + Rn + slice - 1) / slice;
        Pt* Us = (Pt*)ctx->scratch("msm_rc_us", (size_t)Wr * nbits * nslice * sizeof(Pt));
        if constexpr (sizeof(Pt) > 128) {
            if (pair_a) msm_rc_partial_pair<G><<<dim3((2 * (ncol + nrow) + 127) / 128, Wr), 128, 0, st>>>((const G2XYZZ*)B, p.half, s, K, (G2XYZZ*)pc, (G2XYZZ*)pr);
        }
        if (!pair_a) msm_rc_partial<G><<<dim3((ncol + nrow + 127) / 128, Wr), 128, 0, st>>>(B, p.half, s, K, pc, pr);
        if ((ctx->reduce_fold == 2 || (ctx->reduce_fold == 1 && p.half >= (1u << 17))) && Rn / Kc >= 2 && S / Kr >= 2) {
            uint32_t colcur = ncol, rw = S / Kr;
            while (colcur > S || rw > 1) {
                const uint32_t nc2 = colcur > S ? colcur / 2 : 0;
                uint32_t rwl = 0;
                while (rw > 1 && (2u << rwl) < rw) ++rwl;          // log2(rw / 2)
                const uint32_t nr2 = rw > 1 ? Rn << rwl : 0;
                bool pair_f = false;
                if constexpr (sizeof(Pt) > 128) {
                    pair_f = ctx->pair_reduce != 0;
                    if (pair_f) msm_rc_fold_pair<G><<<dim3((2 * (nc2 + nr2) + 127) / 128, Wr), 128, 0, st>>>(
                                    (G2XYZZ*)pc, ncol, colcur, (G2XYZZ*)pr, nrow, S / Kr, rwl, rw > 1, S, Rn, (G2XYZZ*)V);
                }
                if (!pair_f) msm_rc_fold<G><<<dim3((nc2 + nr2 + 127) / 128, Wr), 128, 0, st>>>(pc, ncol, colcur, pr, nrow, S / Kr, rwl, rw > 1, S, Rn, V);
                ctx->launches++;
                if (colcur > S) colcur /= 2;
                if (rw > 1) rw /= 2;
            }
            ctx->launches--;      // the common count below assumes one stage-B launch
        } else {
            if constexpr (sizeof(Pt) > 128) {
                if (pair_t) msm_rc_sum_pair<G><<<dim3(S + Rn, Wr), 32, 0, st>>>((const G2XYZZ*)pc, (const G2XYZZ*)pr, p.half, s, K, (G2XYZZ*)V);
            }
            if (!pair_t) msm_rc_sum<G><<<dim3(S + Rn, Wr), 32, 0, st>>>(pc, pr, p.half, s, K, V);
        }
        if constexpr (sizeof(Pt) > 128) {
            if (pair_t) {
                msm_rc_bits_pair<G><<<dim3(nbits * nslice, 1, Wr), 32, 0, st>>>((const G2XYZZ*)V, p.half, s, nbits, nslice, (G2XYZZ*)Us);
                msm_rc_shift_pair<G><<<dim3(nbits, 1, Wr), 32, 0, st>>>((const G2XYZZ*)Us, nbits, nslice, (G2XYZZ*)U);
                msm_rc_final_pair<G><<<Wr, 32, 0, st>>>((const G2XYZZ*)U, nbits, (G2XYZZ*)(p.table ? (Pt*)d_out : win));
            }
        }
        if (!pair_t) {
            msm_rc_bits<G><<<dim3(nbits * nslice, 1, Wr), 32, 0, st>>>(V, p.half, s, nbits, nslice, Us);
            msm_rc_shift<G><<<dim3(nbits, 1, Wr), 32, 0, st>>>(Us, nbits, nslice, U);
            msm_rc_final<G><<<Wr, 32, 0, st>>>(U, nbits, p.table ? (Pt*)d_out : win);
        }
        ctx->launches += 5;
        if (!p.table) {
            msm_horner<G><<<1, 32, 0, st>>>(win, p.W, p.c, (Pt*)d_out);
            ctx->launches++;
        }
    } else if (p.table) {
        // one bucket set: segment sums, then a tree of block reductions straight into the result
        msm_bucket_reduce<G><<<(p.nseg + 127) / 128, 128, 0, st>>>(B, p.half, p.S, p.nseg, 1, seg);
        ctx->launches++;
        Pt* cur = seg;
        uint32_t cnt = p.nseg;
        Pt* spare = seg + p.nseg;
        while (true) {
            const uint32_t per = cnt > 128 * 8 ? 128 * 8 : cnt;        // inputs per block
            const uint32_t blocks = (cnt + per - 1) / per;
            Pt* dst = blocks == 1 ? (Pt*)d_out : spare;
            msm_window_reduce<G><<<blocks, 128, smem, st>>>(cur, per, cnt, dst);
            ctx->launches++;
            if (blocks == 1) break;
            cur = dst; cnt = blocks; spare = dst + blocks;
        }
    } else {
        msm_bucket_reduce<G><<<(p.nseg * p.W + 127) / 128, 128, 0, st>>>(B, p.half, p.S, p.nseg, p.W, seg);
        msm_window_reduce<G><<<p.W, 128, smem, st>>>(seg, p.nseg, p.nseg * (uint32_t)p.W, win);
        msm_horner<G><<<1, 32, 0, st>>>(win, p.W, p.c, (Pt*)d_out);
        ctx->launches += 3;
    }
}

}  // namespace b200
