#pragma once
// Small group helpers around the MSM: window-table precompute, fixed-base generation, point additions, final affine.
#include "msm_common.cuh"

namespace b200 {

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

// ------------------------------------------------------------------------------------------------- small group helpers
// out = sum(xyzz terms) + sum(affine terms), normalised to affine; layout of out: Aff followed by a u32 infinity flag.
template <typename G>
__global__ void g_finish_kernel(const typename G::Pt* xt, int nx, const typename G::Aff* at, int na, typename G::Aff* out, bool to_std) {
    typedef typename G::Pt Pt;
    if (threadIdx.x || blockIdx.x) return;
    Pt r = Pt::inf();
    for (int i = 0; i < nx; ++i) r.add(xt[i]);
    for (int i = 0; i < na; ++i) r.add_mixed(at[i]);
    typename G::Aff a = r.to_affine_low_latency();
    uint32_t inf = r.is_inf() ? 1u : 0u;
    if (to_std) {
        // limb-wise from_mont of every Fq coordinate
        Fq* f = reinterpret_cast<Fq*>(&a);
        for (int i = 0; i < (int)(sizeof(a) / sizeof(Fq)); ++i) f[i] = f[i].from_mont();
    }
    *out = a;
    *reinterpret_cast<uint32_t*>(out + 1) = inf;
}
// out = sum of n affine points laid out `stride` bytes apart (the gathered per-GPU partial sums of a sharded proof) + an
// optional extra affine term (alpha_g1 / beta_g2); same output layout as g_finish_kernel
template <typename G>
__global__ void g_sum_strided_kernel(const unsigned char* base, size_t stride, int n, const typename G::Aff* extra, typename G::Aff* out, bool to_std) {
    typedef typename G::Pt Pt;
    if (threadIdx.x || blockIdx.x) return;
    Pt r = Pt::inf();
    for (int i = 0; i < n; ++i) r.add_mixed(*reinterpret_cast<const typename G::Aff*>(base + (size_t)i * stride));
    if (extra) r.add_mixed(*extra);
    typename G::Aff a = r.to_affine_low_latency();
    const uint32_t inf = r.is_inf() ? 1u : 0u;
    if (to_std) {
        Fq* f = reinterpret_cast<Fq*>(&a);
        for (int i = 0; i < (int)(sizeof(a) / sizeof(Fq)); ++i) f[i] = f[i].from_mont();
    }
    *out = a;
    *reinterpret_cast<uint32_t*>(out + 1) = inf;
}
template <typename G>
static void g_sum_strided_t(b200_ctx* ctx, const void* base, size_t stride, int n, const void* extra, void* out, bool to_std) {
    g_sum_strided_kernel<G><<<1, 32, 0, ctx->cur()>>>((const unsigned char*)base, stride, n, (const typename G::Aff*)extra, (typename G::Aff*)out, to_std);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

// The whole combine of a sharded proof in ONE launch: block 0 adds the world A partials and alpha_g1, block 1 the B partials
// (G2) and beta_g2, block 2 the C partials; records are 256 bytes apart (A 64 | B 128 | C 64); results in standard form at
// fin + 0 / + 256 / + 512, each followed by its u32 infinity flag.  (Three launches cost three serial inversions: 0.34 ms.)
template <typename G>   // instantiated with G2T only, so that one translation unit emits it
__global__ void g_combine3_kernel(const unsigned char* parts, int world, const Affine<Fq>* alpha, const Affine<Fq2>* beta, unsigned char* fin) {
    if (threadIdx.x) return;
    if (blockIdx.x == 1) {
        XYZZ<Fq2> r = XYZZ<Fq2>::inf();
        for (int i = 0; i < world; ++i) r.add_mixed(*reinterpret_cast<const Affine<Fq2>*>(parts + (size_t)i * 256 + 64));
        r.add_mixed(*beta);
        Affine<Fq2> a = r.to_affine_low_latency();
        a.x.a = a.x.a.from_mont(); a.x.b = a.x.b.from_mont(); a.y.a = a.y.a.from_mont(); a.y.b = a.y.b.from_mont();
        *reinterpret_cast<Affine<Fq2>*>(fin + 256) = a;
        *reinterpret_cast<uint32_t*>(fin + 256 + sizeof(a)) = r.is_inf() ? 1u : 0u;
        return;
    }
    const bool isA = blockIdx.x == 0;
    XYZZ<Fq> r = XYZZ<Fq>::inf();
    for (int i = 0; i < world; ++i) r.add_mixed(*reinterpret_cast<const Affine<Fq>*>(parts + (size_t)i * 256 + (isA ? 0 : 192)));
    if (isA) r.add_mixed(*alpha);
    Affine<Fq> a = r.to_affine_low_latency();
    a.x = a.x.from_mont(); a.y = a.y.from_mont();
    unsigned char* out = fin + (isA ? 0 : 512);
    *reinterpret_cast<Affine<Fq>*>(out) = a;
    *reinterpret_cast<uint32_t*>(out + sizeof(a)) = r.is_inf() ? 1u : 0u;
}

template <typename G>
static void g_finish_t(b200_ctx* ctx, const void* xt, int nx, const void* at, int na, void* out, bool to_std) {
    g_finish_kernel<G><<<1, 32, 0, ctx->cur()>>>((const typename G::Pt*)xt, nx, (const typename G::Aff*)at, na, (typename G::Aff*)out, to_std);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

// out[i] = k_i * generator (affine, (0,0) for infinity); replaces batch_exp (multiexp.tcc:838-905, the windowed fixed-base
// exponentiation the reference's generator uses) for synthetic keys and for the key generator.
// Table: T[w][d] = d * 2^(8 w) * generator, w < 32, d < 256 (affine; 512 KB for G1, 1 MB for G2: L2 resident), built once per
// context; a scalar is then at most 32 mixed additions and one inversion.
template <typename G>
__global__ void __launch_bounds__(128) g_fixed_table_kernel(typename G::Aff* __restrict__ table) {
    typedef typename G::Pt Pt;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= 32 * 256) return;
    const int w = idx >> 8, d = idx & 255;
    const typename G::Aff g = Gen<typename G::F>::get();
    Pt r = Pt::inf();
    for (int b = 7; b >= 0; --b) {
        r.dbl();
        if ((d >> b) & 1) r.add_mixed(g);
    }
    for (int b = 0; b < 8 * w; ++b) r.dbl();
    table[idx] = r.to_affine();
}
template <typename G>
__global__ void __launch_bounds__(128) g_fixed_base_kernel(const Fr* __restrict__ scalars, size_t n, const typename G::Aff* __restrict__ table,
                                                           typename G::Aff* __restrict__ out) {
    typedef typename G::Pt Pt;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Fr k = scalars[i].from_mont();
    Pt r = Pt::inf();
    for (int w = 0; w < 32; ++w) {
        const uint32_t d = (k.v[w >> 2] >> ((w & 3) * 8)) & 255u;
        if (d) r.add_mixed(table[w * 256 + d]);
    }
    out[i] = r.to_affine();
}
template <typename G>
static void g_fixed_base_t(b200_ctx* ctx, const void* d_scalars, size_t n, void* d_out) {
    if (n == 0) return;
    typedef typename G::Aff Aff;
    const int gi = sizeof(Aff) == 64 ? 0 : 1;
    const int lane = ctx->lane;
    ctx->lane = 0;                     // one table per context, whichever lane asks first
    Aff* table = (Aff*)ctx->scratch(gi ? "fixed_table_g2" : "fixed_table_g1", 32 * 256 * sizeof(Aff));
    ctx->lane = lane;
    if (!ctx->fixed_table_ready[gi]) {
        g_fixed_table_kernel<G><<<64, 128, 0, ctx->stream>>>(table);
        ctx->launches++;
        B200_CUDA_CHECK(cudaGetLastError());
        ctx->fixed_table_ready[gi] = true;
    }
    unsigned grid = (unsigned)((n + 127) / 128);
    g_fixed_base_kernel<G><<<grid, 128, 0, ctx->stream>>>((const Fr*)d_scalars, n, table, (Aff*)d_out);
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


#define B200_INSTANTIATE_GROUP_MISC(G, SUFFIX)                                                                                         \
    void g_precompute_##SUFFIX(b200_ctx* ctx, const void* b, size_t n, int c, void* t) { g_precompute_t<G>(ctx, b, n, c, t); }         \
    void g_finish_##SUFFIX(b200_ctx* ctx, const void* xt, int nx, const void* at, int na, void* out, bool to_std) {                   \
        g_finish_t<G>(ctx, xt, nx, at, na, out, to_std);                                                                               \
    }                                                                                                                                  \
    void g_sum_strided_##SUFFIX(b200_ctx* ctx, const void* base, size_t stride, int n, const void* extra, void* out, bool to_std) {    \
        g_sum_strided_t<G>(ctx, base, stride, n, extra, out, to_std);                                                                  \
    }                                                                                                                                  \
    void g_fixed_base_##SUFFIX(b200_ctx* ctx, const void* s, size_t n, void* out) { g_fixed_base_t<G>(ctx, s, n, out); }               \
    void g_add_affine_##SUFFIX(b200_ctx* ctx, const void* a, const void* b, void* out, size_t n) { g_add_affine_t<G>(ctx, a, b, out, n); } \
    void g_scalar_mul_##SUFFIX(b200_ctx* ctx, const void* base, const void* k, void* out) { g_scalar_mul_t<G>(ctx, base, k, out); }

}  // namespace b200
