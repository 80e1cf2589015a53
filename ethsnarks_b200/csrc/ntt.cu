// Fr-side kernels of the Groth16 prove path for sm_100a: radix-2 NTT passes staged in shared memory, the fused
// H-polynomial chain, the witness-times-R1CS sparse products, and batched field ops for the parity tests.
//
// Replaces (FQ = /root/reference/depends/libfqfft/libfqfft, LS = /root/reference/depends/libsnark/libsnark):
//   recursive_domain::{FFT,iFFT,cosetFFT,icosetFFT}   FQ/evaluation_domain/domains/recursive_domain.tcc:102-142
//   _recursive_FFT / butterfly_2 / butterfly_4          FQ/evaluation_domain/domains/recursive_domain_aux.tcc:69-326
//   _multiply_by_coset_and_constant, divide_by_Z…       recursive_domain_aux.tcc:328-346, recursive_domain.tcc:171-183
//   r1cs_to_qap_witness_map                              LS/reductions/r1cs_to_qap/r1cs_to_qap.tcc:217-302
//   linear_combination_light::evaluate                   LS/relations/variable.hpp:296-304
// The transform computed is the same DFT over the same 2^k-th root of unity (get_root_of_unity,
// FF/algebra/fields/field_utils.tcc:35-51), so every output element is bit-identical to the reference's; the
// schedule is different: decimation-in-frequency passes (natural in, bit-reversed out) for the inverse transforms,
// decimation-in-time passes (bit-reversed in, natural out) for the coset-forward ones, so no transform in the H
// chain needs a bit-reversal pass, and the coset / m^-1 scalings ride on the last pass of the inverse transform.
#include "fp.cuh"
#include "internal.hpp"
#include <algorithm>
#include <vector>

namespace b200 {

// 2^28-th primitive root of unity of Fr, Montgomery limbs (mcl_bn128_init.cpp:61)
__device__ __constant__ uint32_t FR_ROOT28[8] = {0xb639feb8u, 0x9632c7c5u, 0x0d0ff299u, 0x985ce340u, 0x01b0ecd8u, 0xb2dd8800u, 0x6d98ce29u, 0x1d69070du};

static B200_DEV Fr fr_from_u32(uint32_t x) {
    Fr r = Fr::zero(); r.v[0] = x;
    return r.to_mont();
}
static B200_DEV uint32_t brev(uint32_t x, uint32_t bits) { return bits ? (__brev(x) >> (32 - bits)) : 0; }

// ---------------------------------------------------------------------------------------- domain tables
// consts: [0] omega_m, [1] m^-1, [2] g^-1, [3] (g^m - 1)^-1, [4] g        (g = multiplicative generator 7)
__global__ void domain_consts_kernel(uint32_t logm, Fr* consts) {
    if (threadIdx.x || blockIdx.x) return;
    Fr w;
#pragma unroll
    for (int i = 0; i < 8; ++i) w.v[i] = FR_ROOT28[i];
    for (uint32_t i = logm; i < 28; ++i) w = w.sqr();
    Fr g = fr_from_u32(7);
    Fr gm = g;
    for (uint32_t i = 0; i < logm; ++i) gm = gm.sqr();
    consts[0] = w;
    consts[1] = fr_from_u32(1u << logm).inv();
    consts[2] = g.inv();
    consts[3] = (gm - Fr::one()).inv();
    consts[4] = g;
}
__global__ void domain_tables_kernel(uint32_t logm, const Fr* __restrict__ consts, Fr* tw_fwd, Fr* tw_inv, Fr* cs_fwd, Fr* cs_inv) {
    const uint32_t m = 1u << logm;
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const Fr w = consts[0], minv = consts[1], ginv = consts[2], g = consts[4];
    if (i < m / 2 || m == 1) {
        tw_fwd[i] = w.pow_u32(i);
        tw_inv[i] = i ? w.pow_u32(m - i) : Fr::one();
    }
    const uint32_t r = brev(i, logm);
    cs_fwd[i] = minv * g.pow_u32(r);
    cs_inv[i] = minv * ginv.pow_u32(r);
}

void domain_create(b200_ctx* ctx, uint32_t logm, Domain* d) {
    if (logm > 28) throw std::runtime_error("domain: log2(m) must be <= 28 (2-adicity of Fr)");
    const size_t m = (size_t)1 << logm;
    d->logm = logm;
    B200_CUDA_CHECK(cudaMalloc(&d->tw_fwd, std::max<size_t>(1, m / 2) * 32));
    B200_CUDA_CHECK(cudaMalloc(&d->tw_inv, std::max<size_t>(1, m / 2) * 32));
    B200_CUDA_CHECK(cudaMalloc(&d->cs_fwd, m * 32));
    B200_CUDA_CHECK(cudaMalloc(&d->cs_inv, m * 32));
    B200_CUDA_CHECK(cudaMalloc(&d->zinv, 5 * 32));
    domain_consts_kernel<<<1, 32, 0, ctx->stream>>>(logm, (Fr*)d->zinv);
    domain_tables_kernel<<<(unsigned)((m + 127) / 128), 128, 0, ctx->stream>>>(logm, (const Fr*)d->zinv, (Fr*)d->tw_fwd, (Fr*)d->tw_inv,
                                                                                 (Fr*)d->cs_fwd, (Fr*)d->cs_inv);
    ctx->launches += 2;
    B200_CUDA_CHECK(cudaGetLastError());
    B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
}
void domain_free(Domain* d) {
    cudaFree(d->tw_fwd); cudaFree(d->tw_inv); cudaFree(d->cs_fwd); cudaFree(d->cs_inv); cudaFree(d->zinv);
    *d = Domain();
}

// ---------------------------------------------------------------------------------------- NTT pass
// One pass = k consecutive radix-2 stages on tiles of 2^k elements, T tiles per block, staged in shared memory as
// eight limb planes (bank-conflict-free 32-bit accesses).  The k index bits the pass works on sit at bit position
// `lo`; blockIdx.x enumerates the remaining (m >> k) index combinations in groups of T, blockIdx.y selects one of up
// to three arrays transformed by the same launch.
//   DIF (decimation in frequency): global stage s pairs (i, i + m>>(s+1)): (a, b) -> (a + b, (a - b) w^(j << s))
//   DIT (decimation in time):      global stage s pairs (i, i + 2^s):      (a, b) -> (a + w b, a - w b), w = tw[j << (logm-1-s)]
struct NttArrays { Fr* p[3]; };

template <bool DIF>
__global__ void __launch_bounds__(512) ntt_pass(NttArrays arrs, uint32_t logm, uint32_t lo, uint32_t k, uint32_t tb,
                                                const Fr* __restrict__ tw, const Fr* __restrict__ post_scale) {
    extern __shared__ uint32_t sm[];   // [8][2^k * T]
    Fr* data = arrs.p[blockIdx.y];
    const uint32_t T = 1u << tb, tile = 1u << k, slots = tile << tb;
    const uint32_t tid = threadIdx.x;   // blockDim.x == slots / 2
    const uint32_t q0 = blockIdx.x << tb;
    const uint32_t lomask = (1u << lo) - 1;

    // ---- load: 2 elements per thread
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const uint32_t e = tid + r * (slots >> 1);
        uint32_t mid, tq;
        if (lo == 0) { mid = e & (tile - 1); tq = e >> k; } else { tq = e & (T - 1); mid = e >> tb; }
        const uint32_t q = q0 + tq;
        const size_t addr = ((size_t)(q >> lo) << (lo + k)) | ((size_t)mid << lo) | (q & lomask);
        const Fr x = data[addr];
        const uint32_t s = (mid << tb) + tq;
#pragma unroll
        for (int l = 0; l < 8; ++l) sm[l * slots + s] = x.v[l];
    }
    __syncthreads();

    // ---- k butterfly stages
    const uint32_t tq = tid & (T - 1), pi = tid >> tb;
    const uint32_t q = q0 + tq, low = q & lomask;
    for (uint32_t l = 0; l < k; ++l) {
        uint32_t hb, mid_a, e;
        if (DIF) {
            hb = k - 1 - l;                                  // log2 of the local half
            mid_a = ((pi >> hb) << (hb + 1)) | (pi & ((1u << hb) - 1));
            const uint32_t s = (logm - lo - k) + l;          // global stage
            e = (((mid_a & ((1u << hb) - 1)) << lo) | low) << s;
        } else {
            hb = l;
            mid_a = ((pi >> hb) << (hb + 1)) | (pi & ((1u << hb) - 1));
            const uint32_t s = lo + l;
            e = (((mid_a & ((1u << hb) - 1)) << lo) | low) << (logm - 1 - s);
        }
        const uint32_t sa = (mid_a << tb) + tq, sb = sa + ((1u << hb) << tb);
        Fr a, b;
#pragma unroll
        for (int i = 0; i < 8; ++i) { a.v[i] = sm[i * slots + sa]; b.v[i] = sm[i * slots + sb]; }
        const Fr w = tw[e];
        Fr x, y;
        if (DIF) { x = a + b; y = (a - b) * w; } else { const Fr t = b * w; x = a + t; y = a - t; }
#pragma unroll
        for (int i = 0; i < 8; ++i) { sm[i * slots + sa] = x.v[i]; sm[i * slots + sb] = y.v[i]; }
        __syncthreads();
    }

    // ---- store
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        const uint32_t e = tid + r * (slots >> 1);
        uint32_t mid, tq2;
        if (lo == 0) { mid = e & (tile - 1); tq2 = e >> k; } else { tq2 = e & (T - 1); mid = e >> tb; }
        const uint32_t qq = q0 + tq2;
        const size_t addr = ((size_t)(qq >> lo) << (lo + k)) | ((size_t)mid << lo) | (qq & lomask);
        const uint32_t s = (mid << tb) + tq2;
        Fr x;
#pragma unroll
        for (int l = 0; l < 8; ++l) x.v[l] = sm[l * slots + s];
        if (post_scale) x = x * post_scale[addr];
        data[addr] = x;
    }
}

// The same pass with TWO stages per shared-memory round trip: a thread holds the four elements of a radix-4 group in
// registers (same four products as two radix-2 stages, three twiddles), so the pass has half the barriers and half the
// shared-memory traffic; ncu on ntt_pass showed mio_throttle + barrier stalls holding the IMAD pipe at 55-62 %
// (profiles/r1_ncu_ntt_pass.txt).  blockDim.x == slots / 4; an odd k ends with one radix-2 stage (two butterflies per thread).
#ifndef B200_NTT4_MINB
#define B200_NTT4_MINB 3
#endif
template <bool DIF>
__global__ void __launch_bounds__(256, B200_NTT4_MINB) ntt_pass4(NttArrays arrs, uint32_t logm, uint32_t lo, uint32_t k, uint32_t tb,
                                                                 const Fr* __restrict__ tw, const Fr* __restrict__ post_scale) {
    extern __shared__ uint32_t sm[];   // [8][2^k * T]
    Fr* data = arrs.p[blockIdx.y];
    const uint32_t T = 1u << tb, tile = 1u << k, slots = tile << tb;
    const uint32_t tid = threadIdx.x;
    const uint32_t q0 = blockIdx.x << tb;
    const uint32_t lomask = (1u << lo) - 1;
    auto ld = [&](uint32_t s) { Fr x; _Pragma("unroll") for (int i = 0; i < 8; ++i) x.v[i] = sm[i * slots + s]; return x; };
    auto st = [&](uint32_t s, const Fr& x) { _Pragma("unroll") for (int i = 0; i < 8; ++i) sm[i * slots + s] = x.v[i]; };

#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t e = tid + r * (slots >> 2);
        uint32_t mid, tq;
        if (lo == 0) { mid = e & (tile - 1); tq = e >> k; } else { tq = e & (T - 1); mid = e >> tb; }
        const uint32_t q = q0 + tq;
        const size_t addr = ((size_t)(q >> lo) << (lo + k)) | ((size_t)mid << lo) | (q & lomask);
        st((mid << tb) + tq, data[addr]);
    }
    __syncthreads();

    const uint32_t tq = tid & (T - 1), g = tid >> tb;      // g in [0, 2^(k-2))
    const uint32_t low = (q0 + tq) & lomask;
    uint32_t l = 0;
    for (; l + 1 < k; l += 2) {
        if (DIF) {
            const uint32_t hb = k - 1 - l, hb2 = hb - 1;
            const uint32_t r = g & ((1u << hb2) - 1);
            const uint32_t mid0 = ((g >> hb2) << (hb + 1)) | r;
            const uint32_t s = (logm - lo - k) + l;
            const uint32_t s0 = (mid0 << tb) + tq, d1 = (1u << hb2) << tb, d2 = (1u << hb) << tb;
            Fr e0 = ld(s0), e1 = ld(s0 + d1), e2 = ld(s0 + d2), e3 = ld(s0 + d2 + d1);
            {   // stage l: (e0, e2), (e1, e3)
                const Fr wa = tw[((r << lo) | low) << s];
                const Fr d = (e0 - e2) * wa;
                e0 = e0 + e2; e2 = d;
                const Fr wb = tw[(((r + (1u << hb2)) << lo) | low) << s];
                const Fr f = (e1 - e3) * wb;
                e1 = e1 + e3; e3 = f;
            }
            {   // stage l + 1: (e0, e1), (e2, e3), one twiddle
                const Fr wc = tw[((r << lo) | low) << (s + 1)];
                const Fr d = (e0 - e1) * wc;
                e0 = e0 + e1; e1 = d;
                const Fr f = (e2 - e3) * wc;
                e2 = e2 + e3; e3 = f;
            }
            st(s0, e0); st(s0 + d1, e1); st(s0 + d2, e2); st(s0 + d2 + d1, e3);
        } else {
            const uint32_t r = g & ((1u << l) - 1);
            const uint32_t mid0 = ((g >> l) << (l + 2)) | r;
            const uint32_t s = lo + l;
            const uint32_t s0 = (mid0 << tb) + tq, d1 = (1u << l) << tb, d2 = (2u << l) << tb;
            Fr e0 = ld(s0), e1 = ld(s0 + d1), e2 = ld(s0 + d2), e3 = ld(s0 + d2 + d1);
            {   // stage l: (e0, e1), (e2, e3), one twiddle
                const Fr wa = tw[((r << lo) | low) << (logm - 1 - s)];
                Fr t = e1 * wa;
                e1 = e0 - t; e0 = e0 + t;
                t = e3 * wa;
                e3 = e2 - t; e2 = e2 + t;
            }
            {   // stage l + 1: (e0, e2), (e1, e3)
                const Fr wb = tw[((r << lo) | low) << (logm - 2 - s)];
                Fr t = e2 * wb;
                e2 = e0 - t; e0 = e0 + t;
                const Fr wc = tw[(((r + (1u << l)) << lo) | low) << (logm - 2 - s)];
                t = e3 * wc;
                e3 = e1 - t; e1 = e1 + t;
            }
            st(s0, e0); st(s0 + d1, e1); st(s0 + d2, e2); st(s0 + d2 + d1, e3);
        }
        __syncthreads();
    }
    if (l < k) {      // odd k: one radix-2 stage, two butterflies per thread
#pragma unroll 1
        for (uint32_t h = 0; h < 2; ++h) {
            const uint32_t pi = g + (h << (k - 2));
            uint32_t hb, e;
            if (DIF) { hb = k - 1 - l; } else { hb = l; }
            const uint32_t mid_a = ((pi >> hb) << (hb + 1)) | (pi & ((1u << hb) - 1));
            if (DIF) e = (((mid_a & ((1u << hb) - 1)) << lo) | low) << ((logm - lo - k) + l);
            else e = (((mid_a & ((1u << hb) - 1)) << lo) | low) << (logm - 1 - (lo + l));
            const uint32_t sa = (mid_a << tb) + tq, sb = sa + ((1u << hb) << tb);
            const Fr a = ld(sa), b = ld(sb), w = tw[e];
            if (DIF) { st(sa, a + b); st(sb, (a - b) * w); } else { const Fr t = b * w; st(sa, a + t); st(sb, a - t); }
        }
        __syncthreads();
    }

#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const uint32_t e = tid + r * (slots >> 2);
        uint32_t mid, tq2;
        if (lo == 0) { mid = e & (tile - 1); tq2 = e >> k; } else { tq2 = e & (T - 1); mid = e >> tb; }
        const uint32_t qq = q0 + tq2;
        const size_t addr = ((size_t)(qq >> lo) << (lo + k)) | ((size_t)mid << lo) | (qq & lomask);
        Fr x = ld((mid << tb) + tq2);
        if (post_scale) x = x * post_scale[addr];
        data[addr] = x;
    }
}

#ifndef B200_NTT_KMAX
#define B200_NTT_KMAX 8
#endif
static const uint32_t NTT_KMAX = B200_NTT_KMAX, NTT_SLOTS_LOG = 10;   // up to 2^8-point tiles, 1024 elements (32 KB) per block

static void plan_passes(uint32_t logm, std::vector<uint32_t>& ks) {
    ks.clear();
    if (logm == 0) return;
    uint32_t np = (logm + NTT_KMAX - 1) / NTT_KMAX;
    uint32_t base = logm / np, extra = logm % np;
    for (uint32_t i = 0; i < np; ++i) ks.push_back(base + (i < extra ? 1 : 0));
}

// runs a full DIF (natural -> bit-reversed) or DIT (bit-reversed -> natural) transform on n_arr arrays in place
template <bool DIF>
static void run_transform(b200_ctx* ctx, uint32_t logm, Fr* a0, Fr* a1, Fr* a2, int n_arr, const Fr* tw, const Fr* post_scale_last) {
    std::vector<uint32_t> ks;
    plan_passes(logm, ks);
    NttArrays arrs;
    arrs.p[0] = a0; arrs.p[1] = a1; arrs.p[2] = a2;
    uint32_t done = 0;
    static bool attr = false;
    if (!attr) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(ntt_pass<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 << NTT_SLOTS_LOG));
        B200_CUDA_CHECK(cudaFuncSetAttribute(ntt_pass<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 << NTT_SLOTS_LOG));
        B200_CUDA_CHECK(cudaFuncSetAttribute(ntt_pass4<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 << NTT_SLOTS_LOG));
        B200_CUDA_CHECK(cudaFuncSetAttribute(ntt_pass4<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32 << NTT_SLOTS_LOG));
        attr = true;
    }
    for (size_t pi = 0; pi < ks.size(); ++pi) {
        const uint32_t k = ks[pi];
        const uint32_t lo = DIF ? (logm - done - k) : done;
        uint32_t tb = std::min<uint32_t>(NTT_SLOTS_LOG - k, logm - k);
        const uint32_t slots = 1u << (k + tb);
        dim3 grid(1u << (logm - k - tb), n_arr);
        const bool last = (pi + 1 == ks.size());
        if (ctx->ntt_radix4 && k >= 2 && slots >= 128)
            ntt_pass4<DIF><<<grid, slots / 4, slots * 32, ctx->stream>>>(arrs, logm, lo, k, tb, tw, last ? post_scale_last : nullptr);
        else
            ntt_pass<DIF><<<grid, slots / 2, slots * 32, ctx->stream>>>(arrs, logm, lo, k, tb, tw, last ? post_scale_last : nullptr);
        ctx->launches++;
        done += k;
    }
    B200_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------- element-wise kernels
__global__ void bitrev_permute_kernel(const uint4* __restrict__ in, uint4* __restrict__ out, uint32_t logm, uint32_t vec_per_elem) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ((size_t)1 << logm)) return;
    size_t j = brev((uint32_t)i, logm);
    for (uint32_t v = 0; v < vec_per_elem; ++v) out[j * vec_per_elem + v] = in[i * vec_per_elem + v];
}
void bitrev_permute(b200_ctx* ctx, const void* d_in, void* d_out, uint32_t logm, size_t elem_bytes) {
    size_t m = (size_t)1 << logm;
    bitrev_permute_kernel<<<(unsigned)((m + 255) / 256), 256, 0, ctx->stream>>>((const uint4*)d_in, (uint4*)d_out, logm, (uint32_t)(elem_bytes / 16));
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}
// data[i] *= c * base^i
__global__ void mul_powers_kernel(Fr* data, size_t n, const Fr* __restrict__ base, const Fr* __restrict__ c) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr f = base->pow_u32((uint32_t)i);
    if (c) f = f * (*c);
    data[i] = data[i] * f;
}
__global__ void mul_const_kernel(Fr* data, size_t n, const Fr* __restrict__ c) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    data[i] = data[i] * (*c);
}
// c = (a * b - c) * zinv      (r1cs_to_qap.tcc:282-293)
__global__ void h_pointwise_kernel(const Fr* __restrict__ a, const Fr* __restrict__ b, const Fr* c, Fr* out, size_t n, const Fr* __restrict__ zinv) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = (a[i] * b[i] - c[i]) * (*zinv);      // out may be c itself
}

void ntt_natural(b200_ctx* ctx, const Domain& dom, void* d_data, int mode) {
    const uint32_t logm = dom.logm;
    const size_t m = (size_t)1 << logm;
    Fr* data = (Fr*)d_data;
    const Fr* consts = (const Fr*)dom.zinv;
    const unsigned grid = (unsigned)((m + 127) / 128);
    if (mode == 2) { mul_powers_kernel<<<grid, 128, 0, ctx->stream>>>(data, m, consts + 4, nullptr); ctx->launches++; }
    const bool inverse = (mode == 1 || mode == 3);
    run_transform<true>(ctx, logm, data, nullptr, nullptr, 1, (const Fr*)(inverse ? dom.tw_inv : dom.tw_fwd), nullptr);
    Fr* tmp = (Fr*)ctx->scratch("ntt_tmp", m * 32);
    bitrev_permute(ctx, data, tmp, logm, 32);
    B200_CUDA_CHECK(cudaMemcpyAsync(data, tmp, m * 32, cudaMemcpyDeviceToDevice, ctx->stream));
    if (mode == 1) { mul_const_kernel<<<grid, 128, 0, ctx->stream>>>(data, m, consts + 1); ctx->launches++; }
    if (mode == 3) { mul_powers_kernel<<<grid, 128, 0, ctx->stream>>>(data, m, consts + 2, consts + 1); ctx->launches++; }
    B200_CUDA_CHECK(cudaGetLastError());
}

void h_chain(b200_ctx* ctx, const Domain& dom, void* d_a, void* d_b, void* d_c) {
    const uint32_t logm = dom.logm;
    const size_t m = (size_t)1 << logm;
    Fr *a = (Fr*)d_a, *b = (Fr*)d_b, *c = (Fr*)d_c;
    // coefficients (bit-reversed order), scaled by m^-1 g^i for the coset
    run_transform<true>(ctx, logm, a, b, c, 3, (const Fr*)dom.tw_inv, (const Fr*)dom.cs_fwd);
    // evaluations on the coset g<omega> (natural order)
    run_transform<false>(ctx, logm, a, b, c, 3, (const Fr*)dom.tw_fwd, nullptr);
    h_pointwise_kernel<<<(unsigned)((m + 127) / 128), 128, 0, ctx->stream>>>(a, b, c, c, m, (const Fr*)dom.zinv + 3);
    ctx->launches++;
    // icosetFFT, output left in bit-reversed order
    run_transform<true>(ctx, logm, c, nullptr, nullptr, 1, (const Fr*)dom.tw_inv, (const Fr*)dom.cs_inv);
    B200_CUDA_CHECK(cudaGetLastError());
}

// The same chain in two halves, for a proof spread over several GPUs: h_chain_front takes ONE of the three evaluation vectors
// (or two) to its evaluations on the coset (iFFT then cosetFFT) - the three vectors are independent up to here, so different GPUs can
// take one each; h_chain_back is the pointwise step and the final icosetFFT once all three are present.
void h_chain_front(b200_ctx* ctx, const Domain& dom, void* const* d_x, int count) {
    if (count < 1) return;
    Fr* x[3] = {(Fr*)d_x[0], count > 1 ? (Fr*)d_x[1] : nullptr, count > 2 ? (Fr*)d_x[2] : nullptr};
    run_transform<true>(ctx, dom.logm, x[0], x[1], x[2], count, (const Fr*)dom.tw_inv, (const Fr*)dom.cs_fwd);
    run_transform<false>(ctx, dom.logm, x[0], x[1], x[2], count, (const Fr*)dom.tw_fwd, nullptr);
    B200_CUDA_CHECK(cudaGetLastError());
}
void h_chain_back(b200_ctx* ctx, const Domain& dom, const void* d_a, const void* d_b, const void* d_c, void* d_out) {
    const size_t m = (size_t)1 << dom.logm;
    h_pointwise_kernel<<<(unsigned)((m + 127) / 128), 128, 0, ctx->stream>>>((const Fr*)d_a, (const Fr*)d_b, (const Fr*)d_c, (Fr*)d_out, m, (const Fr*)dom.zinv + 3);
    ctx->launches++;
    run_transform<true>(ctx, dom.logm, (Fr*)d_out, nullptr, nullptr, 1, (const Fr*)dom.tw_inv, (const Fr*)dom.cs_inv);
    B200_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------- batched field ops (tests)
template <typename F>
__global__ void field_op_kernel(int op, const F* __restrict__ a, const F* __restrict__ b, F* out, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    F x = a[i], r;
    switch (op) {
        case 0: r = x * b[i]; break;
        case 1: r = x + b[i]; break;
        case 2: r = x - b[i]; break;
        case 3: r = x.to_mont(); break;
        case 4: r = x.from_mont(); break;
        case 5: r = x.inv(); break;
        case 6: r = x.sqr(); break;
        case 8: r = x.inv_binary(); break;
        case 9: r = F::dot2(x, b[i], b[i], b[i]); break;   // a b + b^2 under one reduction
        case 10: { const F nx = x.neg_lazy(); r = F::dot4(x, b[i], x, b[i], nx, nx, b[i], b[i]); break; }   // (a + b)^2; a = 0 feeds p * p
        case 11: r = diff_of_products(x, x, b[i], b[i]); break;                                             // a^2 - b^2 as the XYZZ additions use it
        default: r = x.neg(); break;
    }
    out[i] = r;
}
void field_op(b200_ctx* ctx, int field, int op, const void* a, const void* b, void* out, size_t n) {
    if (!n) return;
    unsigned grid = (unsigned)((n + 127) / 128);
    if (field == 0) field_op_kernel<Fr><<<grid, 128, 0, ctx->stream>>>(op, (const Fr*)a, (const Fr*)b, (Fr*)out, n);
    else field_op_kernel<Fq><<<grid, 128, 0, ctx->stream>>>(op, (const Fq*)a, (const Fq*)b, (Fq*)out, n);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------- witness x R1CS (CSR)
static const uint64_t SPMV_LONG_ROW = 256, SPMV_SEG_LEN = 4096;

void csr_upload(b200_ctx* ctx, const uint64_t* row_ptr, const uint32_t* col, const uint32_t* cidx, uint64_t n_rows, DeviceCsr* out) {
    (void)ctx;
    out->n_rows = n_rows;
    out->nnz = n_rows ? row_ptr[n_rows] : 0;
    std::vector<uint32_t> seg_row, long_row, long_seg0;
    std::vector<uint64_t> seg_begin, seg_end;
    for (uint64_t r = 0; r < n_rows; ++r) {
        uint64_t len = row_ptr[r + 1] - row_ptr[r];
        if (len <= SPMV_LONG_ROW) continue;
        long_row.push_back((uint32_t)r);
        long_seg0.push_back((uint32_t)seg_row.size());
        for (uint64_t s = row_ptr[r]; s < row_ptr[r + 1]; s += SPMV_SEG_LEN) {
            seg_row.push_back((uint32_t)r);
            seg_begin.push_back(s);
            seg_end.push_back(std::min(s + SPMV_SEG_LEN, row_ptr[r + 1]));
        }
    }
    long_seg0.push_back((uint32_t)seg_row.size());
    out->n_seg = (uint32_t)seg_row.size();
    out->n_long = (uint32_t)long_row.size();
    auto up = [&](const void* src, size_t bytes) -> void* {
        void* d = nullptr;
        B200_CUDA_CHECK(cudaMalloc(&d, std::max<size_t>(bytes, 16)));
        if (bytes) B200_CUDA_CHECK(cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice));
        return d;
    };
    out->row_ptr = (uint64_t*)up(row_ptr, (n_rows + 1) * 8);
    out->col = (uint32_t*)up(col, out->nnz * 4);
    out->cidx = (uint32_t*)up(cidx, out->nnz * 4);
    out->seg_row = (uint32_t*)up(seg_row.data(), seg_row.size() * 4);
    out->seg_begin = (uint64_t*)up(seg_begin.data(), seg_begin.size() * 8);
    out->seg_end = (uint64_t*)up(seg_end.data(), seg_end.size() * 8);
    out->long_row = (uint32_t*)up(long_row.data(), long_row.size() * 4);
    out->long_seg0 = (uint32_t*)up(long_seg0.data(), long_seg0.size() * 4);
}
void csr_free(DeviceCsr* c) {
    cudaFree(c->row_ptr); cudaFree(c->col); cudaFree(c->cidx); cudaFree(c->seg_row); cudaFree(c->seg_begin); cudaFree(c->seg_end);
    cudaFree(c->long_row); cudaFree(c->long_seg0);
    *c = DeviceCsr();
}

__global__ void __launch_bounds__(128) spmv_rows_kernel(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col,
                                                        const uint32_t* __restrict__ cidx, const Fr* __restrict__ coeffs,
                                                        const Fr* __restrict__ w, Fr* __restrict__ out, uint64_t n_rows, uint64_t long_row) {
    uint64_t r = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n_rows) return;
    const uint64_t s = row_ptr[r], e = row_ptr[r + 1];
    if (e - s > long_row) return;
    Fr acc = Fr::zero();
    for (uint64_t k = s; k < e; ++k) acc = acc + coeffs[cidx[k]] * w[col[k]];
    out[r] = acc;
}
__global__ void __launch_bounds__(128) spmv_segments_kernel(const uint64_t* __restrict__ seg_begin, const uint64_t* __restrict__ seg_end,
                                                            const uint32_t* __restrict__ col, const uint32_t* __restrict__ cidx,
                                                            const Fr* __restrict__ coeffs, const Fr* __restrict__ w, Fr* __restrict__ seg_out) {
    __shared__ uint32_t sh[8][128];
    const uint64_t s = seg_begin[blockIdx.x], e = seg_end[blockIdx.x];
    Fr acc = Fr::zero();
    for (uint64_t k = s + threadIdx.x; k < e; k += blockDim.x) acc = acc + coeffs[cidx[k]] * w[col[k]];
#pragma unroll
    for (int i = 0; i < 8; ++i) sh[i][threadIdx.x] = acc.v[i];
    __syncthreads();
    for (int stride = 64; stride > 0; stride >>= 1) {
        if ((int)threadIdx.x < stride) {
            Fr o;
#pragma unroll
            for (int i = 0; i < 8; ++i) o.v[i] = sh[i][threadIdx.x + stride];
            acc = acc + o;
#pragma unroll
            for (int i = 0; i < 8; ++i) sh[i][threadIdx.x] = acc.v[i];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) seg_out[blockIdx.x] = acc;
}
__global__ void spmv_long_rows_kernel(const uint32_t* __restrict__ long_row, const uint32_t* __restrict__ long_seg0,
                                      const Fr* __restrict__ seg_out, Fr* __restrict__ out, uint32_t n_long) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_long) return;
    Fr acc = Fr::zero();
    for (uint32_t s = long_seg0[i]; s < long_seg0[i + 1]; ++s) acc = acc + seg_out[s];
    out[long_row[i]] = acc;
}

void spmv(b200_ctx* ctx, const DeviceCsr& M, const void* d_coeffs, const void* d_w, void* d_out) {
    if (!M.n_rows) return;
    spmv_rows_kernel<<<(unsigned)((M.n_rows + 127) / 128), 128, 0, ctx->stream>>>(M.row_ptr, M.col, M.cidx, (const Fr*)d_coeffs, (const Fr*)d_w,
                                                                                  (Fr*)d_out, M.n_rows, SPMV_LONG_ROW);
    ctx->launches++;
    if (M.n_seg) {
        Fr* seg_out = (Fr*)ctx->scratch("spmv_seg", (size_t)M.n_seg * 32);
        spmv_segments_kernel<<<M.n_seg, 128, 0, ctx->stream>>>(M.seg_begin, M.seg_end, M.col, M.cidx, (const Fr*)d_coeffs, (const Fr*)d_w, seg_out);
        spmv_long_rows_kernel<<<(M.n_long + 127) / 128, 128, 0, ctx->stream>>>(M.long_row, M.long_seg0, seg_out, (Fr*)d_out, M.n_long);
        ctx->launches += 2;
    }
    B200_CUDA_CHECK(cudaGetLastError());
}

// Instanced rows (r1cs_constraint_light_instance + ITranslator, LS/relations/constraint_satisfaction_problems/r1cs/r1cs.hpp:145-211,
// LS/relations/variable.hpp:279-314): n_inst instances of one master gadget share the master's rows; row j * n_master + r is
// master row r evaluated through translation table j (tables[j * tlen + v] = the instance's index of master variable v;
// index 0, the constant ONE, is never translated - variable.hpp:306-314).  HBM holds the master's CSR once and tlen u32 per
// instance instead of the flattened rows (8 B per nonzero + 8 B per row, per instance, per matrix).
__global__ void __launch_bounds__(128) spmv_instances_kernel(const uint64_t* __restrict__ row_ptr, const uint32_t* __restrict__ col,
                                                             const uint32_t* __restrict__ cidx, const Fr* __restrict__ coeffs,
                                                             const Fr* __restrict__ w, const uint32_t* __restrict__ tables, uint32_t tlen,
                                                             uint64_t n_master, uint64_t n_inst, Fr* __restrict__ out) {
    const uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_master * n_inst) return;
    const uint64_t j = id / n_master, r = id - j * n_master;
    const uint32_t* tab = tables + j * tlen;
    Fr acc = Fr::zero();
    for (uint64_t k = row_ptr[r], e = row_ptr[r + 1]; k < e; ++k) {
        const uint32_t c = col[k];
        acc = acc + coeffs[cidx[k]] * w[c ? tab[c] : 0u];
    }
    out[id] = acc;
}
void spmv_instances(b200_ctx* ctx, const DeviceCsr& master, const void* d_coeffs, const void* d_w, const uint32_t* d_tables, uint32_t tlen,
                    uint64_t n_inst, void* d_out) {
    const uint64_t n = master.n_rows * n_inst;
    if (!n) return;
    spmv_instances_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(master.row_ptr, master.col, master.cidx, (const Fr*)d_coeffs,
                                                                                (const Fr*)d_w, d_tables, tlen, master.n_rows, n_inst, (Fr*)d_out);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

// rows nc .. m-1 of the three evaluation vectors: input-consistency rows then zero padding (r1cs_to_qap.tcc:239-254)
__global__ void witness_rows_kernel(const Fr* __restrict__ w, Fr* a, Fr* b, Fr* c, uint64_t nc, uint64_t ni, uint64_t m) {
    uint64_t i = nc + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const Fr z = Fr::zero();
    a[i] = (i - nc <= ni) ? w[i - nc] : z;
    b[i] = z;
    c[i] = z;
}
void witness_rows(b200_ctx* ctx, const void* d_w, void* d_a, void* d_b, void* d_c, uint64_t nc, uint64_t ni, uint64_t m) {
    if (m <= nc) return;
    witness_rows_kernel<<<(unsigned)((m - nc + 255) / 256), 256, 0, ctx->stream>>>((const Fr*)d_w, (Fr*)d_a, (Fr*)d_b, (Fr*)d_c, nc, ni, m);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------------------- integer-pipe peak probe
// Eight independent IMAD.WIDE.U32 accumulation chains per thread; reports the device's sustained rate of the
// instruction the field multiplication is made of (the roofline denominator bench.py uses).
__global__ void __launch_bounds__(256) imad_peak_kernel(uint64_t* out, uint32_t iters, uint32_t a0, uint32_t b0) {
    uint64_t acc[8];
    uint32_t a = a0 + threadIdx.x;
    const uint32_t b = b0 + blockIdx.x;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = (uint64_t)(threadIdx.x + j) << 17;
    for (uint32_t i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            a ^= i + r;   // keeps the products loop-variant (ALU pipe, not counted)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[j] = (uint64_t)a * (uint32_t)(b + j) + acc[j];
        }
    }
    uint64_t s = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) s ^= acc[j];
    if (s == 0x1234567ull) out[0] = s;   // practically never: keeps the chains live
}
// Same count of multiply-adds issued exactly as the field multiplication issues them: carry chains of four
// mad.lo.cc / madc.hi.cc pairs (IMAD.WIDE.U32.X) on four independent 8-limb accumulators per thread.
__global__ void __launch_bounds__(256) imad_chain_peak_kernel(uint64_t* out, uint32_t iters, uint32_t a0, uint32_t b0) {
    uint32_t acc[4][8];
    const uint32_t a = a0 + threadIdx.x, b = b0 + blockIdx.x;
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[s][j] = threadIdx.x * 977u + s * 31u + j;
    for (uint32_t i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 2; ++r) {
#pragma unroll
            for (int s = 0; s < 4; ++s) chain_mad(acc[s], a + s, a ^ 0x5bd1e995u, a + 0x9e3779b9u, a | 0x80000001u, b + r);
        }
    }
    uint32_t x = 0;
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
        for (int j = 0; j < 8; ++j) x ^= acc[s][j];
    if (x == 0x1234567u) out[0] = x;
}
double imad_peak(b200_ctx* ctx, int reps) {
    uint64_t* out = (uint64_t*)ctx->scratch("imad_peak", 64);
    const uint32_t iters = 8192;
    const unsigned grid = (unsigned)ctx->sm_count * 8;
    cudaEvent_t e0, e1;
    B200_CUDA_CHECK(cudaEventCreate(&e0)); B200_CUDA_CHECK(cudaEventCreate(&e1));
    // warm-up long enough for the clocks to settle
    for (int r = 0; r < 3; ++r) imad_peak_kernel<<<grid, 256, 0, ctx->stream>>>(out, iters, 3, 5);
    double best = 0;
    for (int r = 0; r < reps; ++r) {
        for (int variant = 0; variant < 2; ++variant) {
            B200_CUDA_CHECK(cudaEventRecord(e0, ctx->stream));
            if (variant == 0) imad_peak_kernel<<<grid, 256, 0, ctx->stream>>>(out, iters, 3 + r, 5);
            else imad_chain_peak_kernel<<<grid, 256, 0, ctx->stream>>>(out, iters, 3 + r, 5);
            B200_CUDA_CHECK(cudaEventRecord(e1, ctx->stream));
            B200_CUDA_CHECK(cudaEventSynchronize(e1));
            float ms = 0;
            B200_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
            const double macs = (double)grid * 256 * iters * 32;   // both variants: 32 wide multiply-adds per iteration
            best = std::max(best, macs / (ms * 1e-3));
        }
    }
    ctx->launches += 2 * reps + 3;
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    B200_CUDA_CHECK(cudaGetLastError());
    return best;
}

}  // namespace b200
