// Scalar side of the Groth16 key generator on the GPU.
//
// Replaces, for the scalars only, r1cs_gg_ppzksnark_zok_generator (src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:277-449)
// and r1cs_to_qap_instance_map_with_evaluation (LS/reductions/r1cs_to_qap/r1cs_to_qap.tcc:117-204):
//   u_i   = L_i(t), the Lagrange basis of the size-m domain at t  (basic_radix2_domain_aux.tcc:182-236)
//   At_j  = sum_i A[i][j] u_i  (+ u[nc + j] for the input-consistency rows j <= ni), Bt_j, Ct_j likewise
//   H_i   = t^i Z(t) / delta (i < m - 1),   L_j = (beta At_j + alpha Bt_j + Ct_j) / delta (j > ni),   IC_j = (...) / gamma (j <= ni)
// The group side (scalar -> point) is the fixed-base kernel of msm_misc.cuh; prove.cu assembles the key.
#include "fp.cuh"
#include "internal.hpp"
#include <stdexcept>

namespace b200 {

// consts: [0] Z(t) = t^m - 1, [1] Z(t)/m, [2] delta^-1, [3] gamma^-1, [4] Z(t)/delta
__global__ void gen_consts_kernel(const Fr* __restrict__ toxic /* t, alpha, beta, gamma, delta */, uint32_t logm, Fr* __restrict__ consts) {
    if (threadIdx.x || blockIdx.x) return;
    Fr tm = toxic[0];
    for (uint32_t i = 0; i < logm; ++i) tm = tm.sqr();
    const Fr Z = tm - Fr::one();
    Fr m = Fr::zero();
    m.v[logm >> 5] = 1u << (logm & 31);      // 2^logm as a plain integer (logm <= 28)
    m = m.to_mont();
    consts[0] = Z;
    consts[1] = Z * m.inv();
    consts[2] = toxic[4].inv();
    consts[3] = toxic[3].inv();
    consts[4] = Z * consts[2];
}

// u_i = (Z/m) w^i / (t - w^i)      (t outside the domain: checked by the caller through consts[0] != 0)
__global__ void __launch_bounds__(128) gen_lagrange_kernel(const Fr* __restrict__ toxic, const Fr* __restrict__ consts,
                                                           const Fr* __restrict__ tw_fwd, uint64_t m, Fr* __restrict__ u) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint64_t half = m >> 1;
    Fr w = m == 1 ? Fr::one() : tw_fwd[i < half ? i : i - half];
    if (m > 1 && i >= half) w = w.neg();
    u[i] = consts[1] * w * (toxic[0] - w).inv();
}

__global__ void gen_add_inputs_kernel(Fr* __restrict__ At, const Fr* __restrict__ u, uint64_t nc, uint64_t ni) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i > ni) return;
    At[i] = At[i] + u[nc + i];
}

// out[j] = (beta At_j + alpha Bt_j + Ct_j) * (j <= ni ? gamma^-1 : delta^-1), j <= nv  (IC scalars first, then the L query's)
__global__ void __launch_bounds__(128) gen_abc_kernel(const Fr* __restrict__ toxic, const Fr* __restrict__ consts, const Fr* __restrict__ At,
                                                      const Fr* __restrict__ Bt, const Fr* __restrict__ Ct, uint64_t ni, uint64_t nv,
                                                      Fr* __restrict__ out) {
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j > nv) return;
    const Fr s = toxic[2] * At[j] + toxic[1] * Bt[j] + Ct[j];
    out[j] = s * (j <= ni ? consts[3] : consts[2]);
}

// out[i] = t^i Z(t) / delta
__global__ void __launch_bounds__(128) gen_h_kernel(const Fr* __restrict__ toxic, const Fr* __restrict__ consts, uint64_t n, Fr* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr r = Fr::one();
    const Fr t = toxic[0];
    for (int b = 31; b >= 0; --b) {           // n <= 2^28
        r = r.sqr();
        if ((i >> b) & 1) r = r * t;
    }
    out[i] = r * consts[4];
}

void gen_consts(b200_ctx* ctx, const void* d_toxic, uint32_t logm, void* d_consts) {
    gen_consts_kernel<<<1, 32, 0, ctx->stream>>>((const Fr*)d_toxic, logm, (Fr*)d_consts);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}
void gen_lagrange(b200_ctx* ctx, const void* d_toxic, const void* d_consts, const Domain& dom, void* d_u) {
    const uint64_t m = 1ull << dom.logm;
    gen_lagrange_kernel<<<(unsigned)((m + 127) / 128), 128, 0, ctx->stream>>>((const Fr*)d_toxic, (const Fr*)d_consts, (const Fr*)dom.tw_fwd, m, (Fr*)d_u);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}
void gen_add_inputs(b200_ctx* ctx, void* d_At, const void* d_u, uint64_t nc, uint64_t ni) {
    gen_add_inputs_kernel<<<(unsigned)((ni + 1 + 127) / 128), 128, 0, ctx->stream>>>((Fr*)d_At, (const Fr*)d_u, nc, ni);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}
void gen_abc(b200_ctx* ctx, const void* d_toxic, const void* d_consts, const void* At, const void* Bt, const void* Ct, uint64_t ni, uint64_t nv,
             void* d_out) {
    gen_abc_kernel<<<(unsigned)((nv + 1 + 127) / 128), 128, 0, ctx->stream>>>((const Fr*)d_toxic, (const Fr*)d_consts, (const Fr*)At, (const Fr*)Bt,
                                                                              (const Fr*)Ct, ni, nv, (Fr*)d_out);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}
void gen_h(b200_ctx* ctx, const void* d_toxic, const void* d_consts, uint64_t n, void* d_out) {
    if (!n) return;
    gen_h_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>((const Fr*)d_toxic, (const Fr*)d_consts, n, (Fr*)d_out);
    ctx->launches++;
    B200_CUDA_CHECK(cudaGetLastError());
}

}  // namespace b200
