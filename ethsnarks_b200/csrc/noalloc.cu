// b200_g1_sum_noalloc: the G1 point sum of b200_g1_sum (api.cu: group_sum -> g_finish) for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Inputs and outputs are DEVICE buffers
// of the caller; the only working memory — one XYZZ partial sum per block of the first stage — lives in a file-scope
// __device__ array, so the entry point performs no cudaMalloc / cudaFree and calls no allocating library routine.
// No counterpart in the reference.
//
// The array is shared by every call on a device and keeps the last call's contents.  Nothing in it has to start at zero:
// stage 1 writes slot b from block b for every block it launches (an empty block writes the point at infinity) and stage 2
// reads exactly those slots.  Two calls that overlap on the GPU (different streams / contexts of one device) race on it.
#include "../../include/b200snark.h"
#include "msm_common.cuh"
#include <mutex>

namespace b200 {

#define B200_NOALLOC_BLOCKS 256                 // stage-1 blocks = partial sums held in the static scratch
#define B200_NOALLOC_THREADS 128
#define B200_NOALLOC_SCRATCH_BYTES (B200_NOALLOC_BLOCKS * 128)   // 256 XYZZ G1 points = 32 KB per GPU, held until exit

static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

// tree sum of the block's accumulators through shared memory; the result is valid in thread 0
static __device__ G1XYZZ block_sum(G1XYZZ acc, G1XYZZ* sh) {
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (unsigned s = blockDim.x / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            G1XYZZ a = sh[threadIdx.x];
            a.add(sh[threadIdx.x + s]);
            sh[threadIdx.x] = a;
        }
        __syncthreads();
    }
    return sh[0];
}

// stage 1: block b sums the points b*T + t, b*T + t + grid*T, ... into partial[b]
__global__ void __launch_bounds__(B200_NOALLOC_THREADS) noalloc_sum_partial(const G1Affine* __restrict__ pts, size_t n, G1XYZZ* __restrict__ partial) {
    __shared__ G1XYZZ sh[B200_NOALLOC_THREADS];
    G1XYZZ acc = G1XYZZ::inf();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) acc.add_mixed(pts[i]);
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// stage 2: one block sums the partials and normalises: out = affine point (Montgomery, (0,0) at infinity) + u32 flag
__global__ void __launch_bounds__(B200_NOALLOC_THREADS) noalloc_sum_final(const G1XYZZ* __restrict__ partial, int n_partial, G1Affine* out) {
    __shared__ G1XYZZ sh[B200_NOALLOC_THREADS];
    G1XYZZ acc = G1XYZZ::inf();
    for (int i = threadIdx.x; i < n_partial; i += blockDim.x) acc.add(partial[i]);
    acc = block_sum(acc, sh);
    if (threadIdx.x == 0) {
        *out = acc.to_affine_low_latency();
        *reinterpret_cast<uint32_t*>(out + 1) = acc.is_inf() ? 1u : 0u;
    }
}

// Device address of the scratch array, looked up once per device with that device current.  In host code the name
// g_noalloc_scratch is only a stand-in for the symbol, not a device address: kernels get the pointer from here.
static G1XYZZ* noalloc_scratch_ptr(int device) {
    static std::mutex mu;
    static void* ptr[64] = {};
    if (device < 0 || device >= 64) throw std::invalid_argument("b200_g1_sum_noalloc: device index out of range");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        B200_CUDA_CHECK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        ptr[device] = p;
    }
    return (G1XYZZ*)ptr[device];
}

}  // namespace b200

using namespace b200;

extern "C" int b200_g1_sum_noalloc(b200_ctx* ctx, const uint64_t* d_xy, size_t n, uint64_t* d_out) {
    if (!ctx || !d_out || (n && !d_xy)) return B200_ERR_ARG;
    if (n > B200_G1_SUM_NOALLOC_MAX) {
        ctx->last_error = "b200_g1_sum_noalloc: n = " + std::to_string(n) + " exceeds the maximum of " +
                          std::to_string((size_t)B200_G1_SUM_NOALLOC_MAX) + " points (static scratch: " +
                          std::to_string((size_t)B200_NOALLOC_SCRATCH_BYTES) + " B)";
        return B200_ERR_ARG;
    }
    try {
        B200_CUDA_CHECK(cudaSetDevice(ctx->device));
        G1XYZZ* partial = noalloc_scratch_ptr(ctx->device);
        const size_t want = (n + B200_NOALLOC_THREADS - 1) / B200_NOALLOC_THREADS;
        const int blocks = (int)std::max<size_t>(1, std::min<size_t>(B200_NOALLOC_BLOCKS, want));
        noalloc_sum_partial<<<blocks, B200_NOALLOC_THREADS, 0, ctx->stream>>>((const G1Affine*)d_xy, n, partial);
        noalloc_sum_final<<<1, B200_NOALLOC_THREADS, 0, ctx->stream>>>(partial, blocks, (G1Affine*)d_out);
        ctx->launches += 2;
        B200_CUDA_CHECK(cudaGetLastError());
        return B200_OK;
    } catch (const std::invalid_argument& e) {
        ctx->last_error = e.what();
        return B200_ERR_ARG;
    } catch (const std::exception& e) {
        ctx->last_error = e.what();
        return B200_ERR_CUDA;
    }
}
