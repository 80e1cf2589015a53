// Internal host-side interface between the translation units of libb200snark.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>
#include <map>
#include <stdexcept>
#include <functional>

struct b200_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string last_error;
    int sm_count = 148;
    // growable named scratch buffers (device), reused across calls
    struct Slot { void* p = nullptr; size_t cap = 0; };
    std::map<std::string, Slot> slots;
    // phase timings of the last prove / msm / h call (CUDA events on `stream`)
    std::vector<std::pair<std::string, std::pair<cudaEvent_t, cudaEvent_t>>> timing_events;
    std::vector<std::pair<std::string, float>> timings;
    std::vector<float> timing_starts;   // start of each phase relative to the first phase of the call, ms
    std::vector<unsigned long long> nvtx_ids;   // NVTX range of each phase (same names; no-ops unless a tool is attached)
    bool timing_enabled = true;
    unsigned long long launches = 0;   // kernels launched by this library since init
    int window_override = 0;           // 0 = cost model
    int chunk_override = 0;            // 0 = auto
    int g2_acc_smem = 0;               // dynamic shared memory requested by msm_accumulate<G2> only to cap its blocks per SM,
                                       // so that G1 accumulate blocks of the other lanes are co-resident with it
    int g2_pair = 2;                   // G2 accumulate with one point per lane pair (msm_acc_pair.cuh): 2 = lazy operand forms,
                                       // 1 = strict, 3 = lazy + out-of-line products; 0 = one thread per point
    int finalize_chunks = 1;           // 1 = stitch chunk partial sums with one thread per chunk edge (msm_finalize_chunks)
    int finalize_big_seg = 1;          // 1 = giant buckets (the "scalar == 1" bucket of real witnesses) summed in two levels
    int ntt_radix4 = 1;                // 1 = NTT passes with two stages per shared-memory round trip (ntt_pass4)
    int shared_sort = 1;               // 1 = the witness MSMs of one prove (A, B, L) consume ONE sorted digit list when their
                                       // queries are dense and their tables share a window size
    cudaEvent_t ev_shared = nullptr;
    int pair_reduce = 1;               // 1 = G2 finalize and stages A, B of the bucket reduction with one full addition per lane pair
    int pair_tail = 1;                 // 1 = the latency-bound stages B - E of the G2 bucket reduction with one point per lane pair too
    int reduce_fold = 1;               // stage B of the 2D bucket reduction as halving passes (msm_rc_fold): 1 = for bucket sets of
                                       // >= 2^17 (measured: pays from there on, costs ~20 us of extra launches below), 2 = always, 0 = never
    int reduce_1d = 0;                 // 1 = running-sum bucket reduction instead of the two-dimensional one
    int precompute = 1;                // build fixed-base window tables for resident keys / base sets
    // lanes: extra streams so the independent MSMs of one prove overlap (the latency-bound reduction tail of one MSM
    // hides under the accumulate kernel of the next); each lane has its own scratch buffers
    cudaStream_t lanes[4] = {nullptr, nullptr, nullptr, nullptr};   // lanes[0] == stream
    cudaEvent_t lane_done[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t ev_witness = nullptr;
    int lane = 0;                      // lane the next kernels / scratch requests go to
    int overlap = 1;
    bool fixed_table_ready[2] = {false, false};   // generator window tables of the fixed-base kernel (G1, G2) built
    cudaStream_t cur() const { return lane ? lanes[lane] : stream; }
    // multi-device context (b200_init with n_devices > 1): this is the context of the first device; one more context per
    // further device, each driven by its own host thread so that the devices' launch sequences are issued concurrently
    std::vector<b200_ctx*> peers;
    struct Worker;
    std::vector<Worker*> workers;      // workers[i] drives peers[i]
    cudaEvent_t ev_front = nullptr;    // "this device's share of the transform chains is done" (multi-device proofs)
    int split_chains = 1;              // multi-device / multi-rank proofs: the three evaluation vectors' first two transforms are
                                       // shared out over the devices instead of replicated on each
    cudaEvent_t ev_group = nullptr;    // on a peer: "this device's partial sums have arrived in the first device's gather buffer"

    void* scratch(const char* name, size_t bytes);   // throws std::runtime_error on failure
    void t_begin(const char* name);
    void t_end();
    void t_collect();
};

// one persistent host thread bound to one peer device
struct b200_ctx::Worker {
    explicit Worker(int device);
    ~Worker();
    void start(std::function<void()> job);   // returns at once
    void wait();                             // rethrows what the job threw
    struct Impl;
    Impl* impl;
};

#define B200_CUDA_CHECK(expr)                                                                                   \
    do {                                                                                                        \
        cudaError_t _e = (expr);                                                                                \
        if (_e != cudaSuccess)                                                                                  \
            throw std::runtime_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " @" + __FILE__ + ":" + \
                                     std::to_string(__LINE__));                                                 \
    } while (0)

namespace b200 {

// A sorted digit list that several MSMs over the SAME scalars consume (one prove: A, B in G2 [and G1] and L all multiply the
// witness): digits, sort and offsets run once instead of once per query.  The list is over scalar (variable) indices -
// value = sign << 31 | window << 26 | variable.  Each consumer rewrites the values into its own table indices in one
// streaming pass (msm_translate: 8 B per pair against three 16 B radix passes): a variable maps to a table position through
// MsmPos (inverse index list of a sparse query, or an index range); a variable the query does not hold points at the
// all-zero entry appended to the table, i.e. the point at infinity, which every accumulate kernel skips.
struct MsmShared {
    int c = 0;
    uint32_t n = 0;                 // scalars in the list
    const uint32_t* keys = nullptr;
    const uint32_t* vals = nullptr;
    const uint64_t* offsets = nullptr;
};
struct MsmPos {
    const uint32_t* map;            // variable -> position in the query, 0xffffffff = absent; nullptr = the range [off, off + count)
    uint32_t off, count, stride;    // stride = table stride (entries per window); the zero entry sits at index W * stride
};
#define B200_SHARED_VAR_BITS 26

// --- msm.cu ---------------------------------------------------------------------------------------------------
// Device-resident inputs; result is ONE XYZZ point written to d_out (G1: 4x32 B, G2: 4x64 B).
//   group: 1 = G1, 2 = G2.  d_scalars: Fr elements (Montgomery if scalars_mont, else standard form), optionally
//   gathered: scalar_i = d_scalars[d_gather[i]].  d_bases: affine Montgomery, (0,0) = infinity.
#define B200_DECL_GROUP(SUFFIX)                                                                                                   \
    void msm_run_##SUFFIX(b200_ctx* ctx, const void* d_bases, const void* d_scalars, const uint32_t* d_gather, size_t n,         \
                          bool scalars_mont, void* d_out_xyzz, int table_c, size_t table_stride);                                \
    void g_precompute_##SUFFIX(b200_ctx* ctx, const void* d_bases, size_t n, int c, void* d_table);                              \
    void g_finish_##SUFFIX(b200_ctx* ctx, const void* d_xyzz_terms, int n_terms, const void* d_affine_terms, int n_affine,       \
                           void* d_out_affine, bool to_std);                                                                     \
    void g_sum_strided_##SUFFIX(b200_ctx* ctx, const void* d_base, size_t stride, int n, const void* d_extra_affine,             \
                                void* d_out_affine, bool to_std);                                                                \
    void g_fixed_base_##SUFFIX(b200_ctx* ctx, const void* d_scalars_mont, size_t n, void* d_out_affine);                         \
    void g_add_affine_##SUFFIX(b200_ctx* ctx, const void* d_a, const void* d_b, void* d_out, size_t n);                          \
    void g_scalar_mul_##SUFFIX(b200_ctx* ctx, const void* d_affine_base, const void* d_scalar_mont, void* d_out_xyzz);
B200_DECL_GROUP(g1)
B200_DECL_GROUP(g2)
// one launch: sum of `world` 256-byte partial records + alpha_g1 / beta_g2 -> proof points (standard form) at fin + 0 / 256 / 512
void g_combine3(b200_ctx* ctx, const void* d_parts, int world, const void* d_alpha_g1, const void* d_beta_g2, void* d_fin);
// shared list: digits + sort + offsets of n Montgomery scalars for window size c on the current lane (table mode), then one
// consume per query on any lane that has waited for the prepare
void msm_shared_prepare(b200_ctx* ctx, const void* d_scalars_mont, size_t n, int c, MsmShared* out);
void msm_shared_consume_g1(b200_ctx* ctx, const MsmShared& sh, const void* d_table, MsmPos pos, void* d_out_xyzz);
void msm_shared_consume_g2(b200_ctx* ctx, const MsmShared& sh, const void* d_table, MsmPos pos, void* d_out_xyzz);
// g_finish: out = sum(xyzz terms) + sum(affine terms) normalised to affine (Montgomery, or standard form if to_std),
//           written as the affine point followed by a u32 infinity flag.
// table_c != 0: d_bases is a window table (table[w * table_stride + i] = 2^(table_c w) P_i) built by g_precompute
inline void msm_run(b200_ctx* c, int group, const void* b, const void* s, const uint32_t* g, size_t n, bool mont, void* out,
                    int table_c = 0, size_t table_stride = 0) {
    if (group == 1) msm_run_g1(c, b, s, g, n, mont, out, table_c, table_stride); else msm_run_g2(c, b, s, g, n, mont, out, table_c, table_stride);
}
inline void g_precompute(b200_ctx* c, int group, const void* b, size_t n, int win, void* t) {
    if (group == 1) g_precompute_g1(c, b, n, win, t); else g_precompute_g2(c, b, n, win, t);
}
int msm_table_window(size_t n);   // window size the cost model picks for a table over n bases
inline void g_finish(b200_ctx* c, int group, const void* xt, int nx, const void* at, int na, void* out, bool to_std) {
    if (group == 1) g_finish_g1(c, xt, nx, at, na, out, to_std); else g_finish_g2(c, xt, nx, at, na, out, to_std);
}
inline void g_sum_strided(b200_ctx* c, int group, const void* base, size_t stride, int n, const void* extra, void* out, bool to_std) {
    if (group == 1) g_sum_strided_g1(c, base, stride, n, extra, out, to_std); else g_sum_strided_g2(c, base, stride, n, extra, out, to_std);
}
inline void g_fixed_base(b200_ctx* c, int group, const void* s, size_t n, void* out) {
    if (group == 1) g_fixed_base_g1(c, s, n, out); else g_fixed_base_g2(c, s, n, out);
}
inline void g_add_affine(b200_ctx* c, int group, const void* a, const void* b, void* out, size_t n) {
    if (group == 1) g_add_affine_g1(c, a, b, out, n); else g_add_affine_g2(c, a, b, out, n);
}
inline void g_scalar_mul(b200_ctx* c, int group, const void* base, const void* k, void* out) {
    if (group == 1) g_scalar_mul_g1(c, base, k, out); else g_scalar_mul_g2(c, base, k, out);
}

// --- ntt.cu ---------------------------------------------------------------------------------------------------
struct Domain {
    uint32_t logm = 0;
    void* tw_fwd = nullptr;    // omega^e, e < m/2           (Fr Montgomery)
    void* tw_inv = nullptr;    // omega^-e
    void* cs_fwd = nullptr;    // m^-1 * g^bitrev(p)         (post-scale of iNTT in bit-reversed order, feeds coset NTT)
    void* cs_inv = nullptr;    // m^-1 * g^-bitrev(p)        (post-scale of the final icosetFFT, bit-reversed order)
    void* zinv = nullptr;      // (g^m - 1)^-1, one element
};
void domain_create(b200_ctx* ctx, uint32_t logm, Domain* out);
void domain_free(Domain* d);
// in-place transforms on device data (natural order in/out): mode 0 FFT, 1 iFFT, 2 cosetFFT, 3 icosetFFT
void ntt_natural(b200_ctx* ctx, const Domain& dom, void* d_data, int mode);
// fused H chain: a, b, c hold the evaluations on S (natural order, m entries each); on return c holds the
// coefficients of H in BIT-REVERSED order (c[p] = h[bitrev(p)]), a and b are clobbered.
void h_chain(b200_ctx* ctx, const Domain& dom, void* d_a, void* d_b, void* d_c);
void h_chain_front(b200_ctx* ctx, const Domain& dom, void* const* d_x, int count);      // count of the vectors: S -> coset evaluations
void h_chain_back(b200_ctx* ctx, const Domain& dom, const void* d_a, const void* d_b, const void* d_c, void* d_out);  // pointwise + icosetFFT
void bitrev_permute(b200_ctx* ctx, const void* d_in, void* d_out, uint32_t logm, size_t elem_bytes);
double imad_peak(b200_ctx* ctx, int reps);   // wide 32x32+64 multiply-adds per second, best of reps
void field_op(b200_ctx* ctx, int field, int op, const void* d_a, const void* d_b, void* d_out, size_t n);

// --- spmv (in ntt.cu) -----------------------------------------------------------------------------------------
struct DeviceCsr {
    uint64_t n_rows = 0, nnz = 0;
    uint64_t* row_ptr = nullptr;
    uint32_t* col = nullptr;
    uint32_t* cidx = nullptr;
    // long rows (length > LONG_ROW) are split into segments handled by a block each
    uint32_t n_seg = 0;
    uint32_t* seg_row = nullptr;     // row of each segment
    uint64_t* seg_begin = nullptr;   // [n_seg + 1) nnz ranges (seg_begin[i]..seg_end[i])
    uint64_t* seg_end = nullptr;
    uint32_t n_long = 0;
    uint32_t* long_row = nullptr;    // distinct long rows
    uint32_t* long_seg0 = nullptr;   // [n_long + 1) first segment of each long row
};
void csr_upload(b200_ctx* ctx, const uint64_t* row_ptr, const uint32_t* col, const uint32_t* cidx, uint64_t n_rows, DeviceCsr* out);
void csr_free(DeviceCsr* c);
// out[row] = sum coeffs[cidx[k]] * w[col[k]]  for row < n_rows
void spmv(b200_ctx* ctx, const DeviceCsr& M, const void* d_coeffs, const void* d_w, void* d_out);
// out[j * master.n_rows + r] = master row r through translation table j (r1cs_constraint_light_instance)
void spmv_instances(b200_ctx* ctx, const DeviceCsr& master, const void* d_coeffs, const void* d_w, const uint32_t* d_tables, uint32_t tlen,
                    uint64_t n_inst, void* d_out);
void witness_rows(b200_ctx* ctx, const void* d_w, void* d_a, void* d_b, void* d_c, uint64_t nc, uint64_t ni, uint64_t m);

// --- key generator scalars (gen.cu) -----------------------------------------------------------------------------
void gen_consts(b200_ctx* ctx, const void* d_toxic, uint32_t logm, void* d_consts);
void gen_lagrange(b200_ctx* ctx, const void* d_toxic, const void* d_consts, const Domain& dom, void* d_u);
void gen_add_inputs(b200_ctx* ctx, void* d_At, const void* d_u, uint64_t nc, uint64_t ni);
void gen_abc(b200_ctx* ctx, const void* d_toxic, const void* d_consts, const void* At, const void* Bt, const void* Ct, uint64_t ni, uint64_t nv,
             void* d_out);
void gen_h(b200_ctx* ctx, const void* d_toxic, const void* d_consts, uint64_t n, void* d_out);

}  // namespace b200
