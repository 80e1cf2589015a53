/* b200snark.h — C ABI of libb200snark.so: the B200 (sm_100a) Groth16 proving backend behind ethsnarks' prover.
 *
 * Every entry point replaces a piece of the reference's CPU prove path (paths relative to the reference tree;
 * LS = depends/libsnark/libsnark, FF = depends/libsnark/depends/libff/libff, FQ = depends/libfqfft/libfqfft):
 *
 *   b200_key_load        ethsnarks::load_proving_key                    src/stubs.cpp:36-39, src/utils.hpp:176-185,
 *                        operator>>(istream&, proving_key[_nozk]&)      src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:68-87,124-143
 *   b200_circuit_load    ProverContext.constraint_system + get_domain   …zok.hpp:279-291, src/stubs.cpp:49-75
 *   b200_prove           r1cs_gg_ppzksnark_zok_prover<ppT>               …zok.tcc:451-550  (r = s = NULL), and the r,s form of
 *                        LS/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:389-503
 *   b200_h               r1cs_to_qap_witness_map                         LS/reductions/r1cs_to_qap/r1cs_to_qap.tcc:217-302
 *   b200_ntt             recursive_domain::{FFT,iFFT,cosetFFT,icosetFFT} FQ/evaluation_domain/domains/recursive_domain.tcc:102-142
 *   b200_msm_g1/g2       libff::multi_exp / multi_exp_with_mixed_addition FF/algebra/scalar_multiplication/multiexp.tcc:595-714
 *   b200_field_op        Fp_model mul / add / sub / as_bigint / inverse  FF/algebra/fields/fp.tcc:23-186,228-238,310-510,642-692
 *   b200_g1_add/g2_add   G1/G2 ::add / mixed_add / dbl                   FF/algebra/curves/mcl_bn128/mcl_bn128_g1.cpp:136-164
 *   b200_g1/g2_fixed_base  batch_exp (generator side)                    FF/algebra/scalar_multiplication/multiexp.tcc:838-905
 *
 * Conventions (the reference's existing C ABI precedent is src/verify_dll.cpp:3-10): plain pointers and sizes, all
 * buffers caller-allocated, int status return (0 = ok), no ownership transfer except the opaque handles.
 * Field elements are 4 x uint64 little-endian limbs.  "mont" = Montgomery form with R = 2^256, which is the
 * reference's in-memory and key-file representation (Fp_model::mont_repr, MONTGOMERY_OUTPUT); "std" = the plain
 * integer.  G1 affine = x,y (8 limbs); G2 affine = x.a, x.b, y.a, y.b (16 limbs; x = x.a + x.b*u, mcl order).
 * One handle is bound to one CUDA device (or, from b200_init with n_devices > 1, to one set of devices); calls on one handle
 * must be serialised by the caller (the reference prover is not re-entrant either).  There is no CPU fallback: without a CUDA device b200_init fails with B200_ERR_CUDA.
 */
#ifndef B200SNARK_H_
#define B200SNARK_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200_ctx b200_ctx;
typedef struct b200_key b200_key;
typedef struct b200_bases b200_bases;

enum { B200_OK = 0, B200_ERR_CUDA = 1, B200_ERR_ARG = 2, B200_ERR_FORMAT = 3, B200_ERR_STATE = 4 };
/* proving-key byte layouts: mcl points (default CURVE=MCL_BN128 build) or libff alt_bn128 points; "NOZK" =
 * r1cs_gg_ppzksnark_zok_proving_key_nozk (B query in G2 only), "FULL" = …_proving_key (B query = G2,G1 pairs). */
enum { B200_FMT_MCL_NOZK = 0, B200_FMT_MCL_FULL = 1, B200_FMT_ALT_NOZK = 2, B200_FMT_ALT_FULL = 3,
       B200_FMT_CACHE = 4 /* device-native key cache written by b200_key_export */ };
enum { B200_OP_MUL = 0, B200_OP_ADD = 1, B200_OP_SUB = 2, B200_OP_TO_MONT = 3, B200_OP_FROM_MONT = 4, B200_OP_INV = 5, B200_OP_SQR = 6, B200_OP_NEG = 7, B200_OP_INV_BINARY = 8,
       B200_OP_DOT2 = 9 /* a*b + b*b with one Montgomery reduction (the fused row product of the paired G2 kernel) */,
       B200_OP_DOT4 = 10 /* (a + b)^2 as FOUR products a*b + a*b + (p-a)*(p-a) + b*b under one reduction (the Y3 row set of the paired G2 kernel) */,
       B200_OP_DIFF = 11 /* a*a - b*b as one two-product row set (the Y3 of every G1 XYZZ addition) */ };
enum { B200_NTT_FFT = 0, B200_NTT_IFFT = 1, B200_NTT_COSET_FFT = 2, B200_NTT_ICOSET_FFT = 3 };

/* One R1CS matrix in CSR form: row i holds sum_k coeff_table[coeff_idx[k]] * w[col[k]], k in [row_ptr[i], row_ptr[i+1]).
 * col 0 is the constant ONE (protoboard values[0]); coeff_idx indexes the table given to b200_circuit_load, the
 * flattened form of the reference's ConstantStorage (LS/relations/variable.hpp:47-86). */
typedef struct {
    const uint64_t* row_ptr;   /* n_rows + 1 */
    const uint32_t* col;
    const uint32_t* coeff_idx;
    uint64_t n_rows;
} b200_csr;

/* One handle = one device set.  n_devices == 1 (or 0: device 0): a single-GPU context, the form used with one process per
 * GPU.  n_devices > 1: ONE context owning several GPUs of the box, for callers that are a single process (the reference
 * prover is a library call, src/stubs.cpp:42-47): keys loaded or assembled on it are split 1/N per device, b200_prove runs on
 * all of them (one host thread per device issues its launches; the witness and the per-device partial sums travel over peer
 * memory, so every pair of the devices must have peer access), and returns the same proof as one GPU.  Everything else
 * (standalone MSM / NTT / field ops, the key generator, r,s proving) runs on the first device. */
int b200_init(const int* device_ids, int n_devices, b200_ctx** out);
int b200_device_count(b200_ctx* ctx);
void b200_shutdown(b200_ctx* ctx);
const char* b200_strerror(int code);
const char* b200_last_error(b200_ctx* ctx);
void* b200_stream(b200_ctx* ctx);                        /* the cudaStream_t every kernel of this ctx is launched on */
unsigned long long b200_launch_count(b200_ctx* ctx);     /* kernels launched by this library on ctx since b200_init */
/* Tuning knobs; every setting gives the same results (tests run the alternatives against each other and the oracle).
 *   "window"          c of the windowed MSM and of tables built afterwards; 0 = cost model (libsnark::Config::multi_exp_c)
 *   "chunk"           pairs per accumulate thread; 0 = by size
 *   "precompute"      1 = fixed-base window tables for resident keys / base sets (default), 0 = plain windowed path
 *   "overlap"         1 = the MSMs of one prove on separate lanes (default), 0 = serialised (per-kernel timing)
 *   "timing"          1 = record CUDA-event phase timings (default)
 *   "g2_pair"         G2 accumulate: 2 = one point per lane pair, lazy operand forms (default), 1 = strict forms,
 *                     3 = lazy + out-of-line products, 0 = one thread per point
 *   "finalize_chunks" 1 = stitch chunk partial sums with one thread per chunk edge (default), 0 = one per bucket
 *   "pair_tail"       1 = the latency-bound last stages of the G2 bucket reduction with one point per lane pair (default)
 *   "split_chains"    1 = a multi-device context shares the first two transforms of the three evaluation vectors out over its
 *                     devices (default); 0 = every device runs all of them
 *   "finalize_big_seg" 1 = buckets spanning many chunks (the "scalar == 1" bucket of 0/1-heavy witnesses) summed in two levels,
 *                     64 blocks per bucket then one (default); 0 = one block per bucket
 *   "reduce_fold"     stage B of the bucket reduction as halving passes: 1 = for >= 2^17 buckets (default), 2 = always, 0 = never
 *   "pair_reduce"     1 = G2 finalize and bucket-reduction additions with one full addition per lane pair (default), 0 = per thread
 *   "reduce_1d"       1 = running-sum bucket reduction instead of the two-dimensional one
 *   "ntt_radix4"      1 = NTT passes with two stages per shared-memory round trip (default), 0 = one
 *   "shared_sort"     1 = the witness MSMs of one prove consume ONE sorted digit list when their queries are dense (default)
 */
int b200_set_option(b200_ctx* ctx, const char* name, long value);
int b200_timing_count(b200_ctx* ctx);                    /* phase timings (CUDA events) of the last call */
int b200_timing_get(b200_ctx* ctx, int i, char* name, size_t cap, float* ms);
int b200_timing_start(b200_ctx* ctx, int i, float* ms);  /* start of phase i relative to the first phase of the call */

/* integer-pipe probe: sustained 32x32+64-bit multiply-adds per second of this device (IMAD.WIDE.U32, all SMs) —
 * the measured roofline denominator for the MSM / NTT kernels */
int b200_imad_peak(b200_ctx* ctx, int reps, double* macs_per_s);

/* ---- sub-operations (parity tests, config-5 MSM sweep) ---- */
int b200_field_op(b200_ctx* ctx, int field /*0 Fr, 1 Fq*/, int op, const uint64_t* a, const uint64_t* b, uint64_t* out, size_t n);
int b200_g1_add(b200_ctx* ctx, const uint64_t* a_xy, const uint8_t* a_inf, const uint64_t* b_xy, const uint8_t* b_inf,
                uint64_t* out_xy, uint8_t* out_inf, size_t n);
int b200_g2_add(b200_ctx* ctx, const uint64_t* a_xy, const uint8_t* a_inf, const uint64_t* b_xy, const uint8_t* b_inf,
                uint64_t* out_xy, uint8_t* out_inf, size_t n);
/* sum of n affine points (Montgomery limbs, optional infinity flags) in one kernel, normalised to affine: the gather-and-add
 * "reduce" of a sharded MSM (NCCL has no elliptic-curve reduction) */
int b200_g1_sum(b200_ctx* ctx, const uint64_t* xy_mont, const uint8_t* inf, size_t n, uint64_t* out_xy_mont, uint8_t* out_inf);
int b200_g2_sum(b200_ctx* ctx, const uint64_t* xy_mont, const uint8_t* inf, size_t n, uint64_t* out_xy_mont, uint8_t* out_inf);
/* b200_g1_sum for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * No counterpart in the reference.  The call allocates nothing — no cudaMalloc,
 * cudaFree or allocating library routine: its working memory (one partial sum per block of the first of its two kernels) is a
 * file-scope 32 KB __device__ array that every GPU loading the library holds until exit.  d_xy_mont and d_out are DEVICE
 * pointers of the caller: n affine points (8 limbs each, Montgomery, (0,0) = infinity) in; the affine sum (8 limbs,
 * Montgomery, (0,0) at infinity) followed by a uint32 infinity flag out, i.e. 68 bytes.  The two kernels are queued on
 * b200_stream(ctx) and the call returns without synchronising.  n > B200_G1_SUM_NOALLOC_MAX is rejected with B200_ERR_ARG
 * (text in b200_last_error).  All calls on one device share the array and it keeps the last call's contents (nothing in it
 * needs clearing: every slot that is read was written by the same call); calls that overlap on the GPU race on it.  The first
 * call on a device looks the array's address up (cudaGetSymbolAddress) — make it once before starting a graph capture. */
#define B200_G1_SUM_NOALLOC_MAX (1u << 20)
int b200_g1_sum_noalloc(b200_ctx* ctx, const uint64_t* d_xy_mont, size_t n, uint64_t* d_out);
int b200_g1_fixed_base(b200_ctx* ctx, const uint64_t* scalars_mont, size_t n, uint64_t* out_xy_mont, uint8_t* out_inf);
int b200_g2_fixed_base(b200_ctx* ctx, const uint64_t* scalars_mont, size_t n, uint64_t* out_xy_mont, uint8_t* out_inf);
int b200_ntt(b200_ctx* ctx, uint64_t* data_mont, uint32_t log_m, int mode);
int b200_msm_g1(b200_ctx* ctx, const uint64_t* bases_xy_mont, const uint8_t* inf, const uint64_t* scalars_mont, size_t n,
                uint64_t out_xy_mont[8], uint8_t* out_inf);
int b200_msm_g2(b200_ctx* ctx, const uint64_t* bases_xy_mont, const uint8_t* inf, const uint64_t* scalars_mont, size_t n,
                uint64_t out_xy_mont[16], uint8_t* out_inf);

/* ---- device-resident base sets (what a loaded key holds; used directly by the MSM sweep) ---- */
int b200_bases_upload(b200_ctx* ctx, int group /*1|2*/, const uint64_t* xy_mont, const uint8_t* inf, size_t n, b200_bases** out);
int b200_bases_generate(b200_ctx* ctx, int group, const uint64_t* scalars_mont, size_t n, b200_bases** out); /* k_i * generator */
/* fixed-base window tables in HBM: table[w][i] = 2^(c w) * P_i for every window w, so later MSMs over this set use ONE
 * bucket set and need no window combine.  window = 0: chosen by the cost model.  Costs ceil(255/c) x the base memory. */
int b200_bases_precompute(b200_bases* bases, int window);
int b200_bases_download(b200_bases* bases, uint64_t* xy_mont, uint8_t* inf);
size_t b200_bases_count(b200_bases* bases);
void b200_bases_free(b200_bases* bases);
/* scalars on the host (copied inside the call) or already on the device (scalars_on_device != 0) */
int b200_msm(b200_ctx* ctx, b200_bases* bases, const uint64_t* scalars_mont, size_t n, int scalars_on_device,
             uint64_t* out_xy_mont, uint8_t* out_inf);

/* ---- the prove path ---- */
int b200_key_load(b200_ctx* ctx, const void* pk_bytes, size_t len, int fmt, b200_key** out);
/* Key file converters, host only (no device, no context): the byte-level counterparts of the reference's pk_alt2mcl and
 * pk_mcl2nozk (src/export.cpp:330-408).  Any B200_FMT_{MCL,ALT}_{NOZK,FULL} layout in, any out, except NOZK -> FULL (the stripped
 * key no longer holds what the full one needs: B200_ERR_ARG).  *len_out = bytes needed; out may be NULL / short for a size query. */
int b200_key_convert(const void* pk_bytes, size_t len, int fmt_in, int fmt_out, void* out, size_t cap, size_t* len_out);
/* keep only this rank's 1/world slice of every query (multi-GPU: one process per GPU, see b200_prove_partial) */
int b200_key_load_sharded(b200_ctx* ctx, const void* pk_bytes, size_t len, int fmt, int rank, int world, b200_key** out);
/* a key assembled from resident base sets (synthetic keys for benchmarks); takes ownership of nothing, copies nothing:
 * the base sets must outlive the key.  A_idx / B_idx: variable index of each A / B entry (NULL = identity). */
int b200_key_from_bases(b200_ctx* ctx, b200_bases* A, const uint32_t* A_idx, b200_bases* B_g2, b200_bases* B_g1 /*may be NULL*/,
                        const uint32_t* B_idx, b200_bases* H, b200_bases* L, const uint64_t* alpha_beta_delta_g1_xy /*3 x 8, mont*/,
                        const uint64_t* beta_delta_g2_xy /*2 x 16, mont*/, uint64_t num_variables, uint64_t num_inputs,
                        b200_key** out);
/* multi-GPU with b200_key_from_bases: declare that A / B_g2 / L hold this rank's slice (L starting at entry L_lo of the
 * full L query) and that H is the full query, to be sliced by rank at b200_circuit_load.  Call before circuit_load. */
int b200_key_set_shard(b200_key* key, int rank, int world, uint64_t L_lo);
/* Device-native key cache: writes the key's arrays as they sit in HBM (B200_FMT_CACHE) so that a reload is a few large
 * copies instead of the point-by-point parse of the reference layout (...zok.tcc:124-143, sparse_vector.tcc:292-322).
 * Call with buf = NULL to get the size in *len.  Only for a whole key loaded from bytes, before b200_circuit_load. */
int b200_key_export(b200_key* key, void* buf, size_t cap, size_t* len);
/* Key generator on the device: r1cs_gg_ppzksnark_zok_generator (src/r1cs_gg_ppzksnark_zok/
 * r1cs_gg_ppzksnark_zok.tcc:277-449) with the toxic waste t, alpha, beta, gamma, delta (5 x 4 limbs, Montgomery) supplied by
 * the caller, generators G1::one / G2::one.  Returns a key in the state b200_key_load leaves it in (call
 * b200_circuit_load next) and the verification key pieces in vk_xy: alpha_g1[8] | beta_g2[16] | gamma_g2[16] | delta_g2[16] |
 * IC[(num_inputs + 1) x 8], affine Montgomery, (0,0) = infinity.  full != 0 also builds the B query in G1. */
int b200_generate_key(b200_ctx* ctx, const b200_csr* A, const b200_csr* B, const b200_csr* C, const uint64_t* coeff_table_mont,
                      size_t n_coeffs, uint32_t num_inputs, uint64_t num_variables, const uint64_t* toxic_mont, int full,
                      b200_key** out, uint64_t* vk_xy);
/* A second prove slot on a resident key for pipelined proving (the reference's ProverContext reuse,
 * r1cs_gg_ppzksnark_zok.hpp:279-291): shares every read-only device buffer of `key` (bases, window tables, circuit, domain)
 * and owns only its witness / evaluation buffers, bound to another context (same device), so two proofs can be in flight
 * from two host threads.  Free the slot before the key it was attached to. */
int b200_key_attach(b200_key* key, b200_ctx* ctx, b200_key** out);
void b200_key_free(b200_key* key);
/* out[0..7]: domain_size (= num_variables + 1), |A|, |B|, |H|, |L| (this rank's counts), full(0/1), m, num_inputs */
int b200_key_info(b200_key* key, uint64_t out[8]);
int b200_circuit_load(b200_key* key, const b200_csr* A, const b200_csr* B, const b200_csr* C, const uint64_t* coeff_table_mont,
                      size_t n_coeffs, uint32_t num_inputs, uint64_t num_variables);
/* witness = full assignment, 4 * (num_variables + 1) limbs, w[0] = 1, Montgomery; a host pointer (pageable or pinned)
 * or a device pointer (resident witness) — the copy is cudaMemcpyDefault.  r = s = NULL: the reference's
 * no-ZK prover.  Outputs: affine, STANDARD form (ready for proof_to_json), infinity flags for A, B, C. */
int b200_prove(b200_key* key, const uint64_t* witness_mont, const uint64_t r_mont[4], const uint64_t s_mont[4],
               uint64_t A_xy[8], uint64_t B_xy[16], uint64_t C_xy[8], uint8_t infinity_flags[3]);
/* coefficients of H, natural order, m entries (h[m-1] = 0), Montgomery */
int b200_h(b200_key* key, const uint64_t* witness_mont, uint64_t* h_out_mont);
/* Circuits made of many instances of one master gadget (the reference's r1cs_constraint_light_instance + ITranslator,
 * LS/relations/constraint_satisfaction_problems/r1cs/r1cs.hpp:145-211, LS/relations/variable.hpp:279-314; used by
 * src/gadgets/sha256_many.hpp:75-82 and poseidon.hpp:455-462): the constraint system is the plain rows A, B, C (may be empty)
 * followed, for every instance j in order, by all rows of the master gadget with each variable index v != 0 replaced by
 * tables[j * table_len + v] (index 0, the constant ONE, is never translated).  The device keeps the master rows once and one
 * table per instance instead of the flattened rows.  Same proofs as b200_circuit_load on the flattened system. */
typedef struct {
    const b200_csr *A, *B, *C;      /* the master gadget's rows; col = master-local variable index < table_len */
    const uint32_t* tables;         /* n_instances x table_len */
    uint32_t table_len;
    uint64_t n_instances;
} b200_instances;
int b200_circuit_load_instanced(b200_key* key, const b200_csr* A, const b200_csr* B, const b200_csr* C, const b200_instances* inst,
                                const uint64_t* coeff_table_mont, size_t n_coeffs, uint32_t num_inputs, uint64_t num_variables);
/* evaluations of A, B, C on the domain S (m entries each, Montgomery): the witness-times-R1CS step alone */
int b200_witness_eval(b200_key* key, const uint64_t* witness_mont, uint64_t* a_out, uint64_t* b_out, uint64_t* c_out);
/* multi-GPU: this rank's partial sums over its key shard — evaluation_At, evaluation_Bt (G2), evaluation_Ht + Lt —
 * as affine Montgomery points: partial[0..7] A, [8..23] B, [24..31] C; flags[3]. */
int b200_prove_partial(b200_key* key, const uint64_t* witness_mont, uint64_t partial[32], uint8_t flags[3]);
/* the same, result left ON THE DEVICE: 256 bytes at d_out (A 64 | B 128 | C 64, affine Montgomery, (0,0) = infinity), queued
 * on b200_stream(ctx) without synchronising, so that a collective on that stream (ncclAllGather) can follow at once; and the
 * combine of `world` such records that are already on this device.  With these two the exchange of a sharded proof has no
 * host round trip: partial -> all-gather -> combine -> one 580-byte copy back. */
int b200_prove_partial_dev(b200_key* key, const uint64_t* witness_mont, void* d_out);
int b200_proof_combine_dev(b200_key* key, const void* d_parts /*world x 256 B*/, int world,
                           uint64_t A_xy[8], uint64_t B_xy[16], uint64_t C_xy[8], uint8_t infinity_flags[3]);
/* The staged form, for ranks that also share the TRANSFORM work of one proof: the three evaluation vectors A, B, C go through the
 * witness map and their first two transforms on different ranks (b200_chain_owner(chain, world): 0 = A, 1 = B, 2 = C), the
 * caller broadcasts each vector of coset evaluations from its owner to all ranks (device buffers of m x 32 bytes from
 * b200_key_buffers; ncclBroadcast on b200_stream(ctx)), and stage 2 - pointwise step, last transform, H MSM - leaves the same
 * 256-byte record as b200_prove_partial_dev.  Nothing synchronises.  chains: bit 0 = A, 1 = B, 2 = C (7 = all, no sharing). */
int b200_chain_owner(int chain, int world);
int b200_key_buffers(b200_key* key, void* bufs[3], size_t* bytes_each);
int b200_prove_partial_stage1(b200_key* key, const uint64_t* witness_mont, int chains);
int b200_prove_partial_stage2(b200_key* key, void* d_out);
/* sums `world` partials (after an all-gather), adds alpha_g1 / beta_g2, returns the proof points in standard form */
int b200_proof_combine(b200_key* key, const uint64_t* partials /*world x 32*/, const uint8_t* flags /*world x 3*/, int world,
                       uint64_t A_xy[8], uint64_t B_xy[16], uint64_t C_xy[8], uint8_t infinity_flags[3]);

#ifdef __cplusplus
}
#endif
#endif /* B200SNARK_H_ */
