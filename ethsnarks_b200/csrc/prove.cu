// Proving-key ingest and the Groth16 prove orchestration (C ABI part 2; see include/b200snark.h for the reference map).
//   key layout parsed here:   /root/reference/src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:54-87,109-143,
//                             LS/common/data_structures/sparse_vector.tcc:273-322,
//                             FF/algebra/curves/mcl_bn128/mcl_bn128_g1.cpp:181-295 (mcl points),
//                             FF/algebra/curves/alt_bn128/alt_bn128_g1.cpp:404-465 (libff points)
//   prover:                   …zok.tcc:451-550 (no-ZK) and LS/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:485-493 (r,s)
#include "../../include/b200snark.h"
#include "internal.hpp"
#include <cstring>
#include <algorithm>
#include <memory>

using namespace b200;

struct b200_bases {   // same definition as in api.cu
    b200_ctx* ctx;
    int group;
    size_t n;
    void* d;
    void* table;
    int table_c;
};

struct b200_key {
    b200_ctx* ctx = nullptr;
    bool full = false;          // has the B query in G1 (r,s mode possible)
    bool owns_bases = true;
    int children = 0;           // prove slots attached to this key (b200_key_attach) that are still alive
    bool free_pending = false;  // b200_key_free was called while slots were alive: released when the last slot goes
    b200_key* parent = nullptr; // of an attached slot
    bool attached = false;      // a second prove slot on a resident key: shares every read-only buffer, owns only d_w/d_a/d_b/d_c
    int rank = 0, world = 1;
    uint64_t domain_size = 0;   // num_variables + 1
    uint64_t nA = 0, nB = 0, nH = 0, nL = 0;   // entries held by this rank
    uint64_t L_lo = 0;          // first L entry held by this rank (scalar offset)
    uint64_t H_lo = 0;          // first (bit-reversed) H position held by this rank
    void *A = nullptr, *B2 = nullptr, *B1 = nullptr, *H = nullptr, *L = nullptr;
    uint32_t *A_idx = nullptr, *B_idx = nullptr;
    uint32_t *A_pos = nullptr, *B_pos = nullptr;   // variable -> entry of the A / B query (0xffffffff = absent): shared digit list
    bool A_ident = false, B_ident = false;         // the query's index list is first, first + 1, first + 2, ...: no map needed
    uint32_t A_first = 0, B_first = 0;             // (a whole dense key: first = 0; a shard of one: the start of its slice)
    // fixed-base window tables (built at circuit_load when ctx->precompute): [W][n] affine, window size per query
    void *A_t = nullptr, *B2_t = nullptr, *B1_t = nullptr, *H_t = nullptr, *L_t = nullptr;
    int cA = 0, cB = 0, cH = 0, cL = 0;
    bool H_bitrev = false;      // H bases stored in bit-reversed order (needs the domain size: set at circuit_load)
    std::vector<uint8_t> H_host;   // natural-order H bases kept on the host until the domain size is known
    void* consts_g1 = nullptr;  // alpha_g1, beta_g1, delta_g1 (affine Montgomery)
    void* consts_g2 = nullptr;  // beta_g2, delta_g2
    // circuit
    bool has_circuit = false;
    uint64_t nc = 0, ni = 0, nv = 0, m = 0;
    uint32_t logm = 0;
    DeviceCsr M[3];
    // instanced rows (b200_circuit_load_instanced): master gadget rows + one translation table per instance
    DeviceCsr Mi[3];
    uint32_t* inst_tables = nullptr;
    uint32_t inst_tlen = 0;
    uint64_t n_inst = 0, n_plain = 0;
    void* coeffs = nullptr;
    Domain dom;
    void *d_w = nullptr, *d_a = nullptr, *d_b = nullptr, *d_c = nullptr;
    void* d_h = nullptr;        // staged (multi-GPU) proofs: H's coefficients, so that d_c - which other devices may still be
                                // reading - is not overwritten
    uint64_t H_total = 0;       // |H_query| in the key file
    // a key on a multi-device context: one shard per device (shards[i] on ctx->peers[i - 1], shards[0] on ctx); the
    // wrapper itself holds no device memory
    std::vector<b200_key*> shards;
    void* gather = nullptr;     // on the first device: world x 256 B, the shards' affine partial sums (A 64 | B 128 | C 64)
};

namespace {

struct Guard {
    b200_ctx* ctx;
    template <typename F>
    int run(F&& f) {
        try {
            if (ctx) B200_CUDA_CHECK(cudaSetDevice(ctx->device));
            return f();
        } catch (const std::invalid_argument& e) {
            if (ctx) ctx->last_error = e.what();
            return B200_ERR_ARG;
        } catch (const std::domain_error& e) {
            if (ctx) ctx->last_error = e.what();
            return B200_ERR_FORMAT;
        } catch (const std::logic_error& e) {
            if (ctx) ctx->last_error = e.what();
            return B200_ERR_STATE;
        } catch (const std::exception& e) {
            if (ctx) ctx->last_error = e.what();
            return B200_ERR_CUDA;
        }
    }
};

// ---- byte-stream reader for the reference's mixed ASCII / binary key file
struct Reader {
    const uint8_t* p;
    const uint8_t* end;
    bool alt;
    void need(size_t n) const { if ((size_t)(end - p) < n) throw std::domain_error("proving key: unexpected end of data"); }
    uint64_t number() {   // ASCII decimal terminated by '\n'
        uint64_t v = 0; bool any = false;
        while (p < end && *p >= '0' && *p <= '9') { v = v * 10 + (*p - '0'); ++p; any = true; }
        if (!any || p >= end || *p != '\n') throw std::domain_error("proving key: expected a decimal count followed by newline");
        ++p;
        return v;
    }
    // reads one point into out (affine Montgomery, zeros for infinity); limbs = 4 per Fq coordinate
    void point(int group, uint8_t* out) {
        const size_t ab = group == 1 ? 64 : 128;
        const size_t raw = alt ? ab : ab + ab / 2;   // mcl stores x,y,z
        need(1 + raw);
        const uint8_t flag = *p++;
        if (flag != '0' && flag != '1') throw std::domain_error("proving key: bad is-zero flag byte");
        if (flag == '1') memset(out, 0, ab); else memcpy(out, p, ab);
        p += raw;
    }
};

template <typename T>
T* dev_upload(const std::vector<T>& v) {
    T* d = nullptr;
    B200_CUDA_CHECK(cudaMalloc(&d, std::max<size_t>(v.size() * sizeof(T), 16)));
    if (!v.empty()) B200_CUDA_CHECK(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
    return d;
}
void* dev_upload_bytes(const uint8_t* src, size_t bytes) {
    void* d = nullptr;
    B200_CUDA_CHECK(cudaMalloc(&d, std::max<size_t>(bytes, 16)));
    if (bytes) B200_CUDA_CHECK(cudaMemcpy(d, src, bytes, cudaMemcpyHostToDevice));
    return d;
}
inline void shard_range(uint64_t n, int rank, int world, uint64_t& lo, uint64_t& hi) {
    lo = n * (uint64_t)rank / (uint64_t)world;
    hi = n * (uint64_t)(rank + 1) / (uint64_t)world;
}
inline uint32_t bitrev32(uint32_t x, uint32_t bits) {
    uint32_t r = 0;
    for (uint32_t i = 0; i < bits; ++i) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}

void key_release(b200_key* k) {
    if (!k) return;
    if (!k->shards.empty()) {
        for (b200_key* sh : k->shards) {
            if (!sh) continue;
            cudaSetDevice(sh->ctx->device);
            cudaStreamSynchronize(sh->ctx->stream);
            key_release(sh);
        }
        cudaSetDevice(k->ctx->device);
        cudaFree(k->gather);
        delete k;
        return;
    }
    if (k->attached) {
        cudaFree(k->d_w); cudaFree(k->d_a); cudaFree(k->d_b); cudaFree(k->d_c); cudaFree(k->d_h);
        delete k;
        return;
    }
    if (k->owns_bases) { cudaFree(k->A); cudaFree(k->B2); cudaFree(k->B1); cudaFree(k->L); }
    cudaFree(k->H);
    cudaFree(k->A_t); cudaFree(k->B2_t); cudaFree(k->B1_t); cudaFree(k->H_t); cudaFree(k->L_t);
    cudaFree(k->A_idx); cudaFree(k->B_idx); cudaFree(k->A_pos); cudaFree(k->B_pos); cudaFree(k->consts_g1); cudaFree(k->consts_g2);
    for (int i = 0; i < 3; ++i) { csr_free(&k->M[i]); csr_free(&k->Mi[i]); }
    cudaFree(k->inst_tables);
    cudaFree(k->coeffs);
    if (k->has_circuit) domain_free(&k->dom);
    cudaFree(k->d_w); cudaFree(k->d_a); cudaFree(k->d_b); cudaFree(k->d_c); cudaFree(k->d_h);
    delete k;
}

// place this rank's slice of the H bases on the device in bit-reversed order (position p holds H_query[bitrev(p)];
// the coefficient m-1, which the reference never multiplies, maps to the point at infinity)
void place_h(b200_key* k) {
    const uint64_t m = k->m;
    uint64_t lo, hi;
    shard_range(m, k->rank, k->world, lo, hi);
    std::vector<uint8_t> buf((hi - lo) * 64, 0);
    const uint64_t have = k->H_host.size() / 64;
    for (uint64_t p = lo; p < hi; ++p) {
        const uint64_t i = bitrev32((uint32_t)p, k->logm);
        if (i < have && i + 1 < m) memcpy(&buf[(p - lo) * 64], &k->H_host[i * 64], 64);
    }
    cudaFree(k->H);
    k->H = dev_upload_bytes(buf.data(), buf.size());
    k->H_lo = lo; k->nH = hi - lo;
    k->H_bitrev = true;
}

void build_table(b200_key* k, int group, const void* pts, uint64_t n, void** table, int* c) {
    cudaFree(*table);
    *table = nullptr; *c = 0;
    if (!k->ctx->precompute || !n || !pts) return;
    const int win = k->ctx->window_override ? std::max(4, std::min(24, k->ctx->window_override)) : msm_table_window(n);
    const int W = (255 + win - 1) / win;
    const size_t esz = group == 1 ? 64 : 128;
    const size_t bytes = (size_t)W * n * esz + esz;      // + one all-zero entry (infinity) for the shared digit list
    size_t free_b = 0, total_b = 0;
    B200_CUDA_CHECK(cudaMemGetInfo(&free_b, &total_b));
    if (bytes + (8ull << 30) > free_b) return;   // does not fit beside the MSM scratch: stay on the windowed path
    B200_CUDA_CHECK(cudaMalloc(table, bytes));
    B200_CUDA_CHECK(cudaMemsetAsync((char*)*table + bytes - esz, 0, esz, k->ctx->stream));
    g_precompute(k->ctx, group, pts, n, win, *table);
    *c = win;
}
__global__ void pos_scatter_kernel(const uint32_t* __restrict__ idx, uint32_t n, uint32_t* __restrict__ pos, uint32_t nvar) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && idx[i] < nvar) pos[idx[i]] = i;
}
// inverse of a query's variable index list, for the shared digit list of the witness MSMs (msm_common.cuh: MsmShared)
__global__ void idx_not_contiguous_kernel(const uint32_t* __restrict__ idx, uint32_t n, uint32_t* __restrict__ flag) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n && idx[i] != idx[0] + i) *flag = 1u;
}
void build_pos_map(b200_key* k, const uint32_t* idx, uint64_t n, uint32_t** pos, bool* ident, uint32_t* first) {
    cudaFree(*pos);
    *pos = nullptr;
    *ident = false;
    *first = 0;
    const uint64_t nvar = k->nv + 1;
    if (!idx || !n || nvar >= (1ull << B200_SHARED_VAR_BITS)) return;
    B200_CUDA_CHECK(cudaMalloc(pos, nvar * 4));
    B200_CUDA_CHECK(cudaMemsetAsync(*pos, 0xff, nvar * 4, k->ctx->stream));
    pos_scatter_kernel<<<(unsigned)((n + 255) / 256), 256, 0, k->ctx->stream>>>(idx, (uint32_t)n, *pos, (uint32_t)nvar);
    uint32_t* flag = (uint32_t*)k->ctx->scratch("pos_flag", 4);
    B200_CUDA_CHECK(cudaMemsetAsync(flag, 0, 4, k->ctx->stream));
    idx_not_contiguous_kernel<<<(unsigned)((n + 255) / 256), 256, 0, k->ctx->stream>>>(idx, (uint32_t)n, flag);
    k->ctx->launches += 2;
    uint32_t h = 1;
    B200_CUDA_CHECK(cudaMemcpyAsync(&h, flag, 4, cudaMemcpyDeviceToHost, k->ctx->stream));
    B200_CUDA_CHECK(cudaMemcpyAsync(first, idx, 4, cudaMemcpyDeviceToHost, k->ctx->stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(k->ctx->stream));
    *ident = (h == 0);
}
void build_tables(b200_key* k) {
    build_pos_map(k, k->A_idx, k->nA, &k->A_pos, &k->A_ident, &k->A_first);
    build_pos_map(k, k->B_idx, k->nB, &k->B_pos, &k->B_ident, &k->B_first);
    build_table(k, 1, k->A, k->nA, &k->A_t, &k->cA);
    build_table(k, 2, k->B2, k->nB, &k->B2_t, &k->cB);
    if (k->full) { int cb1 = 0; build_table(k, 1, k->B1, k->nB, &k->B1_t, &cb1); if (cb1 != k->cB) { cudaFree(k->B1_t); k->B1_t = nullptr; } }
    build_table(k, 1, k->H, k->nH, &k->H_t, &k->cH);
    build_table(k, 1, k->L, k->nL, &k->L_t, &k->cL);
    B200_CUDA_CHECK(cudaStreamSynchronize(k->ctx->stream));
}


// ---- device-native key cache (B200_FMT_CACHE): the arrays exactly as they sit in HBM (affine Montgomery, (0,0) = infinity),
// so a reload is a handful of large copies instead of a point-by-point parse of the reference's mixed ASCII/binary layout.
//   "B2KC0001" | u64 full, domain_size, nA, nB, nH, nL | g1 consts 192 B | g2 consts 256 B | A_idx u32[nA] (+pad to 8 B) |
//   B_idx u32[nB] (+pad) | A 64 B x nA | B_g2 128 B x nB | B_g1 64 B x nB (full keys) | H 64 B x nH (natural order) | L 64 B x nL
const char KEY_CACHE_MAGIC[9] = "B2KC0001";
inline size_t pad8(size_t n) { return (n + 7) & ~(size_t)7; }

int key_load_cache(b200_ctx* ctx, const uint8_t* p, size_t len, int rank, int world, b200_key** out) {
    if (len < 8 + 6 * 8 + 192 + 256 || memcmp(p, KEY_CACHE_MAGIC, 8) != 0) throw std::domain_error("key cache: bad magic or truncated header");
    uint64_t hdr[6];
    memcpy(hdr, p + 8, sizeof hdr);
    const uint64_t full = hdr[0], ds = hdr[1], nA = hdr[2], nB = hdr[3], nH = hdr[4], nL = hdr[5];
    if (full > 1) throw std::domain_error("key cache: bad header");
    size_t off = 8 + sizeof hdr;
    const uint8_t* c1 = p + off; off += 192;
    const uint8_t* c2 = p + off; off += 256;
    const size_t need = off + pad8(nA * 4) + pad8(nB * 4) + nA * 64 + nB * 128 + (full ? nB * 64 : 0) + nH * 64 + nL * 64;
    if (nA > (1ull << 31) || nB > (1ull << 31) || nH > (1ull << 31) || nL > (1ull << 31) || len < need) throw std::domain_error("key cache: truncated");
    const uint32_t* a_idx = (const uint32_t*)(p + off); off += pad8(nA * 4);
    const uint32_t* b_idx = (const uint32_t*)(p + off); off += pad8(nB * 4);
    const uint8_t* A = p + off; off += nA * 64;
    const uint8_t* B2 = p + off; off += nB * 128;
    const uint8_t* B1 = p + off; off += full ? nB * 64 : 0;
    const uint8_t* H = p + off; off += nH * 64;
    const uint8_t* L = p + off;
    std::unique_ptr<b200_key, void (*)(b200_key*)> k(new b200_key(), key_release);
    k->ctx = ctx; k->rank = rank; k->world = world; k->full = full != 0; k->domain_size = ds;
    k->consts_g1 = dev_upload_bytes(c1, 192);
    k->consts_g2 = dev_upload_bytes(c2, 256);
    uint64_t lo, hi;
    shard_range(nA, rank, world, lo, hi);
    for (uint64_t i = lo; i < hi; ++i) if (a_idx[i] >= ds) throw std::domain_error("key cache: A index out of range");
    k->nA = hi - lo;
    k->A_idx = (uint32_t*)dev_upload_bytes((const uint8_t*)(a_idx + lo), (hi - lo) * 4);
    k->A = dev_upload_bytes(A + lo * 64, (hi - lo) * 64);
    shard_range(nB, rank, world, lo, hi);
    for (uint64_t i = lo; i < hi; ++i) if (b_idx[i] >= ds) throw std::domain_error("key cache: B index out of range");
    k->nB = hi - lo;
    k->B_idx = (uint32_t*)dev_upload_bytes((const uint8_t*)(b_idx + lo), (hi - lo) * 4);
    k->B2 = dev_upload_bytes(B2 + lo * 128, (hi - lo) * 128);
    if (k->full) k->B1 = dev_upload_bytes(B1 + lo * 64, (hi - lo) * 64);
    k->H_total = nH;
    k->H_host.assign(H, H + nH * 64);
    shard_range(nL, rank, world, lo, hi);
    k->L = dev_upload_bytes(L + lo * 64, (hi - lo) * 64);
    k->nL = hi - lo; k->L_lo = lo;
    *out = k.release();
    return (int)B200_OK;
}

// context of shard i of a multi-device context
inline b200_ctx* shard_ctx(b200_ctx* root, int i) { return i == 0 ? root : root->peers[i - 1]; }
inline int group_world(b200_ctx* root) { return 1 + (int)root->peers.size(); }
// runs f(i) for every device of the context, device i on its own host thread (device 0 on the caller's)
template <typename F>
void for_each_device(b200_ctx* root, F&& f) {
    const int world = group_world(root);
    for (int i = 1; i < world; ++i) root->workers[i - 1]->start([&f, i] { f(i); });
    std::exception_ptr err;
    try { B200_CUDA_CHECK(cudaSetDevice(root->device)); f(0); } catch (...) { err = std::current_exception(); }
    for (int i = 1; i < world; ++i) {
        try { root->workers[i - 1]->wait(); } catch (...) { if (!err) err = std::current_exception(); }
    }
    cudaSetDevice(root->device);
    if (err) std::rethrow_exception(err);
}
b200_key* new_group_key(b200_ctx* root) {
    b200_key* g = new b200_key();
    g->ctx = root;
    g->world = group_world(root);
    g->shards.assign(g->world, nullptr);
    return g;
}

}  // namespace

extern "C" {

static int key_load_one(b200_ctx* ctx, const void* bytes, size_t len, int fmt, int rank, int world, b200_key** out);
int b200_key_load_sharded(b200_ctx* ctx, const void* bytes, size_t len, int fmt, int rank, int world, b200_key** out);

int b200_key_load(b200_ctx* ctx, const void* bytes, size_t len, int fmt, b200_key** out) {
    if (!ctx || ctx->peers.empty()) return b200_key_load_sharded(ctx, bytes, len, fmt, 0, 1, out);
    if (!bytes || !out) return B200_ERR_ARG;
    *out = nullptr;
    // multi-device context: every device parses the byte stream and keeps its 1/N slice of every query
    return Guard{ctx}.run([&] {
        std::unique_ptr<b200_key, void (*)(b200_key*)> g(new_group_key(ctx), key_release);
        std::vector<int> rcs(g->world, 0);
        for_each_device(ctx, [&](int i) { rcs[i] = key_load_one(shard_ctx(ctx, i), bytes, len, fmt, i, g->world, &g->shards[i]); });
        for (int i = 0; i < g->world; ++i)
            if (rcs[i]) { ctx->last_error = shard_ctx(ctx, i)->last_error; return rcs[i]; }
        g->full = g->shards[0]->full; g->domain_size = g->shards[0]->domain_size; g->H_total = g->shards[0]->H_total;
        for (b200_key* sh : g->shards) { g->nA += sh->nA; g->nB += sh->nB; g->nL += sh->nL; }
        *out = g.release();
        return (int)B200_OK;
    });
}

int b200_key_load_sharded(b200_ctx* ctx, const void* bytes, size_t len, int fmt, int rank, int world, b200_key** out) {
    if (!ctx || !bytes || !out || fmt < 0 || fmt > 4 || world < 1 || rank < 0 || rank >= world) return B200_ERR_ARG;
    *out = nullptr;
    if (!ctx->peers.empty() && world != 1) { ctx->last_error = "key_load_sharded on a multi-device context: it shards keys itself (b200_key_load)"; return B200_ERR_ARG; }
    if (!ctx->peers.empty()) return b200_key_load(ctx, bytes, len, fmt, out);
    return key_load_one(ctx, bytes, len, fmt, rank, world, out);
}
static int key_load_one(b200_ctx* ctx, const void* bytes, size_t len, int fmt, int rank, int world, b200_key** out) {
    if (fmt == B200_FMT_CACHE) return Guard{ctx}.run([&] { return key_load_cache(ctx, (const uint8_t*)bytes, len, rank, world, out); });
    return Guard{ctx}.run([&] {
        std::unique_ptr<b200_key, void (*)(b200_key*)> k(new b200_key(), key_release);
        k->ctx = ctx; k->rank = rank; k->world = world;
        k->full = (fmt & 1);
        Reader rd{(const uint8_t*)bytes, (const uint8_t*)bytes + len, fmt >= 2};
        uint8_t c1[3 * 64], c2[2 * 128];
        rd.point(1, c1);            // alpha_g1
        rd.point(1, c1 + 64);       // beta_g1
        rd.point(2, c2);            // beta_g2
        rd.point(1, c1 + 128);      // delta_g1
        rd.point(2, c2 + 128);      // delta_g2
        k->consts_g1 = dev_upload_bytes(c1, sizeof c1);
        k->consts_g2 = dev_upload_bytes(c2, sizeof c2);
        // A query: sparse_vector<G1> in the _nozk key
        const uint8_t* after_consts = rd.p;
        if (!k->full) k->domain_size = rd.number();
        auto read_sparse = [&](bool is_b, std::vector<uint32_t>& idx, std::vector<uint8_t>& v1, std::vector<uint8_t>& v2) {
            const uint64_t n_idx = rd.number();
            std::vector<uint64_t> all_idx(n_idx);
            for (uint64_t i = 0; i < n_idx; ++i) all_idx[i] = rd.number();
            const uint64_t n_val = rd.number();
            if (n_val != n_idx) throw std::domain_error("proving key: sparse vector with different index and value counts");
            // kc_multiexp.tcc:54-84 — entries from the first index >= scalar_size on are ignored
            uint64_t n_use = n_idx;
            for (uint64_t i = 0; i < n_idx; ++i) if (all_idx[i] >= k->domain_size) { n_use = i; break; }
            uint64_t lo, hi;
            shard_range(n_use, rank, world, lo, hi);
            idx.resize(hi - lo);
            if (is_b) v2.resize((hi - lo) * 128);
            if (!is_b || k->full) v1.resize((hi - lo) * 64);
            uint8_t scratch[128];
            for (uint64_t i = 0; i < n_val; ++i) {
                const bool keep = (i >= lo && i < hi);
                if (is_b) {
                    rd.point(2, keep ? &v2[(i - lo) * 128] : scratch);
                    if (k->full) rd.point(1, keep ? &v1[(i - lo) * 64] : scratch);
                } else {
                    rd.point(1, keep ? &v1[(i - lo) * 64] : scratch);
                }
                if (keep) idx[i - lo] = (uint32_t)all_idx[i];
            }
        };
        if (k->full) {
            // un-stripped key: A is a dense libff::G1_vector (…zok.hpp:85, `out << pk.A_query` …zok.tcc:61) - a count and one
            // point per variable, zeros included, no domain size and no index list; its length is the domain size
            rd.p = after_consts;
            const uint64_t n = rd.number();
            k->domain_size = n;
            uint64_t lo, hi;
            shard_range(n, rank, world, lo, hi);
            std::vector<uint32_t> idx(hi - lo);
            std::vector<uint8_t> v1((hi - lo) * 64);
            uint8_t scratch[64];
            for (uint64_t i = 0; i < n; ++i) {
                const bool keep = (i >= lo && i < hi);
                rd.point(1, keep ? &v1[(i - lo) * 64] : scratch);
                if (keep) idx[i - lo] = (uint32_t)i;
            }
            k->nA = idx.size();
            k->A_idx = dev_upload(idx);
            k->A = dev_upload_bytes(v1.data(), v1.size());
        } else {
            std::vector<uint32_t> idx; std::vector<uint8_t> v1, v2;
            read_sparse(false, idx, v1, v2);
            k->nA = idx.size();
            k->A_idx = dev_upload(idx);
            k->A = dev_upload_bytes(v1.data(), v1.size());
        }
        {
            const uint64_t dsB = rd.number();
            if (dsB != k->domain_size) throw std::domain_error("proving key: A and B query domain sizes differ");
            std::vector<uint32_t> idx; std::vector<uint8_t> v1, v2;
            read_sparse(true, idx, v1, v2);
            k->nB = idx.size();
            k->B_idx = dev_upload(idx);
            k->B2 = dev_upload_bytes(v2.data(), v2.size());
            if (k->full) k->B1 = dev_upload_bytes(v1.data(), v1.size());
        }
        {   // H query: kept on the host until circuit_load fixes the domain (bit-reversed placement)
            const uint64_t n = rd.number();
            k->H_total = n;
            k->H_host.resize(n * 64);
            for (uint64_t i = 0; i < n; ++i) rd.point(1, &k->H_host[i * 64]);
        }
        {
            const uint64_t n = rd.number();
            uint64_t lo, hi;
            shard_range(n, rank, world, lo, hi);
            std::vector<uint8_t> v((hi - lo) * 64);
            uint8_t scratch[64];
            for (uint64_t i = 0; i < n; ++i) rd.point(1, (i >= lo && i < hi) ? &v[(i - lo) * 64] : scratch);
            k->L = dev_upload_bytes(v.data(), v.size());
            k->nL = hi - lo; k->L_lo = lo;
        }
        *out = k.release();
        return (int)B200_OK;
    });
}
// Key file converters (host only, no device needed): the byte-level counterparts of the reference's pk_alt2mcl and pk_mcl2nozk
// (src/export.cpp:330-408).  Any of the four reference layouts in, any out - except NOZK -> FULL, which would need the G1 half
// of the B query and the zero entries of A that the stripped key no longer holds.  FULL -> NOZK drops the B query's G1 half
// and the zero entries of the dense A query (…zok.hpp:218-230).
int b200_key_convert(const void* bytes, size_t len, int fmt_in, int fmt_out, void* out, size_t cap, size_t* len_out) {
    if (!bytes || !len_out || fmt_in < 0 || fmt_in > 3 || fmt_out < 0 || fmt_out > 3) return B200_ERR_ARG;
    const bool in_full = fmt_in & 1, out_full = fmt_out & 1, out_alt = fmt_out >= 2;
    if (!in_full && out_full) return B200_ERR_ARG;
    try {
        Reader rd{(const uint8_t*)bytes, (const uint8_t*)bytes + len, fmt_in >= 2};
        std::string o;
        o.reserve(len + len / 2);
        static const uint64_t one[4] = {0xd35d438dc58f0d9dull, 0x0a78eb28f5c70b3dull, 0x666ea36f7879462cull, 0x0e0a77c19a07df2full};   // R mod q
        auto put = [&](int group, const uint8_t* p) {       // p: affine Montgomery, all-zero = infinity
            const size_t ab = group == 1 ? 64 : 128;
            bool inf = true;
            for (size_t i = 0; i < ab; ++i) if (p[i]) { inf = false; break; }
            o.push_back(inf ? '1' : '0');
            if (out_alt) {
                if (!inf) { o.append((const char*)p, ab); return; }
                o.append(ab / 2, '\0');                                   // x = 0
                o.append((const char*)one, 32);                           // y = 1 (G2: 1 + 0 u)
                if (group == 2) o.append(32, '\0');
            } else {
                if (inf) { o.append(ab + ab / 2, '\0'); return; }         // mcl: x, y, z all zero
                o.append((const char*)p, ab);
                o.append((const char*)one, 32);                           // z = 1
                if (group == 2) o.append(32, '\0');
            }
        };
        auto num = [&](uint64_t v) { o += std::to_string(v); o.push_back('\n'); };
        uint8_t pt[128], pt1[64];
        for (int g : {1, 1, 2, 1, 2}) { rd.point(g, pt); put(g, pt); }          // alpha_g1 beta_g1 beta_g2 delta_g1 delta_g2
        // A query
        uint64_t domain = 0;
        if (in_full) {
            const uint64_t n = rd.number();
            domain = n;
            if (out_full) {
                num(n);
                for (uint64_t i = 0; i < n; ++i) { rd.point(1, pt); put(1, pt); }
            } else {
                std::vector<uint64_t> idx;
                std::string vals;
                std::swap(o, vals);                                             // collect the kept points in `vals`
                for (uint64_t i = 0; i < n; ++i) {
                    rd.point(1, pt);
                    bool inf = true;
                    for (int b = 0; b < 64; ++b) if (pt[b]) { inf = false; break; }
                    if (!inf) { idx.push_back(i); put(1, pt); }
                }
                std::swap(o, vals);
                num(n); num(idx.size());
                for (uint64_t i : idx) num(i);
                num(idx.size());
                o += vals;
            }
        } else {
            domain = rd.number();
            const uint64_t n_idx = rd.number();
            num(domain); num(n_idx);
            for (uint64_t i = 0; i < n_idx; ++i) num(rd.number());
            const uint64_t n_val = rd.number();
            if (n_val != n_idx) throw std::domain_error("proving key: sparse vector with different index and value counts");
            num(n_val);
            for (uint64_t i = 0; i < n_val; ++i) { rd.point(1, pt); put(1, pt); }
        }
        // B query: sparse in every layout
        {
            const uint64_t ds = rd.number(), n_idx = rd.number();
            num(ds); num(n_idx);
            for (uint64_t i = 0; i < n_idx; ++i) num(rd.number());
            const uint64_t n_val = rd.number();
            if (n_val != n_idx) throw std::domain_error("proving key: sparse vector with different index and value counts");
            num(n_val);
            for (uint64_t i = 0; i < n_val; ++i) {
                rd.point(2, pt); put(2, pt);
                if (in_full) { rd.point(1, pt1); if (out_full) put(1, pt1); }
            }
        }
        for (int q = 0; q < 2; ++q) {                                            // H, L
            const uint64_t n = rd.number();
            num(n);
            for (uint64_t i = 0; i < n; ++i) { rd.point(1, pt); put(1, pt); }
        }
        *len_out = o.size();
        if (out && cap >= o.size()) memcpy(out, o.data(), o.size());
        return B200_OK;
    } catch (const std::exception&) {
        return B200_ERR_FORMAT;
    }
}

int b200_key_export(b200_key* k, void* buf, size_t cap, size_t* len_out) {
    if (!k || !len_out) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (k->has_circuit || k->world != 1 || !k->owns_bases) {
        ctx->last_error = "key_export: only a whole key loaded from bytes, before circuit_load";
        return B200_ERR_STATE;
    }
    return Guard{ctx}.run([&] {
        const uint64_t nH = k->H_total;
        const size_t need = 8 + 6 * 8 + 192 + 256 + pad8(k->nA * 4) + pad8(k->nB * 4) + k->nA * 64 + k->nB * 128 +
                            (k->full ? k->nB * 64 : 0) + nH * 64 + k->nL * 64;
        *len_out = need;
        if (!buf || cap < need) return (int)B200_OK;      // size query
        uint8_t* p = (uint8_t*)buf;
        memset(p, 0, need);
        memcpy(p, KEY_CACHE_MAGIC, 8);
        const uint64_t hdr[6] = {k->full ? 1ull : 0ull, k->domain_size, k->nA, k->nB, nH, k->nL};
        memcpy(p + 8, hdr, sizeof hdr);
        size_t off = 8 + sizeof hdr;
        auto d2h = [&](const void* d, size_t bytes) { if (bytes) B200_CUDA_CHECK(cudaMemcpy(p + off, d, bytes, cudaMemcpyDeviceToHost)); };
        d2h(k->consts_g1, 192); off += 192;
        d2h(k->consts_g2, 256); off += 256;
        d2h(k->A_idx, k->nA * 4); off += pad8(k->nA * 4);
        d2h(k->B_idx, k->nB * 4); off += pad8(k->nB * 4);
        d2h(k->A, k->nA * 64); off += k->nA * 64;
        d2h(k->B2, k->nB * 128); off += k->nB * 128;
        if (k->full) { d2h(k->B1, k->nB * 64); off += k->nB * 64; }
        if (nH) memcpy(p + off, k->H_host.data(), nH * 64);
        off += nH * 64;
        d2h(k->L, k->nL * 64);
        return (int)B200_OK;
    });
}

int b200_key_from_bases(b200_ctx* ctx, b200_bases* A, const uint32_t* A_idx, b200_bases* B_g2, b200_bases* B_g1, const uint32_t* B_idx,
                        b200_bases* H, b200_bases* L, const uint64_t* g1c, const uint64_t* g2c, uint64_t num_variables, uint64_t num_inputs,
                        b200_key** out) {
    if (!ctx || !A || !B_g2 || !H || !L || !g1c || !g2c || !out) return B200_ERR_ARG;
    if (A->group != 1 || B_g2->group != 2 || H->group != 1 || L->group != 1 || (B_g1 && (B_g1->group != 1 || B_g1->n != B_g2->n))) return B200_ERR_ARG;
    if (num_inputs > num_variables || L->n > num_variables - num_inputs) return B200_ERR_ARG;
    if (!ctx->peers.empty()) {
        // multi-device context: the base sets live on the first device; device i gets a COPY of its index range of A, B and L
        // over peer memory (the shards own their copies, unlike the single-device key, which borrows the caller's sets)
        return Guard{ctx}.run([&] {
            std::unique_ptr<b200_key, void (*)(b200_key*)> g(new_group_key(ctx), key_release);
            const int world = g->world;
            std::vector<uint8_t> H_host(H->n * 64);
            if (H->n) B200_CUDA_CHECK(cudaMemcpy(H_host.data(), H->d, H->n * 64, cudaMemcpyDeviceToHost));
            B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
            for_each_device(ctx, [&](int i) {
                b200_ctx* c = shard_ctx(ctx, i);
                std::unique_ptr<b200_key, void (*)(b200_key*)> k(new b200_key(), key_release);
                k->ctx = c; k->owns_bases = true; k->full = B_g1 != nullptr; k->rank = i; k->world = world;
                k->domain_size = num_variables + 1;
                auto slice = [&](const b200_bases* b, size_t esz, uint64_t lo, uint64_t hi) {
                    void* d = nullptr;
                    B200_CUDA_CHECK(cudaMalloc(&d, std::max<size_t>((hi - lo) * esz, 16)));
                    if (hi > lo) B200_CUDA_CHECK(cudaMemcpyPeer(d, c->device, (const char*)b->d + lo * esz, ctx->device, (hi - lo) * esz));
                    return d;
                };
                auto idx_slice = [&](const uint32_t* idx, uint64_t lo, uint64_t hi) {
                    std::vector<uint32_t> v(hi - lo);
                    for (uint64_t j = lo; j < hi; ++j) {
                        v[j - lo] = idx ? idx[j] : (uint32_t)j;
                        if (v[j - lo] > num_variables) throw std::invalid_argument("key_from_bases: query index beyond num_variables");
                    }
                    return dev_upload(v);
                };
                uint64_t lo, hi;
                shard_range(A->n, i, world, lo, hi);
                k->A = slice(A, 64, lo, hi); k->nA = hi - lo; k->A_idx = idx_slice(A_idx, lo, hi);
                shard_range(B_g2->n, i, world, lo, hi);
                k->B2 = slice(B_g2, 128, lo, hi); k->nB = hi - lo; k->B_idx = idx_slice(B_idx, lo, hi);
                if (B_g1) k->B1 = slice(B_g1, 64, lo, hi);
                shard_range(L->n, i, world, lo, hi);
                k->L = slice(L, 64, lo, hi); k->nL = hi - lo; k->L_lo = lo;
                k->consts_g1 = dev_upload_bytes((const uint8_t*)g1c, 3 * 64);
                k->consts_g2 = dev_upload_bytes((const uint8_t*)g2c, 2 * 128);
                k->H_total = H->n;
                k->H_host = H_host;
                g->shards[i] = k.release();
            });
            g->full = B_g1 != nullptr; g->domain_size = num_variables + 1; g->H_total = H->n;
            g->nA = A->n; g->nB = B_g2->n; g->nL = L->n;
            *out = g.release();
            return (int)B200_OK;
        });
    }
    return Guard{ctx}.run([&] {
        std::unique_ptr<b200_key, void (*)(b200_key*)> k(new b200_key(), key_release);
        k->ctx = ctx; k->owns_bases = false; k->full = B_g1 != nullptr;
        k->domain_size = num_variables + 1;
        k->A = A->d; k->nA = A->n; k->B2 = B_g2->d; k->nB = B_g2->n; k->B1 = B_g1 ? B_g1->d : nullptr;
        k->L = L->d; k->nL = L->n;
        auto idx_or_identity = [&](const uint32_t* idx, size_t n) {
            std::vector<uint32_t> v(n);
            for (size_t i = 0; i < n; ++i) {
                v[i] = idx ? idx[i] : (uint32_t)i;
                if (v[i] > num_variables) throw std::invalid_argument("key_from_bases: query index beyond num_variables");
            }
            return dev_upload(v);
        };
        k->A_idx = idx_or_identity(A_idx, A->n);
        k->B_idx = idx_or_identity(B_idx, B_g2->n);
        k->consts_g1 = dev_upload_bytes((const uint8_t*)g1c, 3 * 64);
        k->consts_g2 = dev_upload_bytes((const uint8_t*)g2c, 2 * 128);
        k->H_total = H->n;
        k->H_host.resize(H->n * 64);
        if (H->n) B200_CUDA_CHECK(cudaMemcpy(k->H_host.data(), H->d, H->n * 64, cudaMemcpyDeviceToHost));
        *out = k.release();
        return (int)B200_OK;
    });
}

int b200_key_set_shard(b200_key* k, int rank, int world, uint64_t L_lo) {
    if (!k || world < 1 || rank < 0 || rank >= world || k->has_circuit || !k->shards.empty()) return B200_ERR_ARG;
    k->rank = rank; k->world = world; k->L_lo = L_lo;
    return B200_OK;
}

int b200_key_attach(b200_key* key, b200_ctx* ctx, b200_key** out) {
    if (!key || !ctx || !out) return B200_ERR_ARG;
    if (!key->shards.empty()) { ctx->last_error = "key_attach: not for keys on a multi-device context"; return B200_ERR_STATE; }
    if (!key->has_circuit) { ctx->last_error = "key_attach before circuit_load"; return B200_ERR_STATE; }
    if (ctx->device != key->ctx->device) { ctx->last_error = "key_attach: the context is on another device"; return B200_ERR_ARG; }
    return Guard{ctx}.run([&] {
        std::vector<uint8_t> keep_h;
        keep_h.swap(key->H_host);                           // do not copy the host copy of the H query into the slot
        std::unique_ptr<b200_key, void (*)(b200_key*)> k(nullptr, key_release);
        try { k.reset(new b200_key(*key)); } catch (...) { keep_h.swap(key->H_host); throw; }
        keep_h.swap(key->H_host);
        k->attached = true;
        k->ctx = ctx;
        k->parent = key; k->children = 0; k->free_pending = false;
        k->d_w = k->d_a = k->d_b = k->d_c = k->d_h = nullptr;
        B200_CUDA_CHECK(cudaMalloc(&k->d_w, (k->nv + 1) * 32));
        B200_CUDA_CHECK(cudaMalloc(&k->d_a, k->m * 32));
        B200_CUDA_CHECK(cudaMalloc(&k->d_b, k->m * 32));
        B200_CUDA_CHECK(cudaMalloc(&k->d_c, k->m * 32));
        B200_CUDA_CHECK(cudaMemset(k->d_a, 0, k->m * 32));
        B200_CUDA_CHECK(cudaMemset(k->d_b, 0, k->m * 32));
        B200_CUDA_CHECK(cudaMemset(k->d_c, 0, k->m * 32));
        key->children++;
        *out = k.release();
        return (int)B200_OK;
    });
}

// A key with live attached slots is not released under them: the release is deferred until the last slot is freed.
void b200_key_free(b200_key* key) {
    if (!key) return;
    cudaSetDevice(key->ctx->device);
    cudaStreamSynchronize(key->ctx->stream);
    if (key->attached) {
        b200_key* parent = key->parent;
        key_release(key);
        if (parent && --parent->children == 0 && parent->free_pending) {
            cudaSetDevice(parent->ctx->device);
            cudaStreamSynchronize(parent->ctx->stream);
            key_release(parent);
        }
        return;
    }
    if (key->children > 0) { key->free_pending = true; return; }
    key_release(key);
}
int b200_key_info(b200_key* k, uint64_t out[8]) {
    if (!k || !out) return B200_ERR_ARG;
    if (!k->shards.empty() && k->has_circuit) { k->nH = 0; for (b200_key* sh : k->shards) k->nH += sh->nH; k->H_bitrev = true; }
    out[0] = k->domain_size; out[1] = k->nA; out[2] = k->nB; out[3] = k->H_bitrev ? k->nH : k->H_total; out[4] = k->nL;
    out[5] = k->full; out[6] = k->m; out[7] = k->ni;
    return B200_OK;
}

static int circuit_load_impl(b200_key* k, const b200_csr* A, const b200_csr* B, const b200_csr* C, const b200_instances* inst,
                             const uint64_t* coeffs, size_t n_coeffs, uint32_t num_inputs, uint64_t num_variables);
int b200_circuit_load(b200_key* k, const b200_csr* A, const b200_csr* B, const b200_csr* C, const uint64_t* coeffs, size_t n_coeffs,
                      uint32_t num_inputs, uint64_t num_variables) {
    return circuit_load_impl(k, A, B, C, nullptr, coeffs, n_coeffs, num_inputs, num_variables);
}
int b200_circuit_load_instanced(b200_key* k, const b200_csr* A, const b200_csr* B, const b200_csr* C, const b200_instances* inst,
                                const uint64_t* coeffs, size_t n_coeffs, uint32_t num_inputs, uint64_t num_variables) {
    if (!inst || !inst->A || !inst->B || !inst->C || (inst->n_instances && !inst->tables) || !inst->table_len) return B200_ERR_ARG;
    if (inst->A->n_rows != inst->B->n_rows || inst->A->n_rows != inst->C->n_rows) return B200_ERR_ARG;
    return circuit_load_impl(k, A, B, C, inst, coeffs, n_coeffs, num_inputs, num_variables);
}
static int circuit_load_impl(b200_key* k, const b200_csr* A, const b200_csr* B, const b200_csr* C, const b200_instances* inst,
                             const uint64_t* coeffs, size_t n_coeffs, uint32_t num_inputs, uint64_t num_variables) {
    if (!k || !A || !B || !C || !coeffs || !n_coeffs) return B200_ERR_ARG;
    if (A->n_rows != B->n_rows || A->n_rows != C->n_rows) return B200_ERR_ARG;
    if (num_variables + 1 != k->domain_size || num_inputs > num_variables) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (!k->shards.empty()) {
        // multi-device key: every device holds the whole circuit (the witness map is small) and its slice of the queries
        return Guard{ctx}.run([&] {
            std::vector<int> rcs(k->world, 0);
            for_each_device(ctx, [&](int i) { rcs[i] = circuit_load_impl(k->shards[i], A, B, C, inst, coeffs, n_coeffs, num_inputs, num_variables); });
            for (int i = 0; i < k->world; ++i)
                if (rcs[i]) { ctx->last_error = shard_ctx(ctx, i)->last_error; return rcs[i]; }
            if (!k->gather) B200_CUDA_CHECK(cudaMalloc(&k->gather, (size_t)k->world * 256 + 1024));
            k->has_circuit = true;
            k->nc = k->shards[0]->nc; k->ni = num_inputs; k->nv = num_variables; k->m = k->shards[0]->m; k->logm = k->shards[0]->logm;
            return (int)B200_OK;
        });
    }
    return Guard{ctx}.run([&] {
        const b200_csr* mats[3] = {A, B, C};
        for (int i = 0; i < 3; ++i) {
            const uint64_t nnz = mats[i]->n_rows ? mats[i]->row_ptr[mats[i]->n_rows] : 0;
            for (uint64_t j = 0; j < nnz; ++j) {
                if (mats[i]->col[j] > num_variables) throw std::invalid_argument("circuit_load: variable index out of range");
                if (mats[i]->coeff_idx[j] >= n_coeffs) throw std::invalid_argument("circuit_load: coefficient index out of range");
            }
        }
        if (inst) {
            const b200_csr* im[3] = {inst->A, inst->B, inst->C};
            for (int i = 0; i < 3; ++i) {
                const uint64_t nnz = im[i]->n_rows ? im[i]->row_ptr[im[i]->n_rows] : 0;
                for (uint64_t j = 0; j < nnz; ++j) {
                    if (im[i]->col[j] >= inst->table_len) throw std::invalid_argument("circuit_load_instanced: master variable index beyond table_len");
                    if (im[i]->coeff_idx[j] >= n_coeffs) throw std::invalid_argument("circuit_load_instanced: coefficient index out of range");
                }
            }
            const uint64_t nt = inst->n_instances * (uint64_t)inst->table_len;
            for (uint64_t j = 0; j < nt; ++j)
                if (inst->tables[j] > num_variables) throw std::invalid_argument("circuit_load_instanced: translated variable index out of range");
        }
        if (k->has_circuit) {
            for (int i = 0; i < 3; ++i) { csr_free(&k->M[i]); csr_free(&k->Mi[i]); }
            cudaFree(k->inst_tables); k->inst_tables = nullptr; k->n_inst = 0; k->inst_tlen = 0;
            cudaFree(k->coeffs); k->coeffs = nullptr;
            domain_free(&k->dom);
            cudaFree(k->d_w); cudaFree(k->d_a); cudaFree(k->d_b); cudaFree(k->d_c); cudaFree(k->d_h);
            k->d_w = k->d_a = k->d_b = k->d_c = k->d_h = nullptr;
            k->has_circuit = false;
        }
        k->n_plain = A->n_rows;
        k->nc = A->n_rows + (inst ? inst->n_instances * inst->A->n_rows : 0); k->ni = num_inputs; k->nv = num_variables;
        // get_domain (src/stubs.cpp:61-75): m = roundUpToNearestPowerOf2(num_constraints + num_inputs + 1)
        uint64_t need = k->nc + k->ni + 1;
        uint32_t logm = 0;
        while (((uint64_t)1 << logm) < need) ++logm;
        if (logm < 1) logm = 1;
        k->logm = logm; k->m = (uint64_t)1 << logm;
        if (k->H_total + 1 < k->m) throw std::invalid_argument("circuit_load: H query shorter than domain size - 1");
        // the L MSM reads nL scalars from w[ni + 1 + L_lo ...]; the A / B gathers index w[0..nv] (…zok.tcc:487-531)
        if (k->L_lo + k->nL > k->nv - k->ni) throw std::invalid_argument("circuit_load: L query longer than num_variables - num_inputs");
        if (k->world == 1 && k->nL != k->nv - k->ni) throw std::invalid_argument("circuit_load: L query length differs from num_variables - num_inputs");
        if (k->nA > k->nv + 1 || k->nB > k->nv + 1) throw std::invalid_argument("circuit_load: A or B query longer than num_variables + 1");
        for (int i = 0; i < 3; ++i) csr_upload(ctx, mats[i]->row_ptr, mats[i]->col, mats[i]->coeff_idx, mats[i]->n_rows, &k->M[i]);
        if (inst) {
            const b200_csr* im[3] = {inst->A, inst->B, inst->C};
            for (int i = 0; i < 3; ++i) csr_upload(ctx, im[i]->row_ptr, im[i]->col, im[i]->coeff_idx, im[i]->n_rows, &k->Mi[i]);
            k->inst_tables = (uint32_t*)dev_upload_bytes((const uint8_t*)inst->tables, inst->n_instances * (size_t)inst->table_len * 4);
            k->inst_tlen = inst->table_len; k->n_inst = inst->n_instances;
        }
        k->coeffs = dev_upload_bytes((const uint8_t*)coeffs, n_coeffs * 32);
        domain_create(ctx, logm, &k->dom);
        k->has_circuit = true;
        B200_CUDA_CHECK(cudaMalloc(&k->d_w, (k->nv + 1) * 32));
        B200_CUDA_CHECK(cudaMalloc(&k->d_a, k->m * 32));
        B200_CUDA_CHECK(cudaMalloc(&k->d_b, k->m * 32));
        B200_CUDA_CHECK(cudaMalloc(&k->d_c, k->m * 32));
        place_h(k);
        build_tables(k);
        return (int)B200_OK;
    });
}

// Key generator: replaces r1cs_gg_ppzksnark_zok_generator (...zok.tcc:277-449) with the toxic waste
// supplied by the caller.  Scalars by gen.cu (Lagrange basis at t, At/Bt/Ct as transposed SpMVs, H/L/IC scalars), points by
// the fixed-base kernel.  The queries are stored dense with identity indices; entries the reference's sparse vectors skip
// (zero scalars) are points at infinity, which every MSM skips - the proofs are the same.  The key comes back in the
// pre-circuit state of b200_key_load: call b200_circuit_load next (or b200_key_export to persist it).
int b200_generate_key(b200_ctx* ctx, const b200_csr* A, const b200_csr* B, const b200_csr* C, const uint64_t* coeffs, size_t n_coeffs,
                      uint32_t num_inputs, uint64_t num_variables, const uint64_t* toxic_mont, int full, b200_key** out, uint64_t* vk_xy) {
    if (!ctx || !A || !B || !C || !coeffs || !n_coeffs || !toxic_mont || !out || !vk_xy) return B200_ERR_ARG;
    if (A->n_rows != B->n_rows || A->n_rows != C->n_rows || num_inputs > num_variables) return B200_ERR_ARG;
    *out = nullptr;
    return Guard{ctx}.run([&] {
        const uint64_t nc = A->n_rows, ni = num_inputs, nv = num_variables;
        uint32_t logm = 0;
        while (((uint64_t)1 << logm) < nc + ni + 1) ++logm;
        if (logm < 1) logm = 1;
        const uint64_t m = (uint64_t)1 << logm;
        struct Tmp {   // device temporaries, freed on every exit path
            Domain dom; DeviceCsr T[3]; std::vector<void*> bufs;
            void* alloc(size_t bytes) { void* d = nullptr; B200_CUDA_CHECK(cudaMalloc(&d, std::max<size_t>(bytes, 16))); bufs.push_back(d); return d; }
            ~Tmp() { for (void* d : bufs) cudaFree(d); for (int i = 0; i < 3; ++i) csr_free(&T[i]); domain_free(&dom); }
        } tmp;
        domain_create(ctx, logm, &tmp.dom);
        void* d_toxic = tmp.alloc(5 * 32);
        B200_CUDA_CHECK(cudaMemcpyAsync(d_toxic, toxic_mont, 5 * 32, cudaMemcpyHostToDevice, ctx->stream));
        void* d_consts = tmp.alloc(5 * 32);
        gen_consts(ctx, d_toxic, logm, d_consts);
        uint64_t z[4];
        B200_CUDA_CHECK(cudaMemcpyAsync(z, d_consts, 32, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        if (!(z[0] | z[1] | z[2] | z[3])) throw std::invalid_argument("generate_key: t lies in the evaluation domain");
        void* d_u = tmp.alloc(m * 32);
        gen_lagrange(ctx, d_toxic, d_consts, tmp.dom, d_u);
        void* d_coeffs = tmp.alloc(n_coeffs * 32);
        B200_CUDA_CHECK(cudaMemcpyAsync(d_coeffs, coeffs, n_coeffs * 32, cudaMemcpyHostToDevice, ctx->stream));
        // At, Bt, Ct: row j of the transposed matrix lists (constraint, coefficient) pairs of variable j
        const b200_csr* mats[3] = {A, B, C};
        void* d_xt[3];
        for (int k3 = 0; k3 < 3; ++k3) {
            const b200_csr* M = mats[k3];
            const uint64_t nnz = nc ? M->row_ptr[nc] : 0;
            std::vector<uint64_t> rp(nv + 2, 0);
            for (uint64_t j = 0; j < nnz; ++j) {
                if (M->col[j] > nv) throw std::invalid_argument("generate_key: variable index out of range");
                if (M->coeff_idx[j] >= n_coeffs) throw std::invalid_argument("generate_key: coefficient index out of range");
                rp[M->col[j] + 1]++;
            }
            for (uint64_t j = 0; j <= nv; ++j) rp[j + 1] += rp[j];
            std::vector<uint32_t> col(std::max<uint64_t>(nnz, 1)), ci(std::max<uint64_t>(nnz, 1));
            std::vector<uint64_t> cur(rp.begin(), rp.end() - 1);
            for (uint64_t r = 0; r < nc; ++r)
                for (uint64_t j = M->row_ptr[r]; j < M->row_ptr[r + 1]; ++j) {
                    const uint64_t pos = cur[M->col[j]]++;
                    col[pos] = (uint32_t)r; ci[pos] = M->coeff_idx[j];
                }
            csr_upload(ctx, rp.data(), col.data(), ci.data(), nv + 1, &tmp.T[k3]);
            d_xt[k3] = tmp.alloc((nv + 1) * 32);
            spmv(ctx, tmp.T[k3], d_coeffs, d_u, d_xt[k3]);
        }
        gen_add_inputs(ctx, d_xt[0], d_u, nc, ni);
        void* d_abc = tmp.alloc((nv + 1) * 32);
        gen_abc(ctx, d_toxic, d_consts, d_xt[0], d_xt[1], d_xt[2], ni, nv, d_abc);
        void* d_hs = tmp.alloc((m - 1) * 32);
        gen_h(ctx, d_toxic, d_consts, m - 1, d_hs);

        std::unique_ptr<b200_key, void (*)(b200_key*)> k(new b200_key(), key_release);
        k->ctx = ctx; k->full = full != 0; k->domain_size = nv + 1;
        k->nA = k->nB = nv + 1; k->nL = nv - ni; k->L_lo = 0; k->H_total = m - 1;
        B200_CUDA_CHECK(cudaMalloc(&k->A, (nv + 1) * 64));
        g_fixed_base(ctx, 1, d_xt[0], nv + 1, k->A);
        B200_CUDA_CHECK(cudaMalloc(&k->B2, (nv + 1) * 128));
        g_fixed_base(ctx, 2, d_xt[1], nv + 1, k->B2);
        if (k->full) {
            B200_CUDA_CHECK(cudaMalloc(&k->B1, (nv + 1) * 64));
            g_fixed_base(ctx, 1, d_xt[1], nv + 1, k->B1);
        }
        B200_CUDA_CHECK(cudaMalloc(&k->L, std::max<uint64_t>((nv - ni) * 64, 16)));
        g_fixed_base(ctx, 1, (const char*)d_abc + (ni + 1) * 32, nv - ni, k->L);
        void* d_ic = tmp.alloc((ni + 1) * 64);
        g_fixed_base(ctx, 1, d_abc, ni + 1, d_ic);
        void* d_h = tmp.alloc((m - 1) * 64);
        g_fixed_base(ctx, 1, d_hs, m - 1, d_h);
        // alpha_g1, beta_g1, delta_g1 | beta_g2, delta_g2, gamma_g2
        void* d_sc = tmp.alloc(6 * 32);
        const int g1_src[3] = {1, 2, 4}, g2_src[3] = {2, 4, 3};
        for (int i = 0; i < 3; ++i) {
            B200_CUDA_CHECK(cudaMemcpyAsync((char*)d_sc + i * 32, (const char*)d_toxic + g1_src[i] * 32, 32, cudaMemcpyDeviceToDevice, ctx->stream));
            B200_CUDA_CHECK(cudaMemcpyAsync((char*)d_sc + (3 + i) * 32, (const char*)d_toxic + g2_src[i] * 32, 32, cudaMemcpyDeviceToDevice, ctx->stream));
        }
        B200_CUDA_CHECK(cudaMalloc(&k->consts_g1, 3 * 64));
        g_fixed_base(ctx, 1, d_sc, 3, k->consts_g1);
        void* d_g2c = tmp.alloc(3 * 128);
        g_fixed_base(ctx, 2, (const char*)d_sc + 3 * 32, 3, d_g2c);
        B200_CUDA_CHECK(cudaMalloc(&k->consts_g2, 2 * 128));
        B200_CUDA_CHECK(cudaMemcpyAsync(k->consts_g2, d_g2c, 2 * 128, cudaMemcpyDeviceToDevice, ctx->stream));
        std::vector<uint32_t> ident(nv + 1);
        for (uint64_t i = 0; i <= nv; ++i) ident[i] = (uint32_t)i;
        k->A_idx = dev_upload(ident);
        k->B_idx = dev_upload(ident);
        k->H_host.resize((m - 1) * 64);
        if (m > 1) B200_CUDA_CHECK(cudaMemcpyAsync(k->H_host.data(), d_h, (m - 1) * 64, cudaMemcpyDeviceToHost, ctx->stream));
        // verification key pieces: alpha_g1 | beta_g2 | gamma_g2 | delta_g2 | IC[0..ni]   (affine Montgomery)
        B200_CUDA_CHECK(cudaMemcpyAsync(vk_xy, k->consts_g1, 64, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaMemcpyAsync(vk_xy + 8, d_g2c, 128, cudaMemcpyDeviceToHost, ctx->stream));                       // beta_g2
        B200_CUDA_CHECK(cudaMemcpyAsync(vk_xy + 24, (const char*)d_g2c + 256, 128, cudaMemcpyDeviceToHost, ctx->stream));   // gamma_g2
        B200_CUDA_CHECK(cudaMemcpyAsync(vk_xy + 40, (const char*)d_g2c + 128, 128, cudaMemcpyDeviceToHost, ctx->stream));   // delta_g2
        B200_CUDA_CHECK(cudaMemcpyAsync(vk_xy + 56, d_ic, (ni + 1) * 64, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        *out = k.release();
        return (int)B200_OK;
    });
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------ prove
namespace {

void run_witness_eval(b200_key* k, const uint64_t* witness, int chains = 7) {
    b200_ctx* ctx = k->ctx;
    ctx->t_begin("h2d.witness");
    if (witness) B200_CUDA_CHECK(cudaMemcpyAsync(k->d_w, witness, (k->nv + 1) * 32, cudaMemcpyDefault, ctx->stream));
    B200_CUDA_CHECK(cudaEventRecord(ctx->ev_witness, ctx->stream));
    ctx->t_end();
    ctx->t_begin("witness_eval");
    void* outs[3] = {k->d_a, k->d_b, k->d_c};
    for (int i = 0; i < 3; ++i) {
        if (!((chains >> i) & 1)) continue;
        spmv(ctx, k->M[i], k->coeffs, k->d_w, outs[i]);
        if (k->n_inst) spmv_instances(ctx, k->Mi[i], k->coeffs, k->d_w, k->inst_tables, k->inst_tlen, k->n_inst, (char*)outs[i] + k->n_plain * 32);
    }
    witness_rows(ctx, k->d_w, k->d_a, k->d_b, k->d_c, k->nc, k->ni, k->m);
    ctx->t_end();
}
void run_h(b200_key* k) {
    b200_ctx* ctx = k->ctx;
    ctx->t_begin("h_chain");
    h_chain(ctx, k->dom, k->d_a, k->d_b, k->d_c);
    ctx->t_end();
}
// the four (five) MSMs over this rank's key slice; results as XYZZ in acc[0..4]: A, B(G2), H, L, B(G1)
struct Acc { void *A, *B2, *H, *L, *B1, *tmp, *fin; };
Acc acc_slots(b200_ctx* ctx) {
    char* base = (char*)ctx->scratch("prove_acc", 8192);
    Acc a;
    a.A = base; a.H = base + 128; a.L = base + 256; a.B1 = base + 384; a.B2 = base + 512; a.tmp = base + 768; a.fin = base + 4096;
    return a;
}
// The A, B(G2), L (and B(G1)) MSMs need only the witness: they are queued on their own lanes right after the witness
// upload and run concurrently with the witness map on the main stream; the H MSM follows the H chain on the main stream.
// finish: 0 = leave the sums in XYZZ; 1 = the no-ZK prover's g_A = alpha + sum A and g_B = beta + sum B (standard form)
// right on the lane that produced the sum, so only g_C is left for the join; 2 = plain affine partial sums (sharded key).
void run_msms_witness_only(b200_key* k, const Acc& a, bool want_b1, int finish = 0) {
    b200_ctx* ctx = k->ctx;
    char* fin = (char*)a.fin;
    const bool ov = ctx->overlap != 0;
    auto lane_begin = [&](int lane, const char* name) {
        ctx->lane = ov ? lane : 0;
        if (ctx->lane) B200_CUDA_CHECK(cudaStreamWaitEvent(ctx->lanes[lane], ctx->ev_witness, 0));
        ctx->t_begin(name);
    };
    auto lane_end = [&]() {
        ctx->t_end();
        if (ctx->lane) B200_CUDA_CHECK(cudaEventRecord(ctx->lane_done[ctx->lane], ctx->lanes[ctx->lane]));
        ctx->lane = 0;
    };
    // One sorted digit list for all of them when every query is dense over the variables and the tables share a window
    // size: digits + sort + offsets run once (on the G2 lane) instead of three or four times.
    // The list covers the variable range [v0, v1) the queries of this key (or of this rank's shard of it) touch; a query is
    // addressed through its variable -> entry map, or through an index range when its index list is contiguous.
    uint64_t v0 = 0, v1 = k->nv + 1;
    if (k->world != 1 && k->A_ident && k->B_ident && k->nA && k->nB && k->nL) {
        const uint64_t l0 = k->ni + 1 + k->L_lo;
        v0 = std::min<uint64_t>(std::min<uint64_t>(k->A_first, k->B_first), l0);
        v1 = std::max<uint64_t>(std::max<uint64_t>(k->A_first + k->nA, k->B_first + k->nB), l0 + k->nL);
    }
    const uint64_t nvar = v1 - v0;
    const bool share = ctx->shared_sort && (k->world == 1 || (k->A_ident && k->B_ident)) && k->A_t && k->B2_t && k->L_t &&
                       k->A_pos && k->B_pos && k->cA == k->cB && k->cA == k->cL && k->cA >= 8 /* <= 32 windows */ && (!want_b1 || k->B1_t) &&
                       k->nv + 1 < (1ull << B200_SHARED_VAR_BITS) && 5 * k->nA >= 4 * nvar && 5 * k->nB >= 4 * nvar && 5 * k->nL >= 4 * nvar;
    if (share) {
        try {
            MsmShared sh;
            lane_begin(1, "msm.B_g2");
            ctx->t_begin("msm.shared_sort");
            msm_shared_prepare(ctx, (const char*)k->d_w + v0 * 32, nvar, k->cA, &sh);
            ctx->t_end();
            if (ctx->lane) B200_CUDA_CHECK(cudaEventRecord(ctx->ev_shared, ctx->lanes[ctx->lane]));
            const MsmPos posA{k->A_ident ? nullptr : k->A_pos + v0, (uint32_t)(k->A_first - v0), (uint32_t)k->nA, (uint32_t)k->nA};
            const MsmPos posB{k->B_ident ? nullptr : k->B_pos + v0, (uint32_t)(k->B_first - v0), (uint32_t)k->nB, (uint32_t)k->nB};
            msm_shared_consume_g2(ctx, sh, k->B2_t, posB, a.B2);
            if (finish == 1) g_finish(ctx, 2, a.B2, 1, k->consts_g2, 1, fin + 256, true);
            if (finish == 2) g_finish(ctx, 2, a.B2, 1, nullptr, 0, fin + 256, false);
            lane_end();
            lane_begin(2, "msm.A");
            if (ctx->lane) B200_CUDA_CHECK(cudaStreamWaitEvent(ctx->lanes[ctx->lane], ctx->ev_shared, 0));
            msm_shared_consume_g1(ctx, sh, k->A_t, posA, a.A);
            if (want_b1) msm_shared_consume_g1(ctx, sh, k->B1_t, posB, a.B1);
            if (finish == 1) g_finish(ctx, 1, a.A, 1, k->consts_g1, 1, fin, true);
            if (finish == 2) g_finish(ctx, 1, a.A, 1, nullptr, 0, fin, false);
            lane_end();
            lane_begin(3, "msm.L");
            if (ctx->lane) B200_CUDA_CHECK(cudaStreamWaitEvent(ctx->lanes[ctx->lane], ctx->ev_shared, 0));
            msm_shared_consume_g1(ctx, sh, k->L_t, MsmPos{nullptr, (uint32_t)(k->ni + 1 + k->L_lo - v0), (uint32_t)k->nL, (uint32_t)k->nL}, a.L);
            lane_end();
        } catch (...) { ctx->lane = 0; throw; }
        return;
    }
    try {
        lane_begin(1, "msm.B_g2");
        msm_run(ctx, 2, k->B2_t ? k->B2_t : k->B2, k->d_w, k->B_idx, k->nB, true, a.B2, k->B2_t ? k->cB : 0, k->nB);
        if (finish == 1) g_finish(ctx, 2, a.B2, 1, k->consts_g2, 1, fin + 256, true);
        if (finish == 2) g_finish(ctx, 2, a.B2, 1, nullptr, 0, fin + 256, false);
        lane_end();
        lane_begin(2, "msm.A");
        msm_run(ctx, 1, k->A_t ? k->A_t : k->A, k->d_w, k->A_idx, k->nA, true, a.A, k->A_t ? k->cA : 0, k->nA);
        if (want_b1) msm_run(ctx, 1, k->B1_t ? k->B1_t : k->B1, k->d_w, k->B_idx, k->nB, true, a.B1, k->B1_t ? k->cB : 0, k->nB);
        if (finish == 1) g_finish(ctx, 1, a.A, 1, k->consts_g1, 1, fin, true);
        if (finish == 2) g_finish(ctx, 1, a.A, 1, nullptr, 0, fin, false);
        lane_end();
        lane_begin(3, "msm.L");
        msm_run(ctx, 1, k->L_t ? k->L_t : k->L, (const char*)k->d_w + (k->ni + 1 + k->L_lo) * 32, nullptr, k->nL, true, a.L, k->L_t ? k->cL : 0, k->nL);
        lane_end();
    } catch (...) { ctx->lane = 0; throw; }
}
void run_msm_h_and_join(b200_key* k, const Acc& a, const void* d_hcoef = nullptr) {
    b200_ctx* ctx = k->ctx;
    if (!d_hcoef) d_hcoef = k->d_c;
    ctx->t_begin("msm.H");
    msm_run(ctx, 1, k->H_t ? k->H_t : k->H, (const char*)d_hcoef + k->H_lo * 32, nullptr, k->nH, true, a.H, k->H_t ? k->cH : 0, k->nH);
    ctx->t_end();
    if (ctx->overlap)
        for (int lane = 1; lane < 4; ++lane) B200_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, ctx->lane_done[lane], 0));
}
// the three finished points sit 256 B apart in one scratch block (A at 0, B at 256, C at 512, each followed by its u32
// infinity flag): one copy, one synchronisation
void fetch3(b200_ctx* ctx, const void* d_fin, uint64_t* A_xy, uint64_t* B_xy, uint64_t* C_xy, uint8_t* flags) {
    uint8_t buf[768];
    B200_CUDA_CHECK(cudaMemcpyAsync(buf, d_fin, 512 + 64 + 4, cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    uint32_t f;
    memcpy(A_xy, buf, 64);        memcpy(&f, buf + 64, 4);        flags[0] = (uint8_t)(f != 0);
    memcpy(B_xy, buf + 256, 128); memcpy(&f, buf + 256 + 128, 4); flags[1] = (uint8_t)(f != 0);
    memcpy(C_xy, buf + 512, 64);  memcpy(&f, buf + 512 + 64, 4);  flags[2] = (uint8_t)(f != 0);
}
// queues one rank's partial proof: on return (no synchronisation) the three affine Montgomery partial sums are being written
// to fin + 0 / + 256 / + 512 of the context's scratch.  witness == nullptr: the witness is already in k->d_w.
char* partial_async(b200_key* k, const uint64_t* witness) {
    b200_ctx* ctx = k->ctx;
    ctx->t_begin("prove.total");
    Acc a = acc_slots(ctx);
    run_witness_eval(k, witness);
    run_msms_witness_only(k, a, false, 2);
    run_h(k);
    run_msm_h_and_join(k, a);
    char* fin = (char*)a.fin;
    g_finish(ctx, 1, a.H, 2, nullptr, 0, fin + 512, false);
    ctx->t_end();
    return fin;
}
// The same in two stages, for a proof spread over several GPUs (or ranks) that share the transform work: stage 1 uploads the
// witness, queues the witness-only MSMs on their lanes, and takes only the evaluation vectors named in `chains` (bit 0 = A,
// 1 = B, 2 = C) through the witness map and the first two transforms; the caller then makes all three vectors of coset
// evaluations present on every device (broadcast from their owners: k->d_a / d_b / d_c); stage 2 is the pointwise step, the
// final transform, the H MSM and the finishes.
void partial_stage1(b200_key* k, const uint64_t* witness, int chains) {
    b200_ctx* ctx = k->ctx;
    ctx->t_begin("prove.total");
    Acc a = acc_slots(ctx);
    run_witness_eval(k, witness, chains);
    run_msms_witness_only(k, a, false, 2);
    void* bufs[3];
    int n = 0;
    if (chains & 1) bufs[n++] = k->d_a;
    if (chains & 2) bufs[n++] = k->d_b;
    if (chains & 4) bufs[n++] = k->d_c;
    ctx->t_begin("h_chain.front");
    h_chain_front(ctx, k->dom, bufs, n);
    ctx->t_end();
}
char* partial_stage2(b200_key* k) {
    b200_ctx* ctx = k->ctx;
    Acc a = acc_slots(ctx);
    if (!k->d_h) B200_CUDA_CHECK(cudaMalloc(&k->d_h, k->m * 32));
    ctx->t_begin("h_chain.back");
    h_chain_back(ctx, k->dom, k->d_a, k->d_b, k->d_c, k->d_h);
    ctx->t_end();
    run_msm_h_and_join(k, a, k->d_h);
    char* fin = (char*)a.fin;
    g_finish(ctx, 1, a.H, 2, nullptr, 0, fin + 512, false);
    ctx->t_end();
    return fin;
}
// which rank / device takes evaluation vector ch (0 = A, 1 = B, 2 = C) through the first two transforms
inline int chain_owner(int ch, int world) { return world >= 3 ? ch : (world == 2 ? (ch == 1 ? 1 : 0) : 0); }
inline int chains_of(int rank, int world) {
    int m = 0;
    for (int ch = 0; ch < 3; ++ch) if (chain_owner(ch, world) == rank) m |= 1 << ch;
    return m;
}
// A 64 B | B 128 B | C 64 B (the point at infinity is (0,0)) to dst, which may be on another device (peer memory)
void pack_partial(b200_ctx* ctx, const char* fin, void* dst) {
    B200_CUDA_CHECK(cudaMemcpyAsync(dst, fin, 64, cudaMemcpyDefault, ctx->stream));
    B200_CUDA_CHECK(cudaMemcpyAsync((char*)dst + 64, fin + 256, 128, cudaMemcpyDefault, ctx->stream));
    B200_CUDA_CHECK(cudaMemcpyAsync((char*)dst + 192, fin + 512, 64, cudaMemcpyDefault, ctx->stream));
}
// sums `world` packed partials (device memory of ctx's device), adds alpha_g1 / beta_g2, leaves the proof in fin (standard form)
char* combine_async(b200_key* k, const void* d_parts, int world) {
    b200_ctx* ctx = k->ctx;
    char* fin = (char*)acc_slots(ctx).fin + 2048;
    g_combine3(ctx, d_parts, world, k->consts_g1, k->consts_g2, fin);
    return fin;
}

}  // namespace

extern "C" {

int b200_witness_eval(b200_key* k, const uint64_t* witness, uint64_t* a_out, uint64_t* b_out, uint64_t* c_out) {
    if (!k || !witness || !a_out || !b_out || !c_out) return B200_ERR_ARG;
    if (!k->has_circuit) { k->ctx->last_error = "witness_eval before circuit_load"; return B200_ERR_STATE; }
    if (!k->shards.empty()) return b200_witness_eval(k->shards[0], witness, a_out, b_out, c_out);
    b200_ctx* ctx = k->ctx;
    return Guard{ctx}.run([&] {
        run_witness_eval(k, witness);
        B200_CUDA_CHECK(cudaMemcpyAsync(a_out, k->d_a, k->m * 32, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaMemcpyAsync(b_out, k->d_b, k->m * 32, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaMemcpyAsync(c_out, k->d_c, k->m * 32, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        ctx->t_collect();
        return (int)B200_OK;
    });
}

int b200_h(b200_key* k, const uint64_t* witness, uint64_t* h_out) {
    if (!k || !witness || !h_out) return B200_ERR_ARG;
    if (!k->has_circuit) { k->ctx->last_error = "h before circuit_load"; return B200_ERR_STATE; }
    if (!k->shards.empty()) return b200_h(k->shards[0], witness, h_out);
    b200_ctx* ctx = k->ctx;
    return Guard{ctx}.run([&] {
        run_witness_eval(k, witness);
        run_h(k);
        bitrev_permute(ctx, k->d_c, k->d_a, k->logm, 32);   // bit-reversed -> natural order for the caller
        B200_CUDA_CHECK(cudaMemcpyAsync(h_out, k->d_a, k->m * 32, cudaMemcpyDeviceToHost, ctx->stream));
        B200_CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        ctx->t_collect();
        return (int)B200_OK;
    });
}

int b200_prove(b200_key* k, const uint64_t* witness, const uint64_t r[4], const uint64_t s[4], uint64_t A_xy[8], uint64_t B_xy[16],
               uint64_t C_xy[8], uint8_t flags[3]) {
    if (!k || !witness || !A_xy || !B_xy || !C_xy || !flags || ((r == nullptr) != (s == nullptr))) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (!k->has_circuit) { ctx->last_error = "prove before circuit_load"; return B200_ERR_STATE; }
    if (!k->shards.empty()) {
        if (r) { ctx->last_error = "r,s proving is not available on a multi-device context"; return B200_ERR_STATE; }
        // One proof on all devices of the context.  The witness goes to the first device once; every other device pulls it
        // over peer memory, runs the whole pipeline on its 1/N slice of the queries (its launch sequence issued by its own
        // host thread) and pushes its three affine partial sums into the first device's gather buffer; the first device adds
        // them and alpha_g1 / beta_g2.  No host round trip between the witness upload and the proof download.
        return Guard{ctx}.run([&] {
            b200_key* k0 = k->shards[0];
            B200_CUDA_CHECK(cudaMemcpyAsync(k0->d_w, witness, (k->nv + 1) * 32, cudaMemcpyDefault, ctx->stream));
            cudaEvent_t ev_w = ctx->ev_group;
            B200_CUDA_CHECK(cudaEventRecord(ev_w, ctx->stream));
            const int world = k->world;
            const size_t vec_bytes = (size_t)k->m * 32;
            // stage 1 everywhere: MSM lanes + this device's share of the three transform chains; "front done" events
            for_each_device(ctx, [&](int i) {
                b200_key* ki = k->shards[i];
                b200_ctx* ci = ki->ctx;
                if (i) B200_CUDA_CHECK(cudaStreamWaitEvent(ci->stream, ev_w, 0));
                partial_stage1(ki, i ? (const uint64_t*)k0->d_w : nullptr, ctx->split_chains ? chains_of(i, world) : 7);
                B200_CUDA_CHECK(cudaEventRecord(ci->ev_front, ci->stream));
            });
            // every device pulls the coset evaluations it did not compute from their owners over peer memory, then stage 2
            for_each_device(ctx, [&](int i) {
                b200_key* ki = k->shards[i];
                b200_ctx* ci = ki->ctx;
                if (ctx->split_chains) {
                    void* mine[3] = {ki->d_a, ki->d_b, ki->d_c};
                    for (int ch = 0; ch < 3; ++ch) {
                        const int o = chain_owner(ch, world);
                        if (o == i) continue;
                        b200_key* ko = k->shards[o];
                        void* theirs[3] = {ko->d_a, ko->d_b, ko->d_c};
                        B200_CUDA_CHECK(cudaStreamWaitEvent(ci->stream, ko->ctx->ev_front, 0));
                        B200_CUDA_CHECK(cudaMemcpyPeerAsync(mine[ch], ci->device, theirs[ch], ko->ctx->device, vec_bytes, ci->stream));
                    }
                }
                char* fin = partial_stage2(ki);
                pack_partial(ci, fin, (char*)k->gather + (size_t)i * 256);
                if (i) B200_CUDA_CHECK(cudaEventRecord(ci->ev_group, ci->stream));
            });
            for (int i = 1; i < k->world; ++i) B200_CUDA_CHECK(cudaStreamWaitEvent(ctx->stream, k->shards[i]->ctx->ev_group, 0));
            ctx->t_begin("combine");
            char* fin = combine_async(k0, k->gather, k->world);
            ctx->t_end();
            fetch3(ctx, fin, A_xy, B_xy, C_xy, flags);
            for (int i = 0; i < k->world; ++i) { cudaSetDevice(k->shards[i]->ctx->device); k->shards[i]->ctx->t_collect(); }
            cudaSetDevice(ctx->device);
            return (int)B200_OK;
        });
    }
    if (k->world != 1) { ctx->last_error = "prove on a sharded key: use prove_partial + proof_combine"; return B200_ERR_STATE; }
    if (r && !k->full) { ctx->last_error = "r,s proving needs a FULL key (B query in G1)"; return B200_ERR_STATE; }
    return Guard{ctx}.run([&] {
        ctx->t_begin("prove.total");
        Acc a = acc_slots(ctx);
        run_witness_eval(k, witness);
        run_msms_witness_only(k, a, r != nullptr, r ? 0 : 1);
        run_h(k);
        run_msm_h_and_join(k, a);
        ctx->t_begin("finish");
        char* fin = (char*)a.fin;
        const char* c1 = (const char*)k->consts_g1;   // alpha, beta, delta (64 B each)
        const char* c2 = (const char*)k->consts_g2;   // beta, delta (128 B each)
        if (!r) {
            // g_A = alpha_g1 + evaluation_At and g_B = beta_g2 + evaluation_Bt were finished on their lanes
            g_finish(ctx, 1, a.H, 2, nullptr, 0, fin + 512, true);       // g_C = evaluation_Ht + evaluation_Lt (H, L adjacent)
        } else {
            // scalars r, s, -(r s) on the device
            char* sc = (char*)ctx->scratch("prove_rs", 4 * 32);
            B200_CUDA_CHECK(cudaMemcpyAsync(sc, r, 32, cudaMemcpyHostToDevice, ctx->stream));
            B200_CUDA_CHECK(cudaMemcpyAsync(sc + 32, s, 32, cudaMemcpyHostToDevice, ctx->stream));
            field_op(ctx, 0, B200_OP_MUL, sc, sc + 32, sc + 64, 1);
            field_op(ctx, 0, B200_OP_NEG, sc + 64, nullptr, sc + 96, 1);
            char* t = (char*)a.tmp;   // XYZZ temporaries, 256 B apart
            // g1_A = alpha + sum A + r delta1   (Montgomery affine kept for the s*g1_A term)
            g_scalar_mul(ctx, 1, c1 + 128, sc, t);
            B200_CUDA_CHECK(cudaMemcpyAsync(t + 128, a.A, 128, cudaMemcpyDeviceToDevice, ctx->stream));
            g_finish(ctx, 1, t, 2, c1, 1, fin + 1024, false);
            g_finish(ctx, 1, t, 2, c1, 1, fin, true);
            // g2_B = beta2 + sum B2 + s delta2
            g_scalar_mul(ctx, 2, c2 + 128, sc + 32, t + 256);
            B200_CUDA_CHECK(cudaMemcpyAsync(t + 512, a.B2, 256, cudaMemcpyDeviceToDevice, ctx->stream));
            g_finish(ctx, 2, t + 256, 2, c2, 1, fin + 256, true);
            // g1_B = beta1 + sum B1 + s delta1
            g_scalar_mul(ctx, 1, c1 + 128, sc + 32, t);
            B200_CUDA_CHECK(cudaMemcpyAsync(t + 128, a.B1, 128, cudaMemcpyDeviceToDevice, ctx->stream));
            g_finish(ctx, 1, t, 2, c1 + 64, 1, fin + 1280, false);
            // g1_C = sum H + sum L + s g1_A + r g1_B - r s delta1
            char* u = t + 768;
            B200_CUDA_CHECK(cudaMemcpyAsync(u, a.H, 256, cudaMemcpyDeviceToDevice, ctx->stream));   // H and L
            g_scalar_mul(ctx, 1, fin + 1024, sc + 32, u + 256);
            g_scalar_mul(ctx, 1, fin + 1280, sc, u + 384);
            g_scalar_mul(ctx, 1, c1 + 128, sc + 96, u + 512);
            g_finish(ctx, 1, u, 5, nullptr, 0, fin + 512, true);
        }
        ctx->t_end();
        ctx->t_end();   // prove.total
        fetch3(ctx, fin, A_xy, B_xy, C_xy, flags);
        ctx->t_collect();
        return (int)B200_OK;
    });
}

int b200_prove_partial(b200_key* k, const uint64_t* witness, uint64_t partial[32], uint8_t flags[3]) {
    if (!k || !witness || !partial || !flags) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (!k->has_circuit) { ctx->last_error = "prove_partial before circuit_load"; return B200_ERR_STATE; }
    if (!k->shards.empty()) { ctx->last_error = "prove_partial on a multi-device context: b200_prove runs all devices"; return B200_ERR_STATE; }
    return Guard{ctx}.run([&] {
        char* fin = partial_async(k, witness);
        fetch3(ctx, fin, partial, partial + 8, partial + 24, flags);
        ctx->t_collect();
        return (int)B200_OK;
    });
}

// The same with the result left on the device: 256 bytes at d_out (A 64 | B 128 | C 64, affine Montgomery, (0,0) = infinity),
// queued on b200_stream(ctx) WITHOUT synchronising, so that a collective on the same stream (ncclAllGather) can follow.
int b200_prove_partial_dev(b200_key* k, const uint64_t* witness, void* d_out) {
    if (!k || !witness || !d_out) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (!k->has_circuit) { ctx->last_error = "prove_partial before circuit_load"; return B200_ERR_STATE; }
    if (!k->shards.empty()) { ctx->last_error = "prove_partial on a multi-device context: b200_prove runs all devices"; return B200_ERR_STATE; }
    return Guard{ctx}.run([&] {
        char* fin = partial_async(k, witness);
        pack_partial(ctx, fin, d_out);
        return (int)B200_OK;
    });
}
// The staged form for ranks that share the transform work (see partial_stage1): stage 1 with this rank's share of the three
// evaluation vectors (chains: bit 0 = A, 1 = B, 2 = C; b200_chain_owner says which rank takes which), then the caller
// broadcasts each vector of coset evaluations from its owner to all ranks (the device buffers come from b200_key_buffers;
// ncclBroadcast on b200_stream(ctx)), then stage 2, which leaves the same 256-byte record as b200_prove_partial_dev.
int b200_chain_owner(int chain, int world) { return (chain < 0 || chain > 2 || world < 1) ? -1 : chain_owner(chain, world); }
int b200_key_buffers(b200_key* k, void* bufs[3], size_t* bytes_each) {
    if (!k || !bufs || !bytes_each || !k->has_circuit || !k->shards.empty()) return B200_ERR_ARG;
    bufs[0] = k->d_a; bufs[1] = k->d_b; bufs[2] = k->d_c;
    *bytes_each = (size_t)k->m * 32;
    return B200_OK;
}
int b200_prove_partial_stage1(b200_key* k, const uint64_t* witness, int chains) {
    if (!k || !witness || chains < 0 || chains > 7) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (!k->has_circuit || !k->shards.empty()) { ctx->last_error = "prove_partial_stage1: needs a single-device key with a circuit"; return B200_ERR_STATE; }
    return Guard{ctx}.run([&] { partial_stage1(k, witness, chains); return (int)B200_OK; });
}
int b200_prove_partial_stage2(b200_key* k, void* d_out) {
    if (!k || !d_out) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (!k->has_circuit || !k->shards.empty()) return B200_ERR_STATE;
    return Guard{ctx}.run([&] {
        char* fin = partial_stage2(k);
        pack_partial(ctx, fin, d_out);
        return (int)B200_OK;
    });
}
// combines `world` packed partials that are already on this device (the all-gathered buffer) into the proof: three small
// kernels on b200_stream(ctx), one copy back
int b200_proof_combine_dev(b200_key* k, const void* d_parts, int world, uint64_t A_xy[8], uint64_t B_xy[16], uint64_t C_xy[8], uint8_t flags[3]) {
    if (!k || !d_parts || world < 1 || !A_xy || !B_xy || !C_xy || !flags) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    if (!k->shards.empty()) return B200_ERR_STATE;
    return Guard{ctx}.run([&] {
        ctx->t_begin("combine");
        char* fin = combine_async(k, d_parts, world);
        ctx->t_end();
        fetch3(ctx, fin, A_xy, B_xy, C_xy, flags);
        ctx->t_collect();
        return (int)B200_OK;
    });
}

int b200_proof_combine(b200_key* k, const uint64_t* partials, const uint8_t* pflags, int world, uint64_t A_xy[8], uint64_t B_xy[16],
                       uint64_t C_xy[8], uint8_t flags[3]) {
    if (!k || !partials || !pflags || world < 1 || !A_xy || !B_xy || !C_xy || !flags) return B200_ERR_ARG;
    b200_ctx* ctx = k->ctx;
    return Guard{ctx}.run([&] {
        std::vector<uint8_t> a1((size_t)(world + 1) * 64, 0), b2((size_t)(world + 1) * 128, 0), c1((size_t)world * 64, 0);
        for (int i = 0; i < world; ++i) {
            const uint64_t* p = partials + (size_t)i * 32;
            if (!pflags[i * 3 + 0]) memcpy(&a1[(size_t)i * 64], p, 64);
            if (!pflags[i * 3 + 1]) memcpy(&b2[(size_t)i * 128], p + 8, 128);
            if (!pflags[i * 3 + 2]) memcpy(&c1[(size_t)i * 64], p + 24, 64);
        }
        char* buf = (char*)ctx->scratch("combine", (size_t)(world + 1) * 256 + 1024);
        char* fin = buf + (size_t)(world + 1) * 256;      // A terms | B terms | C terms | fin (3 x 256 B)
        // alpha_g1 / beta_g2 are appended as one more affine term
        B200_CUDA_CHECK(cudaMemcpyAsync(buf, a1.data(), (size_t)world * 64, cudaMemcpyHostToDevice, ctx->stream));
        B200_CUDA_CHECK(cudaMemcpyAsync(buf + (size_t)world * 64, k->consts_g1, 64, cudaMemcpyDeviceToDevice, ctx->stream));
        g_finish(ctx, 1, nullptr, 0, buf, world + 1, fin, true);
        char* bufB = buf + (size_t)(world + 1) * 64;
        B200_CUDA_CHECK(cudaMemcpyAsync(bufB, b2.data(), (size_t)world * 128, cudaMemcpyHostToDevice, ctx->stream));
        B200_CUDA_CHECK(cudaMemcpyAsync(bufB + (size_t)world * 128, k->consts_g2, 128, cudaMemcpyDeviceToDevice, ctx->stream));
        g_finish(ctx, 2, nullptr, 0, bufB, world + 1, fin + 256, true);
        char* bufC = bufB + (size_t)(world + 1) * 128;
        B200_CUDA_CHECK(cudaMemcpyAsync(bufC, c1.data(), (size_t)world * 64, cudaMemcpyHostToDevice, ctx->stream));
        g_finish(ctx, 1, nullptr, 0, bufC, world, fin + 512, true);
        fetch3(ctx, fin, A_xy, B_xy, C_xy, flags);
        return (int)B200_OK;
    });
}

}  // extern "C"
