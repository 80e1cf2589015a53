// ethsnarks_b200.cpp — the reference's prover entry points on libb200snark.so (see ethsnarks_b200.hpp for the symbol map).
// Host-only C++, compiled with the reference's own flags against the reference's own headers; no CUDA, no torch types.
#include "ethsnarks_b200.hpp"
#include "b200snark.h"
#include "utils.hpp"          // the reference's loadFromFile / writeToFile (src/utils.hpp:166-185)
#include "export.hpp"         // the reference's proof_to_json / vk2json_file declarations (src/export.hpp)
#include "import.hpp"         // the reference's vk_from_json / proof_from_json declarations (src/import.hpp), for stub_verify
#include <libfqfft/evaluation_domain/domains/basic_radix2_domain.hpp>
#include <libfqfft/evaluation_domain/domains/recursive_domain.hpp>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <map>
#include <mutex>
#include <sstream>
#include <stdexcept>

namespace ethsnarks {
namespace {

static_assert(sizeof(FieldT) == 32, "Fr<ppT> must be the bare 4 x u64 Montgomery representation");
static_assert(sizeof(FqT) == 32, "Fq<ppT> must be the bare 4 x u64 Montgomery representation");

struct DeviceKey {            // one proving key + one circuit resident on the device(s)
    b200_bases *A = nullptr, *B2 = nullptr, *H = nullptr, *L = nullptr;
    b200_key* key = nullptr;
    size_t num_constraints = 0, num_variables = 0, num_inputs = 0, n_coeffs = 0;
};
typedef std::pair<const void*, const void*> KeyId;     // (&provingKey, &constraint_system)

std::mutex g_mu;
b200_ctx* g_ctx = nullptr;
std::vector<int> g_devices;
std::map<KeyId, DeviceKey> g_keys;
std::vector<std::pair<std::string, float>> g_phases;
bool g_atexit = false;

void fail(const char* what, int rc) {
    throw std::runtime_error(std::string(what) + ": " + b200_strerror(rc) + (g_ctx ? std::string(" - ") + b200_last_error(g_ctx) : ""));
}

b200_ctx* context() {
    if (g_ctx) return g_ctx;
    if (g_devices.empty()) {
        if (const char* e = getenv("B200_DEVICES")) {
            for (const char* p = e; *p;) {
                char* end = nullptr;
                long v = strtol(p, &end, 10);
                if (end == p) break;
                g_devices.push_back((int)v);
                p = *end == ',' ? end + 1 : end;
            }
        }
        if (g_devices.empty()) g_devices.push_back(0);
    }
    int rc = b200_init(g_devices.data(), (int)g_devices.size(), &g_ctx);
    if (rc != B200_OK) { g_ctx = nullptr; fail("b200_init (no CUDA device? there is no CPU fallback)", rc); }
    if (!g_atexit) { g_atexit = true; atexit(b200_shutdown_all); }
    return g_ctx;
}

void free_device_key(DeviceKey& d) {
    if (d.key) b200_key_free(d.key);
    b200_bases_free(d.A); b200_bases_free(d.B2); b200_bases_free(d.H); b200_bases_free(d.L);
    d = DeviceKey();
}

// ---- reference points -> the C ABI's affine Montgomery limbs ((0,0) flag for the point at infinity)
void put_g1(const G1T& p, uint64_t* xy, uint8_t* inf) {
    if (p.is_zero()) { memset(xy, 0, 64); *inf = 1; return; }
    G1T a = p;
    if (!(a.Z == FqT::one())) a.to_affine_coordinates();
    memcpy(xy, a.X.mont_repr.data, 32);
    memcpy(xy + 4, a.Y.mont_repr.data, 32);
    *inf = 0;
}
void put_g2(const G2T& p, uint64_t* xy, uint8_t* inf) {
    if (p.is_zero()) { memset(xy, 0, 128); *inf = 1; return; }
    G2T a = p;
    if (!(a.Z == libff::Fqe<ppT>::one())) a.to_affine_coordinates();
    memcpy(xy, a.X.c0.mont_repr.data, 32);        // x = c0 + c1 u  ->  x.a, x.b
    memcpy(xy + 4, a.X.c1.mont_repr.data, 32);
    memcpy(xy + 8, a.Y.c0.mont_repr.data, 32);
    memcpy(xy + 12, a.Y.c1.mont_repr.data, 32);
    *inf = 0;
}
FqT fq_from_std(const uint64_t* limbs) {
    libff::bigint<FqT::num_limbs> b;
    memcpy(b.data, limbs, 32);
    return FqT(b);
}
FqT fq_from_mont(const uint64_t* limbs) {
    FqT f;
    memcpy(f.mont_repr.data, limbs, 32);
    return f;
}
G1T g1_from(const uint64_t* xy, bool inf, bool mont) {
    if (inf) return G1T::zero();
    return mont ? G1T(fq_from_mont(xy), fq_from_mont(xy + 4), FqT::one()) : G1T(fq_from_std(xy), fq_from_std(xy + 4), FqT::one());
}
G2T g2_from(const uint64_t* xy, bool inf, bool mont) {
    if (inf) return G2T::zero();
    typedef libff::Fqe<ppT> Fq2T;
    if (mont) return G2T(Fq2T(fq_from_mont(xy), fq_from_mont(xy + 4)), Fq2T(fq_from_mont(xy + 8), fq_from_mont(xy + 12)), Fq2T::one());
    return G2T(Fq2T(fq_from_std(xy), fq_from_std(xy + 4)), Fq2T(fq_from_std(xy + 8), fq_from_std(xy + 12)), Fq2T::one());
}
bool all_zero(const uint64_t* p, size_t n) {
    uint64_t o = 0;
    for (size_t i = 0; i < n; ++i) o |= p[i];
    return o == 0;
}

// ---- the constraint system as three CSR matrices over the flattened ConstantStorage table (variable.hpp:47-86);
// r1cs_constraint_light_instance rows come out translated, through IConstraint::getA/B/C (r1cs.hpp:145-211)
struct HostCsr {
    std::vector<uint64_t> rp[3];
    std::vector<uint32_t> col[3], ci[3];
    std::vector<uint64_t> coeffs;       // 4 limbs per constant
    b200_csr m[3];
};
void export_csr(const libsnark::r1cs_constraint_system<FieldT>& cs, HostCsr* out) {
    const size_t nc = cs.constraints.size();
    for (int k = 0; k < 3; ++k) { out->rp[k].assign(1, 0); out->rp[k].reserve(nc + 1); }
    for (const auto& c : cs.constraints) {
        const libsnark::linear_combination_light<FieldT> lc[3] = {c->getA(), c->getB(), c->getC()};
        for (int k = 0; k < 3; ++k) {
            for (const auto& t : lc[k].terms) { out->col[k].push_back((uint32_t)t.index); out->ci[k].push_back((uint32_t)t.coeff_index); }
            out->rp[k].push_back(out->col[k].size());
        }
    }
    const auto& table = libsnark::ConstantStorage<FieldT>::getInstance().constants;
    out->coeffs.resize(4 * table.size());
    for (size_t i = 0; i < table.size(); ++i) memcpy(&out->coeffs[4 * i], table[i].mont_repr.data, 32);
    for (int k = 0; k < 3; ++k) {
        if (out->col[k].empty()) { out->col[k].push_back(0); out->ci[k].push_back(0); }     // keep the pointers valid
        out->m[k] = b200_csr{out->rp[k].data(), out->col[k].data(), out->ci[k].data(), (uint64_t)nc};
    }
}

// uploads the key's queries and the circuit on first use
DeviceKey& device_key(const ProvingKeyT& pk, const libsnark::r1cs_constraint_system<FieldT>& cs) {
    b200_ctx* ctx = context();
    const KeyId id(&pk, &cs);
    DeviceKey& d = g_keys[id];
    const size_t n_coeffs = libsnark::ConstantStorage<FieldT>::getInstance().constants.size();
    if (d.key && d.num_constraints == cs.num_constraints() && d.num_variables == cs.num_variables() && d.num_inputs == cs.num_inputs() &&
        d.n_coeffs == n_coeffs)
        return d;
    free_device_key(d);
    const size_t nv = cs.num_variables(), ni = cs.num_inputs();
    int rc;
    std::vector<uint64_t> xy;
    std::vector<uint8_t> inf;
    std::vector<uint32_t> a_idx, b_idx;
    // kc_multiexp.tcc:54-84: entries from the first index >= the scalar count (nv + 1) on are never used
    auto usable = [&](const std::vector<size_t>& indices) {
        size_t n = 0;
        while (n < indices.size() && indices[n] <= nv) ++n;
        return n;
    };
    {
        const size_t n = usable(pk.A_query.indices);
        xy.resize(8 * n); inf.resize(n); a_idx.resize(n);
        for (size_t i = 0; i < n; ++i) { put_g1(pk.A_query.values[i], &xy[8 * i], &inf[i]); a_idx[i] = (uint32_t)pk.A_query.indices[i]; }
        if ((rc = b200_bases_upload(ctx, 1, xy.data(), inf.data(), n, &d.A))) fail("upload A query", rc);
    }
    {
        const size_t n = usable(pk.B_query.indices);
        xy.resize(16 * n); inf.resize(n); b_idx.resize(n);
        for (size_t i = 0; i < n; ++i) { put_g2(pk.B_query.values[i], &xy[16 * i], &inf[i]); b_idx[i] = (uint32_t)pk.B_query.indices[i]; }
        if ((rc = b200_bases_upload(ctx, 2, xy.data(), inf.data(), n, &d.B2))) fail("upload B query", rc);
    }
    {
        const size_t n = pk.H_query.size();
        xy.resize(8 * n); inf.resize(n);
        for (size_t i = 0; i < n; ++i) put_g1(pk.H_query[i], &xy[8 * i], &inf[i]);
        if ((rc = b200_bases_upload(ctx, 1, xy.data(), inf.data(), n, &d.H))) fail("upload H query", rc);
    }
    {
        const size_t n = std::min(pk.L_query.size(), nv - ni);
        xy.resize(8 * n); inf.resize(n);
        for (size_t i = 0; i < n; ++i) put_g1(pk.L_query[i], &xy[8 * i], &inf[i]);
        if ((rc = b200_bases_upload(ctx, 1, xy.data(), inf.data(), n, &d.L))) fail("upload L query", rc);
    }
    uint64_t c1[24], c2[32];
    uint8_t dummy;
    put_g1(pk.alpha_g1, c1, &dummy); put_g1(pk.beta_g1, c1 + 8, &dummy); put_g1(pk.delta_g1, c1 + 16, &dummy);
    put_g2(pk.beta_g2, c2, &dummy); put_g2(pk.delta_g2, c2 + 16, &dummy);
    rc = b200_key_from_bases(ctx, d.A, a_idx.data(), d.B2, nullptr, b_idx.data(), d.H, d.L, c1, c2, nv, ni, &d.key);
    if (rc) { free_device_key(d); fail("b200_key_from_bases", rc); }
    HostCsr csr;
    export_csr(cs, &csr);
    rc = b200_circuit_load(d.key, &csr.m[0], &csr.m[1], &csr.m[2], csr.coeffs.data(), csr.coeffs.size() / 4, (uint32_t)ni, nv);
    if (rc) { free_device_key(d); fail("b200_circuit_load", rc); }
    d.num_constraints = cs.num_constraints(); d.num_variables = nv; d.num_inputs = ni; d.n_coeffs = n_coeffs;
    return d;
}

ProofT gpu_prove(const ProvingKeyT& pk, const libsnark::r1cs_constraint_system<FieldT>& cs, const std::vector<FieldT>& full_assignment) {
    std::lock_guard<std::mutex> lock(g_mu);        // the library serialises calls on one handle, as the reference prover does
    if (full_assignment.size() != cs.num_variables() + 1)
        throw std::invalid_argument("prove: the assignment must hold num_variables + 1 values (values[0] is the constant ONE)");
    DeviceKey& d = device_key(pk, cs);
    uint64_t A[8], B[16], C[8];
    uint8_t flags[3];
    int rc = b200_prove(d.key, (const uint64_t*)full_assignment.data(), nullptr, nullptr, A, B, C, flags);
    if (rc) fail("b200_prove", rc);
    g_phases.clear();
    b200_ctx* ctx = context();
    for (int i = 0, n = b200_timing_count(ctx); i < n; ++i) {
        char name[96];
        float ms = 0;
        if (b200_timing_get(ctx, i, name, sizeof name, &ms) == B200_OK) g_phases.push_back({name, ms});
    }
    // the same times under the reference's own block names (libff::enter_block, …zok.tcc:454-544), so that the two phase
    // tables can be laid side by side
    static const struct { const char* block; const char* phases[6]; } kBlocks[] = {
        {"Call to r1cs_gg_ppzksnark_zok_prover", {"prove.total"}},
        {"Compute the polynomial H", {"h2d.witness", "witness_eval", "h_chain", "h_chain.front", "h_chain.back"}},
        {"Compute evaluation to A-query", {"msm.A"}},
        {"Compute evaluation to B-query", {"msm.B_g2"}},
        {"Compute evaluation to H-query", {"msm.H"}},
        {"Compute evaluation to L-query", {"msm.L"}},
    };
    const size_t n_lib = g_phases.size();
    for (const auto& b : kBlocks) {
        float ms = 0;
        for (const char* ph : b.phases)
            if (ph)
                for (size_t i = 0; i < n_lib; ++i)
                    if (g_phases[i].first == ph) ms += g_phases[i].second;
        g_phases.push_back({b.block, ms});
    }
    return ProofT(g1_from(A, flags[0], false), g2_from(B, flags[1], false), g1_from(C, flags[2], false));
}

}  // namespace

void b200_use_devices(const std::vector<int>& device_ids) {
    std::lock_guard<std::mutex> lock(g_mu);
    if (g_ctx) throw std::logic_error("b200_use_devices: the context already exists");
    g_devices = device_ids;
}
void b200_release(const ProvingKeyT& proving_key) {
    std::lock_guard<std::mutex> lock(g_mu);
    for (auto it = g_keys.begin(); it != g_keys.end();) {
        if (it->first.first == &proving_key) { free_device_key(it->second); it = g_keys.erase(it); } else ++it;
    }
}
void b200_shutdown_all() {
    std::lock_guard<std::mutex> lock(g_mu);
    for (auto& kv : g_keys) free_device_key(kv.second);
    g_keys.clear();
    if (g_ctx) { b200_shutdown(g_ctx); g_ctx = nullptr; }
}
std::vector<std::pair<std::string, float>> b200_last_phases() {
    std::lock_guard<std::mutex> lock(g_mu);
    return g_phases;
}

// src/stubs.cpp:36-39
ProvingKeyT load_proving_key(const char* pk_file) { return loadFromFile<ProvingKeyT>(pk_file); }

// src/stubs.cpp:42-47
std::string prove(ProverContextT& context, ProtoboardT& pb) {
    auto primary_input = pb.primary_input();
    ProofT proof = gpu_prove(context.provingKey, pb.constraint_system, pb.values);
    return proof_to_json(proof, primary_input);
}

// src/stubs.cpp:61-75.  The device owns its own copy of the domain; this object is only what callers of the reference expect
// to find in ProverContext::domain.
const std::shared_ptr<libfqfft::evaluation_domain<FieldT>> get_domain(ProtoboardT& pb, const ProvingKeyT&, const libsnark::Config& config) {
    const auto& cs = pb.constraint_system;
    size_t domain_size = 1;
    while (domain_size < cs.num_constraints() + cs.num_inputs() + 1) domain_size <<= 1;
    std::shared_ptr<libfqfft::evaluation_domain<FieldT>> result;
    if (config.fft == "basic_radix2") result.reset(new libfqfft::basic_radix2_domain<FieldT>(domain_size));
    else result.reset(new libfqfft::recursive_domain<FieldT>(domain_size, config));
    return result;
}

// src/stubs.cpp:77-87 with the generator on the device: the five field elements the reference draws inside
// r1cs_gg_ppzksnark_zok_generator (…zok.tcc:289-293) are drawn here with the same FieldT::random_element(); the queries come
// back through the key cache blob (b200_key_export) and are written in the reference's own pk / vk.json formats.
static int generate_keys(const libsnark::r1cs_constraint_system<FieldT>& cs, ProvingKeyT& pk, VerificationKeyT& vkey) {
    std::lock_guard<std::mutex> lock(g_mu);
    b200_ctx* ctx = context();
    const size_t nv = cs.num_variables(), ni = cs.num_inputs();
    FieldT toxic[5];     // t, alpha, beta, gamma, delta
    for (auto& x : toxic) x = FieldT::random_element();
    HostCsr csr;
    export_csr(cs, &csr);
    std::vector<uint64_t> vk(56 + 8 * (ni + 1));
    b200_key* key = nullptr;
    int rc = b200_generate_key(ctx, &csr.m[0], &csr.m[1], &csr.m[2], csr.coeffs.data(), csr.coeffs.size() / 4, (uint32_t)ni, nv,
                               (const uint64_t*)toxic, /*full=*/0, &key, vk.data());
    if (rc) return rc;
    size_t len = 0;
    b200_key_export(key, nullptr, 0, &len);
    std::vector<uint8_t> blob(len);
    rc = b200_key_export(key, blob.data(), len, &len);
    b200_key_free(key);
    if (rc) return rc;
    // cache layout (csrc/prove.cu): "B2KC0001" | u64 full, domain_size, nA, nB, nH, nL | g1 consts 3 x 64 B (alpha, beta, delta) |
    // g2 consts 2 x 128 B (beta, delta) | A_idx u32[nA] (+pad 8) | B_idx u32[nB] (+pad 8) | A | B_g2 | H | L ; (0,0) = infinity
    const uint8_t* p = blob.data();
    uint64_t hdr[6];
    memcpy(hdr, p + 8, sizeof hdr);
    const uint64_t nA = hdr[2], nB = hdr[3], nH = hdr[4], nL = hdr[5];
    auto pad8 = [](size_t n) { return (n + 7) & ~(size_t)7; };
    size_t off = 8 + sizeof hdr;
    const uint64_t* c1 = (const uint64_t*)(p + off); off += 192;
    const uint64_t* c2 = (const uint64_t*)(p + off); off += 256;
    const uint32_t* a_idx = (const uint32_t*)(p + off); off += pad8(nA * 4);
    const uint32_t* b_idx = (const uint32_t*)(p + off); off += pad8(nB * 4);
    const uint64_t* A = (const uint64_t*)(p + off); off += nA * 64;
    const uint64_t* B = (const uint64_t*)(p + off); off += nB * 128;
    const uint64_t* H = (const uint64_t*)(p + off); off += nH * 64;
    const uint64_t* L = (const uint64_t*)(p + off);
    pk = ProvingKeyT();
    pk.alpha_g1 = g1_from(c1, false, true); pk.beta_g1 = g1_from(c1 + 8, false, true); pk.delta_g1 = g1_from(c1 + 16, false, true);
    pk.beta_g2 = g2_from(c2, false, true); pk.delta_g2 = g2_from(c2 + 16, false, true);
    // the reference keeps only the non-zero entries of A and B (…zok.hpp:218-230, kc_batch_exp)
    pk.A_query.domain_size_ = nv + 1;
    for (uint64_t i = 0; i < nA; ++i)
        if (!all_zero(A + 8 * i, 8)) { pk.A_query.indices.push_back(a_idx[i]); pk.A_query.values.push_back(g1_from(A + 8 * i, false, true)); }
    pk.B_query.domain_size_ = nv + 1;
    for (uint64_t i = 0; i < nB; ++i)
        if (!all_zero(B + 16 * i, 16)) { pk.B_query.indices.push_back(b_idx[i]); pk.B_query.values.push_back(g2_from(B + 16 * i, false, true)); }
    pk.H_query.reserve(nH);
    for (uint64_t i = 0; i < nH; ++i) pk.H_query.push_back(g1_from(H + 8 * i, all_zero(H + 8 * i, 8), true));
    pk.L_query.reserve(nL);
    for (uint64_t i = 0; i < nL; ++i) pk.L_query.push_back(g1_from(L + 8 * i, all_zero(L + 8 * i, 8), true));
    // vk pieces: alpha_g1 | beta_g2 | gamma_g2 | delta_g2 | IC[0..ni], affine Montgomery
    std::vector<G1T> ic_rest;
    for (size_t i = 1; i <= ni; ++i) ic_rest.push_back(g1_from(&vk[56 + 8 * i], all_zero(&vk[56 + 8 * i], 8), true));
    G1T ic0 = g1_from(&vk[56], all_zero(&vk[56], 8), true);
    vkey = VerificationKeyT(g1_from(&vk[0], false, true), g2_from(&vk[8], false, true), g2_from(&vk[24], false, true), g2_from(&vk[40], false, true),
                            libsnark::accumulation_vector<G1T>(std::move(ic0), std::move(ic_rest)));
    return 0;
}

int stub_genkeys_from_pb(ProtoboardT& pb, const char* pk_file, const char* vk_file) {
    ProvingKeyT pk;
    VerificationKeyT vkey;
    const int rc = generate_keys(pb.constraint_system, pk, vkey);
    if (rc) return rc;
    vk2json_file(vkey, vk_file);
    writeToFile<ProvingKeyT>(pk_file, pk);
    return 0;
}

// src/stubs.cpp:135-149 - the call every gadget test of the reference ends with (test/test_*.cpp): keys for the protoboard's
// constraint system, a proof for its assignment, the verifier's answer.  Here the generator and the prover run on the device
// and the check is the reference's own verifier on the host; nothing is written to disk.
bool stub_test_proof_verify(const ProtoboardT& in_pb) {
    ProvingKeyT pk;
    VerificationKeyT vkey;
    const int rc = generate_keys(in_pb.constraint_system, pk, vkey);
    if (rc) fail("b200_generate_key", rc);
    struct Release { const ProvingKeyT& k; ~Release() { b200_release(k); } } release{pk};   // the side table is keyed by &pk
    const ProofT proof = gpu_prove(pk, in_pb.constraint_system, in_pb.values);
    return libsnark::r1cs_gg_ppzksnark_zok_verifier_strong_IC<ppT>(vkey, in_pb.primary_input(), proof);
}

// src/stubs.cpp:16-34 - verification stays on the host: the reference's JSON readers (src/import.cpp) and verifier
bool stub_verify(const char* vk_json, const char* proof_json) {
    ppT::init_public_params();
    std::stringstream vk_text(vk_json), proof_text(proof_json);
    const VerificationKeyT vk = vk_from_json(vk_text);
    const InputProofPairType input_and_proof = proof_from_json(proof_text);
    return libsnark::r1cs_gg_ppzksnark_zok_verifier_strong_IC<ppT>(vk, input_and_proof.first, input_and_proof.second);
}

// src/stubs.cpp:90-132 - same usage line, messages and exit codes (1 usage / rejected proof, 2 unreadable file, 0 accepted)
int stub_main_verify(const char* prog_name, int argc, const char** argv) {
    if (argc < 3) {
        std::cerr << "Usage: " << prog_name << " " << argv[0] << " <vk.json> <proof.json>" << std::endl;
        return 1;
    }
    std::string text[2];
    for (int i = 0; i < 2; ++i) {
        std::ifstream in(argv[1 + i]);
        if (!in) {
            std::cerr << "Error: cannot open " << argv[1 + i] << std::endl;
            return 2;
        }
        std::stringstream ss;
        ss << in.rdbuf();
        text[i] = ss.str();
    }
    if (stub_verify(text[0].c_str(), text[1].c_str())) return 0;
    std::cerr << "Error: failed to verify proof!" << std::endl;
    return 1;
}

}  // namespace ethsnarks

namespace libsnark {
template <>
r1cs_gg_ppzksnark_zok_proof<ethsnarks::ppT> r1cs_gg_ppzksnark_zok_prover<ethsnarks::ppT>(
    ProverContext<ethsnarks::ppT>& context, const std::vector<libff::Fr<ethsnarks::ppT>>& full_variable_assignment) {
    if (!context.constraint_system) throw std::invalid_argument("r1cs_gg_ppzksnark_zok_prover: context.constraint_system is not set");
    return ethsnarks::gpu_prove(context.provingKey, *context.constraint_system, full_variable_assignment);
}
}  // namespace libsnark
