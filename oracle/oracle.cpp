// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path; the product (ethsnarks_b200/csrc) never
// links, loads or calls this file.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may use it, and only as the checker / the timed CPU baseline.
//
// CPU restatement of the reference's Groth16 prove path (paths relative to /root/reference;
// LS = depends/libsnark/libsnark, FF = depends/libsnark/depends/libff/libff, FQ = depends/libfqfft/libfqfft):
//   SHA512_rng                       FF/common/rng.tcc:25-71
//   multi_exp_with_mixed_addition    FF/algebra/scalar_multiplication/multiexp.tcc:631-714 (0/1 prefilter)
//   multi_exp_with_density           multiexp.tcc:523-593 (c = 16 default, window combine :583-592)
//   multi_exp_inner_bellman…         multiexp.tcc:469-521, digit extraction get_id :167-181
//   kc_multi_exp_with_mixed_addition LS/knowledge_commitment/kc_multiexp.tcc:21-136 (sparse gather)
//   radix-2 FFT                      FQ/evaluation_domain/domains/basic_radix2_domain_aux.tcc:44-79
//   iFFT / cosetFFT / icosetFFT / Z  FQ/evaluation_domain/domains/recursive_domain.tcc:102-142,171-183
//   get_root_of_unity                FF/algebra/fields/field_utils.tcc:35-51
//   r1cs_to_qap_witness_map          LS/reductions/r1cs_to_qap/r1cs_to_qap.tcc:217-302
//   r1cs_to_qap_instance_map_with_evaluation   r1cs_to_qap.tcc:116-186
//   evaluate_all_lagrange_polynomials          FQ/.../basic_radix2_domain_aux.tcc:182-236
//   generator                        src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:277-449 (seeded toxic waste)
//   no-ZK prover                     …zok.tcc:451-550
//   r,s prover formulas              LS/zk_proof_systems/ppzksnark/r1cs_gg_ppzksnark/r1cs_gg_ppzksnark.tcc:485-493
//   key (de)serialisation            …zok.tcc:109-143, LS/common/data_structures/sparse_vector.tcc:273-322,
//                                    FF/algebra/curves/mcl_bn128/mcl_bn128_g1.cpp:181-295
//   proof JSON                       src/export.cpp:20-33,56-121
//   synthetic R1CS                   LS/relations/constraint_satisfaction_problems/r1cs/examples/r1cs_examples.tcc:25-93,96-161
#include "bn254.hpp"
#include <openssl/sha.h>
#include <cassert>
#include <cstdio>
#include <cstdlib>
#include <sstream>
#include <algorithm>
#include <array>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace orc {

Fr FR_ROOT_OF_UNITY, FR_MULT_GEN;
G1 G1_ONE;
G2 G2_ONE;
static Fq2 TWIST_B;
static bool g_inited = false;

static void dec_to_std(const char* s, u64 out[4]) {
    out[0] = out[1] = out[2] = out[3] = 0;
    for (; *s; ++s) {
        if (*s < '0' || *s > '9') continue;
        u128 c = (u128)(*s - '0');
        for (int i = 0; i < 4; ++i) { c += (u128)out[i] * 10; out[i] = (u64)c; c >>= 64; }
    }
}
static void hex_to_std(const char* s, u64 out[4]) {
    out[0] = out[1] = out[2] = out[3] = 0;
    if (s[0] == '0' && (s[1] == 'x' || s[1] == 'X')) s += 2;
    for (; *s; ++s) {
        int d;
        if (*s >= '0' && *s <= '9') d = *s - '0';
        else if (*s >= 'a' && *s <= 'f') d = *s - 'a' + 10;
        else if (*s >= 'A' && *s <= 'F') d = *s - 'A' + 10;
        else continue;
        u64 c = (u64)d;
        for (int i = 0; i < 4; ++i) { u64 n = (out[i] << 4) | c; c = out[i] >> 60; out[i] = n; }
    }
}
Fr fr_from_dec(const char* s) { u64 t[4]; dec_to_std(s, t); return Fr::from_std(t); }
Fq fq_from_dec(const char* s) { u64 t[4]; dec_to_std(s, t); return Fq::from_std(t); }

std::string hex_std(const u64* s) {
    char buf[80];
    snprintf(buf, sizeof buf, "%016llx%016llx%016llx%016llx", (unsigned long long)s[3], (unsigned long long)s[2],
             (unsigned long long)s[1], (unsigned long long)s[0]);
    const char* p = buf;
    while (*p == '0' && p[1]) ++p;
    return std::string(p);
}

static void setup_params(FpParams& P, const u64 p[4]) {
    memcpy(P.p, p, 32);
    // inv = -p^{-1} mod 2^64 by Newton iteration
    u64 x = 1;
    for (int i = 0; i < 6; ++i) x *= 2 - p[0] * x;
    P.inv = (u64)0 - x;
    // R mod p, R^2 mod p by repeated doubling of 1 (512 doublings)
    u64 t[4] = {1, 0, 0, 0};
    for (int i = 0; i < 512; ++i) {
        u64 c = add4(t, t, t);
        if (c || cmp4(t, p) >= 0) sub4(t, t, p);
        if (i == 255) memcpy(P.one, t, 32);
    }
    memcpy(P.r2, t, 32);
}

void init() {
    if (g_inited) return;
    // moduli: mcl_bn128_init.cpp:41 (r), :67 (q)
    static const u64 R_MOD[4] = {0x43e1f593f0000001ull, 0x2833e84879b97091ull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
    static const u64 Q_MOD[4] = {0x3c208c16d87cfd47ull, 0x97816a916871ca8dull, 0xb85045b68181585dull, 0x30644e72e131a029ull};
    setup_params(Fr::P, R_MOD);
    setup_params(Fq::P, Q_MOD);
    FR_ROOT_OF_UNITY = fr_from_dec("1748695177688661943023146337482803886740723238769601073607632802312037301404");
    FR_MULT_GEN = Fr::from_u64(7);
    G1_ONE = G1::from_affine(Fq::from_u64(1), Fq::from_u64(2));
    Fq2 gx{fq_from_dec("10857046999023057135944570762232829481370756359578518086990519993285655852781"),
           fq_from_dec("11559732032986387107991004021392285783925812861821192530917403151452391805634")};
    Fq2 gy{fq_from_dec("8495653923123431417604973247489272438418190587263600148770280649306958101930"),
           fq_from_dec("4082367875863433681332203403145435568316851327593401208105741076214120093531")};
    G2_ONE = G2::from_affine(gx, gy);
    TWIST_B = Fq2{fq_from_dec("19485874751759354771024239261021720505790618469301721065564631296452457478373"),
                  fq_from_dec("266929791119991161246907387137283842545076965332900288569378510910307636690")};
    g_inited = true;
}

// ------------------------------------------------------------------ SHA512_rng (FF/common/rng.tcc:25-71)
template <typename F>
static F sha512_rng(u64 idx) {
    u64 rv[4];
    u64 iter = 0;
    do {
        u64 hash[8];
        SHA512_CTX c;
        SHA512_Init(&c);
        SHA512_Update(&c, &idx, 8);
        SHA512_Update(&c, &iter, 8);
        SHA512_Final((unsigned char*)hash, &c);
        for (int i = 0; i < 4; ++i) rv[i] = hash[i];
        int bitno = 255;
        while (((F::P.p[bitno / 64] >> (bitno % 64)) & 1) == 0) { rv[bitno / 64] &= ~(1ull << (bitno % 64)); --bitno; }
        ++iter;
    } while (cmp4(rv, F::P.p) >= 0);
    return F::from_std(rv);
}

// ------------------------------------------------------------------ point I/O helpers (Montgomery limbs)
struct AffG1 { Fq x, y; bool inf; };
struct AffG2 { Fq2 x, y; bool inf; };
static AffG1 to_aff(const G1& p) { AffG1 a; p.to_affine(a.x, a.y, a.inf); return a; }
static AffG2 to_aff(const G2& p) { AffG2 a; p.to_affine(a.x, a.y, a.inf); return a; }
static G1 from_aff(const AffG1& a) { return a.inf ? G1::zero() : G1::from_affine(a.x, a.y); }
static G2 from_aff(const AffG2& a) { return a.inf ? G2::zero() : G2::from_affine(a.x, a.y); }

static inline G1 load_g1(const u64* xy, bool inf) {
    if (inf) return G1::zero();
    return G1::from_affine(Fq::from_raw(xy), Fq::from_raw(xy + 4));
}
static inline G2 load_g2(const u64* xy, bool inf) {
    if (inf) return G2::zero();
    return G2::from_affine(Fq2{Fq::from_raw(xy), Fq::from_raw(xy + 4)}, Fq2{Fq::from_raw(xy + 8), Fq::from_raw(xy + 12)});
}
static inline void store_aff(const AffG1& a, u64* xy, uint8_t* inf) {
    memcpy(xy, a.x.v, 32); memcpy(xy + 4, a.y.v, 32); *inf = a.inf;
}
static inline void store_aff(const AffG2& a, u64* xy, uint8_t* inf) {
    memcpy(xy, a.x.a.v, 32); memcpy(xy + 4, a.x.b.v, 32); memcpy(xy + 8, a.y.a.v, 32); memcpy(xy + 12, a.y.b.v, 32); *inf = a.inf;
}
template <typename G> struct AffOf;
template <> struct AffOf<G1> { typedef AffG1 T; };
template <> struct AffOf<G2> { typedef AffG2 T; };

template <typename G, typename A>
static inline G madd(const G& acc, const A& b) { return b.inf ? acc : acc.add_affine(b.x, b.y); }

// ------------------------------------------------------------------ MSM
// digit extraction, multiexp.tcc:167-181 (get_id): c-bit window number k of a standard-form scalar
static inline u64 get_id(const u64* e, unsigned c, unsigned k) {
    unsigned bit = c * k;
    if (bit >= 256) return 0;
    unsigned limb = bit / 64, off = bit % 64;
    u64 v = e[limb] >> off;
    if (off + c > 64 && limb + 1 < 4) v |= e[limb + 1] << (64 - off);
    return v & ((1ull << c) - 1);
}

// multi_exp_with_density (multiexp.tcc:523-593): exps in standard form; density[i] = take part in Pippenger
template <typename G, typename A>
static G pippenger(const std::vector<A>& bases, const std::vector<std::array<u64, 4>>& exps,
                   const std::vector<uint8_t>& density, unsigned c) {
    if (c == 0) c = 16;
    const unsigned num_bits = 254;
    const unsigned chunks = (num_bits + c - 1) / c;
    std::vector<G> partial(chunks, G::zero());
    const size_t n = bases.size();
#pragma omp parallel for schedule(dynamic, 1)
    for (unsigned k = 0; k < chunks; ++k) {
        // multi_exp_inner_bellman_with_density (multiexp.tcc:469-521)
        std::vector<G> buckets((size_t)1 << c, G::zero());
        for (size_t i = 0; i < n; ++i) {
            if (!density[i]) continue;
            u64 id = get_id(exps[i].data(), c, k);
            if (id == 0) continue;
            buckets[id] = madd(buckets[id], bases[i]);
        }
        G running = G::zero(), res = G::zero();
        for (size_t i = ((size_t)1 << c) - 1; i > 0; --i) {
            running = running.add(buckets[i]);
            res = res.add(running);
        }
        partial[k] = res;
    }
    // window combine (multiexp.tcc:583-592)
    G fin = G::zero();
    for (int k = (int)chunks - 1; k >= 0; --k) {
        for (unsigned i = 0; i < c; ++i) fin = fin.dbl();
        fin = fin.add(partial[k]);
    }
    return fin;
}

// multi_exp_with_mixed_addition (multiexp.tcc:631-714): 0 -> skip, 1 -> plain add, else Pippenger
template <typename G, typename A>
static G msm_mixed(const std::vector<A>& bases, const Fr* scalars, unsigned c) {
    const size_t n = bases.size();
    std::vector<std::array<u64, 4>> exps(n);
    std::vector<uint8_t> density(n, 0);
    const Fr one = Fr::one();
    G acc = G::zero();
    for (size_t i = 0; i < n; ++i) {
        if (scalars[i].is_zero()) continue;
        if (scalars[i] == one) { acc = madd(acc, bases[i]); continue; }
        density[i] = 1;
        scalars[i].to_std(exps[i].data());
    }
    return acc.add(pippenger<G, A>(bases, exps, density, c));
}

// ------------------------------------------------------------------ NTT
static inline unsigned bitrev(unsigned x, unsigned bits) {
    unsigned r = 0;
    for (unsigned i = 0; i < bits; ++i) { r = (r << 1) | ((x >> i) & 1); }
    return r;
}
// get_root_of_unity (field_utils.tcc:35-51): omega for a domain of size 2^logm
static Fr root_of_unity(unsigned logm) {
    assert(logm <= 28);
    Fr w = FR_ROOT_OF_UNITY;
    for (unsigned i = logm; i < 28; ++i) w = w.sqr();
    return w;
}
// _basic_serial_radix2_FFT (basic_radix2_domain_aux.tcc:44-79); OpenMP over blocks for the baseline timing
static void radix2_fft(Fr* a, unsigned logm, const Fr& omega) {
    const size_t n = (size_t)1 << logm;
    for (size_t k = 0; k < n; ++k) {
        size_t rk = bitrev((unsigned)k, logm);
        if (k < rk) std::swap(a[k], a[rk]);
    }
    size_t m = 1;
    for (unsigned s = 1; s <= logm; ++s) {
        u64 e = n / (2 * m);
        const Fr w_m = omega.pow_u64(e);
        // per-stage twiddle table, then independent butterflies
        std::vector<Fr> tw(m);
        tw[0] = Fr::one();
        for (size_t j = 1; j < m; ++j) tw[j] = tw[j - 1] * w_m;
#pragma omp parallel for schedule(static)
        for (size_t idx = 0; idx < n / 2; ++idx) {
            size_t k = (idx / m) * 2 * m, j = idx % m;
            Fr t = tw[j] * a[k + j + m];
            a[k + j + m] = a[k + j] - t;
            a[k + j] = a[k + j] + t;
        }
        m *= 2;
    }
}
static void mul_by_coset_and_const(Fr* a, size_t n, const Fr& g, const Fr& c) {   // recursive_domain_aux.tcc:328-346
    Fr u = c;
    for (size_t i = 0; i < n; ++i) { a[i] = a[i] * u; u = u * g; }
}
static void ntt_FFT(Fr* a, unsigned logm) { radix2_fft(a, logm, root_of_unity(logm)); }
static void ntt_iFFT(Fr* a, unsigned logm) {
    size_t n = (size_t)1 << logm;
    radix2_fft(a, logm, root_of_unity(logm).inv());
    Fr sc = Fr::from_u64(n).inv();
#pragma omp parallel for
    for (size_t i = 0; i < n; ++i) a[i] = a[i] * sc;
}
static void ntt_cosetFFT(Fr* a, unsigned logm, const Fr& g) {
    mul_by_coset_and_const(a, (size_t)1 << logm, g, Fr::one());
    ntt_FFT(a, logm);
}
static void ntt_icosetFFT(Fr* a, unsigned logm, const Fr& g) {   // recursive_domain.tcc:136-142
    size_t n = (size_t)1 << logm;
    radix2_fft(a, logm, root_of_unity(logm).inv());
    mul_by_coset_and_const(a, n, g.inv(), Fr::from_u64(n).inv());
}

// ------------------------------------------------------------------ R1CS in CSR form
struct Csr {
    std::vector<u64> row_ptr;      // n_rows + 1
    std::vector<uint32_t> col;     // variable index (0 = constant ONE)
    std::vector<uint32_t> cidx;    // index into coeffs
};
struct Circuit {
    u64 nc = 0, ni = 0, nv = 0;    // constraints, primary inputs, variables (excluding ONE)
    Csr M[3];                      // A, B, C
    std::vector<Fr> coeffs;        // coefficient table (the reference's ConstantStorage, variable.hpp:47-86)
    std::vector<Fr> witness;       // optional embedded full assignment, |w| = nv + 1, w[0] = 1
    unsigned logm() const {
        u64 need = nc + ni + 1; unsigned l = 0;
        while (((u64)1 << l) < need) ++l;
        return l;
    }
    u64 m() const { return (u64)1 << logm(); }
};

static Fr eval_row(const Circuit& c, const Csr& M, u64 row, const Fr* w) {   // variable.hpp:296-304
    Fr acc = Fr::zero();
    for (u64 k = M.row_ptr[row]; k < M.row_ptr[row + 1]; ++k) acc += c.coeffs[M.cidx[k]] * w[M.col[k]];
    return acc;
}

// r1cs_to_qap_witness_map (r1cs_to_qap.tcc:217-302). out has m+1 entries (aH[m] = 0).
static void witness_map(const Circuit& c, const Fr* w, std::vector<Fr>& aH, std::vector<Fr>* evalA = nullptr,
                        std::vector<Fr>* evalB = nullptr, std::vector<Fr>* evalC = nullptr) {
    const unsigned logm = c.logm();
    const u64 m = c.m();
    std::vector<Fr> aA(m + 1, Fr::zero()), aB(m + 1, Fr::zero());
    aH.assign(m + 1, Fr::zero());
    std::vector<Fr>& aC = aH;
#pragma omp parallel for
    for (u64 i = 0; i < c.nc; ++i) {
        aA[i] = eval_row(c, c.M[0], i, w);
        aB[i] = eval_row(c, c.M[1], i, w);
        aC[i] = eval_row(c, c.M[2], i, w);
    }
    for (u64 i = 0; i <= c.ni; ++i) aA[i + c.nc] = w[i];
    if (evalA) *evalA = aA;
    if (evalB) *evalB = aB;
    if (evalC) *evalC = aC;
    const Fr g = FR_MULT_GEN;
    ntt_iFFT(aA.data(), logm); ntt_cosetFFT(aA.data(), logm, g);
    ntt_iFFT(aB.data(), logm); ntt_cosetFFT(aB.data(), logm, g);
    ntt_iFFT(aC.data(), logm); ntt_cosetFFT(aC.data(), logm, g);
#pragma omp parallel for
    for (u64 i = 0; i < m; ++i) aH[i] = aA[i] * aB[i] - aC[i];
    aH[m] = Fr::zero();
    // divide_by_Z_on_coset (recursive_domain.tcc:171-183): Z(g) = g^m - 1
    Fr zinv = (g.pow_u64(m) - Fr::one()).inv();
#pragma omp parallel for
    for (u64 i = 0; i < m; ++i) aH[i] = aH[i] * zinv;
    ntt_icosetFFT(aH.data(), logm, g);
}

// generate_r1cs_example_with_field_input (r1cs_examples.tcc:25-93), a and b from SHA512_rng(seed_a/seed_b)
static Circuit* synth_field(u64 nc, u64 ni, u64 seed_a, u64 seed_b) {
    Circuit* c = new Circuit();
    c->nc = nc; c->ni = ni; c->nv = nc + 2;
    c->coeffs.push_back(Fr::one());
    std::vector<Fr> fa;
    fa.reserve(nc + 2);
    Fr a = sha512_rng<Fr>(seed_a), b = sha512_rng<Fr>(seed_b);
    fa.push_back(a); fa.push_back(b);
    for (int k = 0; k < 3; ++k) c->M[k].row_ptr.push_back(0);
    auto term = [&](int k, u64 var) { c->M[k].col.push_back((uint32_t)var); c->M[k].cidx.push_back(0); };
    auto close_row = [&]() { for (int k = 0; k < 3; ++k) c->M[k].row_ptr.push_back(c->M[k].col.size()); };
    for (u64 i = 0; i + 1 < nc; ++i) {
        Fr tmp;
        if (i % 2) { term(0, i + 1); term(1, i + 2); term(2, i + 3); tmp = a * b; }
        else { term(1, 0); term(0, i + 1); term(0, i + 2); term(2, i + 3); tmp = a + b; }
        fa.push_back(tmp); a = b; b = tmp;
        close_row();
    }
    Fr fin = Fr::zero();
    for (u64 i = 1; i < c->nv; ++i) { term(0, i); term(1, i); fin = fin + fa[i - 1]; }
    term(2, c->nv);
    close_row();
    fa.push_back(fin.sqr());
    c->witness.resize(c->nv + 1);
    c->witness[0] = Fr::one();
    for (u64 i = 0; i < c->nv; ++i) c->witness[i + 1] = fa[i];
    return c;
}

// generate_r1cs_example_with_binary_input (r1cs_examples.tcc:96-161) with std::rand replaced by a seeded LCG
static Circuit* synth_binary(u64 nc, u64 ni, u64 seed) {
    Circuit* c = new Circuit();
    c->nc = nc; c->ni = ni; c->nv = ni + nc;
    c->coeffs.push_back(Fr::one());              // 0: +1
    c->coeffs.push_back(-Fr::one());             // 1: -1
    c->coeffs.push_back(Fr::from_u64(2));        // 2: +2
    c->coeffs.push_back(-Fr::from_u64(2));       // 3: -2
    u64 st = seed * 6364136223846793005ull + 1442695040888963407ull;
    auto rnd = [&]() { st = st * 6364136223846793005ull + 1442695040888963407ull; return (st >> 33); };
    std::vector<u64> bits(c->nv);
    for (u64 i = 0; i < ni; ++i) bits[i] = rnd() % 2;
    for (int k = 0; k < 3; ++k) c->M[k].row_ptr.push_back(0);
    auto term = [&](int k, u64 var, uint32_t ci) { c->M[k].col.push_back((uint32_t)var); c->M[k].cidx.push_back(ci); };
    auto close_row = [&]() { for (int k = 0; k < 3; ++k) c->M[k].row_ptr.push_back(c->M[k].col.size()); };
    size_t lastvar = ni - 1;
    for (u64 i = 0; i < nc; ++i) {
        ++lastvar;
        const u64 u = (i == 0 ? rnd() % ni : rnd() % i);
        const u64 v = (i == 0 ? rnd() % ni : rnd() % i);
        // chose two random bits and XOR them together: z = x + y - 2xy  <=>  (2x) * y = x + y - z
        term(0, u + 1, 2);
        term(1, v + 1, 0);
        if (u == v) term(2, u + 1, 2); else { term(2, u + 1, 0); term(2, v + 1, 0); }
        term(2, lastvar + 1, 1);
        close_row();
        bits[lastvar] = bits[u] ^ bits[v];   // xor of two earlier bits (u,v < lastvar by construction)
    }
    c->witness.resize(c->nv + 1);
    c->witness[0] = Fr::one();
    for (u64 i = 0; i < c->nv; ++i) c->witness[i + 1] = bits[i] ? Fr::one() : Fr::zero();
    return c;
}

static bool is_satisfied(const Circuit& c, const Fr* w) {
    for (u64 i = 0; i < c.nc; ++i) {
        if (eval_row(c, c.M[0], i, w) * eval_row(c, c.M[1], i, w) != eval_row(c, c.M[2], i, w)) return false;
    }
    return true;
}

// ------------------------------------------------------------------ fixed-base multiplication (generator only)
template <typename G, typename A>
struct FixedBase {
    static const unsigned W = 8;
    std::vector<A> table;    // 32 windows x 256 entries, affine
    explicit FixedBase(const G& g) {
        table.resize(32 * 256);
        G base = g;
        for (unsigned w = 0; w < 32; ++w) {
            G cur = G::zero();
            for (unsigned d = 0; d < 256; ++d) {
                table[w * 256 + d] = to_aff(cur);
                cur = cur.add(base);
            }
            base = cur;   // 256 * previous base
        }
    }
    G mul(const Fr& k) const {
        u64 e[4]; k.to_std(e);
        G r = G::zero();
        for (unsigned w = 0; w < 32; ++w) {
            unsigned d = (e[w / 8] >> ((w % 8) * 8)) & 0xff;
            if (d) r = madd(r, table[w * 256 + d]);
        }
        return r;
    }
};

// ------------------------------------------------------------------ key
struct Key {
    // proving key in the layout of r1cs_gg_ppzksnark_zok_proving_key (…zok.hpp:76-153); B_g1 kept for r,s mode
    AffG1 alpha_g1, beta_g1, delta_g1;
    AffG2 beta_g2, delta_g2, gamma_g2;
    u64 domain_size = 0;                       // sparse_vector::domain_size_ = nv + 1
    std::vector<u64> A_idx, B_idx;
    std::vector<AffG1> A_val, B_g1_val, H, L, gamma_ABC;
    std::vector<AffG2> B_val;
    // toxic waste + QAP evaluations (only known for oracle-generated keys)
    bool has_waste = false;
    Fr t, alpha, beta, gamma, delta, Zt;
    std::vector<Fr> At, Bt, Ct;
    u64 m = 0, ni = 0, nv = 0;
};

// evaluate_all_lagrange_polynomials (basic_radix2_domain_aux.tcc:182-236)
static std::vector<Fr> lagrange_at(unsigned logm, const Fr& t) {
    const u64 m = (u64)1 << logm;
    std::vector<Fr> u(m, Fr::zero());
    const Fr omega = root_of_unity(logm);
    Fr tm = t.pow_u64(m);
    if (tm == Fr::one()) {   // t in the domain
        Fr w = Fr::one();
        for (u64 i = 0; i < m; ++i) { if (w == t) { u[i] = Fr::one(); return u; } w = w * omega; }
    }
    const Fr Z = tm - Fr::one();
    Fr l = Z * Fr::from_u64(m).inv();
    Fr r = Fr::one();
    // u[i] = l_i / (t - r_i), l_i = Z/m * omega^i ; batch inversion of (t - omega^i)
    std::vector<Fr> den(m), pre(m);
    for (u64 i = 0; i < m; ++i) { den[i] = t - r; r = r * omega; }
    Fr acc = Fr::one();
    for (u64 i = 0; i < m; ++i) { pre[i] = acc; acc = acc * den[i]; }
    Fr ainv = acc.inv();
    for (u64 i = m; i-- > 0;) { Fr di = ainv * pre[i]; ainv = ainv * den[i]; den[i] = di; }
    for (u64 i = 0; i < m; ++i) { u[i] = l * den[i]; l = l * omega; }
    return u;
}

static Key* genkey(const Circuit& c, u64 seed_base) {
    Key* K = new Key();
    K->has_waste = true;
    K->t = sha512_rng<Fr>(seed_base + 0); K->alpha = sha512_rng<Fr>(seed_base + 1); K->beta = sha512_rng<Fr>(seed_base + 2);
    K->gamma = sha512_rng<Fr>(seed_base + 3); K->delta = sha512_rng<Fr>(seed_base + 4);
    const Fr ginv = K->gamma.inv(), dinv = K->delta.inv();
    const unsigned logm = c.logm();
    const u64 m = c.m();
    K->m = m; K->ni = c.ni; K->nv = c.nv;
    // r1cs_to_qap_instance_map_with_evaluation
    std::vector<Fr> At(c.nv + 1, Fr::zero()), Bt(c.nv + 1, Fr::zero()), Ct(c.nv + 1, Fr::zero());
    K->Zt = K->t.pow_u64(m) - Fr::one();
    std::vector<Fr> u = lagrange_at(logm, K->t);
    for (u64 i = 0; i <= c.ni; ++i) At[i] = u[c.nc + i];
    for (u64 i = 0; i < c.nc; ++i) {
        for (u64 k = c.M[0].row_ptr[i]; k < c.M[0].row_ptr[i + 1]; ++k) At[c.M[0].col[k]] += u[i] * c.coeffs[c.M[0].cidx[k]];
        for (u64 k = c.M[1].row_ptr[i]; k < c.M[1].row_ptr[i + 1]; ++k) Bt[c.M[1].col[k]] += u[i] * c.coeffs[c.M[1].cidx[k]];
        for (u64 k = c.M[2].row_ptr[i]; k < c.M[2].row_ptr[i + 1]; ++k) Ct[c.M[2].col[k]] += u[i] * c.coeffs[c.M[2].cidx[k]];
    }
    FixedBase<G1, AffG1> T1(G1_ONE);
    FixedBase<G2, AffG2> T2(G2_ONE);
    K->alpha_g1 = to_aff(T1.mul(K->alpha)); K->beta_g1 = to_aff(T1.mul(K->beta)); K->delta_g1 = to_aff(T1.mul(K->delta));
    K->beta_g2 = to_aff(T2.mul(K->beta)); K->delta_g2 = to_aff(T2.mul(K->delta)); K->gamma_g2 = to_aff(T2.mul(K->gamma));
    K->domain_size = c.nv + 1;
    for (u64 i = 0; i <= c.nv; ++i) {
        if (!At[i].is_zero()) K->A_idx.push_back(i);
        if (!Bt[i].is_zero()) K->B_idx.push_back(i);
    }
    K->A_val.resize(K->A_idx.size()); K->B_val.resize(K->B_idx.size()); K->B_g1_val.resize(K->B_idx.size());
#pragma omp parallel for schedule(dynamic, 64)
    for (size_t j = 0; j < K->A_idx.size(); ++j) K->A_val[j] = to_aff(T1.mul(At[K->A_idx[j]]));
#pragma omp parallel for schedule(dynamic, 64)
    for (size_t j = 0; j < K->B_idx.size(); ++j) {
        K->B_val[j] = to_aff(T2.mul(Bt[K->B_idx[j]]));
        K->B_g1_val[j] = to_aff(T1.mul(Bt[K->B_idx[j]]));
    }
    // H query: t^i * Zt / delta, i < m - 1
    K->H.resize(m - 1);
    {
        std::vector<Fr> hs(m - 1);
        Fr ti = K->Zt * dinv;
        for (u64 i = 0; i + 1 < m; ++i) { hs[i] = ti; ti = ti * K->t; }
#pragma omp parallel for schedule(dynamic, 64)
        for (u64 i = 0; i < m - 1; ++i) K->H[i] = to_aff(T1.mul(hs[i]));
    }
    K->L.resize(c.nv - c.ni);
#pragma omp parallel for schedule(dynamic, 64)
    for (u64 i = 0; i < c.nv - c.ni; ++i) {
        u64 j = c.ni + 1 + i;
        K->L[i] = to_aff(T1.mul((K->beta * At[j] + K->alpha * Bt[j] + Ct[j]) * dinv));
    }
    K->gamma_ABC.resize(c.ni + 1);
    for (u64 i = 0; i <= c.ni; ++i) K->gamma_ABC[i] = to_aff(T1.mul((K->beta * At[i] + K->alpha * Bt[i] + Ct[i]) * ginv));
    K->At = At; K->Bt = Bt; K->Ct = Ct;
    return K;
}

// ---- serialisation.  fmt 0 = MCL_NOZK (…zok.tcc:109-143 with mcl points), 1 = MCL_FULL (B values = G2 then G1),
//      2 = ALT_NOZK, 3 = ALT_FULL (libff alt_bn128 points: flag char + X + Y Montgomery limbs, alt_bn128_g1.cpp:404-465)
static void put_g1(std::string& o, const AffG1& p, bool alt) {
    o.push_back(p.inf ? '1' : '0');
    if (alt) {
        // libff writes the affine copy of the point; zero is (0, 1, 0) (alt_bn128_g1.cpp:404-417, alt_bn128_init.cpp G1_zero)
        u64 z[4] = {0, 0, 0, 0};
        if (p.inf) { o.append((const char*)z, 32); o.append((const char*)Fq::P.one, 32); }
        else { o.append((const char*)p.x.v, 32); o.append((const char*)p.y.v, 32); }
    } else {
        u64 z[4] = {0, 0, 0, 0};
        if (p.inf) { o.append((const char*)z, 32); o.append((const char*)z, 32); o.append((const char*)z, 32); }
        else { o.append((const char*)p.x.v, 32); o.append((const char*)p.y.v, 32); o.append((const char*)Fq::P.one, 32); }
    }
}
static void put_g2(std::string& o, const AffG2& p, bool alt) {
    o.push_back(p.inf ? '1' : '0');
    if (alt) {
        u64 z[4] = {0, 0, 0, 0};
        if (p.inf) { o.append((const char*)z, 32); o.append((const char*)z, 32); o.append((const char*)Fq::P.one, 32); o.append((const char*)z, 32); }
        else {
            o.append((const char*)p.x.a.v, 32); o.append((const char*)p.x.b.v, 32);
            o.append((const char*)p.y.a.v, 32); o.append((const char*)p.y.b.v, 32);
        }
    } else {
        u64 z[4] = {0, 0, 0, 0};
        if (p.inf) { for (int i = 0; i < 6; ++i) o.append((const char*)z, 32); }
        else {
            o.append((const char*)p.x.a.v, 32); o.append((const char*)p.x.b.v, 32);
            o.append((const char*)p.y.a.v, 32); o.append((const char*)p.y.b.v, 32);
            o.append((const char*)Fq::P.one, 32); o.append((const char*)z, 32);
        }
    }
}
static void put_num(std::string& o, u64 v) { o += std::to_string(v); o.push_back('\n'); }

static std::string serialize_key(const Key& K, int fmt) {
    const bool alt = fmt >= 2, full = (fmt & 1);
    std::string o;
    put_g1(o, K.alpha_g1, alt); put_g1(o, K.beta_g1, alt); put_g2(o, K.beta_g2, alt);
    put_g1(o, K.delta_g1, alt); put_g2(o, K.delta_g2, alt);
    if (full) {
        // the un-stripped key holds A as a dense libff::G1_vector (…zok.hpp:85, built at …zok.tcc:387, written with
        // `out << pk.A_query`, …zok.tcc:61): a count and then one point per variable, zeros included, no index list
        put_num(o, K.domain_size);
        AffG1 zero; zero.inf = true;
        size_t j = 0;
        for (u64 i = 0; i < K.domain_size; ++i) {
            if (j < K.A_idx.size() && K.A_idx[j] == i) put_g1(o, K.A_val[j++], alt); else put_g1(o, zero, alt);
        }
    } else {
        put_num(o, K.domain_size); put_num(o, K.A_idx.size());
        for (u64 i : K.A_idx) put_num(o, i);
        put_num(o, K.A_val.size());
        for (auto& p : K.A_val) put_g1(o, p, alt);
    }
    put_num(o, K.domain_size); put_num(o, K.B_idx.size());
    for (u64 i : K.B_idx) put_num(o, i);
    put_num(o, K.B_val.size());
    for (size_t j = 0; j < K.B_val.size(); ++j) { put_g2(o, K.B_val[j], alt); if (full) put_g1(o, K.B_g1_val[j], alt); }
    put_num(o, K.H.size());
    for (auto& p : K.H) put_g1(o, p, alt);
    put_num(o, K.L.size());
    for (auto& p : K.L) put_g1(o, p, alt);
    return o;
}

// ---- proof
struct Proof { AffG1 A; AffG2 B; AffG1 C; };

template <typename G, typename A>
static G sparse_msm(const std::vector<u64>& idx, const std::vector<A>& vals, const Fr* w, u64 scalar_size, unsigned c) {
    // kc_multi_exp_with_mixed_addition (kc_multiexp.tcc:21-136): gather scalars by index, stop at idx >= scalar_size
    std::vector<A> b; std::vector<Fr> s;
    for (size_t j = 0; j < idx.size(); ++j) {
        if (idx[j] >= scalar_size) break;
        b.push_back(vals[j]); s.push_back(w[idx[j]]);
    }
    return msm_mixed<G, A>(b, s.data(), c);
}

static Proof prove(const Key& K, const Circuit& c, const Fr* w, const Fr* r, const Fr* s, unsigned win) {
    std::vector<Fr> aH;
    witness_map(c, w, aH);
    const u64 m = c.m();
    G1 evA = sparse_msm<G1, AffG1>(K.A_idx, K.A_val, w, c.nv + 1, win);
    G2 evB = sparse_msm<G2, AffG2>(K.B_idx, K.B_val, w, c.nv + 1, win);
    // H: multi_exp (no prefilter), m - 1 terms (…zok.tcc:509-519)
    G1 evH;
    {
        std::vector<std::array<u64, 4>> exps(m - 1);
        std::vector<uint8_t> dens(m - 1, 1);
        for (u64 i = 0; i + 1 < m; ++i) aH[i].to_std(exps[i].data());
        evH = pippenger<G1, AffG1>(K.H, exps, dens, win);
    }
    G1 evL = msm_mixed<G1, AffG1>(K.L, w + c.ni + 1, win);
    G1 gA = from_aff(K.alpha_g1).add(evA);
    G2 gB = from_aff(K.beta_g2).add(evB);
    G1 gC = evH.add(evL);
    if (r && s) {   // r1cs_gg_ppzksnark.tcc:485-493
        u64 rs_[4], ss_[4], rs2[4];
        r->to_std(rs_); s->to_std(ss_); ((*r) * (*s)).to_std(rs2);
        G1 evB1 = sparse_msm<G1, AffG1>(K.B_idx, K.B_g1_val, w, c.nv + 1, win);
        G1 d1 = from_aff(K.delta_g1);
        G2 d2 = from_aff(K.delta_g2);
        gA = gA.add(d1.mul(rs_));
        gB = gB.add(d2.mul(ss_));
        G1 gB1 = from_aff(K.beta_g1).add(evB1).add(d1.mul(ss_));
        gC = gC.add(gA.mul(ss_)).add(gB1.mul(rs_)).add(d1.mul(rs2).neg());
    }
    return Proof{to_aff(gA), to_aff(gB), to_aff(gC)};
}

// Verification "in the exponent" (possible because the oracle generator knows t, alpha, beta, gamma, delta):
//   checks A == a*G1, B == b*G2, C == c*G1 with a,b,c recomputed in Fr from the witness, and the Groth16
//   equation a*b == alpha*beta + (sum_{i<=ni} w_i*gammaABC_i)*gamma + c*delta in Fr.
// Returns a bit mask of failed checks (0 = all pass).
static int check_in_exponent(const Key& K, const Circuit& c, const Fr* w, const Fr* r, const Fr* s, const Proof& P) {
    if (!K.has_waste) return -1;
    std::vector<Fr> aH;
    witness_map(c, w, aH);
    Fr a = K.alpha, b = K.beta, l = Fr::zero(), pub = Fr::zero();
    for (u64 i = 0; i <= c.nv; ++i) {
        a += w[i] * K.At[i]; b += w[i] * K.Bt[i];
        Fr abc = K.beta * K.At[i] + K.alpha * K.Bt[i] + K.Ct[i];
        if (i > c.ni) l += w[i] * abc; else pub += w[i] * abc;
    }
    const Fr dinv = K.delta.inv();
    Fr h = Fr::zero(), ti = Fr::one();
    for (u64 i = 0; i + 1 < c.m(); ++i) { h += aH[i] * ti; ti = ti * K.t; }
    Fr cc = (l + h * K.Zt) * dinv;
    if (r && s) {
        a += (*r) * K.delta; b += (*s) * K.delta;
        cc += (*s) * a + (*r) * b - (*r) * (*s) * K.delta;
    }
    int bad = 0;
    u64 e[4];
    a.to_std(e); if (!G1_ONE.mul(e).eq(from_aff(P.A))) bad |= 1;
    b.to_std(e); if (!G2_ONE.mul(e).eq(from_aff(P.B))) bad |= 2;
    cc.to_std(e); if (!G1_ONE.mul(e).eq(from_aff(P.C))) bad |= 4;
    if (a * b != K.alpha * K.beta + pub + cc * K.delta) bad |= 8;
    return bad;
}

static std::string g1_json(const AffG1& p) {
    u64 x[4], y[4]; p.x.to_std(x); p.y.to_std(y);
    return "\"0x" + hex_std(x) + "\", \"0x" + hex_std(y) + "\"";
}
static std::string g2_json(const AffG2& p) {   // export.cpp:66-96: [x.c1, x.c0], [y.c1, y.c0]
    u64 xa[4], xb[4], ya[4], yb[4];
    p.x.a.to_std(xa); p.x.b.to_std(xb); p.y.a.to_std(ya); p.y.b.to_std(yb);
    return "[\"0x" + hex_std(xb) + "\", \"0x" + hex_std(xa) + "\"],\n [\"0x" + hex_std(yb) + "\", \"0x" + hex_std(ya) + "\"]";
}
static std::string proof_json(const Proof& P, const Fr* inputs, u64 n) {   // export.cpp:99-121
    std::string s = "{\n";
    s += " \"A\" :[" + g1_json(P.A) + "],\n";
    s += " \"B\"  :[" + g2_json(P.B) + "],\n";
    s += " \"C\"  :[" + g1_json(P.C) + "],\n";
    s += " \"input\" :[";
    for (u64 i = 0; i < n; ++i) {
        u64 v[4]; inputs[i].to_std(v);
        s += "\"0x" + hex_std(v) + "\"";
        if (i + 1 < n) s += ", ";
    }
    s += "]\n}";
    return s;
}
static std::string vk_json(const Key& K) {   // export.cpp:124-145
    std::string s = "{\n";
    s += " \"alpha\" :[" + g1_json(K.alpha_g1) + "],\n";
    s += " \"beta\"  :[" + g2_json(K.beta_g2) + "],\n";
    s += " \"gamma\" :[" + g2_json(K.gamma_g2) + "],\n";
    s += " \"delta\" :[" + g2_json(K.delta_g2) + "],\n";
    s += "\"gammaABC\" :[[" + g1_json(K.gamma_ABC[0]) + "]";
    for (size_t i = 1; i < K.gamma_ABC.size(); ++i) s += ",[" + g1_json(K.gamma_ABC[i]) + "]";
    s += "]}";
    return s;
}

}  // namespace orc

// =================================================================== C API (ctypes)
using namespace orc;
extern "C" {

void orc_init() { init(); }
int orc_num_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

#define BATCH2(NAME, F, EXPR)                                                              \
    void NAME(const u64* a, const u64* b, u64* out, size_t n) {                            \
        init();                                                                            \
        _Pragma("omp parallel for") for (size_t i = 0; i < n; ++i) {                       \
            F x = F::from_raw(a + 4 * i), y = F::from_raw(b + 4 * i);                      \
            F z = EXPR; memcpy(out + 4 * i, z.v, 32);                                      \
        }                                                                                  \
    }
BATCH2(orc_fr_mul, Fr, x* y)
BATCH2(orc_fr_add, Fr, x + y)
BATCH2(orc_fr_sub, Fr, x - y)
BATCH2(orc_fq_mul, Fq, x* y)
BATCH2(orc_fq_add, Fq, x + y)
BATCH2(orc_fq_sub, Fq, x - y)

#define BATCH1(NAME, F, BODY)                                                              \
    void NAME(const u64* a, u64* out, size_t n) {                                          \
        init();                                                                            \
        _Pragma("omp parallel for") for (size_t i = 0; i < n; ++i) {                       \
            F x = F::from_raw(a + 4 * i); u64 z[4]; BODY; memcpy(out + 4 * i, z, 32);      \
        }                                                                                  \
    }
BATCH1(orc_fr_from_mont, Fr, x.to_std(z))
BATCH1(orc_fq_from_mont, Fq, x.to_std(z))
BATCH1(orc_fr_to_mont, Fr, { Fr y = Fr::from_std(x.v); memcpy(z, y.v, 32); })
BATCH1(orc_fq_to_mont, Fq, { Fq y = Fq::from_std(x.v); memcpy(z, y.v, 32); })
BATCH1(orc_fr_inv, Fr, { Fr y = x.inv(); memcpy(z, y.v, 32); })
BATCH1(orc_fq_inv, Fq, { Fq y = x.inv(); memcpy(z, y.v, 32); })

void orc_sha512_rng_fr(u64 idx0, size_t n, u64* out_mont) {
    init();
#pragma omp parallel for
    for (size_t i = 0; i < n; ++i) { Fr x = sha512_rng<Fr>(idx0 + i); memcpy(out_mont + 4 * i, x.v, 32); }
}
void orc_constants(u64* out) {   // r, q, Fr.one, Fq.one, Fr.r2, Fq.r2, root_of_unity(mont), [Fr.inv, Fq.inv,0,0]
    init();
    memcpy(out, Fr::P.p, 32); memcpy(out + 4, Fq::P.p, 32); memcpy(out + 8, Fr::P.one, 32); memcpy(out + 12, Fq::P.one, 32);
    memcpy(out + 16, Fr::P.r2, 32); memcpy(out + 20, Fq::P.r2, 32); memcpy(out + 24, FR_ROOT_OF_UNITY.v, 32);
    out[28] = Fr::P.inv; out[29] = Fq::P.inv; out[30] = out[31] = 0;
}
void orc_generators(u64* g1_xy, u64* g2_xy) {
    init();
    uint8_t f;
    store_aff(to_aff(G1_ONE), g1_xy, &f); store_aff(to_aff(G2_ONE), g2_xy, &f);
}
int orc_g1_on_curve(const u64* xy) {
    init();
    Fq x = Fq::from_raw(xy), y = Fq::from_raw(xy + 4);
    return y.sqr() == x.sqr() * x + Fq::from_u64(3);
}
int orc_g2_on_curve(const u64* xy) {
    init();
    Fq2 x{Fq::from_raw(xy), Fq::from_raw(xy + 4)}, y{Fq::from_raw(xy + 8), Fq::from_raw(xy + 12)};
    return y.sqr() == x.sqr() * x + TWIST_B;
}

// k_i * G (scalars Montgomery) -> affine Montgomery + inf flags
void orc_g1_fixed_base(const u64* scalars_mont, size_t n, u64* out_xy, uint8_t* out_inf) {
    init();
    static FixedBase<G1, AffG1>* T = nullptr;
    if (!T) T = new FixedBase<G1, AffG1>(G1_ONE);
#pragma omp parallel for schedule(dynamic, 64)
    for (size_t i = 0; i < n; ++i) store_aff(to_aff(T->mul(Fr::from_raw(scalars_mont + 4 * i))), out_xy + 8 * i, out_inf + i);
}
void orc_g2_fixed_base(const u64* scalars_mont, size_t n, u64* out_xy, uint8_t* out_inf) {
    init();
    static FixedBase<G2, AffG2>* T = nullptr;
    if (!T) T = new FixedBase<G2, AffG2>(G2_ONE);
#pragma omp parallel for schedule(dynamic, 64)
    for (size_t i = 0; i < n; ++i) store_aff(to_aff(T->mul(Fr::from_raw(scalars_mont + 4 * i))), out_xy + 16 * i, out_inf + i);
}
// point addition of two affine points (covers doubling / opposite / infinity cases), for group-law parity tests
void orc_g1_add(const u64* a_xy, uint8_t a_inf, const u64* b_xy, uint8_t b_inf, u64* out_xy, uint8_t* out_inf) {
    init();
    store_aff(to_aff(load_g1(a_xy, a_inf).add(load_g1(b_xy, b_inf))), out_xy, out_inf);
}
void orc_g2_add(const u64* a_xy, uint8_t a_inf, const u64* b_xy, uint8_t b_inf, u64* out_xy, uint8_t* out_inf) {
    init();
    store_aff(to_aff(load_g2(a_xy, a_inf).add(load_g2(b_xy, b_inf))), out_xy, out_inf);
}

// reference-semantics MSM: multi_exp_with_mixed_addition (prefilter = 1) or multi_exp (prefilter = 0)
void orc_msm_g1(const u64* bases_xy, const uint8_t* inf, const u64* scalars_mont, size_t n, unsigned c, int prefilter,
                u64* out_xy, uint8_t* out_inf) {
    init();
    std::vector<AffG1> b(n);
    for (size_t i = 0; i < n; ++i) { b[i].inf = inf && inf[i]; b[i].x = Fq::from_raw(bases_xy + 8 * i); b[i].y = Fq::from_raw(bases_xy + 8 * i + 4); }
    G1 r;
    if (prefilter) r = msm_mixed<G1, AffG1>(b, (const Fr*)scalars_mont, c);
    else {
        std::vector<std::array<u64, 4>> e(n); std::vector<uint8_t> d(n, 1);
        for (size_t i = 0; i < n; ++i) Fr::from_raw(scalars_mont + 4 * i).to_std(e[i].data());
        r = pippenger<G1, AffG1>(b, e, d, c);
    }
    store_aff(to_aff(r), out_xy, out_inf);
}
void orc_msm_g2(const u64* bases_xy, const uint8_t* inf, const u64* scalars_mont, size_t n, unsigned c, int prefilter,
                u64* out_xy, uint8_t* out_inf) {
    init();
    std::vector<AffG2> b(n);
    for (size_t i = 0; i < n; ++i) {
        b[i].inf = inf && inf[i];
        b[i].x = Fq2{Fq::from_raw(bases_xy + 16 * i), Fq::from_raw(bases_xy + 16 * i + 4)};
        b[i].y = Fq2{Fq::from_raw(bases_xy + 16 * i + 8), Fq::from_raw(bases_xy + 16 * i + 12)};
    }
    G2 r;
    if (prefilter) r = msm_mixed<G2, AffG2>(b, (const Fr*)scalars_mont, c);
    else {
        std::vector<std::array<u64, 4>> e(n); std::vector<uint8_t> d(n, 1);
        for (size_t i = 0; i < n; ++i) Fr::from_raw(scalars_mont + 4 * i).to_std(e[i].data());
        r = pippenger<G2, AffG2>(b, e, d, c);
    }
    store_aff(to_aff(r), out_xy, out_inf);
}
// naive sum of scalar multiplications (independent of the Pippenger code), for tiny cases
void orc_msm_g1_naive(const u64* bases_xy, const uint8_t* inf, const u64* scalars_mont, size_t n, u64* out_xy, uint8_t* out_inf) {
    init();
    G1 acc = G1::zero();
    for (size_t i = 0; i < n; ++i) { u64 e[4]; Fr::from_raw(scalars_mont + 4 * i).to_std(e); acc = acc.add(load_g1(bases_xy + 8 * i, inf && inf[i]).mul(e)); }
    store_aff(to_aff(acc), out_xy, out_inf);
}
void orc_msm_g2_naive(const u64* bases_xy, const uint8_t* inf, const u64* scalars_mont, size_t n, u64* out_xy, uint8_t* out_inf) {
    init();
    G2 acc = G2::zero();
    for (size_t i = 0; i < n; ++i) { u64 e[4]; Fr::from_raw(scalars_mont + 4 * i).to_std(e); acc = acc.add(load_g2(bases_xy + 16 * i, inf && inf[i]).mul(e)); }
    store_aff(to_aff(acc), out_xy, out_inf);
}
// sum_i a_i * b_i in Fr (Montgomery in/out) — the "in the exponent" expectation for MSMs over bases k_i*G
void orc_fr_dot(const u64* a, const u64* b, size_t n, u64* out) {
    init();
    Fr acc = Fr::zero();
#pragma omp parallel
    {
        Fr loc = Fr::zero();
#pragma omp for nowait
        for (size_t i = 0; i < n; ++i) loc += Fr::from_raw(a + 4 * i) * Fr::from_raw(b + 4 * i);
#pragma omp critical
        acc += loc;
    }
    memcpy(out, acc.v, 32);
}

// mode: 0 FFT, 1 iFFT, 2 cosetFFT, 3 icosetFFT (coset generator 7)
void orc_ntt(u64* data_mont, unsigned logm, int mode) {
    init();
    Fr* a = (Fr*)data_mont;
    switch (mode) {
        case 0: ntt_FFT(a, logm); break;
        case 1: ntt_iFFT(a, logm); break;
        case 2: ntt_cosetFFT(a, logm, FR_MULT_GEN); break;
        case 3: ntt_icosetFFT(a, logm, FR_MULT_GEN); break;
    }
}
// direct O(n^2) DFT for tiny sizes (independent check of the radix-2 code): out[i] = sum_j a[j] w^(ij)
void orc_dft_naive(const u64* in_mont, u64* out_mont, unsigned logm) {
    init();
    size_t n = (size_t)1 << logm;
    Fr w = root_of_unity(logm);
    for (size_t i = 0; i < n; ++i) {
        Fr wi = w.pow_u64(i), x = Fr::one(), acc = Fr::zero();
        for (size_t j = 0; j < n; ++j) { acc += Fr::from_raw(in_mont + 4 * j) * x; x = x * wi; }
        memcpy(out_mont + 4 * i, acc.v, 32);
    }
}

// ---- circuits
void* orc_circuit_new(u64 nc, u64 ni, u64 nv, const u64* rpA, const uint32_t* colA, const uint32_t* ciA, const u64* rpB,
                      const uint32_t* colB, const uint32_t* ciB, const u64* rpC, const uint32_t* colC, const uint32_t* ciC,
                      const u64* coeffs_mont, u64 n_coeffs) {
    init();
    Circuit* c = new Circuit();
    c->nc = nc; c->ni = ni; c->nv = nv;
    const u64* rp[3] = {rpA, rpB, rpC}; const uint32_t* col[3] = {colA, colB, colC}; const uint32_t* ci[3] = {ciA, ciB, ciC};
    for (int k = 0; k < 3; ++k) {
        c->M[k].row_ptr.assign(rp[k], rp[k] + nc + 1);
        u64 nnz = rp[k][nc];
        c->M[k].col.assign(col[k], col[k] + nnz); c->M[k].cidx.assign(ci[k], ci[k] + nnz);
    }
    c->coeffs.resize(n_coeffs);
    memcpy(c->coeffs.data(), coeffs_mont, 32 * n_coeffs);
    return c;
}
void* orc_synth_field(u64 nc, u64 ni, u64 seed_a, u64 seed_b) { init(); return synth_field(nc, ni, seed_a, seed_b); }
void* orc_synth_binary(u64 nc, u64 ni, u64 seed) { init(); return synth_binary(nc, ni, seed); }
void orc_circuit_free(void* h) { delete (Circuit*)h; }
// out: nc, ni, nv, m, nnzA, nnzB, nnzC, n_coeffs
void orc_circuit_dims(void* h, u64* out) {
    Circuit* c = (Circuit*)h;
    out[0] = c->nc; out[1] = c->ni; out[2] = c->nv; out[3] = c->m();
    for (int k = 0; k < 3; ++k) out[4 + k] = c->M[k].col.size();
    out[7] = c->coeffs.size();
}
void orc_circuit_matrix(void* h, int which, u64* row_ptr, uint32_t* col, uint32_t* cidx) {
    Circuit* c = (Circuit*)h;
    memcpy(row_ptr, c->M[which].row_ptr.data(), 8 * c->M[which].row_ptr.size());
    memcpy(col, c->M[which].col.data(), 4 * c->M[which].col.size());
    memcpy(cidx, c->M[which].cidx.data(), 4 * c->M[which].cidx.size());
}
void orc_circuit_coeffs(void* h, u64* out) { Circuit* c = (Circuit*)h; memcpy(out, c->coeffs.data(), 32 * c->coeffs.size()); }
int orc_circuit_witness(void* h, u64* out) {
    Circuit* c = (Circuit*)h;
    if (c->witness.empty()) return 1;
    memcpy(out, c->witness.data(), 32 * c->witness.size());
    return 0;
}
int orc_is_satisfied(void* h, const u64* w) { return is_satisfied(*(Circuit*)h, (const Fr*)w); }
// witness map: out_h has m + 1 entries; evals (optional, may be NULL) get aA/aB/aC on S (m + 1 entries each)
void orc_h(void* h, const u64* w, u64* out_h, u64* evA, u64* evB, u64* evC) {
    Circuit* c = (Circuit*)h;
    std::vector<Fr> aH, a, b, cc;
    witness_map(*c, (const Fr*)w, aH, evA ? &a : nullptr, evB ? &b : nullptr, evC ? &cc : nullptr);
    memcpy(out_h, aH.data(), 32 * aH.size());
    if (evA) memcpy(evA, a.data(), 32 * a.size());
    if (evB) memcpy(evB, b.data(), 32 * b.size());
    if (evC) memcpy(evC, cc.data(), 32 * cc.size());
}

// ---- keys
void* orc_genkey(void* circ, u64 seed_base) { init(); return genkey(*(Circuit*)circ, seed_base); }
void orc_key_free(void* k) { delete (Key*)k; }
// returns the serialised size; copies at most cap bytes
u64 orc_key_serialize(void* k, int fmt, uint8_t* buf, u64 cap) {
    std::string s = serialize_key(*(Key*)k, fmt);
    if (buf) memcpy(buf, s.data(), std::min<u64>(cap, s.size()));
    return s.size();
}
u64 orc_key_vk_json(void* k, char* buf, u64 cap) {
    std::string s = vk_json(*(Key*)k);
    if (buf && cap) { u64 n = std::min<u64>(cap - 1, s.size()); memcpy(buf, s.data(), n); buf[n] = 0; }
    return s.size();
}
// out: domain_size, nA, nB, nH, nL
void orc_key_dims(void* k, u64* out) {
    Key* K = (Key*)k;
    out[0] = K->domain_size; out[1] = K->A_idx.size(); out[2] = K->B_idx.size(); out[3] = K->H.size(); out[4] = K->L.size();
}

// proof points come back affine in STANDARD form (A_xy[8], B_xy[16] = x.a,x.b,y.a,y.b, C_xy[8]); r,s may be NULL
void orc_prove(void* k, void* circ, const u64* w, const u64* r_mont, const u64* s_mont, unsigned window, u64* A_xy, u64* B_xy,
               u64* C_xy, uint8_t* inf) {
    Proof P = prove(*(Key*)k, *(Circuit*)circ, (const Fr*)w, (const Fr*)r_mont, (const Fr*)s_mont, window);
    P.A.x.to_std(A_xy); P.A.y.to_std(A_xy + 4);
    P.B.x.a.to_std(B_xy); P.B.x.b.to_std(B_xy + 4); P.B.y.a.to_std(B_xy + 8); P.B.y.b.to_std(B_xy + 12);
    P.C.x.to_std(C_xy); P.C.y.to_std(C_xy + 4);
    inf[0] = P.A.inf; inf[1] = P.B.inf; inf[2] = P.C.inf;
}
static Proof proof_from_std(const u64* A_xy, const u64* B_xy, const u64* C_xy, const uint8_t* inf) {
    Proof P;
    P.A = AffG1{Fq::from_std(A_xy), Fq::from_std(A_xy + 4), (bool)inf[0]};
    P.B = AffG2{Fq2{Fq::from_std(B_xy), Fq::from_std(B_xy + 4)}, Fq2{Fq::from_std(B_xy + 8), Fq::from_std(B_xy + 12)}, (bool)inf[1]};
    P.C = AffG1{Fq::from_std(C_xy), Fq::from_std(C_xy + 4), (bool)inf[2]};
    return P;
}
int orc_check_in_exponent(void* k, void* circ, const u64* w, const u64* r_mont, const u64* s_mont, const u64* A_xy,
                          const u64* B_xy, const u64* C_xy, const uint8_t* inf) {
    return check_in_exponent(*(Key*)k, *(Circuit*)circ, (const Fr*)w, (const Fr*)r_mont, (const Fr*)s_mont,
                             proof_from_std(A_xy, B_xy, C_xy, inf));
}
u64 orc_proof_json(const u64* A_xy, const u64* B_xy, const u64* C_xy, const uint8_t* inf, const u64* inputs_mont, u64 n_inputs,
                   char* buf, u64 cap) {
    init();
    std::string s = proof_json(proof_from_std(A_xy, B_xy, C_xy, inf), (const Fr*)inputs_mont, n_inputs);
    if (buf && cap) { u64 n = std::min<u64>(cap - 1, s.size()); memcpy(buf, s.data(), n); buf[n] = 0; }
    return s.size();
}

}  // extern "C"
