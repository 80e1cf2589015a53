// oracle/_ref harness — TEST INFRASTRUCTURE ONLY.  Instantiates the REFERENCE's own prover templates
// (libsnark::r1cs_gg_ppzksnark_zok_prover<libff::alt_bn128_pp>, src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:451-550)
// from the sources under /root/reference, does the context set-up nobody in the reference does, and
// runs one problem file written by tests/orc_ref.py:
//   u64 nc, ni, nv, n_coeffs; coeffs (32 B mont each); for A, B, C: u64 nnz, row_ptr[nc+1] u64, col[nnz] u32, cidx[nnz] u32;
//   witness (nv+1) x 32 B mont; u64 pk_len; pk bytes (ALT_NOZK layout = the reference's own operator<< under these flags)
// Output file: A.x A.y | B.x.c0 B.x.c1 B.y.c0 B.y.c1 | C.x C.y as 32-byte standard-form little-endian integers, then
// u64 m, then aH[0..m) (mont), then double seconds (the fastest prover call, as "Call to r1cs_gg_ppzksnark_zok_prover"),
// then u64 repeats and the duration of every repeat in call order (doubles), so that callers can report mean / median.
#include <fstream>
#include <iostream>
#include <sstream>
#include <chrono>
#include <omp.h>
#include <libff/algebra/curves/alt_bn128/alt_bn128_pp.hpp>
#include <libff/common/rng.hpp>
#include <libff/algebra/scalar_multiplication/multiexp.hpp>
#include <libfqfft/evaluation_domain/domains/basic_radix2_domain_aux.hpp>
#include <libfqfft/evaluation_domain/domains/recursive_domain.hpp>
#include "r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.hpp"

typedef libff::alt_bn128_pp ppT;
typedef libff::Fr<ppT> FieldT;
using namespace libsnark;

template <typename T> static void rd(std::istream& in, T& v) { in.read((char*)&v, sizeof(T)); }
static FieldT rd_fr(std::istream& in) { FieldT f; in.read((char*)f.mont_repr.data, 32); return f; }
template <typename F> static void wr_std(std::ostream& out, const F& f) { auto b = f.as_bigint(); out.write((const char*)b.data, 32); }

// --msm: n bases k_i * G by the reference's fixed-base machinery, then `reps` timed calls of its multi_exp
template <typename G>
static void msm_times(size_t n, const std::vector<FieldT>& ks, const std::vector<FieldT>& ss, int reps, int mix, std::vector<double>& times) {
    const size_t sb = FieldT::size_in_bits();
    libsnark::Config config;
    std::vector<libff::bigint<FieldT::num_limbs>> scratch;
    const size_t window = libff::get_exp_window_size<G>(n);
    auto table = libff::get_window_table<G>(sb, window, G::one());
    std::vector<G> bases = libff::batch_exp<G, FieldT>(sb, window, table, ks);
    libff::batch_to_special<G>(bases);
    for (int r = 0; r < reps; ++r) {
        auto t0 = std::chrono::steady_clock::now();
        G res = mix ? libff::multi_exp_with_mixed_addition<G, FieldT, libff::multi_exp_method_BDLO12>(bases.begin(), bases.end(), ss.begin(), ss.end(), scratch, config)
                    : libff::multi_exp<G, FieldT, libff::multi_exp_method_BDLO12>(bases.begin(), bases.end(), ss.begin(), ss.end(), scratch, config);
        times.push_back(std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count());
        if (res.is_zero()) std::cerr << "zero result\n";
    }
}

int main(int argc, char** argv) {
    if (argc == 4 && std::string(argv[1]) == "--roundtrip-full") {
        // pins the byte layout of the UN-STRIPPED key: the reference's own operator>> for r1cs_gg_ppzksnark_zok_proving_key
        // (…zok.tcc:69-87) reads the file, its operator<< (…zok.tcc:54-66) writes it back; tests compare the bytes
        ppT::init_public_params();
        std::ifstream kin(argv[2], std::ios::binary);
        if (!kin) { std::cerr << "cannot open key file\n"; return 2; }
        r1cs_gg_ppzksnark_zok_proving_key<ppT> full;
        kin >> full;
        std::ofstream kout(argv[3], std::ios::binary);
        kout << full;
        std::cout << "A " << full.A_query.size() << " B " << full.B_query.size() << " of " << full.B_query.domain_size() << " H " << full.H_query.size()
                  << " L " << full.L_query.size() << "\n";
        return 0;
    }
    if (argc >= 5 && std::string(argv[1]) == "--msm") {
        // CPU column of the standalone MSM sweep (BASELINE.json configs[4]): the reference's own multi_exp (multiexp.tcc:595-627,
        // uniform scalars - the H-query call) or multi_exp_with_mixed_addition (:631-714, the 0/1 prefilter - the witness calls)
        // over n distinct bases k_i * G made by the reference's own fixed-base machinery (get_window_table / batch_exp,
        // multiexp.tcc:771-905).  ref_prove --msm <1|2> <log_n> <repeats> [mix: 0 uniform, 1 = 90 % of the scalars in {0,1}]
        const int group = atoi(argv[2]), logn = atoi(argv[3]), reps = atoi(argv[4]), mix = argc > 5 ? atoi(argv[5]) : 0;
        ppT::init_public_params();
        libff::inhibit_profiling_info = true;
        libff::inhibit_profiling_counters = true;
        const size_t n = (size_t)1 << logn;
        std::vector<FieldT> ks(n), ss(n);
        for (size_t i = 0; i < n; ++i) {
            ks[i] = libff::SHA512_rng<FieldT>(((uint64_t)1 << 40) + i);
            ss[i] = (mix && i % 10) ? ((i / 10 + i) & 1 ? FieldT::one() : FieldT::zero()) : libff::SHA512_rng<FieldT>(i);
        }
        std::vector<double> times;
        if (group == 1) msm_times<libff::G1<ppT>>(n, ks, ss, reps, mix, times); else msm_times<libff::G2<ppT>>(n, ks, ss, reps, mix, times);
        std::cout << "{\"group\": \"G" << group << "\", \"log_n\": " << logn << ", \"mix\": " << mix << ", \"threads\": " << omp_get_max_threads()
                  << ", \"seconds\": [";
        for (size_t i = 0; i < times.size(); ++i) std::cout << (i ? ", " : "") << times[i];
        std::cout << "]}\n";
        return 0;
    }
    if (argc < 3) { std::cerr << "usage: ref_prove <problem> <out> [repeats [budget_seconds min_repeats]] | --roundtrip-full <pk_in> <pk_out> | --msm <1|2> <log_n> <repeats> [mix]\n"; return 2; }
    const int repeats = argc > 3 ? atoi(argv[3]) : 1;
    // optional bound for benchmark runs: stop before `repeats` once the prover calls have used budget_seconds, but not
    // before min_repeats of them ran (the number that ran is written to the output)
    const double budget = argc > 5 ? atof(argv[4]) : 0.0;
    const int min_repeats = argc > 5 ? atoi(argv[5]) : repeats;
    ppT::init_public_params();
    libff::inhibit_profiling_info = true;
    libff::inhibit_profiling_counters = true;
    std::ifstream in(argv[1], std::ios::binary);
    if (!in) { std::cerr << "cannot open problem file\n"; return 2; }
    uint64_t nc, ni, nv, ncoef;
    rd(in, nc); rd(in, ni); rd(in, nv); rd(in, ncoef);
    std::vector<unsigned int> cmap(ncoef);
    for (uint64_t i = 0; i < ncoef; ++i) cmap[i] = ConstantStorage<FieldT>::getInstance().add(rd_fr(in));
    r1cs_constraint_system<FieldT> cs;
    cs.primary_input_size = ni;
    cs.auxiliary_input_size = nv - ni;
    std::vector<std::vector<linear_combination_light<FieldT>>> rows(3);
    for (int k = 0; k < 3; ++k) {
        uint64_t nnz; rd(in, nnz);
        std::vector<uint64_t> rp(nc + 1); in.read((char*)rp.data(), 8 * (nc + 1));
        std::vector<uint32_t> col(nnz), ci(nnz);
        in.read((char*)col.data(), 4 * nnz); in.read((char*)ci.data(), 4 * nnz);
        rows[k].resize(nc);
        for (uint64_t r = 0; r < nc; ++r)
            for (uint64_t j = rp[r]; j < rp[r + 1]; ++j) {
                linear_term_light<FieldT> t; t.index = col[j]; t.coeff_index = cmap[ci[j]];
                rows[k][r].terms.push_back(t);
            }
    }
    for (uint64_t r = 0; r < nc; ++r)
        cs.constraints.emplace_back(new r1cs_constraint_light<FieldT>(rows[0][r], rows[1][r], rows[2][r]));
    std::vector<FieldT> w(nv + 1);
    for (auto& x : w) x = rd_fr(in);
    uint64_t pk_len; rd(in, pk_len);
    std::string pk_bytes(pk_len, 0); in.read(&pk_bytes[0], pk_len);
    r1cs_gg_ppzksnark_zok_proving_key_nozk<ppT> pk;
    { std::istringstream ss(pk_bytes); ss >> pk; }

    // the set-up the reference leaves to its (absent) caller: domain, constraint system pointer, m+1 buffers
    libsnark::Config config;
    uint64_t m = 1; while (m < nc + ni + 1) m <<= 1;
    ProverContext<ppT> ctx(pk);
    ctx.config = config;
    ctx.constraint_system = &cs;
    ctx.domain.reset(new libfqfft::recursive_domain<FieldT>(m, config));
    ctx.aA.resize(m + 1); ctx.aB.resize(m + 1); ctx.aH.resize(m + 1);
    ctx.scratch_exponents.resize(m + nv + 8);

    // REF_PHASES=1: keep libff's block timers running (enter_block / leave_block, profiling.cpp) and print the mean time of the
    // prover's own blocks, so that the GPU phase table can be laid beside the reference's under the reference's block names
    const bool want_phases = getenv("REF_PHASES") != nullptr;
    if (want_phases) { libff::inhibit_profiling_counters = false; libff::clear_profiling_counters(); }
    double best = 1e30;
    std::vector<double> times;
    double used = 0;
    r1cs_gg_ppzksnark_zok_proof<ppT> proof;
    for (int i = 0; i < repeats; ++i) {
        auto t0 = std::chrono::steady_clock::now();
        proof = r1cs_gg_ppzksnark_zok_prover<ppT>(ctx, w);
        double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        if (dt < best) best = dt;
        times.push_back(dt);
        used += dt;
        if (budget > 0 && used > budget && i + 1 >= min_repeats) break;
    }
    if (want_phases) {
        const char* blocks[] = {"Call to r1cs_gg_ppzksnark_zok_prover", "Compute the polynomial H", "Compute the proof", "Compute evaluation to A-query",
                                "Compute evaluation to B-query", "Compute evaluation to H-query", "Compute evaluation to L-query"};
        std::cerr << "{";
        bool first = true;
        for (const char* b : blocks) {
            auto it = libff::cumulative_times.find(b);
            if (it == libff::cumulative_times.end()) continue;
            std::cerr << (first ? "" : ", ") << "\"" << b << "\": " << (double)it->second * 1e-6 / (double)times.size();
            first = false;
        }
        std::cerr << "}\n";
    }
    std::ofstream out(argv[2], std::ios::binary);
    auto A = proof.g_A; A.to_affine_coordinates();
    auto B = proof.g_B; B.to_affine_coordinates();
    auto C = proof.g_C; C.to_affine_coordinates();
    wr_std(out, A.X); wr_std(out, A.Y);
    wr_std(out, B.X.c0); wr_std(out, B.X.c1); wr_std(out, B.Y.c0); wr_std(out, B.Y.c1);
    wr_std(out, C.X); wr_std(out, C.Y);
    out.write((const char*)&m, 8);
    for (uint64_t i = 0; i < m; ++i) out.write((const char*)ctx.aH[i].mont_repr.data, 32);
    out.write((const char*)&best, 8);
    const uint64_t nt = times.size();
    out.write((const char*)&nt, 8);
    out.write((const char*)times.data(), 8 * nt);
    return 0;
}
