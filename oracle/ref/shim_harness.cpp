// oracle/_ref/shim_prove — TEST INFRASTRUCTURE ONLY.  Drives the compiled C++ drop-in (shim/ethsnarks_b200.{hpp,cpp}: the
// reference's entry points with the reference's signatures, on libb200snark.so) the way a caller of src/stubs.hpp would, on
// protoboards built by the REFERENCE's own gadget code (circuits.hpp), and checks the result with the REFERENCE's own
// verifier (r1cs_gg_ppzksnark_zok_verifier_strong_IC, …zok.tcc:661-668) and JSON readers (src/import.cpp).
//
//   shim_prove prove   <sha256|mimc_merkle|eddsa> <pk_file> <proof_out.json> [vk.json]
//       load_proving_key(pk_file) -> ProverContextT -> get_domain -> prove(context, pb) -> proof JSON written to proof_out;
//       a second proof through libsnark::r1cs_gg_ppzksnark_zok_prover<ppT>(context, pb.values) must give the same JSON;
//       with vk.json: stub_verify's steps (vk_from_json, proof_from_json, verifier_strong_IC) on that proof, and on the
//       same proof with one input flipped (must be rejected)
//   shim_prove genkeys <sha256|mimc_merkle|eddsa> <pk_out> <vk_out.json> <proof_out.json>
//       stub_genkeys_from_pb on the device, then the reference's loader reads the pk file back, the GPU proves with it and
//       the reference verifier accepts the proof under the generated vk
//   shim_prove verify  <vk.json> <proof.json>
//       the drop-in's stub_main_verify (src/stubs.cpp:90-132) with the same argv convention; host only, no GPU needed
//   shim_prove selftest <sha256|mimc_merkle|eddsa> - -
//       the drop-in's stub_test_proof_verify (src/stubs.cpp:135-149), the call the reference's gadget tests end with: keys and
//       proof from the device, the reference's verifier on the host; also run on a protoboard whose first public input was
//       changed after the witness was made (must be rejected)
// Prints one "key: value" line per fact; exit code 0 only if every check holds.
#include "circuits.hpp"
#include "ethsnarks_b200.hpp"
#include "import.hpp"
#include "export.hpp"
#include <chrono>

using namespace ethsnarks;

static std::string g_mode, g_a, g_b, g_c, g_d;

static std::string slurp(const std::string& path) {
    std::ifstream in(path, std::ios::binary);
    std::stringstream ss;
    ss << in.rdbuf();
    return ss.str();
}
// the steps of ethsnarks::stub_verify (src/stubs.cpp:18-34) on JSON strings
static bool reference_verify(const std::string& vk_json, const std::string& proof_json) {
    std::stringstream vs, ps;
    vs << vk_json;
    ps << proof_json;
    auto vk = vk_from_json(vs);
    auto pair = proof_from_json(ps);
    return libsnark::r1cs_gg_ppzksnark_zok_verifier_strong_IC<ppT>(vk, pair.first, pair.second);
}

static int run(ProtoboardT& pb, const char*) {
    const auto& cs = pb.constraint_system;
    std::cout << "constraints: " << cs.num_constraints() << "\nvariables: " << cs.num_variables() << "\ninputs: " << cs.num_inputs() << "\n";
    int bad = 0;
    if (g_mode == "selftest") {
        const bool ok = stub_test_proof_verify(pb);
        std::cout << "stub_test_proof_verify: " << ok << "\n";
        pb.values[0] += FieldT::one();               // variable 1 = the first public input (a protoboard cannot be copied)
        const bool rejected = !stub_test_proof_verify(pb);
        pb.values[0] -= FieldT::one();
        std::cout << "stub_test_proof_verify_rejects_wrong_input: " << rejected << "\n";
        return !ok + !rejected;
    }
    std::string pk_file = g_a, vk_json, proof_out;
    if (g_mode == "genkeys") {
        auto t0 = std::chrono::steady_clock::now();
        const int rc = stub_genkeys_from_pb(pb, g_a.c_str(), g_b.c_str());
        std::cout << "genkeys_rc: " << rc << "\ngenkeys_seconds: " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() << "\n";
        if (rc) return 1;
        vk_json = slurp(g_b);
        proof_out = g_c;
    } else {
        proof_out = g_b;
        if (!g_c.empty()) vk_json = slurp(g_c);
    }
    ProvingKeyT pk = load_proving_key(pk_file.c_str());              // the reference's operator>> reads the file
    ProverContextT context(pk);
    context.config = libsnark::Config();
    context.constraint_system = &pb.constraint_system;
    context.domain = get_domain(pb, pk, context.config);
    std::cout << "domain_size: " << context.domain->m << "\n";
    auto t0 = std::chrono::steady_clock::now();
    const std::string json = prove(context, pb);                     // first call uploads key + circuit
    std::cout << "first_prove_seconds: " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() << "\n";
    t0 = std::chrono::steady_clock::now();
    const std::string json2 = prove(context, pb);
    std::cout << "prove_seconds: " << std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() << "\n";
    for (const auto& ph : b200_last_phases()) std::cout << "phase_ms " << ph.first << ": " << ph.second << "\n";
    // the template entry point callers of libsnark use directly (…zok.tcc:451-452), specialised in ethsnarks_b200.hpp
    auto proof = libsnark::r1cs_gg_ppzksnark_zok_prover<ppT>(context, pb.values);
    auto primary = pb.primary_input();
    const std::string json3 = proof_to_json(proof, primary);
    std::cout << "repeat_identical: " << (json == json2 && json == json3) << "\n";
    bad += !(json == json2 && json == json3);
    { std::ofstream out(proof_out, std::ios::binary); out << json; }
    if (!vk_json.empty()) {
        const bool ok = reference_verify(vk_json, json);
        std::cout << "reference_verifier_accepts: " << ok << "\n";
        bad += !ok;
        // negative control: the same proof under a changed public input
        auto tampered = primary;
        tampered[0] += FieldT::one();
        const std::string tj = proof_to_json(proof, tampered);
        const bool rejected = !reference_verify(vk_json, tj);
        std::cout << "reference_verifier_rejects_tampered: " << rejected << "\n";
        bad += !rejected;
    }
    b200_release(pk);
    return bad;
}

int main(int argc, char** argv) {
    if (argc == 4 && std::string(argv[1]) == "verify") {
        const char* av[3] = {"verify", argv[2], argv[3]};
        const int rc = stub_main_verify("shim_prove", 3, av);
        std::cout << "stub_main_verify_rc: " << rc << "\n";
        return rc;
    }
    if (argc < 5) { std::cerr << "usage: shim_prove prove <circuit> <pk> <proof_out> [vk.json] | genkeys <circuit> <pk_out> <vk_out> <proof_out>\n"; return 2; }
    g_mode = argv[1];
    g_a = argv[3]; g_b = argv[4];
    if (argc > 5) g_c = argv[5];
    ppT::init_public_params();
    libff::inhibit_profiling_info = true;
    libff::inhibit_profiling_counters = true;
    g_emit = run;
    const int rc = build_circuit(argv[2], "");
    b200_shutdown_all();
    return rc;
}
