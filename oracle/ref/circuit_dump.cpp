// oracle/_ref circuit dumper — TEST INFRASTRUCTURE ONLY.  Builds a circuit with the REFERENCE's own gadget code (circuits.hpp)
// and writes its constraint system and witness in the problem format of harness.cpp / tests/orc_ref.py (with pk_len = 0), so
// that the same R1CS + witness can be proved by the reference prover (ref_prove), the oracle and the GPU library.
//   circuit_dump sha256|mimc_merkle|eddsa <out>
#include "circuits.hpp"

template <typename T> static void wr(std::ostream& out, const T& v) { out.write((const char*)&v, sizeof(T)); }

static int dump(ProtoboardT& pb, const char* path) {
    const auto& cs = pb.constraint_system;
    const uint64_t nc = cs.constraints.size(), ni = pb.num_inputs(), nv = pb.num_variables();
    const auto& consts = ConstantStorage<FieldT>::getInstance().constants;
    const uint64_t ncoef = consts.size();
    std::ofstream out(path, std::ios::binary);
    wr(out, nc); wr(out, ni); wr(out, nv); wr(out, ncoef);
    for (const auto& c : consts) out.write((const char*)c.mont_repr.data, 32);
    for (int k = 0; k < 3; ++k) {
        std::vector<uint64_t> rp(nc + 1, 0);
        std::vector<uint32_t> col, ci;
        for (uint64_t r = 0; r < nc; ++r) {
            const auto lc = k == 0 ? cs.constraints[r]->getA() : k == 1 ? cs.constraints[r]->getB() : cs.constraints[r]->getC();
            for (const auto& t : lc.getTerms()) { col.push_back((uint32_t)t.index); ci.push_back((uint32_t)t.coeff_index); }
            rp[r + 1] = col.size();
        }
        const uint64_t nnz = col.size();
        wr(out, nnz);
        out.write((const char*)rp.data(), 8 * (nc + 1));
        out.write((const char*)col.data(), 4 * nnz);
        out.write((const char*)ci.data(), 4 * nnz);
    }
    // in this fork pb.values[0] is the constant 1 (protoboard.tcc:24-31), so pb.values IS the full assignment (1, x)
    for (const auto& v : pb.values) out.write((const char*)v.mont_repr.data, 32);
    const uint64_t pk_len = 0;
    wr(out, pk_len);
    std::cout << "constraints " << nc << " variables " << nv << " inputs " << ni << " satisfied " << pb.is_satisfied() << "\n";
    return 0;
}

int main(int argc, char** argv) {
    if (argc < 3) { std::cerr << "usage: circuit_dump sha256|mimc_merkle|eddsa <out>\n"; return 2; }
    ppT::init_public_params();
    libff::inhibit_profiling_info = true;
    libff::inhibit_profiling_counters = true;
    g_emit = dump;
    return build_circuit(argv[1], argv[2]);
}
