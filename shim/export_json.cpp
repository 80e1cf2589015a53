// export_json.cpp — the JSON writers declared in the reference's src/export.hpp that the drop-in needs, for builds that
// cannot compile the reference's own src/export.cpp.  That file also holds the mcl key converters (export.cpp:267-408) and
// therefore needs the mcl curve code and gmpxx, which this image does not have; a maintainer whose tree
// builds export.cpp links that file instead and drops this one.  What is restated here, with the same output bytes, is the
// CURVE_ALT_BN128 branch of:
//   HexStringFromBigint            src/export.cpp:20-33      lower-case hex, no leading zeros, "0" for zero
//   outputPointG1AffineAsHex       src/export.cpp:56-68
//   outputPointG2AffineAsHex       src/export.cpp:71-97      printed as [x.c1, x.c0], [y.c1, y.c0]
//   proof_to_json                  src/export.cpp:100-121
//   vk2json / vk2json_file         src/export.cpp:124-155
// tests/golden/*.json (pinned on the reference's VK_STATIC / PROOF_STATIC vectors) fix the format.
#include "export.hpp"
#include <fstream>
#include <sstream>

namespace ethsnarks {

std::string HexStringFromBigint(LimbT x) {
    static const char* digits = "0123456789abcdef";
    std::string s;
    bool started = false;
    for (int limb = (int)LimbT::N - 1; limb >= 0; --limb)
        for (int nib = 15; nib >= 0; --nib) {
            const unsigned d = (unsigned)((x.data[limb] >> (4 * nib)) & 0xf);
            if (d || started) { s.push_back(digits[d]); started = true; }
        }
    return started ? s : std::string("0");
}

std::string outputPointG1AffineAsHex(G1T p) {
    p.to_affine_coordinates();
    return "\"0x" + HexStringFromBigint(p.X.as_bigint()) + "\", \"0x" + HexStringFromBigint(p.Y.as_bigint()) + "\"";
}

std::string outputPointG2AffineAsHex(G2T p) {
    // the reference normalises only when both components of Z are non-zero (export.cpp:76-78): an affine point (Z = 1) is
    // printed as it stands
    if (!p.Z.c0.is_zero() && !p.Z.c1.is_zero()) p.to_affine_coordinates();
    return "[\"0x" + HexStringFromBigint(p.X.c1.as_bigint()) + "\", \"0x" + HexStringFromBigint(p.X.c0.as_bigint()) + "\"],\n [\"0x" +
           HexStringFromBigint(p.Y.c1.as_bigint()) + "\", \"0x" + HexStringFromBigint(p.Y.c0.as_bigint()) + "\"]";
}

std::string proof_to_json(ProofT& proof, PrimaryInputT& input) {
    std::stringstream ss;
    ss << "{\n";
    ss << " \"A\" :[" << outputPointG1AffineAsHex(proof.g_A) << "],\n";
    ss << " \"B\"  :[" << outputPointG2AffineAsHex(proof.g_B) << "],\n";
    ss << " \"C\"  :[" << outputPointG1AffineAsHex(proof.g_C) << "],\n";
    ss << " \"input\" :[";
    for (size_t i = 0; i < input.size(); ++i) {
        ss << "\"0x" << HexStringFromBigint(input[i].as_bigint()) << "\"";
        if (i + 1 < input.size()) ss << ", ";
    }
    ss << "]\n}";
    return ss.str();
}

std::string vk2json(VerificationKeyT& vk) {
    std::stringstream ss;
    ss << "{\n";
    ss << " \"alpha\" :[" << outputPointG1AffineAsHex(vk.alpha_g1) << "],\n";
    ss << " \"beta\"  :[" << outputPointG2AffineAsHex(vk.beta_g2) << "],\n";
    ss << " \"gamma\" :[" << outputPointG2AffineAsHex(vk.gamma_g2) << "],\n";
    ss << " \"delta\" :[" << outputPointG2AffineAsHex(vk.delta_g2) << "],\n";
    ss << "\"gammaABC\" :[[" << outputPointG1AffineAsHex(vk.gamma_ABC_g1.first) << "]";
    for (size_t i = 0; i < vk.gamma_ABC_g1.rest.indices.size(); ++i) ss << ",[" << outputPointG1AffineAsHex(vk.gamma_ABC_g1.rest.values[i]) << "]";
    ss << "]}";
    return ss.str();
}

void vk2json_file(VerificationKeyT& vk, const std::string& path) {
    std::ofstream fh(path, std::ios::binary);
    fh << vk2json(vk);
}

}  // namespace ethsnarks
