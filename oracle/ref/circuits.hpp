// Circuit builders shared by oracle/_ref/circuit_dump and oracle/_ref/shim_prove — TEST INFRASTRUCTURE ONLY.  Each builds one
// BASELINE.json circuit with the REFERENCE's own gadget code, compiled from the sources where they lie under /root/reference,
// generates its witness, and hands the protoboard to g_emit (the dumper writes the problem file, the shim harness proves it).
//   sha256       BASELINE.json configs[1]: sha256_full_gadget_512 (src/gadgets/sha256_full.hpp:69-163) on the
//                input of src/test/test_sha256_full_gadget.cpp:30-99: block = sha256("test") || sha256("test"),
//                digest bits made the primary input; prints the digest (KAT D294F6E5..., :48-52)
//   mimc_merkle  BASELINE.json configs[0]: merkle_path_authenticator<MiMC_e7_hash_gadget> of depth 29
//                (src/gadgets/merkle_tree.hpp:70-192, src/gadgets/mimc.hpp), root public.  First reproduces
//                the reference's depth-1 known answer (src/test/test_merkle_tree.cpp:56-62), which pins the
//                MiMC round constants (and with them the sha3.h stand-in's padding flavour).
//   eddsa        BASELINE.json configs[2], see build_eddsa
#pragma once
#include <fstream>
#include <iostream>
#include <cstring>
#include <initializer_list>
#include <omp.h>
#include "gadgets/sha256_full.hpp"
#include "gadgets/mimc.hpp"
#include "gadgets/merkle_tree.hpp"
#include "jubjub/eddsa.hpp"

using namespace ethsnarks;
using libsnark::ConstantStorage;


static int (*g_emit)(ethsnarks::ProtoboardT& pb, const char* arg) = nullptr;

static int build_sha256(const char* path) {
    using namespace libsnark;
    uint8_t input_buffer[SHA256_block_size_bytes], output_digest[SHA256_digest_size_bytes];
    SHA256_CTX ctx;
    SHA256_Init(&ctx); SHA256_Update(&ctx, "test", 4); SHA256_Final(input_buffer, &ctx);
    memcpy(&input_buffer[SHA256_digest_size_bytes], input_buffer, SHA256_digest_size_bytes);
    SHA256_Init(&ctx); SHA256_Update(&ctx, input_buffer, sizeof(input_buffer)); SHA256_Final(output_digest, &ctx);

    ProtoboardT pb;
    digest_variable<FieldT> full_output(pb, SHA256_digest_size, "full_output");   // allocated first: the primary input
    digest_variable<FieldT> left(pb, SHA256_digest_size, "left");
    digest_variable<FieldT> right(pb, SHA256_digest_size, "right");
    left.generate_r1cs_witness(bytes_to_bv(input_buffer, SHA256_digest_size_bytes));
    right.generate_r1cs_witness(bytes_to_bv(&input_buffer[SHA256_digest_size_bytes], SHA256_digest_size_bytes));
    block_variable<FieldT> full_input(pb, left, right, "full_input");
    sha256_full_gadget_512 full_gadget(pb, full_input, full_output, "full_gadget");
    full_gadget.generate_r1cs_constraints();
    full_gadget.generate_r1cs_witness();
    pb.set_input_sizes(SHA256_digest_size);
    uint8_t got[SHA256_digest_size_bytes];
    bv_to_bytes(full_output.get_digest(), got);
    std::cout << "digest ";
    for (size_t i = 0; i < SHA256_digest_size_bytes; ++i) printf("%02X", got[i]);
    std::cout << (memcmp(got, output_digest, sizeof got) == 0 ? " == openssl-style SHA256" : " MISMATCH") << "\n";
    g_emit(pb, path);
    return pb.is_satisfied() && memcmp(got, output_digest, sizeof got) == 0 ? 0 : 1;
}

// BASELINE.json configs[2]: Baby Jubjub EdDSA verification with the Pedersen-hashed message (jubjub::EdDSA,
// src/jubjub/eddsa.hpp:55-82 is the construction followed) on the reference's known signature of "abc"
// (src/test/jubjub/test_jubjub_eddsa.cpp:44-62); the public key A is the primary input.
static int build_eddsa(const char* path_out) {
    using namespace ethsnarks::jubjub;
    Params params;
    const char* msg = "abc";
    const auto msg_bits = bytes_to_bv((const uint8_t*)msg, strlen(msg));
    const EdwardsPoint A(FieldT("333671881179914989291633188949569309119725676183802886621140166987382124337"),
                         FieldT("4050436616325076046600891135828313078248584449767955905006778857958871314574"));
    const EdwardsPoint R(FieldT("21473010389772475573783051334263374448039981396476357164143587141689900886674"),
                         FieldT("11330590229113935667895133446882512506792533479705847316689101265088791098646"));
    const FieldT sig_s("21807294168737929637405719327036335125520717961882955117047593281820367379946");
    const EdwardsPoint B(params.Gx, params.Gy);
    ProtoboardT pb;
    const auto A_var = A.as_VariablePointT(pb, "A");          // allocated first: the primary input (2 field elements)
    const auto msg_var_bits = make_var_array(pb, msg_bits.size(), "msg_var_bits");
    msg_var_bits.fill_with_bits(pb, msg_bits);
    const auto s_var_bits = make_var_array(pb, FieldT::size_in_bits(), "s_var_bits");
    s_var_bits.fill_with_bits_of_field_element(pb, sig_s);
    EdDSA the_gadget(pb, params, B, A_var, R.as_VariablePointT(pb, "R"), s_var_bits, msg_var_bits, "the_gadget");
    the_gadget.generate_r1cs_constraints();
    the_gadget.generate_r1cs_witness();
    pb.set_input_sizes(2);
    g_emit(pb, path_out);
    return pb.is_satisfied() ? 0 : 1;
}

// depth-1 known answer of the reference's own test; returns true if the gadget reproduces the root
static bool merkle_kat() {
    auto left = FieldT("3703141493535563179657531719960160174296085208671919316200479060314459804651");
    auto right = FieldT("134551314051432487569247388144051420116740427803855572138106146683954151557");
    auto root = FieldT("3075442268020138823380831368198734873612490112867968717790651410945045657947");
    ProtoboardT pb;
    VariableArrayT address_bits; address_bits.allocate(pb, 1, "address_bits"); pb.val(address_bits[0]) = 1;
    VariableArrayT path; path.allocate(pb, 1, "path"); pb.val(path[0]) = left;
    VariableT leaf; leaf.allocate(pb, "leaf"); pb.val(leaf) = right;
    VariableT expected_root; expected_root.allocate(pb, "expected_root"); pb.val(expected_root) = root;
    merkle_path_authenticator<MiMC_e7_hash_gadget> auth(pb, 1, address_bits, merkle_tree_IVs(pb), leaf, expected_root, path, "authenticator");
    auth.generate_r1cs_witness();
    auth.generate_r1cs_constraints();
    return auth.is_valid() && pb.is_satisfied();
}

static int build_mimc_merkle(const char* path_out) {
    if (!merkle_kat()) { std::cerr << "MiMC Merkle known answer NOT reproduced\n"; return 1; }
    std::cout << "merkle KAT (test_merkle_tree.cpp:56-62) reproduced\n";
    const size_t depth = 29;
    ProtoboardT pb;
    VariableT expected_root; expected_root.allocate(pb, "expected_root");       // the primary input
    VariableArrayT address_bits; address_bits.allocate(pb, depth, "address_bits");
    VariableArrayT path; path.allocate(pb, depth, "path");
    VariableT leaf; leaf.allocate(pb, "leaf");
    pb.val(leaf) = FieldT(7) ^ 1000ul;                                          // deterministic, full-width field elements
    for (size_t i = 0; i < depth; ++i) {
        pb.val(path[i]) = FieldT(7) ^ (unsigned long)(2000 + i);
        pb.val(address_bits[i]) = (0x15a3c7e9u >> i) & 1;
    }
    merkle_path_authenticator<MiMC_e7_hash_gadget> auth(pb, depth, address_bits, merkle_tree_IVs(pb), leaf, expected_root, path, "authenticator");
    auth.generate_r1cs_witness();
    pb.val(expected_root) = pb.val(auth.result());
    auth.generate_r1cs_constraints();
    pb.set_input_sizes(1);
    std::cout << "root "; pb.val(expected_root).as_bigint().print();
    g_emit(pb, path_out);
    return auth.is_valid() && pb.is_satisfied() ? 0 : 1;
}


// builds `name` and emits it with `arg`; the padding flavour of the absent SHA3IUF fork is decided by the reference's own
// known answer
static int build_circuit(const char* name, const char* arg) {
    if (!strcmp(name, "sha256")) return build_sha256(arg);
    if (!strcmp(name, "eddsa")) return build_eddsa(arg);
    if (!strcmp(name, "mimc_merkle")) {
        for (unsigned char pad : {(unsigned char)0x01, (unsigned char)0x06}) {
            b200_shim_sha3_pad() = pad;
            if (merkle_kat()) { std::cout << "sha3 padding byte " << (int)pad << "\n"; return build_mimc_merkle(arg); }
        }
        std::cerr << "no padding flavour reproduces the reference's Merkle known answer\n";
        return 1;
    }
    std::cerr << "unknown circuit\n";
    return 2;
}
