// ethsnarks_b200.hpp — the C++ side of the drop-in: the reference's own prover entry points, with the reference's own
// signatures and types, implemented on libb200snark.so (the C ABI in include/b200snark.h).
//
// A caller of the reference links this translation unit INSTEAD of src/stubs.cpp and keeps its code unchanged:
//
//   reference symbol (file:line)                                                  here
//   ethsnarks::load_proving_key(const char*)            src/stubs.hpp:18, stubs.cpp:36-39     same signature, same result
//   ethsnarks::get_domain(pb, pk, config)               src/stubs.hpp:21, stubs.cpp:61-75     same signature, same result
//   ethsnarks::prove(ProverContextT&, ProtoboardT&)     src/stubs.hpp:19, stubs.cpp:42-47     same signature; proof from the GPU
//   ethsnarks::stub_genkeys_from_pb(pb, pk, vk)         src/stubs.hpp:16, stubs.cpp:77-87     same signature; keys from the GPU generator,
//                                                                                           written in the reference's file formats
//   ethsnarks::stub_test_proof_verify(const ProtoboardT&) src/stubs.hpp:14, stubs.cpp:135-149     same signature; generator and prover on
//                                                                                           the GPU, the reference's verifier on the host
//   ethsnarks::stub_verify / stub_main_verify           src/stubs.hpp:10-12, stubs.cpp:16-34,90-132  same signatures; verification stays on
//                                                                                           the host (the reference's import.cpp + verifier)
//   libsnark::r1cs_gg_ppzksnark_zok_prover<ppT>(ProverContext<ppT>&, const std::vector<Fr<ppT>>&)
//                                                       src/r1cs_gg_ppzksnark_zok/r1cs_gg_ppzksnark_zok.tcc:451-550
//                                                       explicit specialisation for ethsnarks::ppT declared below: every
//                                                       translation unit that includes this header calls the GPU prover
//
// Device state hangs off the caller's objects through a side table keyed by (&context.provingKey, constraint system): the
// first prove for a key uploads its queries and the circuit (CSR + the ConstantStorage table) and builds the window tables;
// later proves only move the witness.  ProverContext's aA/aB/aH/scratch_exponents/domain are not touched (the device owns
// its own buffers and domain), so a context that was never "set up"  works too.
//
// Devices: B200_DEVICES="0,1,2,3" (or b200_use_devices) makes the one context span several GPUs of the box - the key's
// queries are split 1/N per GPU and one prove runs on all of them (include/b200snark.h: b200_init with n_devices > 1).
#ifndef ETHSNARKS_B200_HPP_
#define ETHSNARKS_B200_HPP_

#include "ethsnarks.hpp"
#include <libfqfft/evaluation_domain/evaluation_domain.hpp>
#include <memory>
#include <string>
#include <vector>

namespace libsnark {
// the single call the GPU replaces (…zok.tcc:451-452)
template <>
r1cs_gg_ppzksnark_zok_proof<ethsnarks::ppT> r1cs_gg_ppzksnark_zok_prover<ethsnarks::ppT>(
    ProverContext<ethsnarks::ppT>& context, const std::vector<libff::Fr<ethsnarks::ppT>>& full_variable_assignment);
}  // namespace libsnark

namespace ethsnarks {

// ---- src/stubs.hpp:10-21, unchanged signatures (stub_genkeys<GadgetT> / stub_main_genkeys<GadgetT> are templates of that
// header over stub_genkeys_from_pb and work unchanged)
bool stub_verify(const char* vk_json, const char* proof_json);
int stub_main_verify(const char* prog_name, int argc, const char** argv);
bool stub_test_proof_verify(const ProtoboardT& in_pb);
int stub_genkeys_from_pb(ProtoboardT& pb, const char* pk_file, const char* vk_file);
ProvingKeyT load_proving_key(const char* pk_file);
std::string prove(ProverContextT& context, ProtoboardT& pb);
const std::shared_ptr<libfqfft::evaluation_domain<FieldT>> get_domain(ProtoboardT& pb, const ProvingKeyT& proving_key,
                                                                      const libsnark::Config& config);

// ---- additions (no counterpart in the reference)
// GPUs the process-wide context spans; call before the first prove (default: B200_DEVICES, else device 0)
void b200_use_devices(const std::vector<int>& device_ids);
// drops the device copy of a key (all circuits proved with it); the next prove uploads it again
void b200_release(const ProvingKeyT& proving_key);
// frees everything, including the context (also runs at exit)
void b200_shutdown_all();
// phase times of the last prove, named after the reference's enter_block names where one exists
std::vector<std::pair<std::string, float>> b200_last_phases();

}  // namespace ethsnarks

#endif
