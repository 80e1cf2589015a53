"""Regenerates tests/golden/golden.json from the CPU oracle (run here, in the build container; the result is committed).
The reference ships no golden vector for H coefficients, MSM outputs or proofs, so these are the
oracle's outputs on seeded inputs, cross-checked in the exponent and by the pairing verifier that the reference's own
known-answer vector (tests/kat.py) pins."""
import hashlib
import json
import os
import sys
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import orc  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    out = {"circuits": {}, "msm": {}, "ntt": {}, "sha512_rng_fr": [hex(v) for v in orc.arr_to_ints(orc.fr_from_mont(orc.sha512_rng_fr(0, 4)))]}
    for name, mk in (("field_30_3", lambda: orc.Circuit.synth_field(30, 3)), ("field_1000_1", lambda: orc.Circuit.synth_field(1000, 1)),
                     ("binary_1000_10", lambda: orc.Circuit.synth_binary(1000, 10))):
        c = mk(); w = c.witness(); k = orc.Key(c)
        pr = k.prove(w)
        assert k.check_in_exponent(w, pr) == 0
        rs = orc.sha512_rng_fr(1 << 48, 2)
        pr_rs = k.prove(w, rs[0], rs[1])
        assert k.check_in_exponent(w, pr_rs, rs[0], rs[1]) == 0
        out["circuits"][name] = {
            "nc": c.nc, "ni": c.ni, "nv": c.nv, "m": c.m, "witness_sha256": sha(w), "h_sha256": sha(c.h_poly(w)[:c.m]),
            "pk_mcl_nozk_sha256": hashlib.sha256(k.serialize(orc.FMT_MCL_NOZK)).hexdigest(),
            "proof_json": orc.proof_json(pr, w[1:1 + c.ni]), "proof_rs_json": orc.proof_json(pr_rs, w[1:1 + c.ni]),
            "vk_json": k.vk_json(),
        }
    for n in (1, 33, 1000):
        kk = orc.sha512_rng_fr((1 << 40) + n, n); s = orc.sha512_rng_fr(1000 + n, n)
        b1, i1 = orc.g1_fixed_base(kk); b2, i2 = orc.g2_fixed_base(kk)
        r1, _ = orc.msm_g1(b1, i1, s); r2, _ = orc.msm_g2(b2, i2, s)
        out["msm"][str(n)] = {"g1": [hex(v) for v in orc.arr_to_ints(orc.fq_from_mont(r1))], "g2": [hex(v) for v in orc.arr_to_ints(orc.fq_from_mont(r2))]}
    for logm in (3, 10):
        x = orc.sha512_rng_fr(7 * logm, 1 << logm)
        out["ntt"][str(logm)] = [sha(orc.ntt(x, logm, mode)) for mode in range(4)]
    with open(os.path.join(HERE, "golden.json"), "w") as fh:
        json.dump(out, fh, indent=1, sort_keys=True)


REFERENCE_CIRCUITS = {
    "sha256": "BASELINE configs[1]: sha256_full_gadget_512 on sha256('test')||sha256('test') (src/test/test_sha256_full_gadget.cpp:30-99)",
    "mimc_merkle": "BASELINE configs[0]: merkle_path_authenticator<MiMC_e7_hash_gadget> depth 29 (src/gadgets/merkle_tree.hpp:70-192), "
                   "after reproducing the depth-1 known answer of src/test/test_merkle_tree.cpp:56-62",
    "eddsa": "BASELINE configs[2]: jubjub::EdDSA (Pedersen-hashed message) on the known signature of 'abc' "
             "(src/test/jubjub/test_jubjub_eddsa.cpp:44-62)",
}


def make_reference_circuit_goldens():
    """tests/golden/circuit_<name>.json: circuits built by the reference's own gadget code (oracle/_ref/circuit_dump, compiled
    from /root/reference by oracle/ref/Makefile), proved by the oracle under the seeded key; tests/test_oracle.py checks that
    the reference prover binary returns the same proofs."""
    import orc_ref
    for name, what in REFERENCE_CIRCUITS.items():
        circ, w, out = orc_ref.reference_circuit(name)
        key = orc.Key(circ)
        proof = key.prove(w)
        g = {"circuit": what, "dump_stdout": out, "constraints": circ.nc, "variables": circ.nv, "inputs": circ.ni, "m": circ.m,
             "key_seed_base": 1 << 32, "proof_json": orc.proof_json(proof, w[1:1 + circ.ni]),
             "witness_sha256": hashlib.sha256(np.ascontiguousarray(w).tobytes()).hexdigest()}
        json.dump(g, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "circuit_%s.json" % name), "w"), indent=1)


if __name__ == "__main__":
    main()
    make_reference_circuit_goldens()
