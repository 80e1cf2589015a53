"""CPU tests: the oracle against independent Python big-integer arithmetic, the reference's known-answer vector,
and the committed golden fixtures (tests/golden/golden.json, made by tests/golden/make_golden.py)."""
import copy
import hashlib
import json
import os
import numpy as np
import kat
import orc
import pairing
import pytest
import orc_ref

HERE = os.path.dirname(os.path.abspath(__file__))

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_constants_match_reference_init():
    c = orc.constants()   # mcl_bn128_init.cpp:41-89
    assert orc.limbs_to_int(c[0:4]) == orc.R_MOD and orc.limbs_to_int(c[4:8]) == orc.Q_MOD
    assert orc.limbs_to_int(c[16:20]) == 944936681149208446651664254269745548490766851729442924617792859073125903783
    assert orc.limbs_to_int(c[20:24]) == 3096616502983703923843567936837374451735540968419076528771170197431451843209
    assert int(c[28]) == 0xc2e1f593efffffff and int(c[29]) == 0x87d20782e4866389
    w = orc.from_mont(orc.limbs_to_int(c[24:28]), orc.R_MOD)
    assert pow(w, 1 << 28, orc.R_MOD) == 1 and pow(w, 1 << 27, orc.R_MOD) != 1


def test_field_ops_vs_python_ints():
    a = orc.rand_fr_mont(2000, 1); b = orc.rand_fr_mont(2000, 2)
    ai, bi = orc.arr_to_ints(a), orc.arr_to_ints(b)
    for F, p in (("fr", orc.R_MOD), ("fq", orc.Q_MOD)):
        rinv = pow(orc.MONT_R, -1, p)
        assert orc.arr_to_ints(getattr(orc, F + "_mul")(a, b)) == [x * y * rinv % p for x, y in zip(ai, bi)]
        assert orc.arr_to_ints(getattr(orc, F + "_add")(a, b)) == [(x + y) % p for x, y in zip(ai, bi)]
        assert orc.arr_to_ints(getattr(orc, F + "_sub")(a, b)) == [(x - y) % p for x, y in zip(ai, bi)]
        assert orc.arr_to_ints(getattr(orc, F + "_from_mont")(a)) == [x * rinv % p for x in ai]
        inv = orc.arr_to_ints(getattr(orc, F + "_inv")(a[:50]))
        assert all(x * y * rinv * rinv % p == 1 for x, y in zip(ai[:50], inv))


def test_sha512_rng_matches_reference_definition():
    from ethsnarks_b200.host import sha512_rng
    got = orc.arr_to_ints(orc.fr_from_mont(orc.sha512_rng_fr(3, 30)))
    assert got == [sha512_rng(3 + i) for i in range(30)]
    assert [hex(v) for v in got[:0]] == [] and [hex(v) for v in orc.arr_to_ints(orc.fr_from_mont(orc.sha512_rng_fr(0, 4)))] == GOLD["sha512_rng_fr"]


def test_reference_kat_verifies_and_pins_curve_layout():
    """/root/reference/test/test_verify.py:10-33: the static vk/proof must verify; a flipped nibble must not."""
    assert pairing.groth16_verify(kat.VK_STATIC, kat.PROOF_STATIC)
    bad = copy.deepcopy(kat.PROOF_STATIC)
    bad["A"][1] = bad["A"][1][:-1] + ("1" if bad["A"][1][-1] != "1" else "2")
    assert not pairing.groth16_verify(kat.VK_STATIC, bad)
    # the same points are on the oracle's curves, in the oracle's coordinate order (x.a = c0, x.b = c1)
    for k in ("A", "C"):
        xy = orc.fq_to_mont(orc.ints_to_arr([int(v, 16) for v in kat.PROOF_STATIC[k]])).reshape(-1)
        assert orc.lib().orc_g1_on_curve(orc.p64(np.ascontiguousarray(xy)))
    (xa, xb), (ya, yb) = pairing.parse_g2(kat.PROOF_STATIC["B"])
    xy = orc.fq_to_mont(orc.ints_to_arr([xa, xb, ya, yb])).reshape(-1)
    assert orc.lib().orc_g2_on_curve(orc.p64(np.ascontiguousarray(xy)))
    swapped = orc.fq_to_mont(orc.ints_to_arr([xb, xa, yb, ya])).reshape(-1)
    assert not orc.lib().orc_g2_on_curve(orc.p64(np.ascontiguousarray(swapped)))


def test_msm_pippenger_vs_naive_and_exponent():
    k = orc.sha512_rng_fr(100, 60); s = orc.sha512_rng_fr(1000, 60)
    s[3] = 0; s[4] = orc.constants()[8:12]
    for fix, msm in ((orc.g1_fixed_base, orc.msm_g1), (orc.g2_fixed_base, orc.msm_g2)):
        xy, inf = fix(k)
        r1 = msm(xy, None, s); r2 = msm(xy, None, s, naive=True); r3 = msm(xy, None, s, c=5, prefilter=0)
        assert np.array_equal(r1[0], r2[0]) and np.array_equal(r1[0], r3[0])
        assert np.array_equal(fix(orc.fr_dot(k, s).reshape(1, 4))[0][0], r1[0])


def test_ntt_vs_naive_dft_and_roundtrips():
    x = orc.sha512_rng_fr(0, 64)
    assert np.array_equal(orc.ntt(x, 6, 0), orc.dft_naive(x, 6))
    assert np.array_equal(orc.ntt(orc.ntt(x, 6, 0), 6, 1), x)
    assert np.array_equal(orc.ntt(orc.ntt(x, 6, 2), 6, 3), x)


def test_golden_fixtures():
    mk = {"field_30_3": lambda: orc.Circuit.synth_field(30, 3), "field_1000_1": lambda: orc.Circuit.synth_field(1000, 1),
          "binary_1000_10": lambda: orc.Circuit.synth_binary(1000, 10)}
    for name, g in GOLD["circuits"].items():
        c = mk[name](); w = c.witness(); k = orc.Key(c)
        assert (c.nc, c.ni, c.nv, c.m) == (g["nc"], g["ni"], g["nv"], g["m"])
        assert c.is_satisfied(w) and sha(w) == g["witness_sha256"]
        assert sha(c.h_poly(w)[:c.m]) == g["h_sha256"]
        assert hashlib.sha256(k.serialize(orc.FMT_MCL_NOZK)).hexdigest() == g["pk_mcl_nozk_sha256"]
        assert orc.proof_json(k.prove(w), w[1:1 + c.ni]) == g["proof_json"]
        rs = orc.sha512_rng_fr(1 << 48, 2)
        assert orc.proof_json(k.prove(w, rs[0], rs[1]), w[1:1 + c.ni]) == g["proof_rs_json"]
        assert k.vk_json() == g["vk_json"]
    for n, g in GOLD["msm"].items():
        n = int(n)
        kk = orc.sha512_rng_fr((1 << 40) + n, n); s = orc.sha512_rng_fr(1000 + n, n)
        b1, i1 = orc.g1_fixed_base(kk); b2, i2 = orc.g2_fixed_base(kk)
        assert [hex(v) for v in orc.arr_to_ints(orc.fq_from_mont(orc.msm_g1(b1, i1, s)[0]))] == g["g1"]
        assert [hex(v) for v in orc.arr_to_ints(orc.fq_from_mont(orc.msm_g2(b2, i2, s)[0]))] == g["g2"]
    for logm, g in GOLD["ntt"].items():
        x = orc.sha512_rng_fr(7 * int(logm), 1 << int(logm))
        assert [sha(orc.ntt(x, int(logm), mode)) for mode in range(4)] == g


def test_golden_proofs_pass_the_pairing_verifier():
    g = GOLD["circuits"]["field_30_3"]
    assert pairing.groth16_verify(json.loads(g["vk_json"]), json.loads(g["proof_json"]))
    bad = json.loads(g["proof_json"])
    bad["input"][0] = hex(int(bad["input"][0], 16) ^ 1)
    assert not pairing.groth16_verify(json.loads(g["vk_json"]), bad)


def test_in_exponent_check_detects_tampering():
    c = orc.Circuit.synth_field(100, 2); w = c.witness(); k = orc.Key(c)
    pr = k.prove(w)
    assert k.check_in_exponent(w, pr) == 0
    for idx in range(3):
        bad = [np.array(x, copy=True) for x in pr]
        bad[idx][0] ^= np.uint64(1)
        assert k.check_in_exponent(w, tuple(bad)) != 0


# ---- the reference's own prover templates, built by oracle/ref/Makefile into oracle/_ref (when present) ----
import pytest  # noqa: E402
import orc_ref  # noqa: E402

needs_ref = pytest.mark.skipif(not orc_ref.available(), reason="oracle/_ref/ref_prove not built (needs /root/reference)")


@needs_ref
@pytest.mark.parametrize("mk", [lambda: orc.Circuit.synth_field(30, 3), lambda: orc.Circuit.synth_field(1000, 1),
                                lambda: orc.Circuit.synth_binary(1000, 10), lambda: orc.Circuit.synth_field(5000, 2, 7, 8)])
def test_oracle_equals_reference_build(mk):
    """Proof points and all m coefficients of H from libsnark::r1cs_gg_ppzksnark_zok_prover<alt_bn128_pp> itself.
    The reference is run single-threaded: with >= 2 OpenMP threads its own 0/1 prefilter races on a vector<bool>
    (multiexp.tcc:652,692) and its proofs are not reproducible on small inputs."""
    c = mk(); w = c.witness(); k = orc.Key(c)
    (A, B, Cc, _), h, _ = orc_ref.prove(c, w, k, threads=1)
    want = k.prove(w)
    assert np.array_equal(A, want[0]) and np.array_equal(B, want[1]) and np.array_equal(Cc, want[2])
    assert np.array_equal(h, c.h_poly(w)[:c.m])


@needs_ref
def test_reference_build_reads_the_oracle_key_bytes_and_matches_golden():
    """The reference's operator>> accepts the oracle's ALT_NOZK serialisation, and the reference's proof for the golden
    circuit renders to the committed golden JSON."""
    c = orc.Circuit.synth_field(30, 3); w = c.witness(); k = orc.Key(c)
    proof, _, _ = orc_ref.prove(c, w, k, threads=1)
    assert orc.proof_json(proof, w[1:1 + c.ni]) == GOLD["circuits"]["field_30_3"]["proof_json"]


@needs_ref
def test_full_key_layout_is_the_reference_layout():
    """B200_FMT_ALT_FULL is the byte stream of the reference's UN-STRIPPED key: the reference's own operator>> for
    r1cs_gg_ppzksnark_zok_proving_key (dense G1_vector A, knowledge-commitment B) reads the oracle's ALT_FULL bytes and its
    operator<< writes back exactly the same bytes.  A witness that leaves A-query entries at zero makes the dense layout
    differ from the sparse one."""
    import subprocess, tempfile
    c = orc.Circuit.synth_binary(40, 3); k = orc.Key(c)
    blob = k.serialize(orc.FMT_ALT_FULL)
    with tempfile.TemporaryDirectory() as d:
        a, b = os.path.join(d, "full.raw"), os.path.join(d, "back.raw")
        open(a, "wb").write(blob)
        res = subprocess.run([orc_ref.REF_BIN, "--roundtrip-full", a, b], check=True, capture_output=True, text=True)
        assert open(b, "rb").read() == blob
    dims = res.stdout.split()
    assert int(dims[1]) == c.nv + 1          # A dense over all variables + ONE
    assert len(blob) != len(k.serialize(orc.FMT_ALT_NOZK)) + 64 * int(dims[3])   # not the sparse layout with G1 halves added


@pytest.mark.skipif(not (orc_ref.available() and orc_ref.dump_available()), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("name", ["sha256", "mimc_merkle", "eddsa"])
def test_reference_gadget_circuits_oracle_equals_reference_prover(name):
    """BASELINE.json configs[0..2]: circuits built by the reference's OWN gadget code (oracle/_ref/circuit_dump, compiled from
    the reference sources): sha256_full_gadget_512 (digest = the KAT of test_sha256_full_gadget.cpp:48-52), the depth-29 MiMC
    Merkle authenticator (after reproducing the depth-1 root of test_merkle_tree.cpp:56-62) and Baby Jubjub EdDSA on the
    known signature of test_jubjub_eddsa.cpp:44-62.  The oracle and the reference prover binary prove the same R1CS + real
    witness to identical points and H; the proof passes the restated reference verifier."""
    circ, w, out = orc_ref.reference_circuit(name)
    g = json.load(open(os.path.join(HERE, "golden", "circuit_%s.json" % name)))
    assert out == g["dump_stdout"] and "satisfied 1" in out
    if name == "sha256":
        assert "digest D294F6E585874FE640BE4CE636E6EF9E3ADC27620AA3221FDCF5C0A7C11C6F67" in out
    if name == "mimc_merkle":
        assert "merkle KAT (test_merkle_tree.cpp:56-62) reproduced" in out
    assert (circ.nc, circ.nv, circ.ni, circ.m) == (g["constraints"], g["variables"], g["inputs"], g["m"])
    assert circ.is_satisfied(w)
    assert hashlib.sha256(np.ascontiguousarray(w).tobytes()).hexdigest() == g["witness_sha256"]
    key = orc.Key(circ)
    want = key.prove(w)
    assert key.check_in_exponent(w, want) == 0
    got, h_ref, _ = orc_ref.prove(circ, w, key, threads=1)
    for a, b in zip(got[:3], want[:3]):
        assert np.array_equal(np.asarray(a), np.asarray(b))
    assert np.array_equal(h_ref[:circ.m], circ.h_poly(w)[:circ.m])
    pj = orc.proof_json(want, w[1:1 + circ.ni])
    assert pj == g["proof_json"]
    assert pairing.groth16_verify(json.loads(key.vk_json()), json.loads(pj))
