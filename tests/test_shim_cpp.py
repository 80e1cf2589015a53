"""The compiled C++ drop-in (shim/ethsnarks_b200.{hpp,cpp}): the reference's entry points — ethsnarks::prove(ProverContextT&,
ProtoboardT&), load_proving_key, get_domain, stub_genkeys_from_pb (src/stubs.hpp:16-21) and the
libsnark::r1cs_gg_ppzksnark_zok_prover<ppT>(ProverContext&, assignment) template (...zok.tcc:451-452) — built against the
reference's own headers and linked to libb200snark.so.  oracle/_ref/shim_prove drives it on protoboards made by the
reference's own gadget code and checks the proofs with the reference's own verifier."""
import json
import os
import subprocess
import tempfile
import pytest
import orc
import orc_ref

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM_SO = os.path.join(ROOT, "ethsnarks_b200", "libethsnarks_b200.so")
SHIM_BIN = os.path.join(ROOT, "oracle", "_ref", "shim_prove")
needs_shim = pytest.mark.skipif(not (os.path.exists(SHIM_SO) and os.path.exists(SHIM_BIN) and orc_ref.dump_available()),
                                reason="shim / oracle/_ref not built (needs the reference tree at build time)")


def _facts(stdout):
    out = {}
    for line in stdout.splitlines():
        k, sep, v = line.partition(": ")
        if sep:
            out[k] = v
    return out


@needs_shim
def test_shim_exports_the_reference_entry_points():
    syms = subprocess.run("nm -D --defined-only %s | c++filt" % SHIM_SO, shell=True, check=True, capture_output=True, text=True).stdout
    pp = "libff::alt_bn128_pp"
    for want in (
        "ethsnarks::prove[abi:cxx11](libsnark::ProverContext<%s>&, libsnark::protoboard<" % pp,
        "ethsnarks::load_proving_key(char const*)",
        "ethsnarks::get_domain(libsnark::protoboard<",
        "ethsnarks::stub_genkeys_from_pb(libsnark::protoboard<",
        "ethsnarks::stub_test_proof_verify(libsnark::protoboard<",
        "ethsnarks::stub_verify(char const*, char const*)",
        "ethsnarks::stub_main_verify(char const*, int, char const**)",
        "libsnark::r1cs_gg_ppzksnark_zok_prover<%s>(libsnark::ProverContext<%s>&, std::vector<" % (pp, pp),
        "ethsnarks::proof_to_json[abi:cxx11](",
    ):
        assert want in syms, want


@needs_shim
def test_shim_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    circ, w, _ = orc_ref.reference_circuit("eddsa")
    key = orc.Key(circ)
    with tempfile.TemporaryDirectory() as d:
        pk = os.path.join(d, "pk.raw")
        open(pk, "wb").write(key.serialize(orc.FMT_ALT_NOZK))
        res = subprocess.run([SHIM_BIN, "prove", "eddsa", pk, os.path.join(d, "proof.json")], capture_output=True, text=True)
    assert res.returncode != 0
    assert "b200_init" in res.stderr and "no CPU fallback" in res.stderr


@needs_shim
def test_shim_stub_main_verify_on_the_golden_proof():
    """The drop-in's stub_verify / stub_main_verify (src/stubs.cpp:16-34, 90-132; host only): exit code 0 for the committed
    golden proof under the oracle key's vk.json, 1 with the reference's message for a changed public input, 2 for an
    unreadable file, 1 with the usage line for too few arguments."""
    circ, w, _ = orc_ref.reference_circuit("eddsa")
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "circuit_eddsa.json")))
    key = orc.Key(circ)
    with tempfile.TemporaryDirectory() as d:
        vk, proof, bad = (os.path.join(d, n) for n in ("vk.json", "proof.json", "bad.json"))
        open(vk, "w").write(key.vk_json())
        open(proof, "w").write(g["proof_json"])
        tampered = json.loads(g["proof_json"])
        tampered["input"][0] = hex(int(tampered["input"][0], 16) ^ 1)
        open(bad, "w").write(json.dumps(tampered))
        ok = subprocess.run([SHIM_BIN, "verify", vk, proof], capture_output=True, text=True)
        assert ok.returncode == 0 and _facts(ok.stdout)["stub_main_verify_rc"] == "0", ok.stdout + ok.stderr
        rej = subprocess.run([SHIM_BIN, "verify", vk, bad], capture_output=True, text=True)
        assert rej.returncode == 1 and "Error: failed to verify proof!" in rej.stderr
        missing = subprocess.run([SHIM_BIN, "verify", vk, os.path.join(d, "absent.json")], capture_output=True, text=True)
        assert missing.returncode == 2 and "Error: cannot open" in missing.stderr


@pytest.mark.gpu
@needs_shim
def test_shim_stub_test_proof_verify_on_device():
    """stub_test_proof_verify (src/stubs.cpp:135-149), the call the reference's gadget tests end with: generator and prover
    on the device, the reference's verifier on the host; a protoboard with a changed public input is rejected."""
    res = subprocess.run([SHIM_BIN, "selftest", "eddsa", "-", "-"], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    f = _facts(res.stdout)
    assert f["stub_test_proof_verify"] == "1" and f["stub_test_proof_verify_rejects_wrong_input"] == "1"


@pytest.mark.gpu
@needs_shim
@pytest.mark.parametrize("name", ["sha256", "mimc_merkle", "eddsa"])
def test_shim_prove_matches_golden_and_passes_the_reference_verifier(name):
    _prove_and_check(name, {})


@pytest.mark.gpu
@needs_shim
def test_shim_prove_on_a_multi_device_context():
    """B200_DEVICES makes the drop-in's one context span several GPUs (b200_init with n_devices > 1): same golden proof.
    All GPUs of the box when there are several, else the one GPU listed twice (the same code path)."""
    import torch
    n = torch.cuda.device_count()
    devs = ",".join(str(i) for i in range(min(n, 8))) if n >= 2 else "0,0"
    _prove_and_check("mimc_merkle", {"B200_DEVICES": devs})


def _prove_and_check(name, extra_env):
    """BASELINE.json configs[0..2] through the C++ boundary: the JSON string returned by ethsnarks::prove is the committed
    golden proof (the one the reference prover binary reproduces, tests/test_oracle.py) and the reference's own verifier
    accepts it under the oracle key's vk.json, and rejects it under a changed public input."""
    circ, w, _ = orc_ref.reference_circuit(name)
    g = json.load(open(os.path.join(ROOT, "tests", "golden", "circuit_%s.json" % name)))
    key = orc.Key(circ)
    with tempfile.TemporaryDirectory() as d:
        pk, vk, proof = (os.path.join(d, n) for n in ("pk.raw", "vk.json", "proof.json"))
        open(pk, "wb").write(key.serialize(orc.FMT_ALT_NOZK))
        open(vk, "w").write(key.vk_json())
        res = subprocess.run([SHIM_BIN, "prove", name, pk, proof, vk], capture_output=True, text=True, env=dict(os.environ, **extra_env))
        assert res.returncode == 0, res.stdout + res.stderr
        f = _facts(res.stdout)
        assert open(proof).read() == g["proof_json"]
    assert f["constraints"] == str(g["constraints"]) and f["domain_size"] == str(g["m"])
    assert f["repeat_identical"] == "1" and f["reference_verifier_accepts"] == "1" and f["reference_verifier_rejects_tampered"] == "1"


@pytest.mark.gpu
@needs_shim
def test_shim_genkeys_on_device_round_trip():
    """stub_genkeys_from_pb on the device writes the reference's own pk / vk.json formats: the reference's loader reads the
    pk back, the GPU proves with it, the reference verifier accepts under the generated vk."""
    with tempfile.TemporaryDirectory() as d:
        pk, vk, proof = (os.path.join(d, n) for n in ("pk.raw", "vk.json", "proof.json"))
        res = subprocess.run([SHIM_BIN, "genkeys", "eddsa", pk, vk, proof], capture_output=True, text=True)
        assert res.returncode == 0, res.stdout + res.stderr
        f = _facts(res.stdout)
        assert f["genkeys_rc"] == "0" and f["reference_verifier_accepts"] == "1" and f["reference_verifier_rejects_tampered"] == "1"
        assert set(json.load(open(vk)).keys()) == {"alpha", "beta", "gamma", "delta", "gammaABC"}
