"""CPU tests of the host side: the C-ABI library loads and exports every declared symbol, the host mirror of the
reference interface agrees with the oracle, and the N>1 sharding logic (gloo, world_size 2)."""
import os
import re
import subprocess
import sys
import numpy as np
import pytest
import orc
import ethsnarks_b200 as eb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    lib = eb.lib.load()
    header = open(os.path.join(ROOT, "include", "b200snark.h")).read()
    declared = sorted(set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", header)))
    assert declared and set(declared) == set(eb.SYMBOLS)
    for name in declared:
        assert hasattr(lib, name), name


def test_no_gpu_means_loud_failure_not_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(eb.B200Error):
        eb.Context(0)


def test_product_package_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "ethsnarks_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", "Makefile")):
                src = open(os.path.join(dirpath, f)).read()
                assert "liboracle" not in src and "import orc" not in src and "oracle/" not in src, f


def test_synthetic_generators_match_oracle():
    pb = eb.generate_r1cs_example_with_field_input(300, 2)
    c = orc.Circuit.synth_field(300, 2)
    assert eb.host.fr_from_mont(c.witness()) == pb.values
    for which, M in enumerate((pb.cs.A, pb.cs.B, pb.cs.C)):
        rp, col, ci = c.matrix(which)
        assert np.array_equal(rp, M[0]) and np.array_equal(col, M[1]) and np.array_equal(ci, M[2])
    assert pb.cs.is_satisfied(pb.values) and eb.get_domain(pb) == c.m
    pb = eb.generate_r1cs_example_with_binary_input(300, 7, seed=3)
    c = orc.Circuit.synth_binary(300, 7, seed=3)
    assert eb.host.fr_from_mont(c.witness()) == pb.values
    for which, M in enumerate((pb.cs.A, pb.cs.B, pb.cs.C)):
        rp, col, ci = c.matrix(which)
        assert np.array_equal(rp, M[0]) and np.array_equal(col, M[1]) and np.array_equal(ci, M[2])
    assert eb.host.fr_from_mont(c.coeffs()) == pb.cs.coeffs and pb.cs.is_satisfied(pb.values)


def test_proof_to_json_matches_oracle_and_reference_layout():
    c = orc.Circuit.synth_field(30, 3); w = c.witness(); k = orc.Key(c)
    pr = k.prove(w)
    js = eb.proof_to_json(pr, eb.host.fr_from_mont(w[1:4]))
    assert js == orc.proof_json(pr, w[1:4])
    assert js.startswith('{\n "A" :["0x') and '\n "B"  :[["0x' in js and js.endswith(']\n}')


def test_config_defaults_match_reference():
    c = eb.Config()   # src/prover_config.hpp:10-24
    assert (c.fft, c.smt, c.swapAB, c.multi_exp_c, c.multi_exp_prefetch_locality, c.prefetch_stride, c.multi_exp_look_ahead) == \
        ("recursive", False, True, 0, 0, 128, 1)


def test_gloo_world2_shard_ranges_cover_and_gather():
    """The multi-GPU plumbing (all_gather of 32-limb partials, rank 0 combines) with gloo on CPU, world_size 2."""
    script = os.path.join(ROOT, "tests", "_gloo_worker.py")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                          "127.0.0.1", "--master-port", "29533", script], env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "GLOO_OK" in out.stdout
