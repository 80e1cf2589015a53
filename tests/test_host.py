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


def test_constraint_instances_flatten_like_the_reference():
    """append_constraint_instances against a per-row restatement of r1cs_constraint_light_instance::getA/B/C and ::evaluateA/B/C
    (r1cs.hpp:158-200, variable.hpp:306-314): index 0 is never translated, every other index goes through the translator."""
    import random
    H = eb.host
    rng = random.Random(5)
    # master gadget over local variables 1..4 (0 = ONE): (x1 + 3) * x2 = x3 ; x3 * x3 = x4 + 2*x1
    coeffs = [1, 3, 2]
    mA = H._csr([[(1, 0), (0, 1)], [(3, 0)]]); mB = H._csr([[(2, 0)], [(3, 0)]]); mC = H._csr([[(3, 0)], [(4, 0), (1, 2)]])
    master = H.R1CS(0, 4, mA, mB, mC, coeffs)
    # host circuit: one plain constraint x1 * x1 = x2, then 3 instances on disjoint variable blocks
    n_inst, nv = 3, 2 + 4 * 3
    cs = H.R1CS(1, nv, H._csr([[(1, 0)]]), H._csr([[(1, 0)]]), H._csr([[(2, 0)]]), coeffs)
    trs = [H.Translator([0] + [3 + 4 * j + i for i in range(4)]) for j in range(n_inst)]
    H.append_constraint_instances(cs, master, trs)
    assert cs.num_constraints == 1 + 2 * n_inst
    # per-row restatement
    def rows_of(M):
        rp, col, ci = M
        return [[(int(col[k]), int(ci[k])) for k in range(int(rp[i]), int(rp[i + 1]))] for i in range(len(rp) - 1)]
    for name, M, N in (("A", cs.A, mA), ("B", cs.B, mB), ("C", cs.C, mC)):
        want = rows_of({"A": H._csr([[(1, 0)]]), "B": H._csr([[(1, 0)]]), "C": H._csr([[(2, 0)]])}[name])
        for t in trs:
            for r in rows_of(N):
                want.append([(idx if idx == 0 else t.translate(idx), c) for idx, c in r])
        assert rows_of(M) == want, name
    # a satisfying assignment built instance by instance satisfies the flattened system
    vals = [0] * nv
    x = rng.randrange(H.R_MOD); vals[0] = x; vals[1] = x * x % H.R_MOD
    for j in range(n_inst):
        x1, x2 = rng.randrange(H.R_MOD), rng.randrange(H.R_MOD)
        x3 = (x1 + 3) * x2 % H.R_MOD
        x4 = (x3 * x3 - 2 * x1) % H.R_MOD
        vals[2 + 4 * j: 6 + 4 * j] = [x1, x2, x3, x4]
    pb = H.Protoboard(cs, vals)
    assert cs.is_satisfied(pb.values)
    pb.values[4] = (pb.values[4] + 1) % H.R_MOD
    assert not cs.is_satisfied(pb.values)
