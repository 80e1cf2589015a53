"""CPU tests of the host side: the C-ABI library loads and exports every declared symbol, the host mirror of the
reference interface agrees with the oracle, and the N>1 sharding logic (gloo, world_size 2)."""
import json
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


def test_key_converters_match_the_oracle_serialisations():
    """b200_key_convert (host only: pk_alt2mcl / pk_mcl2nozk, src/export.cpp:330-408) between the four reference layouts, each of
    which the oracle writes from the same key; NOZK -> FULL is refused.  The ALT layouts are pinned on the reference's own
    stream operators (tests/test_oracle.py)."""
    from ethsnarks_b200 import lib as L
    import orc
    c = orc.Circuit.synth_binary(60, 3)
    k = orc.Key(c)
    fm = {L.FMT_MCL_NOZK: orc.FMT_MCL_NOZK, L.FMT_MCL_FULL: orc.FMT_MCL_FULL, L.FMT_ALT_NOZK: orc.FMT_ALT_NOZK, L.FMT_ALT_FULL: orc.FMT_ALT_FULL}
    blobs = {f: k.serialize(o) for f, o in fm.items()}
    for fi in fm:
        for fo in fm:
            if not (fi & 1) and (fo & 1):
                with pytest.raises(L.B200Error):
                    L.key_convert(blobs[fi], fi, fo)
            else:
                assert L.key_convert(blobs[fi], fi, fo) == blobs[fo], (fi, fo)
    with pytest.raises(L.B200Error):
        L.key_convert(blobs[L.FMT_ALT_FULL][:-7], L.FMT_ALT_FULL, L.FMT_MCL_NOZK)      # truncated


def test_pk_bellman2ethsnarks_matches_the_oracle_key():
    """host.pk_bellman2ethsnarks (src/export.cpp:267-328): an oracle key written out as a bellman-style JSON (Jacobian points with
    random z, decimal strings) comes back as the bytes of the stripped reference key."""
    import random
    import orc
    H = eb.host
    rng = random.Random(3)
    c = orc.Circuit.synth_binary(40, 2)
    k = orc.Key(c)
    full = k.serialize(orc.FMT_ALT_FULL)
    q = H.Q_MOD
    rinv = pow(H.MONT_R, -1, q)
    pos = [0]

    def take(n):
        b = full[pos[0]:pos[0] + n]; pos[0] += n
        return b

    def num():
        e = full.index(b"\n", pos[0]); v = int(full[pos[0]:e]); pos[0] = e + 1
        return v

    def fqs(b):
        return [int.from_bytes(b[i:i + 32], "little") * rinv % q for i in range(0, len(b), 32)]

    def g1():
        inf = take(1) == b"1"; x, y = fqs(take(64))
        if inf:
            return ["0", "1", "0"]
        z = rng.randrange(1, q)
        return [str(x * z * z % q), str(y * z * z * z % q), str(z)]

    def g2():
        inf = take(1) == b"1"; x0, x1, y0, y1 = fqs(take(128))
        if inf:
            return [["0", "0"], ["1", "0"], ["0", "0"]]
        z = rng.randrange(1, q)                                        # z = z + 0 u keeps the arithmetic in Fq
        return [[str(x0 * z * z % q), str(x1 * z * z % q)], [str(y0 * z ** 3 % q), str(y1 * z ** 3 % q)], [str(z), "0"]]

    d = {"vk_alfa_1": g1(), "vk_beta_1": g1(), "vk_beta_2": g2(), "vk_delta_1": g1(), "vk_delta_2": g2()}
    nA = num()
    d["A"] = [g1() for _ in range(nA)]
    ds, nidx = num(), num()
    idx = [num() for _ in range(nidx)]
    assert num() == nidx
    B1 = [["0", "1", "0"]] * ds; B2 = [[["0", "0"], ["1", "0"], ["0", "0"]]] * ds
    B1, B2 = list(B1), list(B2)
    for i in idx:
        B2[i] = g2(); B1[i] = g1()
    d["B1"], d["B2"] = B1, B2
    d["hExps"] = [g1() for _ in range(num())]
    d["C"] = [["0", "1", "0"], ["0", "1", "0"]] + [g1() for _ in range(num())]
    assert pos[0] == len(full)
    assert H.pk_bellman2ethsnarks(json.dumps(d)) == k.serialize(orc.FMT_ALT_NOZK)
    assert H.pk_bellman2ethsnarks(d, alt=False) == k.serialize(orc.FMT_MCL_NOZK)


def test_multi_product_montgomery_rows_stay_inside_their_accumulators():
    """Big-integer model of the row sets in ethsnarks_b200/csrc/fp.cuh (Fp::dot2 / Fp::dot4): K products summed before each
    row's m*p step.  The accumulators hold 288 bits before the shift and the result must come out below 2p (one conditional
    subtraction).  Checked for both moduli at the largest operands the kernels feed (p itself, from neg_lazy of zero) and on
    random ones; the value is checked too.  The bound FAILS at K = 6 for Fq, which the test also pins."""
    import random
    rng = random.Random(5)
    for p in (orc.R_MOD, orc.Q_MOD):
        inv = (-pow(p, -1, 1 << 32)) % (1 << 32)
        R = 1 << 256

        def rows(pairs):
            T, worst = 0, 0
            for i in range(8):
                for a, b in pairs:
                    T += a * ((b >> (32 * i)) & 0xffffffff)
                T += (((T & 0xffffffff) * inv) & 0xffffffff) * p
                worst = max(worst, T)
                assert T & 0xffffffff == 0
                T >>= 32
            return T, worst

        for K in (2, 4):
            cases = [[(p, p)] * K, [(p - 1, p - 1)] * K, [(p, p - 1)] * K]
            cases += [[(rng.randrange(p + 1), rng.randrange(p + 1)) for _ in range(K)] for _ in range(300)]
            for pairs in cases:
                T, worst = rows(pairs)
                assert worst < 1 << 288
                assert T < 2 * p
                assert T % p == sum(a * b for a, b in pairs) * pow(R, -1, p) % p
        if p == orc.Q_MOD:
            assert rows([(p, p)] * 6)[0] >= 2 * p   # six products would need a second subtraction: dot4 is the limit used
