import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orc
import ethsnarks_b200 as eb
ctx = eb.Context(0)
for name, mk in (("field_30_3", lambda: orc.Circuit.synth_field(30, 3)), ("field_1000_1", lambda: orc.Circuit.synth_field(1000, 1))):
    circ = mk(); w = circ.witness(); key = orc.Key(circ)
    want = key.prove(w)
    for pre in (1, 0):
        for ov in (1, 0):
            ctx.set_option("precompute", pre); ctx.set_option("overlap", ov)
            pk = eb.ProvingKey.load(ctx, key.serialize(orc.FMT_MCL_NOZK), eb.FMT_MCL_NOZK)
            pk.load_circuit(circ.matrix(0), circ.matrix(1), circ.matrix(2), circ.coeffs(), circ.ni, circ.nv)
            for rep in range(3):
                got = pk.prove(w)
                print(name, "pre", pre, "overlap", ov, "rep", rep, [bool(np.array_equal(np.asarray(g), np.asarray(e))) for g, e in zip(got, want)], flush=True)
            pk.free()
