import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import orc
import ethsnarks_b200 as eb
ctx = eb.Context(0)
print("lib", os.environ.get("B200SNARK_LIB"))
for n in (40, 6000):
    k = orc.sha512_rng_fr((1 << 43), n)
    xy, inf = orc.g1_fixed_base(k)
    for kind in ("dense",):
        s = orc.sha512_rng_fr(5000, n)
        if kind == "binary":
            one = eb.host.fr_to_mont(np.array([[1, 0, 0, 0]], np.uint64))[0] if hasattr(eb.host, "fr_to_mont") else None
            s = s.copy(); s[::2] = 0; s[1::2] = one
        want, winf = orc.msm_g1(xy, inf, s)
        b = eb.Bases(ctx, 1, xy=xy, inf=inf)
        for window in (5, 7, 13):
            for r1d in (0, 1):
                b.precompute(window)
                ctx.set_option("reduce_1d", r1d)
                got, gi = b.msm(s)
                ctx.set_option("reduce_1d", 0)
                print("n", n, kind, "table window", window, "reduce_1d", r1d, "ok", bool(gi == winf and np.array_equal(got, want)), flush=True)
        b.free()
