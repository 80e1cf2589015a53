#!/usr/bin/env python
"""Config 5 of BASELINE.json: standalone MSM sweep, G1 and G2, N = 2^16 .. 2^26, device-resident bases and scalars.

Bases are k_i * G (distinct, generated on the GPU), scalars uniform < 2^253; each result is checked "in the exponent"
against the CPU oracle: MSM == (sum k_i s_i) * G.  Prints one JSON line per (group, N).
  python tools/msm_sweep.py [--min 16] [--max 24] [--g2-max 22] [--reps 3] [--no-check]
"""
import argparse
import json
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import ethsnarks_b200 as eb  # noqa: E402
from bench import rand_scalars, MAC_PER_G1_POINT, MAC_PER_G2_POINT  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min", type=int, default=16)
    ap.add_argument("--max", type=int, default=24)
    ap.add_argument("--g2-max", type=int, default=22)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--tables", type=int, default=1)
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--groups", default="1,2")
    ap.add_argument("--affine-rounds", type=int, default=0)
    ap.add_argument("--opt", action="append", default=[], help="library option name=value, repeatable")
    args = ap.parse_args()
    ctx = eb.Context(0)
    peak = ctx.imad_peak(5)
    if args.chunk:
        ctx.set_option("chunk", args.chunk)
    ctx.set_option("affine_rounds", args.affine_rounds)
    for kv in args.opt:
        name, _, val = kv.partition("=")
        ctx.set_option(name, int(val))
    if not args.no_check:
        import orc
    for group in [int(g) for g in args.groups.split(",")]:
        for logn in range(args.min, (args.max if group == 1 else min(args.max, args.g2_max)) + 1):
            n = 1 << logn
            k = rand_scalars(n, 1000 + logn)
            s = rand_scalars(n, 2000 + logn)
            bases = eb.Bases(ctx, group, scalars=k)
            if args.tables:
                bases.precompute(0)
            s_dev = torch.from_numpy(s.view(np.int64)).cuda()
            torch.cuda.synchronize()
            times, sub = [], {}
            for r in range(args.reps + 1):
                got, inf = bases.msm(device_ptr=s_dev.data_ptr(), n=n)
                t = dict(ctx.timings())
                if r:
                    times.append(t["msm.total"])
                    for kk, v in t.items():
                        sub.setdefault(kk, []).append(v)
            ok = None
            if not args.no_check:
                want, _ = (orc.g1_fixed_base if group == 1 else orc.g2_fixed_base)(orc.fr_dot(k, s).reshape(1, 4))
                ok = bool(np.array_equal(got, want[0]))
            ms = float(np.min(times))
            mac = (MAC_PER_G1_POINT if group == 1 else MAC_PER_G2_POINT) * n
            print(json.dumps({"group": "G%d" % group, "log_n": logn, "ms": ms, "mpoints_per_s": n / ms / 1e3,
                              "tmac_s_canonical": mac / (ms * 1e-3) / 1e12, "frac_of_imad_peak": mac / (ms * 1e-3) / peak,
                              "imad_peak_tmac_s": peak / 1e12, "in_exponent_ok": ok, "tables": bool(args.tables), "chunk": args.chunk, "affine_rounds": args.affine_rounds,
                              "phases_ms": {kk: round(float(np.min(v)), 4) for kk, v in sub.items()}}), flush=True)
            bases.free()
            del s_dev
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
