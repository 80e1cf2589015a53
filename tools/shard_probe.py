#!/usr/bin/env python
"""One rank's share of a sharded proof, measured on ONE GPU: builds rank r's 1/N slice of the synthetic key (as bench.py does
under torchrun), times b200_prove_partial_dev with CUDA events on the library's stream and prints the phase table.  The
N-GPU latency of one proof is this plus the exchange (256 B per rank) and the combine.
  python tools/shard_probe.py --log-m 20 --world 8 [--rank 0] [--opt name=value ...]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log-m", type=int, default=20)
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--flavour", default="field")
    ap.add_argument("--opt", action="append", default=[])
    args = ap.parse_args()
    import torch
    import ethsnarks_b200 as eb
    ctx = eb.Context(0)
    for kv in args.opt:
        name, _, val = kv.partition("=")
        ctx.set_option(name, int(val))
    dev = torch.device("cuda", 0)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    key, witness, dims = bench.build_workload(eb, ctx, args.log_m, args.rank, args.world, args.flavour)
    w_dev = torch.from_numpy(witness.view(np.int64)).to(dev)
    out = torch.zeros(32, dtype=torch.int64, device=dev)
    for _ in range(3):
        key.prove_partial_dev(w_dev.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        key.prove_partial_dev(w_dev.data_ptr(), out.data_ptr())
    e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    # serialised phases
    ctx.set_option("overlap", 0)
    key.prove_partial_dev(w_dev.data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    key.prove_partial(w_dev.data_ptr())
    phases = ctx.timings()
    agg, cur = {}, None
    for name, t in phases:
        if name in ("msm.A", "msm.B_g2", "msm.H", "msm.L"):
            cur = name[4:]
            agg[name] = round(t, 4)
        elif name.startswith("msm.") and cur:
            agg["%s@%s" % (name, cur)] = round(t, 4)
        else:
            agg[name] = round(t, 4)
    print(json.dumps({"log_m": args.log_m, "world": args.world, "rank": args.rank, "partial_ms": ms, "info": key.info(), "serialised_phases_ms": agg}))


if __name__ == "__main__":
    main()
