#!/usr/bin/env python
"""ONE process, ONE context over several GPUs (b200_init with n_devices > 1 - the mode the C++ drop-in uses under B200_DEVICES):
latency of b200_prove on the synthetic 2^log_m workload with the key split 1/N per device.  Wall clock around the call (it
returns after the proof is back on the host), median of --steps; the proof is compared with a one-device context's.
  python tools/multi_device_probe.py --devices 0,1,2,3 [--log-m 20] [--steps 20] [--opt split_chains=0]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--devices", default="0,1")
    ap.add_argument("--log-m", type=int, default=20)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--opt", action="append", default=[])
    args = ap.parse_args()
    import torch
    import ethsnarks_b200 as eb
    devices = [int(d) for d in args.devices.split(",")]
    out = {"log_m": args.log_m, "devices": devices}
    proofs = {}
    for devs in ([devices[0]], devices):
        ctx = eb.Context(devs if len(devs) > 1 else devs[0])
        for kv in args.opt:
            name, _, val = kv.partition("=")
            ctx.set_option(name, int(val))
        key, witness, dims = bench.build_workload(eb, ctx, args.log_m, 0, 1)
        w_pin = torch.from_numpy(witness.view(np.int64)).pin_memory()
        for _ in range(3):
            proof = key.prove(w_pin.data_ptr())
        ts = []
        for _ in range(args.steps):
            t0 = time.perf_counter()
            key.prove(w_pin.data_ptr())
            ts.append((time.perf_counter() - t0) * 1e3)
        proofs[len(devs)] = proof
        out["prove_ms_median_%d_dev" % len(devs)] = float(np.median(ts))
        out["prove_ms_min_%d_dev" % len(devs)] = float(np.min(ts))
        out["launches_per_prove_%d_dev" % len(devs)] = None
        l0 = ctx.launches
        key.prove(w_pin.data_ptr())
        out["launches_per_prove_%d_dev" % len(devs)] = ctx.launches - l0
        key.free()
        ctx.close()
        if len(devices) == 1:
            break
    if len(devices) > 1:
        out["identical_to_one_device"] = all(np.array_equal(np.asarray(a), np.asarray(b)) for a, b in zip(proofs[1], proofs[len(devices)]))
        out["speedup"] = out["prove_ms_median_1_dev"] / out["prove_ms_median_%d_dev" % len(devices)]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
