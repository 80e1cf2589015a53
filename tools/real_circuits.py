#!/usr/bin/env python
"""BASELINE.json configs[0..2] - MiMC Merkle depth 29, SHA256 full, Baby Jubjub EdDSA, built by the reference's own gadget code
(oracle/_ref/circuit_dump): GPU prove latency through the host-buffer C ABI beside the reference prover's time on the same R1CS,
witness and key (oracle/_ref/ref_prove, all host threads).  These circuits are tiny for a B200 (2^14 - 2^16 constraints): the
latency is launch- and tail-bound, they are parity workloads first.  One JSON line per circuit.
  python tools/real_circuits.py [--steps 20]"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
import orc  # noqa: E402   (key generator + checker)
import orc_ref  # noqa: E402
import ethsnarks_b200 as eb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--window", type=int, default=0, help="window size c of the fixed-base tables (0 = cost model)")
    ap.add_argument("--opt", action="append", default=[], help="library option name=value, repeatable")
    ap.add_argument("--no-cpu", action="store_true", help="skip the reference CPU prover (window sweeps)")
    args = ap.parse_args()
    ctx = eb.Context(0)
    if args.window:
        ctx.set_option("window", args.window)
    for kv in args.opt:
        ctx.set_option(kv.split("=")[0], int(kv.split("=")[1]))
    for name in ("mimc_merkle", "sha256", "eddsa"):
        circ, w, _ = orc_ref.reference_circuit(name)
        gold = json.load(open(os.path.join(ROOT, "tests", "golden", "circuit_%s.json" % name)))
        key = orc.Key(circ)
        pk = eb.ProvingKey.load(ctx, key.serialize(orc.FMT_MCL_NOZK), eb.FMT_MCL_NOZK)
        pk.load_circuit(circ.matrix(0), circ.matrix(1), circ.matrix(2), circ.coeffs(), circ.ni, circ.nv)
        w = np.ascontiguousarray(w)
        for _ in range(3):
            proof = pk.prove(w)
        assert eb.proof_to_json(proof, eb.host.fr_from_mont(w[1:1 + circ.ni])) == gold["proof_json"], name
        wall, dev = [], []
        for _ in range(args.steps):
            t0 = time.perf_counter()
            pk.prove(w)
            wall.append((time.perf_counter() - t0) * 1e3)
            dev.append(dict(ctx.timings())["prove.total"])
        timeline = [(n, round(a, 3), round(d, 3)) for n, a, d in ctx.timeline()]
        ones = int(np.sum(np.all(w == np.asarray(orc.constants()[8:12], np.uint64), axis=1)))
        zeros = int(np.sum(~w.any(axis=1)))
        cpu = [float("nan")] if args.no_cpu else orc_ref.prove(circ, w, key, repeats=4, threads=os.cpu_count(), all_times=True)[2][1:]
        print(json.dumps({"circuit": name, "window": args.window, "constraints": circ.nc, "variables": circ.nv, "domain": circ.m,
                          "witness_zero_or_one": (zeros + ones) / float(circ.nv + 1),
                          "gpu_prove_wall_ms_median": float(np.median(wall)), "gpu_prove_device_ms_median": float(np.median(dev)),
                          "reference_cpu_prove_ms_median": float(np.median(cpu)) * 1e3, "cpu_threads": os.cpu_count(),
                          "speedup": float(np.median(cpu)) * 1e3 / float(np.median(wall)), "proof_equals_golden": True, "timeline_name_start_ms_dur_ms": timeline}), flush=True)
        pk.free()


if __name__ == "__main__":
    main()
