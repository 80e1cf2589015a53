#!/usr/bin/env python
"""Config 5 of BASELINE.json: standalone MSM sweep, G1 and G2, N = 2^16 .. 2^26, device-resident bases and scalars.

Bases are k_i * G (distinct, generated on the GPU), scalars uniform < 2^253; each result is checked "in the exponent"
against the CPU oracle: MSM == (sum k_i s_i) * G.  Prints one JSON line per (group, N).
  python tools/msm_sweep.py [--min 16] [--max 24] [--g2-max 22] [--reps 5] [--no-check] [--tables 0|1] [--mix uniform|binary90] [--cpu-max 20]
--tables 0 is what a standalone multi_exp caller gets (no precomputed window tables: plain windowed Pippenger); with tables the
row also states what they cost (precompute_ms, table_bytes).  --mix binary90: 90 % of the scalars from {0, 1}, 10 % uniform
(SURVEY 8(d)).  --cpu-max L adds, for sizes up to 2^L, the time of the REFERENCE's own multi_exp on the host cores
(oracle/_ref/ref_prove --msm: multiexp.tcc:595-627 for uniform scalars, :631-714 with the 0/1 prefilter for binary90).
Times are the MEDIAN of --reps timed runs after one warm-up (min also given).

Under torchrun (one process per GPU, NCCL) the (base, scalar) index range is split evenly over the ranks: every
rank runs the whole window pipeline on its slice, the affine partial sums (8 or 16 limbs + flag per rank) are all-gathered
and rank 0 adds them.  Timed on the device: max over ranks of the MSM's own CUDA-event time plus the all-gather + additions.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 tools/msm_sweep.py --from 20 --to 26
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
from bench import rand_scalars, shard_range, max_over_ranks, MAC_PER_G1_POINT, MAC_PER_G2_POINT  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min", "--from", dest="min", type=int, default=16)
    ap.add_argument("--max", "--to", dest="max", type=int, default=24, help="--from/--to: the same under torchrun, whose own parser claims --max")
    ap.add_argument("--g2-max", type=int, default=22)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--mix", default="uniform", choices=["uniform", "binary90"])
    ap.add_argument("--cpu-max", type=int, default=0, help="largest log2 size the CPU reference multi_exp is timed at (0 = off)")
    ap.add_argument("--no-check", action="store_true")
    ap.add_argument("--tables", type=int, default=1)
    ap.add_argument("--chunk", type=int, default=0)
    ap.add_argument("--table-window", type=int, default=0, help="window size c of the fixed-base tables (0 = cost model)")
    ap.add_argument("--groups", default="1,2")
    ap.add_argument("--opt", action="append", default=[], help="library option name=value, repeatable")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = eb.Context(local)
    peak = ctx.imad_peak(5)
    if args.chunk:
        ctx.set_option("chunk", args.chunk)
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
            if args.mix == "binary90":
                rng = np.random.default_rng(3000 + logn)
                small = rng.random(n) < 0.9
                one = np.asarray(eb.host.fr_to_mont([1]), np.uint64).reshape(4)
                s[small] = np.where(rng.integers(0, 2, size=(int(small.sum()), 1)) == 1, one, np.zeros(4, np.uint64))
            lo, hi = shard_range(n, rank, world)
            bases = eb.Bases(ctx, group, scalars=np.ascontiguousarray(k[lo:hi]))
            pre_ms, table_bytes = None, 0
            if args.tables:
                import time
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                bases.precompute(args.table_window)
                torch.cuda.synchronize()
                pre_ms = (time.perf_counter() - t0) * 1e3
            s_dev = torch.from_numpy(np.ascontiguousarray(s[lo:hi]).view(np.int64)).cuda()
            torch.cuda.synchronize()
            times, sub = [], {}
            limbs = 8 if group == 1 else 16
            if world > 1:
                pin = torch.zeros(limbs + 1, dtype=torch.int64).pin_memory()
                pin_np = pin.numpy().view(np.uint64)
                mine = torch.zeros(limbs + 1, dtype=torch.int64, device="cuda")
                allp = torch.zeros(world * (limbs + 1), dtype=torch.int64, device="cuda")
            for r in range(args.reps + 1):
                if world > 1:
                    dist.barrier()
                got, inf = bases.msm(device_ptr=s_dev.data_ptr(), n=hi - lo)
                t = dict(ctx.timings())
                if world > 1:
                    # the one exchange step: all ranks' affine partial sums all-gathered, then added in one kernel
                    part = pin_np
                    part[:limbs] = np.asarray(got).reshape(-1)
                    part[limbs] = int(np.asarray(inf).reshape(-1)[0])
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    mine.copy_(pin, non_blocking=True)
                    dist.all_gather_into_tensor(allp, mine)
                    host = allp.cpu().numpy().view(np.uint64).reshape(world, limbs + 1)
                    pts = np.ascontiguousarray(host[:, :limbs])
                    flg = np.ascontiguousarray(host[:, limbs].astype(np.uint8))
                    got, inf = (ctx.g1_sum if group == 1 else ctx.g2_sum)(pts, flg)      # one kernel: world - 1 additions
                    got = np.asarray(got).reshape(-1)
                    e1.record()
                    torch.cuda.synchronize()
                    t["msm.exchange"] = e0.elapsed_time(e1)
                    t["msm.total"] = max_over_ranks(t["msm.total"]) + max_over_ranks(t["msm.exchange"])
                if r:
                    times.append(t["msm.total"])
                    for kk, v in t.items():
                        sub.setdefault(kk, []).append(v)
            ok = None
            if not args.no_check and rank == 0:
                want, _ = (orc.g1_fixed_base if group == 1 else orc.g2_fixed_base)(orc.fr_dot(k, s).reshape(1, 4))
                ok = bool(np.array_equal(np.asarray(got).reshape(-1), np.asarray(want[0]).reshape(-1)))
            ms = float(np.median(times))
            cpu = None
            if rank == 0 and args.cpu_max and logn <= args.cpu_max:
                import subprocess
                ref = os.path.join(ROOT, "oracle", "_ref", "ref_prove")
                if os.path.exists(ref):
                    env = dict(os.environ, OMP_NUM_THREADS=str(os.cpu_count() or 1))
                    out = subprocess.run([ref, "--msm", str(group), str(logn), "3", "1" if args.mix == "binary90" else "0"], env=env,
                                         capture_output=True, text=True, check=True).stdout
                    cj = json.loads(out.strip().splitlines()[-1])
                    cpu = {"ms": float(np.median(cj["seconds"][1:])) * 1e3, "threads": cj["threads"],
                           "what": "reference multi_exp%s (oracle/_ref/ref_prove --msm), median of 2 after 1 warm-up" % ("_with_mixed_addition" if args.mix == "binary90" else "")}
            mac = (MAC_PER_G1_POINT if group == 1 else MAC_PER_G2_POINT) * n
            if rank == 0:
              print(json.dumps({"group": "G%d" % group, "log_n": logn, "n_gpus": world, "ms": ms, "mpoints_per_s": n / ms / 1e3,
                              "tmac_s_canonical": mac / (ms * 1e-3) / 1e12, "frac_of_imad_peak": mac / (ms * 1e-3) / peak / world,
                              "imad_peak_tmac_s": peak / 1e12, "in_exponent_ok": ok, "tables": bool(args.tables), "mix": args.mix, "ms_min": float(np.min(times)),
                              "reps": len(times), "precompute_ms": pre_ms, "window": int(ctx.last_window()) if hasattr(ctx, "last_window") else None,
                              "cpu_reference": cpu, "speedup_vs_cpu": (cpu["ms"] / ms) if cpu else None,
                              "chunk": args.chunk,
                              "phases_ms": {kk: round(float(np.median(v)), 4) for kk, v in sub.items()}}), flush=True)
            bases.free()
            del s_dev
            torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
