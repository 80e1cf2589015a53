#!/usr/bin/env python
"""bench.py — Groth16 prove throughput on synthetic R1CS (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--log-m 20] [--flavour field|binary] [--impl ours|reference]
                  [--mode both|shard|replicas] [--staged 0|1]

A "step" is one full prove (witness-times-R1CS, the 7-NTT H chain, the A / B(G2) / H / L multi-scalar
multiplications, the final additions) of the synthetic 2^log_m-constraint R1CS of r1cs_examples.tcc:25-93 (or, with
--flavour binary, the 0/1-witness one of :100-157) against a device-resident synthetic proving key of the matching shape.
`value` is measured with the witnesses already in HBM, `e2e` through the host-buffer C-ABI call (pinned witness H2D and proof
D2H inside the timed region); the timed loops alternate between two distinct witnesses of the circuit.
With N > 1 (torchrun) the headline is ONE proof over a key sharded 1/N per rank ("scaling": "strong"): every rank proves its
slice of every query, the ranks share the three transform chains (each vector of coset evaluations is computed by one rank and
broadcast over NCCL), the 256-byte partial sums are all-gathered on the device and rank 0 combines them; the N-rank proof is
compared byte for byte with the one-GPU proof inside the run.  The other arrangement - N independent replicas, weak scaling,
no collective - is measured in the same run and reported under "replicas".  `--mode shard` / `--mode replicas` run one only.
`--impl reference` times the CPU implementation of the same path (oracle/_ref/ref_prove, the reference's own prover built from
its sources; else the oracle port) on the benchmark's own instance, every step one whole prove on all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

MAC_PER_G1_POINT = 23936     # 16 windows x 11 modmul x 136 MAC
MAC_PER_G2_POINT = 63104     # 16 x 29 x 136
# what the kernels execute per mixed addition: G1 8 products x 136 + one two-product row set (Y3) x 200;
# G2 per lane of the pair 6 x 200 + 2 squarings x 136 + one four-product row set x 328, two lanes
MAC_PER_G1_ADD = 8 * 136 + 200
MAC_PER_G2_ADD = 2 * (6 * 200 + 2 * 136 + 328)
BYTES_PER_G1_POINT = 96      # 64 B affine base + 32 B scalar
BYTES_PER_G2_POINT = 160
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the accumulate kernel at 2^20 points, from the ncu --set full
# capture summarised in profiles/r2_ncu_accumulate_fused_y.txt (window tables: 13 table points are read per scalar,
# so the traffic is ~12x the canonical 96/160 B per point; HBM stays under 11 % busy)
NCU_TRAFFIC_BYTES = {("G2", 20): 1.881476e9 + 0.169665e9, ("G1", 20): 1.853902e9 + 0.101182e9}


def table_windows(n, override=0):
    """Windows per scalar of a fixed-base-table MSM over n bases: the cost model of choose_window(n, true)
    (ethsnarks_b200/csrc/msm_impl.cuh), restated here only to annotate the roofline with the executed multiply-adds."""
    best_c, best = 8, float("inf")
    for c in range(6, 23):
        W = (255 + c - 1) // c
        cost = n * W * 11.5 + (2.0 ** (c - 1)) * 34.0
        if cost < best:
            best, best_c = cost, c
    c = max(4, min(24, override)) if override else best_c
    return (255 + c - 1) // c


# ----------------------------------------------------------------------------------------------- multi-rank helpers
def shard_range(n, rank, world):
    return n * rank // world, n * (rank + 1) // world


def all_gather_packed(parts, part, stream=None):
    """rank r's 256-byte record `part` (32 x int64) lands at parts[32 r : 32 r + 32] on every rank; on `stream` when given"""
    import torch
    import torch.distributed as dist
    if stream is None:
        dist.all_gather_into_tensor(parts, part)
    else:
        with torch.cuda.stream(stream):
            dist.all_gather_into_tensor(parts, part)


class DeviceBuffer:
    """a device allocation of the library seen as a torch tensor (int64 words) through __cuda_array_interface__"""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (nbytes // 8,), "typestr": "<i8", "data": (int(ptr), False), "version": 2}


class ShardedProver:
    """One proof over a key sharded 1/N per rank, with the exchange on the device: every rank queues its partial proof
    (b200_prove_partial_dev: 256 B left in HBM, no synchronisation), the records are all-gathered over NCCL on the library's
    own stream, and rank 0 combines them (b200_proof_combine_dev) and copies the proof back - one host synchronisation per
    proof, on rank 0 only."""

    def __init__(self, key, world, rank, stream, dev, staged=True):
        import torch
        from ethsnarks_b200.lib import chain_owner
        self.key, self.world, self.rank, self.stream = key, world, rank, stream
        self.part = torch.zeros(32, dtype=torch.int64, device=dev)             # 256 B: A 64 | B 128 | C 64
        self.parts = torch.zeros(32 * world, dtype=torch.int64, device=dev)
        self.w_stage = None
        # staged form: the ranks also share the transform work - each of the three evaluation vectors goes through the witness
        # map and its first two transforms on ONE rank and is broadcast (NCCL, on the library's stream) to the others
        self.staged = staged and world > 1
        if self.staged:
            self.owner = [chain_owner(ch, world) for ch in range(3)]
            self.chains = sum(1 << ch for ch in range(3) if self.owner[ch] == rank)
            ptrs, nbytes = key.buffers()
            self.vecs = [torch.as_tensor(DeviceBuffer(p, nbytes), device=dev) for p in ptrs]

    def __call__(self, wptr, host=None):
        if host is not None and self.world > 4:
            # the witness starts in (pinned) host memory: rank 0 uploads it ONCE and it is broadcast over NVLink, instead of
            # every rank pulling its own copy through the host's memory system at the same time (ncclBroadcast of w).
            # Measured at 2^20: 8 ranks 6.04 -> 5.68 ms; up to 4 ranks the host sustains the parallel uploads and the extra
            # hop costs more than it saves (N = 4: 7.50 -> 7.89 ms), so they keep uploading in parallel.
            import torch
            import torch.distributed as dist
            if self.w_stage is None or self.w_stage.shape != host.shape:
                self.w_stage = torch.empty(host.shape, dtype=host.dtype, device=self.part.device)
            with torch.cuda.stream(self.stream):
                if self.rank == 0:
                    self.w_stage.copy_(host, non_blocking=True)
                dist.broadcast(self.w_stage, src=0)
            wptr = self.w_stage.data_ptr()
        if self.staged:
            import torch
            import torch.distributed as dist
            self.key.prove_partial_stage1(wptr, self.chains)
            with torch.cuda.stream(self.stream):
                for ch in range(3):
                    dist.broadcast(self.vecs[ch], src=self.owner[ch])
            self.key.prove_partial_stage2(self.part.data_ptr())
        else:
            self.key.prove_partial_dev(wptr, self.part.data_ptr())
        if self.world > 1:
            all_gather_packed(self.parts, self.part, self.stream)
            src = self.parts
        else:
            src = self.part
        return self.key.combine_dev(src.data_ptr(), self.world) if self.rank == 0 else None


def max_over_ranks(x, device="cuda"):
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(x)
    t = torch.tensor([float(x)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def rand_scalars(n, seed):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 1 << 63, size=(n, 4), dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, size=(n, 4), dtype=np.uint64)
    a[:, 3] &= np.uint64((1 << 60) - 1)
    return np.ascontiguousarray(a)


# ----------------------------------------------------------------------------------------------- self-check
Q_MOD = 21888242871839275222246405745257275088696311157297823662689037894645226208583


def _ints(a):
    a = np.asarray(a, np.uint64).reshape(-1, 4).astype(object)
    return a[:, 0] | (a[:, 1] << 64) | (a[:, 2] << 128) | (a[:, 3] << 192)


def verify_in_the_exponent(eb, ctx, key, witness, proof, dims):
    """The synthetic key's bases are k_i * G with known k_i, so the proof has a closed form in the exponent:
    A = (alpha + sum w_i kA_i) G1, B = (beta + sum w_i kB_i) G2, C = (sum h_j kH_j + sum w_l kL_l) G1.  The three
    expected points come from the fixed-base double-and-add kernel, which shares nothing with the MSM/NTT path."""
    from ethsnarks_b200.host import R_MOD, MONT_R, ints_to_limbs
    rinv = pow(MONT_R, -1, R_MOD)
    std = lambda a: (_ints(a) * rinv) % R_MOD                       # Montgomery limbs -> integers
    nv, ni, m = dims["nv"], dims["ni"], dims["m"]
    w = std(witness)
    h = std(key.h_poly(witness))[:m - 1]
    kA, kB, kH, kL = (std(rand_scalars(n, seed)) for n, seed in ((nv + 1, 101), (nv + 1, 102), (m - 1, 103), (nv - ni, 104)))
    c1, c2 = std(rand_scalars(3, 105)), std(rand_scalars(2, 106))
    eA = (int(c1[0]) + int(np.dot(w, kA))) % R_MOD
    eB = (int(c2[0]) + int(np.dot(w, kB))) % R_MOD
    eC = (int(np.dot(h, kH)) + int(np.dot(w[ni + 1:], kL))) % R_MOD
    mont = lambda e: ints_to_limbs([(e * MONT_R) % R_MOD])
    ok = True
    for group, e, got in ((1, eA, proof[0]), (2, eB, proof[1]), (1, eC, proof[2])):
        want = eb.Bases(ctx, group, scalars=mont(e)).download()[0][0]           # Montgomery affine
        got_m = ints_to_limbs([(int(x) * MONT_R) % Q_MOD for x in _ints(got)]).reshape(-1)
        ok = ok and bool(np.array_equal(np.asarray(want).reshape(-1), got_m))
    return ok


# ----------------------------------------------------------------------------------------------- workload
def build_workload(eb, ctx, log_m, rank, world, flavour="field"):
    """Synthetic R1CS (reference generator) + a device-resident synthetic key of the same shape, sliced per rank."""
    m = 1 << log_m
    ni = 1
    nc = m - ni - 1
    pb = eb.generate_r1cs_example_with_binary_input(nc, ni) if flavour == "binary" else eb.generate_r1cs_example_with_field_input(nc, ni)
    cs = pb.cs
    nv = cs.num_variables
    witness = eb.host.fr_to_mont(pb.values)
    lo_a, hi_a = shard_range(nv + 1, rank, world)
    lo_l, hi_l = shard_range(nv - ni, rank, world)
    kA = rand_scalars(nv + 1, 101)[lo_a:hi_a]
    kB = rand_scalars(nv + 1, 102)[lo_a:hi_a]
    kH = rand_scalars(m - 1, 103)
    kL = rand_scalars(nv - ni, 104)[lo_l:hi_l]
    A = eb.Bases(ctx, 1, scalars=kA)
    B2 = eb.Bases(ctx, 2, scalars=kB)
    H = eb.Bases(ctx, 1, scalars=kH)
    Lq = eb.Bases(ctx, 1, scalars=kL)
    c1 = eb.Bases(ctx, 1, scalars=rand_scalars(3, 105)).download()[0]
    c2 = eb.Bases(ctx, 2, scalars=rand_scalars(2, 106)).download()[0]
    idx = np.arange(lo_a, hi_a, dtype=np.uint32)
    key = eb.ProvingKey.from_bases(ctx, A, B2, H, Lq, c1, c2, nv, ni, A_idx=idx, B_idx=idx)
    if world > 1:
        key.set_shard(rank, world, lo_l)
    key.load_circuit(cs.A, cs.B, cs.C, eb.host.fr_to_mont(cs.coeffs), ni, nv)
    return key, witness, dict(m=m, nc=nc, ni=ni, nv=nv, nA=nv + 1, nB=nv + 1, nH=m - 1, nL=nv - ni)


# ----------------------------------------------------------------------------------------------- CPU arm
def host_cpu():
    """(model name, logical cores) of the host the CPU arm runs on"""
    model = "unknown"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    return model, os.cpu_count() or 1


def cpu_prove(log_m, steps, warmup, budget_s, flavour="field", ref_phases=False):
    """Times the CPU implementation of the path on the host cores ON THE BENCHMARK'S OWN INSTANCE (the 2^log_m-constraint
    R1CS, no extrapolation), key generation and key parsing excluded: every step is one whole prove, the scope of the
    reference's "Call to r1cs_gg_ppzksnark_zok_prover" block (...zok.tcc:454-544).
    If oracle/_ref/ref_prove exists (the reference's own prover templates built from /root/reference by oracle/ref/Makefile,
    alt_bn128 build) it is the thing timed (kind "reference"); otherwise the oracle port (kind "port").  All host threads
    are used: OMP_NUM_THREADS is set explicitly, because torchrun exports OMP_NUM_THREADS=1 to its workers.
    The run is bounded: if warmup + steps proves would take longer than budget_s, the warm-up is cut to one prove and the
    timed proves to what fits (never fewer than 5, or `steps` if smaller); the counts actually run are returned.
    Returns dict(times=[seconds per timed prove], warmup=.., threads=.., kind=.., cpu=.., desc=..)."""
    model, cores = host_cpu()
    os.environ["OMP_NUM_THREADS"] = str(cores)        # before liboracle (OpenMP) is loaded: the oracle generates the key
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import orc
    import orc_ref
    use_ref = orc_ref.available()
    m = 1 << log_m
    circ = orc.Circuit.synth_binary(m - 2, 1) if flavour == "binary" else orc.Circuit.synth_field(m - 2, 1)
    w = circ.witness()
    key = orc.Key(circ)
    floor = min(steps, 5)
    if use_ref:
        all_t = orc_ref.prove(circ, w, key, repeats=warmup + steps, threads=cores, all_times=True, budget_s=budget_s,
                              min_repeats=min(warmup, 1) + floor)[2]
    else:
        all_t, used = [], 0.0
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            key.prove(w)
            all_t.append(time.perf_counter() - t0)
            used += all_t[-1]
            if used > budget_s and i + 1 >= min(warmup, 1) + floor:
                break
    wu = warmup if len(all_t) == warmup + steps else min(warmup, 1)
    times = all_t[wu:]
    kind = "reference" if use_ref else "port"
    desc = ("%d whole proves of the 2^%d-constraint %s instance itself (no extrapolation) after %d warm-up, %s, %d OpenMP threads on %s; "
            "the reference MSM uses at most 16 threads (one per window, multiexp.tcc:534-541)" %
            (len(times), log_m, "0/1-witness" if flavour == "binary" else "synthetic", wu,
             "oracle/_ref/ref_prove = the reference's own prover templates, alt_bn128 build (not the reference's default mcl curve code), -O3 -march=x86-64-v3"
             if use_ref else "oracle port", cores, model))
    phases = None
    if use_ref and ref_phases:
        phases = orc_ref.prove_phases(circ, w, key, repeats=2, threads=cores)
    return dict(times=times, warmup=wu, threads=cores, kind=kind, cpu=model, desc=desc, phases=phases)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_prove(args.log_m, args.steps, args.warmup, args.cpu_budget_s, args.flavour, ref_phases=args.ref_phases)
    times = r["times"]
    mean_s, med_s = float(np.mean(times)), float(np.median(times))
    value = 1.0 / mean_s
    line = {
        "impl": "reference", "metric": "groth16_proofs_per_sec", "value": value, "unit": "proofs/s", "n_gpus": args.gpus,
        "steps": len(times), "warmup": r["warmup"], "ms_per_step": mean_s * 1e3, "median_ms_per_step": med_s * 1e3,
        "min_ms_per_step": float(np.min(times)) * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u32x8 (254-bit Montgomery)", "data": "synthetic",
        "config": workload_config(args.log_m, args.flavour),
        "cpu_baseline": {"value": value, "unit": "proofs/s", "cores": r["threads"], "kind": r["kind"], "sample": r["desc"], "cpu": r["cpu"]},
        "e2e": {"value": value, "unit": "proofs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if r["phases"]:
        line["phases_reference_blocks_ms"] = r["phases"]       # the reference's own enter_block timers, mean of 2 extra proves
    print(json.dumps(line))


def workload_config(log_m, flavour):
    if flavour == "binary":
        return {"workload": "0/1-witness R1CS 2^%d constraints (generate_r1cs_example_with_binary_input), full prove (4 MSMs + H NTT chain)" % log_m,
                "log_m": log_m}
    return {"workload": "synthetic R1CS 2^%d constraints, full prove (4 MSMs + H NTT chain)" % log_m, "log_m": log_m}


# ----------------------------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import ethsnarks_b200 as eb

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the GPU arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    shard = world > 1 and args.mode in ("both", "shard")          # headline at N > 1: ONE proof over a key sharded 1/N (strong scaling)
    ctx = eb.Context(local)
    imad_peak = ctx.imad_peak(5)
    if args.window:
        ctx.set_option("window", args.window)
    for kv in args.opt:
        name, _, val = kv.partition("=")
        ctx.set_option(name, int(val))
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed_steps(fn, steps, collect=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = ctx.launches
        e0.record(stream)
        for _ in range(steps):
            fn()
            if collect is not None:
                collect.append(ctx.timings())
        e1.record(stream)
        barrier()
        return max_over_ranks(e0.elapsed_time(e1), dev), ctx.launches - l0

    # the other N > 1 arrangement, measured in the same run and reported beside the headline: N independent replicas, every
    # rank proving its own witnesses against a whole resident key (weak scaling, no data-path collective)
    replicas = None
    if world > 1 and args.mode == "both" and args.log_m <= 22:
        rkey, rwitness, _ = build_workload(eb, ctx, args.log_m, 0, 1, args.flavour)
        rw_dev = torch.from_numpy(rwitness.view(np.int64)).to(dev)
        for _ in range(max(args.warmup, 3)):
            rkey.prove(rw_dev.data_ptr())
        rms, _ = timed_steps(lambda: rkey.prove(rw_dev.data_ptr()), args.steps)
        replicas = {"scaling": "weak", "parallelism": "%d independent replicas, a whole key per GPU, no collective" % world,
                    "ms_per_step": rms / args.steps, "proofs_per_sec": world * args.steps / (rms * 1e-3)}
        rkey.free()
        del rkey, rw_dev
        torch.cuda.synchronize()

    key, witness, dims = build_workload(eb, ctx, args.log_m, rank if shard else 0, world if shard else 1, args.flavour)
    nbytes_w = witness.nbytes
    w_dev = torch.from_numpy(witness.view(np.int64)).to(dev)            # resident witness (`value`)
    w_pin = torch.from_numpy(witness.view(np.int64)).pin_memory()       # pinned host witness (`e2e`)

    # proofs/s is defined over DISTINCT witnesses: the timed loops alternate between two satisfying assignments
    # of the same circuit (the generator's seeds changed); above 2^22 the second one is skipped to bound the set-up time
    w_devs, w_pins = [w_dev], [w_pin]
    if args.log_m <= 22 and args.flavour == "field":      # (the 0/1 generator's seed also changes the circuit: one witness there)
        nc2 = (1 << args.log_m) - 2
        pb2 = eb.generate_r1cs_example_with_field_input(nc2, 1, seed_a=2, seed_b=3)
        w2 = eb.host.fr_to_mont(pb2.values)
        assert w2.shape == witness.shape and not np.array_equal(w2, witness)
        w_devs.append(torch.from_numpy(w2.view(np.int64)).to(dev))
        w_pins.append(torch.from_numpy(w2.view(np.int64)).pin_memory())
        del pb2

    sharded_prove = ShardedProver(key, world, rank, stream, dev, staged=bool(args.staged)) if shard else None

    def step(wptr, host=None):
        """wptr: device or host pointer of the witness; host: the pinned host tensor when the witness starts on the host"""
        return sharded_prove(wptr, host) if shard else key.prove(wptr)

    def timed(bufs, steps, collect):
        it = [0]

        def one():
            b = bufs[it[0] % len(bufs)]
            step(b.data_ptr(), None if b.is_cuda else b)
            it[0] += 1
        return timed_steps(one, steps, collect)

    # pipelined proving: a second prove slot on the same resident key (b200_key_attach), driven from a second host thread,
    # so the start-up (digits + sort) and the latency-bound tail of one proof overlap the next proof's accumulate kernels
    slots = [key]
    ctx2 = None
    if args.in_flight > 1 and not shard:
        ctx2 = eb.Context(local)
        slots.append(key.attach(ctx2))

    def run_steps(wptrs, steps):
        """steps proofs in total over the prove slots, slot j taking witness wptrs[j]"""
        if len(slots) == 1:
            for _ in range(steps):
                step(wptrs[0])
            return
        counts = [steps // len(slots) + (1 if j < steps % len(slots) else 0) for j in range(len(slots))]
        errs = []

        def work(j):
            try:
                for _ in range(counts[j]):
                    slots[j].prove(wptrs[j])
            except Exception as e:        # surfaced after the join
                errs.append(e)
        th = [threading.Thread(target=work, args=(j,)) for j in range(len(slots))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]

    def timed_pipelined(wptrs, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = ctx.launches + (ctx2.launches if ctx2 else 0)
        e0.record(stream)
        run_steps(wptrs, steps)
        torch.cuda.synchronize()          # every slot's stream has drained before the closing event
        e1.record(stream)
        barrier()
        return max_over_ranks(e0.elapsed_time(e1), dev), ctx.launches + (ctx2.launches if ctx2 else 0) - l0

    for _ in range(max(args.warmup, 3)):
        step(w_dev.data_ptr())
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    phases = []
    ms_lat, launches = timed(w_devs, args.steps, None)                      # one proof at a time: latency
    ms_dev, ms_e2e = ms_lat, None
    if len(slots) > 1:
        # each slot proves from its own witness buffer (same values: the MSM/NTT cost does not depend on them)
        w2_dev = w_dev.clone(); w2_pin = w_pin.clone().pin_memory()
        run_steps([w_dev.data_ptr(), w2_dev.data_ptr()], 2 * len(slots))
        ms_dev, launches = timed_pipelined([w_dev.data_ptr(), w2_dev.data_ptr()], args.steps)
        run_steps([w_pin.data_ptr(), w2_pin.data_ptr()], len(slots))
        ms_e2e, _ = timed_pipelined([w_pin.data_ptr(), w2_pin.data_ptr()], args.steps)
    else:
        for wp in w_pins:
            step(wp.data_ptr(), wp)
        ms_e2e, _ = timed(w_pins, args.steps, None)
    # per-kernel durations for the roofline: the same steps with the MSM lanes serialised, so that each kernel's
    # CUDA-event bracket on its launching stream contains that kernel alone
    ctx.set_option("overlap", 0)
    step(w_dev.data_ptr())
    timed(w_devs, min(args.steps, 3), phases)
    ctx.set_option("overlap", 1)
    clocks = sampler.stop() if rank == 0 else None

    verified = None
    if args.verify:
        proof = step(w_dev.data_ptr())
        torch.cuda.synchronize()
        if rank == 0:
            verified = verify_in_the_exponent(eb, ctx, key, witness, proof, dims)
            if not verified:
                raise SystemExit("bench.py: the proof does not match its closed form in the exponent")
    # the N-rank proof must be the proof one GPU gives: rank 0 proves the same witness against the whole key and compares bytes
    identical_to_one_gpu = None
    if shard and args.verify:
        proof_n = step(w_dev.data_ptr())
        torch.cuda.synchronize()
        if rank == 0:
            key1, _, _ = build_workload(eb, ctx, args.log_m, 0, 1, args.flavour)
            proof_1 = key1.prove(w_dev.data_ptr())
            identical_to_one_gpu = all(np.array_equal(np.asarray(a), np.asarray(b)) for a, b in zip(proof_n, proof_1))
            key1.free()
            if not identical_to_one_gpu:
                raise SystemExit("bench.py: the %d-GPU sharded proof differs from the one-GPU proof" % world)
        barrier()
    proofs = args.steps * (1 if (shard or world == 1) else world)
    value = proofs / (ms_dev * 1e-3)
    e2e = proofs / (ms_e2e * 1e-3)

    # per-phase mean over the timed steps (names repeat: the k-th "msm.accumulate" belongs to the k-th MSM: A, B_g2, H, L)
    agg = {}
    outer = ("msm.A", "msm.B_g2", "msm.B_g1", "msm.H", "msm.L")
    for tl in phases:
        cur = None
        for name, ms in tl:
            if name in outer:
                cur = name[4:]
                agg.setdefault(name, []).append(ms)
            elif name.startswith("msm.") and cur:
                agg.setdefault("%s@%s" % (name, cur), []).append(ms)
            else:
                agg.setdefault(name, []).append(ms)
    mean = {k: float(np.mean(v)) for k, v in agg.items()}
    msm_names = ["A", "B_g2", "H", "L"]
    tables_on = not any(kv.split("=")[0] == "tables" and int(kv.split("=")[1]) == 0 for kv in args.opt)
    counts = {"A": key.info()["nA"], "B_g2": key.info()["nB"], "H": key.info()["nH"], "L": key.info()["nL"]}
    kernels = {}
    for nm in msm_names:
        ms = mean.get("msm.accumulate@%s" % nm)
        if ms:
            g2 = nm == "B_g2"
            mac = (MAC_PER_G2_POINT if g2 else MAC_PER_G1_POINT) * counts[nm]
            byts = (BYTES_PER_G2_POINT if g2 else BYTES_PER_G1_POINT) * counts[nm]
            kernels[nm] = {"kernel": "msm_accumulate_g2_pair" if g2 else "msm_accumulate<G1>", "points": counts[nm], "ms": ms,
                           "tmac_s": mac / (ms * 1e-3) / 1e12, "frac_imad": mac / (ms * 1e-3) / imad_peak,
                           "gb_s": byts / (ms * 1e-3) / 1e9}
            if tables_on and counts[nm]:
                # upper bound on the multiply-adds the kernel executes: one mixed addition per (scalar, window) pair; the
                # first point of a bucket run is a copy and zero digits are skipped, so the true count is a little lower
                W = table_windows(counts[nm], args.window)
                mac_x = (MAC_PER_G2_ADD if g2 else MAC_PER_G1_ADD) * W * counts[nm]
                kernels[nm]["executed"] = {"windows": W, "mac_per_point": mac_x // counts[nm], "tmac_s": mac_x / (ms * 1e-3) / 1e12,
                                           "frac_imad": mac_x / (ms * 1e-3) / imad_peak}
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    dom = max(kernels, key=lambda k: kernels[k]["ms"]) if kernels else None
    roofline = None
    if dom:
        kd = kernels[dom]
        roofline = {"bound": "imad", "kernel": kd["kernel"] + " (%s query)" % dom, "achieved": kd["tmac_s"], "peak": imad_peak / 1e12,
                    "unit": "TMAC/s", "frac": kd["frac_imad"],
                    "note": "achieved/frac use SURVEY.md §8(d)'s canonical count (16 windows x 11 / 29 products x 136); the kernels run "
                            "fewer windows (fixed-base tables) and cheaper additions, so frac can exceed 1 - `executed` counts what they run",
                    "executed": kd.get("executed"),
                    "traffic": NCU_TRAFFIC_BYTES.get(("G2" if dom == "B_g2" else "G1", args.log_m)) if world == 1 or not shard else None,
                    "peak_source": "b200_imad_peak microbenchmark (IMAD.WIDE.U32, all SMs), measured in this run",
                    "hbm": {"achieved": kd["gb_s"], "peak": hbm_peak, "unit": "GB/s", "frac": kd["gb_s"] / hbm_peak,
                            "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
                    "g1_msm": kernels.get("H")}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": "groth16_proofs_per_sec", "value": value, "unit": "proofs/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_dev / args.steps, "higher_is_better": True,
        "scaling": "strong" if shard else "weak", "vs_baseline": None, "dtype": "u32x8 (254-bit Montgomery)", "data": "synthetic",
        "config": {**workload_config(args.log_m, args.flavour), "num_inputs": dims["ni"], "msm_lengths": {"A": dims["nA"], "B_g2": dims["nB"], "H": dims["nH"], "L": dims["nL"]},
                   "parallelism": ("ONE proof over a key sharded 1/%d per GPU%s; 256-byte partial sums all-gathered over NCCL on the device" % (world, ", transform chains shared out and broadcast over NCCL" if args.staged else "")) if shard else
                                  (("%d independent replicas" % world if world > 1 else "single GPU") +
                                   (", %d proofs in flight per GPU on one resident key" % len(slots) if len(slots) > 1 else "")),
                   "l2": "inputs larger than L2 (key + sort buffers >> 126 MB per prove)", "distinct_witnesses": len(w_devs), "r_s": "0,0 (reference no-ZK prover)"},
        "latency_ms": ms_lat / args.steps, "proofs_in_flight": len(slots),
        "e2e": {"value": e2e, "unit": "proofs/s", "ms_per_step": ms_e2e / args.steps, "h2d_bytes_per_step": int(nbytes_w),
                "d2h_bytes_per_step": 64 + 128 + 64 + 3 * 4},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
        "phases_ms": {k: round(v, 4) for k, v in sorted(mean.items()) if "@" not in k},
        "phases_reference_blocks_ms": {k: round(v, 4) for k, v in eb.host.phases_as_reference_blocks(mean).items()},
        "msm_subphases_ms": {k: round(v, 4) for k, v in sorted(mean.items()) if "@" in k},
        "msm_kernels": kernels, "verified_in_the_exponent": verified,
    }
    if replicas:
        line["replicas"] = replicas
    if identical_to_one_gpu is not None:
        line["sharded_proof_identical_to_one_gpu"] = identical_to_one_gpu
    if world == 1 and not args.no_cpu_baseline:
        r = cpu_prove(args.log_m, 3, 1, 30.0, args.flavour)        # bounded: 1 warm-up + 3 timed whole proves of this instance
        line["cpu_baseline"] = {"value": 1.0 / float(np.median(r["times"])), "unit": "proofs/s", "cores": r["threads"], "kind": r["kind"],
                                "sample": r["desc"] + "; value = 1 / median", "cpu": r["cpu"]}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--log-m", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="both", choices=["both", "shard", "replicas"])
    ap.add_argument("--cpu-budget-s", type=float, default=120.0, help="--impl reference: bound on the summed prover time")
    ap.add_argument("--ref-phases", action="store_true", help="--impl reference: also report the reference's own block timers")
    ap.add_argument("--flavour", default="field", choices=["field", "binary"],
                    help="R1CS generator: r1cs_examples.tcc:25-93 (uniform field witness) or :100-157 (0/1 witness)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--window", type=int, default=0, help="window size c of the key's fixed-base tables (0 = cost model)")
    ap.add_argument("--opt", action="append", default=[], help="library option name=value (b200_set_option), repeatable")
    ap.add_argument("--in-flight", type=int, default=1, help="prove slots per GPU driven concurrently (1 = one proof at a time)")
    ap.add_argument("--staged", type=int, default=1, help="sharded proofs: 1 = the ranks also share the transform chains (broadcast of the coset evaluations)")
    ap.add_argument("--verify", type=int, default=1, help="check one proof against its closed form in the exponent (untimed)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
