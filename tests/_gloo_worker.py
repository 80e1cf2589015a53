"""world_size-2 gloo worker for tests/test_host.py: exercises bench.py's sharding + gather helpers without a GPU."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# shard ranges partition [0, n) exactly
for n in (0, 1, 7, 1000, (1 << 20) - 1):
    lo, hi = bench.shard_range(n, rank, world)
    t = torch.tensor([lo, hi], dtype=torch.int64)
    allr = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allr, t)
    assert int(allr[0][0]) == 0 and int(allr[-1][1]) == n
    for a, b in zip(allr[:-1], allr[1:]):
        assert int(a[1]) == int(b[0])
# partial gather: every rank contributes one packed 256-byte record (32 x int64), rank order preserved
part = torch.full((32,), rank + 1, dtype=torch.int64)
parts = torch.zeros(32 * world, dtype=torch.int64)
bench.all_gather_packed(parts, part)
assert all(int(parts[32 * r]) == r + 1 and int(parts[32 * r + 31]) == r + 1 for r in range(world))
ms = bench.max_over_ranks(float(rank + 1), device="cpu")
assert ms == float(world)
dist.barrier()
if rank == 0:
    print("GLOO_OK")
dist.destroy_process_group()
