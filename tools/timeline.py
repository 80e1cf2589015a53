#!/usr/bin/env python
"""Prints the phase timeline (start, end on the device clock) of one prove with the MSM lanes overlapped.

  python tools/timeline.py [--log-m 20]
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import ethsnarks_b200 as eb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log-m", type=int, default=20)
    args = ap.parse_args()
    import torch
    ctx = eb.Context(0)
    key, witness, dims = bench.build_workload(eb, ctx, args.log_m, 0, 1)
    w_dev = torch.from_numpy(witness.view(np.int64)).to("cuda:0")
    for overlap in (1, 0):
        ctx.set_option("overlap", overlap)
        for _ in range(3):
            key.prove(w_dev.data_ptr())
        tl = ctx.timeline()
        print("overlap=%d" % overlap)
        for name, at, ms in tl:
            print("  %-18s start %8.3f  end %8.3f  (%7.3f ms)" % (name, at, at + ms, ms))


if __name__ == "__main__":
    main()
