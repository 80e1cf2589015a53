#!/usr/bin/env python
"""Per-proof kernel shares from an ncu launch list (ncu --metrics gpu__time_duration.sum --csv --log-file ...) of bench.py:
keeps the launches of whole proofs (from one h2d/spmv start to the next), prints every kernel's launches and time per proof.
  python tools/ncu_summary.py gpurun_out/r2_launches.csv > profiles/r2_ncu_launches_prove_2p20.txt"""
import collections
import csv
import re
import sys

rows = list(csv.DictReader([l for l in open(sys.argv[1]) if not l.startswith("==")]))
ev = []
for r in rows:
    v = float(r["Metric Value"].replace(",", ""))
    v = v / 1e3 if r["Metric Unit"] == "ns" else v * 1e3 if r["Metric Unit"] == "ms" else v
    ev.append((re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("b200::", ""), v))
starts = [i for i, (n, _) in enumerate(ev) if n.startswith("spmv_rows_kernel") and (i == 0 or not ev[i - 1][0].startswith("spmv_"))]
proofs = [(starts[i], starts[i + 1]) for i in range(len(starts) - 1)]
tot, cnt = collections.Counter(), collections.Counter()
for a, b in proofs:
    for n, v in ev[a:b]:
        tot[n] += v
        cnt[n] += 1
P = len(proofs)
S = sum(tot.values())
print("# %d whole proofs in the list; per proof: %d launches, %.2f ms summed kernel time (cold-cache, serialised under ncu: compare SHARES)" %
      (P, sum(cnt.values()) // P, S / P / 1e3))
print("%-58s %8s %10s %7s" % ("kernel", "launches", "us/proof", "share"))
for n, v in tot.most_common():
    print("%-58s %8.1f %10.1f %6.1f%%" % (n[:58], cnt[n] / P, v / P, 100 * v / S))
