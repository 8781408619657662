"""Aggregates an ncu launch list (`--metrics gpu__time_duration.sum --csv`) per kernel name:
launches, total and mean duration, share of the GPU time of the command.
Usage: python profiles/summarize_launches.py gpurun_out/r2_launches_*.csv > profiles/r2_launches.txt"""
import csv
import sys
from collections import OrderedDict

for path in sys.argv[1:]:
    rows = []
    with open(path, newline="") as f:
        for r in csv.reader(f):
            if len(r) >= 15 and r[12] == "gpu__time_duration.sum":
                rows.append((r[4] + "  grid " + r[8] + " block " + r[7], float(r[14]) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[13], 1e-6)))
    agg = OrderedDict()
    for name, ms in rows:
        a = agg.setdefault(name, [0, 0.0, float("inf"), 0.0])
        a[0] += 1
        a[1] += ms
        a[2] = min(a[2], ms)
        a[3] = max(a[3], ms)
    total = sum(a[1] for a in agg.values()) or 1.0
    print(f"== {path.split('/')[-1]}: {len(rows)} launches, {total:.3f} ms of GPU time (cold-cache, serialised: shares, not absolutes)")
    print("     launches x mean ms = total ms (share)  [min .. max ms: one grid shape may serve batches of several sizes]")
    for name, (n, ms, lo, hi) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        ours = "*" if ("voda" in name or "unnamed" in name or "afsl" in name) else " "
        print(f"  {ours} {n:4d} x {ms / n:10.4f} ms = {ms:10.3f} ms ({100 * ms / total:5.1f} %)  [{lo:8.4f} .. {hi:8.4f}]  {name[:110]}")
    print()
