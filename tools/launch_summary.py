"""Per-kernel summary of an ncu launch list (`--metrics gpu__time_duration.sum --csv`). Usage: launch_summary.py launches.csv"""
import collections
import csv
import sys

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
h = rows[0]
ik, iv, iu = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
t = collections.defaultdict(list)
for r in rows[1:]:
    v = float(r[iv].replace(",", ""))
    v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[iu], 1.0)
    t[r[ik].split("(")[0]].append(v)
engine = ("k_hme", "k_fullsearch", "k_epilogue", "k_downsample_pad", "k_repitch", "k_pad_full", "k_dg_hme")
tot = sum(sum(v) for k, v in t.items() if k in engine)
print(f"ncu launch list {sys.argv[1]} (gpu__time_duration.sum, --clock-control none; cold-cache, serialised)\n")
for k, v in sorted(t.items(), key=lambda kv: -sum(kv[1])):
    share = f"{100 * sum(v) / tot:5.1f}% of the engine's kernels" if k in engine else "(not part of a step)"
    print(f"{k[:60]:60s} n={len(v):4d} mean {sum(v) / len(v):8.1f} us  total {sum(v) / 1e3:8.2f} ms  {share}")
