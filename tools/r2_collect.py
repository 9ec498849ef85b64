"""Copies what tools/r2_run.sh left in gpurun_out/ into profiles/ (tracked): bench lines, the launch list and its
summary, per-kernel section + source-line summaries of the `ncu --set full` captures, and profiles/ncu_traffic.json (the
DRAM bytes per launch bench.py reports as roofline.traffic) from the dram__bytes pass.
Usage: python tools/r2_collect.py [tag]        (tag defaults to r2)"""
import collections
import csv
import json
import os
import shutil
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
O, P = os.path.join(REPO, "gpurun_out"), os.path.join(REPO, "profiles")
TAG = sys.argv[1] if len(sys.argv) > 1 else "r2"
commit = subprocess.run(["git", "-C", REPO, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
dirty = subprocess.run(["git", "-C", REPO, "status", "--porcelain", "--", "svt-av1-psyex_b200/csrc"], capture_output=True, text=True).stdout.strip()
commit += "+uncommitted csrc changes" if dirty else ""

# ---- bench lines ----
for f in sorted(os.listdir(O)):
    if f.startswith(f"{TAG}_bench_") and f.endswith(".json") and os.path.getsize(os.path.join(O, f)):
        shutil.copy(os.path.join(O, f), os.path.join(P, f.replace(f"{TAG}_bench_", f"bench_{TAG}_")))

# ---- multi-GPU lines (tools/r2_multi.sh): bench lines, copy-rate probe, topology, the two-GPU tests' log ----
import re
for f in sorted(os.listdir(O)):
    m = re.match(rf"{TAG}_(\d+)gpu_(bench_.*\.json|h2d.*\.json|topo\.txt|tests\.log)$", f)
    if m and os.path.getsize(os.path.join(O, f)):
        shutil.copy(os.path.join(O, f), os.path.join(P, f"{TAG}_{m.group(1)}gpu_{m.group(2)}"))

# ---- launch list ----
src = os.path.join(O, f"{TAG}_launches.csv")
if os.path.exists(src):
    shutil.copy(src, os.path.join(P, f"launches_{TAG}.csv"))
    out = subprocess.run([sys.executable, os.path.join(REPO, "tools", "launch_summary.py"), src], capture_output=True, text=True).stdout
    open(os.path.join(P, f"launches_{TAG}_summary.txt"), "w").write(out.replace(O + "/", "gpurun_out/"))

# ---- DRAM bytes per launch ----
entries = {}
try:
    entries = json.load(open(os.path.join(P, "ncu_traffic.json")))["entries"]
except Exception:
    pass
for f in sorted(os.listdir(O)):
    if not (f.startswith(f"{TAG}_dram_") and f.endswith(".csv")):
        continue
    wl, kind = f[len(f"{TAG}_dram_"):-4].rsplit("_", 1)
    rows = [r for r in csv.reader(l for l in open(os.path.join(O, f)) if l.startswith('"'))]
    if len(rows) < 2:
        continue
    h = rows[0]
    ik, im, iv, iu, iid = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("ID")
    per = collections.defaultdict(lambda: collections.defaultdict(float))  # kernel -> launch id -> bytes
    for r in rows[1:]:
        if not r[im].startswith("dram__bytes"):
            continue
        v = float(r[iv].replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[r[iu]]
        per[r[ik].split("(")[0]][r[iid]] += v
    shutil.copy(os.path.join(O, f), os.path.join(P, f"ncu_{TAG}_dram_{wl}_{kind}.csv"))
    for k, launches in per.items():
        vals = list(launches.values())
        entries[f"{wl}/{kind}/{k}"] = {
            "dram_bytes_per_launch": sum(vals) / len(vals), "launches": len(vals), "min": min(vals), "max": max(vals),
            "capture": f"ncu_{TAG}_dram_{wl}_{kind}.csv", "commit": commit,
            "how": "ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum --clock-control none over the 16 pictures of one step "
                   "(tools/profile_step.py 16), mean per launch"}
json.dump({"note": "roofline.traffic of bench.py; regenerate with tools/r2_run.sh 3 + tools/r2_collect.py after a kernel change",
           "entries": entries}, open(os.path.join(P, "ncu_traffic.json"), "w"), indent=1, sort_keys=True)

# ---- full captures: sections + source lines ----
for f in sorted(os.listdir(O)):
    if f.startswith(f"{TAG}_k_") and f.endswith(".ncu-rep"):
        name = f[len(TAG) + 1:-8]
        rep = os.path.join(O, f)
        det = subprocess.run(["ncu", "-i", rep, "--page", "details"], capture_output=True, text=True).stdout
        open(os.path.join(P, f"ncu_{TAG}_{name}_details.txt"), "w").write(det)
        lines = subprocess.run([sys.executable, os.path.join(REPO, "tools", "ncu_lines.py"), rep, "40"], capture_output=True, text=True).stdout
        open(os.path.join(P, f"ncu_{TAG}_{name}_lines.txt"), "w").write(lines)
        ops = subprocess.run([sys.executable, os.path.join(REPO, "tools", "ncu_opcodes.py"), rep, "40"], capture_output=True, text=True).stdout
        open(os.path.join(P, f"ncu_{TAG}_{name}_opcodes.txt"), "w").write(ops)
print("\n".join(sorted(f for f in os.listdir(P) if TAG in f)))
