"""Per-source-line instruction and stall-sample summary of an ncu report (needs -lineinfo + --import-source on).
Usage: python tools/ncu_lines.py report.ncu-rep [top_n]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
cur, hdr = None, None
inst, samp, src = collections.Counter(), collections.Counter(), {}
stall = collections.defaultdict(collections.Counter)
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
        ie, isamp = hdr.index("Instructions Executed"), hdr.index("# Samples")
        st = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    elif r[0].isdigit() and hdr:
        k = (cur, int(r[0]))
        src[k] = r[1]
        try:
            inst[k] += int(r[ie])
            samp[k] += int(r[isamp])
            for i, h in st:
                if r[i] not in ("", "-", "0"):
                    stall[k][h] += int(r[i])
        except ValueError:
            pass
ti, ts = sum(inst.values()), sum(samp.values())
print(f"total warp instructions {ti}, samples {ts}")
byf = collections.Counter()
for k, v in samp.items():
    byf[k[0]] += v
print("samples by file:", {k: f"{100 * v / ts:.1f}%" for k, v in byf.most_common(6)})
allst = collections.Counter()
for k in stall:
    allst.update(stall[k])
print("stall reasons:", {k: f"{100 * v / ts:.1f}%" for k, v in allst.most_common(8)})
print("--- top lines by samples")
for k, v in samp.most_common(top):
    print(f"{k[0]}:{k[1]:<4d} samp {100 * v / ts:5.1f}%  inst {100 * inst[k] / ti:5.1f}%  {dict(stall[k].most_common(2))}  {src[k].strip()[:80]}")
print("--- top lines by instructions")
for k, v in inst.most_common(top):
    print(f"{k[0]}:{k[1]:<4d} inst {100 * v / ti:5.1f}%  samp {100 * samp[k] / ts:5.1f}%  {src[k].strip()[:90]}")
