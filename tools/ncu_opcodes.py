"""SASS opcode histogram (executed warp instructions, thread utilisation) of one kernel of an ncu report.
Usage: python tools/ncu_opcodes.py report.ncu-rep [top_n]"""
import collections
import csv
import io
import subprocess
import sys

txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(io.StringIO(txt)))
h = next(r for r in rows if r and r[0] == "Address")
isrc, ie, it = h.index("Source"), h.index("Instructions Executed"), h.index("Thread Instructions Executed")
c, th, full = collections.Counter(), collections.Counter(), collections.Counter()
tot = 0
for r in rows:
    if not r or not r[0].startswith("0x"):
        continue
    n = int(r[ie])
    s = r[isrc].strip()
    if s.startswith("@"):
        s = s.split(" ", 1)[1].strip()
    op = s.split(" ")[0]
    c[op.split(".")[0]] += n
    full[op] += n
    th[op.split(".")[0]] += int(r[it])
    tot += n
print(f"warp instructions executed {tot}")
for k, v in c.most_common(top):
    print(f"{k:14s} {v:12d} {100 * v / tot:5.1f}%   avg active threads {th[k] / max(v, 1):5.1f}")
print("--- with modifiers")
for k, v in full.most_common(top):
    print(f"{k:30s} {v:12d} {100 * v / tot:5.1f}%")
