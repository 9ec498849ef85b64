"""Summary of the per-CTA timeline printed by the ME_HME_TRACE build (tools/trace_hme.sh): CTA durations, time per stage,
co-residency per SM. Usage: python tools/trace_cta.py gpurun_out/trace_cta.txt [b64 per picture]"""
import collections
import statistics
import sys

lines = [l.split() for l in open(sys.argv[1]) if l.startswith("CTA")]
NB = int(sys.argv[2]) if len(sys.argv) > 2 else 510
L = lines[-NB:]  # the last picture
t0 = min(int(l[5]) for l in L)
rows = [(int(l[1]), int(l[3]), int(l[5]) - t0, [int(x) for x in l[7:16]]) for l in L]
dur = [r[3][8] for r in rows]
print("CTA start offsets us: min %.1f max %.1f" % (min(r[2] for r in rows) / 1e3, max(r[2] for r in rows) / 1e3))
print("CTA duration us: min %.1f median %.1f mean %.1f max %.1f" % (min(dur) / 1e3, statistics.median(dur) / 1e3,
                                                                    statistics.mean(dur) / 1e3, max(dur) / 1e3))
print("kernel span us %.1f" % (max(r[2] + r[3][8] for r in rows) / 1e3))
for s in range(9):
    d = [r[3][s] - (r[3][s - 1] if s else 0) for r in rows]
    print("stage %d mean %.1f us  max %.1f" % (s, statistics.mean(d) / 1e3, max(d) / 1e3))
persm = collections.Counter(r[1] for r in rows)
print("CTAs per SM:", dict(collections.Counter(persm.values())))
for n in sorted(set(persm.values())):
    d = [r[3][8] for r in rows if persm[r[1]] == n]
    print(" SMs with %d CTAs: mean CTA duration %.1f us" % (n, statistics.mean(d) / 1e3))
print("-- up to the end of stage 6 (before any CTA prints):")
for n in sorted(set(persm.values())):
    d = [r[3][6] for r in rows if persm[r[1]] == n]
    print(" SMs with %d CTAs: mean time to end of stage 6 %.1f us" % (n, statistics.mean(d) / 1e3))
    for s in (1, 3, 4, 5, 6):
        d = [r[3][s] - r[3][s - 1] for r in rows if persm[r[1]] == n]
        print("    stage %d mean %.1f us" % (s, statistics.mean(d) / 1e3))
