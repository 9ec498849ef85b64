"""Runs a few pictures of a bench workload (for ncu / launch lists). Usage: profile_step.py [npictures] [kind] [workload | w:h:preset:refs0:refs1]"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench  # noqa: E402
from me_b200 import planes  # noqa: E402
from me_b200.engine import Engine  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
kind = sys.argv[2] if len(sys.argv) > 2 else "motion"
wl = sys.argv[3] if len(sys.argv) > 3 else "1080p-m4"
if ":" in wl:  # experiments: width:height:preset:refs0:refs1 (e.g. 3840:2160:4:2:2 = the 1080p workload's controls on 4x the b64s)
    w_, h_, pr_, r0_, r1_ = (int(x) for x in wl.split(":"))
    bench.WORKLOAD, bench.W, bench.H, bench.PRESET, bench.FORCE_REFS = wl, w_, h_, pr_, (r0_, r1_)
else:
    bench.set_workload(wl)
pics = bench.make_pictures(kind, pinned=False)
eng = Engine(bench.W, bench.H, 4, 3, pictures_in_flight=40, devices=[0])
for poc in range(bench.NPIC):
    a, _ = pics[poc]
    eng.register_picture_ptr(poc, a.ctypes.data, a.strides[0], bench.W, bench.H)
eng.set_profiling(True)
for poc in bench.STEP_POCS[:n]:
    _, p, l0, l1 = bench.derive(poc)
    out = planes.Outputs(bench.W, bench.H, p, debug=False)
    eng.submit(poc, p, [l0, l1])
    eng.wait(poc, out)
    t = eng.timings()
    print(f"poc {poc:2d} refs {l0} {l1}: total {t.total_ms:.3f} ms  hme {t.k_search_ms:.3f}  fullsearch {t.k_fullsearch_ms:.3f}  "
          f"epilogue {t.k_epilogue_ms:.3f}  byte-SADs {t.byte_sad_ops/1e9:.2f} G (full search {t.byte_sad_ops_fullsearch/1e9:.2f} G)")
eng.close()
