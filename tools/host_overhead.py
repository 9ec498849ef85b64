"""Wall-clock split of one bench step: submit loop vs wait loop, resident pictures (diagnostic)."""
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import torch  # noqa: E402

import bench  # noqa: E402
from me_b200.engine import Engine  # noqa: E402

pics = bench.make_pictures("motion", pinned=True)
jobs = {poc: bench.derive(poc) for poc in bench.STEP_POCS}
outs = None
eng = Engine(bench.W, bench.H, 4, 3, pictures_in_flight=40, devices=[0])
outs = {poc: bench.pinned_outputs(jobs[poc][1], eng) for poc in bench.STEP_POCS}
for poc in range(bench.NPIC):
    a, _ = pics[poc]
    eng.register_picture_ptr(poc, a.ctypes.data, a.strides[0], bench.W, bench.H)
torch.cuda.synchronize()
for rep in range(6):
    t0 = time.perf_counter()
    for poc in bench.STEP_POCS:
        _, p, l0, l1 = jobs[poc]
        eng.submit_out(poc, p, [l0, l1], outs[poc])
    t1 = time.perf_counter()
    waits = []
    for poc in bench.STEP_POCS:
        tw = time.perf_counter()
        eng.wait(poc, outs[poc])
        waits.append((time.perf_counter() - tw) * 1e3)
    t2 = time.perf_counter()
    print(f"rep {rep}: submit x16 {1e3 * (t1 - t0):.2f} ms, wait x16 {1e3 * (t2 - t1):.2f} ms, total {1e3 * (t2 - t0):.2f} ms; "
          f"waits {[round(w, 2) for w in waits]}")
# one at a time
for poc in bench.STEP_POCS[:4]:
    _, p, l0, l1 = jobs[poc]
    t0 = time.perf_counter()
    eng.submit(poc, p, [l0, l1])
    t1 = time.perf_counter()
    eng.wait(poc, outs[poc])
    t2 = time.perf_counter()
    t = eng.timings()
    print(f"poc {poc}: submit {1e3 * (t1 - t0):.3f} ms wait {1e3 * (t2 - t1):.3f} ms; device total {t.total_ms:.3f} hme {t.hme_ms:.3f} search {t.search_ms:.3f} epi {t.epilogue_ms:.3f}")
eng.close()
