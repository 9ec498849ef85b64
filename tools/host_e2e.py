"""Host-side cost of the end-to-end loop (diagnostic): time inside register / submit / wait / release per picture."""
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
eng = Engine(bench.W, bench.H, 4, 3, pictures_in_flight=60, devices=[0])
outs = {poc: bench.pinned_outputs(jobs[poc][1], eng) for poc in bench.STEP_POCS}
for poc in range(17):
    a, _ = pics[poc]
    eng.register_picture_ptr(poc, a.ctypes.data, a.strides[0], bench.W, bench.H)
torch.cuda.synchronize()
acc = {"reg": 0.0, "submit": 0.0, "wait": 0.0, "release": 0.0}
n = 0
for rep in range(8):
    t_rep = time.perf_counter()
    for poc in bench.STEP_POCS:
        a, _ = pics[poc]
        _, p, l0, l1 = jobs[poc]
        t0 = time.perf_counter()
        eng.register_picture_ptr(poc, a.ctypes.data, a.strides[0], bench.W, bench.H)
        t1 = time.perf_counter()
        eng.submit_out(poc, p, [l0, l1], outs[poc])
        t2 = time.perf_counter()
        if rep >= 2:
            acc["reg"] += t1 - t0
            acc["submit"] += t2 - t1
    for poc in bench.STEP_POCS:
        t0 = time.perf_counter()
        eng.wait(poc, outs[poc])
        t1 = time.perf_counter()
        if rep >= 2:
            acc["wait"] += t1 - t0
    for poc in bench.STEP_POCS:
        t0 = time.perf_counter()
        eng.release_picture(poc)
        t1 = time.perf_counter()
        if rep >= 2:
            acc["release"] += t1 - t0
    if rep >= 2:
        n += 16
    print(f"rep {rep}: {1e3 * (time.perf_counter() - t_rep):.2f} ms")
print({k: f"{1e6 * v / n:.1f} us/picture" for k, v in acc.items()})
eng.close()
