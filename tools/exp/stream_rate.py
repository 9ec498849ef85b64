"""frames/s of pictures streamed through one engine (window of W in flight), any size: stream_rate.py w:h:preset:r0:r1 [window] [reps]"""
import collections, os, sys, time
REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, REPO)
import bench
from me_b200 import planes
from me_b200.engine import Engine
w_, h_, pr_, r0_, r1_ = (int(x) for x in sys.argv[1].split(":"))
window = int(sys.argv[2]) if len(sys.argv) > 2 else 12
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
bench.WORKLOAD, bench.W, bench.H, bench.PRESET, bench.FORCE_REFS = sys.argv[1], w_, h_, pr_, (r0_, r1_)
pics = bench.make_pictures("motion", pinned=False)
eng = Engine(bench.W, bench.H, 4, 3, pictures_in_flight=40, devices=[0])
for poc in range(bench.NPIC):
    a, _ = pics[poc]
    eng.register_picture_ptr(poc, a.ctypes.data, a.strides[0], bench.W, bench.H)
jobs = {poc: bench.derive(poc) for poc in bench.STEP_POCS}
outs = {poc: bench.pinned_outputs(jobs[poc][1], eng) for poc in bench.STEP_POCS}
fl = collections.deque()
def run(n):
    t0 = time.perf_counter()
    for _ in range(n):
        for poc in bench.STEP_POCS:
            while len(fl) >= window:
                q = fl.popleft(); eng.wait(q, outs[q])
            _, p, l0, l1 = jobs[poc]
            eng.submit_out(poc, p, [l0, l1], outs[poc])
            fl.append(poc)
    while fl:
        q = fl.popleft(); eng.wait(q, outs[q])
    return time.perf_counter() - t0
run(3)
dt = run(reps)
nb = ((w_ + 63) // 64) * ((h_ + 63) // 64)
print(f"{sys.argv[1]} window {window}: {16 * reps / dt:.1f} frames/s, {1e6 * dt / (16 * reps):.1f} us per picture, {1e6 * dt / (16 * reps) * 510 / nb:.1f} us per 510 b64")
eng.close()
