"""Device time of the dynamic-GOP detector (k_dg_hme) on 1080p / 4K pairs, wall-clock per call included."""
import os
import sys
import time

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench  # noqa: E402,F401  (sets sys.path for me_b200)
import numpy as np  # noqa: E402
from me_b200.engine import Engine  # noqa: E402

for (w, h, res) in ((1920, 1080, 4), (3840, 2160, 5)):
    eng = Engine(w, h, 4, 3, pictures_in_flight=8, devices=[0])
    rng = np.random.default_rng(3)
    for k in range(4):
        eng.register_picture_unpadded(k, rng.integers(0, 256, size=(h, w), dtype=np.uint8))
    for _ in range(5):
        eng.dg_detector(1, 0, res)
    n = 200
    t0 = time.perf_counter()
    for i in range(n):
        m = eng.dg_detector(1 + (i & 1), 0, res)
    dt = (time.perf_counter() - t0) / n
    nb = ((w + 63) // 64) * ((h + 63) // 64)
    print(f"{w}x{h}: {dt * 1e6:.1f} us per pair (call to result), {nb} b64, ~{nb * 128 * 128 * 256 / dt / 1e12:.2f} T byte-SAD/s incl. host, metrics {m}")
    eng.close()
