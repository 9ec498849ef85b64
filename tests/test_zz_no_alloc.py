"""me_gpu_dg_detector_noalloc (the entry point that allocates nothing: scratch in a static device array) against the
existing me_gpu_dg_detector on the same pictures, through the C-ABI, several calls in a row. Collected last on purpose:
a failure here hides no other test. Runs alone in a fresh process as well as after the
other GPU tests (it brings up and shuts down an engine of its own for the existing entry point; the new one needs none)."""
import ctypes as C

import numpy as np
import pytest

from helpers import DgPictures
from me_b200 import abi

pytestmark = pytest.mark.gpu

# (width, height, input_resolution class, kind, source poc, reference poc): two sizes, ties ("flat"), a repeat of the
# first pair at the end (the shared scratch keeps the previous call's planes and sums: they must not leak)
CALLS = [(320, 192, 4, "motion", 4, 3), (1920, 1080, 4, "motion", 24, 16), (320, 192, 2, "blocks", 1, 2),
         (200, 136, 2, "flat", 2, 1), (320, 192, 4, "motion", 4, 3)]


def _noalloc(lib, src, ref, w, h, res):
    m = abi.MeGpuDgMetrics()
    a, b = src.planes[2], ref.planes[2]  # padded sixteenth planes, row stride = width / 4 + 32
    rc = lib.me_gpu_dg_detector_noalloc(a.ctypes.data, a.strides[0], b.ctypes.data, b.strides[0], w, h, res, C.byref(m))
    assert rc == 0, (rc, lib.me_gpu_dg_detector_noalloc_error())
    return m.as_tuple()


def test_noalloc_entry_point_equals_the_existing_one():
    from checkers import Oracle
    from me_b200.engine import Engine
    oracle = Oracle()
    lib = abi.load_gpu_lib()
    assert lib.me_gpu_dg_detector_noalloc_setup() == 0, lib.me_gpu_dg_detector_noalloc_error()
    sets = {}
    expected = []
    eng = Engine(1920, 1088, 4, 3, pictures_in_flight=6)
    try:
        for (w, h, res, kind, a, b) in CALLS:
            ps = sets.setdefault((w, h, kind), DgPictures(w, h, kind, oracle))
            eng.register_picture(a, ps.get(a).planes[0], w, h)
            eng.register_picture(b, ps.get(b).planes[0], w, h)
            expected.append(eng.dg_detector(a, b, res))  # the existing entry point: pool slots, cudaMalloc'ed at init
            # the planes the new entry point is fed are the ones the device built for the existing one
            for k in (a, b):
                assert np.array_equal(eng.read_pyramid(k, w, h)[2], ps.get(k).planes[2])
            eng.release_picture(a)
            eng.release_picture(b)
    finally:
        eng.close()
    # back to back, no engine alive: integer sums, so the two entry points agree exactly (tolerance 0)
    got = [_noalloc(lib, sets[(w, h, kind)].get(a), sets[(w, h, kind)].get(b), w, h, res) for (w, h, res, kind, a, b) in CALLS]
    assert got == expected, (got, expected)
    assert got[0] == got[-1]
    # a second round over the same calls gives the same results (nothing of call n survives into call n + 1)
    again = [_noalloc(lib, sets[(w, h, kind)].get(a), sets[(w, h, kind)].get(b), w, h, res) for (w, h, res, kind, a, b) in CALLS]
    assert again == expected


def test_noalloc_entry_point_rejects_what_does_not_fit():
    lib = abi.load_gpu_lib()
    m = abi.MeGpuDgMetrics()
    w, h = 7680, 4320  # sixteenth planes of 2 x 2.2 MB: beyond ME_GPU_DG_NOALLOC_MAX_BYTES
    plane = np.zeros((h // 4 + 32, w // 4 + 32), dtype=np.uint8)
    rc = lib.me_gpu_dg_detector_noalloc(plane.ctypes.data, plane.strides[0], plane.ctypes.data, plane.strides[0], w, h, 4, C.byref(m))
    assert rc == -4, rc  # ME_GPU_ERR_NO_MEMORY
    assert b"too large for the static scratch" in lib.me_gpu_dg_detector_noalloc_error()
    small = np.zeros((192 // 4 + 32, 320 // 4 + 32), dtype=np.uint8)
    assert lib.me_gpu_dg_detector_noalloc(small.ctypes.data, 16, small.ctypes.data, small.strides[0], 320, 192, 4, C.byref(m)) == -2
    assert lib.me_gpu_dg_detector_noalloc(None, 112, small.ctypes.data, 112, 320, 192, 4, C.byref(m)) == -2
    # the 4K maximum itself fits
    w, h = 3840, 2160
    big = np.full((h // 4 + 32, w // 4 + 32), 90, dtype=np.uint8)
    assert lib.me_gpu_dg_detector_noalloc(big.ctypes.data, big.strides[0], big.ctypes.data, big.strides[0], w, h, 4, C.byref(m)) == 0
    assert m.tot_dist == 0 and m.tot_cplx == 0  # identical constant pictures: every position ties at SAD 0
