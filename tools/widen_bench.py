"""Throughput of the two SURVEY 8(f) rows built on top of the open-loop path, GPU engine vs the reference's own CPU code
(oracle/_ref, AVX2 kernels, all host threads), 1080p synthetic "motion" pictures:

  tf  temporal-filter ME (me_type = ME_MCTF): (centre, neighbour) pairs per second, tf hme_me_level 0 / 1 / 2
  dg  dynamic-GOP detector HME level 0: (source, reference) pairs per second

Prints one JSON line per measurement. Usage: python tools/widen_bench.py [--no-cpu]
"""
import collections
import ctypes as C
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
import bench  # noqa: E402  (sets sys.path for me_b200)
import torch  # noqa: E402
from helpers import derive_tf  # noqa: E402
import checkers  # noqa: E402
from me_b200 import planes, synth  # noqa: E402
from me_b200.engine import Engine  # noqa: E402

W, H, NPIC = bench.W, bench.H, 16
NO_CPU = "--no-cpu" in sys.argv
# a filtered picture takes its 3 past and 3 future neighbours (tf window of 7); centres 3..12
PAIRS = [(c, c + d) for c in range(3, 13) for d in (-3, -2, -1, 1, 2, 3)]


def gpu_tf(eng, level, reps=6):
    jobs = []
    for (c, nb) in PAIRS:
        _, p = derive_tf(level, W, H, c, me_exit_th=0)
        jobs.append((c, nb, p, planes.Outputs(W, H, p, debug=False)))
    ops = 0

    def run():
        nonlocal ops
        # one centre can only be in flight once: jobs are keyed by the source picture number
        inflight = collections.deque()
        ops = 0
        for (c, nb, p, out) in jobs:
            while any(x[0] == c for x in inflight) or len(inflight) >= 12:
                c0, o0 = inflight.popleft()
                eng.wait(c0, o0)
                ops += eng.timings().byte_sad_ops
            eng.submit_out(c, p, [[nb], []], out)
            inflight.append((c, out))
        while inflight:
            c0, o0 = inflight.popleft()
            eng.wait(c0, o0)
            ops += eng.timings().byte_sad_ops

    run()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        run()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    return len(PAIRS) / dt, ops / dt


def cpu_tf(ref, pics, level, threads):
    work, keep = [], []
    for (c, nb) in PAIRS[:12]:
        _, p = derive_tf(level, W, H, c, me_exit_th=0)
        out = planes.Outputs(W, H, p, debug=False)
        arr, k2 = checkers._refs_array([[pics[nb]], []])
        s = pics[c].c_struct()
        keep.append((out, arr, k2, s, p))
        for first in range(0, 510, 6):
            work.append((s, arr, p, out, first))
    s, arr, p, out, _ = work[0]
    ref.lib.ref_me_b64_range(C.byref(s), C.byref(arr), C.byref(p), C.byref(out.c), 0, 1, 1)

    def one(item):
        s, arr, p, out, first = item
        assert ref.lib.ref_me_b64_range(C.byref(s), C.byref(arr), C.byref(p), C.byref(out.c), first, 6, 1) == 0

    with ThreadPoolExecutor(threads) as pool:
        list(pool.map(one, work[:len(work) // 4]))
        t0 = time.perf_counter()
        list(pool.map(one, work))
        dt = time.perf_counter() - t0
        reps = max(1, int(3.0 / dt))  # about 3 s of wall time
        t0 = time.perf_counter()
        for _ in range(reps):
            list(pool.map(one, work))
        dt = time.perf_counter() - t0
    return 12 * reps / dt, dt


def main():
    eng = Engine(W, H, 4, 3, pictures_in_flight=24, devices=[0])
    seq = synth.Sequence(W, H, NPIC, kind="motion")
    ref = None if NO_CPU else checkers.Reference()
    pics = {}
    for k in range(NPIC):
        p = planes.Picture(W, H, k)
        p.set_luma(seq.luma(k))
        if ref:
            ref.build_pyramid(p)
        pics[k] = p
        eng.register_picture(k, p.planes[0], W, H)
    threads = os.cpu_count() or 1
    peak = 4.0 * eng.sad_peak()
    for level in (0, 1, 2):
        v, ops_s = gpu_tf(eng, level)
        line = {"metric": "temporal-filter ME pairs/sec", "value": v, "unit": "pairs/s", "config": {"workload": f"1080p motion, 7-picture window, tf hme_me_level {level}, pictures resident"},
                "roofline": {"bound": "int_sad", "achieved": ops_s / 1e12, "peak": peak / 1e12, "unit": "Tbyte-SAD/s", "frac": ops_s / peak}}
        if ref:
            cv, dt = cpu_tf(ref, pics, level, threads)
            line["cpu_baseline"] = {"value": cv, "unit": "pairs/s", "cores": threads, "kind": "reference", "sample": f"12 pairs x {round(cv * dt / 12)} repetitions, all 510 b64, {dt:.1f} s wall"}
        print(json.dumps(line), flush=True)
    # dynamic-GOP detector
    pairs = [(a, b) for a in range(8, 16) for b in (0, 4)]
    for (a, b) in pairs[:4]:
        eng.dg_detector(a, b, 4)
    t0 = time.perf_counter()
    reps = 20
    for _ in range(reps):
        for (a, b) in pairs:
            eng.dg_detector(a, b, 4)
    dt = (time.perf_counter() - t0) / (reps * len(pairs))
    ops = 510 * 128 * 128 * 256  # upper bound: unclamped 128 x 128 areas of a 16x16 block
    line = {"metric": "dynamic-GOP detector pairs/sec", "value": 1.0 / dt, "unit": "pairs/s", "config": {"workload": "1080p motion, synchronous calls, pictures resident"},
            "roofline": {"bound": "int_sad", "achieved": ops / dt / 1e12, "peak": peak / 1e12, "unit": "Tbyte-SAD/s", "frac": ops / dt / peak,
                         "note": "call-to-result wall time (launch + 24-byte read-back included); byte-SADs counted for unclamped areas"}}
    if ref:
        def one(pair):
            return ref.dg_detector(pics[pair[0]], pics[pair[1]], 4, simd=1)
        ref.dg_detector(pics[8], pics[0], 4, simd=1)
        with ThreadPoolExecutor(threads) as pool:
            t0 = time.perf_counter()
            list(pool.map(one, pairs * 100))
            cdt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": 100 * len(pairs) / cdt, "unit": "pairs/s", "cores": threads, "kind": "reference", "sample": f"{100 * len(pairs)} pairs, {cdt:.1f} s wall"}
    print(json.dumps(line), flush=True)
    eng.close()


if __name__ == "__main__":
    main()
