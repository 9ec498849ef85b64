"""Where a warp of k_hme spends its cycles: per-phase clock64 sums of the instrumented build
(make -C svt-av1-psyex_b200 trace -> libme_gpu_trace.so), averaged per warp. Experiments only.

    ME_B200_TRACE_LIB=svt-av1-psyex_b200/libme_gpu_trace.so python tools/trace_hme.py [workload] [kind]
"""
import ctypes as C
import os
import sys

REPO = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
os.environ.setdefault("ME_B200_TRACE_LIB", os.path.join(REPO, "svt-av1-psyex_b200", "libme_gpu_trace.so"))
sys.path.insert(0, os.path.join(REPO, "svt-av1-psyex_b200", "python"))
sys.path.insert(0, os.path.join(REPO, "tests"))

import workloads  # noqa: E402
from me_b200 import planes, synth  # noqa: E402
from me_b200.engine import Engine  # noqa: E402

NAMES = ["setup", "zz", "prehme", "cross(barriers)", "L0", "L1", "L2", "centre", "probe", "main", "end-barrier", "  plan", "  decode",
         "  tma-wait", "  search", "  generic", "  ctrl(lane0)", "jobs", "warps", "total"]


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "1080p-m4"
    kind = sys.argv[2] if len(sys.argv) > 2 else "motion"
    wl = workloads.WORKLOADS[name]
    w, h = wl["w"], wl["h"]
    seq = synth.Sequence(w, h, workloads.NPIC, kind=kind)
    eng = Engine(w, h, 4, 3, pictures_in_flight=workloads.NPIC + 2)
    for k in range(workloads.NPIC):
        eng.register_picture_unpadded(k, seq.luma(k))
    buf = (C.c_ulonglong * 64)()
    eng.lib.me_hme_trace_read(buf, 1)
    flush = None
    if os.environ.get("TRACE_COLD"):  # the bench's condition: nothing of the picture in L2 when its kernel starts
        import torch
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for poc in workloads.STEP_POCS:
        _, p, l0, l1 = workloads.derive_for(name, poc)
        if flush is not None:
            flush.zero_()
            torch.cuda.synchronize()
        eng.me_picture(poc, w, h, p, [l0, l1], debug=False)
    n = eng.lib.me_hme_trace_read(buf, 1)
    v = list(buf)[:n]
    warps = max(1, v[18])
    print(f"{name} {kind}: {warps} warps over 16 pictures; cycles per warp (lane 0's clock64)")
    for i, nm in enumerate(NAMES):
        if nm in ("jobs", "warps"):
            print(f"  {nm:18s} {v[i] / (warps if nm == 'jobs' else 1):10.1f}")
        else:
            print(f"  {nm:18s} {v[i] / warps:10.0f}   {100.0 * v[i] / max(1, v[19]):5.1f} %")
    eng.close()


if __name__ == "__main__":
    main()
