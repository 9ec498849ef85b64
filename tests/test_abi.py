"""C-ABI checks that need no GPU: the ctypes mirror matches the C headers, the libraries load and export
every symbol include/me_gpu.h declares (no compute calls)."""
import ctypes as C
import os
import re
import subprocess
import tempfile

import pytest

from me_b200 import abi

REPO = abi.REPO


def _probe(structs):
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "me_gpu.h"', "int main(void){"]
    for name, cls in structs:
        src.append(f'printf("{name} %zu\\n", sizeof({name}));')
        for f, _ in cls._fields_:
            src.append(f'printf("{name}.{f} %zu\\n", offsetof({name}, {f}));')
    src.append("return 0;}")
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "p.c")
        open(c, "w").write("\n".join(src))
        exe = os.path.join(d, "p")
        subprocess.run(["gcc", "-I", os.path.join(REPO, "include"), c, "-o", exe], check=True)
        out = subprocess.run([exe], check=True, capture_output=True, text=True).stdout
    return dict(line.split() for line in out.strip().splitlines())


def test_ctypes_layout_matches_header():
    structs = [("MeGpuParams", abi.MeGpuParams), ("MeGpuEncoderConfig", abi.MeGpuEncoderConfig),
               ("MeGpuOutputs", abi.MeGpuOutputs), ("MeGpuTimings", abi.MeGpuTimings), ("MeGpuDgMetrics", abi.MeGpuDgMetrics)]
    got = _probe(structs)
    for name, cls in structs:
        assert int(got[name]) == C.sizeof(cls), name
        for f, _ in cls._fields_:
            assert int(got[f"{name}.{f}"]) == getattr(cls, f).offset, f"{name}.{f}"


def test_header_symbols_are_listed_and_exported():
    hdr = open(os.path.join(REPO, "include", "me_gpu.h")).read()
    declared = set(re.findall(r"\b(me_gpu_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(abi.GPU_SYMBOLS), declared ^ set(abi.GPU_SYMBOLS)
    if not os.path.exists(abi.GPU_LIB):
        pytest.skip("libme_gpu.so not built (no nvcc)")
    lib = abi.load_gpu_lib()  # loading needs libcudart, not a GPU
    for sym in abi.GPU_SYMBOLS:
        assert hasattr(lib, sym), sym


def test_engine_fails_loudly_without_a_device():
    """No CPU fallback: on a box without a GPU me_gpu_init must fail with ME_GPU_ERR_NO_DEVICE."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    if not os.path.exists(abi.GPU_LIB):
        pytest.skip("libme_gpu.so not built")
    lib = abi.load_gpu_lib()
    rc = lib.me_gpu_init(0, None, 320, 192, 2, 2, 4)
    assert rc == -1, rc
    assert b"no CUDA device" in lib.me_gpu_last_error()


def test_product_does_not_reference_the_oracle():
    """The product sources never include, link or load anything under oracle/ or tests/."""
    pkg = os.path.join(REPO, "svt-av1-psyex_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".cu", ".h", ".c", ".cpp")) or f in ("Makefile", "engine.py", "abi.py", "planes.py", "synth.py"):
                txt = open(os.path.join(root, f), errors="ignore").read()
                assert "me_oracle" not in txt and "libme_ref" not in txt and "me_emul" not in txt and "import checkers" not in txt, os.path.join(root, f)


def test_reference_side_hook_builds_and_fails_loudly_without_the_engine():
    """integration/svt_me_gpu_hook.c compiles against the reference's headers (oracle/Makefile) and, with no engine
    library to dlopen, reports the failure instead of falling back to anything."""
    from helpers import PictureSet, derive
    from checkers import Reference
    if not Reference.available():
        pytest.skip("needs oracle/_ref/libme_ref.so (built where /root/reference exists)")
    ref = Reference()
    assert ref.has_hook()
    w, h, n, poc = 64, 64, 9, 4
    ps = PictureSet(w, h, n, "iid", ref)
    cfg, p, l0, l1 = derive(4, w, h, poc, n)
    with pytest.raises(RuntimeError, match="-10"):
        ref.hook_me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p, "/nonexistent/libme_gpu.so")
