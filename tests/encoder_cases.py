"""In-encoder cases: synthetic clips encoded by SvtAv1EncApp, stock vs. with the B200 ME binding.
Shared by tests/golden/make_encoder_md5.py (stock encoder, build container) and tests/test_gpu_encoder.py (GPU box)."""
import hashlib
import os
import subprocess

import numpy as np

from me_b200 import synth

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
BUILD = os.path.join(REPO, "integration", "_build")
STOCK, GPU = os.path.join(BUILD, "SvtAv1EncApp_stock"), os.path.join(BUILD, "SvtAv1EncApp_gpu")
MD5_FILE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "encoder_md5.json")

# name -> (width, height, frames, preset, extra arguments). Fast presets run with --lp 1: the reference's own multi-threaded
# output at presets >= 5 differs from run to run (two stock runs, same box, same arguments: different md5), so only the
# single-threaded configuration has a bitstream to be equal to. The slower presets are stable with threads.
# "scene" clips switch to an unrelated sequence half way (a scene change: an extra key frame) and run with a short
# intra period and overlay pictures, so that I slices, overlays and short mini-GOPs go through the ME stage.
CASES = {
    "320x192_m8": (320, 192, 17, 8, ["--lp", "1"]),
    "320x192_m6": (320, 192, 17, 6, ["--lp", "1"]),
    "320x192_m4": (320, 192, 17, 4, ["--lp", "4"]),
    "320x192_m2": (320, 192, 17, 2, ["--lp", "4"]),
    "320x192_m0": (320, 192, 9, 0, ["--lp", "4"]),
    "640x360_m8_scene": (640, 360, 25, 8, ["--lp", "1", "--keyint", "12", "--enable-overlays", "1"]),
    "640x360_m4_scene": (640, 360, 17, 4, ["--lp", "4", "--keyint", "8", "--enable-overlays", "1"]),
    # 65 pictures in mini-GOPs of 32 (six hierarchical layers): the dynamic-GOP detector (early_hme) decides 6L vs 5L
    "320x192_m6_6L": (320, 192, 65, 6, ["--lp", "1", "--hierarchical-levels", "5"]),
}


def write_clip(path, w, h, n, scene):
    a = synth.Sequence(w, h, n, kind="motion")
    b = synth.Sequence(w, h, n, kind="motion", seed=4321) if scene else a
    with open(path, "wb") as f:
        for k in range(n):
            f.write((b if scene and k >= n // 2 else a).luma(k).tobytes())
            f.write(bytes([128]) * (w * h // 2))


def encode(binary, name, workdir, env=None):
    """-> (md5 of the .ivf, the encoder's log)"""
    w, h, n, preset, extra = CASES[name]
    clip = os.path.join(workdir, f"{name}.yuv")
    if not os.path.exists(clip):
        write_clip(clip, w, h, n, "scene" in name)
    out = os.path.join(workdir, f"{name}_{os.path.basename(binary)}.ivf")
    cmd = [binary, "-i", clip, "-w", str(w), "-h", str(h), "--fps", "30", "-n", str(n), "--preset", str(preset), "--scm", "0",
           "-b", out] + extra
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run(cmd, capture_output=True, text=True, env=e, timeout=1500)
    log = r.stdout + r.stderr
    if r.returncode != 0 or not os.path.exists(out):
        raise RuntimeError(f"{' '.join(cmd)} failed ({r.returncode}):\n{log[-2000:]}")
    return hashlib.md5(open(out, "rb").read()).hexdigest(), log
