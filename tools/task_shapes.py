"""Task-shape histogram per control stage (CPU emulation of the control logic) for one picture of the bench workload.
Usage: task_shapes.py [poc] [kind]   (tuning aid: shows what k_hme / k_fullsearch have to execute)"""
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))
os.environ["ME_EMUL_DUMP"] = "1"
import bench  # noqa: E402
import helpers  # noqa: E402
from me_b200 import synth  # noqa: E402

poc = int(sys.argv[1]) if len(sys.argv) > 1 else 24
kind = sys.argv[2] if len(sys.argv) > 2 else "motion"
import checkers  # noqa: E402

_, p, l0, l1 = bench.derive(poc)
ps = helpers.PictureSet(bench.W, bench.H, bench.NPIC, kind, checkers.Oracle())
refs = [[ps.get(n) for n in l0], [ps.get(n) for n in l1]]
out = helpers.Emulator().me_picture(ps.get(poc), refs, p)
print("byte-SAD ops", out.byte_sad_ops)
