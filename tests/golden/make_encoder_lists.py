"""Freezes the reference lists of the ENCODER's own prediction structure (tests/golden/encoder_lists.json): the patched
SvtAv1EncApp (integration/build_encoder.sh) encodes a 49-picture clip per preset with SVT_ME_GPU_LOG_LISTS=1 and the hook
prints, per picture, the lists the ME stage received. Needs the GPU box (the hook is only live with the engine):

    gpurun -- 'python tests/golden/make_encoder_lists.py'      ->  gpurun_out/encoder_lists.json, then copy the "lists" over

tests/workloads.py takes the lists of the bench workloads from that file (presets 0, 2, 4, 8; steady-state mini-GOPs)."""
import os, re, sys, json
REPO = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))  # tests/golden -> repo
sys.path.insert(0, os.path.join(REPO, "tests")); sys.path.insert(0, os.path.join(REPO, "svt-av1-psyex_b200", "python"))
import encoder_cases as ec
from me_b200 import abi, synth
out = {}
for preset in (0, 2, 4, 8):
    name = f"lists_m{preset}"
    ec.CASES[name] = (320, 192, 49, preset, ["--lp", "1", "--enable-tf", "0", "--hierarchical-levels", "4"])
    _, log = ec.encode(ec.GPU, name, "/tmp", env={"SVT_ME_GPU_LIB": abi.GPU_LIB, "SVT_ME_GPU_LOG_LISTS": "1"})
    seen = {}
    for m in re.finditer(r"lists poc (\d+) tl (\d+):(.*)", log):
        lists = [[int(x) for x in part.split()[1:]] for part in re.findall(r"L\d[ \d]*", m.group(3))]
        seen[int(m.group(1))] = [int(m.group(2))] + lists
    out[preset] = seen
    for poc in sorted(seen): print(preset, poc, seen[poc])
json.dump(out, open(os.path.join(REPO, "gpurun_out", "encoder_lists.json"), "w"))
