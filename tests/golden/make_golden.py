"""Freezes golden vectors from the REFERENCE ITSELF (oracle/_ref/libme_ref.so = the reference's own ME sources
compiled in place from /root/reference by oracle/Makefile). Run in the build container only:

    python tests/golden/make_golden.py

Writes tests/golden/me_<case>.npz (every MeGpuOutputs field the reference produces) and
tests/golden/derive_params.json (svt_aom_sig_deriv_me outputs) and tests/golden/dg_metrics.json (dg_detector_hme_level0). Inputs are regenerated from the seeded
synthetic sequence (me_b200/synth.py), so only outputs are stored.
"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "svt-av1-psyex_b200", "python"))
sys.path.insert(0, os.path.join(HERE, ".."))

from helpers import DG_CASES, TF_CASES, DgPictures, PictureSet, derive, derive_tf, host_lib  # noqa: E402
from me_b200 import planes  # noqa: E402
from me_b200 import abi  # noqa: E402
from checkers import Reference  # noqa: E402

ME_CASES = [  # (width, height, preset, poc, kind)
    (320, 192, 8, 8, "motion"), (320, 192, 8, 1, "iid"), (320, 192, 4, 4, "motion"), (320, 192, 4, 7, "iid"),
    (320, 192, 2, 2, "motion"), (320, 192, 0, 1, "motion"), (200, 136, 6, 16, "motion"), (200, 136, 10, 7, "iid"),
    (200, 136, -1, 1, "motion"),
]
DERIVE_CASES = [(p, w, h, tl) for p in (-3, -2, -1, 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11)
                for (w, h) in ((320, 192), (1280, 720), (1920, 1080), (3840, 2160)) for tl in (0, 2, 4)]


def case_name(w, h, preset, poc, kind):
    return f"me_{w}x{h}_m{preset}_poc{poc}_{kind}"


TF_DERIVE_CASES = [(lvl, w, h, qp_opt, qp) for lvl in range(5) for (w, h) in ((320, 192), (640, 360), (856, 480), (1920, 1080))
                   for (qp_opt, qp) in ((0, 35), (1, 20), (1, 35), (1, 60))]


def make_tf_golden(ref):
    for (w, h, lvl, kind, c, nb, th) in TF_CASES:
        ps = PictureSet(w, h, 20, kind, ref)
        _, p = derive_tf(lvl, w, h, c, me_exit_th=th)
        out = ref.me_picture(ps.get(c), [[ps.get(nb)], []], p)
        name = f"tf_{w}x{h}_l{lvl}_{kind}_c{c}_n{nb}_th{th}"
        np.savez_compressed(os.path.join(HERE, name + ".npz"), tf_blocks=np.array(out.tf_blocks),
                            **{k: out.arrays[k] for k in planes.TF_FIELDS})
        print("wrote", name)
    table = []
    for (lvl, w, h, qp_opt, qp) in TF_DERIVE_CASES:
        cfg = abi.MeGpuTfConfig(hme_me_level=lvl, qp_opt=qp_opt, qp=qp, is_ref=1, me_exit_th=777, width=w, height=h, picture_number=8)
        table.append({"case": [lvl, w, h, qp_opt, qp], "params": ref.derive_params_tf(cfg).as_dict()})
    json.dump(table, open(os.path.join(HERE, "derive_params_tf.json"), "w"))
    print("wrote derive_params_tf.json:", len(table), "cases")


def main():
    ref = Reference()
    for (w, h, preset, poc, kind) in ME_CASES:
        n = 33
        ps = PictureSet(w, h, n, kind, ref)
        cfg, p, l0, l1 = derive(preset, w, h, poc, n)
        out = ref.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
        arrays = {k: v for k, v in out.arrays.items() if k != "search_area"}  # the reference has no such state
        np.savez_compressed(os.path.join(HERE, case_name(w, h, preset, poc, kind) + ".npz"), l0=np.array(l0), l1=np.array(l1),
                            **arrays)
        print("wrote", case_name(w, h, preset, poc, kind))
    table = []
    for (preset, w, h, tl) in DERIVE_CASES:
        cfg = abi.MeGpuEncoderConfig(enc_mode=preset, qp=35, hierarchical_levels=4, width=w, height=h, picture_number=8,
                                     temporal_layer_index=tl, is_ref=1 if tl < 4 else 0)
        n, z = abi.u8(), abi.u32()
        host_lib().me_gpu_derive_mrp_inputs(C.byref(cfg), C.byref(n), C.byref(z))
        table.append({"preset": preset, "width": w, "height": h, "tl": tl, "params": ref.derive_params(cfg, n.value, z.value).as_dict()})
    json.dump(table, open(os.path.join(HERE, "derive_params.json"), "w"))
    print("wrote derive_params.json:", len(table), "cases")
    # temporal-filter ME (ME_MCTF): tf_* outputs of svt_aom_motion_estimation_b64 + derived controls
    make_tf_golden(ref)
    # dynamic-GOP detector metrics of dg_detector_hme_level0 (pd_process.c:492)
    dg = []
    for (w, h, res, kind, a, b) in DG_CASES:
        ps = DgPictures(w, h, kind, ref)
        dg.append({"case": [w, h, res, kind, a, b], "metrics": list(ref.dg_detector(ps.get(a), ps.get(b), res))})
    json.dump(dg, open(os.path.join(HERE, "dg_metrics.json"), "w"))
    print("wrote dg_metrics.json:", len(dg), "cases")


if __name__ == "__main__":
    main()
