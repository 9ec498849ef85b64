"""me_gpu_derive_params (host C, product) against the reference's svt_aom_sig_deriv_me: frozen golden table
(tests/golden/derive_params.json, generated from the reference) and, in the build container, the reference
itself (oracle/_ref)."""
import ctypes as C
import json
import os

import pytest

from helpers import host_lib
from me_b200 import abi
from checkers import Reference

SKIP = {"max_cand", "max_refs", "max_l0", "num_of_list_to_search", "num_of_ref_pic_to_search", "only_l_bwd"}
GOLD = os.path.join(os.path.dirname(__file__), "golden", "derive_params.json")


def _derive(preset, w, h, tl):
    cfg = abi.MeGpuEncoderConfig(enc_mode=preset, qp=35, hierarchical_levels=4, width=w, height=h, picture_number=8,
                                 temporal_layer_index=tl, is_ref=1 if tl < 4 else 0)
    p = abi.MeGpuParams()
    assert host_lib().me_gpu_derive_params(C.byref(cfg), C.byref(p)) == 0
    return cfg, p


def test_against_golden_table():
    table = json.load(open(GOLD))
    assert len(table) == 180
    for row in table:
        _, p = _derive(row["preset"], row["width"], row["height"], row["tl"])
        got = p.as_dict()
        diff = {k: (got[k], v) for k, v in row["params"].items() if k not in SKIP and got[k] != v}
        assert not diff, (row["preset"], row["width"], row["height"], row["tl"], diff)


@pytest.mark.skipif(not Reference.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("qp", [20, 35, 55])
def test_against_reference_qp_sweep(qp):
    ref = Reference()
    for preset in (-1, 0, 2, 4, 5, 6, 8, 10):
        for (w, h) in ((640, 360), (1920, 1080), (3840, 2160)):
            for tl in (0, 3):
                cfg = abi.MeGpuEncoderConfig(enc_mode=preset, qp=qp, hierarchical_levels=4, width=w, height=h,
                                             picture_number=8, temporal_layer_index=tl, is_ref=1)
                p = abi.MeGpuParams()
                assert host_lib().me_gpu_derive_params(C.byref(cfg), C.byref(p)) == 0
                n, z = abi.u8(), abi.u32()
                host_lib().me_gpu_derive_mrp_inputs(C.byref(cfg), C.byref(n), C.byref(z))
                exp = ref.derive_params(cfg, n.value, z.value).as_dict()
                got = p.as_dict()
                diff = {k: (got[k], exp[k]) for k in exp if k not in SKIP and got[k] != exp[k]}
                assert not diff, (preset, w, h, tl, diff)


def test_reference_counts_follow_mrp_levels():
    """enc_handle.c:3397-3560,4193-4230: M0-M2 level 2 (4/3), M3-M4 level 5 (4/3), M5-M7 level 9 (3/2),
    M8-M9 level 10 (base 3/2, non-base 2/2), M10+ level 0 (1/1); max_cand per pcs.c:91-96."""
    expect = {0: (4, 3, 4, 3), 2: (4, 3, 4, 3), 4: (4, 3, 4, 3), 6: (3, 2, 3, 2), 8: (3, 2, 2, 2), 10: (1, 1, 1, 1)}
    for preset, (b0, b1, n0, n1) in expect.items():
        _, pb = _derive(preset, 1920, 1080, 0)
        _, pn = _derive(preset, 1920, 1080, 3)
        assert tuple(pb.num_of_ref_pic_to_search) == (b0, b1), preset
        assert tuple(pn.num_of_ref_pic_to_search) == (n0, n1), preset
        l0, l1 = max(b0, n0), max(b1, n1)
        assert pb.max_refs == l0 + l1 and pb.max_cand == l0 + l1 + l0 * l1 + (l0 - 1) + (1 if l1 == 3 else 0)


def _derive_tf(lvl, w, h, qp_opt, qp):
    cfg = abi.MeGpuTfConfig(hme_me_level=lvl, qp_opt=qp_opt, qp=qp, is_ref=1, me_exit_th=777, width=w, height=h, picture_number=8)
    p = abi.MeGpuParams()
    assert host_lib().me_gpu_derive_params_tf(C.byref(cfg), C.byref(p)) == 0
    return cfg, p


def test_tf_against_golden_table():
    """me_gpu_derive_params_tf against svt_aom_sig_deriv_me_tf outputs frozen from the reference"""
    table = json.load(open(os.path.join(os.path.dirname(GOLD), "derive_params_tf.json")))
    assert len(table) == 80
    for row in table:
        _, p = _derive_tf(*row["case"])
        assert p.as_dict() == row["params"], row["case"]


@pytest.mark.skipif(not Reference.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_tf_against_reference():
    ref = Reference()
    for lvl in range(5):
        for (w, h) in ((320, 192), (640, 360), (856, 480), (1280, 720), (3840, 2160)):
            for (qp_opt, qp) in ((0, 35), (1, 10), (1, 45), (1, 63)):
                cfg, p = _derive_tf(lvl, w, h, qp_opt, qp)
                assert p.as_dict() == ref.derive_params_tf(cfg).as_dict(), (lvl, w, h, qp_opt, qp)


def test_tf_bad_arguments():
    p = abi.MeGpuParams()
    cfg = abi.MeGpuTfConfig(hme_me_level=5, width=320, height=192)
    assert host_lib().me_gpu_derive_params_tf(C.byref(cfg), C.byref(p)) == -2
    cfg = abi.MeGpuTfConfig(hme_me_level=1, width=322, height=192)
    assert host_lib().me_gpu_derive_params_tf(C.byref(cfg), C.byref(p)) == -2
