"""The CPU oracle (oracle/me_oracle*.c) pinned against the reference: golden vectors frozen from the reference's
own code (tests/golden, see make_golden.py) and - in the build container - the reference itself (oracle/_ref)
on the parameter tables of the reference's unit tests (test/SadTest.cc, test/PictureOperatorTest.cc)."""
import ctypes as C
import glob
import os
import re

import numpy as np
import pytest

import json

from helpers import DG_CASES, TF_CASES, DgPictures, PictureSet, derive, derive_tf, used_refs
from me_b200 import abi, planes
from checkers import Oracle, Reference

GOLD = os.path.join(os.path.dirname(__file__), "golden")
have_ref = pytest.mark.skipif(not Reference.available(), reason="oracle/_ref not built (needs /root/reference)")


@pytest.fixture(scope="module")
def oracle():
    return Oracle()


def _golden_cases():
    out = []
    for f in sorted(glob.glob(os.path.join(GOLD, "me_*.npz"))):
        m = re.match(r"me_(\d+)x(\d+)_m(-?\d+)_poc(\d+)_(\w+)\.npz", os.path.basename(f))
        out.append((f, int(m.group(1)), int(m.group(2)), int(m.group(3)), int(m.group(4)), m.group(5)))
    return out


@pytest.mark.parametrize("case", _golden_cases(), ids=lambda c: os.path.basename(c[0])[:-4])
def test_oracle_reproduces_golden_vectors(oracle, case):
    f, w, h, preset, poc, kind = case
    gold = np.load(f)
    ps = PictureSet(w, h, 33, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, 33)
    assert l0 == list(gold["l0"]) and l1 == list(gold["l1"])
    out = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
    mask = np.zeros((2, 4), dtype=bool)
    mask[0, :len(l0)] = True
    mask[1, :len(l1)] = True
    assert len(gold.files) >= 18
    for name in gold.files:
        if name in ("l0", "l1"):
            continue
        a, b = out.arrays[name], gold[name]
        if name in planes.DEBUG_FIELDS:
            a, b = a[:, mask], b[:, mask]
        assert np.array_equal(a, b), name


@have_ref
@pytest.mark.parametrize("size", [(1920, 1080), (960, 540), (176, 144), (88, 72)])  # PictureOperatorTest.cc:232-234
@pytest.mark.parametrize("step", [2, 4, 8])
def test_downsample_2d_matches_reference(oracle, size, step):
    ref = Reference()
    w, h = size
    rng = np.random.default_rng(w * 10 + step)
    for _ in range(3):
        stride = w + 3
        src = rng.integers(0, 256, size=(h, stride), dtype=np.uint8)
        a = np.zeros((h // step + 1, w // step + 8), dtype=np.uint8)
        b = np.zeros_like(a)
        oracle.lib.oracle_downsample_2d(src.ctypes.data, stride, w, h, a.ctypes.data, a.shape[1], step)
        ref.lib.ref_downsample_2d(src.ctypes.data, stride, w, h, b.ctypes.data, b.shape[1], step)
        assert np.array_equal(a, b)


# TEST_LOOP_AREAS (SadTest.cc:433-444) and a spread of its block sizes (SadTest.cc:62-107)
LOOP_AREAS = [(8, 15), (16, 31), (12, 31), (64, 125), (192, 75), (128, 50), (64, 25), (240, 200), (144, 120), (96, 80),
              (48, 40), (240, 120), (192, 96), (144, 72), (24, 12), (288, 120), (192, 60), (128, 100), (48, 50), (80, 80)]
BLOCKS = [(64, 64), (64, 32), (32, 32), (32, 16), (16, 16), (16, 8), (16, 7), (8, 8), (24, 24), (48, 64), (64, 24), (6, 6),
          (4, 16), (12, 5), (128, 64)]
PATTERNS = ["REF_MAX", "SRC_MAX", "RANDOM", "UNALIGN"]


@have_ref
@pytest.mark.parametrize("pattern", PATTERNS)
def test_sad_loop_matches_reference(oracle, pattern):
    """svt_sad_loop_kernel_c semantics incl. tie-breaking and skip_search_line (SadTest.cc:474-539)."""
    ref = Reference()
    rng = np.random.default_rng(hash(pattern) & 0xFFFF)
    for (bw, bh) in BLOCKS:
        for (sw, sh) in LOOP_AREAS[::3]:
            for skip in (0, 1):
                rstride = 127 if pattern == "UNALIGN" and bw + sw <= 127 else bw + sw + 8
                sstride = bw + 5
                src = np.zeros((bh, sstride), dtype=np.uint8)
                refb = np.zeros((bh + sh + 1, rstride), dtype=np.uint8)
                if pattern == "REF_MAX":
                    refb[...] = 255
                elif pattern == "SRC_MAX":
                    src[...] = 255
                else:
                    src[...] = rng.integers(0, 256, size=src.shape, dtype=np.uint8)
                    # few distinct values: many exact ties, so the scan-order tie-break is exercised
                    refb[...] = rng.integers(0, 4, size=refb.shape, dtype=np.uint8) * 60
                res = []
                for lib, fn in ((oracle.lib, "oracle_sad_loop"), (ref.lib, "ref_sad_loop")):
                    best, x, y = abi.u64(0), abi.i16(-7), abi.i16(-9)
                    getattr(lib, fn)(src.ctypes.data, sstride, refb.ctypes.data, rstride, bh, bw, C.byref(best), C.byref(x),
                                     C.byref(y), rstride, skip, sw, sh)
                    res.append((best.value, x.value, y.value))
                assert res[0] == res[1], (bw, bh, sw, sh, skip, res)


@have_ref
@pytest.mark.parametrize("preset,poc,kind", [(8, 16, "motion"), (8, 2, "iid"), (6, 8, "motion"), (4, 1, "motion"), (4, 8, "iid"),
                                            (2, 7, "motion"), (10, 4, "iid"), (5, 2, "motion"), (7, 1, "iid")])
@pytest.mark.parametrize("size", [(320, 192), (264, 200)])
def test_oracle_matches_reference_me(oracle, size, preset, poc, kind):
    """Whole path, every output and tap, against svt_aom_motion_estimation_b64 of the reference."""
    ref = Reference()
    w, h = size
    ps = PictureSet(w, h, 33, kind, ref)
    ps2 = PictureSet(w, h, 33, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, 33)
    for k in [poc] + l0 + l1:  # pyramids: reference downsampler vs restatement
        for lvl in range(3):
            assert np.array_equal(ps.get(k).planes[lvl], ps2.get(k).planes[lvl])
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    a = oracle.me_picture(ps.get(poc), refs, p)
    b = ref.me_picture(ps.get(poc), refs, p)
    fields = [f for f in a.arrays if f != "search_area"]
    bad = planes.compare_outputs(a, b, fields=fields, used_refs=used_refs(l0, l1))
    assert not bad, bad


@have_ref
@pytest.mark.parametrize("case", __import__("helpers").DEP_CASES, ids=lambda c: f"m{c[0]}_poc{c[1]}_{c[2]}_" + "_".join(sorted(c[3])))
def test_oracle_matches_reference_me_dependent_references(oracle, case):
    """screen-content / low-delay controls (enable_me_sr_adjustment == 2, reduce_hme_l0_sr_th_*): the restatement against
    the reference's own svt_aom_motion_estimation_b64, C and AVX2"""
    ref = Reference()
    preset, poc, kind, kw = case
    w, h = 264, 200
    ps = PictureSet(w, h, 33, kind, ref)
    cfg, p, l0, l1 = derive(preset, w, h, poc, 33, **kw)
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    a = oracle.me_picture(ps.get(poc), refs, p)
    b = ref.me_picture(ps.get(poc), refs, p)
    fields = [f for f in a.arrays if f != "search_area"]
    bad = planes.compare_outputs(a, b, fields=fields, used_refs=used_refs(l0, l1))
    assert not bad, bad
    c = ref.me_picture(ps.get(poc), refs, p, simd=1)
    assert not planes.compare_outputs(b, c, fields=planes.PRODUCT_FIELDS)


@have_ref
def test_reference_simd_equals_reference_c():
    """The timed CPU baseline (AVX2 kernels behind the RTCD pointers) gives the same results as the C path."""
    ref = Reference()
    w, h = 320, 192
    ps = PictureSet(w, h, 33, "motion", ref)
    cfg, p, l0, l1 = derive(4, w, h, 4, 33)
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    a = ref.me_picture(ps.get(4), refs, p, simd=0)
    b = ref.me_picture(ps.get(4), refs, p, simd=1)
    bad = planes.compare_outputs(a, b, fields=planes.PRODUCT_FIELDS)
    assert not bad, bad


def _dg_golden():
    return json.load(open(os.path.join(GOLD, "dg_metrics.json")))


@pytest.mark.parametrize("entry", _dg_golden(), ids=lambda e: "dg_" + "_".join(str(v) for v in e["case"]))
def test_dg_detector_reproduces_golden_metrics(oracle, entry):
    """oracle_dg_detector against DGDetectorMetrics frozen from the reference's dg_detector_hme_level0."""
    w, h, res, kind, a, b = entry["case"]
    ps = DgPictures(w, h, kind, oracle)
    assert list(oracle.dg_detector(ps.get(a), ps.get(b), res)) == entry["metrics"]


def test_dg_golden_covers_every_metric():
    g = _dg_golden()
    assert [e["case"] for e in g] == [list(c) for c in DG_CASES]
    for k in range(4):  # every metric is non-zero somewhere, and sum_in_vectors takes both signs
        assert any(e["metrics"][k] != 0 for e in g)
    assert any(e["metrics"][3] < 0 for e in g) and any(e["metrics"][3] > 0 for e in g)


@have_ref
@pytest.mark.parametrize("case", DG_CASES[:13], ids=lambda c: "dg_" + "_".join(str(v) for v in c))
def test_dg_detector_matches_reference(oracle, case):
    """restatement == the reference's own C == the reference's AVX2 sad_loop behind the RTCD pointer"""
    ref = Reference()
    w, h, res, kind, a, b = case
    ps = DgPictures(w, h, kind, ref)
    want = ref.dg_detector(ps.get(a), ps.get(b), res, simd=0)
    assert oracle.dg_detector(ps.get(a), ps.get(b), res) == want
    assert ref.dg_detector(ps.get(a), ps.get(b), res, simd=1) == want


def _tf_name(case):
    w, h, lvl, kind, c, nb, th = case
    return f"tf_{w}x{h}_l{lvl}_{kind}_c{c}_n{nb}_th{th}"


@pytest.mark.parametrize("case", TF_CASES, ids=_tf_name)
def test_oracle_reproduces_tf_golden_vectors(oracle, case):
    """ME_MCTF: the tf_* results frozen from the reference's svt_aom_motion_estimation_b64 with me_type = ME_MCTF."""
    w, h, lvl, kind, c, nb, th = case
    gold = np.load(os.path.join(GOLD, _tf_name(case) + ".npz"))
    ps = PictureSet(w, h, 20, kind, oracle)
    _, p = derive_tf(lvl, w, h, c, me_exit_th=th)
    out = oracle.me_picture(ps.get(c), [[ps.get(nb)], []], p)
    for name in planes.TF_FIELDS:
        assert np.array_equal(out.arrays[name], gold[name]), name
    assert out.tf_blocks == tuple(gold["tf_blocks"])


def test_tf_golden_covers_both_exits():
    ex = [np.load(os.path.join(GOLD, _tf_name(c) + ".npz"))["tf_early_exit"] for c in TF_CASES]
    assert any(e.all() for e in ex) and any(not e.any() for e in ex) and any(e.any() and not e.all() for e in ex)


@have_ref
@pytest.mark.parametrize("case", TF_CASES, ids=_tf_name)
def test_oracle_matches_reference_tf_me(oracle, case):
    ref = Reference()
    w, h, lvl, kind, c, nb, th = case
    ps = PictureSet(w, h, 20, kind, ref)
    _, p = derive_tf(lvl, w, h, c, me_exit_th=th)
    refs = [[ps.get(nb)], []]
    a = oracle.me_picture(ps.get(c), refs, p)
    b = ref.me_picture(ps.get(c), refs, p)
    b1 = ref.me_picture(ps.get(c), refs, p, simd=1)
    assert not planes.compare_outputs(a, b, fields=planes.TF_FIELDS)
    assert not planes.compare_outputs(b, b1, fields=planes.TF_FIELDS)
    assert a.tf_blocks == b.tf_blocks == b1.tf_blocks
    # taps of the one reference, except the SADs of exited blocks (stale MeContext memory in the reference)
    live = a.arrays["tf_early_exit"] == 0
    for name in ("p_sb_best_mv", "hme_search_centre", "hme_sad"):
        assert np.array_equal(a.arrays[name][:, 0, 0], b.arrays[name][:, 0, 0]), name
    assert np.array_equal(a.arrays["p_sb_best_sad"][live, 0, 0], b.arrays["p_sb_best_sad"][live, 0, 0])
