"""Whole-picture GPU parity on the BASELINE.json configurations themselves, and the leaf-kernel stress patterns of the
reference's SadTest.cc (saturating, all-tie, unaligned stride) driven through the whole path at large search areas.

* every picture of the timed step (POC 17..32) of every bench workload, "motion" and "iid", all b64s, every product
  array: digest == tests/golden/picture_hashes.json, frozen from the REFERENCE's AVX2 path (make_picture_hashes.py,
  which also cross-checks reference AVX2 == reference C == oracle on 30 b64s per picture);
* where oracle/_ref/libme_ref.so travelled with the snapshot, a mismatch is narrowed down to (field, b64) by running the
  reference on the spot; the pattern tests compare with the reference (whole picture) or, without it, with the oracle
  on a sample of b64s.

These engines are sized for one workload each (the C-ABI holds one engine per process at a time)."""
import numpy as np
import pytest

import workloads
from helpers import PictureSet, derive, derive_tf
from me_b200 import planes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle():
    from checkers import Oracle
    return Oracle()


def _reference():
    from checkers import Reference
    return Reference() if Reference.available() else None


def _explain(name, kind, poc, ps, p, l0, l1, got):
    ref = _reference()
    if ref is None:
        return "digest mismatch (no libme_ref.so on this box to narrow it down)"
    for k in set([poc] + l0 + l1):
        ref.build_pyramid(ps.get(k))
    exp = ref.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p, simd=1, debug=False)
    return planes.compare_outputs(exp, got, fields=planes.PRODUCT_FIELDS)


@pytest.mark.parametrize("kind", ["motion", "iid"])
@pytest.mark.parametrize("name", list(workloads.WORKLOADS))
def test_whole_pictures_of_the_timed_step(oracle, name, kind):
    from me_b200.engine import Engine
    gold = workloads.load_hashes()
    wl = workloads.WORKLOADS[name]
    w, h = wl["w"], wl["h"]
    missing = [poc for poc in workloads.STEP_POCS if workloads.hash_key(name, kind, poc) not in gold]
    assert not missing, f"no golden digests for {name}/{kind} pictures {missing}: run tests/golden/make_picture_hashes.py"
    ps = PictureSet(w, h, workloads.NPIC, kind, oracle)
    eng = Engine(w, h, 4, 3, pictures_in_flight=workloads.NPIC + 2)
    bad = {}
    try:
        for k in range(workloads.NPIC):
            eng.register_picture_unpadded(k, ps.seq.luma(k))
        jobs = {poc: workloads.derive_for(name, poc) for poc in workloads.STEP_POCS}
        outs = {poc: planes.Outputs(w, h, jobs[poc][1], debug=False) for poc in workloads.STEP_POCS}
        window = []
        for poc in workloads.STEP_POCS:  # streamed like the bench: several pictures in flight
            if len(window) == 6:
                q = window.pop(0)
                eng.wait(q, outs[q])
            _, p, l0, l1 = jobs[poc]
            eng.submit_out(poc, p, [l0, l1], outs[poc])
            window.append(poc)
        for q in window:
            eng.wait(q, outs[q])
    finally:
        eng.close()
    for poc in workloads.STEP_POCS:
        if workloads.picture_hashes(outs[poc]) != gold[workloads.hash_key(name, kind, poc)]:
            _, p, l0, l1 = jobs[poc]
            bad[poc] = _explain(name, kind, poc, ps, p, l0, l1, outs[poc])
    assert not bad, bad


# --------------------------------------------------------------------------------------------------
# SadTest.cc patterns through the whole path (test/SadTest.cc:163-247: REF_MAX, SRC_MAX, UNALIGN; :525-539 tie-breaks)
# --------------------------------------------------------------------------------------------------
class PatternPictures(PictureSet):
    """pictures of a pattern: "sat" = the searched picture all 255, every other all 0 (every SAD at its maximum:
    the packed keys at their tightest); "tie" = every picture the same constant (every position of every search ties:
    the first one in scan order must win everywhere); "half" = left half 255 / right half 0 in the source, the opposite
    in the references (saturated AND position dependent)."""

    def __init__(self, width, height, pattern, builder, src_poc):
        super().__init__(width, height, 33, "iid", builder)
        self.pattern, self.src_poc = pattern, src_poc

    def luma(self, poc):
        h, w = self.h, self.w
        if self.pattern == "sat":
            return np.full((h, w), 255 if poc == self.src_poc else 0, dtype=np.uint8)
        if self.pattern == "tie":
            return np.full((h, w), 117, dtype=np.uint8)
        a = np.zeros((h, w), dtype=np.uint8)
        if poc == self.src_poc:
            a[:, :w // 2] = 255
        else:
            a[:, w // 2:] = 255
        return a

    def get(self, poc):
        if poc not in self.pics:
            p = planes.Picture(self.w, self.h, poc)
            p.set_luma(self.luma(poc))
            self.builder.build_pyramid(p)
            self.pics[poc] = p
        return self.pics[poc]


def _compare_with_checker(oracle, ps, poc, l0, l1, p, got, fields, sample):
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    ref = _reference()
    if ref is not None:  # the reference's AVX2 path over the whole picture
        exp = ref.me_picture(ps.get(poc), refs, p, simd=1, debug=False)
        return planes.compare_outputs(exp, got, fields=fields)
    bad = []
    for first in sample:  # the plain-C oracle on a few b64s
        exp = oracle.me_picture(ps.get(poc), refs, p, first, 1, debug=False)
        for f in fields:
            if not np.array_equal(exp.arrays[f][first], got.arrays[f][first]):
                bad.append((f, first))
    return bad


@pytest.mark.parametrize("pattern", ["sat", "tie", "half"])
@pytest.mark.parametrize("w,h,preset,poc", [(1920, 1080, 4, 32), (1920, 1080, 8, 24), (3840, 2160, 0, 32), (3840, 2160, 2, 20)])
def test_patterns_open_loop(oracle, w, h, preset, poc, pattern):
    """base-layer / layer-1 pictures: every search area at sa_max (4K preset 0: HME-L0 quadrants of 240x240, integer
    search 256x256: the k_fullsearch tile keys and the hm_job_search8 keys at saturation and with every position tied)"""
    from me_b200.engine import Engine
    cfg, p, l0, l1 = derive(preset, w, h, poc, 33)
    ps = PatternPictures(w, h, pattern, oracle, poc)
    pocs = sorted(set([poc] + l0 + l1))
    eng = Engine(w, h, 4, 3, pictures_in_flight=len(pocs) + 2)
    try:
        for i, k in enumerate(pocs):
            if i & 1:  # UNALIGN: a source stride that is not a multiple of 4 (SadTest.cc:242-247 uses 127)
                lum = np.zeros((h, w + 3), dtype=np.uint8)
                lum[:, :w] = ps.luma(k)
                eng.register_picture_unpadded(k, lum[:, :w])
            else:
                eng.register_picture(k, ps.get(k).planes[0], w, h)
        got = eng.me_picture(poc, w, h, p, [l0, l1], debug=False)
    finally:
        eng.close()
    nb = got.b64_count
    bad = _compare_with_checker(oracle, ps, poc, l0, l1, p, got, planes.PRODUCT_FIELDS, [0, nb // 2 + 7, nb - 1] if w > 1920 else
                                list(range(0, nb, 37)) + [nb - 1])
    assert not bad, bad


@pytest.mark.parametrize("pattern", ["sat", "tie", "half"])
@pytest.mark.parametrize("lvl", [0, 2])
def test_patterns_temporal_filter(oracle, lvl, pattern):
    """ME_MCTF level 0: full-SAD search (no row sub-sampling), 120x120 integer areas, HME with full 16x16 / 32x32 / 64x64
    SADs: a saturated 16x16 block is 65 280, the largest value the 16-bit in-job key field ever sees"""
    from me_b200.engine import Engine
    w, h, c, nb_poc = 1920, 1080, 8, 6
    _, p = derive_tf(lvl, w, h, c, me_exit_th=0)
    ps = PatternPictures(w, h, pattern, oracle, c)
    eng = Engine(w, h, 4, 3, pictures_in_flight=4)
    try:
        eng.register_picture(c, ps.get(c).planes[0], w, h)
        eng.register_picture(nb_poc, ps.get(nb_poc).planes[0], w, h)
        got = eng.me_picture(c, w, h, p, [[nb_poc], []], debug=False)
    finally:
        eng.close()
    bad = _compare_with_checker(oracle, ps, c, [nb_poc], [], p, got, planes.TF_FIELDS, [0, 29, 255, 480, 509])
    assert not bad, bad
