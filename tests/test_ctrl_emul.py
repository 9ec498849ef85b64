"""The product's control logic (csrc/me_ctrl.h: the code k_hme / k_epilogue run on the device) executed on the
CPU by tests/emul with scalar stand-ins for the search kernels, against the oracle. Pins every stage, prune,
search-area rule and the candidate construction without a GPU."""
import pytest

from helpers import TF_CASES, Emulator, PictureSet, derive, derive_tf, used_refs
from me_b200 import planes
from checkers import Oracle

CASES = [(preset, poc, kind) for preset in (8, 4, 2, 6, 10, 5) for poc in (16, 8, 4, 1, 7) for kind in ("motion", "iid")]


@pytest.fixture(scope="module")
def tools():
    return Oracle(), Emulator()


@pytest.mark.parametrize("preset,poc,kind", CASES)
def test_control_logic_matches_oracle(tools, preset, poc, kind):
    oracle, emul = tools
    w, h, n = 200, 136, 33
    ps = PictureSet(w, h, n, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    a = oracle.me_picture(ps.get(poc), refs, p)
    b = emul.me_picture(ps.get(poc), refs, p)
    bad = planes.compare_outputs(a, b, used_refs=used_refs(l0, l1))
    assert not bad, bad
    assert a.byte_sad_ops == b.byte_sad_ops


@pytest.mark.parametrize("force", [(1, 0), (1, 1), (2, 2), (4, 3)])
def test_reference_count_variants(tools, force):
    """single reference / one per list / general candidate construction (motion_estimation.c:2532-2835)"""
    oracle, emul = tools
    w, h, n, poc = 200, 136, 33, 7
    ps = PictureSet(w, h, n, "motion", oracle)
    cfg, p, l0, l1 = derive(4, w, h, poc, n, force_refs=force)
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    a = oracle.me_picture(ps.get(poc), refs, p)
    b = emul.me_picture(ps.get(poc), refs, p)
    bad = planes.compare_outputs(a, b, used_refs=used_refs(l0, l1))
    assert not bad, bad


@pytest.mark.parametrize("case", TF_CASES, ids=lambda c: "tf_" + "_".join(str(v) for v in c))
def test_tf_control_logic_matches_oracle(tools, case):
    """ME_MCTF: unscaled distance, no pruning, early exit before the integer search, tf_* results"""
    oracle, emul = tools
    w, h, lvl, kind, c, nb, th = case
    ps = PictureSet(w, h, 20, kind, oracle)
    _, p = derive_tf(lvl, w, h, c, me_exit_th=th)
    refs = [[ps.get(nb)], []]
    a = oracle.me_picture(ps.get(c), refs, p)
    b = emul.me_picture(ps.get(c), refs, p)
    bad = planes.compare_outputs(a, b, fields=planes.TF_FIELDS + planes.DEBUG_FIELDS, used_refs=[1, 0])
    assert not bad, bad
    assert a.tf_blocks == b.tf_blocks and a.byte_sad_ops == b.byte_sad_ops


def _dep_id(c):
    return f"m{c[0]}_poc{c[1]}_{c[2]}_" + "_".join(sorted(c[3]))


@pytest.mark.parametrize("case", __import__("helpers").DEP_CASES, ids=_dep_id)
def test_searches_that_depend_on_reference_00(tools, case):
    """screen-content / low-delay controls: later references read reference (0, 0)'s HME level 0 centre and its finished
    integer search (two-phase stages in me_ctrl.h)"""
    oracle, emul = tools
    preset, poc, kind, kw = case
    w, h, n = 200, 136, 33
    ps = PictureSet(w, h, n, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n, **kw)
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    a = oracle.me_picture(ps.get(poc), refs, p)
    b = emul.me_picture(ps.get(poc), refs, p)
    bad = planes.compare_outputs(a, b, used_refs=used_refs(l0, l1))
    assert not bad, bad
    assert a.byte_sad_ops == b.byte_sad_ops


def test_dependent_cases_reach_both_branches():
    from helpers import DEP_CASES
    from me_b200 import abi  # noqa: F401
    hits = [0, 0]
    for preset, poc, kind, kw in DEP_CASES:
        _, p, l0, l1 = derive(preset, 200, 136, poc, 33, **kw)
        multi = len(l0) + len(l1) > 1
        hits[0] += bool(multi and p.enable_me_sr_adjustment == 2 and not p.me_early_exit_th)
        hits[1] += bool(multi and p.enable_me_sr_adjustment and p.distance_based_hme_resizing and p.reduce_hme_l0_sr_th_min)
    assert hits[0] >= 2 and hits[1] >= 3, hits
