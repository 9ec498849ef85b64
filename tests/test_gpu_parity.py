"""GPU parity: the CUDA engine (through the C-ABI) against the CPU oracle on the same seeded inputs.
Bit-exact: every candidate, MV, SAD, distortion statistic and parity tap."""
import numpy as np
import pytest

from helpers import PictureSet, derive, used_refs
from me_b200 import planes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle():
    from checkers import Oracle
    return Oracle()


@pytest.fixture(scope="module")
def engine():
    from me_b200.engine import Engine
    eng = Engine(1920, 1088, 4, 3, pictures_in_flight=24)
    yield eng
    eng.close()


def _register(engine, ps, pocs):
    for k in pocs:
        engine.register_picture(k, ps.get(k).planes[0], ps.w, ps.h)


def _release(engine, pocs):
    for k in pocs:
        engine.release_picture(k)


@pytest.mark.parametrize("size", [(320, 192), (200, 136), (64, 64), (1920, 1080)])
def test_pyramid_matches_reference_downsampler(engine, oracle, size):
    w, h = size
    ps = PictureSet(w, h, 2, "motion" if w >= 200 else "iid", oracle)
    pic = ps.get(1)
    engine.register_picture(1, pic.planes[0], w, h)
    got = engine.read_pyramid(1, w, h)
    engine.release_picture(1)
    for lvl in range(3):
        assert np.array_equal(got[lvl], pic.planes[lvl]), f"level {lvl}"
    # unpadded registration: the 68-pixel border is generated on the device
    engine.register_picture_unpadded(2, np.ascontiguousarray(pic.interior(0)))
    got = engine.read_pyramid(2, w, h)
    engine.release_picture(2)
    for lvl in range(3):
        assert np.array_equal(got[lvl], pic.planes[lvl]), f"unpadded level {lvl}"


CASES = [(preset, poc, kind) for preset in (8, 4, 2, 6, 10, 0, -1) for poc in (16, 8, 4, 1, 7) for kind in ("motion", "iid")
         if not (preset <= 0 and kind == "iid" and poc != 1)]


@pytest.mark.parametrize("size", [(320, 192), (200, 136)])
@pytest.mark.parametrize("preset,poc,kind", CASES)
def test_me_parity_small(engine, oracle, size, preset, poc, kind):
    w, h = size
    n = 33
    ps = PictureSet(w, h, n, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    pocs = sorted(set([poc] + l0 + l1))
    _register(engine, ps, pocs)
    try:
        got = engine.me_picture(poc, w, h, p, [l0, l1])
    finally:
        _release(engine, pocs)
    exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
    bad = planes.compare_outputs(exp, got, used_refs=used_refs(l0, l1))
    assert not bad, bad
    assert exp.byte_sad_ops == got.byte_sad_ops


@pytest.mark.parametrize("size", [(320, 192), (200, 136)])
@pytest.mark.parametrize("case", __import__("helpers").DEP_CASES, ids=lambda c: f"m{c[0]}_poc{c[1]}_{c[2]}_" + "_".join(sorted(c[3])))
def test_me_parity_dependent_references(engine, oracle, size, case):
    """screen-content / low-delay controls: references that read reference (0, 0)'s HME level 0 centre (an extra stage inside
    k_hme) and its finished integer search (two k_hme + k_fullsearch passes per picture)"""
    preset, poc, kind, kw = case
    w, h = size
    n = 33
    ps = PictureSet(w, h, n, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n, **kw)
    pocs = sorted(set([poc] + l0 + l1))
    _register(engine, ps, pocs)
    try:
        got = engine.me_picture(poc, w, h, p, [l0, l1])
    finally:
        _release(engine, pocs)
    exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
    bad = planes.compare_outputs(exp, got, used_refs=used_refs(l0, l1))
    assert not bad, bad
    assert exp.byte_sad_ops == got.byte_sad_ops


@pytest.mark.parametrize("preset,poc,kind,first,count", [(8, 4, "motion", 0, 510), (4, 8, "motion", 200, 40),
                                                         (4, 1, "iid", 490, 20), (2, 2, "motion", 29, 3)])
def test_me_parity_1080p_sample(engine, oracle, preset, poc, kind, first, count):
    """Full-size pictures: the GPU does the whole picture, the oracle a sample of b64s (seconds of CPU)."""
    w, h, n = 1920, 1080, 17
    ps = PictureSet(w, h, n, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    pocs = sorted(set([poc] + l0 + l1))
    _register(engine, ps, pocs)
    try:
        got = engine.me_picture(poc, w, h, p, [l0, l1])
    finally:
        _release(engine, pocs)
    exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p, first, count)
    for name in exp.arrays:
        a, b = exp.arrays[name][first:first + count], got.arrays[name][first:first + count]
        if name in planes.DEBUG_FIELDS:
            mask = np.zeros((2, 4), dtype=bool)
            mask[0, :len(l0)] = True
            mask[1, :len(l1)] = True
            a, b = a[:, mask], b[:, mask]
        assert np.array_equal(a, b), name


def test_me_parity_odd_size(engine, oracle):
    """Width and height that are not multiples of 64 (partial b64s on both edges: the per-warp search path)."""
    w, h, n, preset, poc = 136, 72, 9, 4, 4
    ps = PictureSet(w, h, n, "motion", oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    pocs = sorted(set([poc] + l0 + l1))
    _register(engine, ps, pocs)
    try:
        got = engine.me_picture(poc, w, h, p, [l0, l1])
    finally:
        _release(engine, pocs)
    exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
    bad = planes.compare_outputs(exp, got, used_refs=used_refs(l0, l1))
    assert not bad, bad


def test_pictures_in_flight_out_of_order(engine, oracle):
    """Several pictures submitted before any wait, waited for in another order (job contexts, streams, result slabs)."""
    w, h, n, preset = 320, 192, 17, 4
    ps = PictureSet(w, h, n, "motion", oracle)
    jobs = {poc: derive(preset, w, h, poc, n) for poc in (16, 8, 4, 12, 2)}
    pocs = sorted(set(k for poc, (_, _, l0, l1) in jobs.items() for k in [poc] + l0 + l1))
    _register(engine, ps, pocs)
    try:
        for poc, (_, p, l0, l1) in jobs.items():
            engine.submit(poc, p, [l0, l1])
        got = {}
        for poc in (4, 2, 16, 12, 8):
            got[poc] = planes.Outputs(w, h, jobs[poc][1])
            engine.wait(poc, got[poc])
    finally:
        _release(engine, pocs)
    for poc, (_, p, l0, l1) in jobs.items():
        exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
        bad = planes.compare_outputs(exp, got[poc], used_refs=used_refs(l0, l1))
        assert not bad, (poc, bad)


@pytest.mark.parametrize("debug", [True, False])
def test_submit_out_with_and_without_taps(engine, oracle, debug):
    """me_gpu_submit_out: destinations known at submit. With parity taps requested they come back exact; without them the
    engine does not write them at all and the product arrays are unchanged."""
    w, h, n, preset, poc = 320, 192, 17, 4, 8
    ps = PictureSet(w, h, n, "motion", oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    pocs = sorted(set([poc] + l0 + l1))
    _register(engine, ps, pocs)
    try:
        got = planes.Outputs(w, h, p, debug=debug)
        engine.submit_out(poc, p, [l0, l1], got)
        engine.wait(poc, got)
    finally:
        _release(engine, pocs)
    exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
    fields = planes.PRODUCT_FIELDS + (planes.DEBUG_FIELDS if debug else [])
    bad = planes.compare_outputs(exp, got, fields=fields, used_refs=used_refs(l0, l1))
    assert not bad, bad


def test_concurrent_callers(engine, oracle):
    """SURVEY 8(b) threading: the encoder's ME threads (up to 25) call submit / wait concurrently, keyed by picture, while
    picture analysis registers new pictures. Four host threads, each with its own pictures of a different preset."""
    import threading
    w, h, n = 320, 192, 17
    ps = PictureSet(w, h, n, "motion", oracle)
    work = [(4, [16, 8]), (8, [4, 12]), (2, [2, 6]), (0, [1, 10])]
    jobs, pocs = {}, set()
    for preset, lst in work:
        for poc in lst:
            jobs[poc] = derive(preset, w, h, poc, n)
            pocs |= set([poc] + jobs[poc][2] + jobs[poc][3])
    for k in sorted(pocs):
        ps.get(k)  # synthesise up front: the threads below only talk to the engine
    base, got, errors = 5000, {}, []
    barrier = threading.Barrier(len(work))

    def run(tid, lst):
        try:
            mine = sorted(set(k for poc in lst for k in [poc] + jobs[poc][2] + jobs[poc][3]))
            off = base + 100 * tid  # every thread registers its own copies under its own picture numbers
            barrier.wait()
            for k in mine:
                engine.register_picture(off + k, ps.get(k).planes[0], w, h)
            for rep in range(3):
                for poc in lst:
                    _, p, l0, l1 = jobs[poc]
                    q = type(p).from_buffer_copy(p)
                    q.picture_number = off + poc
                    engine.submit(off + poc, q, [[off + k for k in l0], [off + k for k in l1]])
                for poc in reversed(lst):
                    out = planes.Outputs(w, h, jobs[poc][1])
                    engine.wait(off + poc, out)
                    got[(poc, rep)] = out
            for k in mine:
                engine.release_picture(off + k)
        except Exception as e:  # surfaced in the main thread
            errors.append((tid, repr(e)))

    threads = [threading.Thread(target=run, args=(i, lst)) for i, (_, lst) in enumerate(work)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for poc, (_, p, l0, l1) in jobs.items():
        exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
        for rep in range(3):
            bad = planes.compare_outputs(exp, got[(poc, rep)], used_refs=used_refs(l0, l1))
            assert not bad, (poc, rep, bad)


def test_dg_detector_parity(engine, oracle):
    """k_dg_hme through me_gpu_dg_detector against the restatement of dg_detector_hme_level0, every metric exact;
    the golden metrics (frozen from the reference itself) are checked on the way."""
    import json
    import os
    from helpers import DG_CASES, DgPictures
    gold = {tuple(e["case"]): tuple(e["metrics"]) for e in json.load(open(os.path.join(os.path.dirname(__file__), "golden",
                                                                                        "dg_metrics.json")))}
    for case in DG_CASES:
        w, h, res, kind, a, b = case
        ps = DgPictures(w, h, kind, oracle)
        pocs = sorted({a, b})
        _register(engine, ps, pocs)
        got = engine.dg_detector(a, b, res)
        swapped = engine.dg_detector(b, a, res)
        _release(engine, pocs)
        assert got == oracle.dg_detector(ps.get(a), ps.get(b), res) == gold[case], case
        assert swapped == oracle.dg_detector(ps.get(b), ps.get(a), res), case


def test_dg_detector_parity_1080p_and_errors(engine, oracle):
    from helpers import DgPictures
    from me_b200.engine import MeGpuError
    for kind, a, b in (("motion", 24, 16), ("blocks", 1, 2)):
        ps = DgPictures(1920, 1080, kind, oracle)
        _register(engine, ps, [a, b])
        got = engine.dg_detector(a, b, 4)
        with pytest.raises(MeGpuError):
            engine.dg_detector(a, 999, 4)
        _release(engine, [a, b])
        assert got == oracle.dg_detector(ps.get(a), ps.get(b), 4), kind
    small, big = DgPictures(320, 192, "iid", oracle), DgPictures(640, 360, "iid", oracle)
    engine.register_picture(1, small.get(1).planes[0], 320, 192)
    engine.register_picture(2, big.get(2).planes[0], 640, 360)
    with pytest.raises(MeGpuError):
        engine.dg_detector(1, 2, 4)
    _release(engine, [1, 2])


def _tf_cases():
    from helpers import TF_CASES
    return TF_CASES


@pytest.mark.parametrize("case", _tf_cases(), ids=lambda c: "tf_" + "_".join(str(v) for v in c))
def test_tf_me_parity_small(engine, oracle, case):
    """Temporal-filter ME (me_type = ME_GPU_ME_MCTF) through me_gpu_submit: tf_* results, taps and byte-SAD count."""
    from helpers import derive_tf
    w, h, lvl, kind, c, nb, th = case
    ps = PictureSet(w, h, 20, kind, oracle)
    _, p = derive_tf(lvl, w, h, c, me_exit_th=th)
    _register(engine, ps, [c, nb])
    try:
        got = engine.me_picture(c, w, h, p, [[nb], []])
    finally:
        _release(engine, [c, nb])
    exp = oracle.me_picture(ps.get(c), [[ps.get(nb)], []], p)
    bad = planes.compare_outputs(exp, got, fields=planes.TF_FIELDS + planes.DEBUG_FIELDS, used_refs=[1, 0])
    assert not bad, bad
    assert exp.tf_blocks == got.tf_blocks and sum(got.tf_blocks) == got.b64_count
    assert exp.byte_sad_ops == got.byte_sad_ops


@pytest.mark.parametrize("lvl,kind,th,first,count", [(0, "motion", 2000, 0, 60), (1, "motion", 0, 230, 60), (2, "iid", 30000, 450, 60)])
def test_tf_me_parity_1080p_sample(engine, oracle, lvl, kind, th, first, count):
    from helpers import derive_tf
    w, h, c, nb = 1920, 1080, 8, 6
    ps = PictureSet(w, h, 17, kind, oracle)
    _, p = derive_tf(lvl, w, h, c, me_exit_th=th)
    _register(engine, ps, [c, nb])
    try:
        got = engine.me_picture(c, w, h, p, [[nb], []])
    finally:
        _release(engine, [c, nb])
    exp = oracle.me_picture(ps.get(c), [[ps.get(nb)], []], p, first, count)
    for name in planes.TF_FIELDS:
        assert np.array_equal(exp.arrays[name][first:first + count], got.arrays[name][first:first + count]), name
    assert sum(got.tf_blocks) == 510


def test_tf_me_rejects_bad_reference_counts(engine, oracle):
    from helpers import derive_tf
    from me_b200.engine import MeGpuError
    w, h = 320, 192
    ps = PictureSet(w, h, 9, "motion", oracle)
    _register(engine, ps, [4, 5])
    try:
        _, p = derive_tf(1, w, h, 4)
        p.num_of_list_to_search = 2
        p.num_of_ref_pic_to_search[1] = 1
        with pytest.raises(MeGpuError):
            engine.submit(4, p, [[5], [5]])
        _, p = derive_tf(1, w, h, 4)
        p.me_type = 2
        with pytest.raises(MeGpuError):
            engine.submit(4, p, [[5], []])
    finally:
        _release(engine, [4, 5])


def test_wait_protocol_with_several_threads(engine, oracle):
    """SURVEY 8(b) threading: every segment thread of a picture blocks on the same job. One thread retires it
    (me_gpu_wait), the others block in me_gpu_wait_done until the results have been delivered; a second me_gpu_wait while
    the first is blocked is refused with a defined code; me_gpu_cancel retires a job nobody wants; a too small output
    buffer leaves the job in flight; the temporal-filter pair of a picture can be in flight beside its open-loop ME."""
    import threading
    from helpers import derive_tf
    from me_b200.engine import MeGpuError
    w, h, n, preset, poc = 320, 192, 17, 4, 8
    ps = PictureSet(w, h, n, "motion", oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    pocs = sorted(set([poc] + l0 + l1))
    _register(engine, ps, pocs)
    lib = engine.lib
    try:
        exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
        for rep in range(4):
            engine.submit(poc, p, [l0, l1])
            got = planes.Outputs(w, h, p)
            rcs, order = [], []
            gate = threading.Barrier(6)

            def seg(i):
                gate.wait()
                if i == 0:
                    rcs.append(("wait", lib.me_gpu_wait(poc, got.c)))
                    order.append("retired")
                else:
                    rc = lib.me_gpu_wait_done(poc)
                    order.append("done")
                    rcs.append(("done", rc))

            ts = [threading.Thread(target=seg, args=(i,)) for i in range(6)]
            for t in ts:
                t.start()
            for t in ts:
                t.join()
            assert all(rc == 0 for _, rc in rcs), rcs
            bad = planes.compare_outputs(exp, got, used_refs=used_refs(l0, l1))
            assert not bad, (rep, bad)
            assert lib.me_gpu_wait_done(poc) == 0  # already retired: still a defined answer
        assert lib.me_gpu_wait_done(987654) == -3  # ME_GPU_ERR_NOT_FOUND
        # too small a buffer: refused, the job stays; cancel retires it and frees its picture references
        engine.submit(poc, p, [l0, l1])
        small = planes.Outputs(w, h, p)
        small.c.b64_count = 1
        assert lib.me_gpu_wait(poc, small.c) == -2
        assert lib.me_gpu_cancel(poc) == 0
        assert lib.me_gpu_cancel(poc) == -3
        # open-loop ME and a temporal-filter pair of the same picture in flight together, waited for by identity
        _, ptf = derive_tf(1, w, h, poc, me_exit_th=400)
        engine.submit(poc, p, [l0, l1])
        engine.submit(poc, ptf, [[l0[0]], []])
        with pytest.raises(MeGpuError):
            engine.submit(poc, ptf, [[l0[0]], []])  # the same job twice
        tf = planes.Outputs(w, h, ptf)
        assert lib.me_gpu_wait_pair(poc, ptf.me_type, l0[0], tf.c) == 0
        got = planes.Outputs(w, h, p)
        engine.wait(poc, got)
        assert not planes.compare_outputs(exp, got, used_refs=used_refs(l0, l1))
        exp_tf = oracle.me_picture(ps.get(poc), [[ps.get(l0[0])], []], ptf)
        assert not planes.compare_outputs(exp_tf, tf, fields=planes.TF_FIELDS)
    finally:
        _release(engine, pocs)


@pytest.mark.parametrize("preset,poc,kind,kw,first,count", [(2, 4, "motion", dict(sc_class1=1), 215, 24), (0, 1, "iid", dict(sc_class1=1), 29, 1),
                                                            (9, 1, "motion", dict(rtc_tune=1), 0, 510)])
def test_me_parity_1080p_dependent_references(engine, oracle, preset, poc, kind, kw, first, count):
    """the same at 1080p, where the integer searches are tiled: with enable_me_sr_adjustment == 2 reference (0, 0)'s tiles run
    in a k_fullsearch launch of their own before the other references' areas are decided"""
    w, h, n = 1920, 1080, 17
    ps = PictureSet(w, h, n, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n, **kw)
    pocs = sorted(set([poc] + l0 + l1))
    _register(engine, ps, pocs)
    try:
        got = engine.me_picture(poc, w, h, p, [l0, l1])
    finally:
        _release(engine, pocs)
    exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p, first, count)
    for name in exp.arrays:
        a, b = exp.arrays[name][first:first + count], got.arrays[name][first:first + count]
        if name in planes.DEBUG_FIELDS:
            mask = np.zeros((2, 4), dtype=bool)
            mask[0, :len(l0)] = True
            mask[1, :len(l1)] = True
            a, b = a[:, mask], b[:, mask]
        assert np.array_equal(a, b), name
