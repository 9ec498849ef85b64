"""GPU parity cases that need an engine of their own (the C-ABI holds one engine per process: these run after the
module-scoped engine of test_gpu_parity.py has been shut down): 4K geometry, two devices in one process."""
import numpy as np
import pytest

from helpers import PictureSet, derive, used_refs
from me_b200 import planes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle():
    from checkers import Oracle
    return Oracle()


def _register(engine, ps, pocs):
    for k in pocs:
        engine.register_picture(k, ps.get(k).planes[0], ps.w, ps.h)


@pytest.mark.parametrize("preset,poc,n,kind,first,count", [
    (4, 8, 17, "motion", 1000, 6), (2, 4, 17, "motion", 2039, 1),
    # BASELINE configs[2]: preset 2, second mini-GOP (4 + 3 references, 128x128 areas at the far references)
    (2, 18, 33, "iid", 777, 3), (2, 24, 33, "motion", 59, 2),
    # BASELINE configs[4]: preset 0, HME-L0 up to 480x480, integer search up to 256x256, 7 references; the last b64s
    # are the bottom-right corner (48 rows high)
    (0, 18, 33, "iid", 1000, 2), (0, 24, 33, "motion", 2038, 2), (0, 32, 33, "iid", 0, 1)])
def test_me_parity_4k_sample(oracle, preset, poc, n, kind, first, count):
    """BASELINE configs[2] / configs[4] geometry (3840x2160, 2040 b64): whole picture on the GPU, a few b64s on the
    oracle (a preset-0 b64 with 7 references is ~5 s of scalar C)."""
    from me_b200.engine import Engine
    w, h = 3840, 2160
    ps = PictureSet(w, h, n, kind, oracle)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    pocs = sorted(set([poc] + l0 + l1))
    eng = Engine(w, h, 4, 3, pictures_in_flight=10)
    try:
        _register(eng, ps, pocs)
        got = eng.me_picture(poc, w, h, p, [l0, l1])
    finally:
        eng.close()
    exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p, first, count)
    for name in exp.arrays:
        a, b = exp.arrays[name][first:first + count], got.arrays[name][first:first + count]
        if name in planes.DEBUG_FIELDS:
            mask = np.zeros((2, 4), dtype=bool)
            mask[0, :len(l0)] = True
            mask[1, :len(l1)] = True
            a, b = a[:, mask], b[:, mask]
        assert np.array_equal(a, b), name


def test_two_devices_in_one_process(oracle):
    """me_gpu_init over two devices: pyramids replicated over NVLink, jobs dealt round-robin, results identical."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from me_b200.engine import Engine
    w, h, n, preset = 320, 192, 17, 4
    ps = PictureSet(w, h, n, "motion", oracle)
    jobs = {poc: derive(preset, w, h, poc, n) for poc in (16, 8, 4, 12)}
    pocs = sorted(set(k for poc, (_, _, l0, l1) in jobs.items() for k in [poc] + l0 + l1))
    eng = Engine(w, h, 4, 3, pictures_in_flight=24, devices=[0, 1])
    try:
        _register(eng, ps, pocs)
        pyr = eng.read_pyramid(pocs[0], w, h)  # read from the LAST device: the replica
        for lvl in range(3):
            assert np.array_equal(pyr[lvl], ps.get(pocs[0]).planes[lvl]), lvl
        for poc, (_, p, l0, l1) in jobs.items():
            eng.submit(poc, p, [l0, l1])
        got = {}
        for poc in jobs:
            got[poc] = planes.Outputs(w, h, jobs[poc][1])
            eng.wait(poc, got[poc])
    finally:
        eng.close()
    for poc, (_, p, l0, l1) in jobs.items():
        exp = oracle.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
        bad = planes.compare_outputs(exp, got[poc], used_refs=used_refs(l0, l1))
        assert not bad, (poc, bad)


@pytest.mark.parametrize("w,h,preset,poc,kind", [(320, 192, 4, 4, "motion"), (320, 192, 8, 1, "iid"), (640, 360, 2, 8, "motion"),
                                                 (200, 136, 0, 2, "motion")])
def test_reference_side_hook_fills_the_pcs_like_the_cpu_loop(w, h, preset, poc, kind):
    """integration/svt_me_gpu_hook.c (the binding INTEGRATION.md describes, compiled against the reference's own headers
    into oracle/_ref/libme_ref.so) in place of the b64 loop of svt_aom_motion_estimation_kernel: the MeSbResults of
    every b64 and the eight per-b64 pcs arrays equal, byte for byte, what the reference's own
    svt_aom_motion_estimation_b64 loop leaves in an identical pcs."""
    from me_b200 import abi
    from checkers import Reference
    if not Reference.available():
        pytest.skip("oracle/_ref/libme_ref.so is built where /root/reference exists and travels prebuilt")
    ref = Reference()
    assert ref.has_hook()
    n = 17
    ps = PictureSet(w, h, n, kind, ref)
    cfg, p, l0, l1 = derive(preset, w, h, poc, n)
    refs = [[ps.get(k) for k in l0], [ps.get(k) for k in l1]]
    exp = ref.me_picture(ps.get(poc), refs, p, simd=0, debug=False)
    got = ref.hook_me_picture(ps.get(poc), refs, p, abi.GPU_LIB)
    bad = planes.compare_outputs(exp, got, fields=planes.PRODUCT_FIELDS)
    assert not bad, bad
    assert got.c.pu_count == exp.c.pu_count


def test_reference_side_hook_temporal_filter_and_dg_detector():
    """The binding's temporal-filter pair (svt_me_gpu_tf_pair + svt_me_gpu_tf_load_block: what the block loop of
    produce_temporally_filtered_pic then finds in its MeContext) and its dynamic-GOP detector call
    (dg_detector->metrics), against the reference's own code on the same inputs."""
    from helpers import DgPictures, derive_tf
    from me_b200 import abi
    from checkers import Reference
    if not Reference.available():
        pytest.skip("oracle/_ref/libme_ref.so is built where /root/reference exists and travels prebuilt")
    ref = Reference()
    for w, h, lvl, kind, c, nb, th in [(320, 192, 1, "motion", 8, 11, 400), (264, 200, 0, "iid", 4, 5, 20000),
                                       (320, 192, 3, "motion", 8, 9, 6000)]:
        ps = PictureSet(w, h, 20, kind, ref)
        _, p = derive_tf(lvl, w, h, c, me_exit_th=th)
        exp = ref.me_picture(ps.get(c), [[ps.get(nb)], []], p, simd=0, debug=False)
        got = ref.hook_tf_pair(ps.get(c), ps.get(nb), p, abi.GPU_LIB)
        bad = planes.compare_outputs(exp, got, fields=planes.TF_FIELDS)
        assert not bad, (lvl, bad)
        assert exp.tf_blocks == got.tf_blocks
    for w, h, res, kind, a, b in [(320, 192, 4, "motion", 4, 3), (640, 360, 2, "motion", 8, 0), (320, 192, 4, "blocks", 2, 1)]:
        ps = DgPictures(w, h, kind, ref)
        assert ref.hook_dg_detector(ps.get(a), ps.get(b), res, abi.GPU_LIB) == ref.dg_detector(ps.get(a), ps.get(b), res)


@pytest.mark.parametrize("w,h", [(320, 192), (1920, 1080), (200, 136)])
def test_reference_side_hook_pyramid_for_picture_analysis(w, h):
    """picture analysis takes the device-built quarter / sixteenth planes through the binding
    (svt_me_gpu_pyramid_for_analysis) instead of svt_aom_downsample_filtering_input_picture: the planes it gets, padding
    included, and the reference's own histogram of the sixteenth plane (pic_analysis_process.c:166, :1495) equal the CPU ones."""
    from me_b200 import abi
    from checkers import Reference
    if not Reference.available():
        pytest.skip("oracle/_ref/libme_ref.so is built where /root/reference exists and travels prebuilt")
    ref = Reference()
    cpu = PictureSet(w, h, 3, "motion", ref).get(2)  # pyramid by the reference's own downsampler
    dev = planes.Picture(w, h, 2)
    dev.planes[0][...] = cpu.planes[0]
    h_cpu, h_gpu = ref.hook_pyramid_for_analysis(dev, cpu, abi.GPU_LIB)
    for lvl in (1, 2):
        assert np.array_equal(dev.planes[lvl], cpu.planes[lvl]), lvl
    assert np.array_equal(h_cpu, h_gpu) and int(h_cpu.sum()) == (w // 4) * (h // 4)
