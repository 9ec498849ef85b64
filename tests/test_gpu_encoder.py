"""the engine inside the encoder. SvtAv1EncApp built from a copy of the reference with the binding of
integration/ applied (integration/build_encoder.sh; the binaries travel prebuilt) encodes synthetic clips with open-loop
ME on the B200; the bitstream must be byte-identical (md5 of the .ivf) to the one of the unmodified encoder, frozen in
tests/golden/encoder_md5.json by tests/golden/make_encoder_md5.py and, for the quick cases, re-encoded here by the stock
binary on this box as well. Presets 0, 2, 4, 6, 8; I slices, a scene change, overlay pictures and short mini-GOPs in the
"scene" clips; a 65-picture clip in mini-GOPs of 32 for the dynamic-GOP detector; ME runs on every non-I picture of every case."""
import json
import os
import re

import pytest

import encoder_cases as ec
from me_b200 import abi

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def workdir(tmp_path_factory):
    return str(tmp_path_factory.mktemp("enc"))


@pytest.mark.parametrize("name", list(ec.CASES))
def test_bitstream_identical_with_me_on_the_gpu(workdir, name):
    if not (os.path.exists(ec.GPU) and os.path.exists(ec.STOCK)):
        pytest.skip("integration/_build/SvtAv1EncApp_* missing: run integration/build_encoder.sh where /root/reference exists")
    gold = json.load(open(ec.MD5_FILE))["md5"]
    assert name in gold, f"no stock md5 for {name}: run tests/golden/make_encoder_md5.py"
    got, log = ec.encode(ec.GPU, name, workdir, env={"SVT_ME_GPU_LIB": abi.GPU_LIB})
    assert "open-loop ME on the B200 engine" in log, log[-1500:]
    assert got == gold[name], (name, got, gold[name])
    # the pyramid too came from the device: every picture searched had its planes built there for picture analysis
    m = re.search(r"(\d+) pictures searched, (\d+) pyramids built on the device \((\d+) for picture analysis", log)
    assert m, log[-1500:]
    searched, built, for_analysis = (int(x) for x in m.groups())
    assert searched > 0 and for_analysis == built >= searched, (searched, built, for_analysis)
    # mini-GOPs of 32: the dynamic-GOP detector's three picture pairs per decision went to the device too
    dg = int(re.search(r"(\d+) dynamic-GOP detector calls", log).group(1))
    assert dg > 0 if "6L" in name else dg >= 0, dg
    # the temporal filter's ME (key frames and base-layer pictures are filtered in every case) ran on the device as well
    assert int(re.search(r"(\d+) temporal-filter pairs", log).group(1)) > 0, log[-800:]
    if ec.CASES[name][3] >= 6:  # quick: the stock encoder on this box too
        live, _ = ec.encode(ec.STOCK, name, workdir)
        assert live == gold[name], ("stock encoder differs from its golden on this box", name)


def test_encoder_fails_loudly_when_the_engine_library_is_missing(workdir):
    if not os.path.exists(ec.GPU):
        pytest.skip("integration/_build/SvtAv1EncApp_gpu missing")
    with pytest.raises(RuntimeError, match="could not be brought up"):
        ec.encode(ec.GPU, "320x192_m8", workdir, env={"SVT_ME_GPU_LIB": "/nonexistent/libme_gpu.so"})


def test_bench_reference_lists_are_the_encoders_own(workdir):
    """VERDICT missing #8: the bench workloads take their reference lists from tests/golden/encoder_lists.json. That file must be
    what the encoder's own prediction structure (pd_process.c:3675-3693) hands to the ME stage: the patched encoder prints
    the lists it receives (SVT_ME_GPU_LOG_LISTS) for the generator's clip at presets 4 and 8, compared picture by picture."""
    if not os.path.exists(ec.GPU):
        pytest.skip("integration/_build/SvtAv1EncApp_gpu missing")
    import workloads
    frozen = json.load(open(workloads.LISTS_FILE))["lists"]
    for preset in (4, 8):
        name = f"lists_m{preset}"
        ec.CASES[name] = (320, 192, 49, preset, ["--lp", "1", "--enable-tf", "0", "--hierarchical-levels", "4"])
        try:
            _, log = ec.encode(ec.GPU, name, workdir, env={"SVT_ME_GPU_LIB": abi.GPU_LIB, "SVT_ME_GPU_LOG_LISTS": "1"})
        finally:
            del ec.CASES[name]
        seen = {}
        for m in re.finditer(r"lists poc (\d+) tl (\d+):(.*)", log):
            lists = [[int(x) for x in part.split()[1:]] for part in re.findall(r"L\d[ \d]*", m.group(3))]
            seen[m.group(1)] = ([int(m.group(2))] + lists + [[]])[:3]
        assert len(seen) == 48, len(seen)  # every picture but the key frame went through ME
        assert seen == frozen[str(preset)], {k: (seen.get(k), v) for k, v in frozen[str(preset)].items() if seen.get(k) != v}
    # and the workloads use them: 1080p-m4 = preset 4 cut to 2 + 2, the 4K ones = presets 2 / 0 whole
    for wl, poc in (("1080p-m4", 24), ("4k-m2", 20), ("4k-m0", 31), ("1080p-m8", 28)):
        tl, l0, l1 = workloads.lists_for(wl, poc)
        f = frozen[str(workloads.WORKLOADS[wl]["preset"])][str(poc)]
        r = workloads.WORKLOADS[wl]["refs"] or (9, 9)
        assert (tl, l0, l1) == (f[0], f[1][:r[0]], f[2][:r[1]])
