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
    if ec.CASES[name][3] >= 6:  # quick: the stock encoder on this box too
        live, _ = ec.encode(ec.STOCK, name, workdir)
        assert live == gold[name], ("stock encoder differs from its golden on this box", name)


def test_encoder_fails_loudly_when_the_engine_library_is_missing(workdir):
    if not os.path.exists(ec.GPU):
        pytest.skip("integration/_build/SvtAv1EncApp_gpu missing")
    with pytest.raises(RuntimeError, match="could not be brought up"):
        ec.encode(ec.GPU, "320x192_m8", workdir, env={"SVT_ME_GPU_LIB": "/nonexistent/libme_gpu.so"})
