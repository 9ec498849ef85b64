"""Host logic of the encoder binding (integration/svt_me_gpu_hook.c) on the CPU: the hook compiled against the reference's
headers (oracle/_ref/libme_ref.so) bound to a STAND-IN engine (tests/emul/libme_fake_engine.so) that computes nothing but
rejects what a real engine would reject. The scenario (oracle/ref_harness.c: ref_hook_host_logic) covers what only the
GPU-box encoder tests exercised before: registration keyed by the PA reference object, the re-registration after the temporal
filter rewrote a picture, the release when the pool re-uses an object, the identities the detector receives, the
once-per-pair protocol of the temporal filter among eight segment threads (two jobs for sixteen requests, totals handed out
once, pairs freed after the last segment), and the once-per-picture protocol of svt_me_gpu_segment among six segment threads
(one submit, one wait, results scattered into the pcs)."""
import ctypes as C
import os

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "..", "oracle", "_ref", "libme_ref.so")
FAKE = os.path.join(HERE, "emul", "libme_fake_engine.so")


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref is built where /root/reference exists")
def test_binding_host_logic_against_a_stand_in_engine():
    assert os.path.exists(FAKE), "tests/emul/libme_fake_engine.so missing: make -C tests/emul"
    lib = C.CDLL(REF)
    lib.ref_hook_host_logic.argtypes = [C.c_char_p, C.POINTER(C.c_uint64)]
    report = (C.c_uint64 * 9)()
    rc = lib.ref_hook_host_logic(os.path.abspath(FAKE).encode(), report)
    assert rc == 0, f"step {-rc} of ref_hook_host_logic failed; engine counters {list(report)}"
    registers, releases, submits, waits, wait_pairs, read_planes, dg_calls, rejected, held = list(report)
    assert (registers, releases, held) == (5, 2, 3)
    assert (submits, wait_pairs, waits, dg_calls, rejected) == (5, 4, 1, 1, 0)
    assert read_planes == 2 * registers  # a quarter and a sixteenth plane per pyramid
