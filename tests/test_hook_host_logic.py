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


@pytest.mark.skipif(not os.path.isdir("/root/reference/Source"), reason="builds against the reference's headers")
def test_binding_host_logic_is_race_free_under_threadsanitizer():
    """the same scenario as a stand-alone program with the hook, the harness and the stand-in engine built with
    -fsanitize=thread (make -C oracle tsan): the once-per-picture / once-per-pair protocols under a race detector"""
    import subprocess
    oracle = os.path.join(HERE, "..", "oracle")
    b = subprocess.run(["make", "-C", oracle, "tsan"], capture_output=True, text=True)
    if b.returncode != 0:
        pytest.skip("no ThreadSanitizer runtime for this compiler: " + b.stderr[-300:])
    exe, fake = os.path.join(oracle, "_ref", "host_logic_tsan"), os.path.join(oracle, "_ref", "libme_fake_engine_tsan.so")
    r = subprocess.run([exe, os.path.abspath(fake)], capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, TSAN_OPTIONS="halt_on_error=0 exitcode=66"))
    out = r.stdout + r.stderr
    assert "ThreadSanitizer" not in out, out[-3000:]
    assert r.returncode == 0 and "ref_hook_host_logic rc 0" in out, out[-1000:]
