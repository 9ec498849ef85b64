"""bench.py contract checks that need no GPU: the reference arm's JSON line, the workload table against BASELINE.json."""
import json
import os
import subprocess
import sys

import pytest

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))


def test_workloads_cover_the_baseline_configs():
    sys.path.insert(0, REPO)
    import bench
    base = json.load(open(os.path.join(REPO, "BASELINE.json")))
    covered = {wl["baseline_config"] for wl in bench.WORKLOADS.values()}
    # configs[3] (8 independent 1080p streams) is configs[0]'s workload under torchrun, one stream per GPU
    assert covered | {3} == set(range(len(base["configs"])))
    assert bench.WORKLOADS["1080p-m4"]["baseline_config"] == 1 and bench.WORKLOAD == "1080p-m4"
    for name, wl in bench.WORKLOADS.items():
        bench.set_workload(name)
        assert bench.b64_count() == ((wl["w"] + 63) // 64) * ((wl["h"] + 63) // 64)
        _, p, l0, l1 = bench.derive(bench.STEP_POCS[0])
        assert len(l0) >= 1 and p.max_refs >= len(l0) + len(l1)
    bench.set_workload("1080p-m4")


def test_reference_arm_line():
    if not os.path.exists(os.path.join(REPO, "oracle", "_ref", "libme_ref.so")):
        pytest.skip("oracle/_ref is built where /root/reference exists")
    r = subprocess.run([sys.executable, os.path.join(REPO, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, "stdout carries ONE JSON line"
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "open-loop ME frames/sec" and d["unit"] == "frames/s"
    assert d["value"] > 0 and d["higher_is_better"] is True and d["steps"] == 1
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "configs[1]" in d["config"]["workload"]
    assert d["parity"]["equal"] is True  # the arm's own results against the frozen digests
    assert {a["simd"] for a in d["cpu_baseline"]["other_arms"]} == {"avx512", "avx2"}


def test_reference_arm_controls_do_not_need_the_product():
    """workloads.reference_params (the reference's own svt_aom_sig_deriv_me + the mrp table) gives, byte for byte, the
    MeGpuParams the product's host library derives: the reference arm can run without loading libme_host.so."""
    if not os.path.exists(os.path.join(REPO, "oracle", "_ref", "libme_ref.so")):
        pytest.skip("oracle/_ref is built where /root/reference exists")
    import checkers
    import workloads
    ref = checkers.Reference()
    for name in workloads.WORKLOADS:
        for poc in workloads.STEP_POCS:
            _, p, l0, l1 = workloads.derive_for(name, poc)
            q, m0, m1 = workloads.reference_params(name, poc, ref)
            assert bytes(p) == bytes(q) and (l0, l1) == (m0, m1), (name, poc)
    r = subprocess.run([sys.executable, "-c", "import sys, os; sys.argv=['bench.py','--impl','reference','--steps','1','--warmup','0'];"
                        f"sys.path.insert(0, {REPO!r}); import bench; bench.main(); "
                        "print('MAPS', [l.split()[-1] for l in open('/proc/self/maps') if 'libme_' in l])"],
                       capture_output=True, text=True, timeout=600, cwd=REPO)
    assert r.returncode == 0, r.stderr[-2000:]
    maps = [l for l in r.stdout.splitlines() if l.startswith("MAPS")][0]
    assert "libme_ref.so" in maps and "libme_host.so" not in maps and "libme_gpu.so" not in maps, maps
