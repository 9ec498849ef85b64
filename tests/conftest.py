import os
import subprocess
import sys

import pytest

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, os.path.join(REPO, "svt-av1-psyex_b200", "python"))
sys.path.insert(0, os.path.dirname(__file__))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _make(path, target=None):
    cmd = ["make", "-s", "-C", path] + ([target] if target else [])
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


@pytest.fixture(scope="session", autouse=True)
def built_libs():
    """CPU-side checkers and the host library; the CUDA library is only (re)built when missing."""
    pkg = os.path.join(REPO, "svt-av1-psyex_b200")
    if not os.path.exists(os.path.join(REPO, "oracle", "libme_oracle.so")):
        _make(os.path.join(REPO, "oracle"), "libme_oracle.so")
    if os.path.isdir("/root/reference/Source") and not os.path.exists(os.path.join(REPO, "oracle", "_ref", "libme_ref.so")):
        _make(os.path.join(REPO, "oracle"))
    if not os.path.exists(os.path.join(pkg, "libme_host.so")):
        _make(pkg, "libme_host.so")
    if not os.path.exists(os.path.join(REPO, "tests", "emul", "libme_emul.so")):
        _make(os.path.join(REPO, "tests", "emul"))
    if not os.path.exists(os.path.join(pkg, "libme_gpu.so")) and os.path.exists("/usr/local/cuda/bin/nvcc"):
        _make(pkg, "libme_gpu.so")
    yield
