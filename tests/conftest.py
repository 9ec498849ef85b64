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
    """CPU-side checkers, the host library and - where nvcc exists - the CUDA library: `make` every time (incremental), so
    that the tests never run against a library older than the sources. A box without the toolchain or without
    /root/reference (the GPU box for oracle/_ref) uses the prebuilt files that travelled with the snapshot."""
    pkg = os.path.join(REPO, "svt-av1-psyex_b200")
    have_cc = any(os.access(os.path.join(d, "gcc"), os.X_OK) for d in os.environ.get("PATH", "").split(os.pathsep))
    if have_cc:
        _make(os.path.join(REPO, "oracle"), "libme_oracle.so")
        if os.path.isdir("/root/reference/Source"):
            _make(os.path.join(REPO, "oracle"))
        _make(pkg, "libme_host.so")
        _make(os.path.join(REPO, "tests", "emul"))
        if os.path.exists("/usr/local/cuda/bin/nvcc"):
            _make(pkg, "libme_gpu.so")
    yield
