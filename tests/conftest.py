"""pytest configuration: the `gpu` marker, oracle fixtures, golden-fixture loaders.

Only tests/ (here), __graft_entry__.smoke() and bench.py's CPU-baseline legs may touch oracle/.
/root/reference is never read at test time: the reference-built oracle (oracle/_ref/libaqref.so) is a
prebuilt artefact that travels with the snapshot, and tests that need it skip when it is absent.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_sessionstart(session):
    """A fresh checkout has no built artefacts (they are git-ignored): build them once, the way __graft_entry__.build() does, if the
    CUDA library is missing and nvcc is here.  On the GPU box the built files travel with the snapshot and nothing happens."""
    import shutil
    lib = os.path.join(ROOT, "aquery2_b200", "libaqgpu.so")
    if not os.path.exists(lib) and shutil.which("nvcc") and os.environ.get("AQ_TESTS_NO_BUILD") != "1":
        sys.path.insert(0, ROOT)
        import __graft_entry__
        __graft_entry__.build()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def port():
    from oracle import aqo
    return aqo.load_port()


@pytest.fixture(scope="session")
def ref():
    from oracle import aqo
    r = aqo.load_ref()
    if r is None:
        pytest.skip("oracle/_ref/libaqref.so not built (needs /root/reference at build time)")
    return r


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name))
    return load


@pytest.fixture(scope="session")
def gpu():
    """The ctypes binding with the library initialised on device 0.  No fallback: a missing library
    or device is a hard failure of every -m gpu test."""
    from aquery2_b200 import aqgpu
    aqgpu.init(0)
    return aqgpu
