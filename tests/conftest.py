"""Shared fixtures.  `-m "not gpu"` tests run on a CPU-only box (oracle, host mirror, ABI surface,
gloo sharding); `-m gpu` tests are the parity tests proper and call the CUDA path through the C ABI.
Nothing here reads /root/reference (it does not exist on the GPU box)."""
import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    pkg_so = os.path.join(ROOT, "ray-tracer_b200", "librt_b200.so")
    orc_so = os.path.join(ROOT, "oracle", "liboracle.so")
    if not (os.path.exists(pkg_so) and os.path.exists(orc_so)):
        entry = importlib.import_module("__graft_entry__")
        entry.build()


@pytest.fixture(scope="session")
def orc():
    """The f64 CPU oracle (checker)."""
    _ensure_built()
    import oracle_py

    return oracle_py


@pytest.fixture(scope="session")
def rt():
    """ctypes binding of the product library (librt_b200.so)."""
    _ensure_built()
    return importlib.import_module("ray-tracer_b200")


@pytest.fixture(scope="session")
def device(rt):
    """RtContext on cuda:0; fails loudly (no CPU fallback) when there is no device."""
    return rt.Device(0)
