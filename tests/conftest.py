import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


def golden(name):
    import numpy as np
    return np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))


@pytest.fixture(scope="session")
def golden_loader():
    return golden


@pytest.fixture()
def emu_backend(monkeypatch):
    """Route the product's Python host layer to the CPU execution-model
    emulator build of the kernel sources (tests only; see tests/emu)."""
    import pfb_b200._lib as L
    import pfb_b200._dev as D
    from emu.emulib import emu
    import pfb_b200.operators.psf as psf
    monkeypatch.setattr(L, "_LIB", emu())
    monkeypatch.setattr(D, "_EMULATED", True)
    psf.clear_plan_cache()
    yield
    psf.clear_plan_cache()
