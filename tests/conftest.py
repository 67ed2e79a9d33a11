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


def _use_emu(monkeypatch):
    import pfb_b200._lib as L
    import pfb_b200._dev as D
    from emu.emulib import emu
    monkeypatch.setattr(L, "_LIB", emu())
    monkeypatch.setattr(D, "_EMULATED", True)


@pytest.fixture()
def emu_backend(monkeypatch):
    """Route the product's Python host layer to the CPU execution-model
    emulator build of the kernel sources (tests only; see tests/emu)."""
    import pfb_b200.operators.psf as psf
    _use_emu(monkeypatch)
    psf.clear_plan_cache()
    yield
    psf.clear_plan_cache()


@pytest.fixture(params=["emu", pytest.param("cuda", marks=pytest.mark.gpu)])
def backend(request, monkeypatch):
    """The same host-API test runs twice: on the CPU emulator build of the
    kernel sources (`-m "not gpu"`) and on the real libpfb_b200.so / B200
    (`-m gpu`, through ctypes -> the C-ABI of include/pfb_b200.h)."""
    import pfb_b200.operators.psf as psf
    if request.param == "emu":
        _use_emu(monkeypatch)
    else:
        import torch
        import pfb_b200._lib as L
        assert torch.cuda.is_available(), "gpu test needs a CUDA device"
        L.lib()     # raises if libpfb_b200.so is not built: no fallback
    psf.clear_plan_cache()
    yield request.param
    psf.clear_plan_cache()
