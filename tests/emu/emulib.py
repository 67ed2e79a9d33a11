"""Loads the emulated build of the kernel sources (tests only)."""
import ctypes
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
_EMU = None


def emu():
    global _EMU
    if _EMU is None:
        sys.path.insert(0, ROOT)
        from .build_emu import build
        import pfb_b200._lib as L
        _EMU = L.bind(ctypes.CDLL(build()))
    return _EMU


def ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


def dt(a):
    return 0 if a.dtype == np.float32 else 1


def ok(rc):
    assert rc == 0, emu().pfb_last_error().decode()
