"""The drop-in boundary: the sm_100a shared library loads (no GPU needed for
that) and exports every entry point include/pfb_b200.h declares; the ctypes
binding table of the host layer covers the same set.  No compute calls here."""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def declared():
    txt = open(os.path.join(ROOT, "include", "pfb_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(pfb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    import pfb_b200._lib as L
    so = os.path.join(os.path.dirname(L.__file__), "csrc", "libpfb_b200.so")
    if not os.path.exists(so):
        ge.build()
    lib = ctypes.CDLL(so)
    names = declared()
    assert len(names) >= 25, names
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/pfb_b200.h but not exported: {missing}"


def test_ctypes_table_matches_header():
    import pfb_b200._lib as L
    names = set(declared())
    bound = set(L._SIGS)
    assert names == bound, (sorted(names - bound), sorted(bound - names))


def test_product_path_fails_loudly_without_gpu():
    """No CPU fallback: without a CUDA device the host layer raises."""
    import pytest
    import torch
    import pfb_b200._dev as D
    if torch.cuda.is_available() or D._EMULATED:
        pytest.skip("a device (or the test emulator) is active")
    with pytest.raises(RuntimeError):
        D.device()
