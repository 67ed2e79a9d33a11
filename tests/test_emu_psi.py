"""wavelet.cu (Psi.dot / Psi.hdot kernels) on the CPU emulator vs the oracle
and the reference's golden vectors."""
import ctypes

import numpy as np
import pytest

from oracle import ref
from conftest import golden
from emu.emulib import emu, ptr, dt, ok

pmp = pytest.mark.parametrize
IDS = {'self': 0, 'db1': 1, 'db2': 2, 'db3': 3, 'db4': 4, 'db5': 5}


def rel(a, b):
    return np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)


class EmuPsi:
    def __init__(self, nband, nx, ny, bases, nlevel, dtype):
        self.L = emu()
        ids = (ctypes.c_int * len(bases))(*[IDS[b] for b in bases])
        self.plan = ctypes.c_void_p()
        rc = self.L.pfb_psi_plan_create(ctypes.byref(self.plan), nband, nx, ny, len(bases), ids,
                                        nlevel, 0 if dtype == np.float32 else 1)
        self.rc = rc
        if rc == 0:
            a, b = ctypes.c_int(), ctypes.c_int()
            ok(self.L.pfb_psi_plan_query(self.plan, ctypes.byref(a), ctypes.byref(b)))
            self.Nxmax, self.Nymax = a.value, b.value

    def dot(self, x, alpha):
        ok(self.L.pfb_psi_dot(self.plan, ptr(x), ptr(alpha), None))

    def hdot(self, alpha, xo, alpha2=None):
        ok(self.L.pfb_psi_hdot(self.plan, ptr(alpha), ptr(alpha2), ptr(xo), None))

    def __del__(self):
        if self.rc == 0:
            self.L.pfb_psi_plan_destroy(self.plan)


@pmp("tag", ["psi_f64_a", "psi_f64_b", "psi_f64_c", "psi_f32_a"])
def test_psi_golden(tag):
    g = golden(tag)
    x = g["x"]
    dtype = x.dtype.type
    nband, nx, ny = x.shape
    bases = [str(b) for b in g["bases"]]
    psi = EmuPsi(nband, nx, ny, bases, int(g["nlevel"]), dtype)
    assert (psi.Nxmax, psi.Nymax) == (int(g["Nxmax"]), int(g["Nymax"]))
    rng = np.random.default_rng(420)
    rng.standard_normal(x.shape)
    fill = rng.standard_normal(g["alpha"].shape).astype(dtype)
    coef_in = rng.standard_normal(fill.shape).astype(dtype)
    alpha = fill.copy()
    psi.dot(x, alpha)
    tol = 1e-13 if dtype == np.float64 else 1e-6
    assert rel(alpha, g["alpha"]) < tol
    untouched = g["alpha"] == fill
    assert np.array_equal(alpha[untouched], fill[untouched])
    xrec = np.full(x.shape, np.nan, dtype)
    psi.hdot(coef_in, xrec)
    assert rel(xrec, g["xrec"]) < tol


@pmp("dtype", [np.float32, np.float64])
@pmp("nx,ny,nlevel", [(128, 78, 1), (250, 64, 2), (200, 136, 3)])
@pmp("bases", [['self', 'db1', 'db2', 'db3', 'db4', 'db5'],      # db5 -> tile kernels
               ['self', 'db1', 'db2', 'db3', 'db4'],             # streaming kernels
               ['db4', 'db2']], ids=["all6", "stream5", "db42"])
def test_psi_roundtrip_and_oracle(dtype, nx, ny, nlevel, bases):
    """tests/test_psi_operator.py:18 of the reference: hdot(dot(x)) == nbasis*x,
    buffers pre-filled with random numbers; plus direct oracle comparison and
    the fused 2*alpha - alpha2 loader."""
    rng = np.random.default_rng(7)
    nband = 2
    x = rng.standard_normal((nband, nx, ny)).astype(dtype)
    psi = EmuPsi(nband, nx, ny, bases, nlevel, dtype)
    opsi = ref.Psi(nband, nx, ny, bases, nlevel)
    assert (psi.Nxmax, psi.Nymax) == (opsi.Nxmax, opsi.Nymax)
    fill = rng.standard_normal((nband, len(bases), psi.Nymax, psi.Nxmax)).astype(dtype)
    alpha, oalpha = fill.copy(), fill.copy()
    psi.dot(x, alpha)
    opsi.dot(x, oalpha)
    tol = 1e-13 if dtype == np.float64 else 1e-6
    assert rel(alpha, oalpha) < tol
    xrec = rng.standard_normal(x.shape).astype(dtype)
    psi.hdot(alpha, xrec)
    assert rel(xrec, len(bases) * x) < tol
    a2 = rng.standard_normal(alpha.shape).astype(dtype)
    xr2 = np.zeros_like(x)
    psi.hdot(alpha, xr2, a2)
    oxr = np.zeros_like(x)
    opsi.hdot((2 * alpha - a2).astype(dtype), oxr)
    assert rel(xr2, oxr) < 2 * tol


def test_plan_errors():
    p = EmuPsi(1, 64, 63, ['db1'], 1, np.float64)
    assert p.rc == 2       # odd size
    p = EmuPsi(1, 32, 32, ['self', 'db5'], 3, np.float64)
    assert p.rc == 3       # nlevel too large (ValueError in the reference)
    with pytest.raises(ValueError):
        ref.Psi(1, 32, 32, ['self', 'db5'], 3)


@pmp("dtype", [np.float32, np.float64])
@pmp("nx,ny,nlevel", [(128, 78, 1), (250, 64, 2), (200, 136, 3)])
@pmp("bases", [['self', 'db1', 'db2', 'db3', 'db4'], ['db3', 'db1']], ids=["stream5", "db31"])
def test_psi_natural_layout(dtype, nx, ny, nlevel, bases):
    """The natural-orientation layout (pfb_psi_nat_stride & co.) is a pure
    re-arrangement: dot_nat == to_nat(dot), hdot_nat == hdot, from_nat inverts
    to_nat on the pyramid cells and leaves every other slab cell alone."""
    rng = np.random.default_rng(11)
    nband = 2
    L = emu()
    psi = EmuPsi(nband, nx, ny, bases, nlevel, dtype)
    stride = ctypes.c_longlong()
    ok(L.pfb_psi_nat_stride(psi.plan, ctypes.byref(stride)))
    assert stride.value > 0 and stride.value % 32 == 0
    nb = len(bases)
    x = rng.standard_normal((nband, nx, ny)).astype(dtype)
    fill = rng.standard_normal((nband, nb, psi.Nymax, psi.Nxmax)).astype(dtype)
    alpha = fill.copy()
    psi.dot(x, alpha)
    written = alpha != fill
    # to_nat / from_nat
    nat = np.full((nband, nb, stride.value), np.nan, dtype)
    ok(L.pfb_psi_to_nat(psi.plan, nband, ptr(alpha), ptr(nat), None))
    assert np.isfinite(nat).all()
    assert np.count_nonzero(nat) == np.count_nonzero(alpha[written])
    back = fill.copy()
    ok(L.pfb_psi_from_nat(psi.plan, nband, ptr(nat), ptr(back), None))
    assert np.array_equal(back, alpha)
    # analysis straight into the layout (padding cells must stay as they are: zero)
    nat2 = np.zeros_like(nat)
    ok(L.pfb_psi_dot_nat(psi.plan, ptr(x), ptr(nat2), None))
    assert rel(nat2, nat) < (1e-13 if dtype == np.float64 else 1e-6)
    assert np.array_equal(nat2 == 0, nat == 0)
    # synthesis from the layout, plain and with the fused 2a - a2
    c1 = rng.standard_normal(fill.shape).astype(dtype)
    c2 = rng.standard_normal(fill.shape).astype(dtype)
    n1, n2 = np.empty_like(nat), np.empty_like(nat)
    ok(L.pfb_psi_to_nat(psi.plan, nband, ptr(c1), ptr(n1), None))
    ok(L.pfb_psi_to_nat(psi.plan, nband, ptr(c2), ptr(n2), None))
    xa, xb = np.empty_like(x), np.empty_like(x)
    psi.hdot(c1, xa)
    ok(L.pfb_psi_hdot_nat(psi.plan, ptr(n1), None, ptr(xb), None))
    assert np.array_equal(xa, xb)
    psi.hdot(c1, xa, c2)
    ok(L.pfb_psi_hdot_nat(psi.plan, ptr(n1), ptr(n2), ptr(xb), None))
    assert np.array_equal(xa, xb)
    # weights convert as a single "band"
    wn = np.empty((nb, stride.value), dtype)
    ok(L.pfb_psi_to_nat(psi.plan, 1, ptr(c1[1]), ptr(wn), None))
    assert np.array_equal(wn, n1[1])


def test_psi_natural_layout_unsupported():
    L = emu()
    psi = EmuPsi(1, 128, 128, ['db1', 'db5'], 2, np.float64)
    stride = ctypes.c_longlong(-1)
    ok(L.pfb_psi_nat_stride(psi.plan, ctypes.byref(stride)))
    assert stride.value == 0
    a = np.zeros((1, 2, psi.Nymax, psi.Nxmax))
    assert L.pfb_psi_to_nat(psi.plan, 1, ptr(a), ptr(a), None) == 1   # PFB_ERR_BAD_ARG
