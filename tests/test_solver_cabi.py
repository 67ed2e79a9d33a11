"""Solver-level C-ABI (include/pfb_b200.h: pfb_pd_run, pfb_pcg_run, pfb_power_method_run)
against the golden vectors of the unmodified reference (tests/golden/make_golden.py) and
against the Python loops of pfb_b200.opt, on the CPU emulator build and on the B200."""
import ctypes
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
pmp = pytest.mark.parametrize


def golden(tag):
    return np.load(os.path.join(HERE, "golden", f"{tag}.npz"))


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def _setup(g):
    """device copies + plans for the golden solve problem"""
    from pfb_b200 import _dev
    from pfb_b200.operators.psf import get_plan
    from pfb_b200.operators.psi import Psi
    psfhat, dirty = g["psfhat"], g["dirty"]
    nband, nx, ny = dirty.shape
    nyp = 2 * (psfhat.shape[2] - 1)
    bases = [str(b) for b in g["bases"]]
    psi = Psi(nband, nx, ny, bases, int(g["nlevel"]), 1, dtype=dirty.dtype)
    dd = _dev.to_dev(dirty, copy=True)
    code = _dev.dcode(dd)
    plan = get_plan(_dev.to_dev(psfhat), nyp, tuple(dirty.shape), code)
    return psi, plan, dd, code, (nband, nx, ny), bases


def _host(t):
    return t.cpu().numpy() if hasattr(t, "cpu") else np.asarray(t)


@pmp("tag", ["solve_f64", "solve_f32"])
def test_pd_run(backend, tag):
    """pfb_pd_run == primal_dual_optimised (pfb/opt/primal_dual.py:90-171): golden model and
    dual after 10 iterations, positivity 1 and 2, and the tolerance exit."""
    import torch
    from pfb_b200 import _dev, _lib
    g = golden(tag)
    psi, plan, dd, code, (nband, nx, ny), bases = _setup(g)
    f64 = dd.dtype == torch.float64
    tolx = 1e-11 if f64 else 1e-5
    L = _lib.lib()
    nbasis = len(bases)
    for pos in (1, 2):
        x = torch.zeros_like(dd)
        v = torch.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dtype=dd.dtype, device=dd.device)
        w = torch.ones(tuple(v.shape[1:]), dtype=dd.dtype, device=dd.device)
        nit, eps = ctypes.c_int(-1), ctypes.c_double(-1.0)
        _lib.check(L.pfb_pd_run(psi._plan(code), plan._h, _dev.ptr(x), _dev.ptr(v), _dev.ptr(dd), _dev.ptr(w),
                                nband, nx, ny, nbasis, float(g["lam"]), float(g["hessnorm"]), float(nbasis),
                                0.0, 1.0, 0.0, 10, pos, code, ctypes.byref(nit), ctypes.byref(eps),
                                _dev.stream()))
        assert nit.value == 10 and 0.0 < eps.value < 10.0
        assert rel(_host(x), g[f"pdo_x_pos{pos}"]) < tolx
        assert rel(_host(v), g[f"pdo_v_pos{pos}"]) < tolx
    # tolerance exit: same iterate as the Python loop stopped by the same tolerance
    from functools import partial
    from pfb_b200.operators.psf import psf_convolve_cube
    from pfb_b200.opt.primal_dual import primal_dual_optimised, Grad21
    nyp = 2 * (g["psfhat"].shape[2] - 1)
    tol = 3.0 * eps.value
    xr = torch.zeros_like(dd)
    vr = torch.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dtype=dd.dtype, device=dd.device)
    conv = partial(psf_convolve_cube, None, None, torch.empty_like(dd), _dev.to_dev(g["psfhat"]), nyp)
    primal_dual_optimised(xr, vr, float(g["lam"]), psi.hdot, psi.dot, float(g["hessnorm"]), None,
                          torch.ones(tuple(vr.shape[1:]), dtype=dd.dtype, device=dd.device), None,
                          Grad21(conv, dd), nu=nbasis, tol=tol, maxit=40, positivity=1, verbosity=0)
    x = torch.zeros_like(dd)
    v = torch.zeros_like(vr)
    w = torch.ones(tuple(v.shape[1:]), dtype=dd.dtype, device=dd.device)
    _lib.check(L.pfb_pd_run(psi._plan(code), plan._h, _dev.ptr(x), _dev.ptr(v), _dev.ptr(dd), _dev.ptr(w),
                            nband, nx, ny, nbasis, float(g["lam"]), float(g["hessnorm"]), float(nbasis),
                            0.0, 1.0, tol, 40, 1, code, ctypes.byref(nit), ctypes.byref(eps), _dev.stream()))
    assert 1 <= nit.value < 40 and eps.value < tol
    assert np.array_equal(_host(x), _host(xr)) and np.array_equal(_host(v), _host(vr))


@pmp("tag", ["solve_f64", "solve_f32"])
def test_power_method_and_pcg_run(backend, tag):
    """pfb_power_method_run (pfb/opt/power_method.py:11-46) and pfb_pcg_run
    (pfb/opt/pcg.py:58-131, A = hessian_psf_cube) against the reference's results."""
    import torch
    from pfb_b200 import _dev, _lib
    g = golden(tag)
    psi, plan, dd, code, (nband, nx, ny), _ = _setup(g)
    f64 = dd.dtype == torch.float64
    L = _lib.lib()
    b = _dev.to_dev(g["pm_b0"].astype(g["dirty"].dtype), copy=True)
    beta, nit = ctypes.c_double(0.0), ctypes.c_int(-1)
    _lib.check(L.pfb_power_method_run(plan._h, _dev.ptr(b), nband, nx, ny, 1e-3, 40, code, ctypes.byref(beta),
                                      ctypes.byref(nit), _dev.stream()))
    assert abs(beta.value / float(g["pm_beta"]) - 1) < (1e-12 if f64 else 1e-5)
    assert rel(_host(b), g["pm_b"]) < (1e-10 if f64 else 1e-5)
    assert 1 <= nit.value <= 40
    # pcg(hessian_psf_cube(sigmainv=1e-3), dirty, x0=0, M=None, tol=0, maxit=10, minit=10)
    x = torch.zeros_like(dd)
    eps = ctypes.c_double(-1.0)
    _lib.check(L.pfb_pcg_run(plan._h, None, 1e-3, 0, 0.0, _dev.ptr(dd), _dev.ptr(x), 1.0, nband, nx, ny,
                             0.0, 10, 10, 1, code, ctypes.byref(nit), ctypes.byref(eps), _dev.stream()))
    assert nit.value == 10
    assert rel(_host(x), g["pcg_x"]) < (1e-10 if f64 else 2e-5)
    # zero right-hand side: "Initial residual is zero"
    x, zb = torch.zeros_like(dd), torch.zeros_like(dd)
    _lib.check(L.pfb_pcg_run(plan._h, None, 1e-3, 0, 0.0, _dev.ptr(zb), _dev.ptr(x), 1.0,
                             nband, nx, ny, 1e-5, 10, 1, 1, code, ctypes.byref(nit), ctypes.byref(eps),
                             _dev.stream()))
    assert nit.value == 0 and not _host(x).any()


def test_solver_run_errors(backend):
    from pfb_b200 import _dev, _lib
    g = golden("solve_f64")
    psi, plan, dd, code, (nband, nx, ny), bases = _setup(g)
    L = _lib.lib()
    rc = L.pfb_pd_run(psi._plan(code), plan._h, None, None, None, None, nband, nx, ny, len(bases), 1.0, 1.0,
                      1.0, 0.0, 1.0, 0.0, 1, 1, code, None, None, _dev.stream())
    assert rc == 1 and b"pfb_pd_run" in L.pfb_last_error()
    assert L.pfb_pcg_run(plan._h, None, 0.0, 0, 0.0, None, None, 1.0, nband, nx, ny, 0.0, 1, 1, 1, code, None,
                         None, _dev.stream()) == 1
    assert L.pfb_power_method_run(None, None, nband, nx, ny, 1e-3, 1, code, None, None, _dev.stream()) == 1
