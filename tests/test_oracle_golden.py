"""Pin the CPU oracle (oracle/ref.py + oracle/pfb_oracle.c) against golden
vectors produced by the unmodified reference (tests/golden/make_golden.py)."""
from functools import partial

import numpy as np
import pytest

from oracle import ref
from conftest import golden

pmp = pytest.mark.parametrize
BASES6 = ['self', 'db1', 'db2', 'db3', 'db4', 'db5']


def rel(a, b):
    return np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)


@pmp("tag", ["psi_f64_a", "psi_f64_b", "psi_f64_c", "psi_f32_a"])
def test_psi(tag):
    g = golden(tag)
    x = g["x"]
    dt = x.dtype
    nband, nx, ny = x.shape
    nlevel = int(g["nlevel"])
    psi = ref.Psi(nband, nx, ny, BASES6, nlevel, 1)
    assert (psi.Nxmax, psi.Nymax) == (int(g["Nxmax"]), int(g["Nymax"]))
    rng = np.random.default_rng(420)
    x2 = rng.standard_normal(x.shape).astype(dt)
    assert np.array_equal(x2, x)
    fill = rng.standard_normal(g["alpha"].shape).astype(dt)
    coef_in = rng.standard_normal(fill.shape).astype(dt)
    alpha = fill.copy()
    psi.dot(x, alpha)
    tol = 1e-14 if dt == np.float64 else 2e-7
    assert rel(alpha, g["alpha"]) < tol
    # never-written cells keep the pre-fill exactly
    untouched = g["alpha"] == fill
    assert np.array_equal(alpha[untouched], fill[untouched])
    xrec = np.full(x.shape, np.nan, dt)
    psi.hdot(coef_in, xrec)
    assert rel(xrec, g["xrec"]) < tol


@pmp("tag", ["prox_f64", "prox_f32", "prox_f64_1band"])
def test_prox(tag):
    g = golden(tag)
    v, vp, w = g["v"], g["vp"], g["w"]
    tol = 1e-14 if v.dtype == np.float64 else 3e-7
    for i in range(4):
        lam, sigma = g[f"params_{i}"]
        res = np.full(v.shape, np.nan, v.dtype)
        ref.prox_21m_numba(v, res, lam, sigma=sigma, weight=w)
        assert rel(res, g[f"prox21m_{i}"]) < tol
        v2 = v.copy()
        ref.dual_update_numba(vp, v2, lam, sigma=sigma, weight=w)
        assert rel(v2, g[f"dual21m_{i}"]) < tol
        res = np.full(v.shape, np.nan, v.dtype)
        ref.prox_21_numba(v, res, lam, sigma=sigma, weight=w)
        assert rel(res, g[f"prox21_{i}"]) < tol
        v2 = v.copy()
        ref.dual_update_numba_l2(vp, v2, lam, sigma=sigma, weight=w)
        assert rel(v2, g[f"dual21_{i}"]) < tol
    assert rel(ref.prox_21m(v, 0.3, weight=w), g["prox21m_numpy"]) < tol
    assert rel(ref.prox_21(v, 0.3, weight=w), g["prox21_numpy"]) < tol


@pmp("tag", ["hess_f64", "hess_f32", "hess_f64_rect"])
def test_hess(tag):
    g = golden(tag)
    x, psfhat, beam = g["x"], g["psfhat"], g["beam"]
    nband, nx, ny = x.shape
    nxp = psfhat.shape[1]
    nyp = g["psf"].shape[2]
    xpad = np.empty((nband, nxp, nyp), x.dtype)
    xhat = np.empty(psfhat.shape, psfhat.dtype)
    xout = np.empty(x.shape, x.dtype)
    tol = 1e-13 if x.dtype == np.float64 else 5e-7
    assert rel(ref.psf_convolve_cube(xpad, xhat, xout, psfhat, nyp, x), g["conv"]) < tol
    assert rel(ref.psf_convolve_slice(xpad[0], xhat[0], xout[0], psfhat[0], nyp, x[0]),
               g["conv0"]) < tol
    assert rel(ref.hessian_psf_cube(xpad, xhat, xout, None, psfhat, nyp, x,
                                    sigmainv=1e-3), g["h1"]) < tol
    assert rel(ref.hessian_psf_cube(xpad, xhat, xout, beam, psfhat, nyp, x,
                                    sigmainv=0.25, wsum=3.5), g["h2"]) < tol
    assert rel(ref._hessian_psf_slice(xpad[0], xhat[0], xout[0], psfhat[0], beam[0],
                                      nyp, x[0], sigmainv=0.5, wsum=2.0), g["h3"]) < tol


def test_hess_direct_sum_kat():
    """Known-answer test independent of any FFT: direct O(N^4) summation
    out[i,j] = sum_pq x[p,q] psf[nxp//2+i-p, nyp//2+j-q]."""
    rng = np.random.default_rng(5)
    nx, ny, nxp, nyp = 8, 6, 16, 12
    x = rng.standard_normal((1, nx, ny))
    psf = rng.standard_normal((1, nxp, nyp))
    psfhat = np.fft.rfft2(np.fft.ifftshift(psf, axes=(1, 2)), axes=(1, 2))
    out = np.zeros((nx, ny))
    for i in range(nx):
        for j in range(ny):
            for p in range(nx):
                for q in range(ny):
                    out[i, j] += x[0, p, q] * psf[0, nxp // 2 + i - p, nyp // 2 + j - q]
    xpad = np.empty((1, nxp, nyp))
    xhat = np.empty(psfhat.shape, psfhat.dtype)
    xout = np.empty(x.shape)
    assert rel(ref.psf_convolve_cube(xpad, xhat, xout, psfhat, nyp, x)[0], out) < 1e-13


def test_norm_diff():
    g = golden("misc")
    for t, tol in (("f64", 1e-14), ("f32", 1e-6)):
        assert abs(ref.norm_diff(g[f"x_{t}"], g[f"xp_{t}"]) / g[f"nd3_{t}"] - 1) < tol
        assert abs(ref.norm_diff(g[f"x_{t}"][0], g[f"xp_{t}"][0]) / g[f"nd2_{t}"] - 1) < tol


def _solve_setup(g):
    psfhat, dirty = g["psfhat"], g["dirty"]
    dt = dirty.dtype
    nband, nx, ny = dirty.shape
    nxp = psfhat.shape[1]
    nyp = 2 * (psfhat.shape[2] - 1)
    xpad = np.empty((nband, nxp, nyp), dt)
    xhat = np.empty(psfhat.shape, psfhat.dtype)
    xout = np.empty(dirty.shape, dt)
    conv = partial(ref.psf_convolve_cube, xpad, xhat, xout, psfhat, nyp)
    bases = [str(b) for b in g["bases"]]
    psi = ref.Psi(nband, nx, ny, bases, int(g["nlevel"]), 1)
    return conv, psi, (xpad, xhat, xout), nyp


@pmp("tag", ["solve_f64", "solve_f32"])
def test_solvers(tag):
    g = golden(tag)
    conv, psi, (xpad, xhat, xout), nyp = _solve_setup(g)
    dirty, psfhat = g["dirty"], g["psfhat"]
    dt = dirty.dtype
    nband, nx, ny = dirty.shape
    nbasis = psi.nbasis
    f64 = dt == np.float64
    beta, b = ref.power_method(conv, dirty.shape, b0=g["pm_b0"].copy(), tol=1e-3, maxit=40)
    assert abs(beta / g["pm_beta"] - 1) < (1e-12 if f64 else 1e-5)
    assert rel(b, g["pm_b"]) < (1e-11 if f64 else 1e-5)
    hessnorm, lam = float(g["hessnorm"]), float(g["lam"])
    data = dirty.copy()

    def grad21(x):
        return conv(x) - data
    tolx = 1e-11 if f64 else 1e-5
    for pos in (1, 2):
        x = np.zeros(dirty.shape, dt)
        v = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dt)
        w = np.ones(v.shape[1:], dt)
        x, v = ref.primal_dual_optimised(x, v, lam, psi.hdot, psi.dot, hessnorm, None,
                                         w, None, grad21, nu=nbasis, tol=0.0, maxit=10,
                                         positivity=pos)
        assert rel(x, g[f"pdo_x_pos{pos}"]) < tolx
        assert rel(v, g[f"pdo_v_pos{pos}"]) < tolx
    # reweighting branch
    outvar = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dt)
    rew = partial(ref.l1reweight_func, psi.dot, outvar, 1.0, g["rms_comps"])
    x = np.zeros(dirty.shape, dt)
    v = np.zeros(outvar.shape, dt)
    w = np.ones(v.shape[1:], dt)
    x, v = ref.primal_dual_optimised(x, v, lam, psi.hdot, psi.dot, hessnorm, None, w, rew,
                                     grad21, nu=nbasis, tol=0.2, maxit=12, positivity=1,
                                     maxreweight=3)
    assert rel(x, g["pdo_rw_x"]) < tolx
    assert rel(v, g["pdo_rw_v"]) < tolx
    assert rel(ref.l1reweight_func(psi.dot, outvar, 1.0, g["rms_comps"], x),
               g["rw_weight"]) < tolx

    def psiH_f(xx):
        a = np.zeros(outvar.shape, dt)
        psi.dot(xx, a)
        return a

    def psi_f(a):
        xx = np.zeros(dirty.shape, dt)
        psi.hdot(a, xx)
        return xx
    x, v = ref.primal_dual(np.zeros(dirty.shape, dt), np.zeros(outvar.shape, dt), lam,
                           psi_f, psiH_f, hessnorm, partial(ref.prox_21m, weight=1.0),
                           lambda z: grad21(z).copy(), nu=nbasis, tol=0.0, maxit=10,
                           minit=10, positivity=1)
    assert rel(x, g["pd_x"]) < tolx
    assert rel(v, g["pd_v"]) < tolx
    A = partial(ref.hessian_psf_cube, xpad, xhat, xout, None, psfhat, nyp,
                sigmainv=1e-3, wsum=None)
    xs, rs = ref.pcg(A, dirty, np.zeros_like(dirty), M=None, tol=0.0, maxit=10, minit=10,
                     backtrack=True, return_resid=True)
    # CG amplifies rounding differences between FFT backends in fp32
    assert rel(xs, g["pcg_x"]) < (1e-10 if f64 else 2e-4)
    xb = ref._pcg_psf_impl(psfhat, dirty, np.zeros_like(dirty), g["beam"], nyp, 1, 0.05,
                           tol=1e-4, maxit=25, minit=3)
    assert rel(xb, g["pcgpsf_x"]) < (1e-10 if f64 else 2e-4)
