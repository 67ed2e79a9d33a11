"""Full-size parity against the oracle ("one full-size single-band operator
check"): the operators at BASELINE config 3's band size, the dual update
on config 3's whole coefficient cube, and a 10-iteration minor cycle at config 2 (4 x 2048^2)
in both precisions.  Tolerances are the north-star's: operators 1e-6 (fp32) / 1e-11 (fp64)
relative L2, final model 1e-5."""
from functools import partial

import numpy as np
import pytest
import torch

from oracle import ref

pytestmark = pytest.mark.gpu
pmp = pytest.mark.parametrize
BASES = ["self", "db1", "db2", "db3", "db4"]
TOL = {np.float32: 1e-6, np.float64: 1e-11}


def rel(a, b):
    a = a.cpu().numpy() if isinstance(a, torch.Tensor) else a
    b = b.cpu().numpy() if isinstance(b, torch.Tensor) else b
    num = np.linalg.norm((a.astype(np.float64) - b).ravel())
    return num / max(np.linalg.norm(b.ravel().astype(np.float64)), 1e-300)


@pytest.fixture(autouse=True)
def _cuda_lib():
    import os
    import pfb_b200._lib as L
    import pfb_b200.operators.psf as psf
    assert torch.cuda.is_available()
    L.lib()
    ref.lib().orc_set_num_threads(os.cpu_count() or 1)
    psf.clear_plan_cache()
    yield
    psf.clear_plan_cache()
    torch.cuda.empty_cache()


@pmp("dtype", [np.float32, np.float64])
def test_psi_4096_one_band_vs_oracle(dtype):
    """Psi.dot / Psi.hdot on one 4096^2 band (config 3), self+db1..db4, 3 levels."""
    from pfb_b200.operators.psi import Psi
    rng = np.random.default_rng(12)
    nx = ny = 4096
    x = rng.standard_normal((1, nx, ny), dtype=np.float32).astype(dtype)
    psi = Psi(1, nx, ny, BASES, 3, 1)
    opsi = ref.Psi(1, nx, ny, BASES, 3)
    a = np.zeros((1, 5, psi.Nymax, psi.Nxmax), dtype)
    oa = np.zeros_like(a)
    psi.dot(x, a)
    opsi.dot(x, oa)
    assert rel(a, oa) < TOL[dtype]
    coef = rng.standard_normal(a.shape, dtype=np.float32).astype(dtype)
    xr, oxr = np.empty_like(x), np.empty_like(x)
    psi.hdot(coef, xr)
    opsi.hdot(coef, oxr)
    assert rel(xr, oxr) < TOL[dtype]


def test_dual_update_config3_cube_vs_oracle():
    """dual_update_numba on config 3's (8, 5, 4116, 4116) fp32 cube."""
    from pfb_b200.prox.prox_21m import dual_update_numba
    rng = np.random.default_rng(13)
    shape = (8, 5, 4116, 4116)
    v = rng.standard_normal(shape, dtype=np.float32)
    vp = rng.standard_normal(shape, dtype=np.float32)
    w = rng.random(shape[1:], dtype=np.float32)
    lam, sigma = 0.8, 1.7
    vd = torch.from_numpy(v).cuda()
    dual_update_numba(torch.from_numpy(vp).cuda(), vd, lam, sigma=sigma, weight=torch.from_numpy(w).cuda())
    ref.dual_update_numba(vp, v, lam, sigma=sigma, weight=w)       # in place on v
    got = vd.cpu().numpy()
    del vd
    # band by band to bound the fp64 temporaries
    num = sum(float(np.sum((got[b].astype(np.float64) - v[b]) ** 2)) for b in range(8))
    den = sum(float(np.sum(v[b].astype(np.float64) ** 2)) for b in range(8))
    assert np.sqrt(num / den) < 1e-6


@pmp("dtype", [np.float32, np.float64])
@pmp("nx", [2048, 4096])
def test_psf_convolve_random_full_size_vs_oracle(dtype, nx):
    """Random-input psf_convolve_cube, one band: 2048^2 / PSF 4096^2 and 4096^2 / PSF 8192^2."""
    from pfb_b200.operators.psf import psf_convolve_cube
    from pfb_b200.operators.hessian import hessian_psf_cube
    import os
    rng = np.random.default_rng(14)
    nxp = 2 * nx
    cdt = np.complex64 if dtype == np.float32 else np.complex128
    nthr = os.cpu_count() or 1
    # psfhat of a random REAL psf (what gridder.py:712-714 produces): Hermitian half spectrum
    psf = rng.standard_normal((1, nxp, nxp), dtype=np.float32)
    psfhat = ref.sfft.rfft2(psf.astype(np.float64), axes=(1, 2), workers=nthr).astype(cdt)
    del psf
    x = rng.standard_normal((1, nx, nx), dtype=np.float32).astype(dtype)
    xpad, xhat = np.empty((1, nxp, nxp), dtype), np.empty(psfhat.shape, cdt)
    xo = np.empty_like(x)
    exp = ref.psf_convolve_cube(xpad, xhat, xo, psfhat, nxp, x, nthreads=nthr).copy()
    out = np.full(x.shape, np.nan, dtype)
    psf_convolve_cube(None, None, out, psfhat, nxp, x)
    assert rel(out, exp) < TOL[dtype]
    beam = (0.5 + rng.random(x.shape, dtype=np.float32)).astype(dtype)
    exp2 = ref.hessian_psf_cube(xpad, xhat, xo, beam, psfhat, nxp, x, nthreads=nthr, sigmainv=0.25, wsum=3.5)
    got2 = hessian_psf_cube(None, None, None, beam, psfhat, nxp, x, sigmainv=0.25, wsum=3.5)
    assert rel(got2, exp2) < TOL[dtype]


@pmp("dtype", [np.float32, np.float64])
def test_minor_cycle_config2_vs_oracle(dtype):
    """BASELINE config 2 (4 bands 2048^2, PSF 4096^2, self+db1..db4, 3 levels): final model and
    dual after 10 primal-dual iterations against the oracle, both precisions."""
    import os
    from pfb_b200.operators.psf import psf_convolve_cube
    from pfb_b200.operators.psi import Psi
    from pfb_b200.opt.primal_dual import primal_dual_optimised, Grad21
    nthr = os.cpu_count() or 1
    nband, nx = 4, 2048
    nxp = 2 * nx
    rng = np.random.default_rng(420)
    cdt = np.complex64 if dtype == np.float32 else np.complex128
    ii = (np.arange(nxp) - nxp // 2).astype(np.float64)
    psfhat = np.empty((nband, nxp, nxp // 2 + 1), cdt)
    for b in range(nband):
        s = 2.0 * (1 + 0.05 * b)
        g1 = np.exp(-0.5 * (ii / s) ** 2)
        psf = g1[:, None] * g1[None, :]
        g2 = np.exp(-0.5 * (ii / (0.05 * nx)) ** 2)
        psf += 0.01 * np.cos(0.05 * ii)[:, None] * np.cos(0.03 * ii)[None, :] * g2[:, None] * g2[None, :]
        psf /= psf.max() * nband
        psfhat[b] = ref.sfft.rfft2(np.fft.ifftshift(psf), workers=nthr).astype(cdt)
    del psf
    model = np.zeros((nband, nx, nx), dtype)
    for _ in range(60):
        i, j = rng.integers(nx // 8, 7 * nx // 8, 2)
        model[:, i, j] = (1 + np.exp(rng.standard_normal())) * (1 + 0.1 * np.arange(nband)) ** -0.7
    xpad, xhat = np.empty((nband, nxp, nxp), dtype), np.empty(psfhat.shape, cdt)
    xo = np.empty_like(model)
    oconv = partial(ref.psf_convolve_cube, xpad, xhat, xo, psfhat, nxp, nthreads=nthr)
    dirty = oconv(model).copy()
    dirty += (1e-3 * dirty.max() * rng.standard_normal(dirty.shape)).astype(dtype)
    psi = Psi(nband, nx, nx, BASES, 3, 1)
    opsi = ref.Psi(nband, nx, nx, BASES, 3, nthreads=nthr)
    conv = partial(psf_convolve_cube, None, None, np.empty_like(dirty), psfhat, nxp)
    hessnorm = 1.05
    lam = 0.5 * float(np.std(dirty.sum(axis=0)))
    shape_v = (nband, 5, psi.Nymax, psi.Nxmax)
    w = np.ones(shape_v[1:], dtype)
    x, v = np.zeros_like(dirty), np.zeros(shape_v, dtype)
    primal_dual_optimised(x, v, lam, psi.hdot, psi.dot, hessnorm, None, w, None, Grad21(conv, dirty),
                          nu=5, tol=0.0, maxit=10, verbosity=0)
    ox, ov = np.zeros_like(dirty), np.zeros(shape_v, dtype)
    ref.primal_dual_optimised(ox, ov, lam, opsi.hdot, opsi.dot, hessnorm, None, w, None,
                              lambda z: oconv(z) - dirty, nu=5, tol=0.0, maxit=10)
    assert ox.any()
    assert rel(x, ox) < (1e-11 if dtype == np.float64 else 1e-5)
    # The dual is held to the model's bar in fp64.  In fp32 the GPU stencils accumulate in fp32
    # (allowed by the operator bar: Psi.dot is within 1e-6 of the oracle relative to the WHOLE
    # coefficient cube) while the reference accumulates in fp64, and for this sparse,
    # smooth model most detail coefficients are orders of magnitude below the image scale, so the
    # same absolute error is a larger fraction of ||v|| (measured 5.9e-4) than of ||x||.
    ev = rel(v, ov)
    print(f"config 2 {np.dtype(dtype).name}: rel x {rel(x, ox):.3e}  rel v {ev:.3e}")
    assert ev < (1e-11 if dtype == np.float64 else 2e-3)
