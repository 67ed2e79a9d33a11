"""Worker glue between two minor cycles against golden vectors made by
the unmodified reference operators (tests/golden/make_golden.py gen_glue): first the
oracle restatement (CPU), then the product on the emulator build and -- marked gpu -- on
libpfb_b200.so through the C-ABI.  Tolerances: 1e-6 (fp32) / 1e-11 (fp64) relative L2."""
from functools import partial

import numpy as np
import pytest

from conftest import golden
from oracle import ref

pmp = pytest.mark.parametrize
TOL = {"f32": 1e-6, "f64": 1e-11}


def rel(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)


@pmp("t", ["f32", "f64"])
def test_oracle_glue_vs_golden(t):
    g = golden("glue_" + t)
    dt = g["dirty"].dtype
    bases = [str(b) for b in g["bases"]]
    nband, nx, ny = g["dirty"].shape
    nyp = g["psf"].shape[2]
    assert rel(ref.psfhat_from_psf(g["psf"]).view(dt), g["psfhat"].view(dt)) < TOL[t]
    xpad, xhat = np.empty(g["psf"].shape, dt), np.empty(g["psfhat"].shape, g["psfhat"].dtype)
    conv = partial(ref.psf_convolve_cube, xpad, xhat, np.empty_like(g["dirty"]), g["psfhat"], nyp)
    assert rel(ref.make_data(g["residual"], g["model"], conv), g["data"]) < TOL[t]
    rms, rmax, mfs = ref.residual_stats(g["residual"])
    assert abs(rms / g["rms"] - 1) < TOL[t] and abs(rmax / g["rmax"] - 1) < TOL[t]
    psi = ref.Psi(nband, nx, ny, bases, int(g["nlevel"]))
    outvar = np.zeros((nband, len(bases), psi.Nymax, psi.Nxmax), dt)
    rc = ref.rms_comps(psi, g["residual"], g["wsums"].sum(), g["wsums"], float(g["pix_per_beam"]), outvar)
    assert rel(rc, g["rms_comps"]) < TOL[t]
    w = ref.l1reweight_func(psi.dot, outvar, 1.0, g["rms_comps"], g["model"], alpha=4)
    assert rel(w, g["l1weight"]) < 10 * TOL[t]
    assert abs(ref.model_eps(g["model"], g["modelp"]) / g["eps"] - 1) < TOL[t]


@pmp("t", ["f32", "f64"])
def test_glue_vs_golden(backend, t):
    from pfb_b200.operators.psf import psf_convolve_cube, psfhat_from_psf
    from pfb_b200.operators.psi import Psi
    from pfb_b200.workers import glue
    g = golden("glue_" + t)
    dt = g["dirty"].dtype
    bases = [str(b) for b in g["bases"]]
    nband, nx, ny = g["dirty"].shape
    nyp = g["psf"].shape[2]
    # PSFHAT on the device from the real-space PSF (gridder.py:712-714)
    ph = psfhat_from_psf(g["psf"])
    assert ph.dtype == g["psfhat"].dtype and ph.shape == g["psfhat"].shape
    assert rel(ph.view(dt), g["psfhat"].view(dt)) < TOL[t]
    conv = partial(psf_convolve_cube, None, None, np.empty_like(g["dirty"]), ph, nyp)
    assert rel(glue.make_data(g["residual"], g["model"], conv), g["data"]) < 2 * TOL[t]
    rms, rmax, mfs = glue.residual_stats(g["residual"])
    assert rel(mfs, g["residual_mfs"]) < TOL[t]
    assert abs(rms / g["rms"] - 1) < TOL[t] and abs(rmax / g["rmax"] - 1) < TOL[t]
    psi = Psi(nband, nx, ny, bases, int(g["nlevel"]), 1)
    outvar = np.zeros((nband, len(bases), psi.Nymax, psi.Nxmax), dt)
    rc = glue.rms_comps(psi, g["residual"], g["wsums"].sum(), g["wsums"], float(g["pix_per_beam"]), outvar)
    assert rc.shape == g["rms_comps"].shape
    assert rel(rc, g["rms_comps"]) < 2 * TOL[t]
    w = glue.l1reweight_func(psi.dot, outvar, 1.0, g["rms_comps"], g["model"], alpha=4)
    assert rel(w, g["l1weight"]) < 10 * TOL[t]
    assert abs(glue.model_eps(g["model"], g["modelp"]) / g["eps"] - 1) < TOL[t]
    best = glue.BestModel(g["model"], 2.0, 3.0)
    assert not best.update(g["modelp"], 2.5, 1.0) and best.rms == 2.0
    assert best.update(g["modelp"], 1.5, 1.0) and best.rmax == 1.0
    assert rel(best.model.cpu().numpy(), g["modelp"]) == 0


@pmp("shape", [(2, 96, 80), (1, 60, 100), (64, 64)])
def test_psfhat_from_psf_shapes(backend, shape):
    """Power-of-two-free paddings (mixed-radix kernels), rectangular and 2-D inputs."""
    from pfb_b200.operators.psf import psfhat_from_psf
    rng = np.random.default_rng(5)
    for dt, tol in ((np.float32, 1e-6), (np.float64, 1e-11)):
        psf = rng.standard_normal(shape).astype(dt)
        got = psfhat_from_psf(psf)
        exp = ref.psfhat_from_psf(psf if psf.ndim == 3 else psf[None])
        exp = exp if psf.ndim == 3 else exp[0]
        assert got.shape == exp.shape and got.dtype == exp.dtype
        assert rel(got.view(dt), exp.view(dt)) < tol


def test_psi_psib_attribute(backend):
    """Psi.psib (pfb/operators/psi.py:272-274): one bookkeeping object per band."""
    from pfb_b200.operators.psi import Psi
    psi = Psi(2, 64, 48, ["self", "db2"], 2, 1)
    opsi = ref.Psi(2, 64, 48, ["self", "db2"], 2)
    assert len(psi.psib) == 2
    for pb in psi.psib:
        assert (pb.nx, pb.ny, pb.nbasis, pb.nlevel) == (64, 48, 2, 2)
        assert (pb.Nxmax, pb.Nymax) == (opsi.Nxmax, opsi.Nymax)
        assert list(pb.bases) == ["self", "db2"]
        assert pb.tables(1)["Ntotx"] == psi.tables(1)["Ntotx"]


def test_plan_cache_tracks_psf_identity_and_contents(backend):
    """a numpy psfhat updated in place must not keep its stale device copy, and a
    second psfhat of the same shape must not re-bind the plan another caller still holds."""
    from pfb_b200.operators import psf as P
    rng = np.random.default_rng(2)
    nband, nx, ny, nxp, nyp = 1, 24, 20, 64, 64
    mk = lambda: np.fft.rfft2(rng.standard_normal((nband, nxp, nyp)), axes=(1, 2))   # noqa: E731
    pa, pb = mk(), mk()
    x = rng.standard_normal((nband, nx, ny))
    xpad, xhat = np.empty((nband, nxp, nyp)), np.empty(pa.shape, complex)

    def conv(ph):
        return P.psf_convolve_cube(None, None, np.empty_like(x), ph, nyp, x)

    def oconv(ph):
        return ref.psf_convolve_cube(xpad, xhat, np.empty_like(x), ph, nyp, x).copy()
    assert rel(conv(pa), oconv(pa)) < 1e-11
    plan_a = P.get_plan(pa, nyp, x.shape, 1)
    assert rel(conv(pb), oconv(pb)) < 1e-11                 # another PSF, same shape ...
    assert P.get_plan(pa, nyp, x.shape, 1) is plan_a        # ... did not take pa's plan away
    assert rel(conv(pa), oconv(pa)) < 1e-11
    pa *= 0.5                                               # in-place update between major cycles
    pa[0, 3, 4] += 1.0
    assert rel(conv(pa), oconv(pa)) < 1e-11
    assert P.get_plan(pa, nyp, x.shape, 1) is plan_a        # same array object: re-bound, not duplicated
    assert len(next(iter(P._PLANS.values()))) == 2
