"""Device-resident PCG (csrc/pcg.cu, opt/pcg.py): the late-read / speculative-A(p) loop, the
backtracking branch (pfb/opt/pcg.py:96-101), an operator that returns a shared output buffer
(psf_convolve's xout, pfb/operators/psf.py:54) and the batched per-band solve with per-band
termination (pfb/opt/pcg.py:243-291), against the oracle on the golden problem.  Runs on the
emulator build and -- marked gpu -- on libpfb_b200.so."""
from functools import partial

import numpy as np
import pytest

from conftest import golden
from oracle import ref

pmp = pytest.mark.parametrize


def rel(a, b):
    return np.linalg.norm((np.asarray(a, np.float64) - b).ravel()) / np.linalg.norm(np.asarray(b, np.float64).ravel())


def _np(x):
    return x.cpu().numpy() if hasattr(x, "cpu") else np.asarray(x)


@pmp("t", ["f64", "f32"])
def test_pcg_shared_output_and_backtracking(backend, t):
    from pfb_b200.opt.pcg import pcg, ScalarPrecond
    from pfb_b200.operators.psf import psf_convolve_cube
    g = golden("solve_" + t)
    psfhat, dirty = g["psfhat"], g["dirty"]
    dt = dirty.dtype
    nyp = 2 * (psfhat.shape[2] - 1)
    tol = 1e-11 if t == "f64" else 2e-4      # CG feeds fp32 rounding back through alpha every iteration
    xpad, xhat = np.empty((dirty.shape[0], psfhat.shape[1], nyp), dt), np.empty(psfhat.shape, psfhat.dtype)
    xo = np.empty_like(dirty)
    # (a) A returns the same buffer on every call
    xout = np.empty_like(dirty)
    xs = pcg(lambda x: psf_convolve_cube(None, None, xout, psfhat, nyp, x), dirty, np.zeros_like(dirty),
             tol=0.0, maxit=12, minit=12, verbosity=0)
    assert pcg.last_path == "device-scalars"
    oxs = ref.pcg(lambda x: ref.psf_convolve_cube(xpad, xhat, xo, psfhat, nyp, x).copy(), dirty,
                  np.zeros_like(dirty), tol=0.0, maxit=12, minit=12)
    assert rel(xs, oxs) < tol
    # (b) a non-symmetric operator: <r,y> grows on some steps and the line search backtracks
    bm = (0.2 + 1.6 * np.random.default_rng(3).random(dirty.shape)).astype(dt)

    def A(x):
        xn = _np(x)
        return _np(psf_convolve_cube(None, None, np.empty_like(dirty), psfhat, nyp, bm * xn)) + dt.type(1e-3) * xn

    def oA(x):
        return ref.psf_convolve_cube(xpad, xhat, xo, psfhat, nyp, bm * x).copy() + dt.type(1e-3) * x
    for M, oM in ((None, None), (ScalarPrecond(0.05), lambda v: v / dt.type(0.05))):
        o_bt = ref.pcg(oA, dirty, np.zeros_like(dirty), M=oM, tol=0.0, maxit=12, minit=12, backtrack=True)
        o_nb = ref.pcg(oA, dirty, np.zeros_like(dirty), M=oM, tol=0.0, maxit=12, minit=12, backtrack=False)
        assert rel(o_bt, o_nb) > 0.05            # the branch really fires on this problem
        x_bt = pcg(A, dirty, np.zeros_like(dirty), M=M, tol=0.0, maxit=12, minit=12, verbosity=0)
        x_nb = pcg(A, dirty, np.zeros_like(dirty), M=M, tol=0.0, maxit=12, minit=12, verbosity=0, backtrack=False)
        if t == "f64":
            assert rel(x_bt, o_bt) < tol and rel(x_nb, o_nb) < tol
        else:       # fp32: a backtrack decision can flip on the last bit; hold the un-backtracked run
            assert rel(x_nb, o_nb) < 5e-3


def test_pcg_psf_bands_stop_independently(backend):
    """Per-band tolerance: each band leaves the batched solve at its own iteration, like the
    reference's independent per-band calls."""
    from pfb_b200.opt.pcg import _pcg_psf_impl, pcg_psf
    g = golden("solve_f64")
    psfhat, dirty, beam = g["psfhat"].copy(), g["dirty"], g["beam"]
    psfhat[1] *= 0.3                              # bands of different conditioning
    nyp = 2 * (psfhat.shape[2] - 1)
    exp = np.zeros_like(dirty)
    its = []
    for b in range(dirty.shape[0]):
        xpad, xhat = np.empty((psfhat.shape[1], nyp)), np.empty(psfhat.shape[1:], complex)
        A = partial(ref._hessian_psf_slice, xpad, xhat, np.empty_like(dirty[b]), psfhat[b], beam[b], nyp,
                    sigmainv=0.05)
        cnt = [0]

        def Ac(x, A=A, cnt=cnt):
            cnt[0] += 1
            return A(x)
        exp[b] = ref.pcg(Ac, dirty[b], np.zeros_like(dirty[b]), M=lambda v: v / 0.05, tol=3e-2, maxit=40, minit=2)
        its.append(cnt[0])
    assert its[0] != its[1], its                 # the bands do stop at different iterations
    got = _pcg_psf_impl(psfhat, dirty, np.zeros_like(dirty), beam, nyp, 1, 0.05, tol=3e-2, maxit=40, minit=2,
                        verbosity=0)
    assert rel(got, exp) < 1e-11
    got2 = pcg_psf(psfhat, dirty, np.zeros_like(dirty), beam, nyp, 1, 0.05,
                   dict(tol=3e-2, maxit=40, minit=2, verbosity=0))
    assert rel(got2, exp) < 1e-11
