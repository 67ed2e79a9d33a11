"""numpy / C restatement of the reference minor-cycle path (TEST INFRASTRUCTURE
ONLY -- see oracle/__init__.py; parity pinned by tests/test_oracle_golden.py).

Same names, argument order and mutation semantics as the reference so that
parity tests read like the reference's own tests.  Paths cited are relative to
/root/reference.  Heavy loops live in oracle/pfb_oracle.c (built on demand
with gcc); FFTs use scipy.fft (pocketfft), standing in for the un-vendored
ducc0.fft.r2c/c2r the reference calls (pfb/operators/psf.py:47-53).
"""
import ctypes
import os
import subprocess

import numpy as np
import scipy.fft as sfft

from . import filters

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    """Compile oracle/pfb_oracle.c -> oracle/libpfb_oracle.so (gcc, OpenMP)."""
    src = os.path.join(_HERE, "pfb_oracle.c")
    so = os.path.join(_HERE, "libpfb_oracle.so")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-ffp-contract=off",
                               "-fPIC", "-shared", "-o", so, src, "-lm"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = ctypes.CDLL(build())
        L.orc_norm_diff.restype = ctypes.c_double
        L.orc_num_threads.restype = ctypes.c_int
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _is32(a):
    return a.dtype == np.float32


# ---------------------------------------------------------------- wavelets
def coeff_size(nsignal, nfilter):
    """pfb/wavelets/wavelets.py:20-22"""
    return (nsignal + nfilter - 1) // 2


def signal_size(ncoeff, nfilter):
    """pfb/wavelets/wavelets.py:25-27"""
    return 2 * ncoeff - nfilter + 2


def pyramid(nx, ny, wavelet, nlevel):
    """Bookkeeping of pfb/operators/psi.py:49-94 for one wavelet.

    Returns dict with sx, sy, spx, spy (lists per level), ix, iy (dict level ->
    (lo, hi)), Ntotx, Ntoty, F."""
    max_level = filters.dwt_max_level(min(nx, ny), wavelet)
    if nlevel > max_level:
        raise ValueError(f"The requested decomposition level {nlevel} "
                         "is not possible")
    F = int(wavelet[-1]) * 2
    sx, sy, spx, spy = [], [], [], []
    Nx, Ny, Nxm, Nym = nx, ny, 0, 0
    for _ in range(nlevel):
        Cx, Cy = coeff_size(Nx, F), coeff_size(Ny, F)
        Nxm += Cx
        Nym += Cy
        sx.append(Cx)
        sy.append(Cy)
        Nx, Ny = Cx + Cx % 2, Cy + Cy % 2
        spx.append(signal_size(Cx, F))
        spy.append(signal_size(Cy, F))
    Nxm += sx[-1]
    Nym += sy[-1]
    ix, iy = {}, {}
    lowx, lowy = sx[-1], sy[-1]
    ix[nlevel - 1] = (lowx, 2 * lowx)
    iy[nlevel - 1] = (lowy, 2 * lowy)
    lowx *= 2
    lowy *= 2
    for k in reversed(range(nlevel - 1)):
        ix[k] = (lowx, lowx + sx[k])
        iy[k] = (lowy, lowy + sy[k])
        lowx += sx[k]
        lowy += sy[k]
    return dict(F=F, sx=sx, sy=sy, spx=spx, spy=spy, ix=ix, iy=iy,
                Ntotx=Nxm, Ntoty=Nym)


class Psi(object):
    """pfb/operators/psi.py:269-310 (dot = analysis, hdot = synthesis)."""

    def __init__(self, nband, nx, ny, bases, nlevel, nthreads=1):
        self.nband, self.nx, self.ny = nband, nx, ny
        self.bases = list(bases)
        self.nbasis = len(bases)
        self.nlevel = nlevel
        self.nthreads = nthreads
        self.pyr = {}
        Nxmax = Nymax = 0
        nb = self.nbasis
        self._F = np.zeros(nb, np.int32)
        self._taps = np.zeros((nb, 4, 10), np.float64)
        self._tabs = np.zeros((nb, 6, nlevel), np.int32)
        self._Ntotx = np.zeros(nb, np.int32)
        self._Ntoty = np.zeros(nb, np.int32)
        for i, w in enumerate(self.bases):
            if w == "self":
                continue
            p = pyramid(nx, ny, w, nlevel)
            self.pyr[w] = p
            Nxmax, Nymax = max(Nxmax, p["Ntotx"]), max(Nymax, p["Ntoty"])
            self._F[i] = p["F"]
            for k, f in enumerate(filters.filter_bank(w)):
                self._taps[i, k, :p["F"]] = f
            self._tabs[i, 0] = [p["ix"][l][1] for l in range(nlevel)]
            self._tabs[i, 1] = [p["iy"][l][1] for l in range(nlevel)]
            self._tabs[i, 2], self._tabs[i, 3] = p["sx"], p["sy"]
            self._tabs[i, 4], self._tabs[i, 5] = p["spx"], p["spy"]
            self._Ntotx[i], self._Ntoty[i] = p["Ntotx"], p["Ntoty"]
        # the reference leaves Nxmax = Nymax = 0 when bases == ['self'] and
        # then fails on the slice assignment; keep that (callers always add
        # at least one wavelet).
        self.Nxmax, self.Nymax = int(Nxmax), int(Nymax)

    def dot(self, x, alphao):
        """image (nband,nx,ny) -> coeffs (nband,nbasis,Nymax,Nxmax), in place."""
        r32 = int(_is32(alphao))
        for b in range(self.nband):
            xb = _f64(x[b])
            ab = _f64(alphao[b])
            lib().orc_psi_band_dot(_p(xb), self.nx, self.ny, _p(ab), self.nbasis,
                                   self.Nymax, self.Nxmax, _p(self._F),
                                   _p(self._taps), _p(self._tabs), self.nlevel, r32)
            alphao[b] = ab

    def hdot(self, alpha, xo):
        """coeffs -> image, overwriting xo."""
        r32 = int(_is32(xo))
        for b in range(self.nband):
            ab = _f64(alpha[b])
            xb = np.zeros((self.nx, self.ny), np.float64)
            lib().orc_psi_band_hdot(_p(ab), _p(xb), self.nx, self.ny, self.nbasis,
                                    self.Nymax, self.Nxmax, _p(self._F),
                                    _p(self._Ntotx), _p(self._Ntoty),
                                    _p(self._taps), _p(self._tabs), self.nlevel, r32)
            xo[b] = xb


# -------------------------------------------------------------------- prox
def prox_21m(v, sigma, weight=1.0, axis=0):
    """pfb/prox/prox_21m.py:5-27 (numpy form; returns a new array)."""
    l2_norm = np.sum(v, axis=axis)
    l2_soft = np.maximum(np.abs(l2_norm) - sigma * weight, 0.0) * np.sign(l2_norm)
    mask = l2_norm != 0
    ratio = np.zeros(mask.shape, dtype=v.dtype)
    ratio[mask] = l2_soft[mask] / l2_norm[mask]
    return v * np.expand_dims(ratio, axis=axis)


def prox_21(v, sigma, weight=None, axis=0):
    """pfb/prox/prox_21.py:5-22."""
    l2_norm = np.linalg.norm(v, axis=axis)
    l2_soft = np.maximum(l2_norm - sigma * weight, 0.0)
    mask = l2_norm != 0
    ratio = np.zeros(mask.shape, dtype=v.dtype)
    ratio[mask] = l2_soft[mask] / l2_norm[mask]
    return v * np.expand_dims(ratio, axis=axis)


def _prox_numba(v, result, lam, sigma, weight, mode):
    r32 = _is32(result)
    vv, rr = _f64(v), np.zeros(v.shape, np.float64)
    nband = v.shape[0]
    n = int(np.prod(v.shape[1:]))
    w = _f64(np.broadcast_to(weight, v.shape[1:]))
    lib().orc_prox_21(_p(vv), _p(rr), _p(w), ctypes.c_int(nband),
                      ctypes.c_long(n), ctypes.c_double(lam),
                      ctypes.c_double(sigma), ctypes.c_int(mode),
                      ctypes.c_int(int(r32)))
    result[...] = rr


def prox_21m_numba(v, result, lam, sigma=1.0, weight=None):
    """pfb/prox/prox_21m.py:30-61 (writes result)."""
    _prox_numba(v, result, lam, sigma, weight, 0)


def prox_21_numba(v, result, lam, sigma=1.0, weight=None):
    """pfb/prox/prox_21.py:25-52."""
    _prox_numba(v, result, lam, sigma, weight, 1)


def dual_update(v, x, psiH, lam, sigma=1.0, weight=1.0):
    """pfb/prox/prox_21m.py:64-71 (numpy form; psiH is image->coeffs)."""
    vp = v.copy()
    vout = np.zeros_like(v)
    psiH(x, vout)
    vtilde = vp + sigma * vout
    return vtilde - sigma * prox_21m(vtilde / sigma, lam / sigma, weight=weight)


def _dual_numba(vp, v, lam, sigma, weight, mode):
    r32 = _is32(v)
    a, b = _f64(vp), _f64(v)
    if b is v:
        b = v
    nband = v.shape[0]
    n = int(np.prod(v.shape[1:]))
    w = _f64(np.broadcast_to(weight, v.shape[1:]))
    lib().orc_dual_update(_p(a), _p(b), _p(w), ctypes.c_int(nband),
                          ctypes.c_long(n), ctypes.c_double(lam),
                          ctypes.c_double(sigma), ctypes.c_int(mode),
                          ctypes.c_int(int(r32)))
    if b is not v:
        v[...] = b


def dual_update_numba(vp, v, lam, sigma=1.0, weight=None):
    """pfb/prox/prox_21m.py:75-103 (in place on v)."""
    _dual_numba(vp, v, lam, sigma, weight, 0)


def dual_update_numba_l2(vp, v, lam, sigma=1.0, weight=None):
    """pfb/prox/prox_21.py:66-91 (4-D arrays accepted, flattened per band)."""
    _dual_numba(vp, v, lam, sigma, weight, 1)


def prox2(x, gamma):
    """pfb/prox/prox2.py:3"""
    return x / (1 + gamma)


def norm_diff(x, xp):
    """pfb/utils/misc.py:1316-1351."""
    r32 = int(_is32(x))
    a, b = _f64(x), _f64(xp)
    return float(lib().orc_norm_diff(_p(a), _p(b), ctypes.c_long(a.size),
                                     ctypes.c_int(r32)))


# --------------------------------------------------------------- operators
def psf_convolve_slice(xpad, xhat, xout, psfhat, lastsize, x, nthreads=1):
    """pfb/operators/psf.py:11-29."""
    nx, ny = x.shape
    xpad[...] = 0.0
    xpad[0:nx, 0:ny] = x
    xhat[...] = sfft.rfftn(xpad, axes=(0, 1), workers=nthreads)
    xhat *= psfhat
    xpad[...] = sfft.irfftn(xhat, s=(xpad.shape[0], lastsize), axes=(0, 1),
                            workers=nthreads)
    xout[...] = xpad[0:nx, 0:ny]
    return xout


def psf_convolve_cube(xpad, xhat, xout, psfhat, lastsize, x, nthreads=1):
    """pfb/operators/psf.py:32-56.  Returns the shared xout buffer."""
    _, nx, ny = x.shape
    xpad[...] = 0.0
    xpad[:, 0:nx, 0:ny] = x
    xhat[...] = sfft.rfftn(xpad, axes=(1, 2), workers=nthreads)
    xhat *= psfhat
    xpad[...] = sfft.irfftn(xhat, s=(xpad.shape[1], lastsize), axes=(1, 2),
                            workers=nthreads)
    xout[...] = xpad[:, 0:nx, 0:ny]
    return xout


def _hessian_psf_slice(xpad, xhat, xout, psfhat, beam, lastsize, x,
                       nthreads=1, sigmainv=1, wsum=None):
    """pfb/operators/hessian.py:129-158."""
    xin = x * beam if beam is not None else x
    psf_convolve_slice(xpad, xhat, xout, psfhat, lastsize, xin, nthreads=nthreads)
    if beam is not None:
        xout *= beam
    if wsum is not None:
        xout /= wsum
    return xout + x * sigmainv


def hessian_psf_cube(xpad, xhat, xout, beam, psfhat, lastsize, x,
                     nthreads=1, sigmainv=1, wsum=None):
    """pfb/operators/hessian.py:254-281 (note beam/psfhat order)."""
    xin = x * beam if beam is not None else x
    psf_convolve_cube(xpad, xhat, xout, psfhat, lastsize, xin, nthreads=nthreads)
    if beam is not None:
        xout *= beam
    if wsum is not None:
        xout /= wsum
    return xout + x * sigmainv


# ----------------------------------------------------------------- solvers
def pcg(A, b, x0=None, M=None, tol=1e-5, maxit=500, minit=100, verbosity=1,
        report_freq=10, backtrack=True, return_resid=False):
    """pfb/opt/pcg.py:53-136."""
    if x0 is None:
        x0 = np.zeros(b.shape, dtype=b.dtype)
    if M is None:
        def M(x):
            return x
    r = A(x0) - b
    y = M(r)
    if not np.any(y):
        return x0
    p = -y
    rnorm = np.vdot(r, y)
    k = 0
    x = x0
    eps = 1.0
    stall_count = 0
    while (eps > tol or k < minit) and k < maxit and stall_count < 5:
        xp = x.copy()
        rp = r.copy()
        Ap = A(p)
        rnorm = np.vdot(r, y)
        alpha = rnorm / np.vdot(p, Ap)
        x = xp + alpha * p
        r = rp + alpha * Ap
        y = M(r)
        rnorm_next = np.vdot(r, y)
        while rnorm_next > rnorm and backtrack:
            alpha *= 0.75
            x = xp + alpha * p
            r = rp + alpha * Ap
            y = M(r)
            rnorm_next = np.vdot(r, y)
        beta = rnorm_next / rnorm
        p = beta * p - y
        if not np.any(p):
            break
        rnorm = rnorm_next
        k += 1
        eps = norm_diff(x, xp)
    if not return_resid:
        return x
    return x, r


def _pcg_psf_impl(psfhat, b, x0, beam, lastsize, nthreads, sigmainv, tol=1e-5,
                  maxit=500, minit=100, verbosity=1, report_freq=10,
                  backtrack=True):
    """pfb/opt/pcg.py:243-291: independent per-band pcg."""
    from functools import partial
    nband, nx, ny = b.shape
    _, nx_psf, nyo2 = psfhat.shape
    model = np.zeros((nband, nx, ny), dtype=b.dtype, order='C')
    if sigmainv > 0:
        def M(x):
            return x / sigmainv
    else:
        M = None
    for k in range(nband):
        xpad = np.empty((nx_psf, lastsize), dtype=b.dtype)
        xhat = np.empty((nx_psf, nyo2), dtype=psfhat.dtype)
        xout = np.empty((nx, ny), dtype=b.dtype)
        A = partial(_hessian_psf_slice, xpad, xhat, xout, psfhat[k], beam[k],
                    lastsize, nthreads=nthreads, sigmainv=sigmainv)
        model[k] = pcg(A, b[k], x0[k], M=M, tol=tol, maxit=maxit, minit=minit,
                       verbosity=verbosity, report_freq=report_freq,
                       backtrack=backtrack)
    return model


def primal_dual(x, v, lam, psi, psiH, L, prox, grad, nu=1.0, sigma=None,
                mask=None, tol=1e-5, maxit=1000, minit=10, positivity=1,
                report_freq=10, gamma=1.0, verbosity=1):
    """pfb/opt/primal_dual.py:12-87 (functional form: psiH(x) returns coeffs,
    psi(v) returns an image, prox(v, thr) returns an array)."""
    xp = x.copy()
    vp = v.copy()
    if sigma is None:
        sigma = L / (2.0 * gamma) / nu
    tau = 0.9 / (L / (2.0 * gamma) + sigma * nu**2)
    eps = 1.0
    k = 0
    while (eps > tol or k < minit) and k < maxit:
        vtilde = v + sigma * psiH(xp)
        v = vtilde - sigma * prox(vtilde / sigma, lam / sigma)
        x = xp - tau * (psi(2 * v - vp) + grad(xp))
        if positivity == 1:
            x[x < 0.0] = 0.0
        elif positivity == 2:
            msk = np.any(x <= 0, axis=0)
            x[:, msk] = 0.0
        eps = np.linalg.norm(x - xp) / np.linalg.norm(x)
        xp[...] = x[...]
        vp[...] = v[...]
        if np.isnan(eps) or np.isinf(eps):
            raise FloatingPointError("eps is nan/inf")
        k += 1
    return x, v


def primal_dual_optimised(x, v, lam, psiH, psi, L, prox, l1weight, reweighter,
                          grad, nu=1.0, sigma=None, mask=None, tol=1e-5,
                          maxit=1000, positivity=1, report_freq=10, gamma=1.0,
                          verbosity=1, maxreweight=50):
    """pfb/opt/primal_dual.py:91-180.  As in the reference, `psi(xp, v)` is
    image->coeffs (Psi.dot), `psiH(vp, xout)` is coeffs->image (Psi.hdot), the
    `prox` argument is unused and the dual update is the sum-over-band
    dual_update_numba.  pdb traps are replaced by FloatingPointError."""
    xp = x.copy()
    vp = v.copy()
    xout = np.zeros_like(x)
    if sigma is None:
        sigma = L / (2.0 * gamma) / nu
    tau = 0.9 / (L / (2.0 * gamma) + sigma * nu**2)
    eps = 1.0
    numreweight = 0
    for k in range(maxit):
        psi(xp, v)
        dual_update_numba(vp, v, lam, sigma=sigma, weight=l1weight)
        vp[...] = 2.0 * v - vp
        psiH(vp, xout)
        xout += grad(xp)
        x[...] = xp - tau * xout
        if positivity == 1:
            x[x < 0.0] = 0.0
        elif positivity == 2:
            msk = np.any(x <= 0, axis=0)
            x[:, msk] = 0.0
        if x.any():
            eps = norm_diff(x, xp)
        else:
            raise FloatingPointError("model is identically zero")
        if eps < tol:
            if reweighter is not None and numreweight < maxreweight:
                l1weight = reweighter(x)
                numreweight += 1
            else:
                break
        xp[...] = x[...]
        vp[...] = v[...]
        if np.isnan(eps) or np.isinf(eps):
            raise FloatingPointError("eps is nan/inf")
    return x, v


def power_method(A, imsize, b0=None, tol=1e-5, maxit=250, verbosity=1,
                 report_freq=25):
    """pfb/opt/power_method.py:11-49."""
    if b0 is None:
        b = np.random.randn(*imsize)
        b /= np.linalg.norm(b)
    else:
        b = b0 / np.linalg.norm(b0)
    beta = 1.0
    eps = 1.0
    k = 0
    bp = b.copy()
    while eps > tol and k < maxit:
        b = A(bp)
        bnorm = np.linalg.norm(b)
        betap = beta
        beta = np.vdot(bp, b) / np.vdot(bp, bp)
        b /= bnorm
        eps = np.linalg.norm(beta - betap) / betap
        k += 1
        bp[...] = b[...]
    return beta, b


def l1reweight_func(psiH, outvar, rmsfactor, rms_comps, model, alpha=4):
    """pfb/utils/misc.py:1070-1080 (psiH here is image->coeffs, Psi.dot)."""
    psiH(model, outvar)
    mcomps = np.abs(np.sum(outvar, axis=0))
    return (1 + rmsfactor) / (1 + mcomps**alpha / rms_comps**alpha)


# ------------------------------------------------- worker glue
def psfhat_from_psf(psf, nthreads=1):
    """pfb/operators/gridder.py:712-714: psfhat = r2c(iFs(psf, axes=(0,1)), axes=(0,1),
    forward=True, inorm=0) per band; psf (nband, nxp, nyp) real."""
    out = sfft.rfftn(np.fft.ifftshift(psf, axes=(1, 2)), axes=(1, 2), workers=nthreads)
    return out.astype(np.complex64 if _is32(psf) else np.complex128)


def make_data(residual, model, psf_convolve):
    """pfb/workers/spotless.py:259: data = residual + psf_convolve(model)"""
    return residual + psf_convolve(model)


def residual_stats(residual):
    """pfb/workers/spotless.py:153-157, 229-230, 343-344"""
    residual_mfs = np.sum(residual, axis=0)
    return np.std(residual_mfs), np.abs(residual_mfs).max(), residual_mfs


def rms_comps(psi, residual, wsum, wsums, pix_per_beam, psiHoutvar):
    """pfb/workers/spotless.py:214-223 (and :356-363)"""
    fsel = wsums > 0
    tmp2 = residual.copy()
    tmp2[fsel] *= wsum / wsums[fsel, None, None]
    psi.dot((tmp2 / pix_per_beam).astype(residual.dtype), psiHoutvar)
    return np.std(np.sum(psiHoutvar, axis=0), axis=(-1, -2))[:, None, None]


def model_eps(model, modelp):
    """pfb/workers/spotless.py:345"""
    return np.linalg.norm(model - modelp) / np.linalg.norm(model)


# ------------------------------------------------- CLEAN minor loops
def subminor(A, psf, Ip, Iq, model, wsums, gamma=0.05, th=0.0, maxit=10000):
    """pfb/deconv/clark.py:29-82 (+ subtract :12-26).  A (nband, nact) is the active set,
    pixel i of it sitting at (Ip[i], Iq[i]).  The reference masks A by PSF overlap but not
    Ip / Iq (:68-72), which is only consistent when the mask is all True, i.e. when the PSF
    is at least twice the image -- asserted here."""
    nband, nx_psf, ny_psf = psf.shape
    nxo2, nyo2 = nx_psf // 2, ny_psf // 2
    A = A.copy()
    Asearch = np.sum(A, axis=0) ** 2
    pq = Asearch.argmax()
    p, q = Ip[pq], Iq[pq]
    Amax = np.sqrt(Asearch[pq])
    fsel = wsums > 0
    if fsel.sum() == 0:
        raise ValueError("wsums are all zero")
    k = 0
    while Amax > th and k < maxit:
        xhat = A[:, pq].copy()
        # numba evaluates gamma * xhat / wsums in fp64 (gamma, wsums are float64) before the store
        model[fsel, p, q] += gamma * xhat[fsel].astype(np.float64) / wsums[fsel]
        Idelp, Idelq = p - Ip, q - Iq
        mask = (np.abs(Idelp) <= nxo2) & (np.abs(Idelq) <= nyo2)
        assert mask.all(), "PSF smaller than twice the image: the reference's active-set masking is inconsistent"
        A -= xhat[:, None] * psf[:, nxo2 - Idelp, nyo2 - Idelq]
        Asearch = np.sum(A, axis=0) ** 2
        pq = Asearch.argmax()
        p, q = Ip[pq], Iq[pq]
        Amax = np.sqrt(Asearch[pq])
        k += 1
    return model


def clark(ID, PSF, PSFHAT, wsums, threshold=0, gamma=0.05, pf=0.05, maxit=50, subpf=0.5,
          submaxit=1000, report_freq=1, verbosity=1, psfopts=None, sigmathreshold=2, nthreads=1):
    """pfb/deconv/clark.py:84-177.  Returns (model, status)."""
    nband, nx, ny = ID.shape
    _, nx_psf, ny_psf = PSF.shape
    assert np.allclose(wsums.sum(), 1)
    model = np.zeros((nband, nx, ny), dtype=ID.dtype)
    IR = ID.copy()
    xout = np.empty(ID.shape, dtype=ID.dtype)
    xpad = np.empty(PSF.shape, dtype=ID.dtype)
    xhat = np.empty(PSFHAT.shape, dtype=PSFHAT.dtype)
    IRsearch = np.sum(IR, axis=0) ** 2
    pq = IRsearch.argmax()
    p = pq // ny
    q = pq - p * ny
    IRmax = np.sqrt(IRsearch[p, q])
    tol = np.maximum(pf * IRmax, threshold)
    k = 0
    stall_count = 0
    while IRmax > tol and k < maxit and stall_count < 5:
        subth = subpf * IRmax
        Ip, Iq = np.where(IRsearch > subth ** 2)
        model = subminor(IR[:, Ip, Iq], PSF, Ip, Iq, model, wsums, gamma=gamma, th=subth, maxit=submaxit)
        psf_convolve_cube(xpad, xhat, xout, PSFHAT, ny_psf, model, nthreads=nthreads)
        IR = ID - xout
        IRsearch = np.sum(IR, axis=0) ** 2
        pq = IRsearch.argmax()
        p = pq // ny
        q = pq - p * ny
        IRmaxp = IRmax
        IRmax = np.sqrt(IRsearch[p, q])
        k += 1
        if np.abs(IRmaxp - IRmax) / np.abs(IRmaxp) < 1e-3:
            stall_count += stall_count          # (sic) pfb/deconv/clark.py:153: never leaves 0
    return model, (1 if (k >= maxit or stall_count >= 5) else 0)


def hogbom(ID, PSF, threshold=0, gamma=0.1, pf=0.1, maxit=10000, report_freq=1000, verbosity=1):
    """pfb/deconv/hogbom.py:8-72.  Returns (x, status)."""
    nband, nx, ny = ID.shape
    _, nx_psf, ny_psf = PSF.shape
    nx0, ny0 = nx_psf // 2, ny_psf // 2
    x = np.zeros((nband, nx, ny), dtype=ID.dtype)
    IR = ID.copy()
    IRsearch = np.sum(IR, axis=0) ** 2
    pq = IRsearch.argmax()
    p = pq // ny
    q = pq - p * ny
    IRmax = np.sqrt(IRsearch[p, q])
    wsums = np.amax(PSF, axis=(1, 2))
    fsel = wsums > 0
    tol = np.maximum(pf * IRmax, threshold)
    k = 0
    stall_count = 0
    while IRmax > tol and k < maxit and stall_count < 5:
        xhat = IR[fsel, p, q] / wsums[fsel]
        x[:, p, q] += gamma * xhat              # (the reference indexes all bands: needs wsums > 0 everywhere)
        IR[...] = IR - gamma * xhat[:, None, None] * PSF[:, nx0 - p:nx0 + nx - p, ny0 - q:ny0 + ny - q]
        IRsearch = np.sum(IR, axis=0) ** 2
        pq = IRsearch.argmax()
        p = pq // ny
        q = pq - p * ny
        IRmaxp = IRmax
        IRmax = np.sqrt(IRsearch[p, q])
        k += 1
        if np.abs(IRmaxp - IRmax) / np.abs(IRmaxp) < 5e-3:
            stall_count += stall_count          # (sic) pfb/deconv/hogbom.py:55
    return x, (1 if (k >= maxit or stall_count >= 5) else 0)


# ------------------------------------------------- dds2cubes / write-back
def dds2cubes(dds, nband, apparent=False, dual=True, modelname='MODEL'):
    """pfb/utils/misc.py:664-739 with numpy in place of the dask graph.  dds: sequence of
    datasets exposing DIRTY, BEAM, WSUM, bandid and optionally RESIDUAL, PSF, PSFHAT, MODEL,
    DUAL as `.NAME.data` (and `'NAME' in ds`)."""
    real_type = dds[0].DIRTY.dtype
    complex_type = np.result_type(real_type, np.complex64)
    nx, ny = dds[0].DIRTY.shape
    dirty = [np.zeros((nx, ny), real_type) for _ in range(nband)]
    model = [np.zeros((nx, ny), real_type) for _ in range(nband)]
    residual = [np.zeros((nx, ny), real_type) for _ in range(nband)] if 'RESIDUAL' in dds[0] else None
    wsums = [np.zeros(1, real_type) for _ in range(nband)]
    if 'PSF' in dds[0]:
        psf = [np.zeros(dds[0].PSF.shape, real_type) for _ in range(nband)]
        psfhat = [np.zeros(dds[0].PSFHAT.shape, complex_type) for _ in range(nband)]
    else:
        psf = psfhat = None
    mean_beam = [np.zeros((nx, ny), real_type) for _ in range(nband)]
    if dual and 'DUAL' in dds[0]:
        dual = [np.zeros(dds[0].DUAL.shape, real_type) for _ in range(nband)]
    else:
        dual = None
    for ds in dds:
        b = ds.bandid
        if apparent:
            dirty[b] = dirty[b] + ds.DIRTY.data
            if 'RESIDUAL' in ds:
                residual[b] = residual[b] + ds.RESIDUAL.data
        else:
            dirty[b] = dirty[b] + ds.DIRTY.data * ds.BEAM.data
            if 'RESIDUAL' in ds:
                residual[b] = residual[b] + ds.RESIDUAL.data * ds.BEAM.data
        if 'PSF' in ds:
            psf[b] = psf[b] + ds.PSF.data
            psfhat[b] = psfhat[b] + ds.PSFHAT.data
        if modelname in ds:
            model[b] = getattr(ds, modelname).data
        if dual and 'DUAL' in ds:
            dual[b] = ds.DUAL.data
        mean_beam[b] = mean_beam[b] + ds.BEAM.data * ds.WSUM.data[0]
        wsums[b] = wsums[b] + ds.WSUM.data[0]
    wsums = np.stack(wsums).reshape(nband)
    wsum = wsums.sum()
    dirty = np.stack(dirty) / wsum
    model = np.stack(model)
    if residual is not None:
        residual = np.stack(residual) / wsum
    if psf is not None:
        psf = np.stack(psf) / wsum
        psfhat = np.stack(psfhat) / wsum
    if dual:
        dual = np.stack(dual)
    for b in range(nband):
        if wsums[b]:
            mean_beam[b] = mean_beam[b] / wsums[b]
    mean_beam = np.stack(mean_beam)
    return dirty, model, residual, psf, psfhat, mean_beam, wsums, dual


def cubes2dds(dds, residual, model, dual, best_model, wsum):
    """pfb/workers/spotless.py:373-394: the per-dataset write-back of one major iteration:
    RESIDUAL = residual[b] * wsum, MODEL = model[b], DUAL = dual[b], MODEL_BEST = best_model[b]."""
    out = []
    for ds in dds:
        b = ds.bandid
        out.append({'RESIDUAL': residual[b] * wsum, 'MODEL': model[b], 'DUAL': dual[b],
                    'MODEL_BEST': best_model[b]})
    return out
