"""Drop-in for pfb/opt/pcg.py: pcg :53-136 and _pcg_psf_impl :243-291.

All vectors stay on the device; each iteration issues A(p), one fused
two-dot-product kernel, one fused line-update kernel (x, r, <r,y>, eps terms)
and one fused direction kernel (p, any(p)).  xp/rp copies of the reference are
replaced by buffer ping-pong so a backtracking step is just a re-issue of the
line update with a smaller alpha.  With pfb_b200.parallel enabled the scalars
are all-reduced over the band-sharded ranks (global CG over the whole cube).
"""
import math
from functools import partial

import numpy as np
import torch

from .. import _dev, parallel
from ..operators.hessian import _hessian_psf_slice
from . import _vec


class ScalarPrecond:
    """M(x) = x / sigmainv, the preconditioner of _pcg_psf_impl
    (pfb/opt/pcg.py:264-267), in a form the fused kernels can absorb."""

    def __init__(self, sigmainv):
        self.minv = 1.0 / float(sigmainv)

    def __call__(self, x):
        return x * self.minv


def _host_(t, reduce=True):
    """fp64 device scalars -> list of Python floats (all-reduced if sharded)."""
    if reduce:
        parallel.allreduce_sum(t)
    return t.tolist()


def pcg(A, b, x0=None, M=None, tol=1e-5, maxit=500, minit=100, verbosity=1,
        report_freq=10, backtrack=True, return_resid=False, _global=True):
    """`_global=False` (extension) keeps the scalars rank-local when band
    sharding is enabled (independent per-band solves)."""
    _h = partial(_host_, reduce=_global)
    bd = _dev.to_dev(b)
    if x0 is None:
        x = torch.zeros_like(bd)
    else:
        x = _dev.to_dev(x0, copy=True)
        if x.dtype != bd.dtype:
            x = x.to(bd.dtype)
    fused = M is None or isinstance(M, ScalarPrecond)
    minv = 1.0 if M is None else (M.minv if fused else None)
    if M is None:
        def M(v):
            return v
    s2, s3, s1 = _dev.scalars(2), _dev.scalars(3), _dev.scalars(1)

    r = torch.empty_like(bd)
    _vec.axpby(r, 1.0, _dev.to_dev(A(x)), -1.0, bd)
    y = M(r)
    ry, yy = _h(_vec.vdot2(r, y, y, y, out=s2))
    if yy == 0.0:
        if verbosity:
            print("Initial residual is zero")
        res = _dev.like(b, x)
        return (res, _dev.like(b, r)) if return_resid else res
    p = -y
    rnorm = ry
    k = 0
    eps = 1.0
    xn, rn = torch.empty_like(x), torch.empty_like(r)
    while (eps > tol or k < minit) and k < maxit:
        Ap = _dev.to_dev(A(p))
        if fused:   # <r,y> is the previous iteration's rnorm_next: only <p,Ap> is new
            pAp = _h(_vec.vdot2(p, Ap, out=s2))[0]
        else:
            rnorm, pAp = _h(_vec.vdot2(r, y, p, Ap, out=s2))
        alpha = rnorm / pAp
        if fused:
            rnorm_next, num, den = _h(_vec.pcg_update(xn, rn, x, r, p, Ap, alpha, minv, s3))
            while rnorm_next > rnorm and backtrack:
                alpha *= 0.75
                rnorm_next, num, den = _h(_vec.pcg_update(xn, rn, x, r, p, Ap, alpha, minv, s3))
            beta = rnorm_next / rnorm
            nnz, = _h(_vec.pcg_direction(p, rn, beta, minv, s1))
            yn = None
        else:
            _vec.axpby(xn, 1.0, x, alpha, p)
            _vec.axpby(rn, 1.0, r, alpha, Ap)
            yn = M(rn)
            rnorm_next = _h(_vec.vdot2(rn, yn, out=s2))[0]
            while rnorm_next > rnorm and backtrack:
                alpha *= 0.75
                _vec.axpby(xn, 1.0, x, alpha, p)
                _vec.axpby(rn, 1.0, r, alpha, Ap)
                yn = M(rn)
                rnorm_next = _h(_vec.vdot2(rn, yn, out=s2))[0]
            beta = rnorm_next / rnorm
            _vec.axpby(p, beta, p, -1.0, yn)
            nnz = _h(_vec.vdot2(p, p, out=s2))[0]
            from ..utils.misc import norm_diff_terms
            num, den = _h(norm_diff_terms(xn, x, out=s2))
        # x <- xn, r <- rn (ping-pong instead of the reference's copies)
        x, xn = xn, x
        r, rn = rn, r
        y = yn
        if nnz == 0:
            break
        rnorm = rnorm_next
        k += 1
        eps = math.sqrt(num / (1e-12 + den))
        if math.isnan(eps):
            raise FloatingPointError("pcg: eps is NaN")
        if not k % report_freq and verbosity > 1:
            print(f"At iteration {k} epsx = {eps:.3e}")
    if verbosity:
        if k >= maxit:
            print(f"Max iters reached. eps = {eps:.3e}")
        else:
            print(f"Success, converged after {k} iterations")
    res = _dev.like(b, x)
    return (res, _dev.like(b, r)) if return_resid else res


def _pcg_psf_impl(psfhat, b, x0, beam, lastsize, nthreads, sigmainv, tol=1e-5,
                  maxit=500, minit=100, verbosity=1, report_freq=10, backtrack=True):
    """pfb/opt/pcg.py:243-291: independent per-band PCG with
    A = _hessian_psf_slice(psfhat[k], beam[k], sigmainv), M = x/sigmainv."""
    nband = b.shape[0]
    bd, x0d, beamd = _dev.to_dev(b), _dev.to_dev(x0), _dev.to_dev(beam)
    psfd = _dev.to_dev(psfhat)
    model = torch.zeros_like(bd)
    M = ScalarPrecond(sigmainv) if sigmainv > 0 else None
    for k in range(nband):
        A = partial(_hessian_psf_slice, None, None, None, psfd[k],
                    beamd[k] if beamd is not None else None, lastsize,
                    nthreads=nthreads, sigmainv=sigmainv)
        model[k] = pcg(A, bd[k], x0d[k], M=M, tol=tol, maxit=maxit, minit=minit,
                       verbosity=verbosity, report_freq=report_freq, backtrack=backtrack,
                       _global=False)
    return _dev.like(b, model)


def pcg_psf(psfhat, b, x0, beam, lastsize, nthreads, sigmainv, cgopts, compute=True):
    """pfb/opt/pcg.py:310-360.  The reference wraps _pcg_psf_impl in a
    dask.blockwise graph over band chunks; here the bands are solved one after
    the other on the device (eagerly: `compute=False` returns the same array, not
    a lazy graph).  beam: None, (nx, ny), (1, nx, ny) (tiled over the bands) or
    (nband, nx, ny); anything else raises ValueError like the reference."""
    nband = psfhat.shape[0]
    if beam is not None:
        bd = _dev.to_dev(beam)
        if bd.dim() == 2:
            bd = bd[None]
        if bd.shape[0] == 1:
            bd = bd.expand(nband, -1, -1).contiguous()
        elif bd.shape[0] != nband:
            raise ValueError('Beam has incorrect shape')
        beam = bd
    return _pcg_psf_impl(psfhat, b, x0, beam, lastsize, nthreads, sigmainv, **cgopts)
