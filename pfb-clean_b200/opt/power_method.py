"""Drop-in for pfb/opt/power_method.py:11-49 (spectral norm by power
iteration).  The three reductions per iteration are two fused dot kernels with
fp64 accumulation; with band sharding the scalars are all-reduced."""
import math

import numpy as np
import torch

from .. import _dev, parallel
from . import _vec


def power_method(A, imsize, b0=None, tol=1e-5, maxit=250, verbosity=1, report_freq=25):
    if b0 is None:
        b0 = np.random.randn(*imsize)      # same global RNG as the reference
        host = True
    else:
        host = isinstance(b0, np.ndarray)
    bp = _dev.to_dev(b0, copy=True)
    s2 = _dev.scalars(2)
    n0 = parallel.allreduce_sum(_vec.vdot2(bp, bp, out=s2)).tolist()[0]
    _vec.axpby(bp, 1.0 / math.sqrt(n0), bp, 0.0, bp)      # b0 / ||b0||  (power_method.py:20-21)
    beta = 1.0
    eps = 1.0
    k = 0
    b = bp
    while eps > tol and k < maxit:
        b = _dev.to_dev(A(bp))
        if b.dtype != bp.dtype:          # A works in the PSF precision
            bp = bp.to(b.dtype)
        bb = parallel.allreduce_sum(_vec.vdot2(b, b, out=s2)).tolist()[0]
        num, den = parallel.allreduce_sum(_vec.vdot2(bp, b, bp, bp, out=s2)).tolist()
        bnorm = math.sqrt(bb)
        betap = beta
        beta = num / den
        # b /= bnorm ; bp[...] = b   (one fused pass, written straight into bp)
        _vec.axpby(bp, 1.0 / bnorm, b, 0.0, b)
        b = bp
        eps = abs(beta - betap) / betap
        k += 1
        if not k % report_freq and verbosity > 1:
            print(f"At iteration {k} eps = {eps:.3e}")
    if verbosity:
        if k == maxit:
            print(f"Maximum iterations reached. eps = {eps:.3e}, beta = {beta:.3e}")
        else:
            print(f"Success, converged after {k} iterations. beta = {beta:.3e}")
    return beta, (b.cpu().numpy() if host and not _dev._EMULATED else (b.numpy() if host else b))
