"""Drop-in for the image-space parts of pfb/operators/hessian.py:
_hessian_psf_slice :129-158 and hessian_psf_cube :254-281 (note the different
beam/psfhat argument order of the two, kept as in the reference).  The beam
prologue, beam epilogue, /wsum and + sigmainv*x are fused into the first and
last FFT pass.  Returns a NEW array like the reference (`xout + x*sigmainv`).
"""
import numpy as np
import torch

from .. import _dev
from .psf import _convolve


def _new_like(x):
    if isinstance(x, np.ndarray):
        return np.empty_like(x)
    return torch.empty_like(x)


def _hessian_psf_slice(xpad, xhat, xout, psfhat, beam, lastsize, x, nthreads=1,
                       sigmainv=1, wsum=None):
    """pfb/operators/hessian.py:129-158."""
    return _convolve(psfhat, lastsize, x, _new_like(x), beam=beam, add_x=True,
                     sigmainv=sigmainv, wsum=wsum)


def hessian_psf_cube(xpad, xhat, xout, beam, psfhat, lastsize, x, nthreads=1,
                     sigmainv=1, wsum=None):
    """pfb/operators/hessian.py:254-281."""
    return _convolve(psfhat, lastsize, x, _new_like(x), beam=beam, add_x=True,
                     sigmainv=sigmainv, wsum=wsum)
