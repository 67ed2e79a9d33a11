"""Drop-in for pfb/prox/prox_21.py (true l2 norm over the band axis):
prox_21 :5-22, prox_21_numba :25-52, dual_update :55-63, dual_update_numba
:66-91.  Arrays may be (nband, nbasis, ntot) as in the reference or 4-D."""
import torch

from .. import _dev, _lib
from . import _common


def prox_21(v, sigma, weight=None, axis=0):
    if axis != 0:
        raise ValueError("only axis=0 (band axis first) is supported")
    return _common.prox_new(v, sigma, weight, _lib.BAND_L2)


def prox_21_numba(v, result, lam, sigma=1.0, weight=None):
    _common.prox(v, result, lam, sigma, weight, _lib.BAND_L2)


def dual_update_numba(vp, v, lam, sigma=1.0, weight=None):
    _common.dual_update_inplace(vp, v, lam, sigma, weight, _lib.BAND_L2)


def dual_update(v, x, psiH, lam, sigma=1.0, weight=1.0):
    vd = _dev.to_dev(v)
    vout = torch.zeros_like(vd)
    psiH(_dev.to_dev(x), vout)
    _common.dual_update_inplace(vd, vout, lam, sigma, weight, _lib.BAND_L2)
    return _dev.like(v, vout)
