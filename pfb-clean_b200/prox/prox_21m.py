"""Drop-in for pfb/prox/prox_21m.py ("MFS" l21 prox: SUM over the band axis,
then weighted soft threshold): prox_21m :5-27, prox_21m_numba :30-61,
dual_update :64-71, dual_update_numba :75-103."""
import numpy as np
import torch

from .. import _dev, _lib
from . import _common


def prox_21m(v, sigma, weight=1.0, axis=0):
    """numpy form: returns v * soft(|S|, sigma*weight)/|S| as a new array."""
    if axis != 0:
        raise ValueError("only axis=0 (band axis first) is supported, as used by the reference callers")
    return _common.prox_new(v, sigma, weight, _lib.BAND_SUM)


def prox_21m_numba(v, result, lam, sigma=1.0, weight=None):
    """writes `result` (same shape as v) in place."""
    _common.prox(v, result, lam, sigma, weight, _lib.BAND_SUM)


def dual_update_numba(vp, v, lam, sigma=1.0, weight=None):
    """in place on v (pre-loaded with Psi.dot(xp)); vp is read only."""
    _common.dual_update_inplace(vp, v, lam, sigma, weight, _lib.BAND_SUM)


def dual_update(v, x, psiH, lam, sigma=1.0, weight=1.0):
    """numpy form (pfb/prox/prox_21m.py:64-71): returns the updated dual."""
    vd = _dev.to_dev(v)
    vout = torch.zeros_like(vd)
    psiH(_dev.to_dev(x), vout)
    _common.dual_update_inplace(vd, vout, lam, sigma, weight, _lib.BAND_SUM)
    return _dev.like(v, vout)
