"""Shared implementation of the sum-over-band (prox_21m) and l2-over-band
(prox_21) l21 prox / dual update on the streaming kernels of csrc/prox.cu."""
import numpy as np
import torch

from .. import _dev, _lib


def _weight_dev(weight, like, shape):
    """Weights broadcast to the per-coefficient shape (scalars allowed as in the
    numpy reference functions)."""
    if isinstance(weight, (int, float)) or weight is None:
        w = torch.full(shape, 1.0 if weight is None else float(weight), dtype=like.dtype,
                       device=like.device)
        return w
    wd = _dev.to_dev(weight)
    if wd.dtype != like.dtype:
        wd = wd.to(like.dtype)
    if tuple(wd.shape) != tuple(shape):
        wd = wd.expand(shape).contiguous()
    return wd


def prox(v, result, lam, sigma, weight, mode):
    vd = _dev.to_dev(v)
    rd = _dev.out_dev(result)
    n = int(np.prod(vd.shape[1:]))
    wd = _weight_dev(weight, vd, tuple(vd.shape[1:]))
    _lib.check(_lib.lib().pfb_prox21(_dev.ptr(vd), _dev.ptr(rd), _dev.ptr(wd), vd.shape[0], n,
                                     float(lam), float(sigma), mode, _dev.dcode(vd), _dev.stream()))
    return _dev.to_out(rd, result)


def prox_new(v, thr, weight, mode):
    """numpy-form prox: returns a new array, threshold thr*weight."""
    vd = _dev.to_dev(v)
    rd = torch.empty_like(vd)
    n = int(np.prod(vd.shape[1:]))
    wd = _weight_dev(weight, vd, tuple(vd.shape[1:]))
    _lib.check(_lib.lib().pfb_prox21(_dev.ptr(vd), _dev.ptr(rd), _dev.ptr(wd), vd.shape[0], n,
                                     float(thr), 1.0, mode, _dev.dcode(vd), _dev.stream()))
    return _dev.like(v, rd)


def dual_update_inplace(vp, v, lam, sigma, weight, mode, u=None):
    vpd = _dev.to_dev(vp)
    vd = _dev.to_dev(v)
    n = int(np.prod(vd.shape[1:]))
    wd = _weight_dev(weight, vd, tuple(vd.shape[1:]))
    _lib.check(_lib.lib().pfb_dual_update(_dev.ptr(vpd), _dev.ptr(vd), _dev.ptr(wd), _dev.ptr(u),
                                          vd.shape[0], n, float(lam), float(sigma), mode,
                                          _dev.dcode(vd), _dev.stream()))
    return _dev.to_out(vd, v)
