"""Hot-path pieces of pfb/utils/misc.py: norm_diff :1316-1351 and
l1reweight_func :1070-1080."""
import math

import torch

from .. import _dev, _lib


def norm_diff_terms(xd, xpd, out=None):
    """Device-side {sum (x-xp)^2, sum x^2} (fp64, deterministic)."""
    if out is None:
        out = _dev.scalars(2)
    _lib.check(_lib.lib().pfb_norm_diff(_dev.ptr(xd), _dev.ptr(xpd), xd.numel(), _dev.dcode(xd),
                                        _dev.ptr(out), _dev.stream()))
    return out


def norm_diff(x, xp):
    """eps = sqrt(sum (x-xp)^2 / (1e-12 + sum x^2)) -> Python float (syncs)."""
    xd, xpd = _dev.to_dev(x), _dev.to_dev(xp)
    if xd.shape != xpd.shape or xd.dtype != xpd.dtype:
        raise ValueError("norm_diff: x and xp must have the same shape and dtype")
    if xd.dim() not in (2, 3):
        raise ValueError("norm_diff is only implemented for 2D or 3D arrays")
    r = norm_diff_terms(xd, xpd).tolist()
    return math.sqrt(r[0] / (1e-12 + r[1]))


def l1reweight_func(psiH, outvar, rmsfactor, rms_comps, model, alpha=4):
    """pfb/utils/misc.py:1070-1080: weights from the band-summed coefficient
    magnitudes of the current model (psiH is image -> coeffs, i.e. Psi.dot)."""
    psiH(model, outvar)
    od = _dev.to_dev(outvar)
    mcomps = torch.abs(torch.sum(od, dim=0))
    rc = _dev.to_dev(rms_comps) if not isinstance(rms_comps, (int, float)) else rms_comps
    w = (1 + rmsfactor) / (1 + mcomps**alpha / rc**alpha)
    return _dev.like(outvar, w)
