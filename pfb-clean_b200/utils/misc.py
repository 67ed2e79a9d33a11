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


def band_sum(t, out=None):
    """Sum over the leading (band) axis on the device; with band sharding enabled
    the sum runs over ALL ranks' bands (SUM all-reduce of the local band sums)."""
    from .. import parallel
    nb = t.shape[0]
    m = t.numel() // nb
    if out is None:
        out = torch.empty(t.shape[1:], dtype=t.dtype, device=t.device)
    _lib.check(_lib.lib().pfb_dual_partial(_dev.ptr(t), _dev.ptr(t), _dev.ptr(out), nb, m, m, 0.0,
                                           _lib.BAND_SUM, _dev.dcode(t), _dev.stream()))
    return parallel.allreduce_sum(out)


def seg_stats(x, nseg):
    """(nseg, 3) fp64 device tensor {mean, population std, max |x|} of each of the nseg equal
    segments of x (pfb_seg_stats; two-pass std like numpy)."""
    xd = _dev.to_dev(x)
    out = _dev.scalars(3 * nseg)
    _lib.check(_lib.lib().pfb_seg_stats(_dev.ptr(xd), nseg, xd.numel() // nseg, _dev.ptr(out),
                                        _dev.dcode(xd), _dev.stream()))
    return out.view(nseg, 3)


def l1reweight_func(psiH, outvar, rmsfactor, rms_comps, model, alpha=4):
    """pfb/utils/misc.py:1070-1080: weights from the band-summed coefficient
    magnitudes of the current model (psiH is image -> coeffs, i.e. Psi.dot).
    One band-sum kernel + one weight kernel; with band sharding the sum is over
    every rank's bands, so all ranks derive the same weights (the sharded dual
    update relies on that)."""
    psiH(model, outvar)
    od = _dev.to_dev(outvar)
    bs = band_sum(od)
    nbasis = od.shape[1]
    if isinstance(rms_comps, (int, float)):
        rc = torch.full((nbasis,), float(rms_comps), dtype=od.dtype, device=od.device)
    else:
        rc = _dev.to_dev(rms_comps).to(od.dtype).reshape(-1)
        if rc.numel() == 1:
            rc = rc.expand(nbasis)
        if rc.numel() != nbasis:
            raise ValueError("rms_comps must be a scalar or hold one value per basis")
        rc = rc.contiguous()
    w = torch.empty_like(bs)
    _lib.check(_lib.lib().pfb_l1reweight(_dev.ptr(bs), _dev.ptr(rc), nbasis, bs.numel() // nbasis,
                                         float(rmsfactor), float(alpha), _dev.ptr(w), _dev.dcode(od),
                                         _dev.stream()))
    return _dev.like(outvar, w)
