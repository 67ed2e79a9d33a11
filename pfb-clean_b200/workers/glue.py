"""Device-side glue of the spotless major cycle: the lines of
pfb/workers/spotless.py between two calls of primal_dual that touch the whole
cube, as compositions of the operator kernels plus two reductions, so model,
residual, data and dual can stay on the device across major iterations.

Every function takes numpy arrays (reference convention: copied in, result
returned as numpy) or CUDA tensors (no copies, tensor results).  With
pfb_b200.parallel enabled each rank passes its LOCAL bands and the band sums /
statistics are over all ranks' bands.
"""
import numpy as np
import torch

from .. import _dev
from ..utils.misc import band_sum, seg_stats, l1reweight_func  # noqa: F401  (re-exported)
from ..operators.psf import psfhat_from_psf  # noqa: F401  (gridder.py:712-714)
from ..opt import _vec


def make_data(residual, model, psf_convolve):
    """data = residual + psf_convolve(model)   (pfb/workers/spotless.py:259).
    psf_convolve returns its shared xout buffer (psf.py:54); the sum is a new array."""
    conv = _dev.to_dev(psf_convolve(model))
    rd = _dev.to_dev(residual)
    if rd.dtype != conv.dtype:
        rd = rd.to(conv.dtype)
    out = torch.empty_like(conv)
    _vec.axpby(out, 1.0, rd, 1.0, conv)
    return _dev.like(residual, out)


def rms_comps(psi, residual, wsum, wsums, pix_per_beam, psiHoutvar):
    """rms of the band-summed dictionary coefficients of the residual in model units
    (pfb/workers/spotless.py:214-223, 356-363):
        tmp2 = residual.copy(); tmp2[fsel] *= wsum / wsums[fsel]
        psi.dot(tmp2 / pix_per_beam, psiHoutvar)
        rms_comps = np.std(np.sum(psiHoutvar, axis=0), axis=(-1, -2))[:, None, None]
    psiHoutvar (nband, nbasis, Nymax, Nxmax) is overwritten on its pyramid cells like the
    reference's scratch.  Returns (nbasis, 1, 1)."""
    rd = _dev.to_dev(residual)
    ws = np.asarray(wsums, dtype=np.float64).reshape(-1)
    tmp = torch.empty_like(rd)
    for b in range(rd.shape[0]):       # one per-band scale: (wsum / wsums_b) / pix_per_beam
        sc = (float(wsum) / ws[b] if ws[b] > 0 else 1.0) / float(pix_per_beam)
        _vec.axpby(tmp[b], sc, rd[b], 0.0, rd[b])
    psi.dot(tmp, psiHoutvar)
    od = _dev.to_dev(psiHoutvar)
    bs = band_sum(od)
    st = seg_stats(bs, od.shape[1])
    res = st[:, 1].to(od.dtype).reshape(-1, 1, 1)
    return _dev.like(psiHoutvar, res) if isinstance(psiHoutvar, np.ndarray) else res


def residual_stats(residual):
    """rms = np.std(residual_mfs), rmax = np.abs(residual_mfs).max() with
    residual_mfs = np.sum(residual, axis=0)   (pfb/workers/spotless.py:153-157, 229-230,
    338-344).  Returns (rms, rmax, residual_mfs)."""
    rd = _dev.to_dev(residual)
    mfs = band_sum(rd)
    st = seg_stats(mfs, 1).tolist()[0]
    return st[1], st[2], _dev.like(residual, mfs)


def model_eps(model, modelp):
    """eps = ||model - modelp|| / ||model||   (pfb/workers/spotless.py:345)."""
    from ..utils.misc import norm_diff_terms
    from .. import parallel
    md, mp = _dev.to_dev(model), _dev.to_dev(modelp)
    num, den = parallel.allreduce_sum(norm_diff_terms(md, mp)).tolist()
    return float(np.sqrt(num) / np.sqrt(den)) if den > 0 else float("nan")


class BestModel:
    """best_rms / best_rmax / best_model bookkeeping of pfb/workers/spotless.py:231-233,
    347-351 with the copy kept on the device."""

    def __init__(self, model, rms, rmax):
        self.rms, self.rmax = rms, rmax
        self.model = _dev.to_dev(model, copy=True)

    def update(self, model, rms, rmax):
        if rms < self.rms:
            self.rms, self.rmax = rms, rmax
            self.model.copy_(_dev.to_dev(model))
            return True
        return False
