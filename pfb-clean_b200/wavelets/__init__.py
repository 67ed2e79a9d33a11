"""Drop-in for pfb/wavelets (`from pfb.wavelets.wavelets import dwt2d, idwt2d,
coeff_size, signal_size, copyT`, pfb/wavelets/__init__.py:32).

coeff_size / signal_size are the size helpers of wavelets.py:20-27.  dwt2d /
idwt2d keep the reference's functional signatures (wavelets.py:175, :261): the
caller's scratch arrays (`cbuff`, `cbuffT`, `alpha`) and index tables are
accepted and ignored -- the same multi-level CUDA kernels that serve
pfb_b200.operators.psi.Psi are driven through a one-band, one-basis plan, which
is identified from the filter taps (the Daubechies db1..db5 banks compiled into
the library, include/pfb_b200.h pfb_filter_bank).
"""
import ctypes

import numpy as np

from .. import _dev, _lib


def coeff_size(nsignal, nfilter):
    return (nsignal + nfilter - 1) // 2


def signal_size(ncoeff, nfilter):
    return 2 * ncoeff - nfilter + 2


def filter_bank(name):
    """(dec_lo, dec_hi, rec_lo, rec_hi) of 'db1'..'db5' as fp64 arrays -- what
    pywt.Wavelet(name).filter_bank gives the reference (pfb/operators/psi.py:37-41)."""
    if not (isinstance(name, str) and name.startswith("db") and name[2:].isdigit()):
        raise ValueError(f"unknown wavelet {name!r}: expected 'db1'..'db5'")
    order = int(name[2:])
    buf = (ctypes.c_double * 40)()
    _lib.check(_lib.lib().pfb_filter_bank(order, buf))
    t = np.array(buf[:]).reshape(4, 10)[:, :2 * order]
    return tuple(t[i].copy() for i in range(4))


def _basis_of(lo, hi, which):
    """'dbN' whose bank matches the given taps (which = 0: decomposition pair,
    2: reconstruction pair)."""
    lo, hi = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)
    F = lo.shape[0]
    if F % 2 == 0 and 2 <= F <= 10 and hi.shape[0] == F:
        bank = filter_bank(f"db{F // 2}")
        if np.allclose(lo, bank[which], rtol=0, atol=1e-9) and np.allclose(hi, bank[which + 1], rtol=0, atol=1e-9):
            return f"db{F // 2}"
    raise ValueError("only the Daubechies db1..db5 filter banks are compiled into pfb_b200")


_PLANS = {}


def _psi(nx, ny, basis, nlevel):
    from ..operators.psi import Psi
    key = (nx, ny, basis, nlevel)
    p = _PLANS.get(key)
    if p is None:
        p = _PLANS[key] = Psi(1, nx, ny, [basis], nlevel, 1)
    return p


def dwt2d(image, coeffs, cbuff, cbuffT, ix, iy, sx, sy, dec_lo, dec_hi, nlevel):
    """pfb/wavelets/wavelets.py:174-213: multi-level 2-D analysis of the (nx, ny)
    image into the (Ntoty, Ntotx) coefficient slab (transposed w.r.t. the image;
    cells outside the pyramid blocks keep their content)."""
    nx, ny = image.shape
    P = _psi(nx, ny, _basis_of(dec_lo, dec_hi, 0), int(nlevel))
    if tuple(coeffs.shape) != (P.Nymax, P.Nxmax):
        raise ValueError(f"coeffs shape {tuple(coeffs.shape)} != {(P.Nymax, P.Nxmax)}")
    xd, cd = _dev.to_dev(image), _dev.to_dev(coeffs)
    if xd.dtype != cd.dtype:
        raise TypeError("image and coeffs must have the same precision")
    P.dot(xd[None], cd[None, None])
    _dev.to_out(cd, coeffs)


def idwt2d(coeffs, image, alpha, cbuff, cbuffT, ix, iy, sx, sy, spx, spy, rec_lo, rec_hi, nlevel):
    """pfb/wavelets/wavelets.py:260-315: multi-level 2-D synthesis; `image`
    (nx, ny) is overwritten, `coeffs` is not modified."""
    nx, ny = image.shape
    P = _psi(nx, ny, _basis_of(rec_lo, rec_hi, 2), int(nlevel))
    if tuple(coeffs.shape) != (P.Nymax, P.Nxmax):
        raise ValueError(f"coeffs shape {tuple(coeffs.shape)} != {(P.Nymax, P.Nxmax)}")
    cd, od = _dev.to_dev(coeffs), _dev.out_dev(image)
    if od.dtype != cd.dtype:
        raise TypeError("coeffs and image must have the same precision")
    P.hdot(cd[None, None], od[None])
    _dev.to_out(od, image)


def copyT(mat, out):
    """pfb/wavelets/wavelets.py:6-17: out = mat.T -- the `self` basis of Psi.dot
    (pfb/operators/psi.py:196-199) on a one-band plan (even sizes, like Psi)."""
    import torch
    n, m = mat.shape
    if tuple(out.shape) != (m, n):
        raise ValueError(f"out shape {tuple(out.shape)} != {(m, n)}")
    from ..operators.psi import Psi
    key = (n, m, "self+db1", 1)
    P = _PLANS.get(key)
    if P is None:            # a plan needs one wavelet basis next to `self`
        P = _PLANS[key] = Psi(1, n, m, ["self", "db1"], 1, 1)
    md, od = _dev.to_dev(mat), _dev.out_dev(out)
    if od.dtype != md.dtype:
        raise TypeError("mat and out must have the same precision")
    slab = torch.empty((1, 2, P.Nymax, P.Nxmax), dtype=md.dtype, device=md.device)
    P.dot(md[None], slab)
    od.copy_(slab[0, 0, :m, :n])
    _dev.to_out(od, out)
