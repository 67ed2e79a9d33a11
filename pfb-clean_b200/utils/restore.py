"""Restoring convolution on the device (the "cheap f5" next to the minor cycle):
drop-ins for `Gaussian2D`, `get_padding_info` and `convolve2gaussres`
(pfb/utils/misc.py:109-138, 170-184, 186-238), the calls behind pfb/workers/restore.py:200-215 and
pfb/utils/restoration.py:46-52.

`convolve2gaussres` is the PSF-convolution operator of the minor cycle with a Gaussian for a PSF:
the reference pads image and kernel CENTRALLY and brackets the product of their spectra with
ifftshift / fftshift; a circular convolution does not care where the image sits in the padded
array, so this is `psf_convolve_cube(image, psfhat = r2c(iFs(padded kernel)))` on the same padded
size -- the three-pass pruned FFT kernels (only the nx non-zero rows are transformed), with the
kernel's spectrum from `pfb_psfhat_from_psf`.  Arrays are numpy (copied in, numpy out) or CUDA
tensors (tensor out)."""

import numpy as np
import torch

from .. import _dev, _lib
from ..operators.psf import psf_convolve_cube, psfhat_from_psf


def good_size(n, real=True):
    """ducc0.fft.good_size(n, True): smallest 2^a 3^b 5^c >= n (11-smooth when not real)."""
    n = int(n)
    if n <= 6:
        return n
    primes = (2, 3, 5) if real else (2, 3, 5, 7, 11)
    best = 2 * n

    def walk(v, i):
        nonlocal best
        if v >= n:
            best = min(best, v)
            return
        for j in range(i, len(primes)):
            if v * primes[j] < best:
                walk(v * primes[j], j)
    walk(1, 0)
    return best


def get_padding_info(nx, ny, pfrac):
    """pfb/utils/misc.py:170-184."""
    npad_x = int(pfrac * nx)
    nfft = good_size(nx + npad_x, True)
    npad_xl = (nfft - nx) // 2
    npad_xr = nfft - nx - npad_xl
    npad_y = int(pfrac * ny)
    nfft = good_size(ny + npad_y, True)
    npad_yl = (nfft - ny) // 2
    npad_yr = nfft - ny - npad_yl
    padding = ((0, 0), (npad_xl, npad_xr), (npad_yl, npad_yr))
    return padding, slice(npad_xl, -npad_xr), slice(npad_yl, -npad_yr)


def _gauss_dev(xd, yd, GaussPar, normalise, nsigma, dtype):
    """Gaussian2D as a device tensor of `dtype` (fp64 arithmetic like the reference)."""
    S0, S1, PA = (float(g) for g in GaussPar)
    Smaj, Smin = S0, S1
    A = np.array([[1. / Smin ** 2, 0], [0, 1. / Smaj ** 2]])
    t = np.deg2rad(-PA)
    R = np.array([[np.cos(t), -np.sin(t)], [np.sin(t), np.cos(t)]])
    A = np.dot(np.dot(R.T, A), R)
    extent = (nsigma * Smaj) ** 2
    fwhm_conv = 2 * np.sqrt(2 * np.log(2))
    out = torch.empty(tuple(xd.shape), dtype=torch.float64, device=xd.device)
    _lib.check(_lib.lib().pfb_gauss2d(_dev.ptr(xd), _dev.ptr(yd), xd.numel(), float(A[0, 0]),
                                      float(A[0, 1] + A[1, 0]), float(A[1, 1]), float(fwhm_conv),
                                      float(extent), _dev.ptr(out), _lib.F64, _dev.stream()))
    if normalise:
        from .misc import seg_stats
        # np.sum(gausskern): mean * n, accumulated in fp64 in a fixed order
        from ..opt import _vec
        mean = seg_stats(out.reshape(-1), 1).tolist()[0][0]
        _vec.axpby(out, 1.0 / (mean * out.numel()), out, 0.0, out)
    return out if dtype == torch.float64 else out.to(dtype)


def _coords(a):
    d = _dev.to_dev(a)
    return (d if d.dtype == torch.float64 else d.to(torch.float64)).contiguous()


def Gaussian2D(xin, yin, GaussPar=(1., 1., 0.), normalise=True, nsigma=5):
    """pfb/utils/misc.py:109-138 (float64 result of the input's shape)."""
    xd, yd = _coords(xin), _coords(yin)
    sq = xd.squeeze(), yd.squeeze()
    g = _gauss_dev(sq[0].contiguous(), sq[1].contiguous(), GaussPar, normalise, nsigma, torch.float64)
    return _dev.like(xin, g.reshape(tuple(xd.shape)))


def convolve2gaussres(image, xx, yy, gaussparf, nthreads=1, gausspari=None, pfrac=0.5,
                      norm_kernel=False):
    """pfb/utils/misc.py:186-238: convolve the (nband, nx, ny) image to the resolution gaussparf,
    from no resolution (gausspari None) or from the per-band resolutions gausspari."""
    img = _dev.to_dev(image)
    if img.dim() != 3:
        raise ValueError("image must be (nband, nx, ny)")
    nband, nx, ny = img.shape
    padding, _, _ = get_padding_info(nx, ny, pfrac)
    nfx = nx + padding[1][0] + padding[1][1]
    lastsize = ny + padding[2][0] + padding[2][1]
    if lastsize % 2:
        raise ValueError(f"padded size {lastsize} along y is odd: the r2c kernels need an even lastsize "
                         "(choose ny / pfrac so that good_size(ny + int(pfrac*ny)) is even)")
    dt = img.dtype
    xd, yd = _coords(xx).squeeze().contiguous(), _coords(yy).squeeze().contiguous()

    def padded_hat(pars_list):
        """r2c(iFs(pad(Gaussian2D(pars)))) for every entry: (len, nfx, lastsize//2+1)."""
        pad = torch.zeros((len(pars_list), nfx, lastsize), dtype=dt, device=img.device)
        for i, pars in enumerate(pars_list):
            pad[i, padding[1][0]:padding[1][0] + nx, padding[2][0]:padding[2][0] + ny] = \
                _gauss_dev(xd, yd, pars, norm_kernel, 5, dt)
        return psfhat_from_psf(pad)

    ghat = padded_hat([gaussparf])                                    # misc.py:212-215
    if gausspari is None:
        khat = ghat.expand(nband, -1, -1).contiguous()                # imhat *= gausskernhat
    else:
        that = padded_hat([gausspari[i] for i in range(nband)])       # misc.py:224-233
        khat = torch.empty_like(that)
        rv = torch.view_as_real
        _lib.check(_lib.lib().pfb_spec_ratio(_dev.ptr(rv(ghat)), _dev.ptr(rv(that)), _dev.ptr(rv(khat)),
                                             ghat.numel(), that.numel(), _dev.dcode(img), _dev.stream()))
    out = torch.empty_like(img)
    psf_convolve_cube(None, None, out, khat, lastsize, img.contiguous(), nthreads)
    return _dev.like(image, out)
