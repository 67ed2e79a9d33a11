"""Drop-in for pfb/deconv/clark.py: clark :84-177 and subminor :29-82 on the device
(csrc/clean.cu).  Per minor iteration: one active-set compaction, ONE launch for the whole
sub-minor loop (peak search, component, shifted-PSF subtraction, up to submaxit times), the
PSF convolution of the model (the Hessian kernels) and the peak search of the new residual;
the host reads two small scalars (active-set size, peak) per minor iteration."""
import numpy as np
import torch

from .. import _dev, _lib
from ..operators.psf import psf_convolve_cube
from ..opt import _vec


def _np_t(t):
    return np.float32 if t.dtype == torch.float32 else np.float64


def _peak(IR, search, out2):
    nband = IR.shape[0]
    npix = IR.numel() // nband
    _lib.check(_lib.lib().pfb_mfs_peak(_dev.ptr(IR), nband, npix, _dev.ptr(search), _dev.ptr(out2),
                                       _dev.dcode(IR), _dev.stream()))
    pq, val = out2.tolist()
    return int(pq), val


def _active_set(search, thr2, idx, count, counts):
    npix = search.numel()
    _lib.check(_lib.lib().pfb_active_set(_dev.ptr(search), npix, float(thr2), _dev.ptr(idx), idx.numel(),
                                         _dev.ptr(count), _dev.ptr(counts), _dev.dcode(search), _dev.stream()))
    return int(count.item())


def _subminor_dev(IR, psf, idx, nact, model, wsums_dev, gamma, th, maxit, info):
    nband, nx, ny = IR.shape
    _, nxp, nyp = psf.shape
    A = torch.empty((nband, nact), dtype=IR.dtype, device=IR.device)
    _lib.check(_lib.lib().pfb_clark_subminor(_dev.ptr(IR), _dev.ptr(psf), _dev.ptr(idx), nact, _dev.ptr(model),
                                             _dev.ptr(wsums_dev), nband, nx, ny, nxp, nyp, float(gamma),
                                             float(th), int(maxit), _dev.ptr(A), _dev.ptr(info),
                                             _dev.dcode(IR), _dev.stream()))


def subminor(A, psf, Ip, Iq, model, wsums, gamma=0.05, th=0.0, maxit=10000):
    """pfb/deconv/clark.py:29-82.  A = IR[:, Ip, Iq] (nband, nact) is the active set, pixel i
    of it at (Ip[i], Iq[i]); the model is updated in place and returned."""
    md = _dev.to_dev(model)
    nband, nx, ny = md.shape
    Ad, pd = _dev.to_dev(A), _dev.to_dev(psf)
    Ipd = torch.as_tensor(np.asarray(Ip), dtype=torch.int64, device=md.device)
    Iqd = torch.as_tensor(np.asarray(Iq), dtype=torch.int64, device=md.device)
    idx = (Ipd * ny + Iqd).contiguous()
    # the kernel gathers the active set itself: scatter A into a residual-shaped buffer
    IR = torch.zeros((nband, nx * ny), dtype=md.dtype, device=md.device)
    IR[:, idx] = Ad.to(md.dtype)
    wd = torch.as_tensor(np.asarray(wsums, dtype=np.float64), device=md.device)
    info = _dev.scalars(2)
    _subminor_dev(IR.view(nband, nx, ny), pd.to(md.dtype), idx, idx.numel(), md, wd, gamma, th, maxit, info)
    return _dev.to_out(md, model)


def clark(ID, PSF, PSFHAT, wsums, threshold=0, gamma=0.05, pf=0.05, maxit=50, subpf=0.5,
          submaxit=1000, report_freq=1, verbosity=1, psfopts=None, sigmathreshold=2, nthreads=1):
    """pfb/deconv/clark.py:84-177.  Returns (model, status)."""
    IDd = _dev.to_dev(ID)
    PSFd = _dev.to_dev(PSF)
    if PSFd.dtype != IDd.dtype:
        PSFd = PSFd.to(IDd.dtype)
    psfhat = _dev.to_dev(PSFHAT)
    nband, nx, ny = IDd.shape
    _, nx_psf, ny_psf = PSFd.shape
    ws = np.asarray(wsums.cpu() if isinstance(wsums, torch.Tensor) else wsums, dtype=np.float64)
    assert np.allclose(ws.sum(), 1)
    if not (ws > 0).any():
        raise ValueError("wsums are all zero")
    wd = torch.as_tensor(ws, device=IDd.device)
    T = _np_t(IDd).__call__
    model = torch.zeros_like(IDd)
    IR = IDd.clone()
    xout = torch.empty_like(IDd)
    npix = nx * ny
    search = torch.empty(npix, dtype=IDd.dtype, device=IDd.device)
    idx = torch.empty(npix, dtype=torch.int64, device=IDd.device)
    counts = torch.empty((npix + 2047) // 2048, dtype=torch.int32, device=IDd.device)
    count = torch.zeros(1, dtype=torch.int64, device=IDd.device)
    out2, info = _dev.scalars(2), _dev.scalars(2)
    pq, val = _peak(IR, search, out2)
    IRmax = np.sqrt(T(val))
    tol = np.maximum(pf * IRmax, threshold)
    k = 0
    stall_count = 0
    while IRmax > tol and k < maxit and stall_count < 5:
        subth = subpf * IRmax                                  # same scalar dtype as the reference (NEP 50)
        nact = _active_set(search, subth ** 2, idx, count, counts)
        if nact:
            _subminor_dev(IR, PSFd, idx, nact, model, wd, gamma, subth, submaxit, info)
        psf_convolve_cube(None, None, xout, psfhat, ny_psf, model, nthreads=nthreads)
        _vec.axpby(IR, 1.0, IDd, -1.0, xout)                   # IR = ID - xout
        pq, val = _peak(IR, search, out2)
        IRmaxp = IRmax
        IRmax = np.sqrt(T(val))
        k += 1
        if np.abs(IRmaxp - IRmax) / np.abs(IRmaxp) < 1e-3:
            stall_count += stall_count                          # (sic) clark.py:153
        if not k % report_freq and verbosity > 1:
            print(f"At iteration {k} max resid = {IRmax}")
    if k >= maxit or stall_count >= 5:
        if verbosity:
            print(f"{'Max iters reached' if k >= maxit else 'Stalled'}. Max resid = {IRmax:.3e}")
        return _dev.like(ID, model), 1
    if verbosity:
        print(f"Success, converged after {k} iterations. Max resid = {IRmax:.3e}")
    return _dev.like(ID, model), 0
