"""Drop-in for pfb/deconv/hogbom.py:8-72 on the device (csrc/clean.cu): the loop state
(peak, tolerance, iteration count, done flag) lives on the GPU, the host enqueues batches of
iterations and polls the flag once per batch."""
import numpy as np
import torch

from .. import _dev, _lib

_BATCH = 32


def hogbom(ID, PSF, threshold=0, gamma=0.1, pf=0.1, maxit=10000, report_freq=1000, verbosity=1):
    IDd = _dev.to_dev(ID)
    PSFd = _dev.to_dev(PSF)
    if PSFd.dtype != IDd.dtype:
        PSFd = PSFd.to(IDd.dtype)
    nband, nx, ny = IDd.shape
    _, nx_psf, ny_psf = PSFd.shape
    x = torch.zeros_like(IDd)
    IR = IDd.clone()
    wsums = torch.amax(PSFd.reshape(nband, -1), dim=1).contiguous()      # one-off set-up (hogbom.py:30)
    if not bool((wsums > 0).all()):
        raise ValueError("hogbom: every band needs a positive PSF peak")
    state = torch.zeros(8 + nband, dtype=torch.float64, device=IDd.device)
    L = _lib.lib()
    code = _dev.dcode(IDd)
    _lib.check(L.pfb_hogbom_start(_dev.ptr(IR), _dev.ptr(x), _dev.ptr(wsums), nband, nx * ny, float(gamma),
                                  float(pf), float(threshold), int(maxit), _dev.ptr(state), code, _dev.stream()))
    st = state[:8].tolist()
    while st[2] == 0.0:
        _lib.check(L.pfb_hogbom_iterate(_dev.ptr(IR), _dev.ptr(PSFd), _dev.ptr(x), _dev.ptr(wsums), nband, nx, ny,
                                        nx_psf, ny_psf, float(gamma), _BATCH, _dev.ptr(state), code, _dev.stream()))
        st = state[:8].tolist()
        if verbosity > 1:
            print("At iteration %i max residual = %f" % (int(st[3]), st[1]))
    k, IRmax = int(st[3]), st[1]
    if k >= maxit:
        if verbosity:
            print(f"Max iters reached. Max resid = {IRmax:.3e}")
        return _dev.like(ID, x), 1
    if verbosity:
        print(f"Success, converged after {k} iterations. Max resid = {IRmax:.3e}")
    return _dev.like(ID, x), 0
