"""Drop-in for pfb/operators/psf.py (psf_convolve_slice :11-29,
psf_convolve_cube :32-56) on top of the three-pass FFT Hessian kernels
(csrc/hess.cu) through the C-ABI of include/pfb_b200.h.

The reference's caller-owned scratch arrays `xpad` / `xhat` are accepted and
ignored (pass None to save the memory): the padded image and full half-spectrum
are never materialised on the GPU.  As in the reference the result is written
into, and returned as, the caller's `xout`.
"""
import ctypes
import weakref

import torch

from .. import _dev, _lib, parallel


class HessPlan:
    """RAII wrapper of pfb_hess_plan bound to one psfhat."""

    def __init__(self, nband, nx, ny, nxp, nyp, dtype_code):
        self._h = ctypes.c_void_p()
        self._destroy = _lib.lib().pfb_hess_plan_destroy
        _lib.check(_lib.lib().pfb_hess_plan_create(ctypes.byref(self._h), nband, nx, ny,
                                                   nxp, nyp, dtype_code))
        self.shape = (nband, nx, ny, nxp, nyp, dtype_code)
        self.psf_key = None
        self._dyn = None

    def set_psfhat(self, psfhat_dev):
        _lib.check(_lib.lib().pfb_hess_plan_set_psfhat(self._h, _dev.ptr(psfhat_dev), _dev.stream()))

    def set_dynamic(self, on):
        """Dynamic work distribution of the persistent kernels (used by the band-sharded
        solver, whose all-reduce runs concurrently with this operator)."""
        if on != self._dyn:
            _lib.check(_lib.lib().pfb_hess_plan_set_dynamic(self._h, int(on)))
            self._dyn = on

    def apply(self, x, out, beam=None, add_x=False, sigmainv=0.0, wsum=None):
        _lib.check(_lib.lib().pfb_hess_apply(
            self._h, _dev.ptr(x), _dev.ptr(beam), int(add_x), float(sigmainv),
            int(wsum is not None), float(wsum) if wsum is not None else 0.0,
            _dev.ptr(out), _dev.stream()))
        return out

    def can_fuse_primal(self):
        """True when pfb_hess_apply_primal is available (fp32 fast path, TMA pass C)."""
        return bool(_lib.lib().pfb_hess_plan_can_fuse_primal(self._h))

    def apply_primal(self, xp, w, wprev, data, tau, positivity, x, out2):
        """x = P+(xp - tau*(2w - wprev + conv(xp) - data)); out2 = the two eps terms."""
        _lib.check(_lib.lib().pfb_hess_apply_primal(
            self._h, _dev.ptr(xp), _dev.ptr(w), _dev.ptr(wprev), _dev.ptr(data), float(tau),
            int(positivity), _dev.ptr(x), _dev.ptr(out2), _dev.stream()))
        return out2

    def apply_primal_begin(self, xp):
        """The passes of apply_primal that need xp only (result kept in the plan's scratch)."""
        _lib.check(_lib.lib().pfb_hess_apply_primal_begin(self._h, _dev.ptr(xp), _dev.stream()))

    def apply_primal_end(self, xp, w, wprev, data, tau, positivity, x, out2):
        """Last pass + primal step of apply_primal; apply_primal_begin(xp) must precede it."""
        _lib.check(_lib.lib().pfb_hess_apply_primal_end(
            self._h, _dev.ptr(xp), _dev.ptr(w), _dev.ptr(wprev), _dev.ptr(data), float(tau),
            int(positivity), _dev.ptr(x), _dev.ptr(out2), _dev.stream()))
        return out2

    def __del__(self):
        if getattr(self, "_h", None) and self._h.value:
            self._destroy(self._h)
            self._h = ctypes.c_void_p()


_PLANS = {}            # (shape key) -> list of plans, most recently used first
_MAX_PLANS_PER_SHAPE = 2


def _psf_key(psfhat):
    """Identity + cheap content fingerprint of a psfhat.  torch: storage pointer and the
    in-place version counter.  numpy: pointer, layout and a strided sample of the contents
    (about 4096 values), so an array updated in place between major cycles is noticed without
    reading it whole."""
    if isinstance(psfhat, torch.Tensor):
        return ("t", psfhat.data_ptr(), tuple(psfhat.shape), psfhat._version)
    flat = psfhat.reshape(-1) if psfhat.flags.c_contiguous else psfhat.ravel()
    step = max(1, flat.size // 4096)
    sample = flat[::step]
    fp = (complex(sample.sum()), complex(flat[0]), complex(flat[-1]))
    return ("n", psfhat.ctypes.data, psfhat.shape, psfhat.strides, fp)


def clear_plan_cache():
    """Drop cached plans (releases their device memory)."""
    _PLANS.clear()


def get_plan(psfhat, lastsize, x_shape, x_dtype_code):
    """Plan bound to THIS psfhat.  Plans are cached per (shape, psfhat): a second psfhat of the
    same shape (a reweighter or callback convolving with another PSF mid-solve) gets its own
    plan instead of re-binding the one a running solver holds; at most
    _MAX_PLANS_PER_SHAPE live plans per shape, least recently used dropped first."""
    if len(x_shape) == 2:
        nband, (nx, ny) = 1, x_shape
    else:
        nband, nx, ny = x_shape
    nxp = psfhat.shape[-2]
    if psfhat.shape[-1] != lastsize // 2 + 1:
        raise ValueError(f"psfhat last axis {psfhat.shape[-1]} != lastsize//2+1 = {lastsize // 2 + 1}")
    key = (nband, nx, ny, nxp, lastsize, x_dtype_code)
    plans = _PLANS.setdefault(key, [])
    pk = _psf_key(psfhat)
    for i, plan in enumerate(plans):
        if plan.psf_key == pk:
            if i:
                plans.insert(0, plans.pop(i))
            return plan
    # same array object, new contents (updated in place): re-bind its plan rather than keep two
    plan = None
    for i, cand in enumerate(plans):
        if cand.psf_key is not None and cand.psf_key[:2] == pk[:2]:
            plan = plans.pop(i)
            break
    if plan is None:
        if len(plans) >= _MAX_PLANS_PER_SHAPE:
            plans.pop()
        plan = HessPlan(*key)
    plans.insert(0, plan)
    pd = _dev.to_dev(psfhat)
    if _dev.dcode(pd) != x_dtype_code:
        raise TypeError("psfhat precision does not match the image precision")
    plan.set_psfhat(pd)
    plan.psf_key = pk
    plan._keep = psfhat      # keep the pointer in the key valid
    return plan


def _convolve(psfhat, lastsize, x, xout, beam=None, add_x=False, sigmainv=0.0, wsum=None):
    # like the reference (`xpad[...] = x`, xpad/xhat in the precision of the
    # PSF) the transform runs in the precision of psfhat; x is cast if needed
    rdt = torch.float32 if str(psfhat.dtype).endswith("complex64") else torch.float64
    xd = _dev.to_dev(x)
    if xd.dtype != rdt:
        xd = xd.to(rdt)
    plan = get_plan(psfhat, lastsize, tuple(xd.shape), _dev.dcode(xd))
    od = _dev.out_dev(xout)
    direct = od.dtype == rdt
    res = od if direct else torch.empty(xd.shape, dtype=rdt, device=xd.device)
    if res.data_ptr() == xd.data_ptr():
        xd = xd.clone()
    bd = _dev.to_dev(beam) if beam is not None else None
    if bd is not None and bd.dtype != rdt:
        bd = bd.to(rdt)
    plan.set_dynamic(parallel.active())
    plan.apply(xd, res, bd, add_x, sigmainv, wsum)
    if not direct:
        od.copy_(res)
    return _dev.to_out(od, xout)


def psf_convolve_slice(xpad, xhat, xout, psfhat, lastsize, x, nthreads=1):
    """pfb/operators/psf.py:11-29 (2-D arrays)."""
    return _convolve(psfhat, lastsize, x, xout)


def psf_convolve_cube(xpad, xhat, xout, psfhat, lastsize, x, nthreads=1):
    """pfb/operators/psf.py:32-56: xout = crop(irfft2(rfft2(pad(x)) * psfhat))."""
    return _convolve(psfhat, lastsize, x, xout)


def psfhat_from_psf(psf):
    """psfhat = r2c(iFs(psf, axes=(0, 1)), axes=(0, 1), forward=True, inorm=0) per band
    (pfb/operators/gridder.py:712-714) on the device: psf (nband, nxp, nyp) or (nxp, nyp) real
    with its peak at (nxp//2, nyp//2) -> complex (.., nxp, nyp//2 + 1), numpy in -> numpy out,
    CUDA tensor in -> CUDA tensor out.  Forward-only use of the Hessian's own FFT passes."""
    pd = _dev.to_dev(psf)
    if pd.dim() not in (2, 3):
        raise ValueError("psf must be (nxp, nyp) or (nband, nxp, nyp)")
    p3 = pd if pd.dim() == 3 else pd[None]
    nband, nxp, nyp = p3.shape
    cdt = torch.complex64 if p3.dtype == torch.float32 else torch.complex128
    out = torch.empty((nband, nxp, nyp // 2 + 1), dtype=cdt, device=p3.device)
    _lib.check(_lib.lib().pfb_psfhat_from_psf(_dev.ptr(p3), nband, nxp, nyp, _dev.dcode(p3), _dev.ptr(out),
                                              _dev.stream()))
    if pd.dim() == 2:
        out = out[0]
    return _dev.like(psf, out)
