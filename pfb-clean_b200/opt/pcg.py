"""Drop-in for pfb/opt/pcg.py: pcg :53-136, _pcg_psf_impl :243-291, pcg_psf :310-360.

All vectors stay on the device.  With M = None or the scalar preconditioner of
_pcg_psf_impl (ScalarPrecond) the line search is device resident (csrc/pcg.cu): alpha and
beta are computed by a one-thread kernel from sums that never leave the GPU, and an
iteration is enqueued as
    A(p) -> <p,Ap> -> alpha -> fused line update (x, r, <r,y>, eps terms) -> beta ->
    fused direction (p, any(p))
with the NEXT A(p) already queued while the host reads the one small state record it needs
for eps / termination / the rare backtrack.  xp, rp, the old p and the old Ap are kept by
buffer ping-pong, so a backtracking step is a re-issue of the line update with a smaller
alpha.  The same kernels work on nseg independent segments: the whole cube (pcg) or one
segment per band with per-band scalars and per-band termination (_pcg_psf_impl, which the
reference runs concurrently over bands with da.blockwise).  Any other M falls back to the
host-scalar loop (_pcg_host).  With pfb_b200.parallel enabled the sums of the global solve
are all-reduced over the band-sharded ranks.
"""
import math
from functools import partial

import numpy as np
import torch

from .. import _dev, _lib, parallel
from ..operators.hessian import _hessian_psf_slice
from . import _vec


class ScalarPrecond:
    """M(x) = x / sigmainv, the preconditioner of _pcg_psf_impl
    (pfb/opt/pcg.py:264-267), in a form the fused kernels can absorb."""

    def __init__(self, sigmainv):
        self.minv = 1.0 / float(sigmainv)

    def __call__(self, x):
        return x * self.minv


def _host_(t, reduce=True):
    """fp64 device scalars -> list of Python floats (all-reduced if sharded)."""
    if reduce:
        parallel.allreduce_sum(t)
    return t.tolist()


def _pcg_host(A, b, x0=None, M=None, tol=1e-5, maxit=500, minit=100, verbosity=1,
              report_freq=10, backtrack=True, return_resid=False, _global=True):
    """Host-scalar loop (three small device->host reads per iteration): used for
    preconditioners that are arbitrary callables.  `_global=False` keeps the scalars
    rank-local when band sharding is enabled."""
    _h = partial(_host_, reduce=_global)
    bd = _dev.to_dev(b)
    if x0 is None:
        x = torch.zeros_like(bd)
    else:
        x = _dev.to_dev(x0, copy=True)
        if x.dtype != bd.dtype:
            x = x.to(bd.dtype)
    fused = M is None or isinstance(M, ScalarPrecond)
    minv = 1.0 if M is None else (M.minv if fused else None)
    if M is None:
        def M(v):
            return v
    s2, s3, s1 = _dev.scalars(2), _dev.scalars(3), _dev.scalars(1)

    r = torch.empty_like(bd)
    _vec.axpby(r, 1.0, _dev.to_dev(A(x)), -1.0, bd)
    y = M(r)
    ry, yy = _h(_vec.vdot2(r, y, y, y, out=s2))
    if yy == 0.0:
        if verbosity:
            print("Initial residual is zero")
        res = _dev.like(b, x)
        return (res, _dev.like(b, r)) if return_resid else res
    p = -y
    rnorm = ry
    k = 0
    eps = 1.0
    xn, rn = torch.empty_like(x), torch.empty_like(r)
    while (eps > tol or k < minit) and k < maxit:
        Ap = _dev.to_dev(A(p))
        if fused:   # <r,y> is the previous iteration's rnorm_next: only <p,Ap> is new
            pAp = _h(_vec.vdot2(p, Ap, out=s2))[0]
        else:
            rnorm, pAp = _h(_vec.vdot2(r, y, p, Ap, out=s2))
        alpha = rnorm / pAp
        if fused:
            rnorm_next, num, den = _h(_vec.pcg_update(xn, rn, x, r, p, Ap, alpha, minv, s3))
            while rnorm_next > rnorm and backtrack:
                alpha *= 0.75
                rnorm_next, num, den = _h(_vec.pcg_update(xn, rn, x, r, p, Ap, alpha, minv, s3))
            beta = rnorm_next / rnorm
            nnz, = _h(_vec.pcg_direction(p, rn, beta, minv, s1))
            yn = None
        else:
            _vec.axpby(xn, 1.0, x, alpha, p)
            _vec.axpby(rn, 1.0, r, alpha, Ap)
            yn = M(rn)
            rnorm_next = _h(_vec.vdot2(rn, yn, out=s2))[0]
            while rnorm_next > rnorm and backtrack:
                alpha *= 0.75
                _vec.axpby(xn, 1.0, x, alpha, p)
                _vec.axpby(rn, 1.0, r, alpha, Ap)
                yn = M(rn)
                rnorm_next = _h(_vec.vdot2(rn, yn, out=s2))[0]
            beta = rnorm_next / rnorm
            _vec.axpby(p, beta, p, -1.0, yn)
            nnz = _h(_vec.vdot2(p, p, out=s2))[0]
            from ..utils.misc import norm_diff_terms
            num, den = _h(norm_diff_terms(xn, x, out=s2))
        # x <- xn, r <- rn (ping-pong instead of the reference's copies)
        x, xn = xn, x
        r, rn = rn, r
        y = yn
        if nnz == 0:
            break
        rnorm = rnorm_next
        k += 1
        eps = math.sqrt(num / (1e-12 + den))
        if math.isnan(eps):
            raise FloatingPointError("pcg: eps is NaN")
        if not k % report_freq and verbosity > 1:
            print(f"At iteration {k} epsx = {eps:.3e}")
    if verbosity:
        if k >= maxit:
            print(f"Max iters reached. eps = {eps:.3e}")
        else:
            print(f"Success, converged after {k} iterations")
    res = _dev.like(b, x)
    return (res, _dev.like(b, r)) if return_resid else res


def _state_copy(st, pinned, ev):
    """Enqueue the copy of the device state block to pinned host memory (compute stream)."""
    if _dev._EMULATED:
        return st.clone()
    pinned.copy_(st, non_blocking=True)
    ev.record()
    return None


def _state_wait(snap, pinned, ev):
    """Wait for that copy only: whatever was queued behind it keeps the GPU busy."""
    if _dev._EMULATED:
        return snap.numpy()
    ev.synchronize()
    return pinned.numpy().copy()


def _state_read(st, pinned, ev):
    return _state_wait(_state_copy(st, pinned, ev), pinned, ev)


def _pcg_dev(A, bd, x, minv, nseg, tol, maxit, minit, verbosity, report_freq, backtrack,
             reduce, label="band"):
    """Device-resident PCG on nseg independent segments of the contiguous cube bd (see the
    module docstring).  A maps a cube to a cube.  Returns (x, r, k) with k the per-segment
    iteration counts."""
    L = _lib.lib()
    code = _dev.dcode(bd)
    n = bd.numel()
    seglen = n // nseg
    strm = _dev.stream
    st = torch.zeros((nseg, 10), dtype=torch.float64, device=bd.device)
    pinned = torch.empty((nseg, 10), dtype=torch.float64).pin_memory() if not _dev._EMULATED else None
    ev = torch.cuda.Event() if not _dev._EMULATED else None
    ar = (lambda lo, hi: parallel.allreduce_sum(st[0, lo:hi])) if (reduce and parallel.active()) else None
    assert ar is None or nseg == 1

    def dot(a, c, slot):
        _lib.check(L.pfb_pcg_seg_dot(_dev.ptr(a), _dev.ptr(c), nseg, seglen, code, _dev.ptr(st), slot, strm()))
        if ar:
            ar(slot, slot + 1)

    def scalar(step):
        _lib.check(L.pfb_pcg_seg_scalar(_dev.ptr(st), nseg, step, strm()))

    def apply_A(p, own):
        """Ap = A(p) as a device tensor of the working precision."""
        t = _dev.to_dev(A(p))
        if t.dtype != bd.dtype:
            t = t.to(bd.dtype)
        return t if t.is_contiguous() else t.contiguous()

    r = torch.empty_like(bd)
    _vec.axpby(r, 1.0, apply_A(x, None), -1.0, bd)
    if minv == 1.0:
        y = r
    else:
        y = torch.empty_like(r)
        _vec.axpby(y, float(minv), r, 0.0, r)
    dot(r, y, 0)
    dot(y, y, 7)
    h = _state_read(st, pinned, ev)
    active = h[:, 7] != 0.0
    k = np.zeros(nseg, dtype=np.int64)
    eps = np.ones(nseg)
    if not active.any():
        if verbosity:
            print("Initial residual is zero")
        return x, r, k
    st[:, 9] = torch.from_numpy(active.astype(np.float64)).to(st.device)
    p = torch.empty_like(bd)
    _vec.axpby(p, -1.0, y, 0.0, y)
    del y
    xn, rn, pn = torch.empty_like(x), torch.empty_like(r), torch.empty_like(p)
    Ap = apply_A(p, None)
    dot(p, Ap, 1)
    scalar(_lib.PCG_ALPHA)
    running = active.copy()

    def update():
        _lib.check(L.pfb_pcg_seg_update(_dev.ptr(xn), _dev.ptr(rn), _dev.ptr(x), _dev.ptr(r), _dev.ptr(p),
                                        _dev.ptr(Ap), float(minv), nseg, seglen, code, _dev.ptr(st), strm()))
        if ar:
            ar(3, 6)
        scalar(_lib.PCG_BETA)

    def direction():
        _lib.check(L.pfb_pcg_seg_direction(_dev.ptr(pn), _dev.ptr(p), _dev.ptr(rn), float(minv), nseg, seglen,
                                           code, _dev.ptr(st), strm()))
        if ar:
            ar(7, 8)

    def set_active():
        st[:, 9] = torch.from_numpy(running.astype(np.float64)).to(st.device)
    pending_nnz = False          # slot 7 of the state holds the count of the direction used by this iteration
    while running.any():
        # The host needs this iteration's sums for eps / termination / the backtrack test.  They are
        # copied right behind the line update; the (cheap) direction kernel is queued speculatively
        # behind the copy, so by the time the host has decided, A(pn) is queued while the GPU is
        # still busy.  Only a backtrack throws the direction away.
        update()
        snap = _state_copy(st, pinned, ev)
        direction()
        h = _state_wait(snap, pinned, ev)
        if pending_nnz:
            # `if not np.any(p): break` (pcg.py:106-107) of the PREVIOUS iteration, seen one
            # iteration late: the segment's solution is the x this iteration started from
            dead = running & (h[:, 7] == 0.0)
            if dead.any():
                running &= ~dead
                set_active()
                if not running.any():
                    break
                update()                     # pass the dead segments through unchanged
                direction()
                h = _state_read(st, pinned, ev)
        if backtrack and (h[:, 8] != 0.0).any():                 # pfb/opt/pcg.py:96-101
            while (h[:, 8] != 0.0).any():
                scalar(_lib.PCG_SHRINK)
                update()
                h = _state_read(st, pinned, ev)
            direction()
        pending_nnz = True
        changed = False
        for s_ in np.nonzero(running)[0]:
            k[s_] += 1
            eps[s_] = math.sqrt(h[s_, 4] / (1e-12 + h[s_, 5]))
            if math.isnan(eps[s_]):
                raise FloatingPointError("pcg: eps is NaN")
            if not k[s_] % report_freq and verbosity > 1:
                print(f"At iteration {k[s_]} epsx = {eps[s_]:.3e}")
            if not ((eps[s_] > tol or k[s_] < minit) and k[s_] < maxit):
                running[s_] = False
                changed = True
        scalar(_lib.PCG_COMMIT)
        x, xn = xn, x
        r, rn = rn, r
        p, pn = pn, p
        if changed:
            set_active()
        if running.any():
            Ap = apply_A(p, None)
            dot(p, Ap, 1)
            scalar(_lib.PCG_ALPHA)
    if verbosity:
        for s_ in range(nseg):
            tag = f" ({label} {s_})" if nseg > 1 else ""
            if k[s_] >= maxit:
                print(f"Max iters reached{tag}. eps = {eps[s_]:.3e}")
            elif active[s_]:
                print(f"Success, converged after {k[s_]} iterations{tag}")
            else:
                print(f"Initial residual is zero{tag}")
    return x, r, k


def pcg(A, b, x0=None, M=None, tol=1e-5, maxit=500, minit=100, verbosity=1,
        report_freq=10, backtrack=True, return_resid=False, _global=True):
    """pfb/opt/pcg.py:53-136.  `_global=False` (extension) keeps the scalars rank-local when
    band sharding is enabled."""
    if not (M is None or isinstance(M, ScalarPrecond)):
        pcg.last_path = "host-loop"
        return _pcg_host(A, b, x0, M, tol, maxit, minit, verbosity, report_freq, backtrack,
                         return_resid, _global)
    pcg.last_path = "device-scalars"
    bd = _dev.to_dev(b)
    if x0 is None:
        x = torch.zeros_like(bd)
    else:
        x = _dev.to_dev(x0, copy=True)
        if x.dtype != bd.dtype:
            x = x.to(bd.dtype)
    minv = 1.0 if M is None else M.minv
    x, r, _ = _pcg_dev(A, bd.contiguous(), x.contiguous(), minv, 1, tol, maxit, minit, verbosity,
                       report_freq, backtrack, _global)
    res = _dev.like(b, x)
    return (res, _dev.like(b, r)) if return_resid else res


def _pcg_psf_impl(psfhat, b, x0, beam, lastsize, nthreads, sigmainv, tol=1e-5,
                  maxit=500, minit=100, verbosity=1, report_freq=10, backtrack=True):
    """pfb/opt/pcg.py:243-291: independent per-band PCG with
    A = _hessian_psf_slice(psfhat[k], beam[k], sigmainv), M = x/sigmainv -- all bands in one
    launch set, each with its own alpha / beta / eps and its own termination."""
    from ..operators.hessian import hessian_psf_cube
    nband = b.shape[0]
    bd, x0d, beamd = _dev.to_dev(b), _dev.to_dev(x0, copy=True), _dev.to_dev(beam)
    psfd = _dev.to_dev(psfhat)
    minv = 1.0 / float(sigmainv) if sigmainv > 0 else 1.0
    A = partial(hessian_psf_cube, None, None, None, beamd, psfd, lastsize, nthreads=nthreads,
                sigmainv=sigmainv)
    if x0d.dtype != bd.dtype:
        x0d = x0d.to(bd.dtype)
    model, _, _ = _pcg_dev(A, bd.contiguous(), x0d.contiguous(), minv, nband, tol, maxit, minit,
                           verbosity, report_freq, backtrack, False)
    return _dev.like(b, model)


def pcg_psf(psfhat, b, x0, beam, lastsize, nthreads, sigmainv, cgopts, compute=True):
    """pfb/opt/pcg.py:310-360.  The reference wraps _pcg_psf_impl in a
    dask.blockwise graph over band chunks; here the bands are solved one after
    the other on the device (eagerly: `compute=False` returns the same array, not
    a lazy graph).  beam: None, (nx, ny), (1, nx, ny) (tiled over the bands) or
    (nband, nx, ny); anything else raises ValueError like the reference."""
    nband = psfhat.shape[0]
    if beam is not None:
        bd = _dev.to_dev(beam)
        if bd.dim() == 2:
            bd = bd[None]
        if bd.shape[0] == 1:
            bd = bd.expand(nband, -1, -1).contiguous()
        elif bd.shape[0] != nband:
            raise ValueError('Beam has incorrect shape')
        beam = bd
    return _pcg_psf_impl(psfhat, b, x0, beam, lastsize, nthreads, sigmainv, **cgopts)
