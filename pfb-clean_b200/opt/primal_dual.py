"""Drop-in for pfb/opt/primal_dual.py: primal_dual :12-87 (functional form) and
primal_dual_optimised :91-180 (the in-place form the spotless worker runs).

primal_dual_optimised keeps the reference's signature, including its quirks:
`psi(xp, v)` is image -> coeffs (Psi.dot), `psiH(vp, xout)` is coeffs -> image
(Psi.hdot), `prox` is accepted and unused (the dual update is hard-wired to the
sum-over-band dual_update_numba, primal_dual.py:136), x and v are updated in
place and returned.  The pdb traps of the reference become FloatingPointError.

Device-resident loop (default, "nat" path: bases are distinct db1..db4 + self, dual kept
in the natural orientation between iterations, include/pfb_b200.h), per iteration, L levels:
    L   launches  Psi.dot(x_k) -> u                 (issued one iteration ahead, behind the primal step)
    1   launch    dual update, writes v only        (pfb_dual_update; sharded: pfb_dual_update_rec)
    L   launches  w_k = Psi.hdot(v_k)               (synthesis by linearity: 2 w_k - w_{k-1})
    3   launches  PSF convolution, the primal step x = P+(xp - tau(2w_k - w_{k-1} + conv - data))
                  and the eps terms being the epilogue of its last pass (pfb_hess_apply_primal;
                  otherwise pfb_hess_apply + pfb_pd_primal_lin)
and one 16-byte device->host read on a side stream for the convergence test.  The
reference's `xp[...] = x`, `vp[...] = v` copies are buffer swaps.  Fallbacks: "slab" (reference
coefficient layout, `2v - vp` fused into the synthesis loader) and "generic" (arbitrary psi /
psiH callables).  With pfb_b200.parallel enabled the bands are sharded over ranks: the
cross-band sum of the dual update is carried by a recursion that needs the all-reduce of one
band-summed IMAGE per iteration (hidden behind the PSF convolution); `prox_21`-style l2 norms
and PFB_PAR_NOREC=1 use the split partial / all-reduce(slab) / apply kernels; the eps terms
are all-reduced.
"""
import math
import os

import numpy as np
import torch

from .. import _dev, _lib, parallel
from ..operators.psi import Psi
from ..operators.psf import _convolve
from ..prox import _common as _prox
from ..utils.misc import norm_diff_terms
from . import _vec


class Grad21:
    """grad21 = lambda x: psf_convolve(x) - data  (pfb/workers/spotless.py:260)
    as an object, so the solver can fuse the subtraction into the primal
    update instead of materialising the gradient."""

    def __init__(self, psf_convolve, data):
        self.psf_convolve = psf_convolve
        self.data = data
        self._data_dev = None

    def data_dev(self):
        if self._data_dev is None:
            self._data_dev = _dev.to_dev(self.data)
        return self._data_dev

    def bound_psf(self):
        """(psfhat, lastsize) when psf_convolve is the usual
        partial(psf_convolve_cube, xpad, xhat, xout, psfhat, lastsize) of
        pfb/workers/spotless.py:182 over THIS package's operator, else None.
        Lets the solver convolve straight into its own device buffer instead
        of bouncing through the caller's (possibly host) xout every iteration."""
        from functools import partial as _partial
        from ..operators import psf as _psf
        f = self.psf_convolve
        if (isinstance(f, _partial) and f.func in (_psf.psf_convolve_cube, _psf.psf_convolve_slice)
                and len(f.args) == 5 and not f.keywords.get("x")):
            return f.args[3], f.args[4]
        return None

    def __call__(self, x):
        conv = self.psf_convolve(x)
        if isinstance(conv, np.ndarray):
            return conv - (self.data if isinstance(self.data, np.ndarray) else self.data.cpu().numpy())
        return conv - self.data_dev()


def _is_psi_pair(psi, psiH):
    a, b = getattr(psi, "__self__", None), getattr(psiH, "__self__", None)
    return (isinstance(a, Psi) and a is b and getattr(psi, "__func__", None) is Psi.dot
            and getattr(psiH, "__func__", None) is Psi.hdot)


def _off(t, nbytes):
    import ctypes
    return ctypes.c_void_p(t.data_ptr() + nbytes)


def _chunks(nbasis):
    """Basis slabs are exchanged in groups: more groups pipeline the all-reduce
    with the partial / apply kernels of the other groups, fewer groups make
    larger (more bandwidth-efficient) messages.  Measured (profiles/scale_probe2.sh):
    8 x B200: 5 x 68 MB cost 1.40 ms, 1 x 339 MB 0.88 ms, and the step is fastest
    with one group (435 / 424 / 396 it/s for 1 / 2 / 5); 2 x B200: 185 / 194 / 190.
    Default: 2 groups on 2 ranks, 1 beyond; PFB_PAR_CHUNKS overrides."""
    dflt = "2" if parallel.world_size() <= 2 else "1"
    k = max(1, min(nbasis, int(os.environ.get("PFB_PAR_CHUNKS", dflt))))
    edges = [round(i * nbasis / k) for i in range(k + 1)]
    return [(edges[i], edges[i + 1]) for i in range(k) if edges[i + 1] > edges[i]]


def _dual_begin(vp, v, sigma, n, bsum):
    """Band-sharded dual update, first half: local band sums of vp + sigma*v, one
    group of basis slabs at a time, each followed by its asynchronous all-reduce
    (NCCL's own stream).  Returns the work handles.  v is not modified, so
    whatever the caller launches next (the PSF convolution of the primal step)
    overlaps the exchange."""
    L = _lib.lib()
    nb, nbasis = v.shape[0], v.shape[1]
    nslab = n // nbasis
    esz = v.element_size()
    works = []
    for b0, b1 in _chunks(nbasis):
        off = b0 * nslab * esz
        _lib.check(L.pfb_dual_partial(_off(vp, off), _off(v, off), _off(bsum, off), nb, (b1 - b0) * nslab, n,
                                      float(sigma), _lib.BAND_SUM, _dev.dcode(v), _dev.stream()))
        works.append(parallel.allreduce_sum_async(bsum.view(-1)[b0 * nslab:b1 * nslab]))
    return works


def _dual_end(works, vp, v, lam, sigma, w, n, bsum, write_u=True):
    """Second half: v <- (vp + sigma*v) * shrink(all-reduced sums); vp <- 2v - vp."""
    L = _lib.lib()
    nb, nbasis = v.shape[0], v.shape[1]
    nslab = n // nbasis
    esz = v.element_size()
    for (b0, b1), work in zip(_chunks(nbasis), works):
        off = b0 * nslab * esz
        if work is not None:
            work.wait()
        _lib.check(L.pfb_dual_apply(_off(vp, off), _off(v, off), _off(bsum, off), _off(w, off),
                                    _off(vp, off) if write_u else None, nb, (b1 - b0) * nslab, n, float(lam),
                                    float(sigma), _lib.BAND_SUM, _dev.dcode(v), _dev.stream()))


def _dual_step(vp, v, lam, sigma, w, n, bsum, write_u=True):
    """v <- dual update(vp, v); if write_u: vp <- 2 v - vp (else the caller
    fuses 2v - vp into the Psi.hdot loader and vp is left untouched).
    Sharded: split around the all-reduce of the band-sum slab."""
    if parallel.active():
        _dual_end(_dual_begin(vp, v, sigma, n, bsum), vp, v, lam, sigma, w, n, bsum, write_u)
    else:
        _lib.check(_lib.lib().pfb_dual_update(_dev.ptr(vp), _dev.ptr(v), _dev.ptr(w),
                                              _dev.ptr(vp) if write_u else None, v.shape[0], n,
                                              float(lam), float(sigma), _lib.BAND_SUM, _dev.dcode(v),
                                              _dev.stream()))


def _band_sum(t, out):
    """out = sum over the local bands of t (nb, ...) -- pfb_dual_partial with sigma = 0."""
    nb = t.shape[0]
    m = t.numel() // nb
    _lib.check(_lib.lib().pfb_dual_partial(_dev.ptr(t), _dev.ptr(t), _dev.ptr(out), nb, m, m, 0.0,
                                           _lib.BAND_SUM, _dev.dcode(t), _dev.stream()))
    return out


def _dual_rec(vp, v, vbar, ubar, w, n, lam, sigma):
    """pfb_dual_update_rec: local dual update from the recursively carried band sum."""
    _lib.check(_lib.lib().pfb_dual_update_rec(_dev.ptr(vp), _dev.ptr(v), _dev.ptr(vbar), _dev.ptr(ubar),
                                              _dev.ptr(w), v.shape[0], n, float(lam), float(sigma),
                                              _dev.dcode(v), _dev.stream()))


def _dual_factor(vbar, ubar, w, f, n, lam, sigma):
    """pfb_dual_factor: S = vbar + sigma ubar; f = shrink factor; vbar <- f S."""
    _lib.check(_lib.lib().pfb_dual_factor(_dev.ptr(vbar), _dev.ptr(ubar), _dev.ptr(w), _dev.ptr(f), n,
                                          float(lam), float(sigma), _dev.dcode(vbar), _dev.stream()))


def _gap_index(psi_obj, device):
    """Flat indices (into nbasis*Nymax*Nxmax) of the slab cells Psi.dot never
    writes.  Found by running a 1-band analysis over a NaN-filled slab once."""
    idx = getattr(psi_obj, "_gap_idx", None)
    if idx is None:
        one = Psi(1, psi_obj.nx, psi_obj.ny, psi_obj.bases, psi_obj.nlevel, 1)
        a = torch.full((1, one.nbasis, one.Nymax, one.Nxmax), float("nan"), dtype=torch.float64,
                       device=device)
        one.dot(torch.zeros((1, one.nx, one.ny), dtype=torch.float64, device=device), a)
        idx = torch.nonzero(torch.isnan(a.view(-1))).view(-1)
        psi_obj._gap_idx = idx
    return idx


def primal_dual_optimised(x, v, lam, psiH, psi, L, prox, l1weight, reweighter, grad,
                          nu=1.0, sigma=None, mask=None, tol=1e-5, maxit=1000,
                          positivity=1, report_freq=10, gamma=1.0, verbosity=1,
                          maxreweight=50):
    xa = _dev.to_dev(x, copy=_dev._EMULATED and isinstance(x, np.ndarray))
    va = _dev.to_dev(v, copy=_dev._EMULATED and isinstance(v, np.ndarray))
    if xa.dtype != va.dtype:
        raise TypeError("x and v must have the same precision")
    xcur, xprev = xa, xa.clone()
    vcur, vprev = va, None
    xout = torch.zeros_like(xa)
    n = va.numel() // va.shape[0]
    w = _prox._weight_dev(l1weight, va, tuple(va.shape[1:]))
    bsum = torch.empty(va.shape[1:], dtype=va.dtype, device=va.device) if parallel.active() else None
    terms = _dev.scalars(2)

    if sigma is None:
        sigma = L / (2.0 * gamma) / nu
    tau = 0.9 / (L / (2.0 * gamma) + sigma * nu**2)

    # buffer swap instead of `vp[...] = v` needs the never-written slab cells of
    # both dual buffers to agree; they do whenever they are zero (always, for a
    # dual that started at zero).  Otherwise they are patched after each swap.
    fast = _is_psi_pair(psi, psiH)
    gap_idx = None
    if fast:
        gi = _gap_index(psi.__self__, va.device)
        if gi.numel() and bool(torch.any(va.view(va.shape[0], -1)[:, gi] != 0)):
            gap_idx = gi
    # Natural-orientation dual (include/pfb_b200.h): the solver owns v for the
    # whole loop, so it is converted once here and once on exit, and neither
    # Psi.dot nor Psi.hdot has to transpose.  The dual update is element-wise.
    nat = (fast and gap_idx is None and psi.__self__.nat_stride(_dev.dcode(va)) > 0
           and xa.data_ptr() % 16 == 0 and (xa[0].numel() * xa.element_size()) % 16 == 0
           and not os.environ.get("PFB_PD_SLAB"))
    # which loop ran, for the tests: "nat" | "slab" (fused loader) | "generic"
    primal_dual_optimised.last_path = "nat" if nat else ("slab" if fast else "generic")
    if nat:
        P = psi.__self__
        vcur = P.to_nat(va)
        vprev = vcur.clone()
        w = P.to_nat(w).contiguous()
        n = vcur.numel() // vcur.shape[0]
        if bsum is not None:
            bsum = torch.empty(vcur.shape[1:], dtype=va.dtype, device=va.device)
    else:
        vprev = va.clone()
    fused_grad = isinstance(grad, Grad21)
    bound = grad.bound_psf() if fused_grad else None
    conv_buf = torch.empty_like(xa) if bound is not None else None
    data = grad.data_dev() if fused_grad else None
    if data is not None and data.dtype != xa.dtype:
        data = data.to(xa.dtype)
    pos_kernel = positivity if not (positivity == 2 and parallel.active()) else 0

    # Convergence test without draining the GPU: the two eps terms are copied to
    # pinned host memory on a side stream right behind the primal kernel, while
    # the compute stream already runs the analysis Psi(x) of the next iteration
    # (valid whatever the test decides: on a break its output buffer is the dead
    # one).  Only the side stream is synchronised.
    ahead = nat and not _dev._EMULATED and not os.environ.get("PFB_PD_NOAHEAD")
    # Lagged test (no reweighter): eps of iteration k-1 is read after iteration k has been queued,
    # so the host runs a whole iteration ahead of the GPU (at 8 GPUs the per-iteration read left
    # 0.3 of 1.8 ms idle).  A late "converged at k-1" needs no roll-back: it is taken before
    # iteration k's look-ahead analysis, while x_{k-1} and v_{k-1} still sit in xprev / vprev.
    lagged = (ahead and pos_kernel == positivity and reweighter is None
              and not os.environ.get("PFB_PD_NOLAG"))
    if ahead:
        side = torch.cuda.Stream()
        host_terms2 = [torch.empty(2, dtype=torch.float64, pin_memory=True) for _ in range(2)]
        host_terms = host_terms2[0]
        terms2 = [terms, _dev.scalars(2)]
        done2 = [torch.cuda.Event(), torch.cuda.Event()]
        ev = torch.cuda.Event()
    have_dot = False            # Psi(xprev) -> vcur was already issued by the previous iteration
    # Synthesis by linearity: Psi(2 v_k - v_{k-1}) = 2 w_k - w_{k-1} with w_k = Psi(v_k).  The
    # dual kernel then writes v only (no 2v - vp: one pass over the coefficient cube less per
    # iteration) and the primal kernel forms 2 w_k - w_{k-1} from two image-sized arrays.
    lin = nat and not os.environ.get("PFB_PD_NOLIN")
    wprev = None
    if lin:
        wprev = torch.empty_like(xa)
        P.hdot_nat(vprev, wprev)            # w_{-1} = Psi(v_0)
    # PSF convolution with the primal step as the epilogue of its last pass
    # (pfb_hess_apply_primal): no convolution result in memory, no separate primal kernel
    fplan = None
    # Band-sharded ranks take it in two halves (apply_primal_begin / _end): the passes that need
    # xp only are issued where the separate convolution was (they hide the exchange of the
    # band-sum image), the last pass + primal step once the synthesis is there.
    shard_ok = (not parallel.active()) or (nat and not os.environ.get("PFB_PAR_NOREC")
                                             and not os.environ.get("PFB_PAR_NOFUSE"))
    if (lin and bound is not None and data is not None and positivity in (0, 1)
            and shard_ok and xa.shape[-1] % 2 == 0 and xa.data_ptr() % (2 * xa.element_size()) == 0
            and not os.environ.get("PFB_PD_NOFUSE")):
        want = "complex64" if xa.dtype == torch.float32 else "complex128"
        if str(bound[0].dtype).endswith(want):
            from ..operators.psf import get_plan
            cand = get_plan(bound[0], bound[1], tuple(xa.shape), _dev.dcode(xa))
            if cand.can_fuse_primal():
                fplan = cand
    primal_dual_optimised.last_fused = fplan is not None
    # Band-sharded ranks, 21m prox: the cross-band sum of the dual update is carried by a
    # recursion (include/pfb_b200.h, pfb_dual_update_rec) that needs the all-reduce of ONE image,
    # sum_b xp_b, per iteration instead of the (nbasis x larger) coefficient slab; every rank
    # analyses that image itself.  The exchange is issued right behind the primal step and hidden
    # by the next PSF convolution.  The carried sum is re-synchronised exactly every
    # PFB_PAR_RESYNC (default 64) iterations to bound rounding drift.
    # The same recursion lets the per-band half of the dual update ride in the ANALYSIS kernels'
    # epilogue (pfb_psi_dot_nat_dual): the band-independent half (S, the shrink factor f, the new
    # carried sum) is one small kernel over a band-shaped block, and v_b = (vp_b + sigma Psi^T x_b) f
    # is stored straight from the analysis registers -- the analysed coefficients never exist in
    # memory (5.4 GB less traffic per iteration at 8 x 4096^2).  Opt-in (PFB_PD_FUSEDUAL=1): the
    # analysis walks are instruction-issue bound, so the extra loads cost more there than the
    # bytes save in the 0.93-of-peak dual kernel: 2.75 ms (fused walks 2.26 incl. the one-band
    # analysis of the band sum, factor kernel 0.36, band sum 0.13) against 1.08 + 1.41 = 2.49 ms
    # per iteration at 8 x 4096^2 fp32, 140.1 vs 145.6 it/s (profiles/r02/r02o_*).  Numerically it
    # is the rec path: over the reference's 450 iterations every fp32 path stays 1.1e-6 / 2.8e-6
    # from the fp64 trajectory (profiles/r02/r02m_sharded_drift_450its.jsonl).
    norec = parallel.active() and os.environ.get("PFB_PAR_NOREC")
    fusedual = bool(nat and lin and not norec and os.environ.get("PFB_PD_FUSEDUAL"))
    rec = nat and lin and parallel.active() and not norec and not fusedual
    primal_dual_optimised.last_rec = bool(rec or (fusedual and parallel.active()))
    primal_dual_optimised.last_fusedual = bool(fusedual)
    xbar_work = None
    have_xbar = False
    if rec or fusedual:
        P1 = getattr(P, "_one_band", None)
        if P1 is None:
            P1 = P._one_band = Psi(1, P.nx, P.ny, P.bases, P.nlevel, 1,
                                   dtype=np.float32 if va.dtype == torch.float32 else np.float64)
        assert P1.nat_stride(_dev.dcode(va)) == vcur.shape[2]
        xbar = torch.empty_like(xa[0])
        ubar = torch.zeros((1,) + tuple(vcur.shape[1:]), dtype=va.dtype, device=va.device)
        vbar = torch.empty(tuple(vcur.shape[1:]), dtype=va.dtype, device=va.device)
        parallel.allreduce_sum(_band_sum(vprev, vbar))
        resync = int(os.environ.get("PFB_PAR_RESYNC", "64"))
        if fusedual:
            fbuf = torch.empty_like(vbar)

    eps = 1.0
    numreweight = 0
    k = -1
    broke = False
    late = False

    def eps_of(slot):
        done2[slot].synchronize()
        num_, den_ = host_terms2[slot].tolist()
        if den_ == 0.0:
            raise FloatingPointError("primal_dual: model is identically zero "
                                     "(the reference drops into pdb here)")
        e_ = math.sqrt(num_ / (1e-12 + den_))
        if math.isnan(e_) or math.isinf(e_):
            raise FloatingPointError("primal_dual: eps is NaN/Inf")
        return e_
    for k in range(maxit):
        if lagged:
            terms = terms2[k & 1]

        def gradient():
            if bound is not None:
                return _convolve(bound[0], bound[1], xprev, conv_buf)
            return _dev.to_dev(grad.psf_convolve(xprev) if fused_grad else grad(xprev))
        g = None
        begun = False
        if nat:
            # not lin: the dual kernel leaves 2v - vp in the (dead) vp buffer: one coefficient
            # set for the synthesis to read, which is instruction-issue bound
            if (rec or fusedual) and not have_xbar:
                xbar_work = parallel.allreduce_sum_async(_band_sum(xprev, xbar))
            have_xbar = False
            if fusedual:
                if parallel.active():
                    if fplan is None:
                        g = gradient()               # hides the exchange of xbar
                    else:
                        fplan.set_dynamic(True)
                        fplan.apply_primal_begin(xprev)
                        begun = True
                    if xbar_work is not None:
                        xbar_work.wait()
                xbar_work = None
                P1.dot_nat(xbar[None], ubar)
                if resync > 0 and k > 0 and k % resync == 0:
                    parallel.allreduce_sum(_band_sum(vprev, vbar))
                _dual_factor(vbar, ubar, w, fbuf, n, lam, sigma)
                P.dot_nat_dual(xprev, vprev, fbuf, sigma, vcur)
                have_dot = False
            elif not have_dot:
                P.dot_nat(xprev, vcur)
            have_dot = False
            if fusedual:
                pass
            elif rec:
                if fplan is None:
                    g = gradient()               # hides the exchange of xbar
                else:
                    fplan.set_dynamic(True)
                    fplan.apply_primal_begin(xprev)
                    begun = True
                if xbar_work is not None:
                    xbar_work.wait()
                xbar_work = None
                P1.dot_nat(xbar[None], ubar)
                if resync > 0 and k > 0 and k % resync == 0:
                    parallel.allreduce_sum(_band_sum(vprev, vbar))
                _dual_rec(vprev, vcur, vbar, ubar, w, n, lam, sigma)
            elif parallel.active() and not os.environ.get("PFB_PAR_NOOVERLAP"):
                # the PSF convolution hides the all-reduce of the band sums
                works = _dual_begin(vprev, vcur, sigma, n, bsum)
                g = gradient()
                _dual_end(works, vprev, vcur, lam, sigma, w, n, bsum, write_u=not lin)
            else:
                _dual_step(vprev, vcur, lam, sigma, w, n, bsum, write_u=not lin)
            P.hdot_nat(vcur if lin else vprev, xout)
        elif fast:   # `vp = 2v - vp` (primal_dual.py:137) is fused into the synthesis loader
            psi(xprev, vcur)
            _dual_step(vprev, vcur, lam, sigma, w, n, bsum, write_u=False)
            psi.__self__.hdot(vcur, xout, alpha2=vprev)
        else:
            psi(xprev, vcur)
            _dual_step(vprev, vcur, lam, sigma, w, n, bsum)      # vprev now holds 2v - vp
            psiH(vprev, xout)
        if fplan is not None and begun:
            fplan.apply_primal_end(xprev, xout, wprev, data, tau, positivity, xcur, terms)
        elif fplan is not None:
            fplan.set_dynamic(parallel.active())
            fplan.apply_primal(xprev, xout, wprev, data, tau, positivity, xcur, terms)
        else:
            if g is None:
                g = gradient()
            if g.dtype != xa.dtype:
                g = g.to(xa.dtype)
            _vec.pd_primal(xcur, xprev, xout, g, data, tau, pos_kernel, terms, wprev=wprev)
        if lin:
            xout, wprev = wprev, xout
        if pos_kernel != positivity:   # positivity 2 across band-sharded ranks
            neg = (xcur <= 0).any(dim=0).to(torch.int32)
            parallel.allreduce_max(neg)
            xcur[:, neg.bool()] = 0
            norm_diff_terms(xcur, xprev, out=terms)
        if (rec or fusedual) and k + 1 < maxit:
            # next iteration's exchange: sum of the new model over the local bands
            xbar_work = parallel.allreduce_sum_async(_band_sum(xcur, xbar))
            have_xbar = True
        if lagged:
            ev.record()
            if k > 0:
                eps = eps_of((k - 1) & 1)
                if not (k - 1) % report_freq and verbosity > 1:
                    print(f"At iteration {k - 1} eps = {eps:.3e}")
                if eps < tol:                # converged one iteration ago: iteration k is discarded
                    late = True
                    break
            if k + 1 < maxit and not fusedual:
                P.dot_nat(xcur, vprev)
                have_dot = True
            with torch.cuda.stream(side):
                side.wait_event(ev)
                parallel.allreduce_sum(terms)
                host_terms2[k & 1].copy_(terms, non_blocking=True)
                done2[k & 1].record(side)
            # xp <- x, vp <- v as swaps; this iteration's eps is looked at during the next one
            xcur, xprev = xprev, xcur
            vcur, vprev = vprev, vcur
            continue
        if ahead and pos_kernel == positivity:
            ev.record()
            if k + 1 < maxit and not fusedual:
                P.dot_nat(xcur, vprev)       # next iteration: xprev = xcur, vcur = vprev (holds the dead 2v - vp)
                have_dot = True
            with torch.cuda.stream(side):
                side.wait_event(ev)
                parallel.allreduce_sum(terms)
                host_terms.copy_(terms, non_blocking=True)
            side.synchronize()
            num, den = host_terms.tolist()
        else:
            num, den = parallel.allreduce_sum(terms).tolist()
        if den == 0.0:
            raise FloatingPointError("primal_dual: model is identically zero "
                                     "(the reference drops into pdb here)")
        eps = math.sqrt(num / (1e-12 + den))
        if eps < tol:
            if reweighter is not None and numreweight < maxreweight:
                w = _prox._weight_dev(reweighter(xcur), va, tuple(va.shape[1:]))
                if nat:
                    w = P.to_nat(w).contiguous()
                numreweight += 1
            else:
                if numreweight >= maxreweight and verbosity:
                    print("Maximum reweighting steps reached")
                broke = True
                break
        # xp <- x, vp <- v  (pfb/opt/primal_dual.py:164-165) as swaps
        xcur, xprev = xprev, xcur
        if fast:
            vcur, vprev = vprev, vcur
            if gap_idx is not None:
                vcur.view(vcur.shape[0], -1)[:, gap_idx] = vprev.view(vprev.shape[0], -1)[:, gap_idx]
        else:
            vprev.copy_(vcur)
        if math.isnan(eps) or math.isinf(eps):
            raise FloatingPointError("primal_dual: eps is NaN/Inf")
        if not k % report_freq and verbosity > 1:
            print(f"At iteration {k} eps = {eps:.3e}")

    if lagged:
        if late:
            k -= 1                           # the iteration that met the tolerance
        elif k >= 0:
            eps = eps_of(k & 1)              # the last iteration's eps
    if xbar_work is not None:
        xbar_work.wait()
    if verbosity:
        if k == maxit - 1 and not broke and not (lagged and eps < tol):
            print(f"Max iters reached. eps = {eps:.3e}")
        else:
            print(f"Success, converged after {k} iterations")
    xfin = xcur if broke else xprev
    vfin = vcur if (broke or not fast) else vprev
    if nat:
        vfin = P.from_nat(vfin, va)
    return _dev.to_out(xfin, x), _dev.to_out(vfin, v)


def primal_dual(x, v, lam, psi, psiH, L, prox, grad, nu=1.0, sigma=None, mask=None,
                tol=1e-5, maxit=1000, minit=10, positivity=1, report_freq=10, gamma=1.0,
                verbosity=1):
    """pfb/opt/primal_dual.py:12-87, functional form: psiH(x) RETURNS coeffs,
    psi(v) RETURNS an image, prox(v, thr) RETURNS an array.  The operators are the
    caller's; everything between them runs on the fused vector kernels (pfb_axpby for
    vtilde / the Moreau step / 2v - vp, pfb_pd_primal for the primal step with its
    positivity projection and both norms), with x/xp and v/vp as buffer swaps.
    Single rank only (like the reference's non-distributed form)."""
    xd = _dev.to_dev(x, copy=True)
    vd = _dev.to_dev(v, copy=True)
    xp, vp = xd.clone(), vd.clone()
    vt, u = torch.empty_like(vd), torch.empty_like(vd)
    if sigma is None:
        sigma = L / (2.0 * gamma) / nu
    tau = 0.9 / (L / (2.0 * gamma) + sigma * nu**2)
    eps = 1.0
    k = 0
    terms = _dev.scalars(2)

    def dev_like(t, ref):
        t = _dev.to_dev(t)
        return t if t.dtype == ref.dtype else t.to(ref.dtype)
    while (eps > tol or k < minit) and k < maxit:
        # vtilde = v + sigma * psiH(xp)                                  (:50)
        _vec.axpby(vt, 1.0, vp, sigma, dev_like(psiH(xp), vd))
        # v = vtilde - sigma * prox(vtilde / sigma, lam / sigma)         (:51)
        _vec.axpby(u, 1.0 / sigma, vt, 0.0, vt)
        _vec.axpby(vd, 1.0, vt, -sigma, dev_like(prox(u, lam / sigma), vd))
        # x = xp - tau * (psi(2 v - vp) + grad(xp)); positivity; norms   (:54-64)
        _vec.axpby(u, 2.0, vd, -1.0, vp)
        _vec.pd_primal(xd, xp, dev_like(psi(u), xd), dev_like(grad(xp), xd), None, tau, positivity, terms)
        # eps = ||x - xp|| / ||x||
        num, den = terms.tolist()
        eps = math.sqrt(num) / math.sqrt(den) if den > 0 else float("nan")
        xd, xp = xp, xd          # xp <- x, vp <- v as swaps
        vd, vp = vp, vd
        if math.isnan(eps) or math.isinf(eps):
            raise FloatingPointError("primal_dual: eps is NaN/Inf")
        if not k % report_freq and verbosity > 1:
            print(f"At iteration {k} eps = {eps:.3e}")
        k += 1
    if verbosity:
        if k == maxit:
            print(f"Max iters reached. eps = {eps:.3e}")
        else:
            print(f"Success, converged after {k} iterations")
    return _dev.like(x, xp), _dev.like(v, vp)
