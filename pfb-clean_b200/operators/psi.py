"""Drop-in for pfb/operators/psi.py: class Psi (:269-310) with
dot = analysis (image -> coeffs) and hdot = synthesis (coeffs -> image), on the
tiled DWT/IDWT kernels of csrc/wavelet.cu.  Coefficient layout, sizes
(Nxmax, Nymax), the un-normalised convention hdot(dot(x)) = nbasis*x and the
"cells outside the pyramid are never written" behaviour are the reference's.
"""
import ctypes
import weakref

import numpy as np
import torch

from .. import _dev, _lib

_BASIS_ID = {"self": 0, "db1": 1, "db2": 2, "db3": 3, "db4": 4, "db5": 5}


class _PsiBand(object):
    """Per-band bookkeeping view standing in for the reference's psi_band jitclass
    (pfb/operators/psi.py:126-185): sizes and pyramid tables, no arithmetic of its own."""

    def __init__(self, psi):
        self._psi = weakref.ref(psi)     # no cycle: the plan is destroyed with the last user reference
        self.nx, self.ny = psi.nx, psi.ny
        self.bases, self.nbasis, self.nlevel = tuple(psi.bases), psi.nbasis, psi.nlevel
        self.Nxmax, self.Nymax = psi.Nxmax, psi.Nymax

    def tables(self, basis):
        return self._psi().tables(basis)


class Psi(object):
    def __init__(self, nband, nx, ny, bases, nlevel, nthreads=1, dtype=None):
        """Same signature as the reference (nthreads accepted and ignored).
        `dtype` (np.float32/np.float64) may be given up front; otherwise the
        plan is created on first use from the array precision."""
        self.nband, self.nx, self.ny = int(nband), int(nx), int(ny)
        self.bases = list(bases)
        self.nbasis = len(self.bases)
        self.nlevel = int(nlevel)
        self.nthreads = nthreads
        self._plans = {}
        self._destroy = None
        for b in self.bases:
            if b not in _BASIS_ID:
                raise ValueError(f"unknown basis {b!r}: expected 'self' or 'db1'..'db5'")
        # sizes do not depend on precision: build the default plan now so that
        # ValueError (nlevel too large) is raised at construction like the reference
        code = _lib.F32 if dtype is not None and np.dtype(dtype) == np.float32 else _lib.F64
        self._plan(code)
        nxm, nym = ctypes.c_int(), ctypes.c_int()
        _lib.check(_lib.lib().pfb_psi_plan_query(self._plans[code], ctypes.byref(nxm), ctypes.byref(nym)))
        self.Nxmax, self.Nymax = nxm.value, nym.value
        # Psi.psib (pfb/operators/psi.py:272-274): one bookkeeping object per band.  The bands
        # share one device plan here, so the entries are views of the same tables.
        self.psib = [_PsiBand(self) for _ in range(self.nband)]

    def _plan(self, code):
        p = self._plans.get(code)
        if p is None:
            L = _lib.lib()
            ids = (ctypes.c_int * self.nbasis)(*[_BASIS_ID[b] for b in self.bases])
            p = ctypes.c_void_p()
            rc = L.pfb_psi_plan_create(ctypes.byref(p), self.nband, self.nx, self.ny,
                                       self.nbasis, ids, self.nlevel, code)
            if rc in (2, 3):    # odd size / nlevel: the reference raises ValueError
                raise ValueError(L.pfb_last_error().decode())
            _lib.check(rc)
            self._plans[code] = p
            self._destroy = L.pfb_psi_plan_destroy
        return p

    def tables(self, basis):
        """Pyramid bookkeeping of one basis (ix/iy/sx/sy/spx/spy of
        psi_band_maker, pfb/operators/psi.py:49-94) as a dict."""
        L = self.nlevel
        tabs = (ctypes.c_int * (6 * L))()
        ntx, nty = ctypes.c_int(), ctypes.c_int()
        p = next(iter(self._plans.values()))
        _lib.check(_lib.lib().pfb_psi_plan_tables(p, basis, tabs, ctypes.byref(ntx), ctypes.byref(nty)))
        t = np.array(tabs[:]).reshape(6, L)
        return dict(highx=t[0], highy=t[1], sx=t[2], sy=t[3], spx=t[4], spy=t[5],
                    Ntotx=ntx.value, Ntoty=nty.value)

    def _check(self, x, alpha):
        if tuple(x.shape) != (self.nband, self.nx, self.ny):
            raise ValueError(f"image shape {tuple(x.shape)} != {(self.nband, self.nx, self.ny)}")
        if tuple(alpha.shape) != (self.nband, self.nbasis, self.Nymax, self.Nxmax):
            raise ValueError(f"coefficient shape {tuple(alpha.shape)} != "
                             f"{(self.nband, self.nbasis, self.Nymax, self.Nxmax)}")

    def dot(self, x, alphao):
        """image to coeffs (pfb/operators/psi.py:284-296); alphao written in
        place, cells outside the pyramid blocks keep their previous content."""
        self._check(x, alphao)
        xd = _dev.to_dev(x)
        # gaps must survive: a host alphao has to be uploaded first
        ad = _dev.to_dev(alphao)
        if xd.dtype != ad.dtype:
            raise TypeError("x and alphao must have the same precision")
        _lib.check(_lib.lib().pfb_psi_dot(self._plan(_dev.dcode(xd)), _dev.ptr(xd), _dev.ptr(ad),
                                          _dev.stream()))
        _dev.to_out(ad, alphao)

    def hdot(self, alpha, xo, alpha2=None):
        """coeffs to image (pfb/operators/psi.py:298-310); xo overwritten.
        Extension: with alpha2 the coefficients used are 2*alpha - alpha2."""
        self._check(xo, alpha)
        ad = _dev.to_dev(alpha)
        a2 = _dev.to_dev(alpha2) if alpha2 is not None else None
        od = _dev.out_dev(xo)
        if od.dtype != ad.dtype:
            raise TypeError("alpha and xo must have the same precision")
        _lib.check(_lib.lib().pfb_psi_hdot(self._plan(_dev.dcode(ad)), _dev.ptr(ad), _dev.ptr(a2),
                                           _dev.ptr(od), _dev.stream()))
        _dev.to_out(od, xo)

    # ---- natural-orientation layout (include/pfb_b200.h, pfb_psi_nat_stride) ----
    # Device tensors only; used by primal_dual_optimised to keep its dual
    # variable un-transposed between iterations.
    def nat_stride(self, code=None):
        """Elements per (band, basis) block, 0 when the bases do not support it."""
        n = ctypes.c_longlong()
        p = self._plan(code) if code is not None else next(iter(self._plans.values()))
        _lib.check(_lib.lib().pfb_psi_nat_stride(p, ctypes.byref(n)))
        return n.value

    def to_nat(self, alpha):
        """(nb, nbasis, Nymax, Nxmax) or (nbasis, Nymax, Nxmax) slab tensor -> natural layout."""
        nb = alpha.shape[0] if alpha.dim() == 4 else 1
        code = _dev.dcode(alpha)
        out = torch.empty((nb, self.nbasis, self.nat_stride(code)), dtype=alpha.dtype, device=alpha.device)
        _lib.check(_lib.lib().pfb_psi_to_nat(self._plan(code), nb, _dev.ptr(alpha), _dev.ptr(out), _dev.stream()))
        return out if alpha.dim() == 4 else out[0]

    def from_nat(self, nat, alpha):
        nb = alpha.shape[0] if alpha.dim() == 4 else 1
        _lib.check(_lib.lib().pfb_psi_from_nat(self._plan(_dev.dcode(alpha)), nb, _dev.ptr(nat),
                                               _dev.ptr(alpha), _dev.stream()))
        return alpha

    def dot_nat(self, x, nat):
        _lib.check(_lib.lib().pfb_psi_dot_nat(self._plan(_dev.dcode(x)), _dev.ptr(x), _dev.ptr(nat),
                                              _dev.stream()))

    def dot_nat_dual(self, x, vp_nat, f_nat, sigma, v_nat):
        """v_nat = (vp_nat + sigma * Psi^T x) * f_nat (f_nat: one band-shaped block); see
        pfb_psi_dot_nat_dual in include/pfb_b200.h."""
        _lib.check(_lib.lib().pfb_psi_dot_nat_dual(self._plan(_dev.dcode(x)), _dev.ptr(x), _dev.ptr(vp_nat),
                                                   _dev.ptr(f_nat), float(sigma), _dev.ptr(v_nat), _dev.stream()))

    def hdot_nat(self, nat, xo, nat2=None):
        _lib.check(_lib.lib().pfb_psi_hdot_nat(self._plan(_dev.dcode(nat)), _dev.ptr(nat), _dev.ptr(nat2),
                                               _dev.ptr(xo), _dev.stream()))

    def __del__(self):
        if self._destroy is not None:
            for p in self._plans.values():
                self._destroy(p)
            self._plans = {}
