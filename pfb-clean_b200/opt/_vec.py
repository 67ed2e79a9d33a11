"""Thin wrappers over the fused vector kernels of csrc/prox.cu."""
from .. import _dev, _lib


def axpby(out, a, x, b, y):
    _lib.check(_lib.lib().pfb_axpby(_dev.ptr(out), float(a), _dev.ptr(x), float(b), _dev.ptr(y),
                                    out.numel(), _dev.dcode(out), _dev.stream()))
    return out


def vdot2(a, b, c=None, d=None, out=None):
    """device {sum a*b, sum c*d} (fp64)."""
    if out is None:
        out = _dev.scalars(2)
    _lib.check(_lib.lib().pfb_vdot(_dev.ptr(a), _dev.ptr(b), _dev.ptr(c), _dev.ptr(d), a.numel(),
                                   _dev.dcode(a), _dev.ptr(out), _dev.stream()))
    return out


def pcg_update(x, r, xp, rp, p, Ap, alpha, minv, out):
    _lib.check(_lib.lib().pfb_pcg_update(_dev.ptr(x), _dev.ptr(r), _dev.ptr(xp), _dev.ptr(rp),
                                         _dev.ptr(p), _dev.ptr(Ap), float(alpha), float(minv),
                                         x.numel(), _dev.dcode(x), _dev.ptr(out), _dev.stream()))
    return out


def pcg_direction(p, r, beta, minv, out):
    _lib.check(_lib.lib().pfb_pcg_direction(_dev.ptr(p), _dev.ptr(r), float(beta), float(minv),
                                            p.numel(), _dev.dcode(p), _dev.ptr(out), _dev.stream()))
    return out


def pd_primal(x, xp, xout, conv, data, tau, positivity, out, wprev=None):
    """wprev given: xout is w = Psi(v) and the kernel forms 2 w - wprev itself."""
    nband = x.shape[0]
    npix = x.numel() // nband
    if wprev is not None:
        _lib.check(_lib.lib().pfb_pd_primal_lin(_dev.ptr(x), _dev.ptr(xp), _dev.ptr(xout), _dev.ptr(wprev),
                                                _dev.ptr(conv), _dev.ptr(data), float(tau), int(positivity),
                                                nband, npix, _dev.dcode(x), _dev.ptr(out), _dev.stream()))
        return out
    _lib.check(_lib.lib().pfb_pd_primal(_dev.ptr(x), _dev.ptr(xp), _dev.ptr(xout), _dev.ptr(conv),
                                        _dev.ptr(data), float(tau), int(positivity), nband, npix,
                                        _dev.dcode(x), _dev.ptr(out), _dev.stream()))
    return out
