"""Kernel sources of pfb-clean_b200/csrc/prox.cu executed block-by-block on the
CPU emulator (tests/emu) and compared with the oracle.  Checks indexing,
vector/scalar paths, band loops and the two-stage reductions without a GPU."""
import numpy as np
import pytest

from oracle import ref
from emu.emulib import emu, ptr, dt, ok

pmp = pytest.mark.parametrize


def rel(a, b):
    return np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)


@pmp("dtype", [np.float32, np.float64])
@pmp("nband,shape", [(3, (2, 9, 12)), (1, (1, 5, 7)), (10, (2, 8, 8))])
@pmp("mode", [0, 1])
def test_prox_dual(dtype, nband, shape, mode):
    rng = np.random.default_rng(1)
    v = rng.standard_normal((nband,) + shape).astype(dtype)
    v[:, 0, 1, 2] = 0
    vp = rng.standard_normal(v.shape).astype(dtype)
    w = rng.random(shape).astype(dtype)
    n = int(np.prod(shape))
    # the shrink factor is evaluated with one division (1 - max(|r| - lam w, 0)/|r|): same
    # cancellation as the reference's three-division form, different rounding of it
    tol = 1e-12 if dtype == np.float64 else 2e-7
    L = emu()
    for lam, sigma in [(1.0, 1.0), (0.1, 75.0), (0.7, 2.5)]:
        res = np.full(v.shape, np.nan, dtype)
        ok(L.pfb_prox21(ptr(v), ptr(res), ptr(w), nband, n, lam, sigma, mode, dt(v), None))
        exp = np.zeros_like(v)
        (ref.prox_21_numba if mode else ref.prox_21m_numba)(v, exp, lam, sigma=sigma, weight=w)
        assert rel(res, exp) < tol
        # plain dual update
        v2 = v.copy()
        ok(L.pfb_dual_update(ptr(vp), ptr(v2), ptr(w), None, nband, n, lam, sigma, mode, dt(v), None))
        exp = v.copy()
        (ref.dual_update_numba_l2 if mode else ref.dual_update_numba)(vp, exp, lam, sigma=sigma, weight=w)
        assert rel(v2, exp) < tol
        # fused with u = 2v - vp written over vp
        v3, vp3 = v.copy(), vp.copy()
        ok(L.pfb_dual_update(ptr(vp3), ptr(v3), ptr(w), ptr(vp3), nband, n, lam, sigma, mode, dt(v), None))
        assert rel(v3, exp) < tol
        assert rel(vp3, 2 * exp - vp) < 10 * tol
        # split form (band-sharded): two "ranks" holding a band subset each
        if nband >= 2:
            h = nband // 2
            va, vb = v[:h].copy(), v[h:].copy()
            sa, sb = np.zeros(shape, dtype), np.zeros(shape, dtype)
            ok(L.pfb_dual_partial(ptr(vp[:h].copy()), ptr(va), ptr(sa), h, n, 0, sigma, mode, dt(v), None))
            ok(L.pfb_dual_partial(ptr(vp[h:].copy()), ptr(vb), ptr(sb), nband - h, n, 0, sigma, mode, dt(v), None))
            s = sa + sb
            ua, ub = np.zeros_like(va), np.zeros_like(vb)
            ok(L.pfb_dual_apply(ptr(vp[:h].copy()), ptr(va), ptr(s), ptr(w), ptr(ua), h, n, 0, lam, sigma, mode, dt(v), None))
            ok(L.pfb_dual_apply(ptr(vp[h:].copy()), ptr(vb), ptr(s), ptr(w), ptr(ub), nband - h, n, 0, lam, sigma, mode, dt(v), None))
            tol2 = tol if dtype == np.float64 else 5e-6   # fp32 slab sum
            assert rel(np.concatenate([va, vb]), exp) < tol2
            # recursive form (pfb_dual_update_rec): the ranks share only vbar = sum_b vp_b and
            # ubar = sum_b u_b (what Psi^T of the all-reduced band-sum image gives); afterwards
            # the carried sum equals sum_b v_b of the updated dual
            if mode == 0:
                vbar0 = vp.astype(np.float64).sum(axis=0).astype(dtype)
                ubar = v.astype(np.float64).sum(axis=0).astype(dtype)
                va, vb = v[:h].copy(), v[h:].copy()
                ba, bb = vbar0.copy(), vbar0.copy()
                ok(L.pfb_dual_update_rec(ptr(vp[:h].copy()), ptr(va), ptr(ba), ptr(ubar), ptr(w), h, n, lam, sigma, dt(v), None))
                ok(L.pfb_dual_update_rec(ptr(vp[h:].copy()), ptr(vb), ptr(bb), ptr(ubar), ptr(w), nband - h, n, lam, sigma, dt(v), None))
                assert rel(np.concatenate([va, vb]), exp) < tol2
                assert np.array_equal(ba, bb)
                assert rel(ba, exp.astype(np.float64).sum(axis=0)) < (1e-12 if dtype == np.float64 else 5e-6)
            assert rel(np.concatenate([ua, ub]), 2 * exp - vp) < 10 * tol2


@pmp("dtype", [np.float32, np.float64])
@pmp("n", [1, 255, 4096 + 3])
def test_vector_ops(dtype, n):
    rng = np.random.default_rng(2)
    L = emu()
    a, b, c, d = (rng.standard_normal(n).astype(dtype) for _ in range(4))
    out = np.zeros(n, dtype)
    ok(L.pfb_axpby(ptr(out), 2.0, ptr(a), -1.0, ptr(b), n, dt(a), None))
    assert np.array_equal(out, (2 * a - b).astype(dtype))
    res = np.zeros(4)
    ok(L.pfb_vdot(ptr(a), ptr(b), ptr(c), ptr(d), n, dt(a), ptr(res), None))
    tol = 1e-13
    assert abs(res[0] - np.vdot(a.astype(np.float64), b.astype(np.float64))) < tol * n
    assert abs(res[1] - np.vdot(c.astype(np.float64), d.astype(np.float64))) < tol * n
    ok(L.pfb_norm_diff(ptr(a), ptr(b), n, dt(a), ptr(res), None))
    eps = np.sqrt(res[0] / (1e-12 + res[1]))
    assert abs(eps / ref.norm_diff(a, b) - 1) < (1e-12 if dtype == np.float64 else 1e-6)


@pmp("dtype", [np.float32, np.float64])
@pmp("positivity", [0, 1, 2])
def test_pd_primal(dtype, positivity):
    rng = np.random.default_rng(3)
    L = emu()
    nband, npix = 3, 301
    xp, xout, conv, data = (rng.standard_normal((nband, npix)).astype(dtype) for _ in range(4))
    tau = 0.37
    x = np.full_like(xp, np.nan)
    res = np.zeros(2)
    ok(L.pfb_pd_primal(ptr(x), ptr(xp), ptr(xout), ptr(conv), ptr(data), tau, positivity,
                       nband, npix, dt(x), ptr(res), None))
    g = xout + (conv - data)
    exp = (xp - dtype(tau) * g).astype(dtype)
    if positivity == 1:
        exp[exp < 0] = 0
    elif positivity == 2:
        exp[:, np.any(exp <= 0, axis=0)] = 0
    assert rel(x, exp) < (1e-15 if dtype == np.float64 else 1e-7)
    eps = np.sqrt(res[0] / (1e-12 + res[1]))
    assert abs(eps / ref.norm_diff(exp, xp) - 1) < 1e-6


@pmp("dtype", [np.float32, np.float64])
def test_pcg_kernels(dtype):
    rng = np.random.default_rng(4)
    L = emu()
    n = 1000
    xp, rp, p, Ap = (rng.standard_normal(n).astype(dtype) for _ in range(4))
    x, r = np.zeros(n, dtype), np.zeros(n, dtype)
    res = np.zeros(3)
    alpha, minv = 0.3, 2.0
    ok(L.pfb_pcg_update(ptr(x), ptr(r), ptr(xp), ptr(rp), ptr(p), ptr(Ap), alpha, minv, n,
                        dt(x), ptr(res), None))
    ex, er = xp + dtype(alpha) * p, rp + dtype(alpha) * Ap
    tol = 1e-14 if dtype == np.float64 else 1e-6
    assert rel(x, ex) < tol and rel(r, er) < tol
    assert abs(res[0] / np.vdot(er.astype(np.float64), minv * er.astype(np.float64)) - 1) < tol
    assert abs(np.sqrt(res[1] / (1e-12 + res[2])) / ref.norm_diff(ex, xp) - 1) < 10 * tol
    pp = p.copy()
    ok(L.pfb_pcg_direction(ptr(pp), ptr(r), 0.5, minv, n, dt(x), ptr(res), None))
    assert rel(pp, dtype(0.5) * p - dtype(minv) * r) < tol
    assert res[0] == np.count_nonzero(pp)
