"""Generate golden input/output vectors from the UNMODIFIED reference.

Run in the build container only (needs /root/reference):
    python tests/golden/make_golden.py
The reference modules are imported from /root/reference behind the third-party
shims of oracle/shims.py; nothing is copied.  norm_diff and l1reweight_func
live in pfb/utils/misc.py, which cannot be imported whole (it needs ~10 absent
packages), so their *source text* is cut out of that file with ast and executed
as is.  Outputs: small .npz fixtures in tests/golden/.
"""
import ast
import os
import sys
from functools import partial

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import shims  # noqa: E402

shims.install()

from pfb.operators.psi import Psi  # noqa: E402
from pfb.operators.psf import psf_convolve_cube, psf_convolve_slice  # noqa: E402
from pfb.operators.hessian import hessian_psf_cube, _hessian_psf_slice  # noqa: E402
from pfb.prox.prox_21m import (prox_21m, prox_21m_numba, dual_update,  # noqa: E402
                               dual_update_numba)
from pfb.prox import prox_21 as p21  # noqa: E402
from pfb.opt.pcg import pcg, _pcg_psf_impl  # noqa: E402
from pfb.opt.primal_dual import primal_dual_optimised, primal_dual  # noqa: E402
from pfb.opt.power_method import power_method  # noqa: E402
from pfb.wavelets import dwt2d, idwt2d  # noqa: E402


def misc_functions():
    """exec norm_diff / l1reweight_func exactly as written in misc.py."""
    path = os.path.join(shims.REFERENCE, "pfb", "utils", "misc.py")
    src = open(path).read()
    tree = ast.parse(src)
    want = {"norm_diff", "norm_diff_impl", "nb_norm_diff_impl", "l1reweight_func"}
    lines = src.splitlines()
    chunks = []
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in want:
            start = min([node.lineno] + [d.lineno for d in node.decorator_list])
            chunks.append("\n".join(lines[start - 1:node.end_lineno]))
    ns = {}
    exec("import numpy as np\nfrom numba import njit\n"
         "from numba.extending import overload\n"
         "JIT_OPTIONS = {'nogil': True, 'cache': False}\n", ns)
    exec("\n\n".join(chunks), ns)
    return ns["norm_diff"], ns["l1reweight_func"]


norm_diff, l1reweight_func = misc_functions()
import pfb.utils.misc as _misc  # noqa: E402  (the shim module)
_misc.norm_diff = norm_diff
import pfb.opt.pcg as _pcgmod  # noqa: E402
import pfb.opt.primal_dual as _pdmod  # noqa: E402
_pcgmod.norm_diff = norm_diff
_pdmod.norm_diff = norm_diff


def save(name, **arrs):
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrs)
    print("wrote", name, {k: getattr(v, "shape", None) for k, v in arrs.items()})


def synth_psf(rng, nband, nxp, nyp, dtype):
    """Gaussian PSF with weak sidelobes, peak at (nxp//2, nyp//2), sum of band
    peaks = 1 (pfb/workers/spotless.py:148-149)."""
    ii, jj = np.meshgrid(np.arange(nxp) - nxp // 2, np.arange(nyp) - nyp // 2,
                         indexing="ij")
    psf = np.zeros((nband, nxp, nyp))
    for b in range(nband):
        s = 1.5 * (1 + 0.05 * b)
        r2 = ii**2 + jj**2
        psf[b] = np.exp(-r2 / (2 * s * s)) + 0.02 * np.cos(0.3 * ii + 0.1 * b) * \
            np.cos(0.2 * jj) * np.exp(-r2 / (2 * (6 * s)**2))
        psf[b] /= psf[b].max() * nband
    psfhat = np.fft.rfft2(np.fft.ifftshift(psf, axes=(1, 2)), axes=(1, 2))
    cdt = np.complex64 if dtype == np.float32 else np.complex128
    return psf.astype(dtype), psfhat.astype(cdt)


def gen_psi():
    bases = ['self', 'db1', 'db2', 'db3', 'db4', 'db5']
    for tag, (nband, nx, ny, nlevel, dt) in {
            "psi_f64_a": (2, 64, 48, 2, np.float64),
            "psi_f64_b": (1, 64, 122, 1, np.float64),
            "psi_f64_c": (1, 96, 112, 3, np.float64),
            "psi_f32_a": (2, 64, 48, 2, np.float32)}.items():
        rng = np.random.default_rng(420)
        x = rng.standard_normal((nband, nx, ny)).astype(dt)
        psi = Psi(nband, nx, ny, bases, nlevel, 1)
        fill = rng.standard_normal((nband, len(bases), psi.Nymax, psi.Nxmax)).astype(dt)
        alpha = fill.copy()
        psi.dot(x, alpha)
        coef_in = rng.standard_normal(alpha.shape).astype(dt)
        xrec = np.full(x.shape, np.nan, dt)
        psi.hdot(coef_in, xrec)
        # fill / coef_in are re-drawn by the tests from default_rng(420) in the
        # same order (x, fill, coef_in) instead of being stored
        save(tag, x=x, alpha=alpha, xrec=xrec,
             nlevel=nlevel, Nxmax=psi.Nxmax, Nymax=psi.Nymax,
             bases=np.array(bases))


def gen_prox():
    rng = np.random.default_rng(421)
    for tag, (nband, nbasis, nymax, nxmax, dt) in {
            "prox_f64": (3, 2, 37, 45, np.float64),
            "prox_f32": (3, 2, 37, 45, np.float32),
            "prox_f64_1band": (1, 3, 20, 33, np.float64)}.items():
        out = {}
        v = rng.standard_normal((nband, nbasis, nymax, nxmax)).astype(dt)
        v[:, 0, 3, 4] = 0.0          # exact-zero band sum branch
        vp = rng.standard_normal(v.shape).astype(dt)
        w = rng.random((nbasis, nymax, nxmax)).astype(dt)
        out.update(v=v, vp=vp, w=w)
        for i, (lam, sigma) in enumerate([(1.0, 1.0), (0.1, 75.0), (1e-3, 1e-3),
                                          (0.7, 2.5)]):
            res = rng.standard_normal(v.shape).astype(dt)
            prox_21m_numba(v, res, lam, sigma=sigma, weight=w)
            out[f"prox21m_{i}"] = res
            v2 = v.copy()
            dual_update_numba(vp, v2, lam, sigma=sigma, weight=w)
            out[f"dual21m_{i}"] = v2
            # l2 variants operate on (nband, nbasis, ntot)
            v3 = v.reshape(nband, nbasis, -1)
            res3 = np.zeros_like(v3)
            p21.prox_21_numba(v3, res3, lam, sigma=sigma, weight=w.reshape(nbasis, -1))
            out[f"prox21_{i}"] = res3.reshape(v.shape)
            v4 = v.reshape(nband, nbasis, -1).copy()
            p21.dual_update_numba(vp.reshape(nband, nbasis, -1), v4, lam,
                                  sigma=sigma, weight=w.reshape(nbasis, -1))
            out[f"dual21_{i}"] = v4.reshape(v.shape)
            out[f"params_{i}"] = np.array([lam, sigma])
        out["prox21m_numpy"] = prox_21m(v, 0.3, weight=w)
        out["prox21_numpy"] = p21.prox_21(v, 0.3, weight=w)
        save(tag, **out)


def gen_hess():
    for tag, (nband, nx, ny, nxp, nyp, dt) in {
            "hess_f64": (2, 32, 24, 64, 64, np.float64),
            "hess_f32": (2, 32, 24, 64, 64, np.float32),
            "hess_f64_rect": (1, 48, 64, 128, 256, np.float64)}.items():
        rng = np.random.default_rng(422)
        cdt = np.complex64 if dt == np.float32 else np.complex128
        psf, psfhat = synth_psf(rng, nband, nxp, nyp, dt)
        x = rng.standard_normal((nband, nx, ny)).astype(dt)
        beam = (0.5 + rng.random((nband, nx, ny))).astype(dt)
        xpad = np.empty((nband, nxp, nyp), dt)
        xhat = np.empty((nband, nxp, nyp // 2 + 1), cdt)
        xout = np.empty((nband, nx, ny), dt)
        conv = psf_convolve_cube(xpad, xhat, xout, psfhat, nyp, x).copy()
        conv0 = psf_convolve_slice(xpad[0], xhat[0], xout[0], psfhat[0], nyp, x[0]).copy()
        h1 = hessian_psf_cube(xpad, xhat, xout, None, psfhat, nyp, x,
                              sigmainv=1e-3, wsum=None).copy()
        h2 = hessian_psf_cube(xpad, xhat, xout, beam, psfhat, nyp, x,
                              sigmainv=0.25, wsum=3.5).copy()
        h3 = _hessian_psf_slice(xpad[0], xhat[0], xout[0], psfhat[0], beam[0], nyp,
                                x[0], sigmainv=0.5, wsum=2.0).copy()
        save(tag, psf=psf, psfhat=psfhat, x=x, beam=beam, conv=conv, conv0=conv0,
             h1=h1, h2=h2, h3=h3)


def gen_misc():
    rng = np.random.default_rng(423)
    out = {}
    for dt, t in ((np.float64, "f64"), (np.float32, "f32")):
        x = rng.standard_normal((3, 17, 29)).astype(dt)
        xp = (x + 0.01 * rng.standard_normal(x.shape)).astype(dt)
        out[f"x_{t}"], out[f"xp_{t}"] = x, xp
        out[f"nd3_{t}"] = np.float64(norm_diff(x, xp))
        out[f"nd2_{t}"] = np.float64(norm_diff(x[0], xp[0]))
    save("misc", **out)


def synth_problem(seed, nband, nx, ny, nxp, nyp, dt):
    """Small version of the SURVEY.md 8(d) synthetic problem."""
    rng = np.random.default_rng(seed)
    psf, psfhat = synth_psf(rng, nband, nxp, nyp, dt)
    ii, jj = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    model = np.zeros((nband, nx, ny))
    nu = 1.0 + 0.1 * np.arange(nband)
    for _ in range(6):
        cx, cy = rng.uniform(0.15 * nx, 0.85 * nx), rng.uniform(0.15 * ny, 0.85 * ny)
        s = rng.uniform(0.8, 4.0)
        flux = 1 + np.exp(rng.standard_normal())
        alpha = -0.7 + 0.1 * rng.standard_normal()
        g = flux * np.exp(-((ii - cx)**2 + (jj - cy)**2) / (2 * s * s))
        model += g[None] * (nu[:, None, None] ** alpha)
    for _ in range(10):
        px, py = rng.integers(int(0.15 * nx), int(0.85 * nx)), rng.integers(int(0.15 * ny), int(0.85 * ny))
        model[:, px, py] += (1 + np.exp(rng.standard_normal())) * nu ** (-0.7)
    model = model.astype(dt)
    cdt = np.complex64 if dt == np.float32 else np.complex128
    xpad = np.empty((nband, nxp, nyp), dt)
    xhat = np.empty((nband, nxp, nyp // 2 + 1), cdt)
    xout = np.empty((nband, nx, ny), dt)
    conv = partial(psf_convolve_cube, xpad, xhat, xout, psfhat, nyp)
    dirty = conv(model).copy()
    dirty = (dirty + 1e-3 * dirty.max() * rng.standard_normal(dirty.shape)).astype(dt)
    return dict(psf=psf, psfhat=psfhat, model=model, dirty=dirty, conv=conv,
                bufs=(xpad, xhat, xout), rng=rng)


def gen_solvers():
    bases = ['self', 'db1', 'db2', 'db3', 'db4']
    nbasis = len(bases)
    for tag, (nband, nx, ny, nxp, nyp, nlevel, dt) in {
            "solve_f64": (2, 64, 64, 128, 128, 2, np.float64),
            "solve_f32": (2, 64, 64, 128, 128, 2, np.float32)}.items():
        P = synth_problem(424, nband, nx, ny, nxp, nyp, dt)
        conv, psfhat, dirty, rng = P["conv"], P["psfhat"], P["dirty"], P["rng"]
        xpad, xhat, xout = P["bufs"]
        out = dict(psfhat=psfhat, dirty=dirty, model_true=P["model"], nlevel=nlevel,
                   bases=np.array(bases))
        # power method (b0 supplied: the reference draws from global np.random)
        b0 = rng.standard_normal((nband, nx, ny)).astype(dt)
        beta, bvec = power_method(conv, (nband, nx, ny), b0=b0.copy(), tol=1e-3,
                                  maxit=40, verbosity=0)
        out.update(pm_b0=b0, pm_beta=np.float64(beta), pm_b=bvec.copy())  # bvec aliases xout
        hessnorm = 1.05 * float(beta)
        psi = Psi(nband, nx, ny, bases, nlevel, 1)
        lam = 0.5 * float(np.std(np.sum(dirty, axis=0)))
        data = dirty.copy()

        def grad21(x):
            return conv(x) - data
        # --- primal_dual_optimised, 10 fixed iterations
        for pos in (1, 2):
            x = np.zeros((nband, nx, ny), dt)
            v = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dt)
            w = np.ones((nbasis, psi.Nymax, psi.Nxmax), dt)
            x, v = primal_dual_optimised(x, v, lam, psi.hdot, psi.dot, hessnorm,
                                         prox_21m_numba, w, None, grad21,
                                         nu=nbasis, tol=0.0, maxit=10,
                                         positivity=pos, verbosity=0)
            out[f"pdo_x_pos{pos}"], out[f"pdo_v_pos{pos}"] = x.copy(), v.copy()
        # --- with l1 reweighting firing (tol large)
        rms_comps = (0.1 + rng.random(nbasis))[:, None, None].astype(dt)
        outvar = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dt)
        rew = partial(l1reweight_func, psi.dot, outvar, 1.0, rms_comps)
        x = np.zeros((nband, nx, ny), dt)
        v = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dt)
        w = np.ones((nbasis, psi.Nymax, psi.Nxmax), dt)
        x, v = primal_dual_optimised(x, v, lam, psi.hdot, psi.dot, hessnorm,
                                     prox_21m_numba, w, rew, grad21, nu=nbasis,
                                     tol=0.2, maxit=12, positivity=1,
                                     verbosity=0, maxreweight=3)
        out.update(pdo_rw_x=x.copy(), pdo_rw_v=v.copy(), rms_comps=rms_comps,
                   rw_weight=l1reweight_func(psi.dot, outvar, 1.0, rms_comps, x))
        # --- functional primal_dual
        def psiH_f(xx):
            a = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dt)
            psi.dot(xx, a)
            return a

        def psi_f(a):
            xx = np.zeros((nband, nx, ny), dt)
            psi.hdot(a, xx)
            return xx
        x = np.zeros((nband, nx, ny), dt)
        v = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dt)
        x, v = primal_dual(x, v, lam, psi_f, psiH_f, hessnorm,
                           partial(prox_21m, weight=1.0), lambda z: grad21(z).copy(),
                           nu=nbasis, tol=0.0, maxit=10, minit=10, positivity=1,
                           verbosity=0)
        out.update(pd_x=x.copy(), pd_v=v.copy(), lam=np.float64(lam),
                   hessnorm=np.float64(hessnorm))
        # --- pcg on the whole cube, 10 fixed iterations
        A = partial(hessian_psf_cube, xpad, xhat, xout, None, psfhat, nyp,
                    sigmainv=1e-3, wsum=None)
        xs, rs = pcg(A, dirty, np.zeros_like(dirty), M=None, tol=0.0, maxit=10,
                     minit=10, verbosity=0, backtrack=True, return_resid=True)
        out.update(pcg_x=xs.copy(), pcg_r=rs.copy())
        beam = (0.7 + 0.3 * rng.random((nband, nx, ny))).astype(dt)
        xb = _pcg_psf_impl(psfhat, dirty, np.zeros_like(dirty), beam, nyp, 1,
                           0.05, tol=1e-4, maxit=25, minit=3, verbosity=0)
        out.update(beam=beam, pcgpsf_x=xb.copy())
        save(tag, **out)


def gen_glue():
    """Worker glue between two minor cycles.  The spotless worker cannot be
    imported here (it needs an MS, dask, zarr, ...), so its lines are executed on the
    reference's own operators exactly as written there: pfb/workers/spotless.py:153-157
    (residual_mfs), :214-223 / :356-363 (rms_comps), :229-230 / :343-345 (rms, rmax, eps),
    :259 (data), pfb/operators/gridder.py:712-714 (psfhat = r2c(iFs(psf)))."""
    from ducc0.fft import r2c
    iFs = np.fft.ifftshift
    bases = ['self', 'db1', 'db2', 'db3', 'db4']
    nbasis = len(bases)
    for tag, (nband, nx, ny, nxp, nyp, nlevel, dt) in {
            "glue_f64": (3, 64, 48, 128, 128, 2, np.float64),
            "glue_f32": (3, 64, 48, 128, 128, 2, np.float32)}.items():
        P = synth_problem(425, nband, nx, ny, nxp, nyp, dt)
        conv, psf, dirty, rng = P["conv"], P["psf"], P["dirty"], P["rng"]
        out = dict(psf=psf, dirty=dirty, nlevel=nlevel, bases=np.array(bases))
        # gridder.py:712-714, per band slice as the reference does (2-D psf, axes (0, 1))
        psfhat = np.stack([r2c(iFs(psf[b], axes=(0, 1)), axes=(0, 1), nthreads=1, forward=True, inorm=0)
                           for b in range(nband)])
        out["psfhat"] = psfhat
        model = (P["model"] * (rng.random(P["model"].shape) < 0.5)).astype(dt)
        residual = (dirty - conv(model)).astype(dt)
        out.update(model=model, residual=residual)
        # spotless.py:259
        out["data"] = residual + conv(model)
        # spotless.py:153-157, 229-230
        residual_mfs = np.sum(residual, axis=0)
        out.update(residual_mfs=residual_mfs, rms=np.float64(np.std(residual_mfs)),
                   rmax=np.float64(np.abs(residual_mfs).max()))
        # spotless.py:214-223
        wsums = np.array([1.5, 0.0, 2.5])
        wsum = np.sum(wsums)
        pix_per_beam = 11.3
        psi = Psi(nband, nx, ny, bases, nlevel, 1)
        psiHoutvar = np.zeros((nband, nbasis, psi.Nymax, psi.Nxmax), dtype=dirty.dtype)
        fsel = wsums > 0
        tmp2 = residual.copy()
        tmp2[fsel] *= wsum / wsums[fsel, None, None]
        psi.dot(tmp2 / pix_per_beam, psiHoutvar)
        rms_comps = np.std(np.sum(psiHoutvar, axis=0), axis=(-1, -2))[:, None, None]
        out.update(wsums=wsums, pix_per_beam=np.float64(pix_per_beam), rms_comps=rms_comps)
        # spotless.py:364-371
        out["l1weight"] = l1reweight_func(psi.dot, psiHoutvar, 1.0, rms_comps, model, alpha=4)
        # spotless.py:345
        modelp = (model * 0.9).astype(dt)
        out.update(modelp=modelp, eps=np.float64(np.linalg.norm(model - modelp) / np.linalg.norm(model)))
        save(tag, **out)



def gen_clean():
    """CLEAN minor loops: pfb/deconv/clark.py clark / subminor and
    pfb/deconv/hogbom.py hogbom, imported unchanged."""
    _misc.give_edges = None                     # imported by hogbom.py, unused by hogbom()
    from pfb.deconv.clark import clark, subminor
    from pfb.deconv.hogbom import hogbom
    for tag, (nband, nx, ny, nxp, nyp, dt) in {
            "clean_f64": (3, 48, 40, 96, 80, np.float64),
            "clean_f32": (3, 48, 40, 96, 80, np.float32)}.items():
        P = synth_problem(426, nband, nx, ny, nxp, nyp, dt)
        psf, psfhat, dirty = P["psf"], P["psfhat"], P["dirty"]
        wsums = np.amax(psf, axis=(1, 2)).astype(np.float64)
        wsums = wsums / wsums.sum()
        out = dict(psf=psf, psfhat=psfhat, dirty=dirty, wsums=wsums)
        model, status = clark(dirty, psf, psfhat, wsums, threshold=0, gamma=0.1, pf=0.05, maxit=6,
                              subpf=0.5, submaxit=40, verbosity=0)
        out.update(clark_model=model.copy(), clark_status=status)
        # one sub-minor call on its own
        IRs = np.sum(dirty, axis=0) ** 2
        Ip, Iq = np.where(IRs > (0.4 * np.sqrt(IRs.max())) ** 2)
        m0 = np.zeros_like(dirty)
        m1 = subminor(dirty[:, Ip, Iq], psf, Ip, Iq, m0, wsums, gamma=0.1, th=0.4 * np.sqrt(IRs.max()), maxit=25)
        out.update(sub_Ip=Ip, sub_Iq=Iq, sub_model=m1.copy())
        x, st = hogbom(dirty, psf, threshold=0, gamma=0.1, pf=0.1, maxit=60, verbosity=0)
        out.update(hogbom_x=x.copy(), hogbom_status=st)
        save(tag, **out)



class _Var:
    def __init__(self, a):
        self.data = a
        self.dtype = a.dtype
        self.shape = a.shape


class FakeDS(dict):
    """Minimal stand-in for the xarray datasets dds2cubes walks: `'NAME' in ds`, ds.NAME.data,
    ds.bandid."""

    def __init__(self, bandid, **arrs):
        super().__init__({k: _Var(v) for k, v in arrs.items()})
        self.bandid = bandid

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)


def gen_stage():
    """dds2cubes (pfb/utils/misc.py:664-739): the function's source text is cut out of misc.py
    (the module cannot be imported whole) and executed unchanged with a numpy-backed stand-in
    for the three dask calls it makes (da.zeros, da.stack, dask.compute)."""
    import types
    path = os.path.join(shims.REFERENCE, "pfb", "utils", "misc.py")
    src = open(path).read()
    tree = ast.parse(src)
    lines = src.splitlines()
    chunk = [("\n".join(lines[n.lineno - 1:n.end_lineno])) for n in tree.body
             if isinstance(n, ast.FunctionDef) and n.name == "dds2cubes"][0]
    da = types.SimpleNamespace(zeros=lambda shape, chunks=None, dtype=None: np.zeros(shape, dtype=dtype),
                               stack=np.stack)
    dask = types.SimpleNamespace(compute=lambda *a: a)
    ns = {"np": np, "da": da, "dask": dask}
    exec(chunk, ns)
    dds2cubes = ns["dds2cubes"]
    for tag, dt in (("stage_f64", np.float64), ("stage_f32", np.float32)):
        rng = np.random.default_rng(427)
        nband, nx, ny, nxp, nyp, nbasis, nym, nxm = 3, 20, 16, 40, 32, 2, 24, 27
        cdt = np.complex64 if dt == np.float32 else np.complex128
        dds, flat = [], {}
        for i, b in enumerate([0, 1, 1, 2, 0]):          # several time chunks per band
            arrs = dict(DIRTY=rng.standard_normal((nx, ny)).astype(dt),
                        BEAM=(0.5 + rng.random((nx, ny))).astype(dt),
                        RESIDUAL=rng.standard_normal((nx, ny)).astype(dt),
                        PSF=rng.standard_normal((nxp, nyp)).astype(dt),
                        PSFHAT=(rng.standard_normal((nxp, nyp // 2 + 1)) +
                                1j * rng.standard_normal((nxp, nyp // 2 + 1))).astype(cdt),
                        MODEL=rng.standard_normal((nx, ny)).astype(dt),
                        DUAL=rng.standard_normal((nbasis, nym, nxm)).astype(dt),
                        WSUM=np.array([0.5 + rng.random()], dtype=dt))
            dds.append(FakeDS(b, **arrs))
            for k, v in arrs.items():
                flat[f"ds{i}_{k}"] = v
            flat[f"ds{i}_bandid"] = np.int64(b)
        out = {}
        for ap in (False, True):
            r = dds2cubes(dds, nband, apparent=ap)
            for name, a in zip(("dirty", "model", "residual", "psf", "psfhat", "mean_beam", "wsums", "dual"), r):
                out[f"{name}_ap{int(ap)}"] = np.asarray(a)
        save(tag, nds=np.int64(len(dds)), nband=np.int64(nband), **flat, **out)


def gen_restore():
    """convolve2gaussres, Gaussian2D, get_padding_info (pfb/utils/misc.py:109-138, 170-238): the
    functions' source text is cut out of misc.py (the module cannot be imported whole) and executed
    unchanged; r2c / c2r / good_size are the ducc0 stand-ins of oracle/shims.py (scipy.fft)."""
    from ducc0.fft import r2c, c2r, good_size
    path = os.path.join(shims.REFERENCE, "pfb", "utils", "misc.py")
    src = open(path).read()
    tree = ast.parse(src)
    lines = src.splitlines()
    ns = {"np": np, "r2c": r2c, "c2r": c2r, "iFs": np.fft.ifftshift, "Fs": np.fft.fftshift,
          "good_size": lambda n, real=True: good_size(int(n), real)}
    for name in ("Gaussian2D", "get_padding_info", "convolve2gaussres"):
        chunk = [("\n".join(lines[n.lineno - 1:n.end_lineno])) for n in tree.body
                 if isinstance(n, ast.FunctionDef) and n.name == name][0]
        exec(chunk, ns)
    conv, gauss, padinfo = ns["convolve2gaussres"], ns["Gaussian2D"], ns["get_padding_info"]
    rng = np.random.default_rng(429)
    nband, nx, ny = 3, 64, 48
    cell = 1.5
    x = (-(nx // 2) + np.arange(nx)) * cell
    y = (-(ny // 2) + np.arange(ny)) * cell
    xx, yy = np.meshgrid(x, y, indexing="ij")
    image = rng.standard_normal((nband, nx, ny))
    image[:, nx // 3, ny // 4] += 25.0
    gaussparf = (9.0, 6.0, 30.0)
    gausspari = tuple((4.0 + 0.5 * b, 3.0 + 0.25 * b, 10.0 * b) for b in range(nband))
    padding, ux, uy = padinfo(nx, ny, 0.5)
    out = dict(image=image, xx=xx, yy=yy, gaussparf=np.array(gaussparf), gausspari=np.array(gausspari),
               padding=np.array(padding), kern=gauss(xx, yy, gaussparf, normalise=False),
               kern_norm=gauss(xx, yy, gaussparf, normalise=True))
    out["conv"] = conv(image.copy(), xx, yy, gaussparf, 1)
    out["conv_norm"] = conv(image.copy(), xx, yy, gaussparf, 1, norm_kernel=True)
    out["conv_from"] = conv(image.copy(), xx, yy, gaussparf, 1, gausspari=gausspari)
    out["conv_p25"] = conv(image.copy(), xx, yy, gaussparf, 1, pfrac=0.25)
    save("restore_f64", **out)


if __name__ == "__main__" and "--restore-only" in sys.argv:
    gen_restore()
elif __name__ == "__main__" and "--stage-only" in sys.argv:
    gen_stage()
elif __name__ == "__main__" and "--clean-only" in sys.argv:
    gen_clean()
elif __name__ == "__main__" and "--glue-only" in sys.argv:
    gen_glue()
elif __name__ == "__main__" and "--psi-only" in sys.argv:
    gen_psi()
elif __name__ == "__main__" and "--solvers-only" in sys.argv:
    gen_solvers()
elif __name__ == "__main__":
    gen_psi()
    gen_prox()
    gen_hess()
    gen_misc()
    gen_solvers()
    gen_glue()
    gen_clean()
    gen_stage()
    gen_restore()
