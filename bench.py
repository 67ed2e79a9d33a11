#!/usr/bin/env python
"""bench.py -- PFB minor iterations / second (BASELINE.json metric).

A "step" is ONE primal-dual minor iteration (body of primal_dual_optimised,
pfb/opt/primal_dual.py:134-171 of the reference: Psi.dot -> dual update ->
2v-vp -> Psi.hdot -> PSF Hessian -> primal update -> eps) over the WHOLE cube.
Default workload = BASELINE.json configs[2]: 8 bands x 4096^2, PSF 8192^2,
bases self+db1..db4, 3 levels, fp32 -- it fits one B200 (about 14 GB).  With
--gpus N (torchrun) the 8 bands are sharded over the ranks (strong scaling) and
the band-sum slab of the dual update plus the eps scalars are all-reduced (NCCL).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Prints ONE JSON line (rank 0).  `value` = K / device time of K iterations with
every input resident in HBM; `e2e` = the same K iterations through the
reference-facing call primal_dual_optimised(x, v, ...) with HOST (pinned numpy)
buffers, host<->device copies inside the timed region.  `--impl reference`
times the CPU path (the oracle port of the reference, all host threads) on a
bounded sample of the same workload.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BASES = ["self", "db1", "db2", "db3", "db4"]
NLEVEL = 3
METRIC = "pfb_minor_iters_per_sec"
UNIT = "iter/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nband", type=int, default=8)
    ap.add_argument("--nx", type=int, default=4096)
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-sample-bands", type=int, default=1)
    ap.add_argument("--config", type=int, default=0, choices=[0, 1, 2, 3, 4, 5],
                    help="BASELINE.json configuration (sets --nband/--nx): 1 = 1x512^2, 2 = 4x2048^2, "
                         "3 = 8x4096^2 (default workload), 4 = 16x8192^2, 5 = 64x10240^2")
    ap.add_argument("--no-secondary", action="store_true",
                    help="skip the secondary lines (config 2 f32/f64, config 3 f64) of the N=1 run")
    ap.add_argument("--no-parity", action="store_true",
                    help="N>1: skip the sharded-vs-single-GPU parity check (parity_vs_n1)")
    ap.add_argument("--parity-iters", type=int, default=10)
    ap.add_argument("--force-subset-parity", action="store_true",
                    help="testing aid: take the one-band-per-rank parity path used by the large configurations")
    ap.add_argument("--ncu-step", action="store_true",
                    help="profiling aid: run ONE iteration between cudaProfilerStart/Stop and exit "
                         "(use with ncu --profile-from-start off); prints no bench line")
    a = ap.parse_args()
    if a.config:
        a.nband, a.nx = {1: (1, 512), 2: (4, 2048), 3: (8, 4096), 4: (16, 8192), 5: (64, 10240)}[a.config]
    return a


def workload_name(a):
    return (f"{a.nband} bands {a.nx}x{a.nx}, PSF {2 * a.nx}x{2 * a.nx} (psf-oversize 2), "
            f"bases {'+'.join(BASES)}, {NLEVEL} levels, primal-dual minor cycle")


# ------------------------------------------------------------ synthetic data
def make_band_inputs(torch, band, nband, nx, rdt, dev, seed=420):
    """MeerKAT-like PSF (core-dominated, thinned uv density) and a
    sky of Gaussians + point sources for one band, built on `dev` with torch
    (input generation only -- not the product path).
    Returns psfhat (nxp, nyo2) complex, model (nx, nx)."""
    nxp = 2 * nx
    cdt = torch.complex64 if rdt == torch.float32 else torch.complex128
    g = torch.Generator(device="cpu").manual_seed(seed)        # band independent part
    nu = 1.0 + 0.1 * band
    u = torch.fft.fftfreq(nxp, device=dev, dtype=torch.float64)[:, None] * nxp
    vv = torch.fft.rfftfreq(nxp, device=dev, dtype=torch.float64)[None, :] * nxp
    r = torch.sqrt(u * u + vv * vv) / nu
    W = (1.0 + (r / (0.05 * nxp)) ** 2) ** -1.5
    gd = torch.Generator(device=dev).manual_seed(seed + 17)
    W = W * (torch.rand(W.shape, generator=gd, device=dev, dtype=torch.float64) < 0.3)
    psf = torch.fft.irfft2(W, s=(nxp, nxp))                   # peak at (0,0): already ifftshift-ed
    psf = psf / (psf[0, 0] * nband)                           # sum_b psf_b.max() = 1
    psfhat = torch.fft.rfft2(psf).to(cdt)                     # = r2c(ifftshift(psf_centred))
    del W, psf, r
    # sky
    model = torch.zeros((nx, nx), dtype=torch.float64, device=dev)
    ii = torch.arange(nx, device=dev, dtype=torch.float64)
    ngauss, npoint = 25, 50
    cen = 0.15 * nx + 0.7 * nx * torch.rand((ngauss + npoint, 2), generator=g, dtype=torch.float64)
    flux = 1.0 + torch.exp(torch.randn(ngauss + npoint, generator=g, dtype=torch.float64))
    alpha = -0.7 + 0.1 * torch.randn(ngauss + npoint, generator=g, dtype=torch.float64)
    fwhm = 3.0 + (0.02 * nx - 3.0) * torch.rand(ngauss, generator=g, dtype=torch.float64)
    for k in range(ngauss):
        s = float(fwhm[k]) / 2.355
        gx = torch.exp(-0.5 * ((ii - float(cen[k, 0])) / s) ** 2)
        gy = torch.exp(-0.5 * ((ii - float(cen[k, 1])) / (1.5 * s)) ** 2)
        model += float(flux[k]) * nu ** float(alpha[k]) * gx[:, None] * gy[None, :]
    for k in range(ngauss, ngauss + npoint):
        model[int(cen[k, 0]), int(cen[k, 1])] += float(flux[k]) * nu ** float(alpha[k])
    return psfhat, model.to(rdt)


def clock_sampler(idx, stop, rows):
    """nvidia-smi clocks/throttle reasons during the timed region."""
    q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    try:
        p = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits",
                              "-lms", "100", "-i", str(idx)], stdout=subprocess.PIPE, text=True)
    except OSError:
        return
    def pump():
        for line in p.stdout:
            rows.append(line.strip())
    t = threading.Thread(target=pump, daemon=True)
    t.start()
    stop.wait()
    p.terminate()
    t.join(timeout=2)


def summarise_clocks(rows):
    sm, mx, reasons = [], 0.0, set()
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    for r in rows:
        f = [c.strip() for c in r.split(",")]
        if len(f) < 7:
            continue
        try:
            sm.append(float(f[0]))
            mx = max(mx, float(f[1]))
        except ValueError:
            continue
        for n, val in zip(names, f[3:7]):
            if val.lower().startswith("active"):
                reasons.add(n)
    sm.sort()
    med = sm[len(sm) // 2] if sm else None
    return {"sm_mhz": med, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


# -------------------------------------------------------- algorithmic bytes
def byte_model(nb_local, nx, s, Nmax, nbasis, multi, fused=False):
    """Algorithmic (compulsory) bytes, SURVEY.md 8(d) / DESIGN.md, for the bands
    on ONE GPU.  s = bytes per real."""
    nxp = 2 * nx
    nyo2 = nxp // 2 + 1
    P1 = nx * nx * s
    C1 = nbasis * Nmax * Nmax * s
    psfhat = 2 * s * nxp * nyo2
    S = 2 * s * nx * nyo2                      # pruned half spectrum (pipeline intermediate)
    # default loop: synthesis by linearity (opt/primal_dual.py `lin`): the dual kernels do not
    # write 2v - vp, the primal kernel reads w_k and w_{k-1} instead of xout
    lin = 0 if os.environ.get("PFB_PD_NOLIN") else 1
    per_kernel = {                              # inputs + outputs of each kernel, counted once
        "k_dwt_level": None, "k_idwt_level": None,   # per-op figure used instead (3 launches each)
        "k_dwt_stream": None, "k_idwt_stream": None,
        "k_dual": nb_local * (4 - lin) * C1 + C1,     # read vp,v ; write v (and u = 2v-vp if not lin) ; weight once
        "k_dual_partial": nb_local * 3 * C1 + C1,
        "k_dual_apply": nb_local * (4 - lin) * C1 + 2 * C1,   # read vp,v,bsum,w ; write v (,u)
        # band-sharded recursion: read vp_b, u_b ; write v_b ; read Vbar, ubar, w ; write Vbar
        "k_dual_rec": nb_local * 3 * C1 + 4 * C1,
        "k_hess_rows_fwd": nb_local * (P1 + S),
        "k_hess_cols": nb_local * (2 * S + psfhat),
        "k_hess_rows_inv": nb_local * (S + P1),
        "k_pd_primal": nb_local * (5 + lin) * P1,     # read xp,xout(,wprev),conv,data ; write x
    }
    for k in ("rows_fwd", "cols", "rows_inv"):  # fast-path / mixed-radix kernels move the same bytes
        per_kernel["k_hessf_" + k] = per_kernel["k_hess_" + k]
        per_kernel["k_hessm_" + k] = per_kernel["k_hess_" + k]
        per_kernel["k_hessp_" + k] = per_kernel["k_hess_" + k]
    per_op = {                                  # SURVEY 8(d) table, x nb_local
        "psi_dot": nb_local * (P1 + C1),
        "psi_hdot": nb_local * (C1 + P1),       # hdot(u), u = 2v - vp written by the dual update
        "psf_convolve": nb_local * (2 * P1 + psfhat),
        "dual_update": nb_local * (4 - lin) * C1 + C1,
        "primal_update+eps": nb_local * (5 + lin) * P1,
    }
    if fused:
        # pfb_hess_apply_primal: pass C reads S, xp, w_k, w_{k-1}, data and writes x (no conv array,
        # no separate primal kernel)
        for k in ("k_hess_rows_inv", "k_hessf_rows_inv", "k_hessm_rows_inv", "k_hessp_rows_inv"):
            per_kernel[k] = nb_local * (S + 5 * P1)
        per_op["psf_convolve+primal_update"] = nb_local * (6 * P1 + psfhat)
    fused_iter = nb_local * (3 * P1 + psfhat + 2 * C1) + C1     # SURVEY 8(d) headline lower bound
    return per_kernel, per_op, fused_iter


def short_kernel_name(mangled):
    for k in ("k_dwt_level", "k_dwt_stream", "k_self_dot", "k_idwt_level", "k_idwt_stream", "k_slab_to_nat", "k_nat_to_slab", "k_dual_partial", "k_dual_apply", "k_dual_rec", "k_dual_factor", "k_dual",
              "k_hessf_rows_fwd", "k_hessf_cols", "k_hessf_rows_inv",
              "k_hess_rows_fwd", "k_hess_cols", "k_hess_rows_inv",
              "k_hessm_rows_fwd", "k_hessm_cols", "k_hessm_rows_inv",
              "k_hessp_rows_fwd", "k_hessp_cols", "k_hessp_rows_inv", "k_pd_primal", "k_hess_red_final", "k_red_final",
              "k_norm_diff", "k_axpby", "k_dot", "k_prox21", "k_psfhat_prepare"):
        if k in mangled:
            return k
    return mangled[:40]


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ----------------------------------------------------------------- our arm
class Problem:
    """Synthetic minor-cycle problem for `bands` (global band indices) of an
    nband x nx x nx cube on `dev`: psfhat, dirty, Psi plan, hessnorm, lam, weights.
    Set-up only -- nothing here is inside a timed region."""

    def __init__(self, torch, nband, nx, dtype, bands, dev, rank=0, world=1, hessnorm=None, lam=None):
        from functools import partial
        import numpy as np
        import torch.distributed as dist
        from pfb_b200.operators.psf import psf_convolve_cube
        from pfb_b200.operators.psi import Psi
        from pfb_b200.opt.primal_dual import Grad21
        from pfb_b200.opt.power_method import power_method
        self.torch = torch
        self.rdt = rdt = torch.float32 if dtype == "f32" else torch.float64
        self.s = 4 if dtype == "f32" else 8
        self.nband, self.nx, self.nxp = nband, nx, 2 * nx
        self.nbl = nbl = len(bands)
        nxp = 2 * nx
        cdt = torch.complex64 if self.s == 4 else torch.complex128
        self.psfhat = torch.empty((nbl, nxp, nxp // 2 + 1), dtype=cdt, device=dev)
        model = torch.empty((nbl, nx, nx), dtype=rdt, device=dev)
        for i, b in enumerate(bands):
            self.psfhat[i], model[i] = make_band_inputs(torch, b, nband, nx, rdt, dev)
        torch.cuda.empty_cache()
        self.xout = torch.empty_like(model)
        self.conv = partial(psf_convolve_cube, None, None, self.xout, self.psfhat, nxp)
        self.dirty = self.conv(model).clone()
        dmax = self.dirty.max()
        if world > 1:
            dist.all_reduce(dmax, op=dist.ReduceOp.MAX)
        for i, b in enumerate(bands):     # noise seeded per GLOBAL band: same cube whatever the sharding
            gn = torch.Generator(device=dev).manual_seed(1000 + b)
            self.dirty[i] += 1e-3 * dmax * torch.randn(self.dirty[i].shape, generator=gn, device=dev, dtype=rdt)
        del model
        self.psi = Psi(nbl, nx, nx, BASES, NLEVEL, 1, dtype=np.float32 if self.s == 4 else np.float64)
        self.nbasis = len(BASES)
        self.shape_v = (nbl, self.nbasis, self.psi.Nymax, self.psi.Nxmax)
        if hessnorm is None:
            # hessnorm = 1.05 * power_method(psf_convolve)  (pfb/workers/spotless.py:187-193)
            b0 = torch.empty(self.dirty.shape, device=dev, dtype=rdt)
            for i, b in enumerate(bands):
                gb = torch.Generator(device=dev).manual_seed(7 + b)
                b0[i] = torch.randn(b0[i].shape, generator=gb, device=dev, dtype=rdt)
            beta, _ = power_method(self.conv, self.dirty.shape, b0=b0, tol=1e-3, maxit=100, verbosity=0)
            del b0
            hessnorm = 1.05 * float(beta)
        self.hessnorm = hessnorm
        if lam is None:
            mfs = self.dirty.sum(dim=0)
            if world > 1:
                dist.all_reduce(mfs)
            lam = 0.5 * float(mfs.std())
            del mfs
        self.lam = lam
        self.w = torch.ones(self.shape_v[1:], dtype=rdt, device=dev)
        self.grad = Grad21(self.conv, self.dirty)

    def pd(self, x, v, nit, positivity=1):
        from pfb_b200.opt.primal_dual import primal_dual_optimised
        return primal_dual_optimised(x, v, self.lam, self.psi.hdot, self.psi.dot, self.hessnorm, None,
                                     self.w, None, self.grad, nu=self.nbasis, tol=0.0, maxit=nit,
                                     positivity=positivity, verbosity=0)

    def zeros(self):
        t = self.torch
        return (t.zeros_like(self.dirty), t.zeros(self.shape_v, dtype=self.rdt, device=self.dirty.device))


def secondary_config(torch, nband, nx, dtype, dev, steps, warmup):
    """Device-resident PD it/s of another BASELINE.json configuration (N=1, outside the headline's
    timed region): reported under "secondary"."""
    from pfb_b200.operators import psf as _psf
    _psf.clear_plan_cache()
    torch.cuda.empty_cache()
    P = Problem(torch, nband, nx, dtype, list(range(nband)), dev)
    x, v = P.zeros()
    P.pd(x, v, max(3, warmup))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    P.pd(x, v, steps)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    _, _, fused_iter = byte_model(nband, nx, P.s, P.psi.Nxmax, P.nbasis, False)
    hbm, _ = peaks()
    out = {"workload": f"{nband} bands {nx}x{nx}, PSF {2 * nx}x{2 * nx}", "dtype": dtype,
           "value": round(1e3 / ms, 2), "unit": UNIT, "ms_per_step": round(ms, 4), "steps": steps,
           "step_frac_of_hbm_peak": round(fused_iter / (ms * 1e-3) / 1e9 / hbm, 4)}
    del P, x, v
    _psf.clear_plan_cache()
    torch.cuda.empty_cache()
    return out


def parity_vs_n1(torch, dist, a, P, rank, world, dev):
    """N>1, outside the timed region: `iters` iterations from x0 = v0 = 0 on the band-sharded
    path, gathered to rank 0, against the same iterations of the single-GPU path on the whole
    cube on rank 0 (same psfhat / dirty / hessnorm / lam).  Relative L2 of model and dual."""
    from pfb_b200 import parallel
    from pfb_b200.opt.primal_dual import primal_dual_optimised
    nit = a.parity_iters
    subset = None
    if P is None:
        # the whole cube cannot be gathered on one GPU: the same image size and code path with
        # ONE band per rank
        subset = world
        a = argparse.Namespace(**vars(a))
        a.nband = world
        P = Problem(torch, world, a.nx, a.dtype, [rank], dev, rank, world)
    x, v = P.zeros()
    P.pd(x, v, nit)
    path = "rec" if getattr(primal_dual_optimised, "last_rec", False) else "slab"

    def gather(t):
        out = torch.empty((world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=dev) if rank == 0 else None
        parts = list(out.split(t.shape[0])) if rank == 0 else None
        dist.gather(t.contiguous(), parts, dst=0)
        return out
    xs, vs = gather(x), gather(v)
    psf_all = gather(torch.view_as_real(P.psfhat))
    if rank == 0:
        psf_all = torch.view_as_complex(psf_all)
    dirty_all = gather(P.dirty)
    del x, v
    lam_, hn_, s_, nbasis_, shape_v_, rdt_ = P.lam, P.hessnorm, P.s, P.nbasis, P.shape_v, P.rdt
    if subset is not None:          # make room for the whole subset cube on rank 0
        from pfb_b200.operators import psf as _psf0
        del P
        _psf0.clear_plan_cache()
        import gc
        gc.collect()
        torch.cuda.empty_cache()
    res = None
    if rank == 0:
        from functools import partial
        import numpy as np
        from pfb_b200.operators.psf import psf_convolve_cube
        from pfb_b200.operators.psi import Psi
        from pfb_b200.opt.primal_dual import Grad21
        parallel.disable()
        try:
            nb = a.nband
            psi = Psi(nb, a.nx, a.nx, BASES, NLEVEL, 1, dtype=np.float32 if s_ == 4 else np.float64)
            xo = torch.empty_like(dirty_all)
            conv = partial(psf_convolve_cube, None, None, xo, psf_all, 2 * a.nx)
            x1 = torch.zeros_like(dirty_all)
            v1 = torch.zeros((nb,) + tuple(shape_v_[1:]), dtype=rdt_, device=dev)
            w1 = torch.ones(shape_v_[1:], dtype=rdt_, device=dev)
            primal_dual_optimised(x1, v1, lam_, psi.hdot, psi.dot, hn_, None, w1, None,
                                  Grad21(conv, dirty_all), nu=nbasis_, tol=0.0, maxit=nit,
                                  positivity=1, verbosity=0)
            torch.cuda.synchronize()

            def rel(p, q):          # band by band: no cube-sized fp64 temporaries
                num = den = 0.0
                for b_ in range(p.shape[0]):
                    qd = q[b_].double()
                    num += float(torch.sum((p[b_].double() - qd) ** 2))
                    den += float(torch.sum(qd ** 2))
                    del qd
                return math.sqrt(num / den)
            res = {"rel_l2_x": rel(xs, x1), "rel_l2_v": rel(vs, v1), "iters": nit, "path": path,
                   "tol": 1e-5, "x_nonzero": bool(x1.any()), "bands": a.nband,
                   "note": ("sharded primal_dual_optimised vs the single-GPU path on the whole cube (rank 0)"
                            if subset is None else
                            f"the whole cube does not fit one GPU: same image size and code path on a "
                            f"{subset}-band cube (one band per rank) vs the single-GPU path on rank 0")}
        finally:
            parallel.enable()
    dist.barrier()
    return res


def run_ours(a):
    import ctypes
    from functools import partial
    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        raise SystemExit(f"--gpus {a.gpus} but WORLD_SIZE={world}: launch with torchrun for N>1")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)

    import pfb_b200  # noqa: F401
    from pfb_b200 import _lib, parallel
    L = _lib.lib()          # raises if libpfb_b200.so is missing: no fallback
    if world > 1:
        parallel.enable()

    nb, nx = a.nband, a.nx
    nxp = 2 * nx
    if nb % world:
        raise SystemExit("nband must be divisible by the number of GPUs")
    nbl = nb // world
    bands = list(range(rank * nbl, (rank + 1) * nbl))
    # can rank 0 hold the WHOLE cube (psfhat + the plan's copy + scratch, two duals, images) next to
    # its own shard for the parity check?  If not, the check runs first, on a one-band-per-rank cube.
    sz = 4 if a.dtype == "f32" else 8
    per_band = (2 * nx) * (nx + 1) * 2 * sz * 2.5 + 5 * (nx + 24) ** 2 * sz * 5 + nx * nx * sz * 8
    big = world > 1 and (nb * per_band > 60e9 or a.force_subset_parity)
    if nb * per_band > 60e9:
        # pinning the whole dual on the host (tens of GB per rank) is not safe to assume: no e2e leg
        a.no_e2e = True
    parity = None
    if big and not a.no_parity:
        parity = parity_vs_n1(torch, dist, a, None, rank, world, dev)
        from pfb_b200.operators import psf as _psf
        _psf.clear_plan_cache()
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        torch.cuda.reset_peak_memory_stats()
    P = Problem(torch, nb, nx, a.dtype, bands, dev, rank, world)
    rdt, s, psi, nbasis, shape_v = P.rdt, P.s, P.psi, P.nbasis, P.shape_v
    psfhat, dirty, hessnorm, lam = P.psfhat, P.dirty, P.hessnorm, P.lam
    pd = P.pd

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    # ---- device-resident timing (`value`)
    x, v = P.zeros()
    pd(x, v, max(a.warmup, 3))                       # warm-up iterations (untimed)
    if a.ncu_step:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        pd(x, v, 1)
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return None
    rows, stop = [], threading.Event()
    th = threading.Thread(target=clock_sampler, args=(local, stop, rows), daemon=True)
    if rank == 0:
        th.start()
        time.sleep(0.3)
    barrier()
    L.pfb_profile_enable(1)
    n0 = L.pfb_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    pd(x, v, a.steps)
    e1.record()
    barrier()
    launches = L.pfb_launch_count() - n0
    L.pfb_profile_enable(0)
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms)
    stop.set()
    buf = ctypes.create_string_buffer(1 << 16)
    _lib.check(L.pfb_profile_report(buf, len(buf)))
    # ---- secondary metric: PCG iterations/s on the same cube, outside the timed
    # region: A = hessian_psf_cube(sigmainv = 1e-3), b = dirty, x0 = 0, fixed count, backtracking on
    pcg_line = None
    if not a.no_e2e:
        from pfb_b200.operators.hessian import hessian_psf_cube
        from pfb_b200.opt.pcg import pcg
        A = partial(hessian_psf_cube, None, None, None, None, psfhat, nxp, sigmainv=1e-3)
        x0 = torch.zeros_like(dirty)
        npcg = 10
        pcg(A, dirty, x0, tol=0.0, maxit=3, minit=3, verbosity=0)          # warm-up
        barrier()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        pcg(A, dirty, x0, tol=0.0, maxit=npcg, minit=npcg, verbosity=0)
        p1.record()
        barrier()
        pms = torch.tensor([p0.elapsed_time(p1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(pms, op=dist.ReduceOp.MAX)
        pcg_line = {"metric": "pfb_pcg_iters_per_sec", "value": round(npcg / (float(pms) * 1e-3), 2),
                    "unit": "iter/s", "iters": npcg, "ms_per_iter": round(float(pms) / npcg, 4),
                    "path": getattr(pcg, "last_path", "host-loop"),
                    "note": "pcg(hessian_psf_cube sigmainv=1e-3, b=dirty, x0=0), device-resident, incl. setup A(x0)"}
        del x0
    kern = {}
    for line in buf.value.decode().splitlines():
        nm, n, t = line.split("\t")
        k = short_kernel_name(nm)
        c = kern.setdefault(k, [0, 0.0])
        c[0] += int(n)
        c[1] += float(t)
    # ---- memory in use after the timed region
    free_b, total_b = torch.cuda.mem_get_info()
    mem = torch.tensor([float(total_b - free_b), float(torch.cuda.max_memory_allocated())], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(mem, op=dist.ReduceOp.MAX)
    memory = {"device_used_gb_max_over_ranks": round(float(mem[0]) / 1e9, 2),
              "torch_peak_allocated_gb_max_over_ranks": round(float(mem[1]) / 1e9, 2),
              "note": "after the timed region: device memory in use (cudaMemGetInfo: torch pool + the library's "
                      "plan scratch) and torch's peak allocation"}
    # ---- N>1: what the sharded path computes, against the single-GPU path (outside timing)
    if world > 1 and not a.no_parity and not big:
        del x, v
        torch.cuda.empty_cache()
        parity = parity_vs_n1(torch, dist, a, P, rank, world, dev)
        x, v = P.zeros()
    return finish_ours(a, locals())


def finish_ours(a, env):
    """e2e leg, CPU baseline leg and the JSON line (split from run_ours only to
    keep the functions short; `env` is run_ours' namespace)."""
    import numpy as np
    import torch
    import torch.distributed as dist
    g = env
    world, rank, dev, rdt, s = g["world"], g["rank"], g["dev"], g["rdt"], g["s"]
    nb, nbl, nx, psi, nbasis = g["nb"], g["nbl"], g["nx"], g["psi"], g["nbasis"]
    pd, barrier, shape_v, dirty = g["pd"], g["barrier"], g["shape_v"], g["dirty"]
    torch_ = torch
    ms_total, kern, launches = g["ms_total"], g["kern"], g["launches"]
    K = a.steps
    value = K / (ms_total * 1e-3)

    # ---- e2e: same K iterations through the public call with HOST buffers
    e2e = None
    if not a.no_e2e:
        from pfb_b200.opt.primal_dual import Grad21
        from functools import partial
        from pfb_b200.operators.psf import psf_convolve_cube
        del g["x"], g["v"]
        torch.cuda.empty_cache()
        hx = torch.zeros(dirty.shape, dtype=rdt).pin_memory()
        hv = torch.zeros(shape_v, dtype=rdt).pin_memory()
        hd = dirty.cpu().pin_memory()
        hw = torch.ones(shape_v[1:], dtype=rdt).pin_memory()
        hxo = torch.empty(dirty.shape, dtype=rdt).pin_memory()
        # psfhat stays with the operator (uploaded once when the plan is bound,
        # like the reference's one-off PSFHAT set-up): not a per-call input
        hconv = partial(psf_convolve_cube, None, None, hxo.numpy(), g["psfhat"], 2 * nx)

        def pd_host(nit):
            from pfb_b200.opt.primal_dual import primal_dual_optimised
            return primal_dual_optimised(hx.numpy(), hv.numpy(), g["lam"], psi.hdot, psi.dot,
                                         g["hessnorm"], None, hw.numpy(), None,
                                         Grad21(hconv, hd.numpy()), nu=nbasis, tol=0.0, maxit=nit,
                                         positivity=1, verbosity=0)
        pd_host(2)                                    # warm-up of the host path
        hx.zero_(); hv.zero_()
        barrier()
        t0 = time.perf_counter()
        pd_host(K)
        torch.cuda.synchronize()
        dt_ = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt_, op=dist.ReduceOp.MAX)
        h2d = (hx.numel() + hv.numel() + hd.numel() + hw.numel()) * s * world
        d2h = (hx.numel() + hv.numel()) * s * world + 16 * K * world
        e2e = {"value": K / float(dt_), "unit": UNIT,
               "h2d_bytes_per_step": h2d / K, "d2h_bytes_per_step": d2h / K,
               "note": (f"one primal_dual_optimised(x, v, ...) call of {K} iterations with pinned "
                        "numpy x, v, data, l1weight: H2D of all four and D2H of x, v inside the "
                        "timed region (bytes are per call / K); psfhat is operator state bound "
                        "before the call")}

    if rank != 0:
        return None
    hbm, peak_src = peaks()
    fused_primal = "k_pd_primal" not in kern       # pfb_hess_apply_primal ran (primal step inside pass C)
    per_kernel, per_op, fused_iter = byte_model(nbl, nx, s, psi.Nxmax, nbasis, world > 1, fused_primal)
    if "k_dual_rec" in kern or "k_dual_factor" in kern:
        per_kernel["k_dual_partial"] = None     # here only the band-sum helper (no byte model)
    # opt-in fused dual (PFB_PD_FUSEDUAL): band-shaped factor kernel reads Vbar, ubar, w, writes f, Vbar
    per_kernel["k_dual_factor"] = 5 * (psi.Nxmax * psi.Nxmax * nbasis * s) if "k_dual_factor" in kern else None
    # dominant kernel = largest total device time in the timed region
    tot_k = sum(t for _, t in kern.values())
    modelled = {k: v for k, v in kern.items() if per_kernel.get(k) is not None or k in
                ("k_dwt_level", "k_dwt_stream", "k_idwt_level", "k_idwt_stream")}
    dom = max((modelled or kern).items(), key=lambda kv: kv[1][1])
    dname, (dn, dt_ms) = dom[0], dom[1]
    if per_kernel.get(dname) is not None:
        balg = per_kernel[dname]
        what = f"{dname}: inputs+outputs of this kernel once"
    else:   # wavelet level kernels: op-level bytes over the op's launches
        op = "psi_dot" if dname in ("k_dwt_level", "k_dwt_stream") else "psi_hdot"
        balg = per_op[op] / NLEVEL
        what = f"{dname}: {op} algorithmic bytes / {NLEVEL} level launches"
    avg_s = dt_ms * 1e-3 / dn
    ach = balg / avg_s / 1e9
    # DRAM bytes per launch of that kernel from the committed `ncu --set full` capture
    # (profiles/make_traffic.py); only meaningful for the single-GPU workload it was taken on
    traffic, traffic_src = None, None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        tk = tj["kernels"].get(dname) or tj["kernels"].get(dname + "_vec")
        if tk and world == 1 and nb == 8 and nx == 4096 and a.dtype == "f32":
            traffic, traffic_src = tk["dram_bytes_per_launch"], tj["source"]
    except (OSError, ValueError, KeyError):
        pass
    roofline = {"bound": "hbm", "kernel": dname, "achieved": round(ach, 1), "peak": hbm, "unit": "GB/s",
                "frac": round(ach / hbm, 4), "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": balg, "avg_launch_ms": round(avg_s * 1e3, 4),
                "share_of_step": round(dt_ms / tot_k, 4), "basis": what}
    kernels = {k: {"launches": n, "ms_per_step": round(t / K, 4), "share": round(t / tot_k, 4)}
               for k, (n, t) in sorted(kern.items(), key=lambda kv: -kv[1][1])}
    for k, d in kernels.items():
        if per_kernel.get(k) is not None:
            d["GBps"] = round(per_kernel[k] / (d["ms_per_step"] * 1e-3) / 1e9 / (kern[k][0] / K), 1)
            d["frac"] = round(d["GBps"] / hbm, 4)
    ops = {}
    groups = {"psi_dot": ["k_dwt_level", "k_dwt_stream", "k_self_dot"], "psi_hdot": ["k_idwt_level", "k_idwt_stream", "k_slab_to_nat"],
              "psf_convolve": ["k_hess_rows_fwd", "k_hess_cols", "k_hess_rows_inv",
                               "k_hessf_rows_fwd", "k_hessf_cols", "k_hessf_rows_inv",
                               "k_hessm_rows_fwd", "k_hessm_cols", "k_hessm_rows_inv",
                               "k_hessp_rows_fwd", "k_hessp_cols", "k_hessp_rows_inv"],
              "dual_update": (["k_dual", "k_dual_rec"] if "k_dual_rec" in kern
                              else ["k_dual_factor", "k_dual_partial"] if "k_dual_factor" in kern
                              else ["k_dual", "k_dual_partial", "k_dual_apply"]),
              "primal_update+eps": ["k_pd_primal"]}
    if fused_primal:
        groups["psf_convolve+primal_update"] = groups.pop("psf_convolve")
    for op, ks in groups.items():
        t = sum(kern[k][1] for k in ks if k in kern) / K
        if t > 0:
            ops[op] = {"ms_per_step": round(t, 4), "alg_GBps": round(per_op[op] / (t * 1e-3) / 1e9, 1),
                       "frac": round(per_op[op] / (t * 1e-3) / 1e9 / hbm, 4)}
    step_frac = fused_iter / (ms_total * 1e-3 / K) / 1e9 / hbm

    secondary = None
    if world == 1 and not a.no_secondary and (nb, nx, a.dtype) == (8, 4096, "f32"):
        for k_ in ("x", "v", "P", "psfhat", "dirty", "pd", "psi"):
            g.pop(k_, None)
        del pd, dirty, psi
        import gc
        gc.collect()
        secondary = {}
        for name, (nb2, nx2, dt2, st2) in {"config2_f32": (4, 2048, "f32", 50), "config2_f64": (4, 2048, "f64", 50),
                                           "config3_f64": (8, 4096, "f64", 10)}.items():
            try:
                secondary[name] = secondary_config(torch_, nb2, nx2, dt2, dev, st2, 3)
            except Exception as e:      # a secondary line must never cost the headline
                secondary[name] = {"error": f"{type(e).__name__}: {e}"}
    cpu = None
    if not a.no_cpu_baseline and world == 1:
        cpu = cpu_reference(a, steps=2, warmup=1, nbs=a.cpu_sample_bands)

    clocks = summarise_clocks(g["rows"])
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": K,
        "warmup": max(a.warmup, 3), "ms_per_step": round(ms_total / K, 4), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
        "config": {"workload": workload_name(a), "nband": nb, "nx": nx, "ny": nx, "nxp": 2 * nx,
                   "nyp": 2 * nx, "bases": BASES, "nlevels": NLEVEL, "positivity": 1,
                   "bands_per_gpu": nbl, "parallelism": f"band-sharded x{world}",
                   "l2_policy": "inputs larger than L2 (per-iteration working set > 10 GB)",
                   "init": "x0=0, v0=0, l1weight=1, tol=0 (fixed iteration count)"},
        "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline,
        "step_roofline": {"fused_compulsory_bytes_per_step_per_gpu": fused_iter,
                          "frac_of_hbm_peak": round(step_frac, 4)},
        "kernels": kernels, "ops": ops, "pcg": g.get("pcg_line"), "cpu_baseline": cpu,
        "secondary": secondary, "parity_vs_n1": g.get("parity"), "memory": g.get("memory"),
    }
    return line


# ------------------------------------------------------- CPU reference arm
def cpu_reference(a, steps, warmup, nbs=1, full=False):
    """The reference's CPU algorithm for the same minor iteration, timed on the
    host cores: oracle/ref.py (numpy + OpenMP C restatement of the numba
    kernels, scipy.fft/pocketfft standing in for ducc0.fft) with every host
    thread.  `nbs` bands of the workload are timed (the bounded sample of the
    cpu_baseline leg: 1 band, scaled by nband/nbs -- bands are independent apart
    from the cross-band sum); `full` (the --impl reference arm) times ALL bands,
    with the GPU arm's own input generator and a power-method hessnorm."""
    from functools import partial
    import numpy as np
    import torch
    from oracle import ref
    ncores = os.cpu_count() or 1
    # torchrun presets OMP_NUM_THREADS=1: force the thread count of the OpenMP oracle
    # kernels explicitly (main() also re-executes the reference arm with the variable set,
    # because libgomp and pocketfft's pool read it when they are loaded)
    os.environ["OMP_NUM_THREADS"] = str(ncores)
    torch.set_num_threads(ncores)
    nbs = a.nband if full else max(1, min(nbs, a.nband))
    nx, nxp = a.nx, 2 * a.nx
    dt = np.float32 if a.dtype == "f32" else np.float64
    cdt = np.complex64 if a.dtype == "f32" else np.complex128
    rdt = torch.float32 if a.dtype == "f32" else torch.float64
    psfhat = np.empty((nbs, nxp, nxp // 2 + 1), cdt)
    model = np.empty((nbs, nx, nx), dt)
    cpu = torch.device("cpu")
    for b in range(nbs):           # same generator as the GPU arm (make_band_inputs), on the host
        ph, m = make_band_inputs(torch, b, a.nband, nx, rdt, cpu)
        psfhat[b], model[b] = ph.numpy(), m.numpy()
        del ph, m
    # torch's intra-op pool is not needed any more; it shares libgomp's global thread-count
    # setting with the oracle, so the oracle's count is (re)set last
    torch.set_num_threads(1)
    ref.lib().orc_set_num_threads(ncores)
    xpad, xhat = np.empty((nbs, nxp, nxp), dt), np.empty(psfhat.shape, cdt)
    xo = np.empty_like(model)
    conv = partial(ref.psf_convolve_cube, xpad, xhat, xo, psfhat, nxp, nthreads=ncores)
    dirty = conv(model).copy()
    rng = np.random.default_rng(420)
    dirty += (1e-3 * dirty.max() * rng.standard_normal(dirty.shape)).astype(dt)
    del model
    psi = ref.Psi(nbs, nx, nx, BASES, NLEVEL, nthreads=ncores)
    shape_v = (nbs, len(BASES), psi.Nymax, psi.Nxmax)
    x, v = np.zeros_like(dirty), np.zeros(shape_v, dt)
    w = np.ones(shape_v[1:], dt)
    lam = 0.5 * float(np.std(dirty.sum(axis=0)))
    if full:
        # hessnorm = 1.05 * power_method(psf_convolve) as in the GPU arm (pfb/workers/spotless.py:187-193);
        # capped at 20 iterations to bound the CPU time of the arm
        b0 = rng.standard_normal(dirty.shape).astype(dt)
        beta, _ = ref.power_method(lambda z: conv(z).copy(), dirty.shape, b0=b0, tol=1e-3, maxit=20, verbosity=0)
        hessnorm, hn_src = 1.05 * float(beta), "1.05 * power_method(tol 1e-3, maxit 20)"
        del b0
    else:
        hessnorm, hn_src = 1.05 * nbs / a.nband, "1.05 * nbs/nband (PSF peaks at 1/nband per band)"

    def run(nit):
        ref.primal_dual_optimised(x, v, lam, psi.hdot, psi.dot, hessnorm, None, w, None,
                                  lambda z: conv(z) - dirty, nu=len(BASES), tol=0.0, maxit=nit)
    run(max(1, warmup))
    t0 = time.perf_counter()
    run(steps)
    dt_s = (time.perf_counter() - t0) / steps
    val = 1.0 / (dt_s * a.nband / nbs)
    try:
        nthr = int(ref.lib().orc_num_threads())
    except Exception:
        nthr = ncores
    scaled = "" if nbs == a.nband else f", scaled x{nbs}/{a.nband}"
    return {"value": round(val, 5), "unit": UNIT, "cores": nthr, "kind": "port",
            "sample": (f"{steps} timed primal-dual iterations (after {max(1, warmup)} warm-up) on {nbs} of the "
                       f"{a.nband} bands at full size {nx}x{nx}/PSF {nxp}x{nxp}{scaled}; "
                       f"{dt_s:.2f} s per sampled iteration; oracle port (OpenMP C wavelets/prox on {nthr} threads, "
                       f"scipy.fft workers={ncores} for ducc0.fft), host has {ncores} logical cores; "
                       f"inputs from the GPU arm's generator; hessnorm = {hn_src}"),
            "sec_per_sample_iter": round(dt_s, 3), "bands_timed": nbs, "extrapolated": nbs != a.nband}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return None
    steps = max(1, min(a.steps, 2))
    warm = 1
    cpu = cpu_reference(a, steps=steps, warmup=warm, full=True)
    return {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": UNIT,
            "n_gpus": a.gpus, "steps": steps, "warmup": warm,
            "ms_per_step": round(1e3 / cpu["value"], 2), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": a.dtype, "data": "synthetic",
            "config": {"workload": workload_name(a), "nband": a.nband, "nx": a.nx, "ny": a.nx,
                       "nxp": 2 * a.nx, "nyp": 2 * a.nx, "bases": BASES, "nlevels": NLEVEL, "positivity": 1,
                       "init": "x0=0, v0=0, l1weight=1, tol=0 (fixed iteration count)",
                       "note": ("all bands timed (no extrapolation); steps/warmup capped at 2/1: one CPU "
                                "iteration of the full cube takes about 15 s")},
            "cpu_baseline": cpu,
            "e2e": {"value": cpu["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def main():
    a = parse_args()
    if a.impl == "reference":
        # torchrun presets OMP_NUM_THREADS=1, which libgomp, pocketfft's pool and torch read when
        # they are loaded: restart this process with the host's core count before importing them
        if int(os.environ.get("RANK", "0")) != 0:
            return
        ncores = str(os.cpu_count() or 1)
        if os.environ.get("OMP_NUM_THREADS") != ncores and not os.environ.get("PFB_BENCH_REEXEC"):
            env = dict(os.environ, OMP_NUM_THREADS=ncores, PFB_BENCH_REEXEC="1")
            sys.stdout.flush()
            os.execve(sys.executable, [sys.executable] + sys.argv, env)
    # libraries (NCCL prints its version banner) must not pollute the one-line stdout contract:
    # everything written to fd 1 during the run goes to stderr, the JSON line to the real stdout
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    line = run_reference(a) if a.impl == "reference" else run_ours(a)
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    if line is not None:
        print(json.dumps(line), flush=True)
    bad_parity = False
    if line is not None and line.get("parity_vs_n1"):
        pv = line["parity_vs_n1"]
        bad_parity = not (max(pv["rel_l2_x"], pv["rel_l2_v"]) <= pv["tol"])
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
    except Exception:
        pass
    if bad_parity:
        print("bench.py: the band-sharded result differs from the single-GPU result beyond 1e-5", file=sys.stderr)
        sys.exit(3)


if __name__ == "__main__":
    main()
