"""PCG iterations/s and per-kernel device times at a BASELINE configuration (iteration helper)."""
import ctypes, os, sys, time
from functools import partial
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pfb_b200  # noqa
from pfb_b200 import _lib
from pfb_b200.operators.hessian import hessian_psf_cube
from pfb_b200.opt.pcg import pcg, _pcg_psf_impl
nband = int(sys.argv[1]) if len(sys.argv) > 1 else 8
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
nit = int(sys.argv[3]) if len(sys.argv) > 3 else 10
dev = torch.device("cuda")
nxp = 2 * nx
if os.environ.get("PROBE_RANDOM"):
    g = torch.Generator(device=dev).manual_seed(1)
    psf = torch.zeros((nband, nxp, nxp), device=dev)
    ii = torch.arange(nxp, device=dev) - nxp // 2
    g1 = torch.exp(-0.5 * (ii / 2.0) ** 2)
    psf[:] = (g1[:, None] * g1[None, :]) / nband
    psfhat = torch.fft.rfft2(torch.fft.ifftshift(psf, dim=(1, 2)), dim=(1, 2))
    del psf
    b = torch.randn((nband, nx, nx), generator=g, device=dev)
else:           # bench.py's problem
    import bench
    P = bench.Problem(torch, nband, nx, "f32", list(range(nband)), dev)
    psfhat, b = P.psfhat, P.dirty
A = partial(hessian_psf_cube, None, None, None, None, psfhat, nxp, sigmainv=1e-3)
L = _lib.lib()
for mode in ("global", "per-band"):
    def run(n):
        if mode == "global":
            return pcg(A, b, torch.zeros_like(b), tol=0.0, maxit=n, minit=n, verbosity=0)
        return _pcg_psf_impl(psfhat, b, torch.zeros_like(b), None, nxp, 1, 1e-3, tol=0.0, maxit=n, minit=n, verbosity=0)
    run(3)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(nit)
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / nit * 1e3
    L.pfb_profile_enable(1)
    run(nit)
    torch.cuda.synchronize()
    L.pfb_profile_enable(0)
    buf = ctypes.create_string_buffer(1 << 16)
    _lib.check(L.pfb_profile_report(buf, len(buf)))
    print(f"{mode}: {wall:.3f} ms/iteration wall ({1e3 / wall:.1f} it/s), path {getattr(pcg, 'last_path', '?')}")
    tot = 0.0
    for line in buf.value.decode().splitlines():
        nm, n, t = line.split("\t")
        tot += float(t)
        print(f"    {float(t) / nit:8.4f} ms/it  x{int(n) / nit:.1f}  {nm.split('pfb')[-1][:50]}")
    print(f"    kernels total {tot / nit:.3f} ms/it")
