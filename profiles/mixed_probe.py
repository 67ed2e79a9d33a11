"""Timing of the mixed-radix Hessian path (hess_mixed.cuh) on one band of a
non-power-of-two cube; CUDA events, inputs larger than L2."""
import json
import sys
import torch
sys.path.insert(0, ".")
import pfb_b200  # noqa
from pfb_b200.operators.psf import psf_convolve_cube

res = {}
for nband, nx, nxp in [(1, 10240, 20480), (2, 6000, 12000), (4, 3000, 6000)]:
    g = torch.Generator(device="cuda").manual_seed(1)
    psfhat = torch.randn((nband, nxp, nxp // 2 + 1), generator=g, dtype=torch.complex64, device="cuda")
    x = torch.randn((nband, nx, nx), generator=g, dtype=torch.float32, device="cuda")
    out = torch.empty_like(x)
    for _ in range(3):
        psf_convolve_cube(None, None, out, psfhat, nxp, x)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    n = 5
    for _ in range(n):
        psf_convolve_cube(None, None, out, psfhat, nxp, x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    balg = nband * (2 * nx * nx * 4 + 8 * nxp * (nxp // 2 + 1))
    res[f"{nband}x{nx}^2/{nxp}^2"] = {"ms": ms, "alg_GBps": balg / ms / 1e6}
    del psfhat, x, out
print(json.dumps(res))
