"""Per-kernel device times of the PSF Hessian (and optionally of one PD iteration's other
kernels) at a BASELINE configuration, from the library's own launch hook.  Iteration helper:
    python profiles/hess_probe.py [nband nx dtype reps]"""
import ctypes
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import pfb_b200  # noqa: E402,F401
from pfb_b200 import _lib  # noqa: E402
from pfb_b200.operators import psf as P  # noqa: E402

nband = int(sys.argv[1]) if len(sys.argv) > 1 else 8
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
dt = sys.argv[3] if len(sys.argv) > 3 else "f32"
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 20
rdt, cdt = (torch.float32, torch.complex64) if dt == "f32" else (torch.float64, torch.complex128)
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(1)
nxp = 2 * nx
psfhat = torch.randn((nband, nxp, nxp // 2 + 1, 2), generator=g, device=dev, dtype=rdt)
psfhat = torch.view_as_complex(psfhat).contiguous()
x = torch.randn((nband, nx, nx), generator=g, device=dev, dtype=rdt)
out = torch.empty_like(x)
L = _lib.lib()
for _ in range(3):
    P.psf_convolve_cube(None, None, out, psfhat, nxp, x)
torch.cuda.synchronize()
L.pfb_profile_enable(1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    P.psf_convolve_cube(None, None, out, psfhat, nxp, x)
e1.record()
torch.cuda.synchronize()
L.pfb_profile_enable(0)
buf = ctypes.create_string_buffer(1 << 16)
_lib.check(L.pfb_profile_report(buf, len(buf)))
tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("PFB_"))
print(f"[{tag}] {nband}x{nx}^2 {dt}: psf_convolve {e0.elapsed_time(e1) / reps:.4f} ms/call")
for line in buf.value.decode().splitlines():
    nm, n, t = line.split("\t")
    short = nm.split("pfb")[-1][:48]
    print(f"    {float(t) / int(n):8.4f} ms  x{int(n) // reps}  {short}")
