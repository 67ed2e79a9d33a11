"""Drift of the band-sharded primal-dual paths over the reference's default iteration count
(pd_maxit 450, pfb/parser/pdopts.yml:6-8): fp32, 4 bands x 1024^2, on 2 ranks
    torchrun --nproc-per-node 2 profiles/sharded_drift.py [iters]
Compared, all against the fp64 single-GPU trajectory (same inputs rounded to fp32):
  single  fp32 single-GPU path
  rec     recursively carried band sum, re-synchronised every 64 iterations (default)
  rec0    the same without re-synchronisation
  slab    exchange of the coefficient band-sum slab (PFB_PAR_NOREC=1)
Prints one JSON line (rank 0)."""
import json
import os
import sys
from functools import partial

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_sharded import _problem, BASES  # noqa: E402

nit = int(sys.argv[1]) if len(sys.argv) > 1 else 450
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
import pfb_b200  # noqa: E402,F401
from pfb_b200 import parallel  # noqa: E402
from pfb_b200.operators import psf as P  # noqa: E402
from pfb_b200.operators.psi import Psi  # noqa: E402
from pfb_b200.opt.primal_dual import primal_dual_optimised, Grad21  # noqa: E402

nband, nx = 4, 1024
nxp = 2 * nx
nbl = nband // world
sl = slice(rank * nbl, (rank + 1) * nbl)
psfhat64, model64, noise64 = _problem(torch, nband, nx, torch.float64, dev)
dirty64 = P.psf_convolve_cube(None, None, torch.empty_like(model64), psfhat64, nxp, model64).clone()
dirty64 += 1e-3 * dirty64.max() * noise64
lam = 0.5 * float(dirty64.sum(dim=0).std())
P.clear_plan_cache()


def solve(rdt, psf_, dirty_, nb_):
    psi = Psi(nb_, nx, nx, BASES, 3, 1)
    conv = partial(P.psf_convolve_cube, None, None, torch.empty_like(dirty_), psf_, nxp)
    x = torch.zeros_like(dirty_)
    v = torch.zeros((nb_, 5, psi.Nymax, psi.Nxmax), dtype=rdt, device=dev)
    w = torch.ones(tuple(v.shape[1:]), dtype=rdt, device=dev)
    primal_dual_optimised(x, v, lam, psi.hdot, psi.dot, 1.05, None, w, None, Grad21(conv, dirty_), nu=5,
                          tol=0.0, maxit=nit, positivity=1, verbosity=0)
    P.clear_plan_cache()
    return x, v


def rel(a, b):
    return float(torch.linalg.vector_norm((a.double() - b.double()).view(-1)) / torch.linalg.vector_norm(b.double().view(-1)))


parallel.disable()
x64, v64 = solve(torch.float64, psfhat64, dirty64, nband)
psf32, dirty32 = psfhat64.to(torch.complex64), dirty64.float()
out = {"iters": nit, "nband": nband, "nx": nx, "world": world}
xs, vs = solve(torch.float32, psf32, dirty32, nband)
out["single"] = (rel(xs, x64), rel(vs, v64))
parallel.enable()
for name, env in (("rec", {"PFB_PAR_RESYNC": "64"}), ("rec0", {"PFB_PAR_RESYNC": "0"}), ("slab", {"PFB_PAR_NOREC": "1"})):
    os.environ.update(env)
    x, v = solve(torch.float32, psf32[sl].contiguous(), dirty32[sl].contiguous(), nbl)
    for k in env:
        del os.environ[k]
    e = torch.tensor([rel(x, x64[sl]) ** 2 * float((x64[sl] ** 2).sum()), rel(v, v64[sl]) ** 2 * float((v64[sl] ** 2).sum()),
                      float(((x.double() - xs[sl].double()) ** 2).sum()), float(((v.double() - vs[sl].double()) ** 2).sum())],
                     dtype=torch.float64, device=dev)
    dist.all_reduce(e)
    out[name] = {"vs_fp64": ((float(e[0]) / float((x64 ** 2).sum())) ** 0.5, (float(e[1]) / float((v64 ** 2).sum())) ** 0.5),
                 "vs_fp32_single": ((float(e[2]) / float((xs.double() ** 2).sum())) ** 0.5,
                                    (float(e[3]) / float((vs.double() ** 2).sum())) ** 0.5)}
if rank == 0:
    print(json.dumps(out))
dist.barrier()
dist.destroy_process_group()
