"""Band sharding across the GPUs of one box (one process per GPU,
torch.distributed over NCCL/NVLink; gloo on CPU for tests).

The image cube shards over frequency bands: Hessian, Psi.dot and
Psi.hdot are band-local.  Only two things cross ranks:
  * the band-sum slab of the l21 dual update (SUM all-reduce, C1 elements),
  * a handful of fp64 solver scalars (eps terms, <r,y>, <p,Ap>, ...).
The reference's distributed analogue is primal_dual_dist / get_ratio
(pfb/opt/primal_dual.py:187-306), a gather-to-worker-0 + broadcast over dask.

enable() switches the solvers in pfb_b200.opt to the sharded code path; every
rank then passes its LOCAL bands to the solver.
"""
import torch
import torch.distributed as dist

_STATE = {"on": False, "group": None}


def enable(group=None):
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised")
    _STATE["on"] = True
    _STATE["group"] = group


def disable():
    _STATE["on"] = False
    _STATE["group"] = None


def active():
    return _STATE["on"] and dist.get_world_size(_STATE["group"]) > 1


def world_size():
    return dist.get_world_size(_STATE["group"]) if _STATE["on"] else 1


def rank():
    return dist.get_rank(_STATE["group"]) if _STATE["on"] else 0


def allreduce_sum(t):
    """In-place SUM all-reduce (no-op unless sharding is active)."""
    if active():
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=_STATE["group"])
    return t


def allreduce_sum_async(t):
    """SUM all-reduce issued asynchronously (NCCL runs it on its own stream after
    the work already queued on the current stream); returns a work handle whose
    .wait() makes the current stream wait for the result, or None when inactive."""
    if not active():
        return None
    return dist.all_reduce(t, op=dist.ReduceOp.SUM, group=_STATE["group"], async_op=True)


def allreduce_max(t):
    if active():
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=_STATE["group"])
    return t


def band_slice(nband):
    """Contiguous block of bands owned by this rank (nband % world == 0)."""
    w, r = world_size(), rank()
    if nband % w:
        raise ValueError(f"nband={nband} is not divisible by the {w} ranks")
    per = nband // w
    return slice(r * per, (r + 1) * per)
