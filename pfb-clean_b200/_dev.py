"""Device-buffer plumbing shared by the operator modules.

PyTorch is used only for device memory, streams and (in pfb_b200.parallel)
torch.distributed.  Every operator accepts either torch CUDA tensors (device
resident, zero copies) or numpy arrays (the reference's calling convention:
inputs are copied host->device, results are copied back into the caller's
array in place).  There is no CPU execution path: without a CUDA device or
without the built library every call raises.
"""
import ctypes

import numpy as np
import torch

from . import _lib

# tests/emu only: lets the CPU execution-model emulator stand in for the GPU so
# kernel sources and host logic can be checked in a GPU-less container.  Never
# set by product code.
_EMULATED = False


def device():
    if _EMULATED:
        return torch.device("cpu")
    if not torch.cuda.is_available():
        raise RuntimeError("pfb_b200 requires a CUDA device (B200); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def stream():
    if _EMULATED:
        return None
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def sync():
    if not _EMULATED:
        torch.cuda.current_stream().synchronize()


def is_host(a):
    return isinstance(a, np.ndarray)


def to_dev(a, copy=False):
    """numpy / torch -> contiguous tensor on the compute device."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(a))
        if _EMULATED:
            return t.clone() if copy else t
        return t.to(device(), non_blocking=True)
    if not isinstance(a, torch.Tensor):
        raise TypeError(f"expected numpy array or torch tensor, got {type(a)}")
    if a.device != device():
        a = a.to(device(), non_blocking=True)
    if not a.is_contiguous():
        a = a.contiguous()
    return a.clone() if copy else a


def to_out(result, out):
    """Write a device result into the caller's output object (in place)."""
    if isinstance(out, np.ndarray):
        if _EMULATED:
            if result.data_ptr() != out.ctypes.data:
                out[...] = result.numpy()
        else:
            torch.from_numpy(out).copy_(result, non_blocking=False)
        return out
    if out.data_ptr() != result.data_ptr():
        out.copy_(result)
    return out


def out_dev(out):
    """Device tensor to compute into for the caller's output `out`: the tensor
    itself when it already lives on the device (and is contiguous), else a
    fresh device tensor (copied back by to_out)."""
    if isinstance(out, torch.Tensor) and out.device == device() and out.is_contiguous():
        return out
    if isinstance(out, np.ndarray) and _EMULATED and out.flags.c_contiguous:
        return torch.from_numpy(out)
    if isinstance(out, np.ndarray):
        return torch.empty(out.shape, dtype=torch.from_numpy(out[..., :0]).dtype, device=device())
    return torch.empty(out.shape, dtype=out.dtype, device=device())


def like(a, result):
    """Return `result` (device tensor) in the flavour of `a` (numpy or torch)."""
    if isinstance(a, np.ndarray):
        return result.cpu().numpy() if not _EMULATED else result.numpy()
    return result


def dcode(t):
    if t.dtype in (torch.float32, torch.complex64):
        return _lib.F32
    if t.dtype in (torch.float64, torch.complex128):
        return _lib.F64
    raise TypeError(f"unsupported dtype {t.dtype}: float32/float64 only")


def ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def scalars(n):
    """Small fp64 device buffer for reduction results."""
    return torch.zeros(n, dtype=torch.float64, device=device())
