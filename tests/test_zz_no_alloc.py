"""pfb_seg_stats_static: the no-allocation variant of pfb_seg_stats (scratch in a static
device array, include/pfb_b200.h).  Called through the C-ABI and compared with the existing
entry point on two calls in a row that share the static array (the second with other data
and fewer segments, so stale contents of the first would show).  Collected last on
purpose."""
import ctypes

import numpy as np
import pytest

TOL = 1e-12        # same kernels, same order: identical up to nothing; fp64 bar of the glue tests


def _run_pair(L, ptr, dcode, stream, to_dev, scalars):
    rng = np.random.default_rng(31)
    calls = [(5, 4001, np.float32), (3, 777, np.float64), (64, 300, np.float32)]
    for nseg, m, dt in calls:
        x = to_dev((rng.standard_normal((nseg, m)) * (1 + np.arange(nseg))[:, None]).astype(dt))
        a, b = scalars(3 * nseg), scalars(3 * nseg)
        code = 0 if dt == np.float32 else 1
        assert L.pfb_seg_stats(ptr(x), nseg, m, ptr(a), code, stream()) == 0, L.pfb_last_error()
        assert L.pfb_seg_stats_static(ptr(x), nseg, m, ptr(b), code, stream()) == 0, L.pfb_last_error()
        ea, eb = np.asarray(a.cpu() if hasattr(a, "cpu") else a), np.asarray(b.cpu() if hasattr(b, "cpu") else b)
        assert np.all(np.isfinite(eb))
        assert np.max(np.abs(ea - eb) / np.maximum(np.abs(ea), 1e-300)) <= TOL
        xn = np.asarray(x.cpu() if hasattr(x, "cpu") else x, dtype=np.float64)
        ref = np.stack([xn.mean(axis=1), xn.std(axis=1), np.abs(xn).max(axis=1)], axis=1).ravel()
        assert np.max(np.abs(eb - ref) / np.abs(ref)) < 1e-11
    # beyond the static workspace: a clear error, nothing launched
    x = to_dev(np.ones((65, 8), np.float32))
    out = scalars(3 * 65)
    n0 = L.pfb_launch_count()
    rc = L.pfb_seg_stats_static(ptr(x), 65, 8, ptr(out), 0, stream())
    assert rc == 1 and b"at most 64" in L.pfb_last_error()
    assert L.pfb_launch_count() == n0


@pytest.mark.gpu
def test_seg_stats_static_matches_allocating_entry_point():
    import torch
    import pfb_b200._lib as Lm
    import pfb_b200._dev as D
    assert torch.cuda.is_available() and not D._EMULATED
    L = Lm.lib()
    _run_pair(L, D.ptr, D.dcode, D.stream, lambda a: torch.from_numpy(a).cuda(), D.scalars)
    torch.cuda.synchronize()
    # second round in the same process (address cached per device, array reused)
    _run_pair(L, D.ptr, D.dcode, D.stream, lambda a: torch.from_numpy(a).cuda(), D.scalars)


def test_seg_stats_static_emulated(emu_backend):
    """Same check on the CPU execution-model emulator build (host logic, limits)."""
    import torch
    import pfb_b200._lib as Lm
    import pfb_b200._dev as D
    _run_pair(Lm.lib(), D.ptr, D.dcode, D.stream, lambda a: torch.from_numpy(a), D.scalars)
