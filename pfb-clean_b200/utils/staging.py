"""Data formats either side of the minor cycle: dds2cubes
(pfb/utils/misc.py:664-739), which stitches the per-band, per-time-chunk datasets of the
.dds store into the cubes the minor cycle works on, and the per-dataset write-back of a major
iteration (pfb/workers/spotless.py:373-394) -- as pinned-memory staged transfers.

dds2cubes: every array of every dataset crosses PCIe exactly once, through two pinned
staging buffers on a copy stream, and is accumulated on the device (`dirty[b] += DIRTY *
BEAM`, `mean_beam[b] += BEAM * wsum`, ...: pfb_accum_mul) while the next array is in flight;
the cubes come out as device tensors, ready for the operators, without ever existing on the
host.  cubes2dds streams RESIDUAL * wsum, MODEL, DUAL, MODEL_BEST back band by band through
the same pinned buffers.  Datasets are duck-typed (xarray-like): `'NAME' in ds`,
`ds.NAME.data` (or `.values`), `ds.bandid`."""
import numpy as np
import torch

from .. import _dev, _lib
from ..opt import _vec


def _arr(var):
    a = getattr(var, "data", None)
    if a is None:
        a = getattr(var, "values", var)
    return np.ascontiguousarray(np.asarray(a))


class _Stager:
    """Two pinned staging buffers + a copy stream: upload k+1 overlaps the accumulation of k."""

    def __init__(self, nbytes, dev):
        self.dev = dev
        self.emu = _dev._EMULATED
        if not self.emu:
            self.bufs = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
            self.devb = [torch.empty(nbytes, dtype=torch.uint8, device=dev) for _ in range(2)]
            self.copy = torch.cuda.Stream()
            self.done = [torch.cuda.Event(), torch.cuda.Event()]      # H2D finished
            self.free = [torch.cuda.Event(), torch.cuda.Event()]      # consumer finished
            for e in self.free:
                e.record()
        self.k = 0

    def upload(self, a):
        """host array -> device tensor view (valid until two uploads later)."""
        t = torch.from_numpy(a)
        if self.emu:
            return t
        i = self.k & 1
        self.k += 1
        nb = a.nbytes
        self.free[i].synchronize()                 # the kernels that read this slot are done
        self.bufs[i][:nb].copy_(t.view(torch.uint8).view(-1))          # host -> pinned
        with torch.cuda.stream(self.copy):
            self.devb[i][:nb].copy_(self.bufs[i][:nb], non_blocking=True)
            self.done[i].record()
        torch.cuda.current_stream().wait_event(self.done[i])
        self._last = i
        return self.devb[i][:nb].view(t.dtype).view(a.shape)

    def release(self):
        if not self.emu:
            self.free[self._last].record()


def _accum(out, x, y=None, scale=1.0):
    n = x.numel() * (2 if x.is_complex() else 1)
    xr = torch.view_as_real(x) if x.is_complex() else x
    outr = torch.view_as_real(out) if out.is_complex() else out
    _lib.check(_lib.lib().pfb_accum_mul(_dev.ptr(outr), _dev.ptr(xr), _dev.ptr(y), float(scale), n,
                                        _dev.dcode(outr), _dev.stream()))


def dds2cubes(dds, nband, apparent=False, dual=True, modelname='MODEL'):
    """pfb/utils/misc.py:664-739.  Returns (dirty, model, residual, psf, psfhat, mean_beam, wsums,
    dual) with the cubes as DEVICE tensors (wsums as a numpy array of the real type)."""
    d0 = dds[0]
    real_np = _arr(d0.DIRTY).dtype
    rdt = torch.float32 if real_np == np.float32 else torch.float64
    cdt = torch.complex64 if rdt == torch.float32 else torch.complex128
    dev = _dev.device()
    nx, ny = _arr(d0.DIRTY).shape
    z = lambda shape, dt=rdt: torch.zeros((nband,) + tuple(shape), dtype=dt, device=dev)   # noqa: E731
    dirty, model, mean_beam = z((nx, ny)), z((nx, ny)), z((nx, ny))
    residual = z((nx, ny)) if 'RESIDUAL' in d0 else None
    psf = psfhat = None
    biggest = 2 * nx * ny
    if 'PSF' in d0:
        psf, psfhat = z(_arr(d0.PSF).shape), z(_arr(d0.PSFHAT).shape, cdt)
        biggest = max(biggest, psf[0].numel(), 2 * psfhat[0].numel())
    dualc = None
    if dual and 'DUAL' in d0:
        dualc = z(_arr(d0.DUAL).shape)
        biggest = max(biggest, dualc[0].numel())
    stg = _Stager(biggest * np.dtype(real_np).itemsize, dev)
    wsums = np.zeros(nband, dtype=real_np)
    for ds in dds:
        b = int(ds.bandid)
        w = _arr(ds.WSUM).reshape(-1)[0]
        beam = stg.upload(_arr(ds.BEAM))
        _accum(mean_beam[b], beam, None, float(w))            # mean_beam[b] += BEAM * WSUM[0]
        beam = beam.clone() if not apparent else None         # survives the next uploads
        stg.release()
        for name, cube in (("DIRTY", dirty), ("RESIDUAL", residual)):
            if cube is None or name not in ds:
                continue
            x = stg.upload(_arr(getattr(ds, name)))
            _accum(cube[b], x, beam)                          # += DIRTY (* BEAM)
            stg.release()
        if psf is not None and 'PSF' in ds:
            x = stg.upload(_arr(ds.PSF))
            _accum(psf[b], x)
            stg.release()
            x = stg.upload(_arr(ds.PSFHAT))
            _accum(psfhat[b], x)
            stg.release()
        if modelname in ds:
            model[b].copy_(stg.upload(_arr(getattr(ds, modelname))))
            stg.release()
        if dualc is not None and 'DUAL' in ds:
            dualc[b].copy_(stg.upload(_arr(ds.DUAL)))
            stg.release()
        wsums[b] = wsums[b] + w
    wsum = wsums.sum()

    def scale(t, f):                # t *= f with the library's own axpby kernel
        tr = torch.view_as_real(t) if t.is_complex() else t
        _vec.axpby(tr, float(f), tr, 0.0, tr)
    for cube in (dirty, residual, psf, psfhat):
        if cube is not None:
            scale(cube, 1.0 / float(wsum))
    for b in range(nband):
        if wsums[b]:
            scale(mean_beam[b], 1.0 / float(wsums[b]))
    return dirty, model, residual, psf, psfhat, mean_beam, wsums, dualc


def cubes2dds(dds, residual, model, dual, best_model, wsum):
    """The write-back of pfb/workers/spotless.py:373-394: for every dataset a dict of HOST arrays
    {RESIDUAL: residual[b] * wsum, MODEL: model[b], DUAL: dual[b], MODEL_BEST: best_model[b]},
    streamed through pinned memory (one D2H per band and variable, shared by the datasets of a band)."""
    cache = {}

    def band(b):
        if b not in cache:
            r = _dev.to_dev(residual)[b] * float(wsum)
            out = {}
            for name, t in (("RESIDUAL", r), ("MODEL", _dev.to_dev(model)[b]), ("DUAL", _dev.to_dev(dual)[b]),
                            ("MODEL_BEST", _dev.to_dev(best_model)[b])):
                if _dev._EMULATED:
                    out[name] = t.clone().numpy()
                else:
                    h = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                    h.copy_(t, non_blocking=True)
                    out[name] = h
            cache[b] = out
        return cache[b]
    res = [band(int(ds.bandid)) for ds in dds]
    if not _dev._EMULATED:
        torch.cuda.current_stream().synchronize()
        for b, d in cache.items():
            for k in d:
                d[k] = d[k].numpy()
    return res
