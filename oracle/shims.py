"""sys.modules shims so the UNMODIFIED reference hot-path modules import in the
build container (test infrastructure; only tests/golden/make_golden.py uses it,
and only where /root/reference exists -- it does not exist on the GPU box).

Missing third-party packages and what stands in:
  pywt      -> filter taps + dwt_max_level from oracle/filters.py
  ducc0.fft -> scipy.fft (pocketfft, the code base ducc0.fft grew from)
  ducc0.misc.make_noncritical -> identity
  numexpr.evaluate -> numpy evaluation of the two expressions the path uses
  dask / distributed / pyscilog -> inert stubs
  pfb.utils.misc -> stub exporting norm_diff restated from misc.py:1316-1351
"""
import os
import sys
import types

import numpy as np

REFERENCE = "/root/reference"


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install(reference=REFERENCE, numba_cache="/tmp/pfb_numba_cache"):
    if not os.path.isdir(reference):
        raise RuntimeError(f"{reference} not present: the reference cannot be "
                           "imported here (expected on the GPU box)")
    os.environ.setdefault("NUMBA_CACHE_DIR", numba_cache)
    os.makedirs(os.environ["NUMBA_CACHE_DIR"], exist_ok=True)
    here = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    if here not in sys.path:
        sys.path.insert(0, here)
    from oracle import filters
    import scipy.fft as sfft

    # ---- pywt ---------------------------------------------------------
    class Wavelet:
        def __init__(self, name):
            self.name = name
            self.filter_bank = tuple(list(f) for f in filters.filter_bank(name))

    _mod("pywt", Wavelet=Wavelet, dwt_max_level=filters.dwt_max_level)

    # ---- ducc0 --------------------------------------------------------
    def r2c(a, axes, nthreads=1, forward=True, inorm=0, out=None):
        assert forward and inorm == 0
        res = sfft.rfftn(a, axes=axes, workers=nthreads)
        if out is not None:
            out[...] = res
            return out
        return res

    def c2r(a, axes, forward=False, out=None, lastsize=None, inorm=2,
            nthreads=1, allow_overwriting_input=False):
        assert (not forward) and inorm == 2
        s = [a.shape[ax] for ax in axes]
        s[-1] = lastsize
        res = sfft.irfftn(a, s=s, axes=axes, workers=nthreads)
        if out is not None:
            out[...] = res
            return out
        return res

    def c2c(a, axes, forward=True, inorm=0, nthreads=1, out=None):
        f = sfft.fftn if forward else sfft.ifftn
        res = f(a, axes=axes, workers=nthreads,
                norm={0: "backward" if forward else "forward",
                      2: "forward" if forward else "backward"}[inorm])
        if out is not None:
            out[...] = res
            return out
        return res

    ducc0 = _mod("ducc0")
    ducc0.fft = _mod("ducc0.fft", r2c=r2c, c2r=c2r, c2c=c2c,
                     good_size=sfft.next_fast_len)
    ducc0.misc = _mod("ducc0.misc", make_noncritical=lambda x: x,
                      roll_resize_roll=None)
    ducc0.wgridder = _mod("ducc0.wgridder")
    ducc0.wgridder.experimental = _mod("ducc0.wgridder.experimental",
                                       vis2dirty=None, dirty2vis=None)
    ducc0.wgridder.ms2dirty = None
    ducc0.wgridder.dirty2ms = None

    # ---- numexpr ------------------------------------------------------
    def evaluate(expr, out=None, local_dict=None, casting=None):
        frame = sys._getframe(1)
        env = dict(frame.f_globals)
        env.update(frame.f_locals)
        if local_dict:
            env.update(local_dict)
        res = eval(expr, {"__builtins__": {}}, env)
        if out is not None:
            out[...] = res
            return out
        return res

    _mod("numexpr", evaluate=evaluate)

    # ---- dask / distributed / pyscilog -------------------------------
    class _Anything(types.ModuleType):
        def __getattr__(self, k):
            if k.startswith("__"):
                raise AttributeError(k)
            return None

    dask = _Anything("dask")
    sys.modules["dask"] = dask
    da = _Anything("dask.array")
    da.Array = type("Array", (), {})
    sys.modules["dask.array"] = da
    dask.array = da
    sys.modules["dask.graph_manipulation"] = _Anything("dask.graph_manipulation")
    sys.modules["distributed"] = _Anything("distributed")

    class _Log:
        def write(self, *a):
            pass

        def flush(self):
            pass

    _mod("pyscilog", get_logger=lambda *a, **k: _Log(), init=lambda *a: None)

    # ---- pfb.utils.misc (restated norm_diff, misc.py:1316-1351) -------
    def norm_diff(x, xp):
        d = x - xp                      # formed in the input dtype
        num = np.sum(np.square(d), dtype=np.float64)
        den = 1e-12 + np.sum(np.square(x), dtype=np.float64)
        return float(np.sqrt(num / den))

    if reference not in sys.path:
        sys.path.insert(0, reference)
    # the package __init__ of pfb and pfb.utils only define helpers; load the
    # real packages but replace the heavy pfb.utils.misc
    pfb = _mod("pfb")
    pfb.__path__ = [os.path.join(reference, "pfb")]
    utils = _mod("pfb.utils")
    utils.__path__ = [os.path.join(reference, "pfb", "utils")]
    pfb.utils = utils
    utils.misc = _mod("pfb.utils.misc", norm_diff=norm_diff)
    return True
