"""CLEAN minor loops: pfb/deconv/clark.py (clark, subminor) and
pfb/deconv/hogbom.py against golden vectors produced by the unmodified reference
(tests/golden/make_golden.py gen_clean): the oracle restatement first (bit-exact), then the
product on the emulator build and -- marked gpu -- on libpfb_b200.so.  The component
positions are argmax decisions, so parity means the SAME components: the fp64 bar is the
operator bar 1e-11 (the residual goes through the PSF convolution), fp32 1e-5."""
import numpy as np
import pytest

from conftest import golden
from oracle import ref

pmp = pytest.mark.parametrize


def rel(a, b):
    return np.linalg.norm((np.asarray(a, np.float64) - b).ravel()) / np.linalg.norm(np.asarray(b, np.float64).ravel())


@pmp("t", ["f64", "f32"])
def test_oracle_clean_vs_golden(t):
    g = golden("clean_" + t)
    m, s = ref.clark(g["dirty"], g["psf"], g["psfhat"], g["wsums"], threshold=0, gamma=0.1, pf=0.05, maxit=6,
                     subpf=0.5, submaxit=40)
    assert np.array_equal(m, g["clark_model"]) and s == int(g["clark_status"])
    IRs = np.sum(g["dirty"], axis=0) ** 2
    m1 = ref.subminor(g["dirty"][:, g["sub_Ip"], g["sub_Iq"]], g["psf"], g["sub_Ip"], g["sub_Iq"],
                      np.zeros_like(g["dirty"]), g["wsums"], gamma=0.1, th=0.4 * np.sqrt(IRs.max()), maxit=25)
    assert np.array_equal(m1, g["sub_model"])
    x, st = ref.hogbom(g["dirty"], g["psf"], threshold=0, gamma=0.1, pf=0.1, maxit=60)
    assert np.array_equal(x, g["hogbom_x"]) and st == int(g["hogbom_status"])


@pmp("t", ["f64", "f32"])
def test_clean_vs_golden(backend, t):
    from pfb_b200.deconv.clark import clark, subminor
    from pfb_b200.deconv.hogbom import hogbom
    g = golden("clean_" + t)
    tol = 1e-11 if t == "f64" else 1e-5
    # sub-minor loop alone: no FFT involved, the same arithmetic in the same order -> identical
    IRs = np.sum(g["dirty"], axis=0) ** 2
    m0 = np.zeros_like(g["dirty"])
    m1 = subminor(g["dirty"][:, g["sub_Ip"], g["sub_Iq"]], g["psf"], g["sub_Ip"], g["sub_Iq"], m0, g["wsums"],
                  gamma=0.1, th=0.4 * np.sqrt(IRs.max()), maxit=25)
    assert m1 is m0
    assert np.array_equal(np.nonzero(m1), np.nonzero(g["sub_model"]))
    assert np.array_equal(m1, g["sub_model"])
    # hogbom: no FFT either
    x, st = hogbom(g["dirty"], g["psf"], threshold=0, gamma=0.1, pf=0.1, maxit=60, verbosity=0)
    assert st == int(g["hogbom_status"])
    assert np.array_equal(np.nonzero(x), np.nonzero(g["hogbom_x"]))
    assert np.array_equal(x, g["hogbom_x"])
    # clark: the residual of every minor iteration comes from the PSF convolution kernels
    m, s = clark(g["dirty"], g["psf"], g["psfhat"], g["wsums"], threshold=0, gamma=0.1, pf=0.05, maxit=6,
                 subpf=0.5, submaxit=40, verbosity=0)
    assert s == int(g["clark_status"])
    assert np.array_equal(np.nonzero(m), np.nonzero(g["clark_model"]))
    assert rel(m, g["clark_model"]) < tol


def test_hogbom_converges_before_maxit(backend):
    """Threshold exit inside a batch: the queued tail of the batch must be a no-op."""
    from pfb_b200.deconv.hogbom import hogbom
    g = golden("clean_f64")
    x, st = hogbom(g["dirty"], g["psf"], threshold=0, gamma=0.3, pf=0.5, maxit=500, verbosity=0)
    ox, ost = ref.hogbom(g["dirty"], g["psf"], threshold=0, gamma=0.3, pf=0.5, maxit=500)
    assert st == ost == 0
    assert np.array_equal(x, ox)


def test_peak_and_active_set(backend):
    """First-index tie-breaking of the peak search and the ordering of the active set."""
    import torch
    from pfb_b200 import _dev
    from pfb_b200.deconv.clark import _peak, _active_set
    rng = np.random.default_rng(8)
    nband, n = 3, 70001
    IR = rng.standard_normal((nband, n))
    IR[:, 5000] = IR[:, 123]                      # exact ties of the squared band sum ...
    IR[:, [123, 5000, 69999]] *= 50.0             # ... among the largest values
    IR[:, 69999] = IR[:, 123]
    d = _dev.to_dev(IR)
    search = torch.empty(n, dtype=d.dtype, device=d.device)
    pq, val = _peak(d, search, _dev.scalars(2))
    s = np.sum(IR, axis=0) ** 2
    assert pq == int(s.argmax()) == 123 and val == s[123]
    sd = search.cpu().numpy() if hasattr(search, "cpu") else search.numpy()
    assert np.array_equal(sd, s)
    thr = float(np.sort(s)[-4000])
    idx = torch.empty(n, dtype=torch.int64, device=d.device)
    counts = torch.empty((n + 2047) // 2048, dtype=torch.int32, device=d.device)
    cnt = _active_set(search, thr, idx, torch.zeros(1, dtype=torch.int64, device=d.device), counts)
    exp = np.nonzero(s > thr)[0]
    assert cnt == exp.size
    assert np.array_equal(idx[:cnt].cpu().numpy(), exp)
