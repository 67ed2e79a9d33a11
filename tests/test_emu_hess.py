"""hess.cu / fft.cuh (three-pass FFT Hessian) on the CPU emulator vs the
reference's golden vectors, the oracle and a direct-summation known answer."""
import ctypes

import numpy as np
import pytest

from oracle import ref
from conftest import golden
from emu.emulib import emu, ptr, ok

pmp = pytest.mark.parametrize


def rel(a, b):
    return np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300)


class EmuHess:
    def __init__(self, nband, nx, ny, nxp, nyp, dtype):
        self.L = emu()
        self.plan = ctypes.c_void_p()
        self.rc = self.L.pfb_hess_plan_create(ctypes.byref(self.plan), nband, nx, ny, nxp, nyp,
                                              0 if dtype == np.float32 else 1)

    def set_psfhat(self, psfhat):
        ok(self.L.pfb_hess_plan_set_psfhat(self.plan, ptr(np.ascontiguousarray(psfhat)), None))

    def apply(self, x, beam=None, add_x=0, sigmainv=0.0, wsum=None):
        out = np.full(x.shape, np.nan, x.dtype)
        ok(self.L.pfb_hess_apply(self.plan, ptr(x), ptr(beam), add_x, sigmainv,
                                 int(wsum is not None), float(wsum or 0.0), ptr(out), None))
        return out

    def __del__(self):
        if self.rc == 0:
            self.L.pfb_hess_plan_destroy(self.plan)


@pmp("tag", ["hess_f64", "hess_f32", "hess_f64_rect"])
def test_hess_golden(tag):
    g = golden(tag)
    x, psfhat, beam = g["x"], g["psfhat"], g["beam"]
    nband, nx, ny = x.shape
    nxp, nyp = psfhat.shape[1], g["psf"].shape[2]
    H = EmuHess(nband, nx, ny, nxp, nyp, x.dtype.type)
    assert H.rc == 0
    H.set_psfhat(psfhat)
    tol = 1e-13 if x.dtype == np.float64 else 1e-6
    assert rel(H.apply(x), g["conv"]) < tol
    assert rel(H.apply(x, None, 1, 1e-3), g["h1"]) < tol
    assert rel(H.apply(x, beam, 1, 0.25, 3.5), g["h2"]) < tol


@pmp("dtype", [np.float32, np.float64])
@pmp("nx,ny,nxp,nyp", [(8, 6, 16, 32), (20, 40, 64, 128), (100, 64, 256, 128), (30, 200, 32, 512),
                             (6, 10, 4096, 64), (4, 10, 16, 8192), (5, 4, 8192, 32),
                             (5, 4, 16384, 32),                                    # column pass on k_hessf_cols1<8192>
                             (4, 10, 16, 16384),                                   # row passes on the one-group kernels (M = 8192)
                             (10, 12, 1024, 1024), (7, 30, 2048, 1024)])   # fast path (hess_fast.cuh)
def test_hess_oracle(dtype, nx, ny, nxp, nyp):
    if nxp == 16384 and dtype == np.float64:
        pytest.skip("one fp64 line of 16384 points exceeds shared memory (PFB_ERR_FFT_SIZE)")
    rng = np.random.default_rng(11)
    nband = 2
    cdt = np.complex64 if dtype == np.float32 else np.complex128
    psf = rng.standard_normal((nband, nxp, nyp))
    psfhat = np.fft.rfft2(np.fft.ifftshift(psf, axes=(1, 2)), axes=(1, 2)).astype(cdt)
    x = rng.standard_normal((nband, nx, ny)).astype(dtype)
    H = EmuHess(nband, nx, ny, nxp, nyp, dtype)
    assert H.rc == 0
    H.set_psfhat(psfhat)
    xpad = np.empty((nband, nxp, nyp), dtype)
    xhat = np.empty(psfhat.shape, cdt)
    xout = np.empty(x.shape, dtype)
    exp = ref.psf_convolve_cube(xpad, xhat, xout, psfhat, nyp, x)
    assert rel(H.apply(x), exp) < (1e-13 if dtype == np.float64 else 1e-6)


def test_hess_delta_kat():
    """x = delta at (p,q)  =>  out[i,j] = psf[nxp//2+i-p, nyp//2+j-q]."""
    rng = np.random.default_rng(12)
    nx, ny, nxp, nyp = 16, 16, 32, 32
    psf = rng.standard_normal((1, nxp, nyp))
    psfhat = np.fft.rfft2(np.fft.ifftshift(psf, axes=(1, 2)), axes=(1, 2))
    H = EmuHess(1, nx, ny, nxp, nyp, np.float64)
    H.set_psfhat(psfhat)
    x = np.zeros((1, nx, ny))
    p, q = 5, 11
    x[0, p, q] = 1.0
    out = H.apply(x)
    ii, jj = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    assert np.abs(out[0] - psf[0, nxp // 2 + ii - p, nyp // 2 + jj - q]).max() < 1e-13


# padded sizes that are not powers of two (ducc0 good_size lengths made even,
# pfb/workers/grid.py:276-284): mixed-radix path, split and full column pass
MIXED = [(8, 8, 24, 32), (8, 8, 16, 16), (5, 6, 10, 12), (12, 14, 24, 28), (9, 20, 30, 42),
         (20, 18, 44, 36), (13, 10, 26, 52), (30, 50, 60, 100), (40, 24, 48, 50),   # nx > nxp/2: full
         (7, 9, 15, 18), (6, 4, 21, 10),                                           # odd nxp: full
         (48, 70, 96, 140), (75, 66, 150, 132), (10, 600, 20, 1200), (350, 8, 700, 16),
         # (3, 5, 7) x power of two: hess_peel.cuh -- rows (M = 1536, 2560), split columns
         # (H = 1536, 3584), full columns (nxp = 3072), both at once
         (6, 1500, 12, 3072), (2, 2500, 4, 5120), (1500, 6, 3072, 12), (3000, 4, 7168, 8),
         (2000, 6, 3072, 12), (2, 3500, 4, 7168),       # both passes peeled at once: GPU suite
         (5, 1499, 10, 3072), (1537, 3, 3072, 6)]        # odd image sizes on peeled plans


@pmp("dtype", [np.float32, np.float64])
@pmp("nx,ny,nxp,nyp", MIXED)
def test_hess_mixed_radix_oracle(dtype, nx, ny, nxp, nyp):
    rng = np.random.default_rng(13)
    nband = 2
    cdt = np.complex64 if dtype == np.float32 else np.complex128
    psf = rng.standard_normal((nband, nxp, nyp))
    psfhat = np.fft.rfft2(np.fft.ifftshift(psf, axes=(1, 2)), axes=(1, 2)).astype(cdt)
    x = rng.standard_normal((nband, nx, ny)).astype(dtype)
    beam = (0.5 + rng.random(x.shape)).astype(dtype)
    H = EmuHess(nband, nx, ny, nxp, nyp, dtype)
    assert H.rc == 0
    # which passes run as (3|5|7) x 2^k peeled transforms (hess_peel.cuh)
    def peels(n):
        return n <= 10240 and any(n % m == 0 and n // m in (512, 1024, 2048, 4096) for m in (3, 5, 7))
    kern = H.L.pfb_hess_plan_kernels(H.plan)
    assert kern & 2
    assert bool(kern & 4) == peels(nyp // 2)
    assert bool(kern & 8) == peels(nxp // 2 if (nxp % 2 == 0 and nx <= nxp // 2) else nxp)
    H.set_psfhat(psfhat)
    xpad, xhat = np.empty((nband, nxp, nyp), dtype), np.empty(psfhat.shape, cdt)
    xout = np.empty(x.shape, dtype)
    tol = 1e-13 if dtype == np.float64 else 1e-6
    exp = ref.psf_convolve_cube(xpad, xhat, xout, psfhat, nyp, x).copy()
    assert rel(H.apply(x), exp) < tol
    exp2 = ref.hessian_psf_cube(xpad, xhat, xout, beam, psfhat, nyp, x, sigmainv=0.25, wsum=3.5)
    assert rel(H.apply(x, beam, 1, 0.25, 3.5), exp2) < tol


def test_hess_peel_two_buffer_rows(monkeypatch):
    """Row passes of the peeled kernels with separate natural-order / segment buffers (what runs
    when a line's thread groups do not fit one CTA) instead of the in-place lines."""
    monkeypatch.setenv("PFB_HESS_PEEL_NOINPL", "1")
    test_hess_mixed_radix_oracle(np.float32, 6, 1500, 12, 3072)
    test_hess_mixed_radix_oracle(np.float64, 3, 2500, 6, 5120)
    # first form of the peeled column pass (natural-order staging line), split and full
    monkeypatch.setenv("PFB_HESS_PEEL_COLS1", "1")
    test_hess_mixed_radix_oracle(np.float32, 1500, 6, 3072, 12)
    test_hess_mixed_radix_oracle(np.float64, 2000, 6, 3072, 12)


def test_hess_plan_errors():
    assert EmuHess(1, 8, 8, 24, 33, np.float64).rc == 4      # odd lastsize
    assert EmuHess(1, 8, 8, 34, 32, np.float64).rc == 4      # nxp/2 = 17: prime factor > 13
    assert EmuHess(1, 8, 8, 32, 76, np.float64).rc == 4      # nyp/2 = 38 = 2 * 19
    assert EmuHess(1, 8, 8, 32, 2 * 15000, np.float64).rc == 4   # two lines beyond shared memory
    assert EmuHess(1, 40, 8, 32, 32, np.float64).rc == 1     # image larger than PSF


@pmp("dtype", [np.float32, np.float64])
def test_fused_primal_pass_emulated(dtype):
    """pfb_hess_apply_primal (primal step of primal_dual_optimised as the epilogue of pass C,
    pipelined and basic kernels of hess_fast.cuh) against pfb_hess_apply + pfb_pd_primal_lin:
    bit-identical model, eps terms equal to rounding."""
    rng = np.random.default_rng(31)
    nband, nx, ny, nxp, nyp = 1, 6, 400, 1024, 1024
    cdt = np.complex64 if dtype == np.float32 else np.complex128
    psf = rng.standard_normal((nband, nxp, nyp))
    psfhat = np.fft.rfft2(np.fft.ifftshift(psf, axes=(1, 2)), axes=(1, 2)).astype(cdt)
    H = EmuHess(nband, nx, ny, nxp, nyp, dtype)
    assert H.rc == 0 and H.L.pfb_hess_plan_kernels(H.plan) & 1
    H.set_psfhat(psfhat)
    assert H.L.pfb_hess_plan_can_fuse_primal(H.plan) == 1
    xp, w, wprev, data = (rng.standard_normal((nband, nx, ny)).astype(dtype) for _ in range(4))
    code = 0 if dtype == np.float32 else 1
    conv = H.apply(xp)
    for pos in (1,):                             # positivity 0: GPU suite
        x_ref = np.full(xp.shape, np.nan, dtype)
        t_ref = np.zeros(2)
        ok(H.L.pfb_pd_primal_lin(ptr(x_ref), ptr(xp), ptr(w), ptr(wprev), ptr(conv), ptr(data), 0.37, pos, nband,
                                 nx * ny, code, ptr(t_ref), None))
        x = np.full(xp.shape, np.nan, dtype)
        t = np.zeros(2)
        ok(H.L.pfb_hess_apply_primal(H.plan, ptr(xp), ptr(w), ptr(wprev), ptr(data), 0.37, pos, ptr(x), ptr(t), None))
        assert np.array_equal(x, x_ref)
        assert np.allclose(t, t_ref, rtol=1e-12, atol=0)
    # the two halves give the same result as the single call
    x2, t2 = np.full(xp.shape, np.nan, dtype), np.zeros(2)
    ok(H.L.pfb_hess_apply_primal_begin(H.plan, ptr(xp), None))
    ok(H.L.pfb_hess_apply_primal_end(H.plan, ptr(xp), ptr(w), ptr(wprev), ptr(data), 0.37, 1, ptr(x2), ptr(t2), None))
    assert np.array_equal(x2, x) and np.array_equal(t2, t)
    assert H.L.pfb_hess_apply_primal(H.plan, ptr(xp), ptr(w), ptr(wprev), ptr(data), 0.37, 1, ptr(xp), ptr(t), None) == 1
