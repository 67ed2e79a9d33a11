"""Restoring convolution (pfb_b200.utils.restore) against golden vectors produced by the
unmodified reference functions (pfb/utils/misc.py:109-238; tests/golden/make_golden.py
gen_restore), on the CPU emulator build and on the B200."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))


def rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-300)


def test_good_size_and_padding():
    from pfb_b200.utils.restore import good_size, get_padding_info
    import scipy.fft as sfft
    for n in list(range(1, 400)) + [1000, 3360, 4097, 6000, 10241, 15000]:
        assert good_size(n, True) == sfft.next_fast_len(n, real=True), n
        assert good_size(n, False) == sfft.next_fast_len(n, real=False), n
    g = np.load(os.path.join(HERE, "golden", "restore_f64.npz"))
    padding, ux, uy = get_padding_info(64, 48, 0.5)
    assert np.array_equal(np.array(padding), g["padding"])
    assert (ux, uy) == (slice(16, -16), slice(12, -12))


def test_gaussian2d_and_convolve2gaussres(backend):
    from pfb_b200.utils.restore import Gaussian2D, convolve2gaussres
    g = np.load(os.path.join(HERE, "golden", "restore_f64.npz"))
    xx, yy, image = g["xx"], g["yy"], g["image"]
    gf = tuple(g["gaussparf"])
    k = Gaussian2D(xx, yy, gf, normalise=False)
    assert isinstance(k, np.ndarray) and k.dtype == np.float64 and k.shape == xx.shape
    assert rel(k, g["kern"]) < 1e-14 and np.array_equal(k == 0, g["kern"] == 0)
    assert rel(Gaussian2D(xx, yy, gf, normalise=True), g["kern_norm"]) < 1e-14
    tol = 1e-12
    assert rel(convolve2gaussres(image, xx, yy, gf, 1), g["conv"]) < tol
    assert rel(convolve2gaussres(image, xx, yy, gf, 1, norm_kernel=True), g["conv_norm"]) < tol
    assert rel(convolve2gaussres(image, xx, yy, gf, 1, pfrac=0.25), g["conv_p25"]) < tol
    gi = tuple(tuple(r) for r in g["gausspari"])
    # the spectra of the initial kernels fall to round-off level at high frequencies: the quotient
    # amplifies whatever FFT rounding sits there (the reference's result has the same sensitivity)
    assert rel(convolve2gaussres(image, xx, yy, gf, 1, gausspari=gi), g["conv_from"]) < 1e-6
    # fp32 images run in fp32
    out32 = convolve2gaussres(image.astype(np.float32), xx, yy, gf, 1)
    assert out32.dtype == np.float32 and rel(out32, g["conv"]) < 2e-6
    with pytest.raises(ValueError):
        convolve2gaussres(image[:, :, :25], xx[:, :25], yy[:, :25], gf, 1, pfrac=0.1)   # 27: odd lastsize
