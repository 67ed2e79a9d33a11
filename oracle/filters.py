"""Daubechies db1..db5 filter banks (test infrastructure, see oracle/__init__.py).

The reference takes its taps from PyWavelets (``pywt.Wavelet(w).filter_bank``,
/root/reference/pfb/operators/psi.py:37-41); PyWavelets is an un-vendored,
un-pinned dependency (/root/reference/setup.py:11) that is not installed here.
The taps below are the Daubechies minimum-phase scaling filters obtained by
spectral factorisation with mpmath at 60 digits (roots of
P(y) = sum_k C(N-1+k,k) y^k, y = (2 - z - 1/z)/4, zeros inside the unit circle,
normalised to sum sqrt(2)) and rounded to the nearest double.  They agree with
the 17-digit table recalled in SURVEY.md 8(c) to < 4e-12 and satisfy
orthonormality to 1 ulp.  The SAME table is compiled into the CUDA library
(pfb-clean_b200/csrc/filters.h) so parity does not depend on the last digits.

Filter-bank relations (PyWavelets convention, verified by perfect
reconstruction through the reference's own dwt2d/idwt2d):
    rec_lo = dec_lo[::-1];  rec_hi[k] = (-1)**k * dec_lo[k];  dec_hi = rec_hi[::-1]
"""
import numpy as np

DEC_LO = {
    "db1": [0.7071067811865476, 0.7071067811865476],
    "db2": [-0.12940952255126037, 0.2241438680420134,
            0.8365163037378079, 0.48296291314453416],
    "db3": [0.03522629188570953, -0.08544127388202666, -0.13501102001025458,
            0.45987750211849154, 0.8068915093110925, 0.33267055295008263],
    "db4": [-0.010597401785069032, 0.0328830116668852, 0.030841381835560764,
            -0.18703481171909309, -0.027983769416859854, 0.6308807679298589,
            0.7148465705529157, 0.2303778133088965],
    "db5": [0.0033357252854737712, -0.012580751999081999,
            -0.006241490212798274, 0.07757149384004572,
            -0.032244869584638375, -0.24229488706638203,
            0.13842814590132074, 0.7243085284377729, 0.6038292697971896,
            0.16010239797419293],
}


def filter_bank(name):
    """(dec_lo, dec_hi, rec_lo, rec_hi) as float64 arrays, pywt order."""
    dec_lo = np.array(DEC_LO[name], dtype=np.float64)
    rec_lo = dec_lo[::-1].copy()
    sign = (-1.0) ** np.arange(dec_lo.size)
    rec_hi = sign * dec_lo
    dec_hi = rec_hi[::-1].copy()
    return dec_lo, dec_hi, rec_lo, rec_hi


def dwt_max_level(data_len, name):
    """pywt.dwt_max_level: floor(log2(data_len / (filter_len - 1)))."""
    flen = len(DEC_LO[name])
    if flen < 2 or data_len < flen - 1:
        return 0
    return max(int(np.floor(np.log2(data_len / (flen - 1.0)))), 0)
