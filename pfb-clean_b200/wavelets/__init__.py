"""Size helpers of pfb/wavelets/wavelets.py (coeff_size :20-22, signal_size
:25-27).  The transforms themselves are reached through
pfb_b200.operators.psi.Psi (the live callers of the reference only use Psi)."""


def coeff_size(nsignal, nfilter):
    return (nsignal + nfilter - 1) // 2


def signal_size(ncoeff, nfilter):
    return 2 * ncoeff - nfilter + 2
