"""ctypes binding of libpfb_b200.so (the C-ABI declared in include/pfb_b200.h).

There is no CPU fallback: if the shared library has not been built
(``python -c 'import __graft_entry__ as g; g.build()'``) every operator raises.
"""
import ctypes
import os
from ctypes import c_int, c_longlong, c_double, c_void_p, c_char_p, POINTER

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "csrc", "libpfb_b200.so")

F32, F64 = 0, 1
BAND_SUM, BAND_L2 = 0, 1
PCG_ALPHA, PCG_BETA, PCG_SHRINK, PCG_COMMIT = 0, 1, 2, 3

STATUS = {1: "bad argument", 2: "odd image size", 3: "nlevel too large",
          4: "unsupported FFT size", 5: "CUDA error", 6: "allocation failed",
          7: "too many bands"}


class PfbError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"pfb_b200 status {status} ({STATUS.get(status, '?')}): {msg}")
        self.status = status


_P = c_void_p
_SIGS = {
    "pfb_version": (c_int, []),
    "pfb_last_error": (c_char_p, []),
    "pfb_launch_count": (c_longlong, []),
    "pfb_profile_enable": (c_int, [c_int]),
    "pfb_profile_report": (c_int, [c_char_p, c_longlong]),
    "pfb_prox21": (c_int, [_P, _P, _P, c_int, c_longlong, c_double, c_double, c_int, c_int, _P]),
    "pfb_dual_update": (c_int, [_P, _P, _P, _P, c_int, c_longlong, c_double, c_double, c_int, c_int, _P]),
    "pfb_dual_update_rec": (c_int, [_P, _P, _P, _P, _P, c_int, c_longlong, c_double, c_double, c_int, _P]),
    "pfb_dual_factor": (c_int, [_P, _P, _P, _P, c_longlong, c_double, c_double, c_int, _P]),
    "pfb_psi_dot_nat_dual": (c_int, [_P, _P, _P, _P, c_double, _P, _P]),
    "pfb_dual_partial": (c_int, [_P, _P, _P, c_int, c_longlong, c_longlong, c_double, c_int, c_int, _P]),
    "pfb_dual_apply": (c_int, [_P, _P, _P, _P, _P, c_int, c_longlong, c_longlong, c_double, c_double, c_int, c_int, _P]),
    "pfb_axpby": (c_int, [_P, c_double, _P, c_double, _P, c_longlong, c_int, _P]),
    "pfb_vdot": (c_int, [_P, _P, _P, _P, c_longlong, c_int, _P, _P]),
    "pfb_norm_diff": (c_int, [_P, _P, c_longlong, c_int, _P, _P]),
    "pfb_pd_primal": (c_int, [_P, _P, _P, _P, _P, c_double, c_int, c_int, c_longlong, c_int, _P, _P]),
    "pfb_pd_primal_lin": (c_int, [_P, _P, _P, _P, _P, _P, c_double, c_int, c_int, c_longlong, c_int, _P, _P]),
    "pfb_pcg_update": (c_int, [_P, _P, _P, _P, _P, _P, c_double, c_double, c_longlong, c_int, _P, _P]),
    "pfb_psi_plan_create": (c_int, [POINTER(c_void_p), c_int, c_int, c_int, c_int, POINTER(c_int), c_int, c_int]),
    "pfb_psi_plan_destroy": (c_int, [_P]),
    "pfb_filter_bank": (c_int, [c_int, POINTER(ctypes.c_double)]),
    "pfb_hess_plan_can_fuse_primal": (c_int, [_P]),
    "pfb_hess_plan_kernels": (c_int, [_P]),
    "pfb_hess_apply_primal": (c_int, [_P, _P, _P, _P, _P, c_double, c_int, _P, _P, _P]),
    "pfb_hess_apply_primal_begin": (c_int, [_P, _P, _P]),
    "pfb_hess_apply_primal_end": (c_int, [_P, _P, _P, _P, _P, c_double, c_int, _P, _P, _P]),
    "pfb_psi_plan_query": (c_int, [_P, POINTER(c_int), POINTER(c_int)]),
    "pfb_psi_plan_tables": (c_int, [_P, c_int, POINTER(c_int), POINTER(c_int), POINTER(c_int)]),
    "pfb_psi_dot": (c_int, [_P, _P, _P, _P]),
    "pfb_psi_hdot": (c_int, [_P, _P, _P, _P, _P]),
    "pfb_psi_nat_stride": (c_int, [_P, POINTER(c_longlong)]),
    "pfb_psi_to_nat": (c_int, [_P, c_int, _P, _P, _P]),
    "pfb_psi_from_nat": (c_int, [_P, c_int, _P, _P, _P]),
    "pfb_psi_dot_nat": (c_int, [_P, _P, _P, _P]),
    "pfb_psi_hdot_nat": (c_int, [_P, _P, _P, _P, _P]),
    "pfb_hess_plan_create": (c_int, [POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int, c_int]),
    "pfb_hess_plan_destroy": (c_int, [_P]),
    "pfb_hess_plan_set_psfhat": (c_int, [_P, _P, _P]),
    "pfb_hess_plan_set_dynamic": (c_int, [_P, c_int]),
    "pfb_hess_apply": (c_int, [_P, _P, _P, c_int, c_double, c_int, c_double, _P, _P]),
    "pfb_pcg_direction": (c_int, [_P, _P, c_double, c_double, c_longlong, c_int, _P, _P]),
    "pfb_pcg_seg_dot": (c_int, [_P, _P, c_int, c_longlong, c_int, _P, c_int, _P]),
    "pfb_pcg_seg_scalar": (c_int, [_P, c_int, c_int, _P]),
    "pfb_pcg_seg_update": (c_int, [_P, _P, _P, _P, _P, _P, c_double, c_int, c_longlong, c_int, _P, _P]),
    "pfb_pcg_seg_direction": (c_int, [_P, _P, _P, c_double, c_int, c_longlong, c_int, _P, _P]),
    "pfb_gauss2d": (c_int, [_P, _P, c_longlong, c_double, c_double, c_double, c_double, c_double, _P, c_int, _P]),
    "pfb_spec_ratio": (c_int, [_P, _P, _P, c_longlong, c_longlong, c_int, _P]),
    "pfb_pd_run": (c_int, [_P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, c_double, c_double,
                           c_double, c_double, c_double, c_double, c_int, c_int, c_int, _P, _P, _P]),
    "pfb_power_method_run": (c_int, [_P, _P, c_int, c_int, c_int, c_double, c_int, c_int, _P, _P, _P]),
    "pfb_pcg_run": (c_int, [_P, _P, c_double, c_int, c_double, _P, _P, c_double, c_int, c_int, c_int,
                            c_double, c_int, c_int, c_int, c_int, _P, _P, _P]),
    "pfb_mfs_peak": (c_int, [_P, c_int, c_longlong, _P, _P, c_int, _P]),
    "pfb_active_set": (c_int, [_P, c_longlong, c_double, _P, c_longlong, _P, _P, c_int, _P]),
    "pfb_clark_subminor": (c_int, [_P, _P, _P, c_longlong, _P, _P, c_int, c_int, c_int, c_int, c_int,
                                   c_double, c_double, c_int, _P, _P, c_int, _P]),
    "pfb_hogbom_start": (c_int, [_P, _P, _P, c_int, c_longlong, c_double, c_double, c_double, c_int, _P, c_int, _P]),
    "pfb_hogbom_iterate": (c_int, [_P, _P, _P, _P, c_int, c_int, c_int, c_int, c_int, c_double, c_int, _P, c_int, _P]),
    "pfb_accum_mul": (c_int, [_P, _P, _P, c_double, c_longlong, c_int, _P]),
    "pfb_seg_stats": (c_int, [_P, c_int, c_longlong, _P, c_int, _P]),
    "pfb_seg_stats_static": (c_int, [_P, c_int, c_longlong, _P, c_int, _P]),
    "pfb_l1reweight": (c_int, [_P, _P, c_int, c_longlong, c_double, c_double, _P, c_int, _P]),
    "pfb_psfhat_from_psf": (c_int, [_P, c_int, c_int, c_int, c_int, _P, _P]),
}


def bind(cdll):
    """Attach prototypes for every symbol of include/pfb_b200.h."""
    for name, (res, args) in _SIGS.items():
        fn = getattr(cdll, name)      # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    return cdll


_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} not found: build the CUDA library first "
                "(__graft_entry__.build()); pfb_b200 has no CPU fallback")
        _LIB = bind(ctypes.CDLL(SO_PATH))
    return _LIB


def check(status):
    if status != 0:
        raise PfbError(status, lib().pfb_last_error().decode())
