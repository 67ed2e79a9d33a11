"""pfb_b200 -- B200-native minor-cycle hot path of pfb-imaging.

Module layout mirrors the reference so imports can be re-pointed one to one:
    pfb.operators.psf      -> pfb_b200.operators.psf
    pfb.operators.hessian  -> pfb_b200.operators.hessian
    pfb.operators.psi      -> pfb_b200.operators.psi
    pfb.prox.prox_21m      -> pfb_b200.prox.prox_21m
    pfb.prox.prox_21       -> pfb_b200.prox.prox_21
    pfb.opt.pcg            -> pfb_b200.opt.pcg
    pfb.opt.primal_dual    -> pfb_b200.opt.primal_dual
    pfb.opt.power_method   -> pfb_b200.opt.power_method
    pfb.utils.misc         -> pfb_b200.utils.misc   (norm_diff, l1reweight_func)
"""
__version__ = "0.1.0"
