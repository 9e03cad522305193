"""Drop-in optimizer classes of the multi-tensor-apply path (the names apex.optimizers exports,
/root/reference/apex/optimizers/__init__.py), each a thin Python driver over one or two amp_C ops:

    FusedAdam   FusedLAMB   FusedSGD                       SURVEY.md section 8, row a12
    FusedMixedPrecisionLamb                                row f.1 (device-side lr / step / scale; graph-capturable)
    FusedAdagrad   FusedNovoGrad   FusedLARS               row f.3

`_fast_path` holds the steady-state route they share (cached per-group lists, gradients read by the extension).
"""
from . import fused_adagrad, fused_adam, fused_lamb, fused_lars, fused_mixed_precision_lamb, fused_novograd, fused_sgd

FusedAdam = fused_adam.FusedAdam
FusedLAMB = fused_lamb.FusedLAMB
FusedSGD = fused_sgd.FusedSGD
FusedMixedPrecisionLamb = fused_mixed_precision_lamb.FusedMixedPrecisionLamb
FusedAdagrad = fused_adagrad.FusedAdagrad
FusedNovoGrad = fused_novograd.FusedNovoGrad
FusedLARS = fused_lars.FusedLARS

__all__ = ["FusedAdam", "FusedLAMB", "FusedSGD", "FusedMixedPrecisionLamb", "FusedAdagrad", "FusedNovoGrad", "FusedLARS"]
