from .fused_adagrad import FusedAdagrad
from .fused_adam import FusedAdam
from .fused_lamb import FusedLAMB
from .fused_lars import FusedLARS
from .fused_mixed_precision_lamb import FusedMixedPrecisionLamb
from .fused_novograd import FusedNovoGrad
from .fused_sgd import FusedSGD

__all__ = ["FusedAdam", "FusedLAMB", "FusedSGD", "FusedAdagrad", "FusedNovoGrad", "FusedLARS", "FusedMixedPrecisionLamb"]
