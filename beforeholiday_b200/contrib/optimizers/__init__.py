from .distributed_fused_adam import DistributedFusedAdam, FlatLayout, Fragment
from .distributed_fused_lamb import DistributedFusedLAMB

__all__ = ["DistributedFusedAdam", "DistributedFusedLAMB", "FlatLayout", "Fragment"]
