from .distributed_fused_adam import DistributedFusedAdam, FlatLayout, Fragment

__all__ = ["DistributedFusedAdam", "FlatLayout", "Fragment"]
