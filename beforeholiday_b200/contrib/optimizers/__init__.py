from .distributed_fused_adam import DistributedFusedAdam
from .distributed_fused_lamb import DistributedFusedLAMB
from .flat_sharded_adam import FlatLayout, FlatShardedAdam, Fragment

__all__ = ["DistributedFusedAdam", "DistributedFusedLAMB", "FlatLayout", "FlatShardedAdam", "Fragment"]
