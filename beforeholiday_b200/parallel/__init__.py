from .bucket_allreduce import (FlatBucketAllreduce, allreduce_bucket, apply_flat_dist_call, flat_dist_call, flatten,
                               split_by_type, unflatten)

__all__ = ["FlatBucketAllreduce", "allreduce_bucket", "apply_flat_dist_call", "flat_dist_call", "flatten",
           "split_by_type", "unflatten"]
