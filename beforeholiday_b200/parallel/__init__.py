from .bucket_allreduce import FlatBucketAllreduce, allreduce_bucket

__all__ = ["FlatBucketAllreduce", "allreduce_bucket"]
