"""The two process-wide appliers every caller of the path shares (apex/multi_tensor_apply/__init__.py:3-4 in the
reference): `multi_tensor_applier` passes the element-wise ops a chunk size of 8192 elements (ignored by this library's
kernels, which choose their own tile), `multi_tensor_applier_l2norm` passes the norm ops 65536 -- there the chunk is
the granularity of the overflow flag, so it is honoured."""
from .multi_tensor_apply import MultiTensorApply

ELEMENTWISE_CHUNK = 256 * 32
L2NORM_CHUNK = 2048 * 32

multi_tensor_applier = MultiTensorApply(ELEMENTWISE_CHUNK)
multi_tensor_applier_l2norm = MultiTensorApply(L2NORM_CHUNK)

__all__ = ["MultiTensorApply", "multi_tensor_applier", "multi_tensor_applier_l2norm", "ELEMENTWISE_CHUNK", "L2NORM_CHUNK"]
