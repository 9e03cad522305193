"""Same two global appliers as /root/reference/apex/multi_tensor_apply/__init__.py:3-4."""
from .multi_tensor_apply import MultiTensorApply

multi_tensor_applier = MultiTensorApply(256 * 32)
multi_tensor_applier_l2norm = MultiTensorApply(2048 * 32)

__all__ = ["MultiTensorApply", "multi_tensor_applier", "multi_tensor_applier_l2norm"]
