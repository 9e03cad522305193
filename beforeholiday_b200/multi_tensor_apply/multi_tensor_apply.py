"""MultiTensorApply -- the operator-call boundary of the hot path.

Same contract as /root/reference/apex/multi_tensor_apply/multi_tensor_apply.py:3-30: an applier object is called as
``applier(op, noop_flag, tensor_lists, *args)`` and invokes ``op(chunk_size, noop_flag, tensor_lists, *args)``; the class
attribute ``available`` tells callers (the fused optimizers, amp's scaler) whether the CUDA extension can be used, and
``check_avail()`` raises with the original import error when it cannot.  There is no Python fallback for the ops.
"""
from __future__ import annotations

from typing import Optional

_UNAVAILABLE = ("MultiTensorApply cannot be used: beforeholiday_b200.amp_C (the CUDA extension this package is built "
                "around) did not import.  Build it in-tree with `python -m beforeholiday_b200._build`.  Import error: ")


def _probe_extension() -> Optional[BaseException]:
    """None when the extension loads, else the ImportError that stopped it."""
    try:
        from .. import load_amp_C

        load_amp_C()
    except ImportError as err:
        return err
    return None


class MultiTensorApply:
    # class-level, as in the reference: every applier shares one verdict on the extension
    available = False
    warned = False
    import_err: Optional[BaseException] = None

    def __init__(self, chunk_size: int):
        self.chunk_size = chunk_size
        failure = _probe_extension()
        MultiTensorApply.import_err = failure
        MultiTensorApply.available = failure is None

    def check_avail(self) -> None:
        if not MultiTensorApply.available:
            raise RuntimeError(_UNAVAILABLE + repr(MultiTensorApply.import_err))

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        self.check_avail()
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
