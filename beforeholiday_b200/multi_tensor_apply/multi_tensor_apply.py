"""MultiTensorApply -- the operator-call boundary of the hot path.

Mirror of /root/reference/apex/multi_tensor_apply/multi_tensor_apply.py:3-30: a callable that prepends
``chunk_size`` and forwards ``(noop_flag, tensor_lists, *args)`` to an ``amp_C`` op.  ``available`` says whether the
extension imported; there is no Python fallback for the ops themselves.
"""
from __future__ import annotations


class MultiTensorApply(object):
    available = False
    warned = False
    import_err = None

    def __init__(self, chunk_size):
        try:
            from .. import load_amp_C

            load_amp_C()
            MultiTensorApply.available = True
            self.chunk_size = chunk_size
        except ImportError as err:
            MultiTensorApply.available = False
            MultiTensorApply.import_err = err

    def check_avail(self):
        if not MultiTensorApply.available:
            raise RuntimeError(
                "Attempted to call MultiTensorApply method, but MultiTensorApply is not available: the "
                "beforeholiday_b200.amp_C CUDA extension failed to import.  Original import error message:",
                MultiTensorApply.import_err)

    def __call__(self, op, noop_flag_buffer, tensor_lists, *args):
        self.check_avail()
        return op(self.chunk_size, noop_flag_buffer, tensor_lists, *args)
