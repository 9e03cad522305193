"""amp's dynamic loss scaler -- drop-in for apex.amp.scaler.LossScaler (/root/reference/apex/amp/scaler.py:42-226).

Policy (host side, unchanged): start at min(max_loss_scale, 2**16); on overflow halve (floored at min_loss_scale) and
skip the step; after `scale_window` (2000) clean steps double (capped at max_loss_scale = 2**24).

Mechanics (device side): ``unscale`` is one ``amp_C.multi_tensor_scale([model_grads, master_grads], 1/scale)`` launch
that also raises ``_overflow_buf`` when any incoming gradient element is inf/nan; ``unscale_with_stashed`` is one
``amp_C.multi_tensor_axpby`` launch that accumulates into stashed master grads, checking only the incoming grads
(arg_to_check = 0).  ``update_scale`` performs the single device->host read of the flag per iteration.

There is no Python fallback (the reference's scale_check_overflow_python path, scaler.py:6-40, serves as a CPU oracle in
this repository's tests, not as a product path).
"""
from __future__ import annotations

import torch

from ..multi_tensor_apply import multi_tensor_applier


class LossScaler(object):
    warned_no_fused_kernel = False
    warned_unscaling_non_fp32_grad = False
    has_fused_kernel = False

    def __init__(self, loss_scale, init_scale=2.**16, scale_factor=2., scale_window=2000, min_loss_scale=None,
                 max_loss_scale=2.**24):
        if loss_scale == "dynamic":
            self.dynamic = True
            self._loss_scale = min(max_loss_scale, init_scale)
        else:
            self.dynamic = False
            self._loss_scale = loss_scale
        self._max_loss_scale = max_loss_scale
        self._min_loss_scale = min_loss_scale
        self._scale_seq_len = scale_window
        self._unskipped = 0
        self._has_overflow = False
        self._overflow_buf = torch.zeros(1, dtype=torch.int, device="cuda")
        if not multi_tensor_applier.available:
            raise RuntimeError("beforeholiday_b200.amp.LossScaler requires the CUDA extension (no Python fallback): "
                               + repr(multi_tensor_applier.import_err))
        from .. import load_amp_C

        amp_C = load_amp_C()
        LossScaler.has_fused_kernel = True
        LossScaler.multi_tensor_scale_cuda = amp_C.multi_tensor_scale
        LossScaler.multi_tensor_axpby_cuda = amp_C.multi_tensor_axpby

    def loss_scale(self):
        return self._loss_scale

    # unused_scale keeps the reference's positional signature (scaler.py:103)
    def unscale(self, model_grads, master_grads, unused_scale, models_are_masters=False, scale_override=None):
        if self._has_overflow:
            return
        scale = self._loss_scale if scale_override is None else scale_override
        if scale == 1.0 and models_are_masters and not self.dynamic:
            return
        multi_tensor_applier(LossScaler.multi_tensor_scale_cuda, self._overflow_buf, [model_grads, master_grads],
                             1. / scale)

    def unscale_with_stashed(self, model_grads, stashed_master_grads, master_grads, scale_override=None):
        if self._has_overflow:
            return
        grads_have_scale, stashed_have_scale, out_scale = self._loss_scale, 1.0, 1.0
        if scale_override is not None:
            grads_have_scale, stashed_have_scale, out_scale = scale_override
        if not LossScaler.warned_unscaling_non_fp32_grad and master_grads[0].dtype == torch.float16:
            print("Warning:  unscaling grads that are not FP32. Unscaling non-fp32 grads may indicate an error. "
                  "When using Amp, you don't need to call .half() on your model.")
            LossScaler.warned_unscaling_non_fp32_grad = True
        multi_tensor_applier(LossScaler.multi_tensor_axpby_cuda, self._overflow_buf,
                             [model_grads, stashed_master_grads, master_grads],
                             out_scale / grads_have_scale,    # 1./scale
                             out_scale / stashed_have_scale,  # 1.0
                             0)  # check only arg 0, the incoming model grads, for infs

    def clear_overflow_state(self):
        self._has_overflow = False
        self._overflow_buf.zero_()

    # Separate so unscale() can be called more than once before updating.
    def update_scale(self):
        if self.dynamic and not self._has_overflow:
            self._has_overflow = self._overflow_buf.item()  # the one D2H sync per iteration

        if self._has_overflow and self.dynamic:
            should_skip = True
            if self._min_loss_scale:
                self._loss_scale = max(self._min_loss_scale, self._loss_scale / 2.)
            else:
                self._loss_scale = self._loss_scale / 2.
            self._unskipped = 0
        else:
            should_skip = False
            self._unskipped += 1

        if self._unskipped == self._scale_seq_len and self.dynamic:
            self._loss_scale = min(self._max_loss_scale, self._loss_scale * 2.)
            self._unskipped = 0
        return should_skip

    # amp.state_dict() payload of the reference (apex/amp/frontend.py:434-444)
    def state_dict(self):
        return {"loss_scale": self._loss_scale, "unskipped": self._unskipped}

    def load_state_dict(self, sd):
        self._loss_scale = sd["loss_scale"]
        self._unskipped = sd["unskipped"]
