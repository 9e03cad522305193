"""FusedAdam -- drop-in for apex.optimizers.FusedAdam (/root/reference/apex/optimizers/fused_adam.py:4-193).

Same constructor, same ``state`` keys (``exp_avg``, ``exp_avg_sq``), same group key ``step``, same call into
``amp_C.multi_tensor_adam(chunk_size, noop_flag, [g, p, m, v], lr, beta1, beta2, eps, step, mode, bias_correction,
weight_decay)`` -- one call per dtype bucket per param group.  On B200 each call is ONE persistent-grid launch over a
cached device-resident tensor table instead of ~ntensors/36 chunk-packed launches.

Difference from the reference, on purpose: bf16 parameters get their own bucket.  The reference appends them to the
fp16 bucket (fused_adam.py:135 makes its bf16 branch dead) and the kernel then reinterprets bf16 bits as fp16 when a
group mixes both.
"""
from __future__ import annotations

import torch

from ..multi_tensor_apply import multi_tensor_applier


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0., amsgrad=False, set_grad_none=True):
        if amsgrad:
            raise RuntimeError('FusedAdam does not support the AMSGrad variant.')
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.set_grad_none = set_grad_none
        if not multi_tensor_applier.available:
            raise RuntimeError('beforeholiday_b200.optimizers.FusedAdam requires the CUDA extension (no fallback)')
        from .. import load_amp_C

        device = self.param_groups[0]["params"][0].device
        # Skip buffer (the Adam kernel ignores it, as the reference's does)
        self._dummy_overflow_buf = torch.zeros(1, dtype=torch.int, device=device if device.type == "cuda" else "cuda")
        self.multi_tensor_adam = load_amp_C().multi_tensor_adam

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group['params']:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None):
        if any(p is not None for p in [grads, output_params, scale, grad_norms]):
            raise RuntimeError('FusedAdam has been updated.  Simply initialize it identically to torch.optim.Adam, '
                               'and call step() with no arguments.')
        loss = None
        if closure is not None:
            loss = closure()

        for group in self.param_groups:
            bias_correction = 1 if group['bias_correction'] else 0
            beta1, beta2 = group['betas']
            # one step counter per group, as the reference (fused_adam.py:111-114)
            group['step'] = group.get('step', 0) + 1

            buckets = {torch.float16: ([], [], [], []), torch.bfloat16: ([], [], [], []),
                       torch.float32: ([], [], [], [])}
            for p in group['params']:
                if p.grad is None:
                    continue
                if p.grad.data.is_sparse:
                    raise RuntimeError('FusedAdam does not support sparse gradients, please consider SparseAdam instead')
                state = self.state[p]
                if len(state) == 0:
                    state['exp_avg'] = torch.zeros_like(p.data)
                    state['exp_avg_sq'] = torch.zeros_like(p.data)
                if p.dtype not in buckets:
                    raise RuntimeError('FusedAdam only support fp16, bfloat16 and fp32.')
                g, w, m, v = buckets[p.dtype]
                g.append(p.grad.data)
                w.append(p.data)
                m.append(state['exp_avg'])
                v.append(state['exp_avg_sq'])

            for g, w, m, v in buckets.values():
                if g:
                    multi_tensor_applier(self.multi_tensor_adam, self._dummy_overflow_buf, [g, w, m, v],
                                         group['lr'], beta1, beta2, group['eps'], group['step'], self.adam_w_mode,
                                         bias_correction, group['weight_decay'])
        return loss
