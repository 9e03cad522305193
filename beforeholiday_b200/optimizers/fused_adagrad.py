"""FusedAdagrad -- drop-in for apex.optimizers.FusedAdagrad (/root/reference/apex/optimizers/fused_adagrad.py).
State key ``sum``; one ``amp_C.multi_tensor_adagrad(chunk, noop, [g, p, h], lr, eps, mode, weight_decay)`` call per
dtype bucket per group (16-bit and fp32 buckets; bf16 gets its own bucket here, see fused_adam.py)."""
from __future__ import annotations

import torch

from ..multi_tensor_apply import multi_tensor_applier


class FusedAdagrad(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-2, eps=1e-10, weight_decay=0., set_grad_none=True, adagrad_w_mode=False):
        defaults = dict(lr=lr, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self.adagrad_w_mode = 1 if adagrad_w_mode else 0
        self.set_grad_none = set_grad_none
        if not multi_tensor_applier.available:
            raise RuntimeError('beforeholiday_b200.optimizers.FusedAdagrad requires the CUDA extension (no fallback)')
        from .. import load_amp_C

        self._dummy_overflow_buf = torch.zeros(1, dtype=torch.int, device=self.param_groups[0]["params"][0].device)
        self.multi_tensor_adagrad = load_amp_C().multi_tensor_adagrad

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group['params']:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def step(self, closure=None):
        loss = None
        if closure is not None:
            loss = closure()
        for group in self.param_groups:
            buckets = {torch.float16: ([], [], []), torch.bfloat16: ([], [], []), torch.float32: ([], [], [])}
            for p in group['params']:
                if p.grad is None:
                    continue
                if p.grad.data.is_sparse:
                    raise RuntimeError('FusedAdagrad does not support sparse gradients')
                state = self.state[p]
                if len(state) == 0:
                    state['sum'] = torch.zeros_like(p.data)
                if p.dtype not in buckets:
                    raise RuntimeError('FusedAdagrad only support fp16, bfloat16 and fp32.')
                g, w, h = buckets[p.dtype]
                g.append(p.grad.data)
                w.append(p.data)
                h.append(state['sum'])
            for g, w, h in buckets.values():
                if g:
                    multi_tensor_applier(self.multi_tensor_adagrad, self._dummy_overflow_buf, [g, w, h], group['lr'],
                                         group['eps'], self.adagrad_w_mode, group['weight_decay'])
        return loss
