"""FusedNovoGrad -- drop-in for apex.optimizers.FusedNovoGrad (/root/reference/apex/optimizers/fused_novograd.py).

Per-parameter first moment ``exp_avg``; the second moment is ONE scalar per tensor (a running gradient norm) kept per
group in ``group['exp_avg_sq'] = [norms_of_16bit_params, norms_of_fp32_params]`` (device float vectors).  Each step:
``amp_C.multi_tensor_novograd(chunk, noop, [g, p, m], exp_avg_sq, lr, b1, b2, eps, step, bias_correction, wd,
grad_averaging, moment_mode, norm_type)`` blends this step's per-tensor gradient norm into ``exp_avg_sq`` on the device
and applies the update."""
from __future__ import annotations

import torch

from ..multi_tensor_apply import multi_tensor_applier


class FusedNovoGrad(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.,
                 amsgrad=False, reg_inside_moment=False, grad_averaging=True, norm_type=2, init_zero=False,
                 set_grad_none=True):
        if amsgrad:
            raise RuntimeError('FusedNovoGrad does not support the AMSGrad variant.')
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        grad_averaging=grad_averaging, norm_type=norm_type, init_zero=init_zero)
        super().__init__(params, defaults)
        if not multi_tensor_applier.available:
            raise RuntimeError('beforeholiday_b200.optimizers.FusedNovoGrad requires the CUDA extension (no fallback)')
        from .. import load_amp_C

        self._dummy_overflow_buf = torch.zeros(1, dtype=torch.int, device=self.param_groups[0]["params"][0].device)
        self.multi_tensor_novograd = load_amp_C().multi_tensor_novograd
        self.moment_mode = 0 if reg_inside_moment else 1
        self.set_grad_none = set_grad_none

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group['params']:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        # the per-group norm vectors must live on the parameters' device (reference :100-105)
        for group in self.param_groups:
            if len(group['params']) > 0 and 'exp_avg_sq' in group:
                dev = group['params'][0].device
                group['exp_avg_sq'] = [t.to(dev) for t in group['exp_avg_sq']]

    def step(self, closure=None):
        loss = None
        if closure is not None:
            loss = closure()
        for group in self.param_groups:
            bias_correction = 1 if group['bias_correction'] else 0
            beta1, beta2 = group['betas']
            grad_averaging = 1 if group['grad_averaging'] else 0
            group['step'] = group.get('step', 0) + 1
            b16, b32 = ([], [], []), ([], [], [])
            for p in group['params']:
                if p.grad is None:
                    continue
                if p.grad.data.is_sparse:
                    raise RuntimeError('FusedNovoGrad does not support sparse gradients, please consider SparseAdam instead')
                state = self.state[p]
                if len(state) == 0:
                    state['exp_avg'] = torch.zeros_like(p.data)
                if p.dtype in (torch.float16, torch.bfloat16):
                    b = b16
                elif p.dtype == torch.float32:
                    b = b32
                else:
                    raise RuntimeError('FusedNovoGrad only support fp16, bfloat16 and fp32.')
                b[0].append(p.grad.data)
                b[1].append(p.data)
                b[2].append(state['exp_avg'])
            dev = self.param_groups[0]["params"][0].device
            if 'exp_avg_sq' not in group:
                if group['init_zero']:
                    group['exp_avg_sq'] = [torch.zeros(len(b16[0]), device=dev), torch.zeros(len(b32[0]), device=dev)]
                else:  # start from the first step's norm so that the first blend is a no-op (reference :155-166)
                    if group['norm_type'] == 0:
                        f = lambda g: g.float().abs().max()
                    elif group['norm_type'] == 2:
                        f = lambda g: g.float().pow(2).sum().sqrt()
                    else:
                        raise RuntimeError('FusedNovoGrad only support l2/inf norm now.')
                    mk = lambda gs: (torch.stack([f(g) for g in gs]).to(dev) if gs else torch.zeros(0, device=dev))
                    group['exp_avg_sq'] = [mk(b16[0]), mk(b32[0])]
            else:
                assert len(b16[0]) == group['exp_avg_sq'][0].numel()
                assert len(b32[0]) == group['exp_avg_sq'][1].numel()
            for (g, w, m), norms in zip((b16, b32), group['exp_avg_sq']):
                if g:
                    multi_tensor_applier(self.multi_tensor_novograd, self._dummy_overflow_buf, [g, w, m], norms,
                                         group['lr'], beta1, beta2, group['eps'], group['step'], bias_correction,
                                         group['weight_decay'], grad_averaging, self.moment_mode, group['norm_type'])
        return loss
