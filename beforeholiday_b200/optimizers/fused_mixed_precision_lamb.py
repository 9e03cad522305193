"""FusedMixedPrecisionLamb -- drop-in for apex.optimizers.FusedMixedPrecisionLamb
(/root/reference/apex/optimizers/fused_mixed_precision_lamb.py:8-256): the reference's capturable LAMB.

``lr`` and ``step`` are device tensors in each param group; the loss-scale inverse and the found-inf flag come from a
``torch.amp.GradScaler`` (or are 1 / 0); fp32 master copies are kept for parameters of ``reduced_precision_dtype``.
One ``amp_C.multi_tensor_l2norm_mp`` over all gradients gives the global norm, then one
``amp_C.multi_tensor_lamb_mp(chunk, noop, [g, p(32), m, v(, p_model)], lr, b1, b2, eps, step, bias_correction, wd,
grad_averaging, mode, grad_norm, max_grad_norm, use_nvlamb, found_inf, inv_scale)`` per group.  Nothing is read back to
the host, so the step can be captured in a CUDA graph after one eager warm-up step.
"""
from __future__ import annotations

import torch

from ..multi_tensor_apply import multi_tensor_applier, multi_tensor_applier_l2norm


class FusedMixedPrecisionLamb(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, step=0, bias_correction=True, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01,
                 amsgrad=False, adam_w_mode=True, grad_averaging=True, max_grad_norm=1.0, use_nvlamb=False,
                 reduced_precision_dtype=None):
        if amsgrad:
            raise RuntimeError('FusedLAMB does not support the AMSGrad variant.')
        defaults = dict(lr=torch.tensor(lr, dtype=torch.float32), step=torch.tensor([step], dtype=torch.int),
                        bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        grad_averaging=grad_averaging, max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        device = self.param_groups[0]['params'][0].device
        for group in self.param_groups:
            for item in ('lr', 'step'):
                group[item] = group[item].to(device=device)
        if not (multi_tensor_applier.available and multi_tensor_applier_l2norm.available):
            raise RuntimeError('beforeholiday_b200.optimizers.FusedMixedPrecisionLamb requires the CUDA extension')
        from .. import load_amp_C

        amp_C = load_amp_C()
        self.multi_tensor_l2norm = amp_C.multi_tensor_l2norm_mp
        self.multi_tensor_lamb = amp_C.multi_tensor_lamb_mp
        self._dummy_overflow_buf = torch.tensor([0], dtype=torch.int, device=device)
        self.reduced_precision_dtype = reduced_precision_dtype
        self.param_groups_full_precision = []
        self._step_supports_amp_scaling = True
        self.adam_w_mode = 1 if adam_w_mode else 0
        self.use_nvlamb = use_nvlamb

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        for name, default in self.defaults.items():
            if isinstance(default, torch.Tensor):
                self.param_groups[-1][name] = default.clone()

    def _setup_full_precision_params(self):
        for pg in self.param_groups:
            self.param_groups_full_precision.append({'params': [
                p.clone().detach().to(dtype=torch.float32)
                if (self.reduced_precision_dtype is not None and p.dtype == self.reduced_precision_dtype) else None
                for p in pg['params']]})

    @torch.no_grad()
    def step(self, closure=None, grad_scaler=None):
        loss = None
        if closure is not None:
            loss = closure()
        if not self.param_groups_full_precision:
            self._setup_full_precision_params()

        grad_list = []
        for group in self.param_groups:
            for p in group['params']:
                assert group['params'][0].dtype == p.dtype, \
                    "Error: Parameters are not of the identical type: {} != {}".format(group['params'][0].dtype, p.dtype)
                if p.grad is not None:
                    grad_list.append(p.grad)

        device = self.param_groups[0]["params"][0].device
        found_inf = (grad_scaler._check_inf_per_device(self)[device] if grad_scaler is not None
                     else torch.zeros((1,), device=device))
        self._dummy_overflow_buf.copy_(found_inf)
        if grad_scaler:
            scale = grad_scaler._get_scale_async()
            inv_scale = scale.double().reciprocal().float()
        else:
            scale = torch.ones((1,), device=device)
            inv_scale = torch.ones((1,), device=device)
        max_grad_norm = self.defaults['max_grad_norm'] * scale
        grad_norm = multi_tensor_applier_l2norm(self.multi_tensor_l2norm, self._dummy_overflow_buf, [grad_list], False)[0]

        for group, group_full in zip(self.param_groups, self.param_groups_full_precision):
            bias_correction = 1 if group['bias_correction'] else 0
            beta1, beta2 = group['betas']
            grad_averaging = 1 if group['grad_averaging'] else 0
            group['step'] += (self._dummy_overflow_buf != 1).to(torch.int)   # stays on the device
            lists = [[], [], [], []] + ([[]] if self.reduced_precision_dtype is not None else [])
            for p, p_full in zip(group['params'], group_full['params']):
                if p.grad is None:
                    continue
                assert not p.grad.is_sparse
                state = self.state[p]
                if len(state) == 0:
                    dtype = torch.float32 if (self.reduced_precision_dtype is not None
                                              and p.dtype == self.reduced_precision_dtype) else p.dtype
                    state['exp_avg'] = torch.zeros_like(p.data, dtype=dtype)
                    state['exp_avg_sq'] = torch.zeros_like(p.data, dtype=dtype)
                lists[0].append(p.grad.data)
                lists[2].append(state['exp_avg'])
                lists[3].append(state['exp_avg_sq'])
                if self.reduced_precision_dtype is not None:
                    lists[1].append(p_full.data)
                    lists[4].append(p.data)
                else:
                    lists[1].append(p.data)
            if lists[0]:
                multi_tensor_applier(self.multi_tensor_lamb, self._dummy_overflow_buf, lists, group['lr'], beta1, beta2,
                                     group['eps'], group['step'], bias_correction, group['weight_decay'], grad_averaging,
                                     self.adam_w_mode, grad_norm, max_grad_norm, self.use_nvlamb, found_inf, inv_scale)
        return loss
