"""FusedLARS -- drop-in for apex.optimizers.FusedLARS (/root/reference/apex/optimizers/fused_lars.py).

Per group and per dtype set: per-tensor ||w|| and ||g|| (two ``amp_C.multi_tensor_l2norm(..., per_tensor=True)`` calls,
results stay on the device), then ``amp_C.multi_tensor_lars`` computes the layer-wise trust ratio on the device and
applies SGD-with-momentum.  Groups flagged ``is_skipped`` (biases / norms) use the plain learning rate."""
from __future__ import annotations

import torch
from torch.optim.optimizer import required

from ..multi_tensor_apply import multi_tensor_applier
from .fused_sgd import _grad_in_param_layout


class FusedLARS(torch.optim.Optimizer):
    def __init__(self, params, lr=required, momentum=0, dampening=0, weight_decay=0, trust_coefficient=0.001, eps=0.0,
                 nesterov=False, wd_after_momentum=False, materialize_master_grads=True, set_grad_none=False):
        if lr is not required and lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if momentum < 0.0:
            raise ValueError("Invalid momentum value: {}".format(momentum))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov,
                        trust_coefficient=trust_coefficient, eps=eps, is_skipped=False)
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, defaults)
        self.wd_after_momentum = wd_after_momentum
        self.materialize_master_grads = materialize_master_grads
        self.most_recent_scale = 1.0
        self.scale_set_by_backward = False
        self.set_grad_none = set_grad_none
        self.trust_coefficient = trust_coefficient
        self.eps = eps
        if not multi_tensor_applier.available:
            raise RuntimeError('beforeholiday_b200.optimizers.FusedLARS requires the CUDA extension (no fallback)')
        from .. import load_amp_C

        amp_C = load_amp_C()
        self._dummy_overflow_buf = torch.zeros(1, dtype=torch.int, device=self.param_groups[0]["params"][0].device)
        self.multi_tensor_l2norm = amp_C.multi_tensor_l2norm
        self.multi_tensor_lars = amp_C.multi_tensor_lars

    def __setstate__(self, state):
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault('nesterov', False)

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group['params']:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def get_momentums(self, params):
        momentums, first_run = [], True
        for p in params:
            st = self.state[p]
            if 'momentum_buffer' not in st:
                first_run = True
                st['momentum_buffer'] = torch.zeros_like(p.data)
            else:
                first_run = False
            momentums.append(st['momentum_buffer'])
        return momentums, first_run

    def _norms(self, tensors):
        return multi_tensor_applier(self.multi_tensor_l2norm, self._dummy_overflow_buf, [tensors], True)[1]

    def step(self, closure=None):
        loss = None
        if closure is not None:
            loss = closure()
        for group in self.param_groups:
            for dt in (torch.float16, torch.float32):
                ps = [p for p in group['params'] if p.dtype == dt and p.grad is not None]
                if not ps:
                    continue
                gs = [_grad_in_param_layout(p) for p in ps]
                moms, first_run = self.get_momentums(ps)
                w_norms = self._norms([p.data for p in ps])
                g_norms = self._norms([g.data for g in gs])
                multi_tensor_applier(self.multi_tensor_lars, self._dummy_overflow_buf, [gs, ps, moms], g_norms, w_norms,
                                     group['lr'], group['trust_coefficient'], self.eps, group['weight_decay'],
                                     group['momentum'], group['dampening'], group['nesterov'], first_run,
                                     self.wd_after_momentum, 1.0 / self.most_recent_scale, group['is_skipped'])
        self.most_recent_scale = 1.0
        self.scale_set_by_backward = False
        return loss
