"""FusedSGD -- drop-in for apex.optimizers.FusedSGD (/root/reference/apex/optimizers/fused_sgd.py:6-264).

Per param group up to two launch sets are built -- a 16-bit set and an fp32 set -- and each goes to
``amp_C.multi_tensor_sgd(chunk_size, noop_flag, [g, w, mom(, w_fp16)], wd, momentum, dampening, lr, nesterov,
first_run, wd_after_momentum, 1/most_recent_scale)``.  With amp master weights (``self._amp_stash`` present, set by
the amp front-end) the 16-bit set is ``[grads, fp32 masters, fp32 momenta, fp16 model params]`` so the kernel
un-scales the gradient, updates the master and writes the fp16 model copy in one pass (fused_sgd.py:165-212).
"""
from __future__ import annotations

import torch
from torch.optim.optimizer import required

from ..multi_tensor_apply import multi_tensor_applier
from ._fast_path import FastGroups, is_fallback


def _grad_in_param_layout(p, grad=None):
    """The op runs over physical memory order, so the gradient must share the parameter's memory format
    (fused_sgd.py:217-237)."""
    g = p.grad if grad is None else grad
    if p.is_contiguous(memory_format=torch.contiguous_format):
        return g
    if p.is_contiguous(memory_format=torch.channels_last):
        return g.to(memory_format=torch.channels_last)
    if p.is_contiguous(memory_format=torch.channels_last_3d):
        return g.to(memory_format=torch.channels_last_3d)
    raise AssertionError("Unsupported memory format. Supports only contiguous_format, channels_last, or channel_last_3d.")


class FusedSGD(torch.optim.Optimizer):
    def __init__(self, params, lr=required, momentum=0, dampening=0, weight_decay=0, nesterov=False,
                 wd_after_momentum=False, materialize_master_grads=True, set_grad_none=False):
        if lr is not required and lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if momentum < 0.0:
            raise ValueError("Invalid momentum value: {}".format(momentum))
        if weight_decay < 0.0:
            raise ValueError("Invalid weight_decay value: {}".format(weight_decay))
        defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov)
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, defaults)

        self.wd_after_momentum = wd_after_momentum
        self.materialize_master_grads = materialize_master_grads
        self.most_recent_scale = 1.0
        self.scale_set_by_backward = False
        self.set_grad_none = set_grad_none

        if not multi_tensor_applier.available:
            raise RuntimeError('beforeholiday_b200.optimizers.FusedSGD requires the CUDA extension (no fallback)')
        from .. import load_amp_C

        self._dummy_overflow_buf = torch.tensor([0], dtype=torch.int,
                                                device=self.param_groups[0]["params"][0].device)
        self.multi_tensor_sgd = load_amp_C().multi_tensor_sgd
        self._fast = FastGroups()

    def __setstate__(self, state):
        super().__setstate__(state)
        for group in self.param_groups:
            group.setdefault('nesterov', False)
        self._fast = FastGroups()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._fast.clear()       # the momentum buffers are new tensors

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        if hasattr(self, "_fast"):
            self._fast.clear()

    def _fast_step(self, gid, group):
        """Steady-state path (see _fast_path.py): cached parameter / momentum lists, gradients read by the extension.
        Returns the dtypes it has updated; the generic path handles the rest (normally: everything or nothing)."""
        def build(params):
            by_dt = [(dt, [p for p in params if p.dtype == dt]) for dt in (torch.float16, torch.float32)]
            if sum(len(sel) for _, sel in by_dt) != len(params):
                return None                       # other dtypes present: the generic path decides what to do with them
            sets = []
            for dt, sel in by_dt:
                if sel:
                    moms, first_run = self.get_momentums(sel)
                    sets.append({'dtype': dt, 'params': sel, 'moms': moms, 'first_run': first_run})
            return sets

        c = self._fast.get(gid, group, build)
        done = []
        if c is None:
            return done
        for st in c['sets']:
            try:
                self.multi_tensor_sgd(multi_tensor_applier.chunk_size, self._dummy_overflow_buf, [st['params'], st['moms']],
                                      group['weight_decay'], group['momentum'], group['dampening'], group['lr'],
                                      group['nesterov'], st['first_run'], self.wd_after_momentum,
                                      1.0 / self.most_recent_scale, grads_of=0)
            except RuntimeError as e:
                if not is_fallback(e):
                    raise
                return done
            st['first_run'] = False
            done.append(st['dtype'])
        return done

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group['params']:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    def get_momentums(self, params):
        """Momentum buffers for `params`; `first_run` is True when they were just created (the kernel then
        initialises momentum to the incoming gradient, like torch.optim.SGD; fused_sgd.py:121-136)."""
        momentums = []
        first_run = True
        for p in params:
            st = self.state[p]
            if 'momentum_buffer' not in st:
                first_run = True
                st['momentum_buffer'] = torch.zeros_like(p.data)
            else:
                first_run = False
            momentums.append(st['momentum_buffer'])
        return momentums, first_run

    def step(self, closure=None):
        loss = None
        if closure is not None:
            loss = closure()

        explicit_master_params = (hasattr(self, "_amp_stash") and hasattr(self._amp_stash, "fp32_from_fp16_groups"))

        for gid, group in enumerate(self.param_groups):
            first_runs = [True, True]
            if explicit_master_params:
                stash = self._amp_stash
                fp32_params = [p for p in stash.fp32_from_fp32_groups[gid] if p.grad is not None]
                fp32_grads = [p.grad for p in fp32_params]
                fp32_momentums, first_runs[1] = self.get_momentums(fp32_params)
                if self.materialize_master_grads:
                    masters = stash.fp32_from_fp16_groups[gid]
                    fp16_model_params = [p for i, p in enumerate(stash.fp16_groups[gid]) if masters[i].grad is not None]
                    fp32_from_fp16_params = [p for p in masters if p.grad is not None]
                    grads16 = [_grad_in_param_layout(p) for p in fp32_from_fp16_params]
                else:
                    fp16_model_params = [p for p in stash.fp16_groups[gid] if p.grad is not None]
                    grads16 = [_grad_in_param_layout(p) for p in fp16_model_params]
                    fp32_from_fp16_params = [p for i, p in enumerate(stash.fp32_from_fp16_groups[gid])
                                             if stash.fp16_groups[gid][i].grad is not None]
                fp32_from_fp16_momentums, first_runs[0] = self.get_momentums(fp32_from_fp16_params)
                launch_sets = [[grads16, fp32_from_fp16_params, fp32_from_fp16_momentums, fp16_model_params],
                               [fp32_grads, fp32_params, fp32_momentums]]
            else:
                done = self._fast_step(gid, group)
                want = lambda p, dt: p.dtype == dt and dt not in done and p.grad is not None
                fp16_params = [p for p in group['params'] if want(p, torch.float16)] if torch.float16 not in done else []
                fp16_grads = [_grad_in_param_layout(p) for p in fp16_params]
                fp16_momentums, first_runs[0] = self.get_momentums(fp16_params)
                fp32_params = [p for p in group['params'] if want(p, torch.float32)] if torch.float32 not in done else []
                fp32_grads = [_grad_in_param_layout(p) for p in fp32_params]
                fp32_momentums, first_runs[1] = self.get_momentums(fp32_params)
                launch_sets = [[fp16_grads, fp16_params, fp16_momentums], [fp32_grads, fp32_params, fp32_momentums]]

            for launch_set, first_run in zip(launch_sets, first_runs):
                assert len(launch_set[0]) == len(launch_set[1]) == len(launch_set[2])
                if launch_set[0]:
                    multi_tensor_applier(self.multi_tensor_sgd, self._dummy_overflow_buf, launch_set,
                                         group['weight_decay'], group['momentum'], group['dampening'], group['lr'],
                                         group['nesterov'], first_run, self.wd_after_momentum,
                                         1.0 / self.most_recent_scale)

        self.most_recent_scale = 1.0
        self.scale_set_by_backward = False
        return loss
