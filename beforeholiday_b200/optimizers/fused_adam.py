"""FusedAdam -- drop-in for apex.optimizers.FusedAdam (/root/reference/apex/optimizers/fused_adam.py:4-193).

Same constructor, same ``state`` keys (``exp_avg``, ``exp_avg_sq``), same group key ``step``, same call into
``amp_C.multi_tensor_adam(chunk_size, noop_flag, [g, p, m, v], lr, beta1, beta2, eps, step, mode, bias_correction,
weight_decay)`` -- one call per dtype bucket per param group.  On B200 each call is ONE persistent-grid launch over a
cached device-resident tensor table instead of ~ntensors/36 chunk-packed launches.

B200 extensions (off by default; upstream Apex has the same two switches, this fork does not):

  master_weights=True   16-bit parameters get an fp32 master copy (``state['master_param']``) and fp32 moments; the step
                        is ONE pass of ``amp_C.multi_tensor_fused_adam([p32, m, v, g16, p16], ...)`` = 28 B/param: 16-bit
                        gradient in (divided by ``inv_scale``'s reciprocal on the fly), fp32 p/m/v updated, 16-bit model
                        parameter written, skipped entirely on the device when ``found_inf`` is set.  The reference's amp
                        O2 route needs three passes and 40 B/param for the same result.
  capturable=True       ``lr`` and ``step`` live on the device and the kernel reads them there
                        (``amp_C.multi_tensor_fused_adam_capturable``), so ``step()`` can be captured in a CUDA graph.
  Both require decoupled weight decay (adam_w_mode=True) or weight_decay == 0: that is the arithmetic of the reference's
  mixed-precision kernel (DistAdamFunctor).

Difference from the reference, on purpose: bf16 parameters get their own bucket.  The reference appends them to the
fp16 bucket (fused_adam.py:135 makes its bf16 branch dead) and the kernel then reinterprets bf16 bits as fp16 when a
group mixes both.
"""
from __future__ import annotations

import torch

from ..multi_tensor_apply import multi_tensor_applier
from ._fast_path import FastGroups, is_fallback


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True,
                 weight_decay=0., amsgrad=False, set_grad_none=True, capturable=False, master_weights=False):
        if amsgrad:
            raise RuntimeError('FusedAdam does not support the AMSGrad variant.')
        self.capturable = bool(capturable)
        self.master_weights = bool(master_weights)
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
        amp_C = load_amp_C()
        self.multi_tensor_adam = amp_C.multi_tensor_adam
        self.multi_tensor_fused_adam = amp_C.multi_tensor_fused_adam
        self.multi_tensor_fused_adam_capturable = amp_C.multi_tensor_fused_adam_capturable
        self._hp_cache = {}
        self._fast = FastGroups()
        if self.capturable or self.master_weights:
            for group in self.param_groups:
                if not self.adam_w_mode and group['weight_decay'] != 0:
                    raise RuntimeError('FusedAdam(capturable / master_weights) needs adam_w_mode=True or weight_decay=0')
        if self.capturable:
            dev = self._dummy_overflow_buf.device
            for group in self.param_groups:
                group['lr'] = torch.as_tensor(group['lr'], dtype=torch.float32).reshape(1).to(dev)
                group['step'] = torch.zeros(1, dtype=torch.int32, device=dev)
            self._one = torch.ones(1, dtype=torch.float32, device=dev)

    def __setstate__(self, state):
        super().__setstate__(state)
        self._fast = FastGroups()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._fast.clear()       # exp_avg / exp_avg_sq are new tensors

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        if hasattr(self, "_fast"):
            self._fast.clear()

    def _fast_step(self, gid, group, bias_correction, beta1, beta2):
        """Steady-state path of the plain route (see _fast_path.py).  Returns the dtypes it has updated."""
        def build(params):
            dts = (torch.float16, torch.bfloat16, torch.float32)
            if any(p.dtype not in dts for p in params):
                return None
            sets = []
            for dt in dts:
                sel = [p for p in params if p.dtype == dt]
                if not sel:
                    continue
                for p in sel:
                    state = self.state[p]
                    if len(state) == 0:
                        state['exp_avg'] = torch.zeros_like(p.data)
                        state['exp_avg_sq'] = torch.zeros_like(p.data)
                sets.append({'dtype': dt, 'params': sel, 'm': [self.state[p]['exp_avg'] for p in sel],
                             'v': [self.state[p]['exp_avg_sq'] for p in sel]})
            return sets

        c = self._fast.get(gid, group, build)
        done = []
        if c is None:
            return done
        for st in c['sets']:
            try:
                self.multi_tensor_adam(multi_tensor_applier.chunk_size, self._dummy_overflow_buf,
                                       [st['params'], st['m'], st['v']], group['lr'], beta1, beta2, group['eps'],
                                       group['step'], self.adam_w_mode, bias_correction, group['weight_decay'], grads_of=0)
            except RuntimeError as e:
                if not is_fallback(e):
                    raise
                return done
            done.append(st['dtype'])
        return done

    def zero_grad(self, set_to_none=None):
        if self.set_grad_none if set_to_none is None else set_to_none:
            for group in self.param_groups:
                for p in group['params']:
                    p.grad = None
        else:
            super().zero_grad(set_to_none=False)

    # ---- mixed-precision / capturable route ---------------------------------------------------------------------------
    def _hyper(self, group, n, device):
        """Per-tensor hyper-parameter vectors of the mixed-precision kernel, cached per (group values, n)."""
        key = (id(group), n, group['betas'], group['eps'], group['weight_decay'], bool(group['bias_correction']))
        hp = self._hp_cache.get(key)
        if hp is None:
            f = lambda v, dt=torch.float32: torch.full((n,), v, dtype=dt, device=device)
            hp = (f(group['betas'][0]), f(group['betas'][1]), f(1 if group['bias_correction'] else 0, torch.int32),
                  f(group['eps']), f(group['weight_decay']))
            self._hp_cache = {k: v for k, v in self._hp_cache.items() if k[0] != id(group)}
            self._hp_cache[key] = hp
        return hp

    def _step_mixed(self, inv_scale, found_inf):
        flag = self._dummy_overflow_buf
        if found_inf is not None and not self.capturable:
            # Host-driven mode: the step counter lives on the host, so the skip decision must too -- one device->host
            # read per iteration, the same one amp's LossScaler.update_scale() performs (apex/amp/scaler.py:209).
            if bool(found_inf.reshape(-1)[0].item()):
                return
            flag.zero_()
        elif found_inf is not None:
            flag.copy_(found_inf.reshape(-1)[:1].to(torch.int32), non_blocking=True)
        else:
            flag.zero_()
        for group in self.param_groups:
            if self.capturable:
                group['step'] += (flag == 0).to(torch.int32)   # device-side: a skipped step does not count
            else:
                group['step'] = group.get('step', 0) + 1       # the host cannot see the flag without a sync
            by_dtype = {}
            for p in group['params']:
                if p.grad is None:
                    continue
                if p.grad.data.is_sparse:
                    raise RuntimeError('FusedAdam does not support sparse gradients, please consider SparseAdam instead')
                state = self.state[p]
                if len(state) == 0:
                    state['exp_avg'] = torch.zeros_like(p.data, dtype=torch.float32)
                    state['exp_avg_sq'] = torch.zeros_like(p.data, dtype=torch.float32)
                    if p.dtype != torch.float32:
                        if not self.master_weights:
                            raise RuntimeError('FusedAdam(capturable=True) on 16-bit parameters needs master_weights=True')
                        state['master_param'] = p.data.detach().clone().float()
                lists = by_dtype.setdefault(p.dtype, ([], [], [], [], []))
                lists[0].append(state['master_param'] if p.dtype != torch.float32 else p.data)
                lists[1].append(state['exp_avg'])
                lists[2].append(state['exp_avg_sq'])
                lists[3].append(p.grad.data)
                lists[4].append(p.data)
            for dt, lists in by_dtype.items():
                tl = list(lists) if dt != torch.float32 else list(lists[:4])   # fp32 params: no separate model copy
                hp = self._hyper(group, len(tl[0]), tl[0][0].device)
                if self.capturable:
                    gs = self._one if inv_scale is None else inv_scale.reciprocal().reshape(1).float()
                    multi_tensor_applier(self.multi_tensor_fused_adam_capturable, flag, tl, *hp, group['lr'], gs,
                                         group['step'], 1)
                else:
                    gs = 1.0 if inv_scale is None else 1.0 / float(inv_scale)
                    multi_tensor_applier(self.multi_tensor_fused_adam, flag, tl, *hp, group['lr'], gs, group['step'], 1,
                                         True)

    def step(self, closure=None, grads=None, output_params=None, scale=None, grad_norms=None, inv_scale=None,
             found_inf=None):
        if any(p is not None for p in [grads, output_params, scale, grad_norms]):
            raise RuntimeError('FusedAdam has been updated.  Simply initialize it identically to torch.optim.Adam, '
                               'and call step() with no arguments.')
        loss = None
        if closure is not None:
            loss = closure()
        if self.capturable or self.master_weights:
            self._step_mixed(inv_scale, found_inf)
            return loss

        for gid, group in enumerate(self.param_groups):
            bias_correction = 1 if group['bias_correction'] else 0
            beta1, beta2 = group['betas']
            # one step counter per group, as the reference (fused_adam.py:111-114)
            group['step'] = group.get('step', 0) + 1

            done = self._fast_step(gid, group, bias_correction, beta1, beta2)
            if done and len(done) == len(self._fast._cache[gid]['sets']):
                continue
            buckets = {torch.float16: ([], [], [], []), torch.bfloat16: ([], [], [], []),
                       torch.float32: ([], [], [], [])}
            for p in group['params']:
                if p.grad is None or p.dtype in done:
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
