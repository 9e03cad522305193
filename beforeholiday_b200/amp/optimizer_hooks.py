"""Optimizer hooks of amp's dynamic-loss-scaling flow -- the part of apex.amp that sits ON the hot path
(/root/reference/apex/amp/_process_optimizer.py:14-25 master->model copy, :93-202 gradient stash / unscale hooks,
:321-489 `_process_optimizer`).

`process_optimizer(optimizer, master_weights)` gives an optimizer the attributes the reference's `amp.scale_loss`
drives, under the reference's names, so FusedSGD's master-weight path (which reads ``_amp_stash``) and user code that
pokes at the stash keep working:

    optimizer._amp_stash                      fp16_groups / fp32_from_fp16_groups / fp32_from_fp32_groups, all_* lists,
                                              gradient stashes, flags
    optimizer._prepare_amp_backward()         before backward: stash or drop the existing grads
    optimizer._post_amp_backward(scaler)      after backward: fp16 grads -> unscaled fp32 master grads through ONE
                                              amp_C.multi_tensor_scale launch (+ one multi_tensor_axpby launch when
                                              accumulating into existing master grads); the overflow flag is raised by
                                              those kernels
    optimizer._master_params_to_model_params()   ONE amp_C.multi_tensor_scale(masters -> fp16 model params, 1.0)
    optimizer.step / zero_grad                wrapped as in the reference (:353-388)

Only the kernels' callers are rebuilt here; amp's casting policy (opt-level function patching) is out of scope.
"""
from __future__ import annotations

import types

import torch

from ..multi_tensor_apply import multi_tensor_applier
from ..optimizers.fused_sgd import FusedSGD

_LOW = (torch.float16, torch.bfloat16)


class AmpOptimizerState(object):
    """Bag of per-optimizer amp state (same role and name as the reference's class)."""

    def __init__(self):
        self.lazy_init_called = False
        self.already_patched = False
        self.params_have_scaled_gradients = False


def _unscale_models_are_masters(scaler, params, stashed_grads, scale_override=None):
    """Gradients that live in the master dtype already: unscale in place, or merge with the stashed (accumulated)
    gradients (reference :93-135)."""
    grads_have_scale, stashed_have_scale, out_scale = scaler.loss_scale(), 1.0, 1.0
    if scaler.loss_scale() == 1.0 and not scaler.dynamic:
        stashed_grads[:] = [None] * len(stashed_grads)
        return
    if scale_override is not None:
        grads_have_scale, stashed_have_scale, out_scale = scale_override
    fresh, merged, stashed = [], [], []
    for param, old in zip(params, stashed_grads):
        if param.grad is None and old is not None:
            param.grad = old
        elif param.grad is not None and old is None:
            fresh.append(param.grad)
        elif param.grad is not None and old is not None:
            merged.append(param.grad)
            stashed.append(old)
    if fresh:
        scaler.unscale(fresh, fresh, None, models_are_masters=True, scale_override=grads_have_scale / out_scale)
    if merged:
        scaler.unscale_with_stashed(merged, stashed, merged, scale_override=(grads_have_scale, stashed_have_scale, out_scale))
    stashed_grads[:] = [None] * len(stashed_grads)


# ---- with fp32 master weights ---------------------------------------------------------------------------------------
def _lazy_init_with_master_weights(self):
    stash = self._amp_stash
    stash.fp16_groups, stash.fp32_from_fp16_groups, stash.fp32_from_fp32_groups = [], [], []
    for group in self.param_groups:
        low, masters, full = [], [], []
        for i, param in enumerate(group['params']):
            if not param.requires_grad:
                continue
            if param.is_cuda and param.dtype in _LOW:
                low.append(param)
                master = param.detach().clone().float()
                master.requires_grad = True
                group['params'][i] = master          # the optimizer now updates the master
                masters.append(master)
                if param in self.state:
                    self.state[master] = self.state.pop(param)
            elif param.is_cuda and param.dtype == torch.float32:
                full.append(param)
            else:
                raise TypeError("Optimizer's parameters must one of torch.cuda.FloatTensor, torch.cuda.HalfTensor, "
                                "torch.cuda.BFloat16Tensor. Received {}".format(param.type()))
        stash.fp16_groups.append(low)
        stash.fp32_from_fp16_groups.append(masters)
        stash.fp32_from_fp32_groups.append(full)
    stash.all_fp16_params = [p for g in stash.fp16_groups for p in g]
    stash.all_fp32_from_fp16_params = [p for g in stash.fp32_from_fp16_groups for p in g]
    stash.all_fp32_from_fp32_params = [p for g in stash.fp32_from_fp32_groups for p in g]
    stash.all_fp16_grad_stash = [None] * len(stash.all_fp16_params)
    stash.all_fp32_from_fp32_grad_stash = [None] * len(stash.all_fp32_from_fp32_params)
    for p in stash.all_fp32_from_fp16_params + stash.all_fp32_from_fp32_params:
        p.grad = None
    self.load_state_dict(self.state_dict())   # re-key the state after the param swap (reference :89)


def _prepare_backward_with_master_weights(self):
    stash = self._amp_stash
    self._amp_lazy_init()
    for p in stash.all_fp16_params:
        p.grad = None                                      # fp16 grads are consumed each iteration
    for i, p in enumerate(stash.all_fp32_from_fp32_params):
        stash.all_fp32_from_fp32_grad_stash[i] = p.grad    # fp32 grads may be accumulating
        p.grad = None


def _post_backward_with_master_weights(self, scaler):
    stash = self._amp_stash
    self._amp_lazy_init()
    model_new, master_new, model_acc, master_acc = [], [], [], []
    for low, master in zip(stash.all_fp16_params, stash.all_fp32_from_fp16_params):
        if low.grad is None:
            continue
        if master.grad is None:
            master.grad = torch.empty_like(master)
            model_new.append(low.grad)
            master_new.append(master.grad)
        else:
            model_acc.append(low.grad)
            master_acc.append(master.grad)
    if model_new:      # one launch: fp16 -> fp32, x 1/scale, inf/nan -> flag
        scaler.unscale(model_new, master_new, scaler.loss_scale(), models_are_masters=False)
    if model_acc:      # one launch: master += model/scale, checking only the incoming grads
        scaler.unscale_with_stashed(model_acc, master_acc, master_acc)
    _unscale_models_are_masters(scaler, stash.all_fp32_from_fp32_params, stash.all_fp32_from_fp32_grad_stash)


def _master_params_to_model_params(self):
    stash = self._amp_stash
    if stash.all_fp16_params:
        multi_tensor_applier(stash.multi_tensor_scale, stash.dummy_overflow_buf,
                             [stash.all_fp32_from_fp16_params, stash.all_fp16_params], 1.0)


# FusedSGD folds the unscale into its own kernel when materialize_master_grads is False (reference :266-303)
def _prepare_backward_with_master_weights_fused_sgd(self):
    if self.materialize_master_grads:
        return _prepare_backward_with_master_weights(self)
    stash = self._amp_stash
    self._amp_lazy_init()
    for i, p in enumerate(stash.all_fp16_params):
        stash.all_fp16_grad_stash[i] = p.grad
        p.grad = None
    for i, p in enumerate(stash.all_fp32_from_fp32_params):
        stash.all_fp32_from_fp32_grad_stash[i] = p.grad
        p.grad = None


def _post_backward_with_master_weights_fused_sgd(self, scaler):
    if self.materialize_master_grads:
        return _post_backward_with_master_weights(self, scaler)
    stash = self._amp_stash
    self._amp_lazy_init()
    grads_have_scale = scaler.loss_scale()
    stashed_have_scale = self.most_recent_scale
    out_scale = grads_have_scale
    if self.scale_set_by_backward:
        out_scale = min(grads_have_scale, self.most_recent_scale)
    for params, stashed in ((stash.all_fp16_params, stash.all_fp16_grad_stash),
                            (stash.all_fp32_from_fp32_params, stash.all_fp32_from_fp32_grad_stash)):
        _unscale_models_are_masters(scaler, params, stashed, (grads_have_scale, stashed_have_scale, out_scale))
    self.most_recent_scale = out_scale          # the SGD kernel divides by this
    self.scale_set_by_backward = True


# ---- without master weights -----------------------------------------------------------------------------------------
def _lazy_init_no_master_weights(self):
    stash = self._amp_stash
    stash.all_fp16_params, stash.all_fp32_params = [], []
    for group in self.param_groups:
        for param in group['params']:
            if param.is_cuda and param.dtype in _LOW:
                stash.all_fp16_params.append(param)
            elif param.is_cuda and param.dtype == torch.float32:
                stash.all_fp32_params.append(param)
            else:
                raise TypeError("Optimizer's parameters must be one of torch.cuda.FloatTensor, torch.cuda.HalfTensor, "
                                "torch.BFloat16Tensor. Received {}".format(param.type()))
    stash.all_fp16_grad_stash = [None] * len(stash.all_fp16_params)
    stash.all_fp32_grad_stash = [None] * len(stash.all_fp32_params)


def _prepare_backward_no_master_weights(self):
    stash = self._amp_stash
    self._amp_lazy_init()
    for params, stashed in ((stash.all_fp16_params, stash.all_fp16_grad_stash),
                            (stash.all_fp32_params, stash.all_fp32_grad_stash)):
        for i, p in enumerate(params):
            stashed[i] = p.grad
            p.grad = None


def _post_backward_no_master_weights(self, scaler):
    stash = self._amp_stash
    self._amp_lazy_init()
    for params, stashed in ((stash.all_fp16_params, stash.all_fp16_grad_stash),
                            (stash.all_fp32_params, stash.all_fp32_grad_stash)):
        _unscale_models_are_masters(scaler, params, stashed)


def _amp_lazy_init(self):
    stash = self._amp_stash
    if not stash.lazy_init_called:
        self._lazy_init_maybe_master_weights()
        stash.lazy_init_called = True


def process_optimizer(optimizer, master_weights: bool):
    """Attach amp's hooks to `optimizer` (reference `_process_optimizer`, :321-489)."""
    if hasattr(optimizer, "_amp_stash"):
        raise RuntimeError("A given optimizer should only be passed through amp.initialize once.")
    optimizer._amp_stash = AmpOptimizerState()
    for name in ("_lazy_init_maybe_master_weights", "_master_params_to_model_params", "_prepare_amp_backward",
                 "_post_amp_backward", "_amp_lazy_init"):
        if hasattr(optimizer, name):
            raise RuntimeError("Incoming optimizer already has {} defined.".format(name))
    if not multi_tensor_applier.available:
        raise RuntimeError("beforeholiday_b200.amp requires the CUDA extension (no Python fallback)")
    from .. import load_amp_C

    amp_C = load_amp_C()
    stash = optimizer._amp_stash
    stash.multi_tensor_scale = amp_C.multi_tensor_scale
    stash.multi_tensor_l2norm = amp_C.multi_tensor_l2norm
    stash.dummy_overflow_buf = torch.zeros(1, dtype=torch.int, device="cuda")
    bind = lambda fn: types.MethodType(fn, optimizer)
    optimizer._amp_lazy_init = bind(_amp_lazy_init)
    is_sgd = isinstance(optimizer, FusedSGD)

    if master_weights:
        optimizer._lazy_init_maybe_master_weights = bind(_lazy_init_with_master_weights)
        optimizer._master_params_to_model_params = bind(_master_params_to_model_params)
        old_step = optimizer.step

        def new_step(self, closure=None):
            if closure is not None:
                raise RuntimeError("Currently, Amp does not support closure use with optimizers.")
            retval = old_step()
            if not is_sgd:                      # FusedSGD writes the fp16 copy inside its kernel
                self._master_params_to_model_params()
            for p in self._amp_stash.all_fp32_from_fp16_params:
                p.grad = None                   # master grads are rebuilt from the model grads every iteration
            return retval

        def new_zero_grad(self, set_to_none=None):
            st = self._amp_stash
            self._amp_lazy_init()
            for p in st.all_fp16_params + st.all_fp32_from_fp32_params:
                if p.grad is not None:
                    p.grad.detach_()
                    p.grad.zero_()
            for p in st.all_fp32_from_fp16_params:
                p.grad = None

        optimizer.step = bind(new_step)
        optimizer.zero_grad = bind(new_zero_grad)
        optimizer._prepare_amp_backward = bind(_prepare_backward_with_master_weights_fused_sgd if is_sgd
                                               else _prepare_backward_with_master_weights)
        optimizer._post_amp_backward = bind(_post_backward_with_master_weights_fused_sgd if is_sgd
                                            else _post_backward_with_master_weights)
    else:
        optimizer._lazy_init_maybe_master_weights = bind(_lazy_init_no_master_weights)
        optimizer._prepare_amp_backward = bind(_prepare_backward_no_master_weights)
        optimizer._post_amp_backward = bind(_post_backward_no_master_weights)
    return optimizer
