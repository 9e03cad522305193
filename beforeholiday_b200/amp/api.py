"""The user-facing slice of amp that drives the hot path: `initialize` (pure-fp16/bf16 model + fp32 master weights,
i.e. opt levels O0 / O2 / O5 only), `scale_loss`, `state_dict` / `load_state_dict`.

Mirrors /root/reference/apex/amp/frontend.py:259-400 (initialize, restricted), handle.py:17-158 (scale_loss) and
frontend.py:434-474 (state dict).  What is deliberately NOT here: O1/O3/O4 -- torch-function patching and the casting
cache are Python policy with no kernel behind them ("OUT").

Flow per iteration:

    with amp.scale_loss(loss, optimizer) as scaled_loss:      # loss.float() * loss_scale
        scaled_loss.backward()
    # on exit: clear flag -> optimizer._post_amp_backward(scaler): ONE multi_tensor_scale launch unscales the fp16 grads
    # into fp32 master grads and raises the flag on inf/nan -> scaler.update_scale(): the ONE D2H read of the
    # iteration -> on overflow optimizer.step is replaced by a one-shot no-op and the scale is halved
    optimizer.step()                                          # FusedAdam/LAMB/SGD on the masters, then masters -> model
"""
from __future__ import annotations

import contextlib

import torch

from .optimizer_hooks import process_optimizer
from .scaler import LossScaler


class _AmpState(object):
    def __init__(self):
        self.loss_scalers = []
        self.enabled = False
        self.master_weights = False
        self.verbosity = 1


_amp_state = _AmpState()

_LEVELS = {
    # opt_level: (cast model to, master_weights, default loss_scale)      reference frontend.py:119-257
    "O0": (None, False, 1.0),
    "O2": (torch.float16, True, "dynamic"),
    "O5": (torch.bfloat16, True, 1.0),
}


def initialize(models, optimizers=None, enabled=True, opt_level="O2", loss_scale=None, master_weights=None,
               num_losses=1, min_loss_scale=None, max_loss_scale=2. ** 24, verbosity=1):
    """Prepare model(s) and optimizer(s) for mixed-precision training with fp32 master weights.

    Supported opt levels: "O0" (fp32, nothing changes), "O2" (fp16 model, fp32 masters, dynamic loss scale),
    "O5" (bf16 model, fp32 masters, loss scale 1).  Unlike the reference, batch-norm modules are not kept in fp32
    (keep_batchnorm_fp32) and forward inputs are not auto-cast: those belong to amp's casting policy, which this
    package does not rebuild.  Returns what it was given (model(s), optimizer(s)).
    """
    if opt_level not in _LEVELS:
        raise RuntimeError(f"opt_level {opt_level!r} is not available here: only O0, O2 and O5 (master-weight levels) "
                           "touch the multi-tensor optimizer path; O1/O3/O4 are casting policies (see DESIGN.md)")
    _amp_state.verbosity = verbosity
    _amp_state.enabled = bool(enabled)
    if not enabled:
        return (models, optimizers) if optimizers is not None else models
    cast_to, mw_default, ls_default = _LEVELS[opt_level]
    mw = mw_default if master_weights is None else bool(master_weights)
    ls = ls_default if loss_scale is None else loss_scale
    _amp_state.master_weights = mw

    model_list = models if isinstance(models, (list, tuple)) else [models]
    if cast_to is not None:
        for m in model_list:
            m.to(dtype=cast_to)
    opt_list = [] if optimizers is None else (optimizers if isinstance(optimizers, (list, tuple)) else [optimizers])
    for opt in opt_list:
        process_optimizer(opt, mw)
    _amp_state.loss_scalers = [LossScaler(ls, min_loss_scale=min_loss_scale, max_loss_scale=max_loss_scale)
                               for _ in range(num_losses)]
    return (models, optimizers) if optimizers is not None else models


@contextlib.contextmanager
def scale_loss(loss, optimizers, loss_id=0, model=None, delay_unscale=False, delay_overflow_check=False):
    """Yield ``loss.float() * loss_scale``; on exit unscale the gradients, check for overflow, and arrange for the
    next ``optimizer.step()`` to be skipped if one occurred (reference handle.py:17-158)."""
    if not _amp_state.loss_scalers:
        raise RuntimeError("Invoked 'with amp.scale_loss`, but internal Amp state has not been initialized.  "
                           "model, optimizer = amp.initialize(model, optimizer, opt_level=...) must be called "
                           "before `with amp.scale_loss`.")
    if not _amp_state.enabled:
        yield loss
        return
    if isinstance(optimizers, torch.optim.Optimizer):
        optimizers = [optimizers]
    scaler = _amp_state.loss_scalers[loss_id]
    loss_scale = scaler.loss_scale()
    if (not _amp_state.master_weights) and (not scaler.dynamic) and loss_scale == 1.0:
        yield loss.float()
        return
    if not delay_unscale:
        for opt in optimizers:
            if not opt._amp_stash.params_have_scaled_gradients:
                opt._prepare_amp_backward()

    yield loss.float() * loss_scale

    if delay_unscale:
        for opt in optimizers:
            opt._amp_stash.params_have_scaled_gradients = True
        return
    scaler.clear_overflow_state()
    for opt in optimizers:
        opt._post_amp_backward(scaler)
        opt._amp_stash.params_have_scaled_gradients = False
    should_skip = False if delay_overflow_check else scaler.update_scale()
    if not should_skip:
        return
    for opt in optimizers:
        if opt._amp_stash.already_patched:
            continue
        real_step = opt.step

        def skip_step(closure=None, opt=opt, real_step=real_step):
            if closure is not None:
                raise RuntimeError("Currently, Amp does not support closure use with optimizers.")
            if _amp_state.verbosity:
                print("Gradient overflow.  Skipping step, loss scaler {} reducing loss scale to {}".format(
                    loss_id, scaler.loss_scale()))
            for p in getattr(opt._amp_stash, "all_fp32_from_fp16_params", []):
                p.grad = None
            if hasattr(opt, "most_recent_scale"):
                opt.most_recent_scale = 1.0
                opt.scale_set_by_backward = False
            opt.step = real_step                   # one-shot: the next call is the real step again
            opt._amp_stash.already_patched = False

        opt.step = skip_step
        opt._amp_stash.already_patched = True


def master_params(optimizer):
    """Generator over the parameters the optimizer actually updates (the fp32 masters under O2/O5)."""
    for group in optimizer.param_groups:
        for p in group['params']:
            yield p


def state_dict(destination=None):
    """{'loss_scaler%d': {'loss_scale', 'unskipped'}} -- the reference's checkpoint payload (frontend.py:434-444)."""
    out = {} if destination is None else destination
    for i, s in enumerate(_amp_state.loss_scalers):
        out['loss_scaler%d' % i] = {'loss_scale': s.loss_scale(), 'unskipped': s._unskipped}
    return out


def load_state_dict(sd):
    if len(sd) != len(_amp_state.loss_scalers):
        print('Warning: state_dict contains {} entries, while {} loss_scalers are used'.format(
            len(sd), len(_amp_state.loss_scalers)))
    for i, s in enumerate(_amp_state.loss_scalers):
        entry = sd.get('loss_scaler%d' % i)
        if entry is not None:
            s._loss_scale = entry['loss_scale']
            s._unskipped = entry['unskipped']
