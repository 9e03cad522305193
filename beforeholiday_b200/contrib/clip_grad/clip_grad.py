"""clip_grad_norm_ -- drop-in for apex.contrib.clip_grad.clip_grad_norm_
(/root/reference/apex/contrib/clip_grad/clip_grad.py:16-128): torch.nn.utils.clip_grad_norm_ with the 2-norm of CUDA
fp32 / fp16 / bf16 gradients computed by ``amp_C.multi_tensor_l2norm`` (one launch pair per dtype) and the scaling done
by ``amp_C.multi_tensor_scale`` in place.  Other norm types, and CPU-only parameter sets, defer to torch.

The clip coefficient is kept on the device until ``multi_tensor_scale`` needs it as a Python float (its signature takes
``float scale``), exactly like the reference: that conversion is the one host sync of the call.
"""
from __future__ import annotations

from typing import Iterable, Union

import torch

from ...multi_tensor_apply import multi_tensor_applier

_tensor_or_tensors = Union[torch.Tensor, Iterable[torch.Tensor]]


def clip_grad_norm_(parameters: _tensor_or_tensors, max_norm: float, norm_type: float = 2.0,
                    error_if_nonfinite: bool = False) -> torch.Tensor:
    if isinstance(parameters, torch.Tensor):
        parameters = [parameters]
    parameters = [p for p in parameters if p.grad is not None]
    max_norm = float(max_norm)
    norm_type = float(norm_type)
    if len(parameters) == 0:
        return torch.tensor(0.)
    if not (multi_tensor_applier.available and norm_type == 2.0 and any(p.is_cuda for p in parameters)):
        return torch.nn.utils.clip_grad_norm_(parameters, max_norm, norm_type=norm_type,
                                              error_if_nonfinite=error_if_nonfinite)
    from ... import load_amp_C

    amp_C = load_amp_C()
    device = next(p.device for p in parameters if p.is_cuda)
    fused = {torch.float32: [], torch.float16: [], torch.bfloat16: []}
    misc = []
    for p in parameters:
        g = p.grad.detach()
        if p.dtype in fused and p.device == device and g.is_contiguous():
            fused[p.dtype].append(g)
        else:
            misc.append(g)
    flag = torch.zeros([1], dtype=torch.int32, device=device)
    norms = [multi_tensor_applier(amp_C.multi_tensor_l2norm, flag, [gs], False)[0] for gs in fused.values() if gs]
    norms += [torch.linalg.norm(g).unsqueeze(0).to(device) for g in misc]
    total_norm = torch.linalg.norm(torch.cat(norms))
    if error_if_nonfinite and torch.logical_or(total_norm.isnan(), total_norm.isinf()):
        raise RuntimeError(f'The total norm of order {norm_type} for gradients from `parameters` is non-finite, so it '
                           'cannot be clipped. To disable this error and scale the gradients by the non-finite norm '
                           'anyway, set `error_if_nonfinite=False`')
    clip_coef_clamped = torch.clamp(max_norm / (total_norm + 1e-6), max=1.0)
    for gs in fused.values():
        if gs:
            multi_tensor_applier(amp_C.multi_tensor_scale, flag, [gs, gs], clip_coef_clamped)
    for g in misc:
        g.mul_(clip_coef_clamped.to(g.device))
    return total_norm
