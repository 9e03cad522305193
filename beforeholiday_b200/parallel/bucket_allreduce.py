"""Bucketed gradient all-reduce: the step on the optimizer path's upstream side
(/root/reference/apex/parallel/distributed.py:426-478, `allreduce_bucket`; csrc/flatten_unflatten.cpp).

The reference flattens a bucket with `torch._utils._flatten_dense_tensors` (one copy kernel per tensor via cat),
optionally pre-divides (`mul_`), all-reduces, averages (`mul_`) and un-flattens with `amp_C.multi_tensor_scale(..., 1.0)`:
up to four full passes over the bucket besides the collective.  Here both scalings ride on the copies, each of which
is ONE table-driven launch:

    multi_tensor_scale([grads] -> [views of a persistent flat buffer], 1/predivide)      flatten  (+ pre-divide, + cast)
    all_reduce(flat)                                                                     NCCL
    multi_tensor_scale([views of flat] -> [grads], predivide/world)                      unflatten (+ average, + cast)

The flat buffer (optionally fp32: `allreduce_always_fp32`) is kept between calls, so the tensor tables stay cached and
a steady-state call uploads nothing.  The overflow flag raised by the second launch tells the caller whether the
reduced gradients contain inf/nan -- for free.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.distributed as dist

from ..multi_tensor_apply import multi_tensor_applier


class FlatBucketAllreduce:
    def __init__(self, tensors: Sequence[torch.Tensor], process_group=None, gradient_average: bool = True,
                 gradient_predivide_factor: float = 1.0, allreduce_always_fp32: bool = False):
        if not multi_tensor_applier.available:
            raise RuntimeError("beforeholiday_b200.parallel requires the CUDA extension (no fallback)")
        from .. import load_amp_C

        self.tensors = list(tensors)
        assert self.tensors and all(t.is_cuda and t.dtype == self.tensors[0].dtype for t in self.tensors), \
            "a bucket holds CUDA tensors of one dtype (the reference splits buckets by type, distributed.py:60-76)"
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_initialized() else 1
        self.gradient_average = gradient_average
        self.predivide = float(gradient_predivide_factor)
        dev = self.tensors[0].device
        flat_dtype = torch.float32 if allreduce_always_fp32 else self.tensors[0].dtype
        # 128-byte aligned slots so that every view takes the vector path
        align = 128 // torch.empty((), dtype=flat_dtype).element_size()
        offs, o = [], 0
        for t in self.tensors:
            offs.append(o)
            o += (t.numel() + align - 1) // align * align
        self.flat = torch.zeros(o, dtype=flat_dtype, device=dev)
        self.views = [self.flat[a:a + t.numel()] for a, t in zip(offs, self.tensors)]
        self.overflow_buf = torch.zeros(1, dtype=torch.int32, device=dev)
        self._scale = load_amp_C().multi_tensor_scale

    def __call__(self, tensors: Optional[Sequence[torch.Tensor]] = None) -> torch.Tensor:
        """All-reduce (and average) the bucket in place; returns the flat reduced buffer.  `tensors` may be passed when
        the gradient tensors were re-allocated since construction (same shapes)."""
        ts = self.tensors if tensors is None else list(tensors)
        src = [t.reshape(-1) if t.is_contiguous() else t for t in ts]
        self.overflow_buf.zero_()
        multi_tensor_applier(self._scale, self.overflow_buf, [src, self.views], 1.0 / self.predivide)
        if self.world > 1:
            dist.all_reduce(self.flat, group=self.group)
        post = (self.predivide / self.world) if self.gradient_average else self.predivide
        multi_tensor_applier(self._scale, self.overflow_buf, [self.views, src], post)
        return self.flat


def allreduce_bucket(bucket: Sequence[torch.Tensor], process_group=None, gradient_average: bool = True,
                     gradient_predivide_factor: float = 1.0, allreduce_always_fp32: bool = False) -> torch.Tensor:
    """One-shot form with the reference's argument meaning (a fresh flat buffer per call)."""
    return FlatBucketAllreduce(bucket, process_group, gradient_average, gradient_predivide_factor,
                               allreduce_always_fp32)()


# ---- apex_C.flatten / unflatten and the flat_dist_call helpers (apex/parallel/distributed.py:13-76) -------------------
def _amp_scale():
    from .. import load_amp_C

    return load_amp_C().multi_tensor_scale


def flatten(bucket: Sequence[torch.Tensor]) -> torch.Tensor:
    """One contiguous 1-D tensor holding the bucket's tensors back to back (csrc/flatten_unflatten.cpp:5-8), written by
    ONE multi_tensor_scale launch instead of one copy kernel per tensor."""
    bucket = list(bucket)
    flat = torch.empty(sum(t.numel() for t in bucket), dtype=bucket[0].dtype, device=bucket[0].device)
    if flat.numel():
        flag = torch.zeros(1, dtype=torch.int32, device=flat.device)
        multi_tensor_applier(_amp_scale(), flag, [[t.reshape(-1) for t in bucket], list(_views(flat, bucket))], 1.0)
    return flat


def _views(flat, like):
    o = 0
    for t in like:
        yield flat[o:o + t.numel()]
        o += t.numel()


def unflatten(coalesced: torch.Tensor, bucket: Sequence[torch.Tensor]):
    """Views of `coalesced` shaped like the bucket's tensors (csrc/flatten_unflatten.cpp:10-13)."""
    return [v.view(t.shape) for v, t in zip(_views(coalesced, bucket), bucket)]


def apply_flat_dist_call(bucket, call, extra_args=None):
    """flatten -> call(coalesced, *extra_args) -> (average if call is all_reduce) -> copy back (:36-49).  The average
    rides on the copy back, which is one launch for every tensor that is contiguous in the default layout.  A tensor in
    another dense layout (channels_last weights, transposed views) was flattened in LOGICAL order by `flatten`, and
    `reshape(-1)` of it is a temporary, so it is written back through `copy_` on a shaped view of the coalesced buffer,
    as the reference does for every tensor (`buf.copy_(synced)`, :48-49)."""
    bucket = list(bucket)
    coalesced = flatten(bucket)
    call(coalesced, *(extra_args or ()))
    post = 1.0 / dist.get_world_size() if call is dist.all_reduce else 1.0
    views = list(_views(coalesced, bucket))
    src = [v for v, t in zip(views, bucket) if t.is_contiguous()]
    dst = [t.view(-1) for t in bucket if t.is_contiguous()]
    if src:
        flag = torch.zeros(1, dtype=torch.int32, device=coalesced.device)
        multi_tensor_applier(_amp_scale(), flag, [src, dst], post)
    for v, t in zip(views, bucket):
        if not t.is_contiguous():
            t.copy_(v.view(t.shape) if post == 1.0 else v.view(t.shape) * post)


def split_by_type(tensors):
    buckets = {}
    for t in tensors:
        buckets.setdefault(t.type(), []).append(t)
    return buckets


def flat_dist_call(tensors, call, extra_args=None):
    """Organise `tensors` by type and run `call` on each type's flattened bucket (:68-73): how the reference's
    DistributedDataParallel broadcasts the parameters from rank 0 at construction (:252)."""
    for bucket in split_by_type(tensors).values():
        apply_flat_dist_call(bucket, call, extra_args)
