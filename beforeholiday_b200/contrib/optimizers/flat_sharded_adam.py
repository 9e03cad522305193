"""FlatShardedAdam -- the engine under DistributedFusedAdam: ZeRO-2 style sharded Adam over NCCL/NVLink on one flat
parameter space (the multi-GPU row of the hot path).  The drop-in optimizer class with the reference's constructor,
step(grad_scaler=...), clip_grad_norm, no_sync, checkpointing and backward-overlapped gradient sync lives in
distributed_fused_adam.py and drives this engine; bench.py drives it directly (parameter sizes only, no nn.Module).

Follows the structure of the reference's apex.contrib.optimizers.DistributedFusedAdam
(/root/reference/apex/contrib/optimizers/distributed_fused_adam.py:19, bucket layout :279-288, grad sync :668-742,
shard update + param all-gather :977-1115) re-laid-out for one NVSwitch node:

  flat parameter space   every parameter gets a 128-byte-aligned offset in one flat space, padded to
                         nbuckets x (world x sub_shard) elements
  bucket b               flat[b*B, (b+1)*B), B = world * sub_shard; rank r owns sub-shard [b*B + r*S, b*B + (r+1)*S)
  per rank, resident     full flat gradients (grad dtype, e.g. bf16), full flat model params (param-sync dtype),
                         fp32 p / m / v only for the owned sub-shards (1/world of the optimizer state)
  step(), per bucket     reduce_scatter(grads, in place) -> ONE mixed-precision Adam launch over the bucket's parameter
                         fragments (amp_C.multi_tensor_fused_adam: fp32 p/m/v, 16-bit g in, 16-bit params out,
                         g / grad_scale fused) -> all_gather(params, in place)
                         The three stages of successive buckets run on three streams (two NCCL communicators), so the
                         update and the two collectives overlap; NVLink carries both directions at once.
  step_fused()           (fused_p2p=True) ONE kernel does all three: it sums the gradient shard straight out of every
                         peer's buffer and stores the updated parameters straight into every peer's buffer
                         (amp_C.multi_tensor_fused_adam_p2p over symmetric memory), a barrier on each side.
  grad norm / clipping   local sum of squares over the owned fragments (amp_C.multi_tensor_l2norm) -> scalar
                         all_reduce(SUM) -> sqrt (reference :808-887); the overflow flag is all_reduce(MAX)-ed.

The collectives are `torch.distributed` (NCCL) calls issued from Python, as in the reference; all arithmetic is in
libmt_b200.so.  There is no CPU path: the update raises without the CUDA extension.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch
import torch.distributed as dist


def _ceil_to(x: int, a: int) -> int:
    return (x + a - 1) // a * a


def second_communicator(group, backend="nccl"):
    """Another communicator over the ranks of `group` (None = the whole job), so that two collectives of one step can
    be in flight at once.  Collective over the WHOLE job: when `group` is a sub-group (the state sharded over part of
    the job, the rest holding replicas), every rank creates every such group's communicator, in the same order, as
    torch.distributed.new_group requires -- creating only one's own would give disjoint groups the same name."""
    n = dist.get_world_size()
    ranks = tuple(range(n)) if group is None else tuple(dist.get_process_group_ranks(group))
    if len(ranks) == n:
        return dist.new_group(ranks=list(ranks), backend=backend)
    every = [None] * n
    dist.all_gather_object(every, ranks)
    mine = None
    for rs in sorted(set(every)):
        g = dist.new_group(ranks=list(rs), backend=backend)
        if rs == ranks:
            mine = g
    return mine


@dataclass
class Fragment:
    """Piece of parameter `param` that lies inside one of this rank's sub-shards."""
    param: int      # index into the flat parameter list
    bucket: int
    flat_lo: int    # range in the flat space
    flat_hi: int
    state_lo: int   # offset in this rank's compact fp32 state
    param_lo: int   # offset inside the parameter


class FlatLayout:
    """Pure host logic (no GPU): offsets, buckets, shard ownership, fragments."""

    ALIGN_BYTES = 128  # per-parameter alignment, as the reference (distributed_fused_adam.py:281)

    def __init__(self, numels: Sequence[int], world: int, rank: int, elem_bytes: int = 2,
                 bucket_cap_elems: int = 512 * 1024 * 1024):
        assert world >= 1 and 0 <= rank < world
        self.numels = [int(n) for n in numels]
        self.world, self.rank = world, rank
        align = max(1, self.ALIGN_BYTES // elem_bytes)
        self.offsets = []
        o = 0
        for n in self.numels:
            self.offsets.append(o)
            o += _ceil_to(n, align)
        self.total = o
        unit = world * 256   # every sub-shard starts on a 256-element (>= 512-byte) boundary
        nb = max(1, -(-self.total // max(unit, _ceil_to(bucket_cap_elems, unit))))
        self.bucket_elems = _ceil_to(-(-self.total // nb), unit)
        self.nbuckets = -(-self.total // self.bucket_elems)
        self.padded = self.nbuckets * self.bucket_elems
        self.sub_shard = self.bucket_elems // world
        self.state_elems = self.nbuckets * self.sub_shard

    def bucket_range(self, b: int):
        return b * self.bucket_elems, (b + 1) * self.bucket_elems

    def owned_range(self, b: int, rank: Optional[int] = None):
        r = self.rank if rank is None else rank
        lo = b * self.bucket_elems + r * self.sub_shard
        return lo, lo + self.sub_shard

    def fragments(self, rank: Optional[int] = None) -> List[Fragment]:
        out = []
        import bisect

        ends = [o + n for o, n in zip(self.offsets, self.numels)]
        for b in range(self.nbuckets):
            lo, hi = self.owned_range(b, rank)
            i = bisect.bisect_right(ends, lo)  # first param whose end is > lo
            while i < len(self.numels) and self.offsets[i] < hi:
                flo, fhi = max(lo, self.offsets[i]), min(hi, ends[i])
                if fhi > flo:
                    out.append(Fragment(i, b, flo, fhi, b * self.sub_shard + (flo - lo), flo - self.offsets[i]))
                i += 1
        return out


class FlatShardedAdam:
    """Sharded mixed-precision AdamW over one flat parameter space, constructed from parameter SIZES: the flat gradient
    and parameter buffers and this rank's fp32 p / m / v shard are allocated here.  Hyper-parameters are per GROUP
    (`groups`: dicts with lr, betas, eps, weight_decay, bias_correction; `group_of[i]` = group of parameter i), as in
    torch.optim param groups; lr is read from the group dict at every step, so schedulers work."""

    def __init__(self, numels: Sequence[int], lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8,
                 adam_w_mode=True, weight_decay=0.0, grad_dtype=torch.bfloat16, param_dtype=torch.bfloat16,
                 process_group=None, device=None, bucket_cap_mb: float = 1024.0, average_grad_sync=True, overlap=True,
                 per_param_weight_decay: Optional[Sequence[float]] = None, fused_p2p: bool = False,
                 groups: Optional[List[dict]] = None, group_of: Optional[Sequence[int]] = None, redundant_group=None):
        self.group = process_group
        # optimizer state is SHARDED over `process_group` and REPLICATED over `redundant_group` (reference: a
        # distributed_size x redundant_size grid, distributed_fused_adam.py:62-74); gradients are reduce-scattered over
        # the former, then all-reduced over the latter (:717-732)
        self.redundant_group = redundant_group
        self.redundant_size = dist.get_world_size(redundant_group) if redundant_group is not None else 1
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        numels = [int(n) for n in numels]
        self.device = torch.device(device if device is not None else "cuda")
        self.grad_dtype, self.param_dtype = grad_dtype, param_dtype
        self.adam_w_mode = adam_w_mode
        self.average_grad_sync = average_grad_sync
        self.overlap = overlap
        self.fused_p2p = bool(fused_p2p) and self.world > 1 and self.redundant_size == 1
        # fused multicast route: sixteenths of the flat buffer whose gradient loads / parameter stores go peer-to-peer
        # instead of through the switch (0, 0 = everything through the switch); see DistAdamP2POp in csrc/mt_ops.cuh
        self.hybrid_split = (0, 0)
        self.step_count = 0
        if groups is None:
            # one group -- or, with per_param_weight_decay, one group per distinct decay value
            wds = list(per_param_weight_decay) if per_param_weight_decay is not None else [weight_decay] * len(numels)
            distinct = sorted(set(wds))
            groups = [dict(lr=lr, betas=betas, eps=eps, weight_decay=w, bias_correction=bias_correction) for w in distinct]
            group_of = [distinct.index(w) for w in wds]
        assert group_of is not None and len(group_of) == len(numels)
        self.groups, self.group_of = groups, [int(g) for g in group_of]
        esz = torch.empty((), dtype=grad_dtype).element_size()
        self.layout = FlatLayout(numels, self.world, self.rank, esz, int(bucket_cap_mb * 1024 * 1024 / esz))
        self.frags = self.layout.fragments()
        self._alloc()
        self._per_bucket = self._build_launch_sets()
        self._init_comm()

    # legacy single-group accessors (bench.py / tools set and read these)
    @property
    def lr(self):
        return self.groups[0]["lr"]

    @lr.setter
    def lr(self, v):
        for g in self.groups:
            g["lr"] = v

    # ---- buffers ---------------------------------------------------------------------------------------------------
    def _alloc(self):
        L = self.layout
        if self.fused_p2p:
            # Symmetric allocations: every rank maps every other rank's buffers, so the update kernel can read the
            # gradient shard from, and write the updated parameters to, all peers directly over NVLink.
            import torch.distributed._symmetric_memory as symm_mem

            assert self.grad_dtype == self.param_dtype, "fused_p2p needs one 16-bit dtype for gradients and parameters"
            grp = dist.group.WORLD if self.group is None else self.group
            self.grads_flat = symm_mem.empty(L.padded, dtype=self.grad_dtype, device=self.device)
            self.params_flat = symm_mem.empty(L.padded, dtype=self.param_dtype, device=self.device)
            self.grads_flat.zero_()
            self.params_flat.zero_()
            self._hdl_g = symm_mem.rendezvous(self.grads_flat, group=grp)
            self._hdl_p = symm_mem.rendezvous(self.params_flat, group=grp)
            self._grad_ptrs = [int(x) for x in self._hdl_g.buffer_ptrs]
            self._param_ptrs = [int(x) for x in self._hdl_p.buffer_ptrs]
            # NVSwitch multicast addresses (0 when the fabric has no NVLS): in-switch reduction and broadcast
            self._grad_mc = int(getattr(self._hdl_g, "multicast_ptr", 0) or 0)
            self._param_mc = int(getattr(self._hdl_p, "multicast_ptr", 0) or 0)
            self.has_multicast = self._grad_mc != 0 and self._param_mc != 0
            assert self._grad_ptrs[self.rank] == self.grads_flat.data_ptr()
            assert self._param_ptrs[self.rank] == self.params_flat.data_ptr()
        else:
            self.grads_flat = torch.zeros(L.padded, dtype=self.grad_dtype, device=self.device)
            self.params_flat = torch.zeros(L.padded, dtype=self.param_dtype, device=self.device)
        self.p32 = torch.zeros(L.state_elems, dtype=torch.float32, device=self.device)
        self.m = torch.zeros(L.state_elems, dtype=torch.float32, device=self.device)
        self.v = torch.zeros(L.state_elems, dtype=torch.float32, device=self.device)
        self.noop = torch.zeros(1, dtype=torch.int32, device=self.device)

    @property
    def copy_in_kernel(self) -> bool:
        """The update kernel writes the rounded parameter copy itself when gradients and synced parameters share a
        dtype (its p_copy list has the gradient's type); otherwise the owned shard is cast from the fp32 masters."""
        return self.grad_dtype == self.param_dtype

    def _hp_tensors(self, gid: int, k: int):
        g = self.groups[gid]
        t = lambda val, dt: torch.full((k,), val, dtype=dt, device=self.device)
        return (t(g["betas"][0], torch.float32), t(g["betas"][1], torch.float32),
                t(1 if g.get("bias_correction", True) else 0, torch.int32), t(g["eps"], torch.float32),
                t(g["weight_decay"], torch.float32))

    @staticmethod
    def _hp_key(g):
        return (g["betas"][0], g["betas"][1], bool(g.get("bias_correction", True)), g["eps"], g["weight_decay"])

    def _build_launch_sets(self):
        """Per bucket, per group: the tensor lists [p32, m, v, g(, p_copy)] of fragment views + per-fragment
        hyper-parameter arrays.  Entry = dict(gid, lists, hp, key)."""
        L = self.layout
        per_bucket = []
        nl = 5 if self.copy_in_kernel else 4
        for b in range(L.nbuckets):
            by_group = {}
            for f in self.frags:
                if f.bucket != b:
                    continue
                lists = by_group.setdefault(self.group_of[f.param], [[] for _ in range(nl)])
                n = f.flat_hi - f.flat_lo
                lists[0].append(self.p32[f.state_lo:f.state_lo + n])
                lists[1].append(self.m[f.state_lo:f.state_lo + n])
                lists[2].append(self.v[f.state_lo:f.state_lo + n])
                lists[3].append(self.grads_flat[f.flat_lo:f.flat_hi])
                if nl == 5:
                    lists[4].append(self.params_flat[f.flat_lo:f.flat_hi])
            per_bucket.append([dict(gid=gid, lists=lists, hp=self._hp_tensors(gid, len(lists[0])),
                                    key=self._hp_key(self.groups[gid])) for gid, lists in sorted(by_group.items())])
        return per_bucket

    def _refresh_hp(self, entry):
        """Group hyper-parameters other than lr live in device arrays; rebuild them when the group dict changed."""
        key = self._hp_key(self.groups[entry["gid"]])
        if key != entry["key"]:
            entry["hp"] = self._hp_tensors(entry["gid"], len(entry["lists"][0]))
            entry["key"] = key

    def _init_comm(self):
        self.rs_group = self.group
        self.ag_group = self.group
        self.rs_stream = self.ag_stream = None
        if self.world > 1 and self.overlap and self.device.type == "cuda":
            # a second communicator so that all-gathers run concurrently with reduce-scatters
            self.ag_group = second_communicator(self.group)
            self.rs_stream = torch.cuda.Stream(self.device)
            self.ag_stream = torch.cuda.Stream(self.device)

    def init_master_from_params(self):
        """fp32 master copy of the owned fragments from the (replicated) flat model params."""
        for f in self.frags:
            n = f.flat_hi - f.flat_lo
            self.p32[f.state_lo:f.state_lo + n].copy_(self.params_flat[f.flat_lo:f.flat_hi])

    def zero_grad(self):
        self.grads_flat.zero_()

    # ---- collectives (overridable: the CPU tests swap in gloo-compatible ones) ------------------------------------
    def _reduce_scatter(self, b: int):
        L = self.layout
        lo, hi = L.bucket_range(b)
        olo, ohi = L.owned_range(b)
        if self.world > 1:
            dist.reduce_scatter_tensor(self.grads_flat[olo:ohi], self.grads_flat[lo:hi], op=dist.ReduceOp.SUM,
                                       group=self.rs_group)

    def _all_gather(self, b: int):
        L = self.layout
        lo, hi = L.bucket_range(b)
        olo, ohi = L.owned_range(b)
        if self.world > 1:
            dist.all_gather_into_tensor(self.params_flat[lo:hi], self.params_flat[olo:ohi], group=self.ag_group)

    def _reduce_redundant(self, b: int):
        """Sum the owned sub-shard over the replicas of this shard (no-op without a redundant group)."""
        if self.redundant_size > 1:
            olo, ohi = self.layout.owned_range(b)
            dist.all_reduce(self.grads_flat[olo:ohi], op=dist.ReduceOp.SUM, group=self.redundant_group)

    def _all_reduce_scalar(self, t, op):
        if self.world > 1:
            dist.all_reduce(t, op=op, group=self.group)

    def broadcast_params(self, src_rank_in_group: int = 0):
        """Every rank takes the flat parameter buffer of one rank (construction time)."""
        if self.world > 1:
            src = src_rank_in_group if self.group is None else dist.get_global_rank(self.group, src_rank_in_group)
            dist.broadcast(self.params_flat, src=src, group=self.group)

    # ---- compute -------------------------------------------------------------------------------------------------------
    def _amp_C(self):
        from ... import load_amp_C

        return load_amp_C()

    def _launch_update(self, entry, lr, grad_scale, step):
        """One launch of the mixed-precision Adam kernel over one (bucket, group) fragment set.  lr / grad_scale / step
        are Python numbers, or 1-element CUDA tensors (all three then: the kernel reads them on the device, so a
        clipping coefficient or a loss scale computed on the GPU never visits the host).  Overridable (CPU tests)."""
        mode = 1 if self.adam_w_mode else 0
        if isinstance(grad_scale, torch.Tensor):
            self._amp_C().multi_tensor_fused_adam_capturable(8192, self.noop, entry["lists"], *entry["hp"], lr, grad_scale,
                                                             step, mode)
        else:
            self._amp_C().multi_tensor_fused_adam(8192, self.noop, entry["lists"], *entry["hp"], lr, grad_scale, step,
                                                  mode, True)

    def _cast_owned_shard(self, b: int):
        """grad dtype != synced-parameter dtype: the kernel wrote no copy; cast the bucket's owned masters instead."""
        L = self.layout
        olo, ohi = L.owned_range(b)
        self.params_flat[olo:ohi].copy_(self.p32[b * L.sub_shard:(b + 1) * L.sub_shard])

    def cast_out(self, src, dst):
        """One converting multi-tensor copy src[i] -> dst[i] (fused_adam_cuda.maybe_cast_mt in the reference); does
        nothing when the step was skipped (the flag is set)."""
        self._amp_C().maybe_cast_mt(8192, self.noop, [src, dst])

    def _dev_scalars(self):
        """Device copies of every group's lr and of the step count (capturable route)."""
        if not hasattr(self, "_lr_dev"):
            self._lr_dev = torch.zeros(len(self.groups), dtype=torch.float32, device=self.device)
            self._step_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
        return self._lr_dev, self._step_dev

    def _update_bucket(self, b: int, grad_scale, step=None):
        """grad_scale: Python float, or a 1-element CUDA tensor (then lr / step are also passed on the device)."""
        on_dev = isinstance(grad_scale, torch.Tensor)
        for entry in self._per_bucket[b]:
            self._refresh_hp(entry)
            gid = entry["gid"]
            if on_dev:
                lr_dev, step_dev = self._dev_scalars()
                self._launch_update(entry, lr_dev[gid:gid + 1], grad_scale, step_dev if step is None else step)
            else:
                self._launch_update(entry, self.groups[gid]["lr"], grad_scale, self.step_count)
        if not self.copy_in_kernel:
            self._cast_owned_shard(b)

    def _group_launch_sets(self):
        """All buckets' fragments of one group in one launch set (fused route: one kernel per group per step)."""
        if getattr(self, "_all_sets", None) is None:
            merged = {}
            for entries in self._per_bucket:
                for e in entries:
                    d = merged.setdefault(e["gid"], [[] for _ in e["lists"]])
                    for i, l in enumerate(e["lists"]):
                        d[i] += l
            self._all_sets = [dict(gid=gid, lists=lists, hp=self._hp_tensors(gid, len(lists[0])),
                                   key=self._hp_key(self.groups[gid])) for gid, lists in sorted(merged.items())]
        return self._all_sets

    def step_fused(self, grad_scale: float = 1.0, multicast: Optional[bool] = None, ld_p2p_of16: Optional[int] = None,
                   st_p2p_of16: Optional[int] = None):
        """One step with BOTH collectives folded into the update kernel (amp_C.multi_tensor_fused_adam_p2p): each rank
        reads its gradient shard from every peer's buffer and writes its updated parameters into every peer's buffer.
        A cross-GPU barrier on each side makes the peers' gradients complete before and the parameter writes visible
        after (symmetric-memory signal pads; stream-ordered, no host sync).  With NVSwitch multicast available
        (`has_multicast`) the sum and the broadcast happen inside the switch (multimem.ld_reduce / multimem.st);
        ld_p2p_of16 / st_p2p_of16 (default: self.hybrid_split) send that share of the blocks peer-to-peer instead."""
        assert self.fused_p2p, "construct with fused_p2p=True"
        self.step_count += 1
        self.noop.zero_()
        scale = grad_scale * (self.world if self.average_grad_sync else 1)
        # default route: through the switch from 4 ranks on; at 2 ranks plain peer access moves S/2 per direction where
        # the switch moves 1.5 S (it also reads / writes the requester's own copy) -- 23.9 vs 40.0 ms on Llama-3-8B
        use_mc = (self.has_multicast and self.world > 2) if multicast is None else (bool(multicast) and self.has_multicast)
        self._hdl_g.barrier(channel=0)
        for e in self._group_launch_sets():
            self._refresh_hp(e)
            ld16 = self.hybrid_split[0] if ld_p2p_of16 is None else int(ld_p2p_of16)
            st16 = self.hybrid_split[1] if st_p2p_of16 is None else int(st_p2p_of16)
            kw = {"ld_p2p_of16": ld16, "st_p2p_of16": st16} if use_mc else {}
            self._amp_C().multi_tensor_fused_adam_p2p(8192, self.noop, e["lists"], *e["hp"], self.groups[e["gid"]]["lr"],
                                                      scale, self.step_count, 1 if self.adam_w_mode else 0, self.rank,
                                                      self._grad_ptrs, self._param_ptrs, self._grad_mc if use_mc else 0,
                                                      self._param_mc if use_mc else 0, **kw)
        self._hdl_p.barrier(channel=1)

    def owned_grad_fragments(self, params: Optional[Sequence[int]] = None):
        """Views of this rank's (reduced) gradient fragments, all or those of the given parameter indices."""
        sel = None if params is None else set(params)
        return [self.grads_flat[f.flat_lo:f.flat_hi] for f in self.frags if sel is None or f.param in sel]

    def local_grad_sq(self, params: Optional[Sequence[int]] = None, flag: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Sum of squares over this rank's reduced gradient fragments (1-element tensor); raises `flag` (default: the
        step's skip flag self.noop) when a chunk's sum is inf/nan."""
        key = None if params is None else tuple(sorted(params))
        cache = self.__dict__.setdefault("_norm_lists", {})
        if key not in cache:
            cache[key] = self.owned_grad_fragments(params)
        gl = cache[key]
        if not gl:
            return torch.zeros(1, dtype=torch.float32, device=self.device)
        n, _ = self._amp_C().multi_tensor_l2norm(65536, self.noop if flag is None else flag, [gl], False)
        return n * n

    def grad_norm(self, grads_are_reduced: bool = True, params: Optional[Sequence[int]] = None) -> torch.Tensor:
        """Global L2 norm of the (reduced, summed) gradients: local shard sum of squares -> all_reduce(SUM) -> sqrt
        (reference :808-887).  Also raises self.noop on every rank if any shard holds inf/nan (all_reduce MAX)."""
        assert grads_are_reduced
        sq = self.local_grad_sq(params)
        self._all_reduce_scalar(sq, dist.ReduceOp.SUM)
        self._all_reduce_scalar(self.noop, dist.ReduceOp.MAX)
        return sq.sqrt()

    # ---- the step, in its two halves -----------------------------------------------------------------------------------
    def reduce_gradients(self, buckets: Optional[Sequence[int]] = None, side_stream: bool = True):
        """Reduce-scatter the given buckets (default: all), in place.  With side_stream the collectives run on the
        reduce-scatter stream, ordered behind the work already queued on the current stream, and an event per bucket
        is left in self._rs_events for the update to wait on -- the caller's stream is NOT blocked (this is what a
        post-backward hook calls).  Otherwise they run on the current stream."""
        L = self.layout
        buckets = range(L.nbuckets) if buckets is None else buckets
        if not hasattr(self, "_rs_events"):
            self._rs_events = {}
        if self.rs_stream is None or not side_stream:
            for b in buckets:
                self._reduce_scatter(b)
                self._reduce_redundant(b)
            return
        main = torch.cuda.current_stream(self.device)
        self.rs_stream.wait_stream(main)
        with torch.cuda.stream(self.rs_stream):
            for b in buckets:
                self._reduce_scatter(b)
                self._reduce_redundant(b)
                e = torch.cuda.Event()
                e.record(self.rs_stream)
                self._rs_events[b] = e

    def wait_reductions(self):
        """Order the current stream behind every outstanding side-stream reduce-scatter."""
        evs = getattr(self, "_rs_events", {})
        if evs:
            main = torch.cuda.current_stream(self.device)
            for e in evs.values():
                main.wait_event(e)
            evs.clear()

    def update_and_gather(self, grad_scale, step=None):
        """Shard update + parameter all-gather, bucket by bucket; bucket b's all-gather (second communicator, own
        stream) overlaps bucket b+1's update.  Waits per bucket on its reduce-scatter event when one is pending."""
        L = self.layout
        evs = getattr(self, "_rs_events", {})
        if self.ag_stream is None:
            for b in range(L.nbuckets):
                self._update_bucket(b, grad_scale, step)
                self._all_gather(b)
            return
        main = torch.cuda.current_stream(self.device)
        self.ag_stream.wait_stream(main)
        for b in range(L.nbuckets):
            e = evs.pop(b, None)
            if e is not None:
                main.wait_event(e)
            self._update_bucket(b, grad_scale, step)
            e = torch.cuda.Event()
            e.record(main)
            with torch.cuda.stream(self.ag_stream):
                self.ag_stream.wait_event(e)
                self._all_gather(b)
        main.wait_stream(self.ag_stream)

    def _step_clipped(self, scale: float, max_grad_norm: float):
        """reduce-scatter everything -> global norm of the averaged gradients (shard sums + scalar all-reduce, reference
        :808-887) -> update with the clip factor folded into the kernel's gradient divisor (kept on the device) ->
        all-gather.  Returns the unclipped global norm as a device tensor."""
        self.reduce_gradients(side_stream=False)
        norm = self.grad_norm() / scale
        divisor = (scale * torch.clamp(norm / (max_grad_norm + 1e-6), min=1.0)).float().reshape(1)
        lr_dev, step_dev = self._dev_scalars()
        lr_dev.copy_(torch.tensor([g["lr"] for g in self.groups], dtype=torch.float32))
        step_dev.fill_(self.step_count)
        self.update_and_gather(divisor)
        return norm

    def step(self, grad_scale: float = 1.0, grads_already_reduced: bool = False, max_grad_norm: Optional[float] = None):
        """One optimizer step.  Gradients are expected in self.grads_flat (local, un-reduced unless told otherwise).
        With max_grad_norm the averaged gradients are clipped to that global L2 norm (and the norm is returned).
        The skip flag (self.noop) starts every step at zero: an overflow seen by one step's norm pass skips that step
        only."""
        self.step_count += 1
        self.noop.zero_()
        scale = grad_scale * (self.world * self.redundant_size if (self.average_grad_sync and not grads_already_reduced) else 1)
        if max_grad_norm is not None:
            assert not grads_already_reduced
            return self._step_clipped(scale, float(max_grad_norm))
        if not grads_already_reduced:
            self.reduce_gradients()
        self.update_and_gather(scale)

    # ---- checkpoint -----------------------------------------------------------------------------------------------------
    def state_dict(self):
        """Per-rank shard state (a sharded checkpoint: 12 bytes/param never funnel through one GPU)."""
        return {"step": self.step_count, "rank": self.rank, "world": self.world, "p32": self.p32.clone(),
                "exp_avg": self.m.clone(), "exp_avg_sq": self.v.clone()}

    def load_state_dict(self, sd):
        assert sd["world"] == self.world and sd["rank"] == self.rank, "sharded checkpoint belongs to another rank/world"
        self.step_count = sd["step"]
        self.p32.copy_(sd["p32"])
        self.m.copy_(sd["exp_avg"])
        self.v.copy_(sd["exp_avg_sq"])
