"""DistributedFusedAdam -- ZeRO-2 style sharded Adam over NCCL/NVLink, the multi-GPU row of the hot path.

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


class DistributedFusedAdam:
    """Sharded mixed-precision AdamW.  Construct either from parameter sizes (``numels``; buffers are allocated here --
    what bench.py does for an 8 B-parameter set) or from ``params`` (existing tensors are copied into the flat buffer
    and re-pointed at views of it, like the reference's contiguous param/grad buffers, :333-345)."""

    def __init__(self, params=None, numels: Optional[Sequence[int]] = None, lr=1e-3, bias_correction=True,
                 betas=(0.9, 0.999), eps=1e-8, adam_w_mode=True, weight_decay=0.0, grad_dtype=torch.bfloat16,
                 param_dtype=torch.bfloat16, process_group=None, device=None, bucket_cap_mb: float = 1024.0,
                 average_grad_sync=True, overlap=True, per_param_weight_decay: Optional[Sequence[float]] = None,
                 fused_p2p: bool = False):
        assert (params is None) != (numels is None), "give exactly one of params / numels"
        self.group = process_group
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0
        if params is not None:
            params = list(params)
            if params and isinstance(params[0], dict):
                # torch-style param groups: weight_decay may differ per group (it is a per-tensor input of the kernel);
                # lr / betas / eps are shared by the whole optimizer
                groups, params, per_param_weight_decay = params, [], []
                for g in groups:
                    ps = list(g['params'])
                    params += ps
                    per_param_weight_decay += [g.get('weight_decay', weight_decay)] * len(ps)
            numels = [p.numel() for p in params]
            grad_dtype = param_dtype = params[0].dtype
            device = params[0].device if device is None else device
        self.params = params
        self.device = torch.device(device if device is not None else "cuda")
        self.grad_dtype, self.param_dtype = grad_dtype, param_dtype
        self.lr, self.betas, self.eps = lr, betas, eps
        self.bias_correction, self.adam_w_mode, self.weight_decay = bias_correction, adam_w_mode, weight_decay
        self.average_grad_sync = average_grad_sync
        self.overlap = overlap
        self.fused_p2p = bool(fused_p2p) and self.world > 1
        self.step_count = 0
        esz = torch.empty((), dtype=grad_dtype).element_size()
        self.layout = FlatLayout(numels, self.world, self.rank, esz, int(bucket_cap_mb * 1024 * 1024 / esz))
        L = self.layout
        self.frags = L.fragments()
        wd_of = list(per_param_weight_decay) if per_param_weight_decay is not None else [weight_decay] * len(numels)
        self._alloc()
        self._per_bucket = self._build_launch_sets(wd_of)
        self._init_comm()
        if params is not None:
            self._adopt_params()

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

    def _build_launch_sets(self, wd_of):
        """Per bucket: the five tensor lists [p32, m, v, g, p_copy] of fragment views + per-fragment hyper-parameters."""
        L = self.layout
        per_bucket = []
        for b in range(L.nbuckets):
            fr = [f for f in self.frags if f.bucket == b]
            lists = [[], [], [], [], []]
            for f in fr:
                n = f.flat_hi - f.flat_lo
                lists[0].append(self.p32[f.state_lo:f.state_lo + n])
                lists[1].append(self.m[f.state_lo:f.state_lo + n])
                lists[2].append(self.v[f.state_lo:f.state_lo + n])
                lists[3].append(self.grads_flat[f.flat_lo:f.flat_hi])
                lists[4].append(self.params_flat[f.flat_lo:f.flat_hi])
            k = len(fr)
            hp = None
            if k:
                t = lambda vals, dt: torch.tensor(vals, dtype=dt, device=self.device)
                hp = (t([self.betas[0]] * k, torch.float32), t([self.betas[1]] * k, torch.float32),
                      t([1 if self.bias_correction else 0] * k, torch.int32), t([self.eps] * k, torch.float32),
                      t([wd_of[f.param] for f in fr], torch.float32))
            per_bucket.append((lists, hp))
        return per_bucket

    def _init_comm(self):
        self.rs_group = self.group
        self.ag_group = self.group
        self.rs_stream = self.ag_stream = None
        if self.world > 1 and self.overlap and self.device.type == "cuda":
            ranks = list(range(dist.get_world_size())) if self.group is None else dist.get_process_group_ranks(self.group)
            # a second communicator so that all-gathers run concurrently with reduce-scatters
            self.ag_group = dist.new_group(ranks=ranks, backend="nccl")
            self.rs_stream = torch.cuda.Stream(self.device)
            self.ag_stream = torch.cuda.Stream(self.device)

    def _adopt_params(self):
        L = self.layout
        with torch.no_grad():
            for p, o in zip(self.params, L.offsets):
                n = p.numel()
                self.params_flat[o:o + n].copy_(p.detach().reshape(-1))
                p.data = self.params_flat[o:o + n].view(p.shape)
                p.grad = self.grads_flat[o:o + n].view(p.shape)
        self.init_master_from_params()

    def init_master_from_params(self):
        """fp32 master copy of the owned fragments from the (replicated) model params."""
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

    def _all_reduce_scalar(self, t, op):
        if self.world > 1:
            dist.all_reduce(t, op=op, group=self.group)

    # ---- compute -------------------------------------------------------------------------------------------------------
    def _amp_C(self):
        from ... import load_amp_C

        return load_amp_C()

    def _update_bucket(self, b: int, grad_scale):
        """grad_scale: Python float, or a 1-element CUDA tensor (then lr / step are also passed on the device and the
        kernel reads all three there -- no host round trip for a clipping coefficient computed on the GPU)."""
        lists, hp = self._per_bucket[b]
        if not lists[0]:
            return
        mode = 1 if self.adam_w_mode else 0
        if isinstance(grad_scale, torch.Tensor):
            self._amp_C().multi_tensor_fused_adam_capturable(8192, self.noop, lists, *hp, self._lr_dev, grad_scale,
                                                             self._step_dev, mode)
        else:
            self._amp_C().multi_tensor_fused_adam(8192, self.noop, lists, *hp, self.lr, grad_scale, self.step_count,
                                                  mode, True)

    def _all_fragments_launch_set(self, wd_of=None):
        if getattr(self, "_all_set", None) is None:
            lists = [[], [], [], [], []]
            hps = [[], [], [], [], []]
            for (bl, hp) in self._per_bucket:
                if not bl[0]:
                    continue
                for d in range(5):
                    lists[d] += bl[d]
                    hps[d].append(hp[d])
            self._all_set = (lists, tuple(torch.cat(h) for h in hps))
        return self._all_set

    def step_fused(self, grad_scale: float = 1.0, multicast: Optional[bool] = None):
        """One step with BOTH collectives folded into the update kernel (amp_C.multi_tensor_fused_adam_p2p): each rank
        reads its gradient shard from every peer's buffer and writes its updated parameters into every peer's buffer.
        A cross-GPU barrier on each side makes the peers' gradients complete before and the parameter writes visible
        after (symmetric-memory signal pads; stream-ordered, no host sync).  With NVSwitch multicast available
        (`has_multicast`) the sum and the broadcast happen inside the switch (multimem.ld_reduce / multimem.st)."""
        assert self.fused_p2p, "construct with fused_p2p=True"
        self.step_count += 1
        scale = grad_scale * (self.world if self.average_grad_sync else 1)
        lists, hp = self._all_fragments_launch_set()
        use_mc = self.has_multicast if multicast is None else (bool(multicast) and self.has_multicast)
        self._hdl_g.barrier(channel=0)
        if lists[0]:
            self._amp_C().multi_tensor_fused_adam_p2p(8192, self.noop, lists, *hp, self.lr, scale, self.step_count,
                                                      1 if self.adam_w_mode else 0, self.rank, self._grad_ptrs,
                                                      self._param_ptrs, self._grad_mc if use_mc else 0,
                                                      self._param_mc if use_mc else 0)
        self._hdl_p.barrier(channel=1)

    def grad_norm(self, grads_are_reduced: bool = True) -> torch.Tensor:
        """Global L2 norm of the (reduced) gradients: local shard sum of squares -> all_reduce(SUM) -> sqrt.
        Also raises self.noop on any rank if any shard holds inf/nan (all_reduce MAX)."""
        assert grads_are_reduced
        amp_C = self._amp_C()
        gl = [l for (lists, _) in self._per_bucket for l in lists[3]]
        sq = torch.zeros(1, dtype=torch.float32, device=self.device)
        if gl:
            n, _ = amp_C.multi_tensor_l2norm(65536, self.noop, [gl], False)
            sq = n * n
        self._all_reduce_scalar(sq, dist.ReduceOp.SUM)
        self._all_reduce_scalar(self.noop, dist.ReduceOp.MAX)
        return sq.sqrt()

    def _step_clipped(self, scale: float, max_grad_norm: float):
        """reduce-scatter everything -> global norm of the averaged gradients (shard sums + scalar all-reduce, reference
        :808-887) -> update with the clip factor folded into the kernel's gradient divisor (kept on the device) ->
        all-gather.  Returns the unclipped global norm as a device tensor."""
        L = self.layout
        for b in range(L.nbuckets):
            self._reduce_scatter(b)
        norm = self.grad_norm() / scale
        divisor = (scale * torch.clamp(norm / (max_grad_norm + 1e-6), min=1.0)).float().reshape(1)
        if not hasattr(self, "_lr_dev"):
            self._lr_dev = torch.zeros(1, dtype=torch.float32, device=self.device)
            self._step_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._lr_dev.fill_(self.lr)
        self._step_dev.fill_(self.step_count)
        for b in range(L.nbuckets):
            self._update_bucket(b, divisor)
            self._all_gather(b)
        return norm

    def step(self, grad_scale: float = 1.0, grads_already_reduced: bool = False, max_grad_norm: Optional[float] = None):
        """One optimizer step.  Gradients are expected in self.grads_flat (local, un-reduced unless told otherwise).
        With max_grad_norm the averaged gradients are clipped to that global L2 norm (and the norm is returned)."""
        self.step_count += 1
        L = self.layout
        scale = grad_scale * (self.world if (self.average_grad_sync and not grads_already_reduced) else 1)
        if max_grad_norm is not None:
            assert not grads_already_reduced
            return self._step_clipped(scale, float(max_grad_norm))
        if self.rs_stream is None:  # serial: single GPU, or overlap disabled
            for b in range(L.nbuckets):
                if not grads_already_reduced:
                    self._reduce_scatter(b)
                self._update_bucket(b, scale)
                self._all_gather(b)
            return
        main = torch.cuda.current_stream(self.device)
        self.rs_stream.wait_stream(main)
        self.ag_stream.wait_stream(main)
        ev_rs = []
        if not grads_already_reduced:
            with torch.cuda.stream(self.rs_stream):
                for b in range(L.nbuckets):
                    self._reduce_scatter(b)
                    e = torch.cuda.Event()
                    e.record(self.rs_stream)
                    ev_rs.append(e)
        for b in range(L.nbuckets):
            if ev_rs:
                main.wait_event(ev_rs[b])
            self._update_bucket(b, scale)
            e = torch.cuda.Event()
            e.record(main)
            with torch.cuda.stream(self.ag_stream):
                self.ag_stream.wait_event(e)
                self._all_gather(b)
        main.wait_stream(self.ag_stream)

    # ---- checkpoint -----------------------------------------------------------------------------------------------------
    def state_dict(self):
        """Per-rank shard state (the reference gathers to rank 0, :1123-1262; a sharded checkpoint is the B200-scale
        choice: 12 bytes/param never funnel through one GPU)."""
        return {"step": self.step_count, "rank": self.rank, "world": self.world, "p32": self.p32.clone(),
                "exp_avg": self.m.clone(), "exp_avg_sq": self.v.clone()}

    def load_state_dict(self, sd):
        assert sd["world"] == self.world and sd["rank"] == self.rank, "sharded checkpoint belongs to another rank/world"
        self.step_count = sd["step"]
        self.p32.copy_(sd["p32"])
        self.m.copy_(sd["exp_avg"])
        self.v.copy_(sd["exp_avg_sq"])
