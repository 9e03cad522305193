"""DistributedFusedAdam -- drop-in for apex.contrib.optimizers.DistributedFusedAdam (ZeRO-2 AdamW), B200 edition.

Same class surface as the reference (/root/reference/apex/contrib/optimizers/distributed_fused_adam.py): a
torch.optim.Optimizer with the constructor of :182-203, `_step_supports_amp_scaling` (:180), `step(closure=None, *,
grad_scaler=None)` (:914), `grad_norm` / `clip_grad_norm` (:855-912), `no_sync` (:767), `grad_sync` (:795),
`zero_grad(set_to_none)` (:495), `init_params` (:382), `grad_buffer_view` (:570), `state_dict(gather_on_root=True)` /
`load_state_dict` (:1123-1280), and post-backward hooks that start a bucket's reduce-scatter as soon as every
gradient of the bucket has been produced (:325-362, 612-726).  What differs is underneath (flat_sharded_adam.py):

  * ONE flat gradient buffer and ONE flat parameter buffer per rank, laid out at construction in param-group order
    (the reference grows its buckets lazily in backward order).  Parameters whose dtype equals `param_sync_dtype` are
    re-pointed at views of the flat parameter buffer, gradients whose dtype equals `grad_sync_dtype` at views of the flat
    gradient buffer -- backward accumulates straight into the bucket, the all-gather lands straight in the parameters,
    and neither the reference's per-fragment `add_` copies (:548-553) nor its copy-out launches (:1075-1104) exist.
    Other dtypes take a converting copy on the way in (hook) and out (one multi-tensor cast per step).
  * reduce-scatter and all-gather run IN PLACE on those buffers (reduce_scatter_tensor / all_gather_into_tensor).
  * The update is one launch per (bucket, param group) of the mixed-precision Adam kernel; learning rate, step count,
    loss scale and clipping coefficient are read by the kernel from device memory, so `step()` never synchronises with
    the host (the reference does two `.item()` calls per step, :944,948) and a step skipped by the GradScaler is
    skipped on the device (found_inf becomes the kernel's no-op flag; the step counter advances only on clean steps).
  * `fused_p2p=True` (extension; needs symmetric memory, 16-bit sync dtypes): when a step needs neither clipping nor a
    grad scaler and no bucket has been reduced early (overlap_grad_sync=False), reduce-scatter + update + all-gather
    run as ONE kernel over NVLink peer memory / NVSwitch multicast.
  * `numels=[...]` (extension) instead of `params`: build the optimizer over bare sizes; parameters are created as
    views of the flat buffer.

As in the reference, every parameter of the groups has optimizer state and takes part in every step (:395-405,
:977-1115): one that received no gradient is stepped with a zero gradient, i.e. it decays by lr x weight_decay
(torch.optim.AdamW would skip it).  Keep parameters that must not move out of the optimizer.

`process_group` may be a `distributed_process_group` x `redundant_process_group` grid as in the reference (:62-74): the
state is sharded over the former and replicated over the latter; gradients are reduce-scattered over the shards'
group, then all-reduced over the replicas' group (:717-732).  `pipeline_size` and `contiguous_grad_buffer` are
accepted (the buffers here are always contiguous, the pipeline is the three-stream one of the engine).
"""
from __future__ import annotations

import contextlib
import io
import threading
from typing import Optional, Sequence

import torch
import torch.distributed as dist

from .flat_sharded_adam import FlatLayout, FlatShardedAdam, Fragment  # noqa: F401  (re-exported)

_SYNC_DTYPES = (torch.float32, torch.float16, torch.bfloat16)


class DistributedFusedAdam(torch.optim.Optimizer):
    _step_supports_amp_scaling = True
    _engine_cls = FlatShardedAdam          # the CPU tests substitute a gloo + oracle engine here

    def __init__(self, params=None, lr=1e-3, bias_correction=True, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.,
                 amsgrad=False, dtype=torch.float32, grad_sync_dtype=None, param_sync_dtype=None, device='cuda',
                 process_group=None, distributed_process_group=None, redundant_process_group=None,
                 average_grad_sync=True, overlap_grad_sync=True, bucket_cap_mb=100, pipeline_size=2,
                 contiguous_grad_buffer=False, *, numels: Optional[Sequence[int]] = None, adam_w_mode=True,
                 fused_p2p=False):
        if amsgrad:
            raise RuntimeError('DistributedFusedAdam does not support the AMSGrad variant.')
        grad_sync_dtype = dtype if grad_sync_dtype is None else grad_sync_dtype
        param_sync_dtype = dtype if param_sync_dtype is None else param_sync_dtype
        if dtype != torch.float32 or grad_sync_dtype not in _SYNC_DTYPES or param_sync_dtype not in _SYNC_DTYPES:
            raise RuntimeError('Invalid dtypes for DistributedFusedAdam '
                               f'(dtype={dtype}, grad_sync_dtype={grad_sync_dtype}, param_sync_dtype={param_sync_dtype}))')
        self.dtype, self.grad_sync_dtype, self.param_sync_dtype = dtype, grad_sync_dtype, param_sync_dtype
        self.device = torch.device(device)
        # process_group = distributed_process_group x redundant_process_group (reference :62-74, :232-262): the state is
        # sharded over the former and replicated over the latter
        self.process_group = process_group
        self.distributed_process_group = process_group if distributed_process_group is None else distributed_process_group
        self.redundant_process_group = redundant_process_group
        inited = dist.is_initialized()
        self.process_group_size = dist.get_world_size(process_group) if inited else 1
        self.distributed_size = dist.get_world_size(self.distributed_process_group) if inited else 1
        self.distributed_rank = dist.get_rank(self.distributed_process_group) if inited else 0
        self.redundant_size = dist.get_world_size(redundant_process_group) if redundant_process_group is not None else 1
        if self.process_group_size != self.distributed_size * self.redundant_size:
            raise RuntimeError('Invalid process group configuration '
                               f'(process group size = {self.process_group_size}, '
                               f'distributed process group size = {self.distributed_size}, '
                               f'redundant process group size = {self.redundant_size})')
        self.average_grad_sync = average_grad_sync
        self.overlap_grad_sync = overlap_grad_sync
        self.greedy_grad_copy = True
        self.pipeline_size = pipeline_size
        self.contiguous_grad_buffer = True
        self.adam_w_mode = adam_w_mode

        # ---- parameters: given, or created over bare sizes as views of the flat buffer (filled in below) ----------------
        created = numels is not None
        if created:
            assert params is None, "give exactly one of params / numels"
            params = [torch.nn.Parameter(torch.empty(0, dtype=param_sync_dtype, device=self.device), requires_grad=True)
                      for _ in numels]
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay)
        super().__init__(params, defaults)
        self._params = [p for g in self.param_groups for p in g['params']]
        self._index = {p: i for i, p in enumerate(self._params)}
        group_of = [gi for gi, g in enumerate(self.param_groups) for _ in g['params']]
        sizes = [int(n) for n in numels] if created else [p.numel() for p in self._params]

        # ---- engine: layout, flat buffers, fp32 shard state, collectives, kernels ------------------------------------------
        # bucket_cap_mb counts bytes of the gradient sync dtype, as in the reference (:262-270)
        self._eng = self._engine_cls(sizes, adam_w_mode=adam_w_mode, grad_dtype=grad_sync_dtype,
                                     param_dtype=param_sync_dtype, process_group=self.distributed_process_group,
                                     device=self.device, bucket_cap_mb=bucket_cap_mb,
                                     average_grad_sync=average_grad_sync, overlap=True, fused_p2p=fused_p2p,
                                     groups=self.param_groups, group_of=group_of,
                                     redundant_group=redundant_process_group if self.redundant_size > 1 else None)
        E, L = self._eng, self._eng.layout
        self.bucket_size, self.shard_size = L.bucket_elems, L.sub_shard
        # buckets that hold a piece of parameter i: its flat range [o, o + n) cut at multiples of the bucket size
        self._param_buckets = [list(range(o // L.bucket_elems, (o + max(n, 1) - 1) // L.bucket_elems + 1))
                               for o, n in zip(L.offsets, sizes)]
        self._bucket_params = [[] for _ in range(L.nbuckets)]
        for i, bs in enumerate(self._param_buckets):
            for b in bs:
                self._bucket_params[b].append(i)

        # ---- adopt the parameters ----------------------------------------------------------------------------------------
        self._flat_param = []        # parameter i lives in the flat parameter buffer (no copy-out needed)
        with torch.no_grad():
            for i, (p, o, n) in enumerate(zip(self._params, L.offsets, sizes)):
                view = E.params_flat[o:o + n]
                if created:
                    p.data = view
                    self._flat_param.append(True)
                    continue
                view.copy_(p.detach().reshape(-1))
                flat = p.dtype == param_sync_dtype and p.device == E.params_flat.device and p.is_contiguous()
                if flat:
                    p.data = view.view(p.shape)
                self._flat_param.append(flat)
            if self.process_group_size > 1 and not created:
                # every rank starts from rank 0's values (the reference broadcasts parameter by parameter, :336-340)
                self._broadcast_flat_params()
                for i, p in enumerate(self._params):
                    if not self._flat_param[i]:
                        o, n = L.offsets[i], sizes[i]
                        p.copy_(E.params_flat[o:o + n].view(p.shape))
            self._init_masters()
        self._cast_sets = self._build_cast_sets()

        # ---- gradient bookkeeping ----------------------------------------------------------------------------------------
        self._lock = threading.Lock()
        self._grads_generated = set()
        self._in_backward = False
        self._bucket_reduced = [False] * L.nbuckets      # reduce-scatter issued since the last zero_grad / step
        self._bucket_dirty = [False] * L.nbuckets        # local gradients added since the bucket's last reduce-scatter
        self._num_grads = sum(1 for p in self._params if p.requires_grad)
        self._inv_grad_scale = torch.ones(1, dtype=torch.float32, device=self.device)
        self._clipped = False
        self._grad_norm = None
        self._step_dev = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._step_host_stale = False
        self.state['step'] = 0
        self._point_grads_at_flat_buffer()
        self._grad_accs = []
        self._register_post_backward_hooks()

    def _broadcast_flat_params(self):
        """Rank 0 of the WHOLE process group (all shards, all replicas) is the source."""
        E = self._eng
        if self.redundant_size == 1:
            E.broadcast_params()
        else:
            src = 0 if self.process_group is None else dist.get_global_rank(self.process_group, 0)
            buf = E.params_flat if E.params_flat.is_cuda else E.params_flat.float()
            dist.broadcast(buf, src=src, group=self.process_group)
            if buf is not E.params_flat:
                E.params_flat.copy_(buf)

    # flat buffers and layout (bench.py and the tests fill gradients directly)
    @property
    def grads_flat(self):
        return self._eng.grads_flat

    @property
    def params_flat(self):
        return self._eng.params_flat

    @property
    def layout(self):
        return self._eng.layout

    @property
    def noop(self):
        return self._eng.noop

    # ------------------------------------------------------------------------------------------------------------------
    # construction helpers
    def _init_masters(self):
        """fp32 master shard from the MODEL parameters (not from the possibly narrower sync copy), reference :481-489."""
        E = self._eng
        for f in E.frags:
            n = f.flat_hi - f.flat_lo
            src = self._params[f.param].detach().reshape(-1)[f.param_lo:f.param_lo + n]
            E.p32[f.state_lo:f.state_lo + n].copy_(src)

    def _build_cast_sets(self):
        """[synced flat views], [model parameters] for the parameters that do not live in the flat buffer."""
        L = self._eng.layout
        src, dst, slow = [], [], []
        for i, p in enumerate(self._params):
            if self._flat_param[i]:
                continue
            view = self._eng.params_flat[L.offsets[i]:L.offsets[i] + p.numel()]
            if p.is_contiguous() and p.device == view.device and p.dtype in _SYNC_DTYPES:
                src.append(view)
                dst.append(p.detach().view(-1))
            else:
                slow.append((view, p))
        return src, dst, slow

    def _grad_view(self, i):
        L = self._eng.layout
        p = self._params[i]
        return self._eng.grads_flat[L.offsets[i]:L.offsets[i] + p.numel()].view(p.shape)

    def _grad_is_flat(self, i):
        p = self._params[i]
        return p.dtype == self.grad_sync_dtype and p.device == self._eng.grads_flat.device

    def _point_grads_at_flat_buffer(self):
        for i, p in enumerate(self._params):
            if p.requires_grad and self._grad_is_flat(i):
                p.grad = self._grad_view(i)

    def _register_post_backward_hooks(self):
        """One hook per parameter, run right after its gradient has been accumulated (reference :325-362)."""
        for i, p in enumerate(self._params):
            if not p.requires_grad:
                continue

            def hook(param, i=i):
                self._on_grad_ready(i)

            self._grad_accs.append(p.register_post_accumulate_grad_hook(hook))

    # ------------------------------------------------------------------------------------------------------------------
    # gradient flow: hook -> (copy into the flat buffer) -> bucket full -> reduce-scatter on the side stream
    def _on_grad_ready(self, i):
        with self._lock:
            if not self._in_backward:
                # first gradient of this backward pass: have the autograd engine call us back when the pass is over
                self._in_backward = True
                self._grads_generated.clear()
                torch.autograd.Variable._execution_engine.queue_callback(self._on_backward_done)
            if self.greedy_grad_copy or self._is_in_flat_buffer(i):
                self._grad_copy(i)
                self._grads_generated.add(i)
                if self.overlap_grad_sync:
                    self._try_start_bucket_grad_sync([i])

    def _is_in_flat_buffer(self, i):
        p = self._params[i]
        return p.grad is not None and self._grad_is_flat(i) and p.grad.data_ptr() == self._grad_view(i).data_ptr()

    def _grad_copy(self, i):
        """Make the flat buffer hold parameter i's gradient contribution (reference :517-556).  A gradient that already
        IS a view of the flat buffer has nothing to move; anything else is added in (converting) and released."""
        p = self._params[i]
        if p.grad is None:
            return
        view = self._grad_view(i)
        if p.grad.data_ptr() != view.data_ptr():
            view.add_(p.grad.detach().to(view.dtype) if p.grad.dtype != view.dtype else p.grad.detach())
            p.grad = view if self._grad_is_flat(i) else None
        for b in self._param_buckets[i]:
            self._bucket_dirty[b] = True
        self._grad_norm = None

    def _try_start_bucket_grad_sync(self, params):
        """Start the reduce-scatter of every bucket all of whose gradients have been produced (reference :612-664)."""
        ready = []
        for i in params:
            for b in self._param_buckets[i]:
                if self._bucket_dirty[b] and all((j in self._grads_generated) or not self._params[j].requires_grad
                                                 for j in self._bucket_params[b]):
                    ready.append(b)
        if ready:
            self._start_bucket_grad_sync(sorted(set(ready)), during_backward=True)

    def _start_bucket_grad_sync(self, buckets, during_backward):
        """In-place reduce-scatter of `buckets` on the engine's side stream.  During backward the rest of each bucket is
        cleared behind its reduce-scatter: a later micro-batch accumulates into the same buffer, and a second
        reduce-scatter must then add only NEW contributions to the owner's running sum."""
        E, L = self._eng, self._eng.layout
        if self.process_group_size > 1:
            E.reduce_gradients(buckets, side_stream=True)
            if during_backward:
                ctx = torch.cuda.stream(E.rs_stream) if E.rs_stream is not None else contextlib.nullcontext()
                with ctx:
                    for b in buckets:
                        lo, hi = L.bucket_range(b)
                        olo, ohi = L.owned_range(b)
                        E.grads_flat[lo:olo].zero_()
                        E.grads_flat[ohi:hi].zero_()
                        if E.rs_stream is not None:
                            E._rs_events[b].record(E.rs_stream)
        for b in buckets:
            self._bucket_dirty[b] = False
            self._bucket_reduced[b] = True
        self._grad_norm = None

    def _on_backward_done(self):
        """End of a backward pass (autograd engine callback): the next pass writes into the same flat buffer, so the
        compute stream must not run ahead of the reduce-scatters still in flight."""
        with self._lock:
            self._in_backward = False
            self._grads_generated.clear()
            E = self._eng
            if E.rs_stream is not None and getattr(E, "_rs_events", None):
                main = torch.cuda.current_stream(self.device)
                for e in E._rs_events.values():
                    main.wait_event(e)

    @contextlib.contextmanager
    def no_sync(self, greedy_grad_copy=False):
        """Disable overlapped gradient synchronization inside the context (reference :767-793), like DDP.no_sync."""
        old = (self.greedy_grad_copy, self.overlap_grad_sync)
        self.greedy_grad_copy, self.overlap_grad_sync = greedy_grad_copy, False
        try:
            yield
        finally:
            self.greedy_grad_copy, self.overlap_grad_sync = old

    def grad_sync(self):
        """Ensure that all gradients are synchronized (reference :795-807): move whatever still sits in p.grad into the
        flat buffer, reduce-scatter every bucket that holds unreduced contributions, wait."""
        with self._lock:
            for i, p in enumerate(self._params):
                if p.grad is not None and not self._is_in_flat_buffer(i):
                    self._grad_copy(i)      # (gradients accumulated in place were accounted for by their hooks)
            E = self._eng
            pending = [b for b in range(E.layout.nbuckets) if self._bucket_dirty[b] or not self._bucket_reduced[b]]
            if pending:
                self._start_bucket_grad_sync(pending, during_backward=False)
            E.wait_reductions()
            self._grads_generated.clear()

    def zero_grad(self, set_to_none=True):
        """Clear parameter gradients (reference :495-515).  The flat buffer is zeroed in one memset; gradients that live
        in it stay attached as views (there is nothing to free), others are dropped or zeroed as asked."""
        with self._lock:
            E = self._eng
            E.wait_reductions()
            E.grads_flat.zero_()
            for i, p in enumerate(self._params):
                if p.requires_grad and self._grad_is_flat(i):
                    p.grad = self._grad_view(i)
                elif p.grad is None or set_to_none:
                    p.grad = None
                else:
                    p.grad.zero_()
            n = E.layout.nbuckets
            self._bucket_dirty, self._bucket_reduced = [False] * n, [False] * n
            self._grads_generated.clear()
            self._inv_grad_scale = torch.ones(1, dtype=torch.float32, device=self.device)
            self._clipped = False
            self._grad_norm = None

    def init_params(self, params=None):
        """Reference :382-408 initializes optimizer state lazily; here every parameter's state exists from construction."""

    def grad_buffer_view(self, param):
        """View of the flat gradient buffer that corresponds to `param` (reference :570-590)."""
        return self._grad_view(self._index[param]).detach()

    # ------------------------------------------------------------------------------------------------------------------
    # norms and clipping
    def _local_grad_norm(self, parameters=(), norm_type=2.0):
        """Square of the 2-norm of this rank's shard of the (averaged) gradients (reference :808-853)."""
        assert float(norm_type) == 2.0
        self.grad_sync()
        E = self._eng
        idx = None
        if parameters and len(parameters) != self._num_grads:
            idx = [self._index[p] for p in parameters]
        if not hasattr(self, "_norm_flag"):
            self._norm_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        sq = E.local_grad_sq(idx, flag=self._norm_flag)
        if self.average_grad_sync and self.process_group_size > 1:
            sq = sq / float(self.process_group_size) ** 2   # the buffer holds SUMS; the reference's holds averages
        return sq.detach().view([])

    def grad_norm(self, parameters=(), norm_type=2.0, force=False):
        """Gradient norm of the parameters in the optimizer, cached until the gradients change (reference :855-887)."""
        if force or self._grad_norm is None:
            sq = self._local_grad_norm(list(parameters), norm_type).reshape(1)
            self._eng._all_reduce_scalar(sq, dist.ReduceOp.SUM)
            self._grad_norm = sq.sqrt().view([])
        return self._grad_norm.detach()

    def clip_grad_norm(self, max_norm, parameters=(), norm_type=2.0):
        """Clip the gradient norm; the scaling is deferred into the optimizer step (reference :889-912)."""
        assert max_norm > 0
        total_norm = self.grad_norm(parameters=parameters, norm_type=norm_type)
        self._inv_grad_scale = torch.clamp((total_norm + 1e-6) / max_norm, min=1.0).view(1).float()
        self._clipped = True
        return total_norm

    # ------------------------------------------------------------------------------------------------------------------
    # the step
    def _lr_to_device(self):
        """All groups' learning rates in one pinned-to-device copy (schedulers change them between steps)."""
        lr_dev, _ = self._eng._dev_scalars()
        lrs = [float(g['lr']) for g in self.param_groups]
        if getattr(self, "_lr_cached", None) != lrs:
            if self.device.type == "cuda":
                if not hasattr(self, "_lr_host"):
                    self._lr_host = torch.zeros(len(lrs), dtype=torch.float32, pin_memory=True)
                self._lr_host.copy_(torch.tensor(lrs, dtype=torch.float32))
                lr_dev.copy_(self._lr_host, non_blocking=True)
            else:
                lr_dev.copy_(torch.tensor(lrs, dtype=torch.float32))
            self._lr_cached = lrs
        return lr_dev

    def _can_fuse(self, grad_scaler):
        E = self._eng
        return (E.fused_p2p and grad_scaler is None and not self._clipped and not any(self._bucket_reduced)
                and self.distributed_size > 1)

    def step(self, closure=None, *, grad_scaler=None):
        """Apply the AdamW step (reference :914-1121).  No host synchronisation: with a grad scaler, found_inf is
        computed on the device from the gradient norm and becomes the update kernels' no-op flag."""
        loss = None
        if closure is not None:
            loss = closure()
        E = self._eng
        if self._can_fuse(grad_scaler):
            # one kernel: reduce-scatter + update + all-gather over NVLink peer memory
            with self._lock:
                for i, p in enumerate(self._params):
                    if p.grad is not None and not self._is_in_flat_buffer(i):
                        self._grad_copy(i)
            E.step_fused()
            self._step_dev.add_(1)
            if not self._step_host_stale:
                self.state['step'] += 1
            self._after_update()
            return loss

        self.grad_sync()
        world_div = float(self.process_group_size) if (self.average_grad_sync and self.process_group_size > 1) else 1.0
        E.noop.zero_()
        if grad_scaler is not None:
            grad_norm = self.grad_norm()
            found_inf = torch.logical_not(torch.isfinite(grad_norm)).reshape(1)
            scaler_state = grad_scaler._per_optimizer_states[id(self)]
            scaler_state['found_inf_per_device'] = {found_inf.device: found_inf.float()}
            assert grad_scaler._scale is not None
            E.noop.copy_(found_inf)                                     # the kernels skip themselves when it is set
            self._step_dev.add_(1 - found_inf.to(torch.int32))          # bias correction counts clean steps only
            divisor = self._inv_grad_scale * grad_scaler._scale.reshape(1).float() * world_div
            self._step_host_stale = True
        else:
            self._step_dev.add_(1)
            divisor = self._inv_grad_scale * world_div
            if not self._step_host_stale:
                self.state['step'] += 1
        E.step_count += 1
        self._lr_to_device()
        E.update_and_gather(divisor.reshape(1), step=self._step_dev)
        self._after_update()
        return loss

    def _after_update(self):
        """Parameters that do not live in the flat buffer get their values from it (one converting multi-tensor copy;
        reference :1075-1104 launches one per bucket and dtype); bucket bookkeeping starts over."""
        src, dst, slow = self._cast_sets
        E = self._eng
        if src:
            E.cast_out(src, dst)
        for view, p in slow:
            p.detach().copy_(view.view(p.shape))
        n = E.layout.nbuckets
        self._bucket_dirty, self._bucket_reduced = [False] * n, [False] * n
        self._grad_norm = None

    # ------------------------------------------------------------------------------------------------------------------
    # checkpointing
    def _host_step(self):
        if self._step_host_stale:
            self.state['step'] = int(self._step_dev.item())
            self._step_host_stale = False
        return self.state['step']

    def _bucket_state(self):
        E = self._eng
        return {'params_shard': E.p32, 'exp_avg_shard': E.m, 'exp_avg_sq_shard': E.v, 'world': self.distributed_size,
                'rank': self.distributed_rank, 'engine_step': E.step_count}

    def state_dict(self, gather_on_root=True):
        """Optimizer state (reference :1123-1262).  gather_on_root=True: every rank must call; the root rank of the
        process group gets {'gathered_states': [one serialized state per rank]}, the others None.  False: this rank's
        own shard (p32 / exp_avg / exp_avg_sq of its 1/world of the flat space) -- the sharded checkpoint."""
        self._host_step()
        self.state['buckets'] = self._bucket_state()
        state_dict = super().state_dict()
        if not gather_on_root:
            return state_dict
        buf = io.BytesIO()
        torch.save(state_dict, buf)
        mine = buf.getvalue()
        if self.distributed_size == 1:
            return {'gathered_states': [mine]}
        gathered = [None] * self.distributed_size if self.distributed_rank == 0 else None
        # (with replicas every shard group gathers on its own first rank; the replicas' states are identical)
        g = self.distributed_process_group
        root = dist.get_global_rank(g, 0) if g is not None else 0
        dist.gather_object(mine, gathered, dst=root, group=g)
        return {'gathered_states': gathered} if self.distributed_rank == 0 else None

    def load_state_dict(self, state_dict):
        """Load optimizer state (reference :1264-1280): a gathered dict (each rank takes its own entry) or a local one."""
        if 'gathered_states' in state_dict:
            blob = state_dict['gathered_states'][self.distributed_rank]
            state_dict = torch.load(io.BytesIO(blob), map_location=self.device)
        super().load_state_dict(state_dict)
        E = self._eng
        b = self.state['buckets']
        if b['world'] != self.distributed_size or b['rank'] != self.distributed_rank:
            raise RuntimeError(f"optimizer state was saved by rank {b['rank']} of {b['world']}; this is rank "
                               f"{self.distributed_rank} of {self.distributed_size} (resharding is not supported)")
        with torch.no_grad():
            E.p32.copy_(b['params_shard'])
            E.m.copy_(b['exp_avg_shard'])
            E.v.copy_(b['exp_avg_sq_shard'])
        E.step_count = int(b['engine_step'])
        self._step_dev.fill_(int(self.state['step']))
        self._step_host_stale = False
        self._lr_cached = None
        E.groups = self.param_groups     # load_state_dict rebuilt the group dicts; the launch sets follow them
        self.state['buckets'] = self._bucket_state()
