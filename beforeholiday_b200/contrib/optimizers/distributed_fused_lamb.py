"""DistributedFusedLAMB -- ZeRO-2 style sharded LAMB over NCCL/NVLink.

Mirrors the public surface of the reference's apex.contrib.optimizers.DistributedFusedLAMB
(/root/reference/apex/contrib/optimizers/distributed_fused_lamb.py:83-135 constructor, :590-671 gradient reductions +
global norm, :694-748 step pipeline, :799-870 global_scale / L2_grad_norm / complete_reductions / step, :872-910
state_dict) on the flat layout of DistributedFusedAdam (one NVSwitch node, `FlatLayout`):

  per rank, resident     full flat gradients and full flat model parameters (the model's dtype; the model's tensors ARE
                         views of these buffers, so the reference's flatten copy (:750-781) and its maybe_cast_mt copy
                         back into the model (:850-866) do not exist), fp32 p / m / v / u for the owned sub-shards only
  complete_reductions()  per bucket reduce_scatter(AVG, in place)  ->  sum of squares over the owned shards
                         (amp_C.multi_tensor_l2norm)  ->  scalar all_reduce  ->  L2_grad_norm (stays on the device)
  step()                 overflow flag and step counter on the device (:702-714)
                         ONE stage-1 launch over all owned fragments (amp_C.multi_tensor_lamb_compute_update_term) that
                         also yields, per parameter, this shard's sum of round_model(p)^2 and u^2: the two extra norm
                         sweeps of the reference (:673-692) -- one of which reads every FULL parameter on every rank
                         -- become one [2, nparams] all_reduce
                         per bucket: stage 2 (amp_C.multi_tensor_lamb_update_weights) writes the fp32 master and the
                         model-dtype copy, then all_gather(in place) on a second stream/communicator
All arithmetic is in libmt_b200.so; there is no CPU path.

Differences from the reference, on purpose: (1) the learning rate is read from ``param_groups[0]['lr']`` every step
(the reference keeps ``self._lr`` at its initial 0.0 unless the caller pokes it, :132,744); (2) a parameter with several
fragments on one rank contributes the SUM of their update norms (the reference's ``scatter_`` keeps one of them,
:689); (3) ``dwu_group_size`` / ``dwu_num_*_pg`` describe a multi-node hierarchy and are accepted and ignored on one
node; (4) with ``clip_after_ar=False`` the overflow decision, which the reference takes per rank from its LOCAL gradient
norm (:697-703 with :626-632), is all-reduced so that every rank skips or applies the same step.

Options honoured as in the reference:
  overlap_reductions     post-backward hooks start a bucket's reduce-scatter (own stream) as soon as every gradient of
                         the bucket has been produced in a non-accumulation backward pass (:249-301, :783-797);
                         complete_reductions() reduces what is left and forms the norm
  clip_after_ar=False    the LOCAL gradients are clipped by their own norm and pre-divided by the world size before a SUM
                         reduce-scatter; stage 1 then does not clip (:629-646, :697)
  e5m2_allgather=True    stage 2 writes the new parameters as one e5m2 byte each (reference packing, :34-53 of
                         multi_tensor_distopt_lamb_kernel.cu), the all-gather moves half the bytes, and one converting
                         launch per bucket unpacks them into the model parameters (:270, :320, :850-866)
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from .flat_sharded_adam import FlatLayout, second_communicator


class DistributedFusedLAMB(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, bias_correction=True, grad_averaging=True, betas=(0.9, 0.999), eps=1e-8,
                 weight_decay=0., max_grad_norm=0., adam_w_mode=True, use_nvlamb=False,
                 step_supports_amp_scaling=True, overlap_reductions=True, dwu_group_size=0, dwu_num_blocks=4,
                 dwu_num_chunks=4, dwu_num_rs_pg=1, dwu_num_ar_pg=4, dwu_num_ag_pg=0, e5m2_allgather=False,
                 verbose=False, clip_after_ar=True, process_group=None, device=None):
        defaults = dict(lr=lr, bias_correction=bias_correction, betas=betas, eps=eps, weight_decay=weight_decay,
                        grad_averaging=grad_averaging, max_grad_norm=max_grad_norm)
        super().__init__(params, defaults)
        self._e5m2_allgather = bool(e5m2_allgather)
        self._clip_after_ar = bool(clip_after_ar)
        self._adam_w_mode = 1 if adam_w_mode else 0
        self._use_nvlamb = bool(use_nvlamb)
        self._step_supports_amp_scaling = step_supports_amp_scaling
        self._overlap = bool(overlap_reductions)
        self._is_accumulation_step = False
        self._last_step = False
        self._verbose = verbose
        self.group = process_group
        self.world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(self.group) if dist.is_initialized() else 0

        self._model_params: List[torch.nn.Parameter] = []
        self._group_props = []
        for g in self.param_groups:
            for p in g['params']:
                if not p.requires_grad:
                    continue
                self._model_params.append(p)
                self._group_props.append((g['weight_decay'], 1 if g['bias_correction'] else 0, g['betas'][0], g['betas'][1],
                                          g['eps']))
        assert self._model_params, "no parameters to optimize"
        dt = self._model_params[0].dtype
        assert all(p.dtype == dt for p in self._model_params), \
            "all parameters must share one dtype (they live in one flat buffer that NCCL reduces / gathers in place)"
        assert dt in (torch.float16, torch.bfloat16, torch.float32)
        self.model_dtype = dt
        self.device = torch.device(device) if device is not None else self._model_params[0].device
        numels = [p.numel() for p in self._model_params]
        esz = torch.empty((), dtype=dt).element_size()
        nb = max(1, int(dwu_num_blocks) * int(dwu_num_chunks))
        self.layout = FlatLayout(numels, self.world, self.rank, esz, max(1, -(-sum(numels) // nb)))
        self.frags = self.layout.fragments()
        self._alloc()
        self._build_launch_sets()
        self.rs_stream = None
        self._init_comm()
        self._adopt_params()
        self._bucket_reduced = [False] * self.layout.nbuckets
        self._rs_events = {}
        self._grads_generated = set()
        self._in_backward = False
        self._hooks = []
        if self._overlap and self._clip_after_ar and self.world > 1:
            self._register_hooks()

    # ---- buffers --------------------------------------------------------------------------------------------------
    def _alloc(self):
        L, dev = self.layout, self.device
        z = lambda n, dt: torch.zeros(n, dtype=dt, device=dev)
        self.grads_flat = z(L.padded, self.model_dtype)
        self.params_flat = z(L.padded, self.model_dtype)
        self._fp32_p, self._fp32_m = z(L.state_elems, torch.float32), z(L.state_elems, torch.float32)
        self._fp32_v, self._fp32_u = z(L.state_elems, torch.float32), z(L.state_elems, torch.float32)
        self._overflow_buf = z(1, torch.int32)
        self._step = z(1, torch.int32)
        self._lr = z(1, torch.float32)
        self._lr_host = None
        self._global_scale = None
        self._L2_grad_norm = None
        self._one = torch.ones(1, dtype=torch.float32, device=dev)
        # e5m2_allgather: what the all-gather moves is this byte buffer; params_flat is unpacked from it afterwards
        self._params_e5m2 = z(L.padded, torch.uint8) if self._e5m2_allgather else None

    def _build_launch_sets(self):
        """Tensor lists over ALL owned fragments (fragment order = flat order, so the fragments of one parameter are
        consecutive and `seg_begin` can describe them as runs), the per-fragment hyper-parameter arrays, and the per-bucket
        slices stage 2 is launched on."""
        fr, dev = self.frags, self.device
        g, p, m, v, u, c = [], [], [], [], [], []
        for f in fr:
            n = f.flat_hi - f.flat_lo
            s = slice(f.state_lo, f.state_lo + n)
            g.append(self.grads_flat[f.flat_lo:f.flat_hi])
            c.append((self._params_e5m2 if self._e5m2_allgather else self.params_flat)[f.flat_lo:f.flat_hi])
            p.append(self._fp32_p[s]); m.append(self._fp32_m[s]); v.append(self._fp32_v[s]); u.append(self._fp32_u[s])
        self._stage1_lists = [g, p, m, v, u]
        self._stage2_lists = [u, p, c]
        t = lambda vals, dt: torch.tensor(vals, dtype=dt, device=dev)
        props = [self._group_props[f.param] for f in fr]
        self._decay = t([q[0] for q in props], torch.float32)
        self._bias_correction = t([q[1] for q in props], torch.int32)
        self._beta1 = t([q[2] for q in props], torch.float32)
        self._beta2 = t([q[3] for q in props], torch.float32)
        self._beta3 = t([1.0 - q[2] for q in props], torch.float32)   # carried for the signature; the kernel recomputes it
        self._eps = t([q[4] for q in props], torch.float32)
        self._offsets = t([f.param for f in fr], torch.int64)          # fragment -> model parameter (reference :545)
        seg, seg_param = [], []
        for i, f in enumerate(fr):
            if i == 0 or f.param != fr[i - 1].param:
                seg.append(i)
                seg_param.append(f.param)
        self._seg_begin = t(seg + [len(fr)], torch.int32)
        self._seg_param = t(seg_param, torch.int64)
        assert len(set(seg_param)) == len(seg_param), "fragments of one parameter must be consecutive"
        self._need_norms = self._use_nvlamb or any(q[0] != 0 for q in props)
        # some rank needs the norms <=> every rank must join the all_reduce: decide from the GLOBAL property list
        self._any_rank_needs_norms = self._use_nvlamb or any(q[0] != 0 for q in self._group_props)
        self._bucket_frag_range = []
        i = 0
        for b in range(self.layout.nbuckets):
            j = i
            while j < len(fr) and fr[j].bucket == b:
                j += 1
            self._bucket_frag_range.append((i, j))
            i = j

    def _init_comm(self):
        self.rs_group = self.ag_group = self.group
        self.ag_stream = None
        if self.world > 1 and self._overlap and self.device.type == "cuda":
            self.ag_group = second_communicator(self.group)   # all-gathers overlap stage 2 of the next bucket
            self.ag_stream = torch.cuda.Stream(self.device)
            self.rs_stream = torch.cuda.Stream(self.device)               # reduce-scatters started from backward hooks

    def _adopt_params(self):
        """Model tensors become views of the flat buffers (parameters AND gradients), masters are initialised from them
        (reference :523-524)."""
        L = self.layout
        with torch.no_grad():
            for p, o in zip(self._model_params, L.offsets):
                n = p.numel()
                self.params_flat[o:o + n].copy_(p.detach().reshape(-1))
                p.data = self.params_flat[o:o + n].view(p.shape)
                p.grad = self.grads_flat[o:o + n].view(p.shape)
        for f in self.frags:
            n = f.flat_hi - f.flat_lo
            self._fp32_p[f.state_lo:f.state_lo + n].copy_(self.params_flat[f.flat_lo:f.flat_hi])

    # ---- collectives (overridable: the CPU tests swap in gloo-compatible ones) ---------------------------------------
    def _reduce_scatter(self, b: int):
        """Average the bucket over the ranks into the owned sub-shard, in place.  NCCL's AVG pre-multiplies by 1/world
        before summing -- the reference's `_flatten_grad_mt(1/world)` followed by a SUM reduce-scatter (:591-603)."""
        L = self.layout
        lo, hi = L.bucket_range(b)
        olo, ohi = L.owned_range(b)
        if self.world > 1:
            # clip_after_ar=False pre-divided the local gradients by the world size itself (:642-643)
            op = dist.ReduceOp.AVG if self._clip_after_ar else dist.ReduceOp.SUM
            dist.reduce_scatter_tensor(self.grads_flat[olo:ohi], self.grads_flat[lo:hi], op=op, group=self.rs_group)

    def _all_gather(self, b: int):
        L = self.layout
        lo, hi = L.bucket_range(b)
        olo, ohi = L.owned_range(b)
        buf = self._params_e5m2 if self._e5m2_allgather else self.params_flat
        if self.world > 1:
            dist.all_gather_into_tensor(buf[lo:hi], buf[olo:ohi], group=self.ag_group)
        if self._e5m2_allgather:
            self._unpack_e5m2(b)

    def _unpack_e5m2(self, b: int):
        """Gathered e5m2 bytes of bucket b -> model parameters (one converting launch; skipped on overflow)."""
        lo, hi = self.layout.bucket_range(b)
        self._amp_C().maybe_cast_mt(8192, self._overflow_buf, [[self._params_e5m2[lo:hi]], [self.params_flat[lo:hi]]])

    def _all_reduce(self, t, op=dist.ReduceOp.SUM):
        if self.world > 1:
            dist.all_reduce(t, op=op, group=self.group)

    # ---- compute (overridable for the CPU tests) -------------------------------------------------------------------------
    def _amp_C(self):
        from ... import load_amp_C

        return load_amp_C()

    def _shard_sq_norm(self) -> torch.Tensor:
        """Sum of squares of the reduced gradients in this rank's sub-shards (float32 [1], on the device)."""
        L = self.layout
        views = [self.grads_flat[slice(*L.owned_range(b))] for b in range(L.nbuckets)]
        flag = torch.zeros(1, dtype=torch.int32, device=self.device)   # overflow is decided from the norm itself (:702-703)
        n, _ = self._amp_C().multi_tensor_l2norm(65536, flag, [views], False)
        return n * n

    def _stage1(self, global_scale, global_grad_norm, max_grad_norm) -> Optional[torch.Tensor]:
        if not self.frags:
            return None
        # the dtype whose rounding of the master IS the model parameter: its norm is taken from the masters
        norm_dt = torch.uint8 if self._e5m2_allgather else self.model_dtype
        extra = (self._seg_begin, norm_dt) if self._any_rank_needs_norms else ()
        return self._amp_C().multi_tensor_lamb_compute_update_term(
            8192, self._overflow_buf, self._stage1_lists, self._beta1, self._beta2, self._beta3, self._bias_correction,
            self._step, self._eps, self._adam_w_mode, self._decay, global_scale, global_grad_norm, max_grad_norm, *extra)

    def _stage2(self, b: int, param_norm, update_norm, global_grad_norm):
        i, j = self._bucket_frag_range[b]
        if i == j:
            return
        self._amp_C().multi_tensor_lamb_update_weights(
            8192, self._overflow_buf, [l[i:j] for l in self._stage2_lists], param_norm[i:j], update_norm,
            self._offsets[i:j], self._lr, self._decay[i:j], global_grad_norm, self._use_nvlamb)

    # ---- the reference's public surface ----------------------------------------------------------------------------------
    def set_is_accumulation_step(self, is_accumulation_step):
        self._is_accumulation_step = is_accumulation_step

    def set_last_step(self, last_step):
        self._last_step = last_step

    def set_global_scale(self, global_scale):
        """Loss scale the gradients carry (a float or a 1-element tensor, e.g. GradScaler's)."""
        self._global_scale = global_scale

    @property
    def global_scale(self):
        return self._global_scale

    @property
    def L2_grad_norm(self):
        return self._L2_grad_norm

    def zero_grad(self, set_to_none: bool = False):
        # the gradients are views of the flat buffer: they are cleared, never dropped
        self.grads_flat.zero_()
        for p, o in zip(self._model_params, self.layout.offsets):
            if p.grad is None or p.grad.data_ptr() != self.grads_flat[o:o + 1].data_ptr():
                p.grad = self.grads_flat[o:o + p.numel()].view(p.shape)

    # ---- gradient reductions ---------------------------------------------------------------------------------------------
    def _register_hooks(self):
        """overlap_reductions: one post-accumulate hook per parameter (reference :249-301)."""
        L = self.layout
        B = L.bucket_elems
        sizes = [p.numel() for p in self._model_params]
        self._param_buckets = [list(range(o // B, (o + max(n, 1) - 1) // B + 1)) for o, n in zip(L.offsets, sizes)]
        self._bucket_params = [[] for _ in range(L.nbuckets)]
        for i, bs in enumerate(self._param_buckets):
            for b in bs:
                self._bucket_params[b].append(i)
        for i, p in enumerate(self._model_params):
            self._hooks.append(p.register_post_accumulate_grad_hook(lambda param, i=i: self._on_grad_ready(i)))

    def _on_grad_ready(self, i):
        """Reference `_do_overlapped_reduction` (:783-797): nothing happens on accumulation steps; otherwise a bucket
        whose gradients have all arrived in this backward pass is reduce-scattered at once, on its own stream."""
        if self._is_accumulation_step or self._last_step:
            return
        if not self._in_backward:
            self._in_backward = True
            self._grads_generated.clear()
            torch.autograd.Variable._execution_engine.queue_callback(self._on_backward_done)
        p = self._model_params[i]
        o = self.layout.offsets[i]
        if p.grad is None or p.grad.data_ptr() != self.grads_flat[o:o + 1].data_ptr():
            return          # the gradient was re-allocated by the caller: complete_reductions() will pick it up
        self._grads_generated.add(i)
        ready = [b for b in self._param_buckets[i] if not self._bucket_reduced[b]
                 and all(j in self._grads_generated for j in self._bucket_params[b])]
        if ready and self.rs_stream is None:        # no side stream (CPU stand-in): reduce in line
            for b in ready:
                self._reduce_scatter(b)
                self._bucket_reduced[b] = True
        elif ready:
            main = torch.cuda.current_stream(self.device)
            self.rs_stream.wait_stream(main)
            with torch.cuda.stream(self.rs_stream):
                for b in ready:
                    self._reduce_scatter(b)
                    e = torch.cuda.Event()
                    e.record(self.rs_stream)
                    self._rs_events[b] = e
                    self._bucket_reduced[b] = True

    def _on_backward_done(self):
        self._in_backward = False
        self._grads_generated.clear()

    def _adopt_stray_grads(self):
        """Gradients that are no longer views of the flat buffer (the caller replaced p.grad) are added back in."""
        for p, o in zip(self._model_params, self.layout.offsets):
            view = self.grads_flat[o:o + p.numel()]
            if p.grad is not None and p.grad.data_ptr() != view.data_ptr():
                view.add_(p.grad.detach().reshape(-1).to(view.dtype))
                p.grad = view.view(p.shape)

    def _clip_before_reduction(self):
        """clip_after_ar=False (reference :629-646): L2 norm of the LOCAL flat gradients; coefficient
        min(1, loss_scale * max_grad_norm / (1e-6 + norm)), replaced by 1 when it is not finite; the gradients are
        multiplied by round_fp16(coefficient) / world so that the SUM reduce-scatter yields clipped averages."""
        flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        norm = self._local_norm(flag)
        gs = self._device_scalar(1.0 if self._global_scale is None else self._global_scale, "_gs_cache")
        coeff = gs * float(self.defaults['max_grad_norm']) / (1e-6 + norm)
        coeff = torch.where(coeff > 1, torch.ones_like(coeff), coeff)
        coeff = torch.where(coeff + 1 > coeff, coeff, torch.ones_like(coeff))       # nan / inf -> 1 (:639-641)
        scale = (coeff.half() / self.world).to(self.grads_flat.dtype)
        self.grads_flat.mul_(scale)
        return norm

    def _local_norm(self, flag) -> torch.Tensor:
        n, _ = self._amp_C().multi_tensor_l2norm(65536, flag, [[self.grads_flat]], False)
        return n

    def complete_reductions(self):
        """Reduce-scatter every bucket that has not been reduced during backward and form the L2 gradient norm
        (reference :590-671, :813-834): the global norm of the averaged, still loss-scaled gradients -- or, with
        clip_after_ar=False, this rank's local norm, taken before the clipped gradients are reduced."""
        self._adopt_stray_grads()
        local_norm = None
        if not self._clip_after_ar:
            local_norm = self._clip_before_reduction()
        main = torch.cuda.current_stream(self.device) if self.device.type == "cuda" else None
        for e in self._rs_events.values():
            main.wait_event(e)
        self._rs_events.clear()
        for b in range(self.layout.nbuckets - 1, -1, -1):
            if not self._bucket_reduced[b]:
                self._reduce_scatter(b)
        self._bucket_reduced = [False] * self.layout.nbuckets
        if local_norm is not None:
            # every rank must take the same skip decision: share whether ANY local norm was non-finite
            bad = (~(local_norm + 1 > local_norm)).to(torch.float32)
            self._all_reduce(bad, op=dist.ReduceOp.MAX)
            self._L2_grad_norm = torch.where(bad > 0, torch.full_like(local_norm, float("inf")), local_norm)
            return
        sq = self._shard_sq_norm()
        self._all_reduce(sq)
        self._L2_grad_norm = sq.sqrt()

    def _device_scalar(self, x, cache_name):
        if isinstance(x, torch.Tensor):
            return x.detach().to(device=self.device, dtype=torch.float32).reshape(1)
        buf = getattr(self, cache_name, None)
        if buf is None or buf[0] != float(x):
            buf = (float(x), torch.full((1,), float(x), dtype=torch.float32, device=self.device))
            setattr(self, cache_name, buf)
        return buf[1]

    def _pipeline_step(self):
        gs = self._device_scalar(1.0 if self._global_scale is None else self._global_scale, "_gs_cache")
        # clipping before the reduction already happened in complete_reductions(): stage 1 must not clip again (:697)
        max_grad_norm = float(self.defaults['max_grad_norm']) if self._clip_after_ar else 0.0
        gnorm = self._L2_grad_norm
        assert gnorm is not None, "call complete_reductions() before step()"
        # overflow <=> the global norm is inf/nan; identical on every rank because the norm was all-reduced (:702-710)
        is_finite = (gnorm + 1 > gnorm).to(torch.int32)
        self._overflow_buf.copy_(1 - is_finite)
        self._step += is_finite                      # the step counter only advances on clean steps (:713)
        lr = float(self.param_groups[0]['lr'])
        if lr != self._lr_host:
            self._lr.fill_(lr)
            self._lr_host = lr
        seg_sq = self._stage1(gs, gnorm, max_grad_norm)
        nparams = len(self._model_params)
        if self._any_rank_needs_norms:
            full = torch.zeros(2, nparams, dtype=torch.float32, device=self.device)
            if seg_sq is not None:
                full.index_copy_(1, self._seg_param, seg_sq)
            self._all_reduce(full)
            full.sqrt_()
            param_norm, update_norm = full[0].index_select(0, self._offsets), full[1]
        else:
            param_norm = torch.zeros(max(1, len(self.frags)), dtype=torch.float32, device=self.device)
            update_norm = torch.zeros(max(1, nparams), dtype=torch.float32, device=self.device)
        self._param_norms, self._update_norms = param_norm, update_norm
        main = torch.cuda.current_stream(self.device) if self.ag_stream is not None else None
        for b in range(self.layout.nbuckets):
            self._stage2(b, param_norm, update_norm, gnorm)
            if self.ag_stream is None:
                self._all_gather(b)
            else:
                ev = torch.cuda.Event()
                ev.record(main)
                with torch.cuda.stream(self.ag_stream):
                    self.ag_stream.wait_event(ev)
                    self._all_gather(b)
        if self.ag_stream is not None:
            main.wait_stream(self.ag_stream)

    def step(self, closure=None, grad_scaler=None):
        loss = closure() if closure is not None else None
        self._pipeline_step()
        if grad_scaler is not None:   # reference :843-847
            found_inf = self._overflow_buf.float()
            optimizer_state = grad_scaler._per_optimizer_states[id(self)]
            optimizer_state["found_inf_per_device"][self.device] = found_inf
        return loss

    # ---- checkpoint (reference :872-910: this rank's shard) ------------------------------------------------------------
    def state_dict(self):
        return {'step': self._step, 'fp32_p': self._fp32_p, 'fp32_m': self._fp32_m, 'fp32_v': self._fp32_v}

    def load_state_dict(self, state_dict):
        self._step.copy_(state_dict['step'].to(self.device))
        for name in ('fp32_p', 'fp32_m', 'fp32_v'):
            getattr(self, '_' + name).copy_(state_dict[name].to(self.device))
        # the model-dtype copy of the owned fragments follows the masters; the next step's all-gather publishes it
        for f in self.frags:
            n = f.flat_hi - f.flat_lo
            self.params_flat[f.flat_lo:f.flat_hi].copy_(self._fp32_p[f.state_lo:f.state_lo + n])
