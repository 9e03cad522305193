"""World-size-2 tests of the sharded optimizer's HOST logic on CPU (gloo): layout, fragment ownership, and the
reduce-scatter -> shard update -> all-gather orchestration.  The shard update itself is CUDA-only in the product;
here it is replaced by the CPU oracle (test infrastructure) through the class's overridable hooks, so what is under
test is everything around the kernel."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import sys

from beforeholiday_b200.contrib.optimizers import FlatLayout

sys.path.insert(0, os.path.dirname(__file__))
from cpu_engine import CpuDistributedFusedAdam, CpuShardedAdam  # noqa: E402  (gloo + oracle stand-ins)

SIZES = [1000, 7, 4096, 33333, 1, 129]


def test_layout_partitions_flat_space_exactly():
    for world in (1, 2, 4, 8):
        lay0 = FlatLayout(SIZES, world, 0, elem_bytes=2, bucket_cap_elems=8192)
        assert lay0.padded % (world * 256) == 0 and lay0.padded >= lay0.total
        assert all(o % 64 == 0 for o in lay0.offsets)  # 128-byte aligned bf16 offsets
        covered = torch.zeros(lay0.padded, dtype=torch.int32)
        elems = 0
        for r in range(world):
            lay = FlatLayout(SIZES, world, r, elem_bytes=2, bucket_cap_elems=8192)
            for b in range(lay.nbuckets):
                lo, hi = lay.owned_range(b)
                covered[lo:hi] += 1
            for f in lay.fragments():
                n = f.flat_hi - f.flat_lo
                assert n > 0 and lay.offsets[f.param] + f.param_lo == f.flat_lo
                olo, ohi = lay.owned_range(f.bucket)
                assert olo <= f.flat_lo and f.flat_hi <= ohi
                assert f.state_lo == f.bucket * lay.sub_shard + (f.flat_lo - olo)
                elems += n
        assert bool((covered == 1).all())          # every flat element owned by exactly one rank
        assert elems == sum(SIZES)                 # every parameter element lands in exactly one fragment


def _worker(rank, world, port, outdir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        full = [torch.randn(n) * 0.02 for n in SIZES]
        wds = [0.1 if i % 2 else 0.0 for i in range(len(SIZES))]
        opt = CpuShardedAdam(SIZES, lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.0, device="cpu",
                             grad_dtype=torch.bfloat16, param_dtype=torch.bfloat16, bucket_cap_mb=8192 * 2 / 2 ** 20,
                             per_param_weight_decay=wds)
        L = opt.layout
        assert L.nbuckets > 1
        for p, o in zip(full, L.offsets):
            opt.params_flat[o:o + p.numel()].copy_(p)
        opt.init_master_from_params()
        for it in range(3):
            # every rank contributes its own gradient; the step averages them
            g = torch.Generator().manual_seed(100 * it + rank)
            for n, o in zip(SIZES, L.offsets):
                opt.grads_flat[o:o + n].copy_(torch.randn(n, generator=g) * 1e-2)
            opt.step()
        out = [opt.params_flat[o:o + n].clone() for n, o in zip(SIZES, L.offsets)]
        torch.save(out, os.path.join(outdir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(300)
def test_sharded_step_world2_matches_single_process(tmp_path):
    world = 2
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    results = {r: torch.load(tmp_path / f"rank{r}.pt") for r in range(world)}
    # every rank ends with the same full bf16 parameters
    for a, b in zip(results[0], results[1]):
        assert torch.equal(a.view(torch.int16), b.view(torch.int16))

    # single-process replay with the oracle: bf16 master init, summed bf16 grads / world
    from oracle import mt_oracle as O

    torch.manual_seed(0)
    full = [(torch.randn(n) * 0.02) for n in SIZES]
    p32 = [p.bfloat16().float() for p in full]
    m = [torch.zeros(n) for n in SIZES]
    v = [torch.zeros(n) for n in SIZES]
    pc = [torch.zeros(n, dtype=torch.bfloat16) for n in SIZES]
    wds = [0.1 if i % 2 else 0.0 for i in range(len(SIZES))]
    k = len(SIZES)
    for it in range(3):
        gs = []
        for n in SIZES:
            gs.append(torch.zeros(n))
        for r in range(world):
            g = torch.Generator().manual_seed(100 * it + r)
            for i, n in enumerate(SIZES):
                gs[i] += (torch.randn(n, generator=g) * 1e-2).bfloat16().float()
        gs = [x.bfloat16() for x in gs]
        O.multi_tensor_fused_adam(8192, torch.zeros(1, dtype=torch.int32), [p32, m, v, gs, pc], [0.9] * k, [0.95] * k,
                                  [1] * k, [1e-8] * k, wds, 1e-2, float(world), it + 1, 1)
    for a, b in zip(results[0], pc):
        assert torch.equal(a.view(torch.int16), b.view(torch.int16))


# ---------------------------------------------------------------------------------------------------------------------
# Sharded LAMB: host logic of DistributedFusedLAMB (fragment runs, per-parameter norm exchange, overflow/step handling)
LAMB_SIZES = [1000, 7, 40000, 33333, 1, 129]      # 40000 and 33333 straddle several sub-shards


def _lamb_cls():
    from beforeholiday_b200.contrib.optimizers import DistributedFusedLAMB

    class CpuShardedLamb(DistributedFusedLAMB):
        """gloo + oracle stand-ins for the NCCL collectives and the CUDA kernels."""

        def _init_comm(self):
            self.rs_group = self.ag_group = self.group
            self.ag_stream = None

        def _reduce_scatter(self, b):
            lo, hi = self.layout.bucket_range(b)
            buf = self.grads_flat[lo:hi].float() / (self.world if self._clip_after_ar else 1)   # AVG, or SUM of pre-divided
            dist.all_reduce(buf, group=self.group)
            olo, ohi = self.layout.owned_range(b)
            self.grads_flat[olo:ohi].copy_(buf[olo - lo:ohi - lo])

        def _all_gather(self, b):
            lo, hi = self.layout.bucket_range(b)
            olo, ohi = self.layout.owned_range(b)
            buf = self._params_e5m2 if self._e5m2_allgather else self.params_flat
            parts = [torch.empty(self.layout.sub_shard, dtype=torch.float32) for _ in range(self.world)]
            dist.all_gather(parts, buf[olo:ohi].float().contiguous(), group=self.group)
            buf[lo:hi].copy_(torch.cat(parts))
            if self._e5m2_allgather:
                self._unpack_e5m2(b)

        def _shard_sq_norm(self):
            L = self.layout
            return sum((self.grads_flat[slice(*L.owned_range(b))].double() ** 2).sum() for b in range(L.nbuckets)).float().reshape(1)

        def _stage1(self, gs, gnorm, max_grad_norm):
            from oracle import mt_oracle as O

            if not self.frags:
                return None
            sq = O.multi_tensor_lamb_compute_update_term(
                8192, self._overflow_buf, self._stage1_lists, self._beta1.tolist(), self._beta2.tolist(), None,
                self._bias_correction.tolist(), int(self._step), self._eps.tolist(), self._adam_w_mode,
                self._decay.tolist(), float(gs), float(gnorm), max_grad_norm,
                torch.uint8 if self._e5m2_allgather else self.model_dtype)
            seg = self._seg_begin.tolist()
            return torch.stack([torch.stack([sq[s, seg[k]:seg[k + 1]].sum() for k in range(len(seg) - 1)])
                                for s in range(2)]).float()

        def _local_norm(self, flag):
            return self.grads_flat.double().norm().float().reshape(1)

        def _unpack_e5m2(self, b):
            from oracle import mt_oracle as O

            if int(self._overflow_buf) == 0:
                lo, hi = self.layout.bucket_range(b)
                self.params_flat[lo:hi].copy_(O.convert(O.convert(self._params_e5m2[lo:hi], torch.float32), self.model_dtype))

        def _stage2(self, b, param_norm, update_norm, gnorm):
            from oracle import mt_oracle as O

            i, j = self._bucket_frag_range[b]
            if i < j:
                O.multi_tensor_lamb_update_weights(8192, self._overflow_buf, [l[i:j] for l in self._stage2_lists],
                                                   param_norm[i:j].tolist(), update_norm.tolist(), self._offsets[i:j].tolist(),
                                                   float(self._lr), self._decay[i:j].tolist(), float(gnorm), self._use_nvlamb)

    return CpuShardedLamb


LAMB_OPTIONS = {"default": {}, "e5m2_allgather": {"e5m2_allgather": True}, "clip_before_ar": {"clip_after_ar": False},
                "overlap_hooks": {"overlap_reductions": True}}


def _lamb_worker(rank, world, port, outdir, option="default"):
    import sys

    sys.path.insert(0, os.path.dirname(__file__))
    from lamb_replay import make_grad

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        params = [torch.nn.Parameter((torch.randn(n) * 0.05).bfloat16()) for n in LAMB_SIZES]
        groups = [dict(params=params[0::2], weight_decay=0.01), dict(params=params[1::2], weight_decay=0.0)]
        opt = _lamb_cls()(groups, lr=2e-3, betas=(0.9, 0.999), eps=1e-6, max_grad_norm=1.0, use_nvlamb=False,
                          dwu_num_blocks=2, dwu_num_chunks=2, device="cpu", **LAMB_OPTIONS[option])
        assert opt.layout.nbuckets > 1
        assert any(sum(1 for f in opt.frags if f.param == q) > 1 for q in range(len(LAMB_SIZES)))  # multi-fragment runs
        opt.set_global_scale(64.0)
        order = params[0::2] + params[1::2]      # the optimizer's flat order
        for it in range(4):
            gs = [make_grad(1000 * it + 10 * rank + i, p.numel(), torch.bfloat16, 64.0).view(p.shape)
                  for i, p in enumerate(order)]
            if it == 2 and rank == 1:
                gs[3].view(-1)[5] = float("inf")               # one rank overflows: every rank must skip the step
            if option == "overlap_hooks":
                # a real backward pass (d/dp sum(p * g) = g): the hooks reduce-scatter each bucket as it fills up
                opt.zero_grad()
                sum((p * g).sum() for p, g in zip(order, gs)).backward()
                assert any(opt._bucket_reduced)
            else:
                for p, g in zip(order, gs):
                    p.grad.copy_(g)
            opt.complete_reductions()
            opt.step()
            assert int(opt._overflow_buf) == (1 if it == 2 else 0)
        assert int(opt._step) == 3
        torch.save([p.detach().clone() for p in order], os.path.join(outdir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("world,option", [(2, "default"), (4, "default"), (2, "e5m2_allgather"), (2, "clip_before_ar"),
                                          (2, "overlap_hooks")])
def test_sharded_lamb_matches_single_process(tmp_path, world, option):
    import sys

    sys.path.insert(0, os.path.dirname(__file__))
    from lamb_replay import replay
    from oracle import mt_oracle as O

    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_lamb_worker, args=(r, world, port, str(tmp_path), option)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    res = {r: torch.load(tmp_path / f"rank{r}.pt") for r in range(world)}
    for r in range(1, world):
        for a, b in zip(res[0], res[r]):
            assert torch.equal(a.view(torch.int16), b.view(torch.int16))
    torch.manual_seed(0)
    init = [(torch.randn(n) * 0.05).bfloat16() for n in LAMB_SIZES]
    order = init[0::2] + init[1::2]
    sizes = [x.numel() for x in order]
    props = [(0.01, 1, 0.9, 0.999, 1e-6)] * len(init[0::2]) + [(0.0, 1, 0.9, 0.999, 1e-6)] * len(init[1::2])
    _, pc, step = replay(O, sizes, order, props, world, 4, torch.bfloat16, 2e-3, 64.0, 1.0, 1, False,
                         lambda it, r, i: 1000 * it + 10 * r + i, overflow_at=2, e5m2=(option == "e5m2_allgather"),
                         clip_after_ar=(option != "clip_before_ar"))
    assert step == 3
    for a, b in zip(res[0], pc):
        ia, ib = a.view(torch.int16).to(torch.int32), b.view(torch.int16).to(torch.int32)
        assert int((ia - ib).abs().max()) <= 1          # bf16 copies within 1 ulp (norm summation order)
        assert float((ia != ib).float().mean()) < 0.01


# ---------------------------------------------------------------------------------------------------------------------
# The drop-in use with a real module (the scenario of tests/test_gpu_dist.py::test_dist_adam_module_vs_ddp_adamw, modelled
# on the reference's apex/contrib/test/optimizers/test_dist_adam.py:84-156) on CPU: param groups with different weight
# decay, parameters and gradients re-pointed at views of the flat buffers, tiny buckets so that fragments straddle
# sub-shards -- against torch.optim.AdamW on the rank-averaged gradients.
def _module_cpu_worker(rank, world, port, outdir, clip=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        mk = lambda: torch.nn.Sequential(*[torch.nn.Linear(7, 7) for _ in range(11)])
        model, ref = mk(), mk()
        ref.load_state_dict(model.state_dict())
        decay = [p for n, p in model.named_parameters() if n.endswith("weight")]
        no_decay = [p for n, p in model.named_parameters() if n.endswith("bias")]
        opt = CpuDistributedFusedAdam([{"params": decay, "weight_decay": 0.1}, {"params": no_decay, "weight_decay": 0.0}],
                                      lr=1e-2, betas=(0.9, 0.95), eps=1e-8, bucket_cap_mb=71 * 4 / 2 ** 20, device="cpu")
        assert opt.layout.nbuckets > 1 and opt.grads_flat.dtype == torch.float32
        assert all(p.data_ptr() == opt.params_flat.data_ptr() + 4 * o for p, o in zip(decay + no_decay, opt.layout.offsets))
        rdecay = [p for n, p in ref.named_parameters() if n.endswith("weight")]
        rno = [p for n, p in ref.named_parameters() if n.endswith("bias")]
        ropt = torch.optim.AdamW([{"params": rdecay, "weight_decay": 0.1}, {"params": rno, "weight_decay": 0.0}],
                                 lr=1e-2, betas=(0.9, 0.95), eps=1e-8)
        for it in range(4):
            opt.zero_grad()
            ropt.zero_grad()
            g = torch.Generator().manual_seed(1000 * it + rank)
            x = torch.randn(16, 7, generator=g)
            model(x).pow(2).mean().backward()          # accumulates straight into the flat gradient buffer's views
            assert all(p.grad.data_ptr() == opt.grads_flat.data_ptr() + 4 * o
                       for p, o in zip(decay + no_decay, opt.layout.offsets))
            ref(x).pow(2).mean().backward()
            for p in ref.parameters():
                dist.all_reduce(p.grad)
                p.grad /= world
            if clip:
                rn = torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.05)
                n = opt.clip_grad_norm(0.05)
                opt.step()
                assert abs(float(n) - float(rn)) <= 1e-5 * float(rn) and float(rn) > 0.05   # it really clips
            else:
                opt.step()
            ropt.step()
        torch.save({"mine": [p.detach().clone() for p in model.parameters()],
                    "ref": [p.detach().clone() for p in ref.parameters()]}, os.path.join(outdir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
@pytest.mark.parametrize("clip", [False, True], ids=["plain", "clip_grad_norm"])
def test_sharded_adam_on_a_real_module_world2_vs_adamw(tmp_path, clip):
    world = 2
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_module_cpu_worker, args=(r, world, port, str(tmp_path), clip)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    res = [torch.load(tmp_path / f"rank{r}.pt") for r in range(world)]
    for a, b in zip(res[0]["mine"], res[1]["mine"]):
        assert torch.equal(a, b)
    for a, b in zip(res[0]["mine"], res[0]["ref"]):
        assert torch.allclose(a, b, rtol=2e-5, atol=2e-6)


def test_dropin_over_bare_sizes_matches_the_engine():
    """`numels=` (extension): the optimizer creates its parameters as views of the flat buffer.  One process, no process
    group: two steps on gradients written straight into the flat buffer must equal the engine's result bit for bit."""
    sizes = [1000, 7, 4096, 333]
    torch.manual_seed(3)
    init = torch.randn(sum(sizes)) * 0.02
    grads = [torch.randn(sum(sizes)) * 1e-2 for _ in range(2)]
    kw = dict(lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)

    def fill(flat, layout, src):
        o = 0
        for n, off in zip(sizes, layout.offsets):
            flat[off:off + n].copy_(src[o:o + n])
            o += n

    opt = CpuDistributedFusedAdam(numels=sizes, device="cpu", bucket_cap_mb=4096 * 4 / 2 ** 20, **kw)
    assert [p.numel() for g in opt.param_groups for p in g["params"]] == sizes
    assert all(p.data_ptr() == opt.params_flat.data_ptr() + 4 * o
               for p, o in zip((p for g in opt.param_groups for p in g["params"]), opt.layout.offsets))
    eng = CpuShardedAdam(sizes, device="cpu", grad_dtype=torch.float32, param_dtype=torch.float32,
                         bucket_cap_mb=4096 * 4 / 2 ** 20, **kw)
    for o in (opt, eng):
        fill(o.params_flat, o.layout, init)
    opt._init_masters()
    eng.init_master_from_params()
    for g in grads:
        opt.zero_grad()
        eng.zero_grad()
        fill(opt.grads_flat, opt.layout, g)
        fill(eng.grads_flat, eng.layout, g)
        opt.step()
        eng.step()
    assert torch.equal(opt.params_flat, eng.params_flat) and not torch.equal(opt.params_flat[:1000], init[:1000])
    assert opt.state_dict(gather_on_root=False)["state"]["step"] == 2


# ---------------------------------------------------------------------------------------------------------------------
# second_communicator: the extra communicator the sharded optimizers open for their all-gathers
def _second_comm_worker(rank, world, port, outdir):
    from beforeholiday_b200.contrib.optimizers.flat_sharded_adam import second_communicator

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # whole job
        g = second_communicator(None, backend="gloo")
        t = torch.tensor([float(rank + 1)])
        dist.all_reduce(t, group=g)
        assert float(t) == world * (world + 1) / 2
        # state sharded over disjoint halves of the job (the distributed x redundant grid): each rank asks for a second
        # communicator over ITS half; every rank must take part in creating both
        halves = [dist.new_group([0, 1]), dist.new_group([2, 3])]
        mine = halves[rank // 2]
        g2 = second_communicator(mine, backend="gloo")
        assert dist.get_process_group_ranks(g2) == dist.get_process_group_ranks(mine)
        t = torch.tensor([float(rank + 1)])
        dist.all_reduce(t, group=g2)
        assert float(t) == (3.0 if rank < 2 else 7.0)
        # both communicators of a half are usable side by side
        a, b = torch.ones(1), torch.ones(1)
        w1 = dist.all_reduce(a, group=mine, async_op=True)
        w2 = dist.all_reduce(b, group=g2, async_op=True)
        w1.wait()
        w2.wait()
        assert float(a) == float(b) == 2.0
        open(os.path.join(outdir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_second_communicator_over_sub_groups_world4(tmp_path):
    world = 4
    ctx = mp.get_context("spawn")
    port = _free_port()
    procs = [ctx.Process(target=_second_comm_worker, args=(r, world, port, str(tmp_path))) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(240)
        assert p.exitcode == 0
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
