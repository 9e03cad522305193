"""The reference's own test scenarios for DistributedFusedAdam
(/root/reference/apex/contrib/test/optimizers/test_dist_adam.py), restated as plain functions that run inside one rank
of an initialised process group -- on GPUs over NCCL (tests/test_gpu_dist_adam.py) and on CPU over gloo with the oracle
standing in for the kernels (tests/test_dist_adam_cpu.py).  `make` builds the optimizer under test (the product class on
GPU, the gloo/oracle-engined subclass on CPU); the reference side is torch.optim.AdamW on gradients averaged over ranks
(what its DDP-wrapped reference model computes, :41-49).

Each case returns a dict of plain results; a failed expectation raises AssertionError inside the rank.
"""
import contextlib
import io

import torch
import torch.distributed as dist


class SimpleModel(torch.nn.Module):          # reference :10-24
    def __init__(self, num_layers, size):
        super().__init__()
        self.layers = torch.nn.ModuleList([torch.nn.Linear(size, size, bias=(i % 3 == 0)) for i in range(num_layers)])

    def forward(self, x):
        y = 0
        for i, l in enumerate(self.layers):
            y = y + (i + 1) * l(x)
        return y


def _allreduce_avg(params):
    w = dist.get_world_size()
    for p in params:
        if p.grad is not None:
            g = p.grad.float()
            dist.all_reduce(g)
            p.grad.copy_(g / w)


def make_models(make, num_layers, size, dtype=torch.float32, param_sync_dtype=None, device="cpu",
                overlap_communication=True, **extra):
    """reference :26-74 -- same hyper-parameters, param groups with different lr."""
    ref_model = SimpleModel(num_layers, size).to(dtype=dtype, device=device)
    dist_model = SimpleModel(num_layers, size).to(dtype=dtype, device=device)
    with torch.no_grad():
        for rp, dp in zip(ref_model.parameters(), dist_model.parameters()):
            dp.copy_(rp)
        for rp in ref_model.parameters():        # DDP broadcasts rank 0's parameters at construction
            buf = rp.float()
            dist.broadcast(buf, src=0)
            rp.copy_(buf)
    optim_args = dict(lr=0.1, betas=(0.1, 0.2), eps=0.25, weight_decay=0.1)
    rps = list(ref_model.parameters())
    ref_optim = torch.optim.AdamW([{'params': rps[1::2], 'lr': 0.2}, {'params': rps[0::2]}], **optim_args)
    dps = list(dist_model.parameters())
    dist_optim = make([{'params': dps[1::2], 'lr': 0.2}, {'params': dps[0::2]}],
                      overlap_grad_sync=overlap_communication, bucket_cap_mb=71 / (4 * 1024 * 1024),
                      dtype=torch.float32, param_sync_dtype=param_sync_dtype, device=device, **optim_args, **extra)
    return ref_model, ref_optim, dist_model, dist_optim


def matches_pytorch(make, rank, device, num_layers=11, layer_size=7, batch_size=3, num_steps=3, micro_batch_steps=3,
                    overlap_communication=True, use_nosync=True, dtype=torch.float32, param_sync_dtype=None,
                    rtol=None, atol=None, **extra):
    """reference :84-156"""
    torch.manual_seed(1234 + rank)
    ref_model, ref_optim, dist_model, dist_optim = make_models(make, num_layers, layer_size, dtype, param_sync_dtype,
                                                               device, overlap_communication, **extra)
    # (sub-shards are 256-element aligned: the reference's 11 x Linear(7, 7) spans several buckets up to world 3)
    assert dist_optim.layout.nbuckets > 1 or dist.get_world_size() > 3
    for step in range(num_steps):
        ref_optim.zero_grad()
        dist_optim.zero_grad()
        for micro_step in range(micro_batch_steps):
            x = (torch.rand(batch_size, layer_size) - 0.5).to(dtype=dtype, device=device)
            dy = (torch.rand(batch_size, layer_size) - 0.5).to(dtype=dtype, device=device)
            x_ref = x.detach().clone().requires_grad_(True)
            y_ref = ref_model(x_ref)
            y_ref.backward(dy)
            x_dist = x.detach().clone().requires_grad_(True)
            y_dist = dist_model(x_dist)
            ctx = dist_optim.no_sync if (use_nosync and micro_step < micro_batch_steps - 1) else contextlib.nullcontext
            with ctx():
                y_dist.backward(dy)
            torch.testing.assert_close(y_dist, y_ref, rtol=rtol, atol=atol)
            torch.testing.assert_close(x_dist.grad, x_ref.grad, rtol=rtol, atol=atol)
        _allreduce_avg(ref_model.parameters())
        ref_optim.step()
        dist_optim.step()
        for rp, dp in zip(ref_model.parameters(), dist_model.parameters()):
            torch.testing.assert_close(dp, rp, rtol=rtol, atol=atol)
    return {"params": [p.detach().float().cpu() for p in dist_model.parameters()]}


def raises_on_mismatch(make, rank, device):
    """reference :183-211: a step on one side only must make the models differ (the comparison is not vacuous)."""
    torch.manual_seed(1234 + rank)
    ref_model, ref_optim, dist_model, dist_optim = make_models(make, 11, 7, device=device)
    dist_optim.zero_grad()
    x = (torch.rand(3, 7) + 0.5).to(device)
    dy = (torch.rand(3, 7) + 0.5).to(device)
    dist_model(x).backward(dy)
    dist_optim.step()
    for rp, dp in zip(ref_model.parameters(), dist_model.parameters()):
        try:
            torch.testing.assert_close(dp, rp)
        except AssertionError:
            continue
        raise AssertionError("parameters still match after a one-sided step")
    return {}


def clip_grad_norm(make, rank, device):
    """reference :213-246"""
    torch.manual_seed(1234 + rank)
    ref_model, ref_optim, dist_model, dist_optim = make_models(make, 1, 1, device=device)
    norms = []
    for x, dy in zip([3, 1, 4, 1, 5, 9], [1, -1, 1, -1, 1, -1]):
        x = torch.tensor([[float(x)]], device=device)
        dy = torch.tensor([[float(dy)]], device=device)
        ref_optim.zero_grad()
        ref_model(x.detach()).backward(dy.detach())
        _allreduce_avg(ref_model.parameters())
        ref_grad_norm = torch.nn.utils.clip_grad_norm_(ref_model.parameters(), 3.5)
        ref_optim.step()
        dist_optim.zero_grad()
        dist_model(x.detach()).backward(dy.detach())
        dist_grad_norm = dist_optim.clip_grad_norm(3.5)
        dist_optim.step()
        torch.testing.assert_close(dist_grad_norm, ref_grad_norm)
        for rp, dp in zip(ref_model.parameters(), dist_model.parameters()):
            torch.testing.assert_close(dp, rp)
        norms.append(float(dist_grad_norm))
    return {"norms": norms}


def grad_scaler(make, rank, device):
    """reference :248-287: inf / nan gradients make the scaler skip the step and back off; both sides stay equal."""
    torch.manual_seed(1234 + rank)
    ref_model, ref_optim, dist_model, dist_optim = make_models(make, 1, 1, device=device)
    args = dict(init_scale=3.21, growth_factor=1.23, backoff_factor=0.876, growth_interval=1)
    dev_type = torch.device(device).type
    ref_scaler = torch.amp.GradScaler(dev_type, **args)
    dist_scaler = torch.amp.GradScaler(dev_type, **args)
    scales = []
    for x, dy in zip([3, 1, 4, 1, 5, 9], [1, float('inf'), 1, 1, float('nan'), -1]):
        x = torch.tensor([[float(x)]], device=device)
        dy = torch.tensor([[float(dy)]], device=device)
        ref_optim.zero_grad()
        ref_scaler.scale(ref_model(x.detach())).backward(dy.detach())
        _allreduce_avg(ref_model.parameters())
        ref_scaler.step(ref_optim)
        ref_scaler.update()
        dist_optim.zero_grad()
        dist_scaler.scale(dist_model(x.detach())).backward(dy.detach())
        dist_scaler.step(dist_optim)
        dist_scaler.update()
        for rp, dp in zip(ref_model.parameters(), dist_model.parameters()):
            torch.testing.assert_close(dp, rp)
        assert float(dist_scaler.get_scale()) == float(ref_scaler.get_scale())
        scales.append(float(dist_scaler.get_scale()))
    # two of the six steps were skipped: bias correction must have counted four
    assert dist_optim.state_dict(gather_on_root=False)['state']['step'] == 4
    return {"scales": scales}


def checkpoint(make, rank, device):
    """reference :289-380: save on the root rank, load everywhere, keep training in lockstep."""
    torch.manual_seed(1234 + rank)
    num_layers, layer_size = 5, 2
    _, _, model_save, optim_save = make_models(make, num_layers, layer_size, device=device)
    _, _, model_load, optim_load = make_models(make, num_layers, layer_size, device=device)
    for step in range(3):
        optim_save.zero_grad()
        for micro_step in range(2):
            x = (torch.rand(4, layer_size) - 0.5).to(device)
            dy = (torch.rand(4, layer_size) - 0.5).to(device)
            model_save(x).backward(dy)
        optim_save.step()
    differ = False
    for ps, pl in zip(model_save.parameters(), model_load.parameters()):
        differ = differ or not torch.allclose(ps, pl)
    assert differ
    state_dict = {'model': model_save.state_dict(), 'optim': optim_save.state_dict()}
    if rank == 0:
        buf = io.BytesIO()
        torch.save(state_dict, buf)
        blob = [buf.getvalue()]
    else:
        assert state_dict['optim'] is None
        blob = [None]
    dist.broadcast_object_list(blob, src=0)
    state_dict = torch.load(io.BytesIO(blob[0]), map_location=device)
    model_load.load_state_dict(state_dict['model'])
    optim_load.load_state_dict(state_dict['optim'])
    for ps, pl in zip(model_save.parameters(), model_load.parameters()):
        torch.testing.assert_close(pl, ps)
    for step in range(3):
        optim_save.zero_grad()
        optim_load.zero_grad()
        for micro_step in range(3):
            x = (torch.rand(5, layer_size) - 0.5).to(device)
            dy = (torch.rand(5, layer_size) - 0.5).to(device)
            x_save = x.detach().clone().requires_grad_(True)
            y_save = model_save(x_save)
            y_save.backward(dy)
            x_load = x.detach().clone().requires_grad_(True)
            y_load = model_load(x_load)
            y_load.backward(dy)
            torch.testing.assert_close(y_load, y_save)
            torch.testing.assert_close(x_load.grad, x_save.grad)
        optim_save.step()
        optim_load.step()
        for ps, pl in zip(model_save.parameters(), model_load.parameters()):
            torch.testing.assert_close(pl, ps)
    return {}


def overflow_does_not_stick(make, rank, device):
    """Round-1 defect (ADVICE, high): the skip flag raised by one overflowing step was never cleared, so every later
    step was silently skipped.  One step with an inf gradient under a grad scaler, then a clean one that must move
    the parameters."""
    torch.manual_seed(1234 + rank)
    _, _, model, optim = make_models(make, 3, 4, device=device)
    scaler = torch.amp.GradScaler(torch.device(device).type, init_scale=4.0, growth_interval=100)
    before = [p.detach().clone() for p in model.parameters()]
    x = torch.ones(2, 4, device=device)
    optim.zero_grad()
    scaler.scale(model(x)).backward(torch.full((2, 4), float('inf'), device=device))
    scaler.step(optim)
    scaler.update()
    for b, p in zip(before, model.parameters()):
        assert torch.equal(b, p.detach())                      # skipped
    optim.zero_grad()
    scaler.scale(model(x)).backward(torch.ones(2, 4, device=device))
    scaler.step(optim)
    scaler.update()
    moved = [not torch.equal(b, p.detach()) for b, p in zip(before, model.parameters())]
    assert all(moved), moved
    # and without any scaler
    optim.zero_grad()
    model(x).backward(torch.ones(2, 4, device=device))
    snap = [p.detach().clone() for p in model.parameters()]
    optim.step()
    assert all(not torch.equal(s, p.detach()) for s, p in zip(snap, model.parameters()))
    return {}


def grid_2x2(make, rank, device):
    """A distributed x redundant process-group grid (reference constructor :62-74, :232-262; gradient path :717-732):
    world 4 as two replicas of a state sharded over two ranks.  Every rank must end with the same parameters as
    AdamW on the gradients averaged over all four.  (Other world sizes: nothing to do.)"""
    if dist.get_world_size() != 4:
        return {}
    shard_groups = [dist.new_group([0, 1]), dist.new_group([2, 3])]
    replica_groups = [dist.new_group([0, 2]), dist.new_group([1, 3])]
    out = matches_pytorch(make, rank, device, distributed_process_group=shard_groups[rank // 2],
                          redundant_process_group=replica_groups[rank % 2])
    # and the grouping is checked, as in the reference
    try:
        make([torch.nn.Parameter(torch.zeros(4, device=device))], distributed_process_group=shard_groups[rank // 2],
             device=device)
    except RuntimeError as e:
        assert "Invalid process group configuration" in str(e)
    else:
        raise AssertionError("a 2-rank shard group without replicas inside a 4-rank process group must be refused")
    return out


def frozen_param_and_lr_schedule(make, rank, device):
    """Not in the reference's file, but what its class does as a torch.optim.Optimizer: a parameter that never gets a
    gradient (requires_grad=False) must not block its bucket's reduce-scatter; like the reference -- which creates state
    for every parameter of the groups (distributed_fused_adam.py:395-405) and updates every fragment of every bucket
    (:977-1115) -- it is stepped with a ZERO gradient, i.e. it only decays (torch.optim.AdamW would skip it; the
    comparison below hands AdamW explicit zero gradients).  A scheduler that rewrites param_groups[i]['lr'] between steps
    must be honoured by the next step (the learning rates live in device memory here), and
    zero_grad(set_to_none=False) must keep accumulating correctly."""
    torch.manual_seed(4321 + rank)
    ref_model = SimpleModel(7, 5).to(device)
    dist_model = SimpleModel(7, 5).to(device)
    with torch.no_grad():
        for rp in ref_model.parameters():
            buf = rp.float()
            dist.broadcast(buf, src=0)
            rp.copy_(buf)
        for rp, dp in zip(ref_model.parameters(), dist_model.parameters()):
            dp.copy_(rp)
    for m in (ref_model, dist_model):
        m.layers[2].weight.requires_grad_(False)
        m.layers[3].bias.requires_grad_(False)
    args = dict(lr=0.05, betas=(0.8, 0.9), eps=1e-3, weight_decay=0.05)
    rps, dps = list(ref_model.parameters()), list(dist_model.parameters())
    ref_optim = torch.optim.AdamW([{'params': rps[:5], 'lr': 0.1}, {'params': rps[5:]}], **args)
    dist_optim = make([{'params': dps[:5], 'lr': 0.1}, {'params': dps[5:]}], bucket_cap_mb=71 / (4 * 1024 * 1024),
                      device=device, **args)
    ref_frozen = [ref_model.layers[2].weight, ref_model.layers[3].bias]
    scheds = [torch.optim.lr_scheduler.StepLR(o, step_size=1, gamma=0.5) for o in (ref_optim, dist_optim)]
    for step in range(4):
        ref_optim.zero_grad(set_to_none=False)
        dist_optim.zero_grad(set_to_none=False)
        for micro in range(2):
            x = (torch.rand(3, 5) - 0.5).to(device)
            dy = (torch.rand(3, 5) - 0.5).to(device)
            ref_model(x).backward(dy)
            dist_model(x).backward(dy)
        _allreduce_avg(ref_model.parameters())
        for p in ref_frozen:
            p.grad = torch.zeros_like(p)
        ref_optim.step()
        dist_optim.step()
        for sch in scheds:
            sch.step()
        assert dist_optim.param_groups[0]['lr'] == ref_optim.param_groups[0]['lr'] == 0.1 * 0.5 ** (step + 1)
        for rp, dp in zip(ref_model.parameters(), dist_model.parameters()):
            torch.testing.assert_close(dp, rp)
    assert dist_model.layers[2].weight.grad is None and dist_model.layers[3].bias.grad is None
    return {}


CASES = {
    "matches_pytorch": matches_pytorch,
    "matches_pytorch_no_overlap": lambda make, rank, device, **kw: matches_pytorch(
        make, rank, device, overlap_communication=False, use_nosync=False, **kw),
    "matches_pytorch_sync_every_step": lambda make, rank, device, **kw: matches_pytorch(
        make, rank, device, use_nosync=False, **kw),
    "matches_pytorch_fp64": lambda make, rank, device, **kw: matches_pytorch(
        make, rank, device, dtype=torch.float64, rtol=1.3e-6, atol=1e-5, **kw),
    "matches_pytorch_fp16": lambda make, rank, device, **kw: matches_pytorch(
        make, rank, device, dtype=torch.float16, rtol=1e-2, atol=1e-2, **kw),
    "matches_pytorch_allgather_fp16": lambda make, rank, device, **kw: matches_pytorch(
        make, rank, device, dtype=torch.float32, param_sync_dtype=torch.float16, rtol=1e-2, atol=1e-2, **kw),
    "raises_on_mismatch": raises_on_mismatch,
    "clip_grad_norm": clip_grad_norm,
    "grad_scaler": grad_scaler,
    "checkpoint": checkpoint,
    "overflow_does_not_stick": overflow_does_not_stick,
    "grid_2x2": grid_2x2,
    "frozen_param_and_lr_schedule": frozen_param_and_lr_schedule,
}
