"""Optimizer-level GPU tests, modelled on the reference's tests/L0/run_optimizers/*.py and tests/L0/run_amp/*.py:
the drop-in classes against torch.optim (the reference tests' own oracle) with the reference's seeds, sizes and
hyper-parameters, at the reference's tolerance AND the north-star one (1e-6 relative)."""
import copy

import pytest
import torch

from util import bits_equal, rel_close

pytestmark = pytest.mark.gpu


def gen_param_and_grads(sizes, dtype, device="cuda"):
    params = [torch.nn.Parameter(torch.rand(s, dtype=dtype, device=device)) for s in sizes]
    return params


def run_pair(ref_opt_ctor, tst_opt_ctor, sizes, iters=7, dtype=torch.float32, max_abs=1e-3, rtol=1e-6, atol=2e-7):
    torch.manual_seed(9876)  # test_fused_optimizer.py:17
    ref_params = gen_param_and_grads(sizes, dtype)
    tst_params = [torch.nn.Parameter(p.detach().clone()) for p in ref_params]
    ref_opt, tst_opt = ref_opt_ctor(ref_params), tst_opt_ctor(tst_params)
    for _ in range(iters):
        for rp, tp in zip(ref_params, tst_params):
            rp.grad = torch.rand_like(rp)
            tp.grad = rp.grad.clone()
        ref_opt.step()
        tst_opt.step()
    for rp, tp in zip(ref_params, tst_params):
        assert (rp - tp).abs().max().item() <= max_abs
        if rtol is not None:
            assert rel_close(tp, rp, rtol, atol)
    return ref_opt, tst_opt


NELEM = 278011  # test_fused_optimizer.py:65 (odd: unaligned tail)
MULTI = [[4096, 1024], [4096], [4096, 2048], [32320, 1024], [1]]  # :122


@pytest.mark.parametrize("adam_w_mode,wd", [(False, 0.0), (False, 0.1), (True, 0.01)])
@pytest.mark.parametrize("sizes", [[NELEM], MULTI, [1]])
def test_fused_adam_vs_torch(adam_w_mode, wd, sizes):
    from beforeholiday_b200.optimizers import FusedAdam

    opts = dict(lr=5e-4, betas=(0.9, 0.999), eps=1e-08, weight_decay=wd)
    ref = (torch.optim.AdamW if adam_w_mode else torch.optim.Adam)
    run_pair(lambda ps: ref(ps, amsgrad=False, **opts), lambda ps: FusedAdam(ps, adam_w_mode=adam_w_mode, **opts), sizes)


def test_fused_adam_bf16_and_fp16_buckets():
    from beforeholiday_b200.optimizers import FusedAdam

    torch.manual_seed(1)
    ps = [torch.nn.Parameter(torch.rand(5000, device="cuda").to(dt)) for dt in (torch.float16, torch.bfloat16, torch.float32)]
    refs = [torch.nn.Parameter(p.detach().float().clone()) for p in ps]
    opt = FusedAdam(ps, lr=1e-2, adam_w_mode=False)
    ref = torch.optim.Adam(refs, lr=1e-2)
    for p, r in zip(ps, refs):
        p.grad = torch.rand_like(p)
        r.grad = p.grad.float()
    opt.step()
    ref.step()
    for p, r in zip(ps, refs):
        assert torch.allclose(p.float(), r, rtol=1e-2 if p.dtype != torch.float32 else 1e-6, atol=1e-2 if p.dtype != torch.float32 else 1e-7)
    assert set(opt.state[ps[0]].keys()) == {"exp_avg", "exp_avg_sq"} and opt.param_groups[0]["step"] == 1


def test_fused_adam_state_dict_roundtrip():
    from beforeholiday_b200.optimizers import FusedAdam

    torch.manual_seed(2)
    p = torch.nn.Parameter(torch.rand(1000, device="cuda"))
    opt = FusedAdam([p], lr=1e-3)
    p.grad = torch.rand_like(p)
    opt.step()
    sd = copy.deepcopy(opt.state_dict())
    q = torch.nn.Parameter(p.detach().clone())
    opt2 = FusedAdam([q], lr=1e-3)
    opt2.load_state_dict(sd)
    g = torch.rand_like(p)
    p.grad, q.grad = g.clone(), g.clone()
    opt.step()
    opt2.step()
    assert bits_equal(p, q) and opt2.param_groups[0]["step"] == 2


def test_fused_adam_multi_device_message():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from beforeholiday_b200.optimizers import FusedAdam

    # test_fused_optimizer.py:115-120: current device != tensor device must still work
    p = torch.nn.Parameter(torch.rand(1000, device="cuda:1"))
    with torch.cuda.device(0):
        opt = FusedAdam([p], lr=1e-3)
        p.grad = torch.rand_like(p)
        opt.step()
    torch.cuda.synchronize(1)


@pytest.mark.parametrize("kw", [dict(lr=0.25, momentum=0.125), dict(lr=0.1, momentum=0.9, nesterov=True, weight_decay=1e-4),
                                dict(lr=0.1, momentum=0.0, weight_decay=1e-4)])
@pytest.mark.parametrize("sizes", [[NELEM], MULTI])
def test_fused_sgd_vs_torch(kw, sizes):
    from beforeholiday_b200.optimizers import FusedSGD

    # atol: torch's CUDA SGD contracts g + wd*p and buf*mu + g into FMAs; the two trajectories differ by rounding
    # RELATIVE TO THE MOMENTUM BUFFER (|buf| ~ 5-10 x |grad| at momentum 0.9), i.e. ~1e-6 absolute on O(1) params.
    run_pair(lambda ps: torch.optim.SGD(ps, **kw), lambda ps: FusedSGD(ps, **kw), sizes, atol=2e-6)


def test_fused_sgd_half():
    from beforeholiday_b200.optimizers import FusedSGD

    kw = dict(lr=0.25, momentum=0.125)  # test_fused_optimizer.py:278-289
    run_pair(lambda ps: torch.optim.SGD(ps, **kw), lambda ps: FusedSGD(ps, **kw), [NELEM], dtype=torch.float16,
             max_abs=1e-2, rtol=None)


def test_fused_sgd_channels_last():
    from beforeholiday_b200.optimizers import FusedSGD

    # test_fused_optimizer_channels_last.py:77-89
    torch.manual_seed(3)
    w = torch.rand(3, 4, 2, 3, device="cuda").contiguous(memory_format=torch.channels_last)
    p = torch.nn.Parameter(w.clone().contiguous(memory_format=torch.channels_last))
    r = torch.nn.Parameter(w.clone())
    opt, ref = FusedSGD([p], lr=0.25, momentum=0.125), torch.optim.SGD([r], lr=0.25, momentum=0.125)
    for _ in range(3):
        g = torch.rand(3, 4, 2, 3, device="cuda")  # contiguous grad
        p.grad, r.grad = g.clone(), g.clone()
        opt.step()
        ref.step()
    assert torch.allclose(p, r, rtol=1e-6, atol=1e-7)


class TorchLamb:
    """RefLAMB of the reference's test (tests/L0/run_optimizers/test_lamb.py:93-143), on float64 for a clean oracle."""

    def __init__(self, params, lr, betas, eps, wd):
        self.ps = [p.detach().double().clone() for p in params]
        self.m = [torch.zeros_like(p) for p in self.ps]
        self.v = [torch.zeros_like(p) for p in self.ps]
        self.lr, self.betas, self.eps, self.wd, self.t = lr, betas, eps, wd, 0

    def step(self, grads):
        self.t += 1
        grads = [g.double() for g in grads]
        gn = torch.sqrt(sum((g * g).sum() for g in grads))
        clip = 1.0 / max(gn.item(), 1.0)
        b1, b2 = self.betas
        for p, m, v, g in zip(self.ps, self.m, self.v, grads):
            g = g * clip
            m.mul_(b1).add_(g, alpha=1 - b1)
            v.mul_(b2).addcmul_(g, g, value=1 - b2)
            upd = (m / (1 - b1 ** self.t)) / ((v / (1 - b2 ** self.t)).sqrt() + self.eps)
            if self.wd != 0:
                upd = upd + self.wd * p
            wn, un = p.norm(), upd.norm()
            ratio = (wn / un).item() if wn > 0 and un > 0 else 1.0
            p.add_(upd, alpha=-self.lr * ratio)


@pytest.mark.parametrize("wd", [0.0, 0.01])
@pytest.mark.parametrize("sizes", [[NELEM], MULTI])
def test_fused_lamb_vs_reflamb(wd, sizes):
    from beforeholiday_b200.optimizers import FusedLAMB

    torch.manual_seed(9876)
    ps = [torch.nn.Parameter(torch.rand(s, device="cuda")) for s in sizes]
    ref = TorchLamb(ps, 5e-4, (0.9, 0.999), 1e-6, wd)
    opt = FusedLAMB(ps, lr=5e-4, betas=(0.9, 0.999), eps=1e-6, weight_decay=wd, use_nvlamb=True)
    for _ in range(7):
        gs = [torch.rand_like(p) for p in ps]
        for p, g in zip(ps, gs):
            p.grad = g.clone()
        ref.step(gs)
        opt.step()
    for p, r in zip(ps, ref.ps):
        assert (p.double() - r).abs().max().item() <= 1e-3           # reference tolerance (test_lamb.py:149)
        assert rel_close(p, r.float(), 2e-6, 2e-7)


def test_loss_scaler_dynamic_policy_and_skip():
    """amp dynamic loss scaling around a step: unscale + overflow check, skip + halve on overflow,
    double after `scale_window` clean steps; skipped steps leave params and moments untouched."""
    from beforeholiday_b200.amp import LossScaler
    from beforeholiday_b200.optimizers import FusedAdam

    torch.manual_seed(4)
    masters = [torch.nn.Parameter(torch.rand(n, device="cuda")) for n in (70001, 333)]
    opt = FusedAdam(masters, lr=1e-3)
    scaler = LossScaler("dynamic", scale_window=3)
    assert scaler.loss_scale() == 2.0 ** 16
    history = []
    for it in range(8):
        scale = scaler.loss_scale()
        model_grads = [(torch.rand_like(p) * 1e-3 * scale).half() for p in masters]
        if it in (1, 6):
            model_grads[it % 2][-1 if it == 1 else 0] = float("inf") if it == 1 else float("nan")
        master_grads = [torch.empty_like(p) for p in masters]
        scaler.clear_overflow_state()
        scaler.unscale(model_grads, master_grads, scale)
        before = [p.detach().clone() for p in masters]
        skip = scaler.update_scale()
        if not skip:
            for p, g in zip(masters, master_grads):
                p.grad = g
            opt.step()
            ref = [(mg.float() * (1.0 / scale)) for mg in model_grads]
            assert all(bits_equal(a, b) for a, b in zip(master_grads, ref))
        else:
            assert all(bits_equal(p, b) for p, b in zip(masters, before))
        history.append((skip, scaler.loss_scale()))
    S = 2.0 ** 16
    assert history == [(False, S), (True, S / 2), (False, S / 2), (False, S / 2), (False, S), (False, S), (True, S / 2), (False, S / 2)]
    sd = scaler.state_dict()
    s2 = LossScaler("dynamic")
    s2.load_state_dict(sd)
    assert s2.loss_scale() == scaler.loss_scale()


def test_loss_scaler_unscale_with_stashed():
    from beforeholiday_b200.amp import LossScaler

    torch.manual_seed(5)
    scaler = LossScaler("dynamic")
    model = [torch.randn(5001, device="cuda").half()]
    stashed = [torch.randn(5001, device="cuda")]
    master = [torch.empty(5001, device="cuda")]
    scaler.clear_overflow_state()
    scaler.unscale_with_stashed(model, stashed, master)
    a = 1.0 / scaler.loss_scale()
    assert bits_equal(master[0], model[0].float() * a + stashed[0] * 1.0)
    stashed[0][5] = float("inf")  # only arg 0 (incoming grads) is checked
    scaler.unscale_with_stashed(model, stashed, master)
    assert scaler.update_scale() is False
    model[0][5] = float("inf")
    scaler.clear_overflow_state()
    scaler.unscale_with_stashed(model, stashed, master)
    assert scaler.update_scale() is True


def test_fused_adam_master_weights_and_capturable():
    """The 28 B/param route through the optimizer class: bf16 params + fp32 masters, loss-scaled grads, found_inf skip;
    and the capturable variant replayed from a CUDA graph, both against torch.optim.AdamW on fp32 masters."""
    from beforeholiday_b200.optimizers import FusedAdam

    torch.manual_seed(6)
    sizes = [70001, 333, 4096]
    base = [torch.rand(n, device="cuda").bfloat16() for n in sizes]
    ref_p = [torch.nn.Parameter(b.float().clone()) for b in base]
    ref = torch.optim.AdamW(ref_p, lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    kw = dict(lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, master_weights=True)
    a = [torch.nn.Parameter(b.clone()) for b in base]
    b_ = [torch.nn.Parameter(b.clone()) for b in base]
    oa, ob = FusedAdam(a, **kw), FusedAdam(b_, capturable=True, **kw)
    static_g = [torch.zeros(n, device="cuda", dtype=torch.bfloat16) for n in sizes]
    for p, g in zip(b_, static_g):
        p.grad = g
    inv_scale = torch.tensor([1.0 / 1024], device="cuda")
    found = torch.zeros(1, device="cuda")
    graph = None
    for it in range(5):
        grads = [(torch.rand(n, device="cuda") * 1e-2) for n in sizes]
        overflow = it == 2
        found.fill_(1.0 if overflow else 0.0)
        for p, g in zip(a, grads):
            p.grad = (g * 1024).bfloat16()
        for s, g in zip(static_g, grads):
            s.copy_((g * 1024).bfloat16())
        if not overflow:
            for p, g in zip(ref_p, grads):
                p.grad = (g * 1024).bfloat16().float() / 1024
            ref.step()
        oa.step(inv_scale=inv_scale, found_inf=found)
        if it == 1:      # capture after one eager warm-up step
            graph = torch.cuda.CUDAGraph()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                with torch.cuda.graph(graph, stream=side):
                    ob.step(inv_scale=inv_scale, found_inf=found)
            torch.cuda.current_stream().wait_stream(side)
        if graph is None:
            ob.step(inv_scale=inv_scale, found_inf=found)
        else:
            graph.replay()
    torch.cuda.synchronize()
    assert int(ob.param_groups[0]["step"].item()) == 4          # the overflow step was not counted on the device
    for r, x, y in zip(ref_p, a, b_):
        mx, my = oa.state[x]["master_param"], ob.state[y]["master_param"]
        assert bits_equal(mx, my) and bits_equal(x, y)           # eager == graph replay
        assert bits_equal(x.detach(), mx.bfloat16())             # model copy is the rounded master
        assert rel_close(mx, r, 2e-6, 2e-7)


# ------------------------------------------------------------------------------------------------ steady-state fast path
def _mk(sizes, dtype=torch.float32, seed=3):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(s, device="cuda", generator=g).to(dtype) * 0.1) for s in sizes]


def _set_grads(params, it, skip=()):
    for i, p in enumerate(params):
        if i in skip:
            p.grad = None
            continue
        g = torch.Generator(device="cuda").manual_seed(1000 * it + i)
        p.grad = (torch.randn(p.shape, device="cuda", generator=g) * 1e-2).to(p.dtype)


@pytest.mark.parametrize("which", ["sgd", "adam", "lamb"])
def test_fast_path_is_bit_identical_to_generic_path_and_falls_back(which):
    """The cached-list / grads_of route must produce the same bits as the generic per-step list building, take over after
    the first step, and hand a step back to the generic route when a gradient is missing (then resume)."""
    from beforeholiday_b200.optimizers import FusedAdam, FusedLAMB, FusedSGD

    sizes = [(33, 17), (1000,), (4097,), (64, 3, 3, 3), (1,)]
    ctor = {"sgd": lambda ps: FusedSGD(ps, lr=0.1, momentum=0.9, dampening=0.1, weight_decay=1e-4),
            "adam": lambda ps: FusedAdam(ps, lr=1e-3, weight_decay=0.01),
            "lamb": lambda ps: FusedLAMB(ps, lr=1e-3, weight_decay=0.01, max_grad_norm=0.05)}[which]
    dts = [torch.float32, torch.float16, torch.float32, torch.float32, torch.float16]
    pa = [torch.nn.Parameter(p.detach().clone()) for p in (_mk([s], dt)[0] for s, dt in zip(sizes, dts))]
    pb = [torch.nn.Parameter(p.detach().clone()) for p in pa]
    # channels_last parameter: its gradient arrives contiguous -> layout differs -> that step uses the generic path
    fast, slow = ctor(pa), ctor(pb)
    slow._fast.enabled = False
    amp_C = __import__("beforeholiday_b200").load_amp_C()
    for it in range(6):
        skip = (2,) if it == 3 else ()
        _set_grads(pa, it, skip)
        _set_grads(pb, it, skip)
        fast.step()
        slow.step()
        for a, b in zip(pa, pb):
            assert bits_equal(a, b), f"{which}: step {it}"
    assert fast._fast._cache and not slow._fast._cache
    for a, b in zip(pa, pb):
        sa, sb = fast.state[a], slow.state[b]
        for k in sa:
            if torch.is_tensor(sa[k]):
                assert bits_equal(sa[k], sb[k]), k
    # checkpoint round trip drops the cached lists (the state tensors are replaced)
    fast.load_state_dict(fast.state_dict())
    assert not fast._fast._cache
    _set_grads(pa, 9)
    _set_grads(pb, 9)
    fast.step()
    slow.step()
    for a, b in zip(pa, pb):
        assert bits_equal(a, b)


def test_grads_of_rejects_what_it_cannot_take(amp_C):
    p = [torch.nn.Parameter(torch.zeros(8, device="cuda")), torch.nn.Parameter(torch.zeros(8, device="cuda"))]
    m = [torch.zeros(8, device="cuda") for _ in p]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    p[0].grad = torch.ones(8, device="cuda")
    with pytest.raises(RuntimeError, match="grads_of: a parameter has no gradient"):
        amp_C.multi_tensor_sgd(8192, f, [p, m], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0, grads_of=0)
    assert float(p[0].detach().abs().sum()) == 0.0   # nothing was launched
    p[1].grad = torch.ones(16, device="cuda")[::2]
    with pytest.raises(RuntimeError, match="grads_of: gradient layout differs"):
        amp_C.multi_tensor_sgd(8192, f, [p, m], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0, grads_of=0)
    p[1].grad = torch.full((8,), 2.0, device="cuda")
    amp_C.multi_tensor_sgd(8192, f, [p, m], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0, grads_of=0)
    assert torch.equal(p[0].detach(), torch.full((8,), -0.1, device="cuda")) and torch.equal(m[1], p[1].grad)
    n, _ = amp_C.multi_tensor_l2norm(65536, f, [p], False, grads_of=0)
    assert abs(float(n) - (8 * 1 + 8 * 4) ** 0.5) < 1e-6


def test_grads_of_on_non_leaf_tensors_does_not_touch_python_warnings(amp_C):
    """Non-leaf tensors have no .grad and Tensor::grad() would WARN through Python -- from a worker thread that would
    need the GIL the caller holds.  The extension must read the gradient slot silently, also on the threaded walk."""
    base = torch.zeros(40000 * 4, device="cuda", requires_grad=True)
    views = list((base * 1.0).split(4))            # 40 000 non-leaf tensors: the threaded walk (>= 32768 entries)
    moms = [torch.zeros(4, device="cuda") for _ in views]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    with pytest.raises(RuntimeError, match="grads_of: a parameter has no gradient"):
        amp_C.multi_tensor_sgd(8192, f, [views, moms], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0, grads_of=0)
