"""GPU parity for the rows SURVEY.md section 8(f) marks "next": multi_tensor_{l2norm_mp, lamb_mp, adagrad, novograd,
lars}, FusedMixedPrecisionLamb / FusedAdagrad / FusedNovoGrad / FusedLARS and contrib.clip_grad -- against the CPU
oracle (bit-exact where no norm is involved), the reference's CUDA kernels (oracle/_ref) and torch."""
import pytest
import torch

from test_gpu_ops import BF16, F16, F32, F64, make_lists, run_both, to_gpu
from util import ulp_diff_16, bits_equal, cpu_clone, max_rel, n_mismatch, rel_close

pytestmark = pytest.mark.gpu

SIZES = [0, 1, 9, 4097, 70001, 300000]


# ------------------------------------------------------------------------------------------------ adagrad
@pytest.mark.parametrize("dt", [F32, F16, BF16, F64])
@pytest.mark.parametrize("mode,wd", [(0, 0.0), (0, 0.1), (1, 0.01)])
def test_adagrad(amp_C, oracle, dt, mode, wd):
    lists = make_lists([dt] * 3, SIZES, seed=21, scale=0.1)
    lists[2] = [t.abs() for t in lists[2]]
    run_both(amp_C, oracle, "multi_tensor_adagrad", lists, (1e-2, 1e-10, mode, wd))


def test_fused_adagrad_vs_torch():
    from beforeholiday_b200.optimizers import FusedAdagrad

    torch.manual_seed(9876)
    sizes = [[4096, 1024], [4096], [4096, 2048], [32320, 1024], [1]]  # test_fused_optimizer.py:122 (live for Adagrad)
    ref_p = [torch.nn.Parameter(torch.rand(s, device="cuda")) for s in sizes]
    tst_p = [torch.nn.Parameter(p.detach().clone()) for p in ref_p]
    ref, tst = torch.optim.Adagrad(ref_p, lr=5e-4, eps=1e-8, weight_decay=1e-3), FusedAdagrad(tst_p, lr=5e-4, eps=1e-8, weight_decay=1e-3)
    for _ in range(7):
        for a, b in zip(ref_p, tst_p):
            a.grad = torch.rand_like(a)
            b.grad = a.grad.clone()
        ref.step()
        tst.step()
    for a, b in zip(ref_p, tst_p):
        assert rel_close(b, a, 1e-6, 2e-7)


# ------------------------------------------------------------------------------------------------ novograd
@pytest.mark.parametrize("dt", [F32, F16, BF16])
@pytest.mark.parametrize("norm_type", [0, 2])
@pytest.mark.parametrize("mode", [0, 1])
def test_novograd(amp_C, oracle, dt, norm_type, mode):
    sizes = [1, 9, 4097, 70001, 300000]
    lists = make_lists([dt] * 3, sizes, seed=22, scale=0.1)
    lists_gpu, lists_cpu = to_gpu(lists), cpu_clone(lists)
    norms_cpu = torch.tensor([0.5, 0.0, 1.5, 2.0, 0.25])
    norms_gpu = norms_cpu.clone().cuda()
    fg = torch.zeros(1, dtype=torch.int32, device="cuda")
    fc = torch.zeros(1, dtype=torch.int32)
    args = (1e-3, 0.95, 0.98, 1e-8, 3, 1, 0.01, 1, mode, norm_type)
    amp_C.multi_tensor_novograd(8192, fg, lists_gpu, norms_gpu, *args)
    oracle.multi_tensor_novograd(8192, fc, lists_cpu, norms_cpu, *args)
    assert fg.item() == fc.item() == 0
    assert max_rel(norms_gpu, norms_cpu) < 1e-5          # blended per-tensor norms (L-inf is exact)
    if norm_type == 0:
        assert bits_equal(norms_gpu, norms_cpu)
    # the update divides by the blended norm: bit-exact for L-inf, 1e-5-relative in the norm otherwise
    for d in (1, 2):
        for a, b in zip(lists_gpu[d], lists_cpu[d]):
            if norm_type == 0:
                assert bits_equal(a, b), f"list {d}: {n_mismatch(a, b)}"
            elif dt == F32:
                assert torch.allclose(a.cpu(), b, rtol=2e-6, atol=1e-9)      # the divisor (blended norm) moved by <= 1e-7
            else:
                # a 1e-7 relative change of the divisor can flip a 16-bit rounding, never more: <= 1 ulp, and rarely
                assert ulp_diff_16(a, b) <= 1, f"list {d}: {ulp_diff_16(a, b)} ulp"
                assert n_mismatch(a, b) <= max(2, a.numel() // 200)


def test_fused_novograd_runs_and_tracks_norms():
    from beforeholiday_b200.optimizers import FusedNovoGrad

    torch.manual_seed(1)
    ps = [torch.nn.Parameter(torch.rand(n, device="cuda")) for n in (5000, 33)]
    opt = FusedNovoGrad(ps, lr=1e-2, betas=(0.95, 0.98), weight_decay=1e-3, norm_type=2)
    gs = [torch.rand_like(p) for p in ps]
    for p, g in zip(ps, gs):
        p.grad = g.clone()
    before = [p.detach().clone() for p in ps]
    opt.step()
    v = opt.param_groups[0]["exp_avg_sq"][1]
    # first step: exp_avg_sq initialised with ||g||, the blend sqrt(b2*n^2 + (1-b2)*n^2) leaves it unchanged
    assert torch.allclose(v, torch.stack([g.norm() for g in gs]), rtol=1e-5)
    # step 1 with bias correction: m = (1-b1)*(g/(v/bc2+eps) + wd*p), bc2 = sqrt(1-b2); p -= lr * m/(1-b1)
    bc2 = (1 - 0.98) ** 0.5
    for p, p0, g, n in zip(ps, before, gs, v):
        upd = g / (n / bc2 + 1e-8) + 1e-3 * p0
        assert torch.allclose(p, p0 - 1e-2 * upd, rtol=1e-5, atol=1e-7)


# ------------------------------------------------------------------------------------------------ lars
@pytest.mark.parametrize("g_t,w_t,copy", [(F16, F16, False), (F32, F32, False), (F16, F32, True), (F32, F32, True),
                                         (BF16, BF16, False), (BF16, F32, True)])
@pytest.mark.parametrize("nesterov,is_skipped", [(False, False), (True, False), (False, True)])
def test_lars(amp_C, oracle, g_t, w_t, copy, nesterov, is_skipped):
    sizes = [1, 9, 4097, 70001]
    dts = [g_t, w_t, w_t] + ([F16] if copy else [])
    lists = make_lists(dts, sizes, seed=23, scale=0.5)
    gn, pn = [0.7, 0.0, 3.0, 11.0], [1.0, 2.0, 0.0, 9.0]
    lists_gpu, lists_cpu = to_gpu(lists), cpu_clone(lists)
    fg = torch.zeros(1, dtype=torch.int32, device="cuda")
    fc = torch.zeros(1, dtype=torch.int32)
    args = (0.1, 0.001, 1e-8, 1e-4, 0.9, 0.0, nesterov, False, False, 1.0, is_skipped)
    amp_C.multi_tensor_lars(8192, fg, lists_gpu, torch.tensor(gn, device="cuda"), torch.tensor(pn, device="cuda"), *args)
    oracle.multi_tensor_lars(8192, fc, lists_cpu, gn, pn, *args)
    for d in range(1, len(dts)):
        for a, b in zip(lists_gpu[d], lists_cpu[d]):
            assert bits_equal(a, b), f"list {d}: {n_mismatch(a, b)}"
    # flag set -> nothing happens
    before = [t.clone() for t in lists_gpu[1]]
    amp_C.multi_tensor_lars(8192, fg + 1, lists_gpu, torch.tensor(gn, device="cuda"), torch.tensor(pn, device="cuda"), *args)
    assert all(bits_equal(a, b) for a, b in zip(lists_gpu[1], before))


def test_fused_lars_step():
    from beforeholiday_b200.optimizers import FusedLARS

    torch.manual_seed(2)
    p = torch.nn.Parameter(torch.rand(70001, device="cuda"))
    p0 = p.detach().clone()
    opt = FusedLARS([p], lr=0.1, momentum=0.9, weight_decay=1e-4, trust_coefficient=0.001)
    g = torch.rand_like(p)
    p.grad = g.clone()
    opt.step()
    trust = 0.001 * p0.norm() / (g.norm() + p0.norm() * 1e-4)
    mom = -(0.1 * trust) * (g + 1e-4 * p0)
    assert torch.allclose(p, p0 + mom, rtol=1e-5, atol=1e-7)
    assert torch.allclose(opt.state[p]["momentum_buffer"], mom, rtol=1e-4, atol=1e-9)


# ------------------------------------------------------------------------------------------------ l2norm_mp / lamb_mp
def test_l2norm_mp(amp_C, oracle):
    xs = [torch.randn(n, device="cuda") for n in (70001, 9)]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    n1, pt1 = amp_C.multi_tensor_l2norm_mp(65536, f, [xs], True)
    n2, pt2 = amp_C.multi_tensor_l2norm(65536, f, [xs], True)
    assert bits_equal(n1, n2) and bits_equal(pt1, pt2) and f.item() == 0
    # flag already set: the kernels return at once and the outputs are left untouched
    f.fill_(1)
    launches = amp_C.launch_count()
    n3, pt3 = amp_C.multi_tensor_l2norm_mp(65536, f, [xs], True)
    assert amp_C.launch_count() - launches == 1 and f.item() == 1      # one launch: the finalize rides in it
    # ... and the completion counters were handed back at zero: the next call on the same lists is right again
    f.zero_()
    n4, pt4 = amp_C.multi_tensor_l2norm_mp(65536, f, [xs], True)
    assert bits_equal(n4, n2) and bits_equal(pt4, pt2) and f.item() == 0


@pytest.mark.parametrize("g_t,masters", [(F32, False), (F16, False), (F16, True), (BF16, True)])
@pytest.mark.parametrize("mode,wd,use_nvlamb", [(1, 0.01, False), (1, 0.0, True), (0, 0.01, True), (1, 0.0, False)])
def test_lamb_mp(amp_C, oracle, g_t, masters, mode, wd, use_nvlamb):
    sizes = [1, 9, 4097, 70001, 300000]
    p_t = F32 if masters else g_t
    dts = [g_t, p_t, p_t, p_t] + ([g_t] if masters else [])
    lists = make_lists(dts, sizes, seed=24, scale=0.1)
    lists[0] = [t * 64 for t in lists[0]]          # loss-scaled gradients
    lists[3] = [t.abs() for t in lists[3]]
    lists_gpu, lists_cpu = to_gpu(lists), cpu_clone(lists)
    gnorm = torch.cat([g.double().reshape(-1) for g in lists[0]]).norm().float().reshape(1)
    dev = lambda t: t.cuda()
    lr, step = torch.tensor(5e-4), torch.tensor([4], dtype=torch.int32)
    maxn, found_inf, inv_scale = torch.tensor([64.0]), torch.zeros(1), torch.tensor([1.0 / 64])
    fg = torch.zeros(1, dtype=torch.int32, device="cuda")
    fc = torch.zeros(1, dtype=torch.int32)
    amp_C.multi_tensor_lamb_mp(8192, fg, lists_gpu, dev(lr), 0.9, 0.999, 1e-6, dev(step), 1, wd, 1, mode, dev(gnorm),
                               dev(maxn), use_nvlamb, dev(found_inf), dev(inv_scale))
    oracle.multi_tensor_lamb_mp(8192, fc, lists_cpu, lr, 0.9, 0.999, 1e-6, step, 1, wd, 1, mode, gnorm, maxn, use_nvlamb,
                                found_inf, inv_scale)
    for d in (0, 2, 3):   # u, m, v: no norm involved -> bit-exact
        for a, b in zip(lists_gpu[d], lists_cpu[d]):
            assert bits_equal(a, b), f"list {d}: {n_mismatch(a, b)}"
    for d in [1] + ([4] if masters else []):
        for a, b in zip(lists_gpu[d], lists_cpu[d]):
            if a.dtype == F32:
                assert torch.allclose(a.cpu(), b, rtol=1e-6, atol=1e-7)
            else:
                from util import ulp_diff_16
                assert ulp_diff_16(a, b) <= 1
    # flag set: nothing is touched
    before = [t.clone() for t in lists_gpu[1]]
    amp_C.multi_tensor_lamb_mp(8192, fg + 1, lists_gpu, dev(lr), 0.9, 0.999, 1e-6, dev(step), 1, wd, 1, mode, dev(gnorm),
                               dev(maxn), use_nvlamb, dev(found_inf), dev(inv_scale))
    assert all(bits_equal(a, b) for a, b in zip(lists_gpu[1], before))


def test_fused_mixed_precision_lamb_matches_fused_lamb_and_captures():
    """fp32, no scaler: FusedMixedPrecisionLamb must walk the same trajectory as FusedLAMB (same math, device-side
    scalars), and its step must be capturable in a CUDA graph after a warm-up step."""
    from beforeholiday_b200.optimizers import FusedLAMB, FusedMixedPrecisionLamb

    torch.manual_seed(3)
    sizes = [70001, 333, 4096]
    a = [torch.nn.Parameter(torch.rand(n, device="cuda")) for n in sizes]
    b = [torch.nn.Parameter(p.detach().clone()) for p in a]
    oa = FusedLAMB(a, lr=5e-4, weight_decay=0.01, max_grad_norm=1.0)
    ob = FusedMixedPrecisionLamb(b, lr=5e-4, weight_decay=0.01, max_grad_norm=1.0)
    grads = [[torch.rand(n, device="cuda") for n in sizes] for _ in range(4)]
    static_g = [torch.zeros(n, device="cuda") for n in sizes]
    for p, g in zip(b, static_g):
        p.grad = g
    for it in range(2):   # eager
        for p, g in zip(a, grads[it]):
            p.grad = g.clone()
        for s, g in zip(static_g, grads[it]):
            s.copy_(g)
        oa.step()
        ob.step()
    for x, y in zip(a, b):
        assert torch.allclose(x, y, rtol=1e-6, atol=1e-7)
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for s, g in zip(static_g, grads[2]):
            s.copy_(g)
        with torch.cuda.graph(graph, stream=side):
            ob.step()
    torch.cuda.current_stream().wait_stream(side)
    # capture does not run the step; replay it for iterations 2 and 3
    for it in (2, 3):
        for p, g in zip(a, grads[it]):
            p.grad = g.clone()
        for s, g in zip(static_g, grads[it]):
            s.copy_(g)
        oa.step()
        graph.replay()
    torch.cuda.synchronize()
    assert int(ob.param_groups[0]["step"].item()) == 4
    for x, y in zip(a, b):
        assert torch.allclose(x, y, rtol=2e-6, atol=2e-7)


# ------------------------------------------------------------------------------------------------ clip_grad
@pytest.mark.parametrize("dt", [F32, F16, BF16])
def test_clip_grad_norm(dt):
    from beforeholiday_b200.contrib.clip_grad import clip_grad_norm_

    torch.manual_seed(4)
    ps = [torch.nn.Parameter(torch.rand(n, device="cuda").to(dt)) for n in (70001, 9, 4096)]
    qs = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    for p, q in zip(ps, qs):
        p.grad = torch.rand_like(p)
        q.grad = p.grad.clone()
    n1 = clip_grad_norm_(ps, max_norm=1.0)
    n2 = torch.nn.utils.clip_grad_norm_(qs, max_norm=1.0)
    # torch returns the norm in the gradients' dtype (bf16: 8 bits); this returns fp32
    assert torch.allclose(n1.float().cpu(), n2.float().cpu(), rtol={F32: 1e-5, F16: 2e-3, BF16: 1e-2}[dt])
    for p, q in zip(ps, qs):
        assert torch.allclose(p.grad.float(), q.grad.float(), rtol=1e-2 if dt != F32 else 1e-5, atol=1e-3 if dt != F32 else 1e-7)


# ------------------------------------------------------------------------------------------------ vs reference CUDA
def test_more_ops_vs_reference_cuda(amp_C):
    from test_gpu_vs_reference import _load_ref, check_close, clone, flag, rnd

    ref = _load_ref("amp_C")
    sizes = [9, 4097, 70001, 278011]
    # adagrad
    lists = [[rnd(n, F32, 10 * d + i, 0.1) for i, n in enumerate(sizes)] for d in range(3)]
    lists[2] = [t.abs() for t in lists[2]]
    l1, l2 = clone(lists), clone(lists)
    amp_C.multi_tensor_adagrad(8192, flag(), l1, 1e-2, 1e-10, 0, 0.1)
    ref.multi_tensor_adagrad(8192, flag(), l2, 1e-2, 1e-10, 0, 0.1)
    for d in (1, 2):
        for a, b in zip(l1[d], l2[d]):
            check_close(a, b, f"adagrad {d}")
    # novograd, both norm types
    for norm_type in (0, 2):
        l1, l2 = clone(lists), clone(lists)
        n1 = torch.tensor([0.5, 1.0, 2.0, 0.1], device="cuda")
        n2 = n1.clone()
        args = (1e-3, 0.95, 0.98, 1e-8, 3, 1, 0.01, 1, 1, norm_type)
        amp_C.multi_tensor_novograd(8192, flag(), l1, n1, *args)
        ref.multi_tensor_novograd(8192, flag(), l2, n2, *args)
        assert max_rel(n1, n2) < 1e-5
        for d in (1, 2):
            for a, b in zip(l1[d], l2[d]):
                assert rel_close(a, b, 2e-5, 1e-6 * float(b.abs().max())), f"novograd {d}"
    # lars
    l1, l2 = clone(lists), clone(lists)
    gn = torch.tensor([0.7, 1.0, 3.0, 11.0], device="cuda")
    pn = torch.tensor([1.0, 2.0, 0.5, 9.0], device="cuda")
    args = (0.1, 0.001, 1e-8, 1e-4, 0.9, 0.0, True, False, False, 1.0, False)
    amp_C.multi_tensor_lars(8192, flag(), l1, gn, pn, *args)
    ref.multi_tensor_lars(8192, flag(), l2, gn, pn, *args)
    for d in (1, 2):
        for a, b in zip(l1[d], l2[d]):
            check_close(a, b, f"lars {d}")
    # lamb_mp (fp16 grads, fp32 masters, fp16 model copy)
    dts = [F16, F32, F32, F32, F16]
    lists = [[rnd(n, dt, 10 * d + i, 0.1) for i, n in enumerate(sizes)] for d, dt in enumerate(dts)]
    lists[3] = [t.abs() for t in lists[3]]
    l1, l2 = clone(lists), clone(lists)
    t = lambda v, dt=F32: torch.tensor(v, dtype=dt, device="cuda")
    gnorm, _ = ref.multi_tensor_l2norm_mp(65536, flag(), [lists[0]], False)
    args = (t(5e-4), 0.9, 0.999, 1e-6, t([3], torch.int32), 1, 0.01, 1, 1, gnorm, t([1.0]), False, t([0.0]), t([1.0]))
    amp_C.multi_tensor_lamb_mp(8192, flag(), l1, *args)
    ref.multi_tensor_lamb_mp(8192, flag(), l2, *args)
    for d in (0, 2, 3, 4):
        for a, b in zip(l1[d], l2[d]):
            check_close(a, b, f"lamb_mp {d}")
    # p -= ratio * u with u STORED IN FP16 by stage 1: wherever the two implementations' u differ by the permitted
    # 1 fp16 ulp (2^-10 relative), that element's step differs by 1e-3 of itself -- bound p by a tenth of a percent of
    # the largest step instead of 1e-6 of p.
    for a, b, p0 in zip(l1[1], l2[1], lists[1]):
        step_max = float((b - p0).abs().max())
        assert float((a - b).abs().max()) <= 2e-3 * step_max + 1e-7 * float(b.abs().max())
