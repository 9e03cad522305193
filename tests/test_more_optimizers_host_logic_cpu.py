"""Python drivers of FusedNovoGrad / FusedAdagrad / FusedLARS on CPU, with the CUDA ops replaced by oracle stand-ins
(test infrastructure): driver + restatement together must reproduce what the reference's own tests pair the fused
optimizers with -- the pure-torch Novograd of tests/L0/run_optimizers/test_fused_novograd.py (golden fixture),
torch.optim.Adagrad -- and, for LARS, SGD-with-momentum scaled by the layer-wise ratio."""
from pathlib import Path

import pytest
import torch

GOLD_NOVO = torch.load(Path(__file__).parent / "golden" / "ref_novograd.pt", weights_only=False)


def _detached(lists):
    return [[t.detach() for t in l] for l in lists]


@pytest.mark.parametrize("case", sorted(GOLD_NOVO["cases"]))
def test_fused_novograd_driver_reproduces_reference_python(oracle, case):
    from beforeholiday_b200.optimizers import FusedNovoGrad

    gold = GOLD_NOVO["cases"][case]
    o = gold["opts"]
    gen = torch.Generator()
    params = [torch.nn.Parameter(torch.randn(n, generator=gen.manual_seed(300 + i)) * 0.1) for i, n in enumerate(GOLD_NOVO["SIZES"])]
    # the option set under which the reference test expects FusedNovoGrad == Novograd (test_fused_novograd.py:128-131)
    opt = FusedNovoGrad(params, lr=o["lr"], betas=o["betas"], eps=o["eps"], weight_decay=o["weight_decay"],
                        grad_averaging=o["grad_averaging"], amsgrad=False, bias_correction=False, reg_inside_moment=True,
                        norm_type=2, init_zero=False)
    opt.multi_tensor_novograd = lambda chunk, flag, lists, norms, *a: oracle.multi_tensor_novograd(chunk, flag, _detached(lists), norms, *a)
    for it in range(GOLD_NOVO["ITERS"]):
        for i, p in enumerate(params):
            p.grad = torch.randn(p.numel(), generator=gen.manual_seed(4000 + 10 * it + i)) * 0.05
        opt.step()
    for p, gp in zip(params, gold["params"]):
        assert (p.detach() - gp).abs().max() <= 1e-3                       # the reference test's bound
        assert torch.allclose(p.detach(), gp, rtol=2e-5, atol=1e-7)
    assert opt.param_groups[0]["step"] == GOLD_NOVO["ITERS"]
    norms32 = opt.param_groups[0]["exp_avg_sq"][1]
    for n, gv in zip(norms32, gold["exp_avg_sq"]):
        assert abs(float(n) ** 2 - float(gv)) <= 2e-5 * float(gv)


def test_fused_adagrad_driver_vs_torch(oracle):
    from beforeholiday_b200.optimizers import FusedAdagrad

    torch.manual_seed(9876)
    sizes = [(33, 17), (1000,), (1,)]
    ref_p = [torch.nn.Parameter(torch.rand(s)) for s in sizes]
    tst_p = [torch.nn.Parameter(p.detach().clone()) for p in ref_p]
    ref = torch.optim.Adagrad(ref_p, lr=5e-4, eps=1e-8, weight_decay=1e-3)
    tst = FusedAdagrad(tst_p, lr=5e-4, eps=1e-8, weight_decay=1e-3)
    tst.multi_tensor_adagrad = lambda chunk, flag, lists, *a: oracle.multi_tensor_adagrad(chunk, flag, _detached(lists), *a)
    for _ in range(7):
        for a, b in zip(ref_p, tst_p):
            a.grad = torch.rand_like(a)
            b.grad = a.grad.clone()
        ref.step()
        tst.step()
    for a, b in zip(ref_p, tst_p):
        assert torch.allclose(b.detach(), a.detach(), rtol=1e-6, atol=2e-7)


@pytest.mark.parametrize("is_skipped", [False, True])
def test_fused_lars_driver_applies_the_layerwise_ratio(oracle, is_skipped):
    from beforeholiday_b200.optimizers import FusedLARS

    torch.manual_seed(1)
    ps = [torch.nn.Parameter(torch.rand(257) + 0.5), torch.nn.Parameter(torch.rand(4, 33) - 0.5)]
    w0 = [p.detach().clone() for p in ps]
    lr, tc, wd, mom = 0.1, 0.02, 1e-4, 0.9
    opt = FusedLARS([{"params": ps, "is_skipped": is_skipped}], lr=lr, momentum=mom, weight_decay=wd, trust_coefficient=tc, eps=1e-8)
    opt.multi_tensor_l2norm = lambda chunk, flag, lists, per: oracle.multi_tensor_l2norm(chunk, flag, _detached(lists), per)
    opt.multi_tensor_lars = lambda chunk, flag, lists, gn, pn, *a: oracle.multi_tensor_lars(chunk, flag, _detached(lists), gn, pn, *a)
    gs = [torch.randn_like(p) * 0.1 for p in ps]
    for p, g in zip(ps, gs):
        p.grad = g.clone()
    opt.step()
    for p, w, g in zip(ps, w0, gs):
        gn, pn = float(g.double().norm()), float(w.double().norm())
        ratio = 1.0 if is_skipped else tc * pn / (gn + pn * wd + 1e-8)
        want = w - lr * ratio * (g + wd * w)                   # first step from zero momentum
        assert torch.allclose(p.detach(), want, rtol=2e-6, atol=1e-8)
        assert torch.allclose(opt.state[p]["momentum_buffer"], -(lr * ratio) * (g + wd * w), rtol=2e-6, atol=1e-9)


def test_fused_mixed_precision_lamb_driver_matches_lamb_and_keeps_masters(oracle):
    """FusedMixedPrecisionLamb's driver (device-side lr / step tensors, fp32 masters for reduced-precision parameters,
    model copy written by the kernel) over the "_mp" restatements == the RefLAMB-pinned multi_tensor_lamb on fp32 copies
    of the same parameters; the 16-bit model parameters are the rounded masters."""
    from beforeholiday_b200.optimizers import FusedMixedPrecisionLamb

    gen = torch.Generator()
    sizes = [4099, 33, 257]
    init = [(torch.randn(n, generator=gen.manual_seed(50 + i)) * 0.1).half() for i, n in enumerate(sizes)]
    params = [torch.nn.Parameter(x.clone()) for x in init]
    opt = FusedMixedPrecisionLamb(params, lr=2e-3, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01, max_grad_norm=1.0,
                                  reduced_precision_dtype=torch.float16)
    opt.multi_tensor_l2norm = lambda chunk, flag, lists, per: oracle.multi_tensor_l2norm_mp(chunk, flag, _detached(lists), per)
    opt.multi_tensor_lamb = lambda chunk, flag, lists, *a: oracle.multi_tensor_lamb_mp(chunk, flag, _detached(lists), *a)
    p32 = [x.float() for x in init]
    m, v = [torch.zeros(n) for n in sizes], [torch.zeros(n) for n in sizes]
    for it in range(4):
        grads = [(torch.randn(n, generator=gen.manual_seed(900 + 10 * it + i)) * 0.05).half() for i, n in enumerate(sizes)]
        for p, g in zip(params, grads):
            p.grad = g.clone()
        opt.step()
        # replay: the kernel sees fp16 gradients, fp32 masters; the update term is stored in the gradient dtype (fp16)
        gs = [g.clone() for g in grads]
        f = torch.zeros(1, dtype=torch.int32)
        gn, _ = oracle.multi_tensor_l2norm(65536, f, [gs], False)
        copies = [torch.zeros(n, dtype=torch.float16) for n in sizes]
        oracle.multi_tensor_lamb_mp(8192, f, [gs, p32, m, v, copies], torch.tensor(2e-3), 0.9, 0.999, 1e-6,
                                    torch.tensor([it + 1], dtype=torch.int32), 1, 0.01, 1, 1, gn, torch.tensor([1.0]), False,
                                    torch.zeros(1), torch.ones(1))
        for p, c in zip(params, copies):
            assert torch.equal(p.detach().view(torch.int16), c.view(torch.int16)), f"step {it}"
    assert int(opt.param_groups[0]['step']) == 4
    for full, want, p in zip(opt.param_groups_full_precision[0]['params'], p32, params):
        assert torch.equal(full, want)                                         # fp32 masters carried by the optimizer
        assert torch.equal(p.detach(), full.half())
    # a raised found-inf flag: nothing moves and the step counter stays
    class Scaler:
        def _check_inf_per_device(self, o):
            return {params[0].device: torch.ones(1)}

        def _get_scale_async(self):
            return torch.tensor([1024.0])
    before = [p.detach().clone() for p in params]
    for p in params:
        p.grad = torch.ones_like(p)
    opt.step(grad_scaler=Scaler())
    assert all(torch.equal(a, b.detach()) for a, b in zip(before, params)) and int(opt.param_groups[0]['step']) == 4


def test_clip_grad_norm_defers_to_torch_off_gpu():
    from beforeholiday_b200.contrib.clip_grad import clip_grad_norm_

    torch.manual_seed(0)
    a = [torch.nn.Parameter(torch.randn(50)), torch.nn.Parameter(torch.randn(3, 7))]
    b = [torch.nn.Parameter(p.detach().clone()) for p in a]
    for p, q in zip(a, b):
        p.grad = torch.randn_like(p) * 3
        q.grad = p.grad.clone()
    n1 = clip_grad_norm_(a, 1.0)
    n2 = torch.nn.utils.clip_grad_norm_(b, 1.0)
    assert torch.equal(n1, n2) and all(torch.equal(p.grad, q.grad) for p, q in zip(a, b))
    assert float(clip_grad_norm_([torch.nn.Parameter(torch.zeros(3))], 1.0)) == 0.0   # no gradients at all
