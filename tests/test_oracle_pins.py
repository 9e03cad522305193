"""Pin the CPU oracle (oracle/mt_oracle.c) before anything trusts it.

Pins, in order of authority:
  1. the known answers of the reference's own unit tests (scale -> 1.0, axpby -> 136.0, l2norm -> torch.norm, the
     nan/inf flag semantics): /root/reference/tests/L0/run_amp/test_multi_tensor_{scale,axpby,l2norm}.py;
  2. outputs of the reference's own CPU-capable Python (apex/amp/scaler.py fallbacks, RefLAMB from the reference's
     test_lamb.py) stored in tests/golden/ref_cpu.pt by tests/golden/make_golden_cpu.py;
  3. torch.optim.Adam/AdamW/SGD trajectories -- the oracle the reference's optimizer tests use
     (tests/L0/run_optimizers/test_fused_optimizer.py), same fixture file;
  4. (GPU box) golden vectors from the reference's CUDA kernels: tests/test_gpu_vs_reference.py.
All run on CPU (`-m "not gpu"`).
"""
import itertools
from pathlib import Path

import pytest
import torch

from util import bits_equal, max_rel, rel_close

GOLD = torch.load(Path(__file__).parent / "golden" / "ref_cpu.pt", weights_only=False)
N, ITERS = GOLD["N"], GOLD["ITERS"]


def gen_inputs(seed, n, dtype=torch.float32, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n, generator=g) * scale).to(dtype)


def flag():
    return torch.zeros(1, dtype=torch.int32)


# ---------------------------------------------------------------------------------------- 1. reference KATs
DTYPES3 = [torch.float32, torch.float16, torch.bfloat16]


@pytest.mark.parametrize("in_t,out_t", list(itertools.product(DTYPES3, DTYPES3)))
@pytest.mark.parametrize("sizea,sizeb", [(7777, 555), (777, 555), (555, 2048 * 32 + 1), (555, 333)])
def test_scale_kat(oracle, in_t, out_t, sizea, sizeb):
    # test_multi_tensor_scale.py:36-56: fill 4.0, scale 1/4 -> exactly 1.0, flag stays 0
    ins = [torch.full((sizea,), 4.0, dtype=in_t), torch.full((sizeb,), 4.0, dtype=in_t)]
    outs = [torch.zeros(sizea, dtype=out_t), torch.zeros(sizeb, dtype=out_t)]
    f = flag()
    oracle.multi_tensor_scale(2048 * 32, f, [ins, outs], 1.0 / 4.0)
    assert all(bool((o == 1.0).all()) for o in outs)
    assert f.item() == 0
    # :119-124: nan at [0][0], inf at the end of the last tensor -> flag
    for t, i, val in ((0, 0, float("nan")), (1, sizeb - 1, float("inf")), (0, sizea // 2, float("-inf"))):
        bad = [x.clone() for x in ins]
        bad[t][i] = val
        f = flag()
        oracle.multi_tensor_scale(2048 * 32, f, [bad, outs], 1.0 / 4.0)
        assert f.item() == 1


@pytest.mark.parametrize("x_t,y_t,o_t", list(itertools.product(DTYPES3, DTYPES3, DTYPES3)))
def test_axpby_kat(oracle, x_t, y_t, o_t):
    # test_multi_tensor_axpby.py:39,75-80: a=2, b=8, x=4, y=16 -> 136.0 in every dtype
    n = 7777
    x, y, o = torch.full((n,), 4.0, dtype=x_t), torch.full((n,), 16.0, dtype=y_t), torch.zeros(n, dtype=o_t)
    f = flag()
    oracle.multi_tensor_axpby(2048 * 32, f, [[x], [y], [o]], 2.0, 8.0, -1)
    assert bool((o == 136.0).all()) and f.item() == 0
    for which, arg, expect in (("x", 0, 1), ("x", 1, 0), ("y", 1, 1), ("y", 0, 0), ("x", -1, 1), ("y", -1, 1)):
        xb, yb = x.clone(), y.clone()
        (xb if which == "x" else yb)[n // 3] = float("inf")
        f = flag()
        oracle.multi_tensor_axpby(2048 * 32, f, [[xb], [yb], [o]], 2.0, 8.0, arg)
        assert f.item() == expect, (which, arg)


@pytest.mark.parametrize("dt", [torch.float32, torch.float16])
@pytest.mark.parametrize("chunk", [2048 * 32, 333, 33333])
@pytest.mark.parametrize("per_tensor", [False, True])
def test_l2norm_kat(oracle, dt, chunk, per_tensor):
    # test_multi_tensor_l2norm.py:35-56: constant 4.0, allclose to torch.norm (global and per tensor), flag 0
    sizes = [7777, 555, 2048 * 32 + 1]
    xs = [torch.full((n,), 4.0, dtype=dt) for n in sizes]
    f = flag()
    norm, norm_pt = oracle.multi_tensor_l2norm(chunk, f, [xs], per_tensor)
    ref = torch.cat([x.float() for x in xs]).norm()
    assert torch.allclose(norm, ref.reshape(1))
    if per_tensor:
        assert torch.allclose(norm_pt, torch.stack([x.float().norm() for x in xs]))
    assert f.item() == 0
    # random data against a float64 sum: the north-star tolerance for norms is 1e-5 relative
    xs = [gen_inputs(5 + i, n, dt) for i, n in enumerate(sizes)]
    norm, norm_pt = oracle.multi_tensor_l2norm(chunk, f, [xs], True)
    assert max_rel(norm, torch.cat([x.double() for x in xs]).norm().reshape(1)) < 1e-5
    assert max_rel(norm_pt, torch.stack([x.double().norm() for x in xs])) < 1e-5


def test_l2norm_flag_is_per_chunk_partial(oracle):
    # the flag is raised by a non-finite CHUNK PARTIAL (l2norm_kernel.cu:103-104): finite inputs whose squares
    # overflow fp32 trip it, and a finite global overflow of finite partials does not.
    x = torch.full((1000,), 3e19, dtype=torch.float32)  # x^2 = 9e38 > FLT_MAX
    f = flag()
    oracle.multi_tensor_l2norm(65536, f, [[x]], False)
    assert f.item() == 1
    y = torch.full((8,), 1.5e19, dtype=torch.float32)  # each chunk (1 element): 2.25e38, finite; the sum is not
    f = flag()
    n, _ = oracle.multi_tensor_l2norm(1, f, [[y]], False)
    assert f.item() == 0 and torch.isinf(n).all()


def test_l2norm_scale(oracle):
    x = gen_inputs(3, 70001)
    o = torch.zeros_like(x, dtype=torch.float16)
    f = flag()
    n, npt = oracle.multi_tensor_l2norm_scale(65536, f, [[x], [o]], 0.5, True)
    assert bits_equal(o, (x * 0.5).half())
    assert max_rel(n, x.double().norm().reshape(1)) < 1e-5 and max_rel(npt, x.double().norm().reshape(1)) < 1e-5


# ---------------------------------------------------------------------------------------- 2. reference python
def test_scale_python_fallback_golden(oracle):
    for case in GOLD["scale_python"]:
        model = gen_inputs(11, N, case["in_dtype"], 3.0)
        if case["inject"] == "inf":
            model[N // 2] = float("inf")
        if case["inject"] == "nan":
            model[-1] = float("nan")
        master = torch.zeros(N, dtype=torch.float32)
        f = flag()
        oracle.multi_tensor_scale(65536, f, [[model], [master]], 1.0 / 65536.0)
        assert bool(f.item()) == case["overflow"]
        if not case["overflow"]:  # the python fallback returns before writing on overflow
            assert bits_equal(master, case["master"])


def test_axpby_python_fallback_golden(oracle):
    for case in GOLD["axpby_python"]:
        model = gen_inputs(12, N, torch.float16, 2.0)
        if case["inject"] == "inf":
            model[7] = float("inf")
        stashed = gen_inputs(13, N, torch.float32)
        master = torch.zeros(N, dtype=torch.float32)
        f = flag()
        oracle.multi_tensor_axpby(65536, f, [[model], [stashed], [master]], 1.0 / 128.0, 1.0, 0)
        assert bool(f.item()) == case["overflow"]
        if not case["overflow"]:
            assert bits_equal(master, case["master"])


@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_lamb_vs_reflamb_golden(oracle, wd):
    # FusedLAMB(use_nvlamb=True) is what the reference test pairs with RefLAMB (test_lamb.py:163-170)
    gold = GOLD["ref_lamb"][wd]
    sizes = gold["sizes"]
    ps = [gen_inputs(100 + i, n).abs() * 0.5 + 0.01 for i, n in enumerate(sizes)]
    ms = [torch.zeros(n) for n in sizes]
    vs = [torch.zeros(n) for n in sizes]
    f = flag()
    for it in range(ITERS):
        gs = [gen_inputs(1000 + 10 * it + i, n) * 0.1 for i, n in enumerate(sizes)]
        gnorm, _ = oracle.multi_tensor_l2norm(65536, f, [gs], False)
        oracle.multi_tensor_lamb(8192, f, [gs, ps, ms, vs], 5e-4, 0.9, 0.999, 1e-6, it + 1, 1, wd, 1, 1, gnorm, 1.0, True)
    for p, gp in zip(ps, gold["params"]):
        # reference test tolerance: max_abs_diff 1e-3 (test_lamb.py:149); the two LAMBs differ only by rounding
        assert (p - gp).abs().max() <= 1e-3
        assert rel_close(p, gp, 2e-5, 1e-7)
    for m, gm in zip(ms, gold["m"]):
        assert rel_close(m, gm, 1e-5, 1e-7)  # moments are signed sums: atol covers cancellation


# ---------------------------------------------------------------------------------------- 3. torch.optim
def _run_oracle_adam(oracle, mode, wd):
    torch.manual_seed(9876)
    p = torch.rand(N)
    m, v = torch.zeros(N), torch.zeros(N)
    for it in range(ITERS):
        g = gen_inputs(2000 + it, N).abs() % 1.0
        oracle.multi_tensor_adam(8192, flag(), [[g], [p], [m], [v]], 5e-4, 0.9, 0.999, 1e-8, it + 1, mode, 1, wd)
    return p


@pytest.mark.parametrize("name,mode,wd", [("adam_l2", 0, 0.0), ("adam_l2_wd", 0, 0.1), ("adamw", 1, 0.01)])
def test_adam_vs_torch_optim_golden(oracle, name, mode, wd):
    p = _run_oracle_adam(oracle, mode, wd)
    gold = GOLD["torch_optim"][name]
    assert (p - gold).abs().max() <= 1e-3                 # the reference's own bound (test_fused_optimizer.py:13)
    assert rel_close(p, gold, 1e-6, 2e-7)                 # north-star: 1e-6 relative (atol covers params near 0)


@pytest.mark.parametrize("name,kw", [
    ("sgd_mom", dict(lr=0.25, momentum=0.125, wd=0.0, nesterov=False)),
    ("sgd_nesterov_wd", dict(lr=0.1, momentum=0.9, wd=1e-4, nesterov=True)),
])
def test_sgd_vs_torch_optim_golden(oracle, name, kw):
    torch.manual_seed(9876)
    p = torch.rand(N)
    mom = torch.zeros(N)
    for it in range(ITERS):
        g = gen_inputs(2000 + it, N).abs() % 1.0
        oracle.multi_tensor_sgd(8192, flag(), [[g], [p], [mom]], kw["wd"], kw["momentum"], 0.0, kw["lr"], kw["nesterov"],
                                it == 0, False, 1.0)
    gold = GOLD["torch_optim"][name]
    assert rel_close(p, gold, 1e-6, 2e-7)


def test_sgd_noop_skips(oracle):
    p0 = gen_inputs(1, 1000)
    p, mom, g = p0.clone(), torch.zeros(1000), gen_inputs(2, 1000)
    f = torch.ones(1, dtype=torch.int32)
    oracle.multi_tensor_sgd(8192, f, [[g], [p], [mom]], 0.0, 0.9, 0.0, 0.1, False, True, False, 1.0)
    assert bits_equal(p, p0) and bool((mom == 0).all())


def test_sgd_fp16_master_copy(oracle):
    # case 3 (fp16 grad, fp32 master, fp16 model copy): the copy is the rounded master
    g, w, mom = gen_inputs(1, 999, torch.float16), gen_inputs(2, 999), torch.zeros(999)
    out = torch.zeros(999, dtype=torch.float16)
    oracle.multi_tensor_sgd(8192, flag(), [[g], [w], [mom], [out]], 1e-4, 0.9, 0.0, 0.1, False, True, False, 1.0 / 128)
    assert bits_equal(out, w.half())


def test_dist_adam_matches_adamw_and_copy(oracle):
    # DistAdamFunctor mode 1 == AdamW on fp32 masters with bf16 grads; p_copy is the rounded master
    n = 5001
    p = gen_inputs(1, n) * 0.02
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([ref], lr=1e-3, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    m, v = torch.zeros(n), torch.zeros(n)
    pc = torch.zeros(n, dtype=torch.bfloat16)
    for it in range(3):
        g = gen_inputs(50 + it, n, torch.bfloat16, 1e-3 * 1024)
        ref.grad = g.float() / 1024
        opt.step()
        oracle.multi_tensor_fused_adam(8192, flag(), [[p], [m], [v], [g], [pc]], [0.9], [0.95], [1], [1e-8], [0.1], 1e-3,
                                       1024.0, it + 1, 1)
    assert rel_close(p, ref.data, 1e-6, 1e-8)
    assert bits_equal(pc, p.bfloat16())


def test_dist_lamb_stages_reproduce_pinned_lamb(oracle):
    """The sharded-LAMB restatement (apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu) run as
    stage 1 -> per-tensor norms -> stage 2 must reproduce, bit for bit, the multi_tensor_lamb restatement that is pinned
    against the reference's RefLAMB above, when nothing clips and the loss scale is 1 (same functor arithmetic)."""
    sizes = [5001, 33, 70001]
    n = len(sizes)
    for mode, wd, nv in ((1, 0.01, False), (0, 0.01, True), (1, 0.0, True)):
        ps = [gen_inputs(100 + i, k).abs() * 0.5 + 0.01 for i, k in enumerate(sizes)]
        ms, vs = [torch.zeros(k) for k in sizes], [torch.zeros(k) for k in sizes]
        ps2, ms2, vs2 = [p.clone() for p in ps], [m.clone() for m in ms], [v.clone() for v in vs]
        us = [torch.zeros(k) for k in sizes]
        copies = [torch.zeros(k, dtype=torch.float16) for k in sizes]
        for it in range(3):
            gs = [gen_inputs(1000 + 10 * it + i, k) * 0.1 for i, k in enumerate(sizes)]
            gs2 = [g.clone() for g in gs]
            f = flag()
            gnorm, _ = oracle.multi_tensor_l2norm(65536, f, [gs], False)
            oracle.multi_tensor_lamb(8192, f, [gs, ps, ms, vs], 5e-4, 0.9, 0.999, 1e-6, it + 1, 1, wd, 1, mode, gnorm, 1e30, nv)
            _, pn = oracle.multi_tensor_l2norm(8192, f, [ps2], True)    # multi_tensor_lamb norms with ITS chunk size
            oracle.multi_tensor_lamb_compute_update_term(8192, f, [gs2, ps2, ms2, vs2, us], [0.9] * n, [0.999] * n, [0.0] * n,
                                                         [1] * n, it + 1, [1e-6] * n, mode, [wd] * n, 1.0, gnorm, 0.0)
            _, un = oracle.multi_tensor_l2norm(8192, f, [us], True)
            oracle.multi_tensor_lamb_update_weights(8192, f, [us, ps2, copies], pn, un, list(range(n)), 5e-4, [wd] * n, gnorm, nv)
        for a, b in zip(ps + ms + vs, ps2 + ms2 + vs2):
            assert bits_equal(a, b)
        for c, p in zip(copies, ps2):
            assert bits_equal(c, p.half())


def test_dist_lamb_combined_scale(oracle):
    """combined_scale = global_scale / min(1, max_norm / (gnorm/global_scale + 1e-6)) (:133-137): unscaling and
    clipping in one divisor."""
    g = gen_inputs(7, 4099) * 64
    p, m, v, u = gen_inputs(8, 4099), torch.zeros(4099), torch.zeros(4099), torch.zeros(4099)
    gnorm = g.double().norm().float()
    oracle.multi_tensor_lamb_compute_update_term(8192, flag(), [[g], [p], [m], [v], [u]], [0.9], [0.999], [0.0], [0], 1,
                                                 [1e-6], 1, [0.0], 64.0, gnorm, 1.0)
    cs = 64.0 / min(1.0, 1.0 / (float(gnorm) / 64.0 + 1e-6))
    want_m = (1 - torch.tensor(0.9)) * (g / torch.tensor(cs, dtype=torch.float32))
    assert rel_close(m, want_m, 1e-6, 1e-9)
    assert abs(float((g / cs).double().norm()) - 1.0) < 1e-4      # the clipped, unscaled gradient has norm max_norm


# ---------------------------------------------------------------------------------------- 5. the "next" optimizers
GOLD_NOVO = torch.load(Path(__file__).parent / "golden" / "ref_novograd.pt", weights_only=False)


@pytest.mark.parametrize("case", sorted(GOLD_NOVO["cases"]))
def test_novograd_vs_reference_python_golden(oracle, case):
    """multi_tensor_novograd restatement (csrc/multi_tensor_novograd.cu:66-121 + the norm blend of
    multi_tensor_l2norm_kernel.cu:272-298) against the pure-torch Novograd the reference's own test compares
    FusedNovoGrad with (tests/L0/run_optimizers/test_fused_novograd.py:10-114; fixture made by
    tests/golden/make_golden_novograd.py), driven the way apex/optimizers/fused_novograd.py:118-214 drives the kernel:
    norms initialised with the first step's gradient norm, no bias correction, regularisation inside the moment, L2
    norm.  Tolerances: the reference test's 1e-3 absolute / relative (:test_fused_optimizer.py:13-16) and 2e-5."""
    gold = GOLD_NOVO["cases"][case]
    opts, sizes = gold["opts"], GOLD_NOVO["SIZES"]
    g0 = torch.Generator()
    ps = [(torch.randn(n, generator=g0.manual_seed(300 + i)) * 0.1) for i, n in enumerate(sizes)]
    ms = [torch.zeros(n) for n in sizes]
    norms = None
    for it in range(GOLD_NOVO["ITERS"]):
        gs = [(torch.randn(n, generator=g0.manual_seed(4000 + 10 * it + i)) * 0.05) for i, n in enumerate(sizes)]
        if norms is None:   # init_zero=False: "init with first step norm, so first blend have no effect" (:167-176)
            norms = torch.tensor([float(torch.sum(torch.pow(g, 2)).sqrt()) for g in gs])
        oracle.multi_tensor_novograd(8192, flag(), [gs, ps, ms], norms, opts["lr"], opts["betas"][0], opts["betas"][1],
                                     opts["eps"], it + 1, 0, opts["weight_decay"], 1 if opts["grad_averaging"] else 0, 0, 2)
    for p, gp in zip(ps, gold["params"]):
        assert (p - gp).abs().max() <= 1e-3
        assert rel_close(p, gp, 2e-5, 1e-7)
    for m, gm in zip(ms, gold["exp_avg"]):
        assert rel_close(m, gm, 2e-5, 1e-7)
    # the fused op keeps the NORM, the python reference its square
    for n, gv in zip(norms, gold["exp_avg_sq"]):
        assert abs(float(n) ** 2 - float(gv)) <= 2e-5 * float(gv)


@pytest.mark.parametrize("wd", [0.0, 1e-3])
def test_adagrad_vs_torch_optim(oracle, wd):
    """AdagradFunctor restatement, mode 0 (L2), against torch.optim.Adagrad -- the oracle of the reference's
    TestFusedAdagrad (tests/L0/run_optimizers/test_fused_optimizer.py:92-139: lr 5e-4, eps 1e-8, weight_decay 1e-3)."""
    torch.manual_seed(9876)
    p = torch.rand(N)
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.Adagrad([ref], lr=5e-4, eps=1e-8, weight_decay=wd)
    h = torch.zeros(N)
    for it in range(ITERS):
        g = gen_inputs(6000 + it, N).abs() % 1.0
        ref.grad = g.clone()
        opt.step()
        oracle.multi_tensor_adagrad(8192, flag(), [[g], [p], [h]], 5e-4, 1e-8, 0, wd)
    assert (p - ref.data).abs().max() <= 1e-3
    assert rel_close(p, ref.data, 1e-6, 2e-7)
    assert rel_close(h, opt.state[ref]["sum"], 1e-6, 1e-9)


@pytest.mark.parametrize("mode,wd,nv", [(1, 0.01, False), (0, 0.01, True), (1, 0.0, False)])
def test_lamb_mp_reproduces_pinned_lamb(oracle, mode, wd, nv):
    """The "_mp" LAMB restatement (csrc/multi_tensor_lamb_mp.cu) with inv_scale = 1 and fp32 everywhere is the same
    arithmetic as multi_tensor_lamb, which is pinned against the reference's RefLAMB above: bit for bit, incl. norms."""
    sizes = [5001, 33, 70001]
    ps = [gen_inputs(100 + i, k).abs() * 0.5 + 0.01 for i, k in enumerate(sizes)]
    ms, vs = [torch.zeros(k) for k in sizes], [torch.zeros(k) for k in sizes]
    ps2, ms2, vs2 = [x.clone() for x in ps], [x.clone() for x in ms], [x.clone() for x in vs]
    for it in range(3):
        gs = [gen_inputs(1000 + 10 * it + i, k) * 0.1 for i, k in enumerate(sizes)]
        gs2 = [g.clone() for g in gs]
        gnorm, _ = oracle.multi_tensor_l2norm(65536, flag(), [gs], False)
        oracle.multi_tensor_lamb(8192, flag(), [gs, ps, ms, vs], 5e-4, 0.9, 0.999, 1e-6, it + 1, 1, wd, 1, mode, gnorm, 1.0, nv)
        oracle.multi_tensor_lamb_mp(8192, flag(), [gs2, ps2, ms2, vs2], torch.tensor(5e-4), 0.9, 0.999, 1e-6,
                                    torch.tensor([it + 1], dtype=torch.int32), 1, wd, 1, mode, gnorm, torch.tensor([1.0]), nv,
                                    torch.zeros(1), torch.tensor([1.0]))
        for a, b in zip(gs + ps + ms + vs, gs2 + ps2 + ms2 + vs2):
            assert bits_equal(a, b)


@pytest.mark.parametrize("nesterov", [False, True])
def test_lars_without_trust_ratio_is_torch_sgd_and_ratio_formula(oracle, nesterov):
    """LARSFunctor restatement (csrc/multi_tensor_lars.cu:84-135).  With the layer-wise ratio switched off (is_skipped, the
    reference's route for biases / norm layers) it is SGD with momentum in the "velocity carries -lr" form, i.e.
    torch.optim.SGD at constant lr; with it, the step is scaled by tc*|w| / (|g| + wd*|w| + eps) (:96-104)."""
    torch.manual_seed(9876)
    w = torch.rand(N)
    ref = torch.nn.Parameter(w.clone())
    opt = torch.optim.SGD([ref], lr=0.05, momentum=0.9, weight_decay=1e-4, nesterov=nesterov)
    mom = torch.zeros(N)
    for it in range(ITERS):
        g = gen_inputs(7000 + it, N) * 0.1
        ref.grad = g.clone()
        opt.step()
        oracle.multi_tensor_lars(8192, flag(), [[g], [w], [mom]], [1.0], [1.0], 0.05, 0.001, 0.0, 1e-4, 0.9, 0.0, nesterov,
                                 False, False, 1.0, True)
    assert rel_close(w, ref.data, 1e-6, 2e-7)
    # one step WITH the ratio, from zero momentum: w' = w - lr*trust*(g + wd*w)
    w2, g2, m2 = torch.rand(N), gen_inputs(7100, N) * 0.1, torch.zeros(N)
    gn, pn = float(g2.double().norm()), float(w2.double().norm())
    want = w2 - 0.05 * (0.001 * pn / (gn + pn * 1e-4 + 1e-8)) * (g2 + 1e-4 * w2) * (1.9 if nesterov else 1.0)
    oracle.multi_tensor_lars(8192, flag(), [[g2], [w2], [m2]], [gn], [pn], 0.05, 0.001, 1e-8, 1e-4, 0.9, 0.0, nesterov, False,
                             False, 1.0, False)
    assert rel_close(w2, want, 2e-6, 1e-8)


@pytest.mark.parametrize("wd,nv", [(0.01, False), (0.0, True)])
def test_legacy_split_lamb_follows_pinned_lamb(oracle, wd, nv):
    """multi_tensor_lamb_stage1_cuda / stage2_cuda restatements (csrc/multi_tensor_lamb_stage_1.cu:81-100,126-130,
    _stage_2.cu:41-47,77) in fp32: the same LAMB as multi_tensor_lamb (decoupled decay, averaged gradients) except that
    the legacy host code corrects the bias with step+1 through float pow -- so it must follow the RefLAMB-pinned
    restatement run at step+1: moments bit for bit, parameters to 1e-6."""
    sizes = [5001, 33, 20001]
    n = len(sizes)
    ps = [gen_inputs(100 + i, k).abs() * 0.5 + 0.01 for i, k in enumerate(sizes)]
    ms, vs = [torch.zeros(k) for k in sizes], [torch.zeros(k) for k in sizes]
    ps2, ms2, vs2 = [x.clone() for x in ps], [x.clone() for x in ms], [x.clone() for x in vs]
    us = [torch.zeros(k) for k in sizes]
    for it in range(3):
        gs = [gen_inputs(1000 + 10 * it + i, k) * 0.1 for i, k in enumerate(sizes)]
        gs2 = [g.clone() for g in gs]
        gnorm, _ = oracle.multi_tensor_l2norm(65536, flag(), [gs], False)
        oracle.multi_tensor_lamb(8192, flag(), [gs, ps, ms, vs], 5e-4, 0.9, 0.999, 1e-6, it + 1, 1, wd, 1, 1, gnorm, 1.0, nv)
        _, pn = oracle.multi_tensor_l2norm(8192, flag(), [ps2], True)
        oracle.multi_tensor_lamb_stage1_cuda(8192, flag(), [gs2, ps2, ms2, vs2, us], [wd] * n, it, 0.9, 0.999, 1e-6, gnorm, 1.0)
        _, un = oracle.multi_tensor_l2norm(8192, flag(), [us], True)
        oracle.multi_tensor_lamb_stage2_cuda(8192, flag(), [ps2, us], pn, un, 5e-4, wd, nv)
        # multi_tensor_lamb leaves its update term in g.  1 - powf(0.999, t) cancels: the float pow's 6e-8 error becomes
        # ~3e-5 relative in bias-correction 2, hence ~1e-4 in the update term -- a property of the reference's legacy host
        # code, not of the restatement (moments, which take no bias correction, must agree exactly)
        for a, b in zip(gs, us):
            assert rel_close(b, a, 3e-4, 1e-7)
    for a, b in zip(ms + vs, ms2 + vs2):
        assert bits_equal(a, b)
    for a, b in zip(ps, ps2):
        assert rel_close(b, a, 1e-6, 1e-9)


def test_e5m2_packing_is_the_reference_formula():
    """apex/contrib/csrc/optimizers/multi_tensor_distopt_lamb_kernel.cu:34-66: byte = high byte of
    half(v + 2^floor(log2|v|) / 8); unpack = that byte over a zero low byte.  Restated here with torch ops (fp32 add,
    round-to-nearest fp16 cast) and compared with the oracle's C over random and edge values."""
    import torch
    from oracle import mt_oracle as O

    g = torch.Generator().manual_seed(5)
    v = torch.cat([torch.randn(20000, generator=g) * 10 ** torch.randint(-6, 5, (20000,), generator=g).float(),
                   torch.tensor([0.0, -0.0, 1.125, -1.125, 1.0, 65504.0, 70000.0, 6e-8, float("inf"), -float("inf")])])
    pow2 = (v.view(torch.int32) & -8388608).view(torch.float32)          # 0xFF800000: sign + exponent
    want = ((v + pow2 / 8.0).half().view(torch.int16).to(torch.int32) >> 8) & 0xFF
    got = O.convert(v, torch.uint8).to(torch.int32)
    assert torch.equal(got, want)
    back = O.convert(got.to(torch.uint8), torch.float32)
    assert torch.equal(back, (want << 8).to(torch.int16).view(torch.float16).float())
