"""GPU parity for the sharded-LAMB kernels: multi_tensor_lamb_compute_update_term,
multi_tensor_lamb_update_weights and maybe_cast_mt -- against the CPU oracle (bit-exact: no norm is computed inside the
reference kernels), against the fused per-parameter norm sums' double-precision restatement (1e-5 relative), and against
the reference's own CUDA kernels (oracle/_ref/distributed_lamb_cuda, compiled unmodified)."""
import pytest
import torch

from test_gpu_ops import BF16, F16, F32, make_lists, to_gpu
from util import bits_equal, cpu_clone, n_mismatch

pytestmark = pytest.mark.gpu

SIZES = [1, 9, 4097, 70001, 300000, 0, 16384, 33]


def _hyper(n, uniform=False):
    """Per-fragment hyper-parameters; non-uniform so that a wrong tensor index shows."""
    k = torch.arange(n, dtype=torch.float32)
    if uniform:
        b1, b2, eps, wd = torch.full((n,), 0.9), torch.full((n,), 0.999), torch.full((n,), 1e-6), torch.full((n,), 0.01)
        bc = torch.ones(n, dtype=torch.int32)
    else:
        b1 = 0.9 - 0.01 * (k % 3)
        b2 = 0.999 - 0.001 * (k % 2)
        eps = 1e-6 * (1 + k % 4)
        wd = torch.where(k % 3 == 1, torch.zeros(n), 0.01 * (1 + k % 5))
        bc = (k % 4 != 2).to(torch.int32)
    return b1, b2, torch.zeros(n), bc, eps, wd


def _stage1_lists(g_t, seed, offsets=None):
    lists = make_lists([g_t, F32, F32, F32, F32], SIZES, seed=seed, scale=0.1, offsets=offsets)
    lists[0] = [t * 128 for t in lists[0]]   # loss-scaled gradients
    lists[3] = [t.abs() for t in lists[3]]
    return lists


@pytest.mark.parametrize("g_t", [F16, BF16, F32])
@pytest.mark.parametrize("mode,max_norm,uniform", [(1, 1.0, False), (0, 1.0, False), (1, 0.0, True), (0, 0.0, False)])
def test_compute_update_term_matches_oracle(amp_C, oracle, g_t, mode, max_norm, uniform):
    lists = _stage1_lists(g_t, 31)
    n = len(SIZES)
    b1, b2, b3, bc, eps, wd = _hyper(n, uniform)
    lg, lc = to_gpu(lists), cpu_clone(lists)
    gs = torch.tensor([128.0])
    gnorm = torch.cat([g.double().reshape(-1) for g in lists[0]]).norm().float().reshape(1)   # > max_norm * scale: clips
    step = torch.tensor([5], dtype=torch.int32)
    fg, fc = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32)
    d = lambda t: t.cuda()
    r = amp_C.multi_tensor_lamb_compute_update_term(8192, fg, lg, d(b1), d(b2), d(b3), d(bc), d(step), d(eps), mode, d(wd),
                                                    d(gs), d(gnorm), max_norm)
    assert r is None
    oracle.multi_tensor_lamb_compute_update_term(8192, fc, lc, b1, b2, b3, bc, step, eps, mode, wd, gs, gnorm, max_norm)
    assert fg.item() == 0
    for li in (2, 3, 4):
        for i, (a, b) in enumerate(zip(lg[li], lc[li])):
            assert bits_equal(a, b), f"list {li} tensor {i}: {n_mismatch(a, b)} mismatches"
    for li in (0, 1):   # g and p are inputs only
        for a, b in zip(lg[li], lists[li]):
            assert bits_equal(a, b)


@pytest.mark.parametrize("offsets", [(1, 1, 1, 1, 1), (2, 4, 4, 4, 4), (3, 0, 1, 2, 5)])
def test_compute_update_term_misaligned_fragments(amp_C, oracle, offsets):
    lists = _stage1_lists(BF16, 32, offsets)
    n = len(SIZES)
    b1, b2, b3, bc, eps, wd = _hyper(n)
    lg, lc = to_gpu(lists, offsets), cpu_clone(lists)
    gs, gnorm, step = torch.tensor([1.0]), torch.tensor([0.5]), torch.tensor([1], dtype=torch.int32)
    fg, fc = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32)
    d = lambda t: t.cuda()
    amp_C.multi_tensor_lamb_compute_update_term(8192, fg, lg, d(b1), d(b2), d(b3), d(bc), d(step), d(eps), 1, d(wd), d(gs),
                                                d(gnorm), 1.0)
    oracle.multi_tensor_lamb_compute_update_term(8192, fc, lc, b1, b2, b3, bc, step, eps, 1, wd, gs, gnorm, 1.0)
    for li in (2, 3, 4):
        for i, (a, b) in enumerate(zip(lg[li], lc[li])):
            assert bits_equal(a, b), f"list {li} tensor {i}: {n_mismatch(a, b)} mismatches"


@pytest.mark.parametrize("model_t", [F16, BF16, F32])
def test_fused_norm_sums(amp_C, oracle, model_t):
    """Extension: per-parameter-run sums of round_model(p)^2 and u^2, checked against double-precision sums."""
    lists = _stage1_lists(F16, 33)
    n = len(SIZES)
    b1, b2, b3, bc, eps, wd = _hyper(n)
    lg, lc = to_gpu(lists), cpu_clone(lists)
    gs, gnorm, step = torch.tensor([128.0]), torch.tensor([3.0]), torch.tensor([2], dtype=torch.int32)
    seg = torch.tensor([0, 1, 3, 4, 7, 8], dtype=torch.int32)   # runs of fragments: {0} {1,2} {3} {4,5,6} {7}
    fg, fc = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32)
    d = lambda t: t.cuda()
    out = amp_C.multi_tensor_lamb_compute_update_term(8192, fg, lg, d(b1), d(b2), d(b3), d(bc), d(step), d(eps), 1, d(wd),
                                                      d(gs), d(gnorm), 1.0, d(seg), model_t)
    sq = oracle.multi_tensor_lamb_compute_update_term(8192, fc, lc, b1, b2, b3, bc, step, eps, 1, wd, gs, gnorm, 1.0, model_t)
    assert out.shape == (2, 5) and out.dtype == F32 and fg.item() == 0
    want = torch.stack([torch.stack([sq[s, seg[k]:seg[k + 1]].sum() for k in range(5)]) for s in range(2)])
    assert torch.allclose(out.cpu().double(), want, rtol=1e-5, atol=0), (out.cpu(), want)
    for li in (2, 3, 4):   # the update itself is unchanged by the extension
        for a, b in zip(lg[li], lc[li]):
            assert bits_equal(a, b)
    # the model-dtype rounding matters: it is the norm of the 16-bit model parameter that the reference takes
    if model_t != F32:
        plain = torch.stack([(p.double() ** 2).sum() for p in lists[1]])
        rounded = torch.stack([(p.to(model_t).double() ** 2).sum() for p in lists[1]])
        assert torch.allclose(sq[0], rounded, rtol=1e-12) and not torch.equal(plain, rounded)
    # a non-finite update raises the flag, as the reference's l2norm sweep over u does
    lg2 = to_gpu(lists)
    lg2[0][3][5] = float("inf")
    amp_C.multi_tensor_lamb_compute_update_term(8192, fg, lg2, d(b1), d(b2), d(b3), d(bc), d(step), d(eps), 1, d(wd), d(gs),
                                                d(gnorm), 1.0, d(seg), model_t)
    assert fg.item() == 1


@pytest.mark.parametrize("copy_t", [F16, BF16, F32])
@pytest.mark.parametrize("use_nvlamb", [False, True])
def test_update_weights_matches_oracle(amp_C, oracle, copy_t, use_nvlamb):
    lists = make_lists([F32, F32, copy_t], SIZES, seed=34, scale=0.1)
    n = len(SIZES)
    wd = _hyper(n)[5]
    lg, lc = to_gpu(lists), cpu_clone(lists)
    pn = torch.tensor([1.5, 0.0, 2.0, 3.0, 0.7, 1.0, 9.0, 4.0])
    un = torch.tensor([0.5, 0.25, 0.0, 2.0, 1.0])
    off = torch.tensor([0, 1, 1, 2, 3, 3, 3, 4], dtype=torch.int64)
    lr = torch.tensor([2e-3])
    fg, fc = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32)
    d = lambda t: t.cuda()
    amp_C.multi_tensor_lamb_update_weights(8192, fg, lg, d(pn), d(un), d(off), d(lr), d(wd), d(lr), use_nvlamb)
    oracle.multi_tensor_lamb_update_weights(8192, fc, lc, pn, un, off, lr, wd, lr, use_nvlamb)
    for li in (0, 1, 2):
        for i, (a, b) in enumerate(zip(lg[li], lc[li])):
            assert bits_equal(a, b), f"list {li} tensor {i}: {n_mismatch(a, b)} mismatches"


@pytest.mark.parametrize("use_nvlamb", [False, True])
def test_update_weights_e5m2_copy_matches_oracle(amp_C, oracle, use_nvlamb):
    """e5m2_allgather: the copy list is one byte per element, packed as the reference does
    (multi_tensor_distopt_lamb_kernel.cu:34-53, dispatch :489) -- NOT torch's round-to-nearest-even float8_e5m2."""
    lists = make_lists([F32, F32, F32], SIZES, seed=37, scale=0.1)
    lists[2] = [torch.zeros(t.numel(), dtype=torch.uint8) for t in lists[2]]
    n = len(SIZES)
    wd = _hyper(n)[5]
    lg, lc = to_gpu(lists), cpu_clone(lists)
    pn = torch.tensor([1.5, 0.0, 2.0, 3.0, 0.7, 1.0, 9.0, 4.0])
    un = torch.tensor([0.5, 0.25, 0.0, 2.0, 1.0])
    off = torch.tensor([0, 1, 1, 2, 3, 3, 3, 4], dtype=torch.int64)
    lr = torch.tensor([2e-3])
    fg, fc = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32)
    d = lambda t: t.cuda()
    amp_C.multi_tensor_lamb_update_weights(8192, fg, lg, d(pn), d(un), d(off), d(lr), d(wd), d(lr), use_nvlamb)
    oracle.multi_tensor_lamb_update_weights(8192, fc, lc, pn, un, off, lr, wd, lr, use_nvlamb)
    for i, (a, b) in enumerate(zip(lg[1], lc[1])):
        assert bits_equal(a, b)
    for i, (a, b) in enumerate(zip(lg[2], lc[2])):
        assert torch.equal(a.cpu(), b), f"tensor {i}: {int((a.cpu() != b).sum())} packed bytes differ"
    # and it is the reference's round-half-up packing: differs from torch's float8_e5m2 on ties
    t = torch.tensor([1.125, -1.125, 1.0, 3.0], device="cuda")
    out = torch.zeros(4, dtype=torch.uint8, device="cuda")
    amp_C.multi_tensor_lamb_update_weights(8192, fg, [[torch.zeros(4, device="cuda")], [t.clone()], [out]], d(torch.ones(1)),
                                           d(torch.ones(1)), d(torch.zeros(1, dtype=torch.int64)), d(lr), d(torch.zeros(1)),
                                           d(lr), False)
    assert out.tolist() == [0x3D, 0xBD, 0x3C, 0x42]


@pytest.mark.parametrize("out_t", [F16, BF16, F32])
def test_maybe_cast_mt_unpacks_e5m2(amp_C, oracle, out_t):
    g = torch.Generator().manual_seed(38)
    packed = [torch.randint(0, 256, (n,), dtype=torch.uint8, generator=g) for n in SIZES]
    outs = [torch.zeros(n, dtype=out_t) for n in SIZES]
    lg = [[t.cuda() for t in packed], [t.cuda() for t in outs]]
    fg = torch.zeros(1, dtype=torch.int32, device="cuda")
    amp_C.maybe_cast_mt(8192, fg, lg)
    for a, b in zip(lg[1], packed):
        want = oracle.convert(oracle.convert(b, F32), out_t)      # byte -> high byte of an fp16 -> out_t
        assert bits_equal(a, want)


def test_flag_set_touches_nothing(amp_C):
    lists = _stage1_lists(F16, 35)
    n = len(SIZES)
    b1, b2, b3, bc, eps, wd = [t.cuda() for t in _hyper(n)]
    lg = to_gpu(lists)
    before = [[t.clone() for t in l] for l in lg]
    one = torch.ones(1, dtype=torch.int32, device="cuda")
    t = lambda v, dt=F32: torch.tensor(v, dtype=dt, device="cuda")
    seg = t([0, 8], torch.int32)
    out = amp_C.multi_tensor_lamb_compute_update_term(8192, one, lg, b1, b2, b3, bc, t([1], torch.int32), eps, 1, wd, t([1.0]),
                                                      t([1.0]), 1.0, seg, F16)
    assert float(out.abs().sum()) == 0.0
    amp_C.multi_tensor_lamb_update_weights(8192, one, [lg[4], lg[1], lg[0]], t([1.0] * n), t([1.0] * n),
                                           torch.arange(n, device="cuda"), t([1e-3]), wd, t([1.0]), True)
    amp_C.maybe_cast_mt(8192, one, [lg[1], lg[0]])
    for la, lb in zip(lg, before):
        assert all(bits_equal(a, b) for a, b in zip(la, lb))


@pytest.mark.parametrize("in_t,out_t", [(F16, F32), (F32, F16), (BF16, BF16), (F32, BF16), (F16, F16)])
def test_maybe_cast_mt(amp_C, oracle, in_t, out_t):
    lists = make_lists([in_t, out_t], SIZES, seed=36, scale=3.0)
    lists[0][2][7] = float("inf")   # non-finite values are copied, not flagged
    lg, lc = to_gpu(lists), cpu_clone(lists)
    fg, fc = torch.zeros(1, dtype=torch.int32, device="cuda"), torch.zeros(1, dtype=torch.int32)
    amp_C.maybe_cast_mt(8192, fg, lg)
    oracle.maybe_cast_mt(8192, fc, lc)
    assert fg.item() == 0
    for a, b in zip(lg[1], lc[1]):
        assert bits_equal(a, b)


def test_vs_reference_cuda(amp_C):
    """The reference's own kernels (--use_fast_math build) on the same inputs."""
    from test_gpu_vs_reference import _load_ref, check_close, clone, flag, rnd

    ref = _load_ref("distributed_lamb_cuda")
    sizes = [9, 4097, 70001, 278011]
    n = len(sizes)
    t = lambda v, dt=F32: torch.tensor(v, dtype=dt, device="cuda")
    for g_t, mode, max_norm in ((F16, 1, 1.0), (F32, 0, 1.0), (F16, 1, 0.0)):
        dts = [g_t, F32, F32, F32, F32]
        lists = [[rnd(m, dt, 10 * d + i, 0.1) for i, m in enumerate(sizes)] for d, dt in enumerate(dts)]
        lists[0] = [g * 128 for g in lists[0]]
        lists[3] = [v.abs() for v in lists[3]]
        b1, b2, b3, bc, eps, wd = [x.cuda() for x in _hyper(n)]
        gnorm = torch.cat([g.double().reshape(-1) for g in lists[0]]).norm().float().reshape(1)
        args = (b1, b2, b3, bc, t([3], torch.int32), eps, mode, wd, t([128.0]), gnorm, max_norm)
        l1, l2 = clone(lists), clone(lists)
        amp_C.multi_tensor_lamb_compute_update_term(8192, flag(), l1, *args)
        ref.multi_tensor_lamb_compute_update_term(8192, flag(), l2, *args)
        for d in (2, 3, 4):
            for a, b in zip(l1[d], l2[d]):
                check_close(a, b, f"update_term list {d} ({g_t}, mode {mode}, max_norm {max_norm})")
        # stage 2 on the reference's update term, fp16 model copy
        pn, un = t([1.5, 0.0, 2.0, 3.0]), t([0.5, 0.25, 1.0])
        off = t([0, 1, 1, 2], torch.int64)
        for nv in (False, True):
            copies1 = [torch.empty_like(p, dtype=F16) for p in lists[1]]
            copies2 = [torch.empty_like(p, dtype=F16) for p in lists[1]]
            p1, p2 = [p.clone() for p in lists[1]], [p.clone() for p in lists[1]]
            amp_C.multi_tensor_lamb_update_weights(8192, flag(), [l2[4], p1, copies1], pn, un, off, t([2e-3]), wd, gnorm, nv)
            ref.multi_tensor_lamb_update_weights(8192, flag(), [l2[4], p2, copies2], pn, un, off, t([2e-3]), wd, gnorm, nv)
            for a, b in zip(p1 + copies1, p2 + copies2):
                check_close(a, b, f"update_weights nvlamb={nv}")


# ------------------------------------------------------------------------------------------------ the optimizer class
LAMB_SIZES = [1000, 7, 40000, 333333, 1, 129]


def _check_vs_replay(got, want_pc, dtype):
    for a, b in zip(got, want_pc):
        a = a.detach().cpu().reshape(-1)
        if dtype == F32:
            assert torch.allclose(a, b, rtol=2e-6, atol=1e-8)
        else:
            ia, ib = a.view(torch.int16).to(torch.int32), b.view(torch.int16).to(torch.int32)
            assert int((ia - ib).abs().max()) <= 1
            assert float((ia != ib).float().mean()) < 0.01


@pytest.mark.parametrize("dtype", [BF16, F16, F32])
@pytest.mark.parametrize("mode,use_nvlamb,max_norm", [(1, False, 1.0), (0, True, 0.0)])
def test_distributed_fused_lamb_single_gpu_matches_oracle_replay(oracle, dtype, mode, use_nvlamb, max_norm):
    """world = 1 (no process group): the whole pipeline -- norm, overflow/step bookkeeping on the device, stage 1 with the
    fused per-parameter norm sums, stage 2 per bucket -- against the oracle applied to whole parameters."""
    from beforeholiday_b200.contrib.optimizers import DistributedFusedLAMB
    from lamb_replay import make_grad, replay

    torch.manual_seed(0)
    init = [(torch.randn(n) * 0.05).to(dtype) for n in LAMB_SIZES]
    params = [torch.nn.Parameter(x.clone().cuda()) for x in init]
    groups = [dict(params=params[0::2], weight_decay=0.01), dict(params=params[1::2], weight_decay=0.0)]
    opt = DistributedFusedLAMB(groups, lr=2e-3, betas=(0.9, 0.999), eps=1e-6, max_grad_norm=max_norm,
                               adam_w_mode=bool(mode), use_nvlamb=use_nvlamb, dwu_num_blocks=2, dwu_num_chunks=2)
    assert opt.layout.nbuckets == 4
    opt.set_global_scale(torch.tensor([64.0], device="cuda"))
    order = params[0::2] + params[1::2]
    for it in range(4):
        for i, p in enumerate(order):
            p.grad.copy_(make_grad(1000 * it + i, p.numel(), dtype, 64.0).view(p.shape))
        if it == 2:
            order[3].grad.view(-1)[5] = float("nan")
        opt.complete_reductions()
        opt.step()
        assert int(opt._overflow_buf) == (1 if it == 2 else 0)
    assert int(opt._step) == 3 and int(opt.state_dict()['step']) == 3
    oi = init[0::2] + init[1::2]
    props = [(0.01, 1, 0.9, 0.999, 1e-6)] * 3 + [(0.0, 1, 0.9, 0.999, 1e-6)] * 3
    p32, pc, _ = replay(oracle, [x.numel() for x in oi], oi, props, 1, 4, dtype, 2e-3, 64.0, max_norm, mode, use_nvlamb,
                        lambda it, r, i: 1000 * it + i, overflow_at=2)
    _check_vs_replay(order, pc, dtype)
    # the fp32 masters of the shard (= everything at world 1) follow the replay to norm-rounding accuracy
    for f in opt.frags:
        n = f.flat_hi - f.flat_lo
        a = opt._fp32_p[f.state_lo:f.state_lo + n].cpu()
        b = p32[f.param][f.param_lo:f.param_lo + n]
        assert torch.allclose(a, b, rtol=2e-6, atol=1e-8)
    # checkpoint round trip
    sd = {k: v.clone() for k, v in opt.state_dict().items()}
    opt._fp32_p.zero_()
    opt.load_state_dict(sd)
    assert torch.equal(opt._fp32_p, sd['fp32_p'])


@pytest.mark.parametrize("option", ["e5m2_allgather", "clip_before_ar", "overlap_hooks"])
@pytest.mark.parametrize("dtype", [F16, BF16])
def test_distributed_fused_lamb_options_single_gpu(oracle, dtype, option):
    """The reference's e5m2_allgather / clip_after_ar=False / overlap_reductions options (distributed_fused_lamb.py:88-91,
    :629-646, :783-797) at world 1 against the oracle replay with the same option: e5m2 parameters are the packed
    masters (bit-level: the replay packs with the oracle's restatement of the reference formula); clipping before the
    reduction uses the local norm; with hooks the gradients come from a real backward pass."""
    from beforeholiday_b200.contrib.optimizers import DistributedFusedLAMB
    from lamb_replay import make_grad, replay

    torch.manual_seed(0)
    init = [(torch.randn(n) * 0.05).to(dtype) for n in LAMB_SIZES]
    params = [torch.nn.Parameter(x.clone().cuda()) for x in init]
    groups = [dict(params=params[0::2], weight_decay=0.01), dict(params=params[1::2], weight_decay=0.0)]
    kw = {"e5m2_allgather": dict(e5m2_allgather=True), "clip_before_ar": dict(clip_after_ar=False),
          "overlap_hooks": dict(overlap_reductions=True)}[option]
    opt = DistributedFusedLAMB(groups, lr=2e-3, betas=(0.9, 0.999), eps=1e-6, max_grad_norm=1.0, dwu_num_blocks=2,
                               dwu_num_chunks=2, **kw)
    opt.set_global_scale(torch.tensor([64.0], device="cuda"))
    order = params[0::2] + params[1::2]
    for it in range(3):
        opt.zero_grad()
        gs = [make_grad(1000 * it + i, p.numel(), dtype, 64.0).view(p.shape).cuda() for i, p in enumerate(order)]
        if option == "overlap_hooks":
            # d/dp sum(p * g) = g: a real backward pass, gradients accumulate into the flat buffer's views
            sum((p * g).sum() for p, g in zip(order, gs)).backward()
            assert all(p.grad.data_ptr() == opt.grads_flat[o:o + 1].data_ptr() for p, o in zip(order, opt.layout.offsets))
        else:
            for p, g in zip(order, gs):
                p.grad.copy_(g)
        opt.complete_reductions()
        opt.step()
        assert int(opt._overflow_buf) == 0
    oi = init[0::2] + init[1::2]
    props = [(0.01, 1, 0.9, 0.999, 1e-6)] * 3 + [(0.0, 1, 0.9, 0.999, 1e-6)] * 3
    _, pc, _ = replay(oracle, [x.numel() for x in oi], oi, props, 1, 3, dtype, 2e-3, 64.0, 1.0, 1, False,
                      lambda it, r, i: 1000 * it + i, e5m2=(option == "e5m2_allgather"),
                      clip_after_ar=(option != "clip_before_ar"))
    if option == "e5m2_allgather":
        # parameters carry two mantissa bits: compare the packed bytes (one e5m2 step = the reference's own granularity)
        for a, b in zip(order, pc):
            pa = oracle.convert(a.detach().cpu().reshape(-1).float(), torch.uint8).to(torch.int32)
            pb = oracle.convert(b.float(), torch.uint8).to(torch.int32)
            assert torch.equal(oracle.convert(oracle.convert(a.detach().cpu().reshape(-1).float(), torch.uint8), dtype),
                               a.detach().cpu().reshape(-1))       # really e5m2-representable
            assert int((pa - pb).abs().max()) <= 1 and float((pa != pb).float().mean()) < 0.01
    else:
        _check_vs_replay(order, pc, dtype)
