"""GPU parity: every amp_C op on CUDA tensors against the CPU oracle on clones of the same inputs.

Bar (north_star): overflow flags bit-exact; element-wise fp32 results bit-exact with the oracle (same IEEE operation
order, no FMA) -- stronger than the 1e-6 the spec asks for; 16-bit outputs bit-exact (<= 1 ulp asked); norms <= 1e-5
relative (summation order differs).  All calls go amp_C -> C ABI (include/mt_b200.h) -> kernels.
"""
import itertools

import pytest
import torch

from util import bits_equal, cpu_clone, max_rel, n_mismatch

pytestmark = pytest.mark.gpu

F32, F16, BF16, F64 = torch.float32, torch.float16, torch.bfloat16, torch.float64
DT3 = [F32, F16, BF16]

# sizes chosen to hit: empty, sub-vector, exactly one group, tile edges (4096 / 16384), odd tails, multi-tile
SIZES = [0, 1, 7, 8, 9, 4095, 4096, 4097, 16384 + 5, 278011, 1 << 20]


def rnd(n, dtype, seed, scale=1.0, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n, generator=g) * scale).to(dtype).to(device)


def make_lists(dtypes, sizes, seed=0, scale=1.0, offsets=None):
    """One list per dtype.  offsets[d] (elements) slices every tensor of list d out of a larger buffer so that its
    address is misaligned by that many elements."""
    lists = []
    for d, dt in enumerate(dtypes):
        off = 0 if offsets is None else offsets[d]
        l = []
        for i, n in enumerate(sizes):
            base = rnd(n + off, dt, seed * 1000 + d * 100 + i, scale)
            l.append(base[off:off + n])
        lists.append(l)
    return lists


def to_gpu(lists, offsets=None):
    out = []
    for d, l in enumerate(lists):
        off = 0 if offsets is None else offsets[d]
        gl = []
        for t in l:
            buf = torch.empty(t.numel() + off, dtype=t.dtype, device="cuda")
            v = buf[off:off + t.numel()]
            v.copy_(t)
            gl.append(v)
        out.append(gl)
    return out


def run_both(amp_C, oracle, name, lists_cpu, args, chunk=2048 * 32, offsets=None, flag_init=0, check=None):
    lists_gpu = to_gpu(lists_cpu, offsets)
    lists_cpu = cpu_clone(lists_cpu)
    fg = torch.full((1,), flag_init, dtype=torch.int32, device="cuda")
    fc = torch.full((1,), flag_init, dtype=torch.int32)
    rg = getattr(amp_C, name)(chunk, fg, lists_gpu, *args)
    rc = getattr(oracle, name)(chunk, fc, lists_cpu, *args)
    torch.cuda.synchronize()
    assert fg.item() == fc.item(), f"{name}: flag {fg.item()} != oracle {fc.item()}"
    for d in (check if check is not None else range(len(lists_cpu))):
        for i, (a, b) in enumerate(zip(lists_gpu[d], lists_cpu[d])):
            assert bits_equal(a, b), f"{name}: list {d} tensor {i} (n={b.numel()}): {n_mismatch(a, b)} mismatches"
    return rg, rc


# ------------------------------------------------------------------------------------------------ scale
@pytest.mark.parametrize("in_t,out_t", list(itertools.product(DT3, DT3)))
def test_scale_random(amp_C, oracle, in_t, out_t):
    lists = make_lists([in_t, out_t], SIZES, seed=1, scale=3.0)
    run_both(amp_C, oracle, "multi_tensor_scale", lists, (1.0 / 65536.0,))


@pytest.mark.parametrize("offsets", [(1, 1), (3, 3), (5, 5), (1, 2), (0, 3)])
@pytest.mark.parametrize("in_t,out_t", [(F16, F32), (F32, F32), (F32, BF16)])
def test_scale_misaligned(amp_C, oracle, in_t, out_t, offsets):
    lists = make_lists([in_t, out_t], [5, 4099, 70001], seed=2, offsets=offsets)
    run_both(amp_C, oracle, "multi_tensor_scale", lists, (0.125,), offsets=offsets)


@pytest.mark.parametrize("sizea,sizeb", [(7777 * 77, 555 * 555), (777, 555), (555, 2048 * 32 + 1), (2048 * 32 + 1, 555), (555, 33333)])
@pytest.mark.parametrize("chunk", [2048 * 32, 333, 33333])
@pytest.mark.parametrize("repeat", [1, 55])
@pytest.mark.parametrize("in_t,out_t", [(F16, F32), (F32, F16), (BF16, F32), (F32, F32)])
def test_scale_reference_grid(amp_C, sizea, sizeb, chunk, repeat, in_t, out_t):
    # /root/reference/tests/L0/run_amp/test_multi_tensor_scale.py:36-124 (known answers, flag semantics)
    if repeat == 55 and sizea * repeat > 10 ** 7:
        pytest.skip("bounded")
    val = 4.0
    a = torch.full((sizea,), val, dtype=in_t, device="cuda")
    b = torch.full((sizeb,), val, dtype=in_t, device="cuda")
    ins = [a.clone() if i % 2 == 0 else b.clone() for i in range(2 * repeat)]
    outs = [torch.zeros_like(t, dtype=out_t) for t in ins]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    amp_C.multi_tensor_scale(chunk, f, [ins, outs], 1.0 / val)
    assert all(bool((o == 1.0).all()) for o in outs) and f.item() == 0
    # in place
    if in_t == out_t:
        amp_C.multi_tensor_scale(chunk, f, [ins, ins], 1.0 / val)
        assert all(bool((o == 1.0).all()) for o in ins) and f.item() == 0
        for t in ins:
            t.fill_(val)
    for t, ind, v in ((0, 0, float("nan")), (2 * repeat - 1, sizeb - 1, float("inf")), (2 * (repeat // 2), sizea // 2, float("inf"))):
        f.zero_()
        ins[t][ind] = v
        amp_C.multi_tensor_scale(chunk, f, [ins, outs], 1.0 / val)
        assert f.item() == 1
        ins[t][ind] = val


# ------------------------------------------------------------------------------------------------ axpby
@pytest.mark.parametrize("x_t,y_t,o_t", list(itertools.product(DT3, DT3, DT3)))
def test_axpby_random(amp_C, oracle, x_t, y_t, o_t):
    lists = make_lists([x_t, y_t, o_t], [0, 3, 4097, 100003], seed=3)
    run_both(amp_C, oracle, "multi_tensor_axpby", lists, (2.0, 8.0, -1))


@pytest.mark.parametrize("arg_to_check", [-1, 0, 1])
@pytest.mark.parametrize("which", ["x", "y"])
def test_axpby_flag(amp_C, oracle, arg_to_check, which):
    lists = make_lists([F16, F32, F32], [9000, 33], seed=4)
    (lists[0] if which == "x" else lists[1])[0][4321] = float("inf")
    run_both(amp_C, oracle, "multi_tensor_axpby", lists, (1.0 / 128, 1.0, arg_to_check))


def test_axpby_reference_kat(amp_C):
    # test_multi_tensor_axpby.py:39,75-80
    for x_t, y_t, o_t in itertools.product(DT3, DT3, DT3):
        x = [torch.full((n,), 4.0, dtype=x_t, device="cuda") for n in (7777 * 7, 555)]
        y = [torch.full((n,), 16.0, dtype=y_t, device="cuda") for n in (7777 * 7, 555)]
        o = [torch.zeros(n, dtype=o_t, device="cuda") for n in (7777 * 7, 555)]
        f = torch.zeros(1, dtype=torch.int32, device="cuda")
        amp_C.multi_tensor_axpby(333, f, [x, y, o], 2.0, 8.0, -1)
        assert all(bool((t == 136.0).all()) for t in o) and f.item() == 0


# ------------------------------------------------------------------------------------------------ l2norm
@pytest.mark.parametrize("dt", DT3)
@pytest.mark.parametrize("chunk", [2048 * 32, 8192, 333, 33333])
def test_l2norm(amp_C, oracle, dt, chunk):
    sizes = [0, 1, 555, 7777, 2048 * 32 + 1, 300001]
    lists = make_lists([dt], sizes, seed=5)
    (ng, ptg), (nc, ptc) = run_both(amp_C, oracle, "multi_tensor_l2norm", lists, (True,), chunk=chunk)
    exact = torch.stack([x.double().norm() for x in lists[0]])
    assert max_rel(ng, nc) < 1e-5 and max_rel(ng, torch.cat([x.double() for x in lists[0]]).norm().reshape(1)) < 1e-5
    assert torch.allclose(ptg.cpu().double(), exact, rtol=1e-5, atol=0)
    assert torch.allclose(ptg.cpu(), ptc, rtol=1e-5, atol=0)
    (ng2, pt2), _ = run_both(amp_C, oracle, "multi_tensor_l2norm", lists, (False,), chunk=chunk)
    assert pt2.numel() == 0 and bits_equal(ng2, ng)  # deterministic: same order with or without per_tensor


def test_l2norm_reference_kat(amp_C):
    # test_multi_tensor_l2norm.py:35-56
    for dt, chunk, per_tensor in itertools.product([F32, F16], [2048 * 32, 333, 33333], [False, True]):
        xs = [torch.full((n,), 4.0, dtype=dt, device="cuda") for n in (7777 * 7, 555 * 5) * 3]
        f = torch.zeros(1, dtype=torch.int32, device="cuda")
        norm, norm_pt = amp_C.multi_tensor_l2norm(chunk, f, [xs], per_tensor)
        ref = torch.cat([x.float() for x in xs]).norm()
        assert torch.allclose(norm, ref.reshape(1))
        if per_tensor:
            assert torch.allclose(norm_pt, torch.stack([x.float().norm() for x in xs]))
        assert f.item() == 0


def test_l2norm_flag_semantics(amp_C, oracle):
    # inf / nan inputs; finite inputs whose chunk partial overflows; finite partials whose total overflows (no flag)
    for case in ("inf", "nan", "sq_overflow", "total_overflow"):
        x = rnd(70000, F32, 6)
        chunk = 65536
        if case == "inf":
            x[69999] = float("inf")
        elif case == "nan":
            x[0] = float("nan")
        elif case == "sq_overflow":
            x[123] = 3e19
        else:
            x = torch.full((8,), 1.5e19)
            chunk = 1
        (ng, _), (nc, _) = run_both(amp_C, oracle, "multi_tensor_l2norm", [[x]], (False,), chunk=chunk)
        assert bool(ng.isfinite().cpu().item()) == bool(nc.isfinite().item())


def test_l2norm_flag_when_only_the_chunk_sum_overflows(amp_C, oracle):
    # 65536-element chunk whose four 16384-element quarters each sum to 1.6e38 (finite) but whose total is 6.5e38:
    # the reference's chunk partial is inf -> flag.  (One tile per reference chunk is what gives the product the same
    # granularity; a finer tiling would miss this case.)
    x = torch.full((65536 + 100,), 1e17)
    (ng, _), (nc, _) = run_both(amp_C, oracle, "multi_tensor_l2norm", [[x]], (False,), chunk=65536)
    assert torch.isinf(ng).all() and torch.isinf(nc).all()
    y = torch.full((65536 + 100,), 0.4e17)   # chunk sum 1.0e38: finite everywhere -> no flag
    (ng, _), (nc, _) = run_both(amp_C, oracle, "multi_tensor_l2norm", [[y]], (False,), chunk=65536)
    assert max_rel(ng, nc) < 1e-5


@pytest.mark.parametrize("in_t,out_t", [(F32, F32), (F16, F32), (F32, F16), (F16, F16)])
def test_l2norm_scale(amp_C, oracle, in_t, out_t):
    lists = make_lists([in_t, out_t], [0, 5, 4097, 200001], seed=7)
    (ng, ptg), (nc, ptc) = run_both(amp_C, oracle, "multi_tensor_l2norm_scale", lists, (0.25, True), chunk=8192, check=[1])
    assert max_rel(ng, nc) < 1e-5 and torch.allclose(ptg.cpu(), ptc, rtol=1e-5, atol=0)


# ------------------------------------------------------------------------------------------------ adam
@pytest.mark.parametrize("dt", [F32, F16, BF16, F64])
@pytest.mark.parametrize("mode,wd", [(0, 0.0), (0, 0.1), (1, 0.01)])
@pytest.mark.parametrize("bias_correction", [0, 1])
def test_adam(amp_C, oracle, dt, mode, wd, bias_correction):
    sizes = SIZES if dt == F32 else [0, 1, 9, 4097, 70001]
    lists = make_lists([dt] * 4, sizes, seed=8, scale=0.1)
    lists[3] = [t.abs() for t in lists[3]]  # v >= 0
    run_both(amp_C, oracle, "multi_tensor_adam", lists, (5e-4, 0.9, 0.999, 1e-8, 3, mode, bias_correction, wd))


@pytest.mark.parametrize("offsets", [(1,) * 4, (7,) * 4, (1, 2, 3, 4)])
def test_adam_misaligned(amp_C, oracle, offsets):
    lists = make_lists([F32] * 4, [3, 4099, 33333], seed=9, scale=0.1, offsets=offsets)
    lists[3] = [t.abs() for t in lists[3]]
    run_both(amp_C, oracle, "multi_tensor_adam", lists, (1e-3, 0.9, 0.999, 1e-8, 10, 1, 1, 0.01), offsets=offsets)


def test_adam_propagates_nonfinite_and_ignores_flag(amp_C, oracle):
    # the Adam functor neither reads nor writes the flag (multi_tensor_adam.cu:39-41)
    lists = make_lists([F32] * 4, [5000], seed=10, scale=0.1)
    lists[3] = [t.abs() for t in lists[3]]
    lists[0][0][17] = float("inf")
    lists[0][0][18] = float("nan")
    run_both(amp_C, oracle, "multi_tensor_adam", lists, (1e-3, 0.9, 0.999, 1e-8, 1, 0, 1, 0.0), flag_init=1)


def test_adam_many_tensors(amp_C, oracle):
    sizes = [64, 128, 256, 512, 1024, 2048, 768, 3072, 4096, 16384, 36864, 65536, 147456] * 40  # 520 tensors
    lists = make_lists([F32] * 4, sizes, seed=11, scale=0.05)
    lists[3] = [t.abs() for t in lists[3]]
    run_both(amp_C, oracle, "multi_tensor_adam", lists, (1e-3, 0.9, 0.999, 1e-8, 2, 1, 1, 0.01))


# ------------------------------------------------------------------------------------------------ mixed / sharded adam
@pytest.mark.parametrize("g_t", DT3)
@pytest.mark.parametrize("copy", [False, True])
@pytest.mark.parametrize("mode", [0, 1])
def test_fused_adam_mixed(amp_C, oracle, g_t, copy, mode):
    sizes = [0, 1, 9, 4097, 70001, 1 << 20]
    dts = [F32, F32, F32, g_t] + ([g_t] if copy else [])
    lists = make_lists(dts, sizes, seed=12, scale=0.05)
    lists[2] = [t.abs() for t in lists[2]]
    n = len(sizes)
    hp_cpu = ([0.9] * n, [0.95] * n, [1] * n, [1e-8] * n, [0.1 if i % 2 else 0.0 for i in range(n)])
    dev = lambda x, dt: torch.tensor(x, dtype=dt, device="cuda")
    hp_gpu = (dev(hp_cpu[0], F32), dev(hp_cpu[1], F32), dev(hp_cpu[2], torch.int32), dev(hp_cpu[3], F32), dev(hp_cpu[4], F32))
    lists_gpu = to_gpu(lists)
    lists_cpu = cpu_clone(lists)
    fg = torch.zeros(1, dtype=torch.int32, device="cuda")
    fc = torch.zeros(1, dtype=torch.int32)
    amp_C.multi_tensor_fused_adam(8192, fg, lists_gpu, *hp_gpu, 1e-3, 1024.0, 7, mode)
    oracle.multi_tensor_fused_adam(8192, fc, lists_cpu, *hp_cpu, 1e-3, 1024.0, 7, mode)
    for d in range(len(dts)):
        if d == 3:
            continue
        for a, b in zip(lists_gpu[d], lists_cpu[d]):
            assert bits_equal(a, b), f"list {d}: {n_mismatch(a, b)} mismatches"


def test_fused_adam_honours_noop(amp_C):
    lists = to_gpu(make_lists([F32, F32, F32, BF16, BF16], [10000], seed=13))
    before = [l[0].clone() for l in lists]
    hp = [torch.tensor([v], dtype=dt, device="cuda") for v, dt in ((0.9, F32), (0.95, F32), (1, torch.int32), (1e-8, F32), (0.0, F32))]
    f = torch.ones(1, dtype=torch.int32, device="cuda")
    amp_C.multi_tensor_fused_adam(8192, f, lists, *hp, 1e-3, 1.0, 1, 1, True)
    assert all(bits_equal(l[0], b) for l, b in zip(lists, before))
    amp_C.multi_tensor_fused_adam(8192, f, lists, *hp, 1e-3, 1.0, 1, 1, False)  # reference behaviour: flag ignored
    assert not bits_equal(lists[0][0], before[0])


# ------------------------------------------------------------------------------------------------ sgd
SGD_CASES = [(F16, F16, False), (F32, F32, False), (F16, F32, True), (F32, F32, True), (BF16, BF16, False), (BF16, F32, True)]


@pytest.mark.parametrize("g_t,w_t,copy", SGD_CASES)
@pytest.mark.parametrize("momentum,nesterov,first_run,wd,wd_after", [
    (0.125, False, True, 0.0, False), (0.125, False, False, 0.0, False), (0.9, True, False, 1e-4, False),
    (0.9, False, False, 1e-4, True), (0.0, False, True, 1e-4, False)])
def test_sgd(amp_C, oracle, g_t, w_t, copy, momentum, nesterov, first_run, wd, wd_after):
    dts = [g_t, w_t, w_t] + ([F16] if copy else [])
    lists = make_lists(dts, [0, 1, 9, 4097, 70001], seed=14, scale=0.5)
    run_both(amp_C, oracle, "multi_tensor_sgd", lists, (wd, momentum, 0.0, 0.25, nesterov, first_run, wd_after, 1.0 / 128))


def test_sgd_noop_early_exit(amp_C, oracle):
    lists = make_lists([F32, F32, F32], [5000, 17], seed=15)
    run_both(amp_C, oracle, "multi_tensor_sgd", lists, (0.0, 0.9, 0.0, 0.1, False, False, False, 1.0), flag_init=1)


def test_sgd_rejects_unsupported_combo(amp_C):
    lists = to_gpu(make_lists([F32, F16, F16], [100], seed=16))
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    with pytest.raises(RuntimeError, match="only supports some combinations"):
        amp_C.multi_tensor_sgd(8192, f, lists, 0.0, 0.9, 0.0, 0.1, False, False, False, 1.0)


# ------------------------------------------------------------------------------------------------ lamb
@pytest.mark.parametrize("dt", DT3)
@pytest.mark.parametrize("mode,wd,use_nvlamb", [(1, 0.01, False), (1, 0.0, False), (1, 0.0, True), (0, 0.01, True)])
def test_lamb(amp_C, oracle, dt, mode, wd, use_nvlamb):
    sizes = [1, 9, 4097, 70001, 300000]
    lists = make_lists([dt] * 4, sizes, seed=17, scale=0.1)
    lists[3] = [t.abs() for t in lists[3]]
    lists_gpu = to_gpu(lists)
    lists_cpu = cpu_clone(lists)
    gnorm = torch.cat([g.double().reshape(-1) for g in lists[0]]).norm().float().reshape(1)
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    args = (5e-4, 0.9, 0.999, 1e-6, 4, 1, wd, 1, mode)
    amp_C.multi_tensor_lamb(8192, f, lists_gpu, *args, gnorm.cuda(), 1.0, use_nvlamb)
    oracle.multi_tensor_lamb(8192, torch.zeros(1, dtype=torch.int32), lists_cpu, *args, gnorm, 1.0, use_nvlamb)
    # u (stored over g), m, v do not depend on any norm: bit-exact
    for d in (0, 2, 3):
        for a, b in zip(lists_gpu[d], lists_cpu[d]):
            assert bits_equal(a, b), f"list {d}: {n_mismatch(a, b)} mismatches"
    # p depends on the trust ratio = lr * ||p|| / ||u||, whose norms differ in summation order (<= 1e-5 rel)
    for a, b in zip(lists_gpu[1], lists_cpu[1]):
        if dt == F32:
            assert torch.allclose(a.cpu(), b, rtol=1e-6, atol=1e-7)
        else:
            from util import ulp_diff_16
            assert ulp_diff_16(a, b) <= 1


def test_lamb_legacy_stages(amp_C, oracle):
    for g_t, p_t, u_t in ((F32, F32, F32), (F16, F32, F16), (F16, F16, F16), (BF16, F32, F32)):
        sizes = [9, 4097, 70001]
        lists = make_lists([g_t, p_t, p_t, p_t, u_t], sizes, seed=18, scale=0.1)
        lists[3] = [t.abs() for t in lists[3]]
        decay = [0.0, 0.01, 0.1]
        lists_gpu, lists_cpu = to_gpu(lists), cpu_clone(lists)
        f = torch.zeros(1, dtype=torch.int32, device="cuda")
        gn = torch.tensor([2.5])
        amp_C.multi_tensor_lamb_stage1_cuda(8192, f, lists_gpu, torch.tensor(decay, device="cuda"), 3, 0.9, 0.999, 1e-6, gn.cuda(), 1.0)
        oracle.multi_tensor_lamb_stage1_cuda(8192, None, lists_cpu, decay, 3, 0.9, 0.999, 1e-6, gn, 1.0)
        for d in (2, 3, 4):
            for a, b in zip(lists_gpu[d], lists_cpu[d]):
                assert bits_equal(a, b), (g_t, p_t, u_t, d, n_mismatch(a, b))
        pn, un = [1.5, 0.0, 2.0], [0.5, 1.0, 4.0]
        amp_C.multi_tensor_lamb_stage2_cuda(8192, f, [lists_gpu[1], lists_gpu[4]], torch.tensor(pn, device="cuda"),
                                            torch.tensor(un, device="cuda"), 1e-3, 0.01, False)
        oracle.multi_tensor_lamb_stage2_cuda(8192, None, [lists_cpu[1], lists_cpu[4]], pn, un, 1e-3, 0.01, False)
        for a, b in zip(lists_gpu[1], lists_cpu[1]):
            assert bits_equal(a, b)


# ------------------------------------------------------------------------------------------------ boundary behaviour
def test_errors_match_reference_messages(amp_C):
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    a = torch.zeros(10, device="cuda")
    with pytest.raises(RuntimeError, match="expected input to be on cuda"):
        amp_C.multi_tensor_scale(8192, f, [[torch.zeros(10)], [torch.zeros(10)]], 1.0)
    with pytest.raises(RuntimeError, match="Size mismatch"):
        amp_C.multi_tensor_scale(8192, f, [[a], [torch.zeros(11, device="cuda")]], 1.0)
    with pytest.raises(RuntimeError, match="not contiguous"):
        amp_C.multi_tensor_scale(8192, f, [[torch.zeros(10, 10, device="cuda").t()], [torch.zeros(10, 10, device="cuda")]], 1.0)
    with pytest.raises(RuntimeError, match="tensor_lists.size\\(\\) != depth"):
        amp_C.multi_tensor_scale(8192, f, [[a]], 1.0)
    with pytest.raises(RuntimeError, match="is not > 0"):
        amp_C.multi_tensor_scale(8192, f, [[], []], 1.0)


def test_channels_last_runs_over_physical_order(amp_C):
    # multi_tensor_apply.cuh:64-68 accepts channels_last; the op is applied over physical memory order
    x = torch.randn(3, 4, 2, 3, device="cuda").contiguous(memory_format=torch.channels_last)
    o = torch.empty_like(x)
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    amp_C.multi_tensor_scale(8192, f, [[x], [o]], 2.0)
    assert torch.equal(o, x * 2.0)


def test_table_cache_and_graph_capture(amp_C):
    torch.manual_seed(0)
    g = [torch.randn(n, device="cuda", dtype=BF16) for n in (70001, 4097)]
    p = [torch.randn(n, device="cuda") for n in (70001, 4097)]
    m = [torch.zeros_like(t) for t in p]
    v = [torch.zeros_like(t) for t in p]
    pc = [torch.zeros_like(t, dtype=BF16) for t in p]
    hp = [torch.tensor([val] * 2, dtype=dt, device="cuda") for val, dt in ((0.9, F32), (0.95, F32), (1, torch.int32), (1e-8, F32), (0.01, F32))]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    state0 = [[t.clone() for t in l] for l in (p, m, v, pc)]

    def step():
        amp_C.multi_tensor_fused_adam(8192, f, [p, m, v, g, pc], *hp, 1e-3, 1.0, 1, 1, True)

    s0 = amp_C.table_cache_stats(0)
    step()
    s1 = amp_C.table_cache_stats(0)
    step()
    s2 = amp_C.table_cache_stats(0)
    assert s1["misses"] == s0["misses"] + 1 and s2["misses"] == s1["misses"] and s2["hits"] == s1["hits"] + 1
    assert s2["uploads_bytes"] == s1["uploads_bytes"]  # steady state uploads nothing
    torch.cuda.synchronize()
    eager2 = [[t.clone() for t in l] for l in (p, m, v, pc)]

    # same two steps through a captured graph
    for l, l0 in zip((p, m, v, pc), state0):
        for t, t0 in zip(l, l0):
            t.copy_(t0)
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        with torch.cuda.graph(graph, stream=side):
            step()
    torch.cuda.current_stream().wait_stream(side)
    for l, l0 in zip((p, m, v, pc), state0):   # capture does not execute; reset anyway
        for t, t0 in zip(l, l0):
            t.copy_(t0)
    graph.replay()
    graph.replay()
    torch.cuda.synchronize()
    for l, le in zip((p, m, v, pc), eager2):
        for a, b in zip(l, le):
            assert bits_equal(a, b)


def test_capture_miss_is_a_clear_error(amp_C):
    x = [torch.randn(1234, device="cuda")]
    o = [torch.empty(1234, device="cuda")]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    graph = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with pytest.raises(RuntimeError, match="warm-up"):
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                amp_C.multi_tensor_scale(8192, f, [x, o], 2.0)
    torch.cuda.synchronize()
