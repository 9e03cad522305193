"""Parity against the REFERENCE'S OWN CUDA kernels ("O-A" in SURVEY 8c): oracle/_ref/amp_C*.so is the reference's
csrc/amp_C_frontend.cpp + multi_tensor_*.cu compiled unmodified for sm_100a by oracle/build_ref.py.  Both modules run
on the same device buffers' clones.

Tolerances (north_star): flags bit-exact; fp32 updates within 1e-6 relative (the reference build is --use_fast_math:
approximate div/sqrt, FTZ, FMA -- a few ulp from IEEE); 16-bit outputs within 1 ulp; norms within 1e-5 relative.
These tests also pin the CPU oracle against the real reference (oracle vs reference at the same tolerances).
They calls the C ABI directly through ctypes as well as through amp_C.
"""
import ctypes as C
import importlib.machinery
import importlib.util
import sysconfig
from pathlib import Path

import pytest
import torch

from util import bits_equal, cpu_clone, max_rel, rel_close, ulp_diff_16

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent
F32, F16, BF16 = torch.float32, torch.float16, torch.bfloat16


def _load_ref(name):
    p = ROOT / "oracle" / "_ref" / (name + sysconfig.get_config_var("EXT_SUFFIX"))
    if not p.exists():
        pytest.skip(f"{p} not built (oracle/build_ref.py runs in the build container only)")
    loader = importlib.machinery.ExtensionFileLoader(name, str(p))
    spec = importlib.util.spec_from_loader(name, loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    return mod


@pytest.fixture(scope="module")
def ref_amp_C():
    return _load_ref("amp_C")


@pytest.fixture(scope="module")
def ref_dist_adam():
    return _load_ref("distributed_adam_cuda")


def rnd(n, dtype, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n, generator=g) * scale).to(dtype).cuda()


def clone(lists):
    return [[t.clone() for t in l] for l in lists]


def flag():
    return torch.zeros(1, dtype=torch.int32, device="cuda")


SIZES = [1, 9, 4097, 70001, 278011, 1 << 20]


def check_close(a, b, what):
    if a.dtype == F32:
        # atol: an absolute floor for values that are differences of larger quantities (cancellation)
        assert rel_close(a, b, 1e-6, 1e-7 * float(b.abs().max().clamp_min(1e-30))), f"{what}: max rel {max_rel(a, b)}"
    else:
        # 16-bit outputs: within 1 ulp.  Where the fp32 value being rounded is itself a near-total cancellation (e.g.
        # beta1*m + (1-beta1)*g with opposite signs) its fp32 rounding error -- FMA vs separate multiply/add -- can span
        # several ulps of the tiny 16-bit result; those elements are held to an absolute bound of 2^-20 of the tensor's
        # largest magnitude instead (4 fp32 ulps of the operands).
        af, bf_ = a.float().cpu(), b.float().cpu()
        ia = a.cpu().view(torch.int16).to(torch.int32)
        ib = b.cpu().view(torch.int16).to(torch.int32)
        key = lambda i: torch.where(i < 0, -(i & 0x7fff), i)
        ulp = (key(ia) - key(ib)).abs()
        ok = (ulp <= 1) | ((af - bf_).abs() <= float(bf_.abs().max()) * 2.0 ** -20) | (af.isnan() & bf_.isnan())
        assert bool(ok.all()), f"{what}: {int((~ok).sum())} elements beyond 1 ulp, worst {int(ulp[~ok].max())} ulp"


@pytest.mark.parametrize("in_t,out_t", [(F16, F32), (F32, F16), (F32, F32), (BF16, F32), (F32, BF16)])
def test_scale_vs_reference(amp_C, ref_amp_C, in_t, out_t):
    ins = [rnd(n, in_t, i, 100.0) for i, n in enumerate(SIZES)]
    for inject in (None, "inf", "nan"):
        if inject:
            ins[3][777] = float(inject)
        o1 = [torch.zeros(n, dtype=out_t, device="cuda") for n in SIZES]
        o2 = clone([o1])[0]
        f1, f2 = flag(), flag()
        amp_C.multi_tensor_scale(65536, f1, [ins, o1], 1.0 / 65536)
        ref_amp_C.multi_tensor_scale(65536, f2, [ins, o2], 1.0 / 65536)
        assert f1.item() == f2.item() == (1 if inject else 0)
        for a, b in zip(o1, o2):
            assert bits_equal(a, b)  # one exact multiply: must be identical even under fast-math


def test_axpby_vs_reference(amp_C, ref_amp_C):
    x = [rnd(n, F16, i, 10.0) for i, n in enumerate(SIZES)]
    y = [rnd(n, F32, 50 + i) for i, n in enumerate(SIZES)]
    for arg, bad in ((0, "x"), (1, "x"), (-1, "y"), (0, None)):
        xs, ys = clone([x])[0], clone([y])[0]
        if bad == "x":
            xs[2][5] = float("inf")
        if bad == "y":
            ys[2][5] = float("nan")
        o1 = [torch.zeros(n, device="cuda") for n in SIZES]
        o2 = clone([o1])[0]
        f1, f2 = flag(), flag()
        amp_C.multi_tensor_axpby(65536, f1, [xs, ys, o1], 1.0 / 128, 1.0, arg)
        ref_amp_C.multi_tensor_axpby(65536, f2, [xs, ys, o2], 1.0 / 128, 1.0, arg)
        assert f1.item() == f2.item()
        for a, b in zip(o1, o2):
            ok = a.isfinite() & b.isfinite()
            check_close(a[ok], b[ok], "axpby")


@pytest.mark.parametrize("dt", [F32, F16, BF16])
@pytest.mark.parametrize("chunk", [65536, 8192])
def test_l2norm_vs_reference(amp_C, ref_amp_C, oracle, dt, chunk):
    xs = [rnd(n, dt, i) for i, n in enumerate(SIZES)]
    f1, f2 = flag(), flag()
    n1, pt1 = amp_C.multi_tensor_l2norm(chunk, f1, [xs], True)
    n2, pt2 = ref_amp_C.multi_tensor_l2norm(chunk, f2, [xs], True)
    assert f1.item() == f2.item() == 0
    assert max_rel(n1, n2) < 1e-5 and max_rel(pt1, pt2) < 1e-5
    fo = torch.zeros(1, dtype=torch.int32)
    no, pto = oracle.multi_tensor_l2norm(chunk, fo, [[x.cpu() for x in xs]], True)
    assert max_rel(no, n2) < 1e-5 and max_rel(pto, pt2) < 1e-5   # pins the oracle against the real reference
    xs[1][3] = float("inf")
    amp_C.multi_tensor_l2norm(chunk, f1, [xs], False)
    ref_amp_C.multi_tensor_l2norm(chunk, f2, [xs], False)
    assert f1.item() == f2.item() == 1


def test_l2norm_scale_vs_reference(amp_C, ref_amp_C):
    xs = [rnd(n, F32, i) for i, n in enumerate(SIZES)]
    o1 = [torch.zeros(n, dtype=F16, device="cuda") for n in SIZES]
    o2 = clone([o1])[0]
    f1, f2 = flag(), flag()
    n1, pt1 = amp_C.multi_tensor_l2norm_scale(65536, f1, [xs, o1], 0.5, True)
    n2, pt2 = ref_amp_C.multi_tensor_l2norm_scale(65536, f2, [xs, o2], 0.5, True)
    assert max_rel(n1, n2) < 1e-5 and max_rel(pt1, pt2) < 1e-5 and f1.item() == f2.item()
    assert all(bits_equal(a, b) for a, b in zip(o1, o2))


@pytest.mark.parametrize("dt", [F32, F16, BF16])
@pytest.mark.parametrize("mode,wd", [(0, 0.0), (0, 0.1), (1, 0.01)])
def test_adam_vs_reference(amp_C, ref_amp_C, oracle, dt, mode, wd):
    lists = [[rnd(n, dt, 10 * d + i, 0.1) for i, n in enumerate(SIZES)] for d in range(4)]
    lists[3] = [t.abs() for t in lists[3]]
    l1, l2, lo = clone(lists), clone(lists), cpu_clone(lists)
    args = (5e-4, 0.9, 0.999, 1e-8, 5, mode, 1, wd)
    # fp32: a few steps.  16-bit state: one step -- a 1-ulp difference stored in 16-bit m / v legitimately feeds back
    # into later steps, so only the single-step result can be held to 1 ulp.
    for _ in range(3 if dt == F32 else 1):
        amp_C.multi_tensor_adam(8192, flag(), l1, *args)
        ref_amp_C.multi_tensor_adam(8192, flag(), l2, *args)
        oracle.multi_tensor_adam(8192, torch.zeros(1, dtype=torch.int32), lo, *args)
    for d in (1, 2, 3):
        for a, b, c in zip(l1[d], l2[d], lo[d]):
            check_close(a, b, f"adam list {d}")
            check_close(c.cuda(), b, f"oracle vs reference, adam list {d}")


@pytest.mark.parametrize("g_t,w_t,copy", [(F32, F32, False), (F16, F32, True), (F16, F16, False), (BF16, F32, True)])
def test_sgd_vs_reference(amp_C, ref_amp_C, g_t, w_t, copy):
    dts = [g_t, w_t, w_t] + ([F16] if copy else [])
    lists = [[rnd(n, dt, 10 * d + i, 0.5) for i, n in enumerate(SIZES)] for d, dt in enumerate(dts)]
    l1, l2 = clone(lists), clone(lists)
    for first_run in (True, False):
        args = (1e-4, 0.9, 0.0, 0.1, True, first_run, False, 1.0 / 128)
        amp_C.multi_tensor_sgd(8192, flag(), l1, *args)
        ref_amp_C.multi_tensor_sgd(8192, flag(), l2, *args)
    for d in range(1, len(dts)):
        for a, b in zip(l1[d], l2[d]):
            check_close(a, b, f"sgd list {d}")
    # noop flag honoured by both
    f1, f2 = flag() + 1, flag() + 1
    before = clone(l1)
    amp_C.multi_tensor_sgd(8192, f1, l1, *args)
    ref_amp_C.multi_tensor_sgd(8192, f2, l2, *args)
    assert all(bits_equal(a, b) for a, b in zip(l1[1], before[1]))


@pytest.mark.parametrize("dt", [F32, F16])
@pytest.mark.parametrize("wd,use_nvlamb", [(0.01, False), (0.0, True), (0.0, False)])
def test_lamb_vs_reference(amp_C, ref_amp_C, dt, wd, use_nvlamb):
    lists = [[rnd(n, dt, 10 * d + i, 0.1) for i, n in enumerate(SIZES)] for d in range(4)]
    lists[3] = [t.abs() for t in lists[3]]
    l1, l2 = clone(lists), clone(lists)
    f = flag()
    gn, _ = ref_amp_C.multi_tensor_l2norm(65536, f, [lists[0]], False)
    args = (5e-4, 0.9, 0.999, 1e-6, 3, 1, wd, 1, 1, gn, 1.0, use_nvlamb)
    amp_C.multi_tensor_lamb(8192, flag(), l1, *args)
    ref_amp_C.multi_tensor_lamb(8192, flag(), l2, *args)
    for d in range(4):
        for a, b in zip(l1[d], l2[d]):
            check_close(a, b, f"lamb list {d}")


@pytest.mark.parametrize("g_t", [F32, F16, BF16])
@pytest.mark.parametrize("mode", [0, 1])
def test_dist_adam_vs_reference(amp_C, ref_dist_adam, g_t, mode):
    dts = [F32, F32, F32, g_t, g_t]
    lists = [[rnd(n, dt, 10 * d + i, 0.05) for i, n in enumerate(SIZES)] for d, dt in enumerate(dts)]
    lists[2] = [t.abs() for t in lists[2]]
    l1, l2 = clone(lists), clone(lists)
    n = len(SIZES)
    hp = [torch.tensor([v] * n, dtype=dt, device="cuda") for v, dt in ((0.9, F32), (0.95, F32), (1, torch.int32), (1e-8, F32), (0.1, F32))]
    for step in (1, 2):
        amp_C.multi_tensor_fused_adam(8192, flag(), l1, *hp, 1e-3, 1024.0, step, mode)
        ref_dist_adam.multi_tensor_fused_adam(8192, flag(), l2, *hp, 1e-3, 1024.0, step, mode)
    for d in (0, 1, 2, 4):
        for a, b in zip(l1[d], l2[d]):
            check_close(a, b, f"dist adam list {d}")


def test_c_abi_direct_ctypes(oracle):
    """Drive libmt_b200.so with raw device pointers (what a non-torch host binds): scale, adam and l2norm."""
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    launches0 = L.mt_launch_count()
    x = [rnd(70001, F16, 1, 100.0), rnd(9, F16, 2)]
    o = [torch.zeros(70001, device="cuda"), torch.zeros(9, device="cuda")]
    f = flag()
    tl = _lib.Lists([x, o])
    _lib.check(L.mt_multi_tensor_scale(8192, _lib.vp(f), tl.ref, C.c_float(0.25), _lib.stream_ptr()))
    xc, oc = [t.cpu() for t in x], [torch.zeros(70001), torch.zeros(9)]
    oracle.multi_tensor_scale(8192, torch.zeros(1, dtype=torch.int32), [xc, oc], 0.25)
    assert all(bits_equal(a, b) for a, b in zip(o, oc)) and f.item() == 0

    lists = [[rnd(33333, F32, 10 + d, 0.1)] for d in range(4)]
    lists[3] = [lists[3][0].abs()]
    lc = cpu_clone(lists)
    tl = _lib.Lists(lists)
    _lib.check(L.mt_multi_tensor_adam(8192, _lib.vp(f), tl.ref, C.c_float(1e-3), C.c_float(0.9), C.c_float(0.999),
                                      C.c_float(1e-8), 2, 1, 1, C.c_float(0.01), _lib.stream_ptr()))
    oracle.multi_tensor_adam(8192, None, lc, 1e-3, 0.9, 0.999, 1e-8, 2, 1, 1, 0.01)
    for d in range(4):
        assert bits_equal(lists[d][0], lc[d][0])

    xs = [rnd(300001, F32, 20), rnd(5, F32, 21)]
    tl = _lib.Lists([xs])
    nbytes = L.mt_norm_scratch_bytes(65536, tl.ref)
    assert nbytes == 4 * (5 + 1)
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    ret, ret_pt = torch.empty(1, device="cuda"), torch.empty(2, device="cuda")
    _lib.check(L.mt_multi_tensor_l2norm(65536, _lib.vp(f), tl.ref, 1, _lib.vp(ret), _lib.vp(ret_pt), _lib.vp(scratch),
                                        C.c_size_t(nbytes), _lib.stream_ptr()))
    assert max_rel(ret, torch.cat([t.double() for t in xs]).norm().reshape(1)) < 1e-5
    assert max_rel(ret_pt, torch.stack([t.double().norm() for t in xs])) < 1e-5
    # scratch too small -> MT_ESCRATCH with a message, nothing launched
    rc = L.mt_multi_tensor_l2norm(65536, _lib.vp(f), tl.ref, 1, _lib.vp(ret), _lib.vp(ret_pt), _lib.vp(scratch),
                                  C.c_size_t(4), _lib.stream_ptr())
    assert rc == -4 and b"scratch too small" in L.mt_last_error()
    assert L.mt_launch_count() - launches0 == 1 + 1 + 1      # scale, adam, l2norm: one launch each


def test_c_abi_rejects_bad_entries_then_serves_cached_sets():
    """Per-entry validation (NULL pointer, negative numel) happens when a table is built; a set that is already cached
    is served from the most-recently-used list without being re-walked."""
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    L.mt_last_error.restype = C.c_char_p
    x = [rnd(5000, F32, 1), rnd(7, F32, 2)]
    o = [torch.zeros(5000, device="cuda"), torch.zeros(7, device="cuda")]
    f = flag()
    tl = _lib.Lists([x, o])
    tl.ptr_arr[3] = None                                    # list 1, tensor 1
    rc = L.mt_multi_tensor_scale(8192, _lib.vp(f), tl.ref, C.c_float(1.0), _lib.stream_ptr())
    assert rc != 0 and b"NULL pointer" in L.mt_last_error()
    tl.ptr_arr[3] = o[1].data_ptr()
    tl.numel_arr[0] = -5
    rc = L.mt_multi_tensor_scale(8192, _lib.vp(f), tl.ref, C.c_float(1.0), _lib.stream_ptr())
    assert rc != 0 and b"negative numel" in L.mt_last_error()
    tl.numel_arr[0] = 5000
    st0 = _lib.cache_stats(0)
    for _ in range(3):
        _lib.check(L.mt_multi_tensor_scale(8192, _lib.vp(f), tl.ref, C.c_float(2.0), _lib.stream_ptr()))
    st1 = _lib.cache_stats(0)
    assert st1["misses"] - st0["misses"] == 1 and st1["hits"] - st0["hits"] == 2
    assert torch.equal(o[0], x[0] * 2) and torch.equal(o[1], x[1] * 2)
