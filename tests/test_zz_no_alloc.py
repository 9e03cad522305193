"""mt_multi_tensor_l2norm_noalloc -- the L2-norm entry point for callers that cannot allocate device memory.

Why it exists: the regular op needs a scratch buffer from the caller (and cudaMalloc()s its tensor table on a cache
miss).  A host bound to the bare C ABI from another runtime has no device allocator to make that scratch with; a
process that handed all of HBM to a single pool at start-up would see cudaMalloc fail; and a latency-critical step
cannot afford cudaMalloc's device-wide synchronisation.  The variant keeps its working memory in a fixed array inside
the library instead, so it has size limits and must not overlap with itself on the GPU.

This file is named so that pytest collects it last: if it fails, a `-x` run has already executed every other test.
It drives the C ABI directly (ctypes), exactly as such a host would, and checks the new entry point against the
existing one on two calls in a row -- the second deliberately smaller, so stale contents of the shared array from the
first would show up.  Tolerance: the norm tests' 1e-5 relative; the flag must be identical.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _call_pair(L, _lib, xs, chunk, per_tensor=True):
    tl = _lib.Lists([xs])
    n = len(xs)
    f1 = torch.zeros(1, dtype=torch.int32, device="cuda")
    f2 = torch.zeros(1, dtype=torch.int32, device="cuda")
    r1, p1 = torch.full((1,), -1.0, device="cuda"), torch.full((n,), -1.0, device="cuda")
    r2, p2 = torch.full((1,), -2.0, device="cuda"), torch.full((n,), -2.0, device="cuda")
    nbytes = L.mt_norm_scratch_bytes(chunk, tl.ref)
    scratch = torch.empty(max(nbytes, 4), dtype=torch.uint8, device="cuda")
    _lib.check(L.mt_multi_tensor_l2norm(chunk, _lib.vp(f1), tl.ref, int(per_tensor), _lib.vp(r1), _lib.vp(p1),
                                        _lib.vp(scratch), C.c_size_t(nbytes), _lib.stream_ptr()))
    _lib.check(L.mt_multi_tensor_l2norm_noalloc(chunk, _lib.vp(f2), tl.ref, int(per_tensor), _lib.vp(r2), _lib.vp(p2),
                                                _lib.stream_ptr()))
    torch.cuda.synchronize()
    return (r1.cpu(), p1.cpu(), int(f1.item())), (r2.cpu(), p2.cpu(), int(f2.item()))


def test_noalloc_l2norm_matches_regular_entry_point_twice_in_a_row():
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    assert 0 < L.mt_noalloc_resident_bytes() < 4 * 1024 * 1024  # "modest": the array every GPU carries
    g = torch.Generator().manual_seed(7)
    # call 1: many chunks (bf16 grads, chunk 8192); call 2: fewer tensors, fewer chunks, fp32, with an inf
    first = [torch.randn(n, generator=g).bfloat16().cuda() for n in (300001, 1, 0, 70000, 8192 * 5)]
    second = [torch.randn(n, generator=g).cuda() for n in (1234, 65536 + 1)]
    (ra, pa, fa), (rb, pb, fb) = _call_pair(L, _lib, first, 8192)
    assert fa == fb == 0
    assert torch.allclose(rb, ra, rtol=1e-5, atol=0) and torch.allclose(pb, pa, rtol=1e-5, atol=0)
    exact = torch.stack([t.double().norm() for t in first]).float().cpu()
    assert torch.allclose(pb, exact, rtol=1e-5, atol=0)

    second[1][40000] = float("inf")
    (ra, pa, fa), (rb, pb, fb) = _call_pair(L, _lib, second, 65536)
    assert fa == fb == 1
    assert torch.isinf(rb).all() and torch.isinf(ra).all()
    assert torch.allclose(pb[:1], pa[:1], rtol=1e-5, atol=0) and torch.isinf(pb[1]) and torch.isinf(pa[1])

    # and once more without the inf and without per-tensor output: global norm only
    second[1][40000] = 0.5
    (ra, _, fa), (rb, _, fb) = _call_pair(L, _lib, second, 65536, per_tensor=False)
    assert fa == fb == 0 and torch.allclose(rb, ra, rtol=1e-5, atol=0)


def test_noalloc_l2norm_refuses_sizes_beyond_its_built_in_limits():
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    x = torch.zeros(1 << 20, device="cuda")
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    r = torch.zeros(1, device="cuda")
    launches = L.mt_launch_count()
    tl = _lib.Lists([[x]])
    rc = L.mt_multi_tensor_l2norm_noalloc(1, _lib.vp(f), tl.ref, 0, _lib.vp(r), None, _lib.stream_ptr())  # 2^20 chunks
    assert rc == -1 and b"MT_NOALLOC_MAX_CHUNKS" in L.mt_last_error()
    many = [x[i:i + 1] for i in range(16385)]
    tl = _lib.Lists([many])
    rc = L.mt_multi_tensor_l2norm_noalloc(65536, _lib.vp(f), tl.ref, 0, _lib.vp(r), None, _lib.stream_ptr())
    assert rc == -1 and b"MT_NOALLOC_MAX_TENSORS" in L.mt_last_error()
    assert L.mt_launch_count() == launches  # nothing was launched


def test_noalloc_l2norm_allocates_nothing_and_refuses_a_capturing_stream():
    """Backs two statements of the header: the entry point makes no device allocation of its own (free memory as seen
    by the driver is the same before and after a series of calls, also for list shapes it has not met before), and a
    stream that is being captured is turned away with MT_ECAPTURE -- without spoiling the capture -- because the
    table is staged through one host buffer that every later call overwrites."""
    from beforeholiday_b200 import _lib

    L = _lib.lib()
    g = torch.Generator().manual_seed(11)
    sets = [[torch.randn(n, generator=g).cuda() for n in ns] for ns in ((5000, 3), (70001,), (1, 2, 3, 4, 100000))]
    f = torch.zeros(1, dtype=torch.int32, device="cuda")
    r = torch.zeros(1, device="cuda")
    pts = [torch.zeros(len(xs), device="cuda") for xs in sets]
    tls = [_lib.Lists([xs]) for xs in sets]
    _lib.check(L.mt_multi_tensor_l2norm_noalloc(4096, _lib.vp(f), tls[0].ref, 1, _lib.vp(r), _lib.vp(pts[0]),
                                                _lib.stream_ptr()))  # first call on the device: address lookup
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        for tl, pt, xs in zip(tls, pts, sets):
            _lib.check(L.mt_multi_tensor_l2norm_noalloc(4096, _lib.vp(f), tl.ref, 1, _lib.vp(r), _lib.vp(pt),
                                                        _lib.stream_ptr()))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "the no-alloc entry point changed the device's free memory"
    for pt, xs in zip(pts, sets):
        exact = torch.stack([t.double().norm() for t in xs]).float()
        assert torch.allclose(pt, exact, rtol=1e-5, atol=0)

    # on a capturing stream: refused, nothing enqueued, the capture survives and replays what WAS recorded
    y = torch.zeros(4, device="cuda")
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    launches = L.mt_launch_count()
    with torch.cuda.stream(s):
        graph.capture_begin()
        y.add_(1.0)
        rc = L.mt_multi_tensor_l2norm_noalloc(4096, _lib.vp(f), tls[1].ref, 1, _lib.vp(r), _lib.vp(pts[1]),
                                              C.c_void_p(s.cuda_stream))
        graph.capture_end()
    assert rc == -3 and b"captured" in L.mt_last_error()
    assert L.mt_launch_count() == launches
    graph.replay()
    torch.cuda.synchronize()
    assert float(y[0]) == 1.0
    # a different call afterwards, eagerly, still works
    _lib.check(L.mt_multi_tensor_l2norm_noalloc(4096, _lib.vp(f), tls[2].ref, 1, _lib.vp(r), _lib.vp(pts[2]),
                                                _lib.stream_ptr()))
    torch.cuda.synchronize()
    assert torch.allclose(pts[2], torch.stack([t.double().norm() for t in sets[2]]).float(), rtol=1e-5, atol=0)
