"""Out-of-bounds WRITE detection without compute-sanitizer (closed on this pool): every tensor of every list is carved
out of one arena with sentinel bands on both sides, at element offsets that exercise the vector path, the head/tail
peel and the mutually-misaligned scalar fallback; after each op every band must still hold the sentinel bit pattern,
and read-only lists must be bit-identical to what was put in."""
import pytest
import torch

pytestmark = pytest.mark.gpu

F32, F16, BF16 = torch.float32, torch.float16, torch.bfloat16
SIZES = [1, 7, 8, 9, 255, 2048, 2049, 4097, 16384, 16385, 70001]
BAND = 40            # elements; not a multiple of 8 so that tensors start at odd alignments
SENT = 1234.5        # compared by bits after rounding to the arena dtype


class Arena:
    def __init__(self, dtype, sizes, skew):
        self.dtype, self.sizes = dtype, sizes
        total = BAND + sum(n + BAND + skew for n in sizes) + 64
        self.buf = torch.full((total,), SENT, dtype=dtype, device="cuda")
        self.views, self.bands = [], []
        o = BAND + skew
        self.bands.append((0, o))
        for n in sizes:
            self.views.append(self.buf[o:o + n])
            nxt = o + n + BAND + skew
            self.bands.append((o + n, min(nxt, total)))
            o = nxt
        self.bands.append((o, total))
        self.sent_bits = torch.full((1,), SENT, dtype=dtype).view(torch.int16 if dtype != F32 else torch.int32)[0].item()

    def fill(self, seed, scale=0.1, positive=False):
        g = torch.Generator(device="cuda").manual_seed(seed)
        for v in self.views:
            x = torch.randn(v.numel(), device="cuda", generator=g) * scale
            v.copy_((x.abs() if positive else x).to(self.dtype))
        return self

    def check(self, what):
        bits = self.buf.view(torch.int16 if self.dtype != F32 else torch.int32)
        for lo, hi in self.bands:
            if hi > lo:
                bad = (bits[lo:hi] != self.sent_bits).nonzero()
                assert bad.numel() == 0, f"{what}: guard band [{lo},{hi}) overwritten at {int(bad[0]) + lo}"


def _run(op, dtypes, skews, read_only=(), positive=()):
    arenas = [Arena(dt, SIZES, sk).fill(10 + d, positive=d in positive) for d, (dt, sk) in enumerate(zip(dtypes, skews))]
    before = {d: [v.clone() for v in arenas[d].views] for d in read_only}
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    op(flag, [a.views for a in arenas])
    torch.cuda.synchronize()
    for d, a in enumerate(arenas):
        a.check(f"list {d}")
    for d, old in before.items():
        for a, b in zip(arenas[d].views, old):
            assert torch.equal(a.view(torch.int16 if a.dtype != F32 else torch.int32),
                               b.view(torch.int16 if b.dtype != F32 else torch.int32)), f"read-only list {d} modified"
    return flag


SKEWS = [(0, 0, 0, 0, 0), (1, 1, 1, 1, 1), (3, 3, 3, 3, 3), (1, 2, 3, 4, 5), (0, 4, 0, 4, 0), (7, 7, 7, 7, 7)]


@pytest.mark.parametrize("skews", SKEWS)
def test_scale_axpby_cast_stay_in_bounds(amp_C, skews):
    _run(lambda f, l: amp_C.multi_tensor_scale(8192, f, l, 0.5), [F16, F32], skews, read_only=(0,))
    _run(lambda f, l: amp_C.multi_tensor_scale(8192, f, l, 0.5), [F32, BF16], skews, read_only=(0,))
    _run(lambda f, l: amp_C.multi_tensor_axpby(8192, f, l, 0.5, 2.0, 0), [F16, F32, BF16], skews, read_only=(0, 1))
    _run(lambda f, l: amp_C.maybe_cast_mt(8192, f, l), [F32, F16], skews, read_only=(0,))
    _run(lambda f, l: amp_C.multi_tensor_l2norm_scale(8192, f, l, 0.5, True), [F16, F32], skews, read_only=(0,))
    _run(lambda f, l: amp_C.multi_tensor_l2norm(8192, f, l, True), [BF16], skews, read_only=(0,))


@pytest.mark.parametrize("skews", SKEWS)
def test_optimizers_stay_in_bounds(amp_C, skews):
    _run(lambda f, l: amp_C.multi_tensor_adam(8192, f, l, 1e-3, 0.9, 0.999, 1e-8, 2, 1, 1, 0.01), [F32] * 4, skews,
         read_only=(0,), positive=(3,))
    _run(lambda f, l: amp_C.multi_tensor_sgd(8192, f, l, 1e-4, 0.9, 0.0, 0.1, False, False, False, 1.0),
         [F16, F32, F32, F16], skews, read_only=(0,))
    gn = torch.tensor([1.0], device="cuda")
    _run(lambda f, l: amp_C.multi_tensor_lamb(8192, f, l, 1e-3, 0.9, 0.999, 1e-6, 2, 1, 0.01, 1, 1, gn, 1.0, False),
         [F32] * 4, skews, positive=(3,))
    _run(lambda f, l: amp_C.multi_tensor_adagrad(8192, f, l, 1e-2, 1e-10, 0, 0.1), [F16] * 3, skews, read_only=(0,),
         positive=(2,))
    n = len(SIZES)
    T = lambda v, dt=F32: torch.full((n,), v, dtype=dt, device="cuda")
    _run(lambda f, l: amp_C.multi_tensor_fused_adam(8192, f, l, T(0.9), T(0.95), T(1, torch.int32), T(1e-8), T(0.1), 1e-3,
                                                    1.0, 2, 1, True),
         [F32, F32, F32, BF16, BF16], skews, read_only=(3,), positive=(2,))


@pytest.mark.parametrize("skews", SKEWS)
def test_sharded_lamb_stays_in_bounds(amp_C, skews):
    n = len(SIZES)
    T = lambda v, dt=F32: torch.full((n,), v, dtype=dt, device="cuda")
    one = lambda v, dt=F32: torch.tensor([v], dtype=dt, device="cuda")
    seg = torch.arange(n + 1, dtype=torch.int32, device="cuda")
    _run(lambda f, l: amp_C.multi_tensor_lamb_compute_update_term(8192, f, l, T(0.9), T(0.999), T(0.1), T(1, torch.int32),
                                                                  one(2, torch.int32), T(1e-6), 1, T(0.01), one(1.0), one(1.0),
                                                                  1.0, seg, F16),
         [F16, F32, F32, F32, F32], skews, read_only=(0, 1), positive=(3,))
    _run(lambda f, l: amp_C.multi_tensor_lamb_update_weights(8192, f, l, T(1.0), T(2.0), torch.arange(n, device="cuda"),
                                                             one(1e-3), T(0.01), one(1.0), True),
         [F32, F32, BF16], skews, read_only=(0,))
