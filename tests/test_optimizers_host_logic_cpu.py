"""Host logic of FusedSGD / FusedAdam / FusedLAMB on CPU: the Python around the kernels -- list building, first-run
handling, step counters, the steady-state route with its fall-back and its invalidation -- with the CUDA ops replaced by
oracle-backed stand-ins (test infrastructure) that honour the same `grads_of=` contract as the extension.  The launches
themselves are covered by the -m gpu tests; this file keeps the host logic under test where there is no GPU."""
import pytest
import torch

from util import bits_equal


class OracleOps:
    """amp_C look-alike on CPU tensors.  `grads_of=k`: prepend the gradients of tensor_lists[k], refusing exactly what
    the extension refuses (missing / sparse / differently laid out gradient) BEFORE touching anything."""

    def __init__(self, oracle):
        self.O = oracle
        self.calls = []

    def _with_grads(self, lists, grads_of, only=False):
        if grads_of < 0:
            return lists
        grads = []
        for p in lists[grads_of]:
            g = p.grad
            if g is None:
                raise RuntimeError("grads_of: a parameter has no gradient")
            if g.is_sparse:
                raise RuntimeError("grads_of: sparse gradient")
            if g.stride() != p.stride():
                raise RuntimeError("grads_of: gradient layout differs from its parameter")
            grads.append(g)
        return [grads] if only else [grads] + list(lists)

    @staticmethod
    def _raw(lists):
        return [[t.detach() for t in l] for l in lists]

    def multi_tensor_sgd(self, chunk, flag, lists, *args, grads_of=-1):
        self.calls.append(("sgd", grads_of))
        self.O.multi_tensor_sgd(chunk, flag, self._raw(self._with_grads(lists, grads_of)), *args)

    def multi_tensor_adam(self, chunk, flag, lists, *args, grads_of=-1):
        self.calls.append(("adam", grads_of))
        self.O.multi_tensor_adam(chunk, flag, self._raw(self._with_grads(lists, grads_of)), *args)

    def multi_tensor_lamb(self, chunk, flag, lists, *args, grads_of=-1):
        self.calls.append(("lamb", grads_of))
        *head, gnorm, max_norm, nvlamb = args
        self.O.multi_tensor_lamb(chunk, flag, self._raw(self._with_grads(lists, grads_of)), *head, gnorm, max_norm, nvlamb)

    def multi_tensor_l2norm(self, chunk, flag, lists, per_tensor=False, grads_of=-1):
        self.calls.append(("l2norm", grads_of))
        return self.O.multi_tensor_l2norm(chunk, flag, self._raw(self._with_grads(lists, grads_of, only=grads_of >= 0)), per_tensor)


SHAPES = [(33, 17), (1000,), (257,), (4, 3, 3, 3), (1,)]


def _make(which, ops, dtypes, fast):
    from beforeholiday_b200.optimizers import FusedAdam, FusedLAMB, FusedSGD

    g = torch.Generator().manual_seed(3)
    params = [torch.nn.Parameter((torch.randn(s, generator=g) * 0.1).to(dt)) for s, dt in zip(SHAPES, dtypes)]
    if which == "sgd":
        opt = FusedSGD(params, lr=0.1, momentum=0.9, dampening=0.1, weight_decay=1e-4)
        opt.multi_tensor_sgd = ops.multi_tensor_sgd
    elif which == "adam":
        # like the reference (fused_adam.py:77) the constructor puts its flag buffer on CUDA whatever the parameters'
        # device: redirect that one allocation for the CPU run
        orig = torch.zeros
        torch.zeros = lambda *a, **k: orig(*a, **{**k, "device": "cpu"})
        try:
            opt = FusedAdam(params, lr=1e-3, weight_decay=0.01)
        finally:
            torch.zeros = orig
        opt.multi_tensor_adam = ops.multi_tensor_adam
    else:
        opt = FusedLAMB(params, lr=1e-3, weight_decay=0.01, max_grad_norm=0.05)
        opt.multi_tensor_lamb, opt.multi_tensor_l2norm = ops.multi_tensor_lamb, ops.multi_tensor_l2norm
    opt._fast.enabled = fast
    return params, opt


def _grads(params, it, skip=()):
    for i, p in enumerate(params):
        if i in skip:
            p.grad = None
            continue
        g = torch.Generator().manual_seed(1000 * it + i)
        p.grad = (torch.randn(p.shape, generator=g) * 1e-2).to(p.dtype)


@pytest.mark.parametrize("which", ["sgd", "adam", "lamb"])
def test_steady_state_route_equals_generic_route(oracle, which):
    dts = [torch.float32, torch.float16, torch.float32, torch.float32, torch.float16]
    fops, sops = OracleOps(oracle), OracleOps(oracle)
    pa, fast = _make(which, fops, dts, True)
    pb, slow = _make(which, sops, dts, False)
    for it in range(6):
        skip = (2,) if it == 3 else ()            # step 3: one gradient missing -> that step must use the generic route
        _grads(pa, it, skip)
        _grads(pb, it, skip)
        n0 = len(fops.calls)
        fast.step()
        slow.step()
        used_fast = any(k >= 0 for _, k in fops.calls[n0:])
        if which == "lamb":
            # the missing-gradient step: the norm sweep raises, nothing else has run, the generic route takes over
            assert used_fast == (it != 3) or (it == 3 and all(name == "l2norm" for name, k in fops.calls[n0:] if k >= 0))
        else:
            assert used_fast, "the steady-state route must at least be attempted"
        for a, b in zip(pa, pb):
            assert bits_equal(a.detach(), b.detach()), f"{which}: parameters differ after step {it}"
    assert all(k < 0 for _, k in sops.calls)                    # the generic route never uses grads_of
    for a, b in zip(pa, pb):
        for key, va in fast.state[a].items():
            if torch.is_tensor(va):
                assert bits_equal(va, slow.state[b][key]), key
    if which != "sgd":
        assert fast.param_groups[0]['step'] == slow.param_groups[0]['step'] == 6


@pytest.mark.parametrize("which", ["sgd", "adam", "lamb"])
def test_cached_lists_follow_state_dict_and_new_groups(oracle, which):
    dts = [torch.float32] * 5
    ops = OracleOps(oracle)
    ps, opt = _make(which, ops, dts, True)
    _grads(ps, 0)
    opt.step()
    assert opt._fast._cache
    sd = opt.state_dict()
    opt.load_state_dict(sd)
    assert not opt._fast._cache                                 # state tensors were replaced
    _grads(ps, 1)
    opt.step()
    key = 'momentum_buffer' if which == "sgd" else 'exp_avg'
    cached = opt._fast._cache[0]['sets'][0]['moms' if which == "sgd" else 'm']
    assert all(c is opt.state[p][key] for c, p in zip(cached, ps))   # the route writes the LIVE state tensors
    extra = torch.nn.Parameter(torch.zeros(7))
    opt.add_param_group({'params': [extra]})
    assert not opt._fast._cache
    _grads(ps + [extra], 2)
    before = extra.detach().clone()
    opt.step()
    assert not bits_equal(extra.detach(), before)               # the new group is updated too


def test_sgd_first_run_semantics_survive_the_route(oracle):
    """momentum buffers start as the first gradient (first_run), also when the first step already uses the cached route,
    and a parameter that joins later gets ITS first-run treatment through the generic route."""
    ops = OracleOps(oracle)
    ps, opt = _make("sgd", ops, [torch.float32] * 5, True)
    ref_ops = OracleOps(oracle)
    pr, ref = _make("sgd", ref_ops, [torch.float32] * 5, False)
    for it in range(3):
        skip = (4,) if it == 0 else ()             # parameter 4 has no gradient on the very first step
        _grads(ps, it, skip)
        _grads(pr, it, skip)
        opt.step()
        ref.step()
        for a, b in zip(ps, pr):
            assert bits_equal(a.detach(), b.detach())
        for a, b in zip(ps, pr):
            if 'momentum_buffer' in ref.state[b]:
                assert bits_equal(opt.state[a]['momentum_buffer'], ref.state[b]['momentum_buffer'])
