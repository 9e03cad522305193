"""Host logic of the optimizers' steady-state route (beforeholiday_b200/optimizers/_fast_path.py): when the cached lists
are built, reused and dropped.  Pure Python: runs without a GPU (the launches themselves are covered by -m gpu tests)."""
import torch

from beforeholiday_b200.optimizers._fast_path import FastGroups, is_fallback


def _params(n=4, with_grads=True):
    ps = [torch.nn.Parameter(torch.zeros(3, 2)) for _ in range(n)]
    if with_grads:
        for p in ps:
            p.grad = torch.ones(3, 2)
    return ps


def test_cache_is_built_once_and_reused():
    fg = FastGroups()
    group = {'params': _params()}
    calls = []

    def build(params):
        calls.append(len(params))
        return [{'params': list(params)}]

    a = fg.get(0, group, build)
    b = fg.get(0, group, build)
    assert a is b and calls == [4] and a['n'] == 4 and a['sets'][0]['params'][0] is group['params'][0]


def test_not_built_unless_every_gradient_is_dense_and_laid_out_like_its_parameter():
    fg = FastGroups()
    built = []
    build = lambda params: built.append(1) or [{}]
    ps = _params()
    ps[2].grad = None
    assert fg.get(0, {'params': ps}, build) is None and not built          # missing gradient
    ps = _params()
    ps[1].grad = torch.ones(2, 3).t()                                        # same shape, different strides
    assert fg.get(0, {'params': ps}, build) is None and not built
    ps = _params()
    ps[0].grad = torch.ones(3, 2).to_sparse()
    assert fg.get(0, {'params': ps}, build) is None and not built
    assert fg.get(0, {'params': _params()}, lambda params: None) is None     # the optimizer's own veto (e.g. dtypes)
    fg.enabled = False
    assert fg.get(0, {'params': _params()}, build) is None and not built


def test_cache_is_dropped_when_the_group_changes_or_on_clear():
    fg = FastGroups()
    n = []
    build = lambda params: n.append(1) or [{'params': list(params)}]
    group = {'params': _params()}
    c1 = fg.get(0, group, build)
    group['params'] = group['params'] + _params(1)                           # add a parameter: rebuilt
    c2 = fg.get(0, group, build)
    assert c2 is not c1 and c2['n'] == 5 and len(n) == 2
    fg.clear()
    assert fg.get(0, group, build) is not c2 and len(n) == 3
    # a later step that lacks a gradient does NOT rebuild or drop anything here: the extension raises "grads_of: ..."
    group['params'][0].grad = None
    assert fg.get(0, group, build)['n'] == 5 and len(n) == 3


def test_fallback_marker():
    assert is_fallback(RuntimeError("grads_of: a parameter has no gradient"))
    assert not is_fallback(RuntimeError("A tensor was not contiguous."))
