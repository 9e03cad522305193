"""Dynamic loss-scale POLICY of amp's LossScaler (apex/amp/scaler.py:42-226) on CPU: start value, halving and skipping on
overflow, floor, doubling after `scale_window` clean steps, cap, the single flag read per iteration, state_dict.  The two
device ops are replaced by the oracle's restatements (test infrastructure); the launches themselves are -m gpu tests."""
import pytest
import torch


@pytest.fixture
def Scaler(oracle, monkeypatch):
    from beforeholiday_b200.amp import scaler as S

    orig = torch.zeros
    monkeypatch.setattr(torch, "zeros", lambda *a, **k: orig(*a, **{**k, "device": "cpu"}))

    def make(*a, **k):
        sc = S.LossScaler(*a, **k)
        S.LossScaler.multi_tensor_scale_cuda = staticmethod(oracle.multi_tensor_scale)
        S.LossScaler.multi_tensor_axpby_cuda = staticmethod(oracle.multi_tensor_axpby)
        return sc

    yield make
    import beforeholiday_b200

    amp_C = beforeholiday_b200.load_amp_C()
    S.LossScaler.multi_tensor_scale_cuda = amp_C.multi_tensor_scale
    S.LossScaler.multi_tensor_axpby_cuda = amp_C.multi_tensor_axpby


def _step(sc, grads16, overflow=False):
    """One iteration as amp drives it: clear, unscale into fp32 masters, update."""
    sc.clear_overflow_state()
    g = [t.clone() for t in grads16]
    if overflow:
        g[1][3] = float("inf")
    masters = [torch.zeros(t.shape) for t in g]
    scale = sc.loss_scale()
    sc.unscale(g, masters, scale)
    skipped = sc.update_scale()
    return skipped, masters


def test_policy(Scaler):
    sc = Scaler("dynamic", scale_window=4, min_loss_scale=2.0 ** 13)
    assert sc.loss_scale() == 2.0 ** 16 and sc.dynamic
    grads = [torch.full((5,), 3.0, dtype=torch.float16), torch.full((9,), -1.5, dtype=torch.float16)]
    skipped, masters = _step(sc, grads)
    assert not skipped and torch.equal(masters[0], torch.full((5,), 3.0 / 65536)) and torch.equal(masters[1], torch.full((9,), -1.5 / 65536))
    # overflow: skip, halve, window restarts
    skipped, _ = _step(sc, grads, overflow=True)
    assert skipped and sc.loss_scale() == 2.0 ** 15 and sc._unskipped == 0
    # floor
    for _ in range(5):
        _step(sc, grads, overflow=True)
    assert sc.loss_scale() == 2.0 ** 13
    # four clean steps double it once; the window restarts
    for i in range(4):
        skipped, _ = _step(sc, grads)
        assert not skipped
    assert sc.loss_scale() == 2.0 ** 14 and sc._unskipped == 0
    # cap
    sc2 = Scaler("dynamic", init_scale=2.0 ** 30, scale_window=1, max_loss_scale=2.0 ** 24)
    assert sc2.loss_scale() == 2.0 ** 24
    _step(sc2, grads)
    assert sc2.loss_scale() == 2.0 ** 24
    # state round trip
    sd = sc.state_dict()
    sc3 = Scaler("dynamic")
    sc3.load_state_dict(sd)
    assert sc3.loss_scale() == sc.loss_scale() and sc3._unskipped == sc._unskipped


def test_static_scale_never_skips_and_unscale_is_bypassed_at_one(Scaler):
    sc = Scaler(128.0)
    grads = [torch.full((4,), 256.0, dtype=torch.float16)]
    skipped, masters = _step(sc, grads, overflow=False)
    assert not skipped and torch.equal(masters[0], torch.full((4,), 2.0)) and sc.loss_scale() == 128.0
    sc1 = Scaler(1.0)
    m = [torch.zeros(4)]
    sc1.unscale([torch.ones(4)], m, 1.0, models_are_masters=True)      # scale 1, same tensors: nothing to do
    assert float(m[0].sum()) == 0.0


def test_once_overflowed_later_unscales_of_the_iteration_are_skipped(Scaler):
    sc = Scaler("dynamic")
    sc.clear_overflow_state()
    bad = [torch.tensor([float("nan"), 1.0], dtype=torch.float16)]
    m = [torch.zeros(2)]
    sc.unscale(bad, m, sc.loss_scale())
    assert sc.update_scale() is True
    m2 = [torch.zeros(2)]
    sc.unscale([torch.ones(2, dtype=torch.float16)], m2, sc.loss_scale())   # _has_overflow latched: no launch
    assert float(m2[0].sum()) == 0.0
    stashed, out = [torch.ones(2)], [torch.zeros(2)]
    sc.unscale_with_stashed([torch.ones(2, dtype=torch.float16)], stashed, out)
    assert float(out[0].sum()) == 0.0
    sc.clear_overflow_state()
    sc.unscale_with_stashed([torch.full((2,), 4.0, dtype=torch.float16)], stashed, out)
    assert torch.equal(out[0], torch.full((2,), 4.0 / sc.loss_scale() + 1.0))
