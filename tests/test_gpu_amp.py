"""amp's dynamic-loss-scaling flow around the fused optimizers, modelled on the reference's
tests/L0/run_amp/test_fused_sgd.py:55-185: toy model, O2 (fp16 params + fp32 masters), an inf injected into a gradient
at a chosen iteration; the skipped step must leave parameters untouched and halve the scale, and the trajectory must
follow torch.optim on fp32 masters (test_o2_adam_matches_manual_master_weight_loop)."""
import pytest
import torch

from util import bits_equal

pytestmark = pytest.mark.gpu


class Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.w0 = torch.nn.Parameter(torch.arange(2 * 33, dtype=torch.float32).reshape(2, 33) / 64 - 0.3)
        self.w1 = torch.nn.Parameter(torch.arange(33 * 4099, dtype=torch.float32).reshape(33, 4099) % 17 / 32 - 0.2)

    def forward(self, x):                      # x: [B, 2]
        return ((x @ self.w0) @ self.w1).float().pow(2).mean()


def _run(opt_name, inject_at, iters=5):
    from beforeholiday_b200 import amp
    from beforeholiday_b200.optimizers import FusedAdam, FusedSGD

    torch.manual_seed(0)
    model = Toy().cuda()
    if opt_name == "sgd":
        opt = FusedSGD(model.parameters(), lr=0.05, momentum=0.125, materialize_master_grads=True)
    elif opt_name == "sgd_inkernel_unscale":
        opt = FusedSGD(model.parameters(), lr=0.05, momentum=0.125, materialize_master_grads=False)
    else:
        opt = FusedAdam(model.parameters(), lr=1e-2, adam_w_mode=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", loss_scale="dynamic", verbosity=0)
    assert model.w0.dtype == torch.float16
    scales, skipped = [], []
    for it in range(iters):
        x = torch.randn(8, 2, device="cuda")
        loss = model(x.half())
        with amp.scale_loss(loss, opt) as scaled_loss:
            scaled_loss.backward()
            if it == inject_at:
                model.w1.grad[3, 7] = float("inf")
        before = [p.detach().clone() for p in model.parameters()]
        patched = opt._amp_stash.already_patched
        opt.step()
        opt.zero_grad()
        skipped.append(patched)
        scales.append(amp.state_dict()["loss_scaler0"]["loss_scale"])
        if patched:   # a skipped step must leave the parameters untouched (test_fused_sgd.py:165-182)
            assert all(bits_equal(a, b) for a, b in zip(before, model.parameters()))
        else:
            assert not all(bits_equal(a, b) for a, b in zip(before, model.parameters()))
    return model, opt, scales, skipped


@pytest.mark.parametrize("opt_name", ["sgd", "sgd_inkernel_unscale", "adam"])
@pytest.mark.parametrize("inject_at", [None, 0, 2])
def test_scale_loss_skip_and_scale_policy(opt_name, inject_at):
    model, opt, scales, skipped = _run(opt_name, inject_at)
    S = 2.0 ** 16
    expect_skip = [i == inject_at for i in range(5)]
    assert skipped == expect_skip
    cur, expect_scales = S, []
    for s in expect_skip:
        cur = cur / 2 if s else cur
        expect_scales.append(cur)
    assert scales == expect_scales
    # model params are the fp16 rounding of the fp32 masters the optimizer updates
    from beforeholiday_b200 import amp

    for low, master in zip(opt._amp_stash.all_fp16_params, opt._amp_stash.all_fp32_from_fp16_params):
        assert bits_equal(low.detach(), master.detach().half())


def test_o2_adam_matches_manual_master_weight_loop():
    """amp O2 + FusedAdam == manual loop: unscale fp16 grads into fp32, torch.optim.Adam on fp32 masters, cast back."""
    from beforeholiday_b200 import amp
    from beforeholiday_b200.optimizers import FusedAdam

    torch.manual_seed(1)
    model = Toy().cuda()
    opt = FusedAdam(model.parameters(), lr=1e-2, adam_w_mode=False)
    model, opt = amp.initialize(model, opt, opt_level="O2", loss_scale=128.0, verbosity=0)
    masters = [p.detach().float().clone().requires_grad_(True) for p in model.parameters()]
    ref = torch.optim.Adam(masters, lr=1e-2)
    for it in range(4):
        x = torch.randn(8, 2, device="cuda").half()
        loss = model(x)
        with amp.scale_loss(loss, opt) as scaled_loss:
            scaled_loss.backward()
        for m, p in zip(masters, model.parameters()):
            m.grad = p.grad.float() * (1.0 / 128.0)
        ref.step()
        opt.step()
        opt.zero_grad()
        for m, p, fm in zip(masters, model.parameters(), opt._amp_stash.all_fp32_from_fp16_params):
            assert torch.allclose(fm, m, rtol=2e-6, atol=1e-7)
            assert bits_equal(p.detach(), fm.detach().half())
    sd = amp.state_dict()
    assert sd == {"loss_scaler0": {"loss_scale": 128.0, "unskipped": 4}}
    amp.load_state_dict({"loss_scaler0": {"loss_scale": 64.0, "unskipped": 1}})
    assert amp.state_dict()["loss_scaler0"] == {"loss_scale": 64.0, "unskipped": 1}
