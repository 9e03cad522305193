#!/usr/bin/env python
"""Golden NovoGrad trajectories from the REFERENCE's own pure-torch optimizer, in the build container.

    python tests/golden/make_golden_novograd.py

The reference's test for FusedNovoGrad carries a plain-PyTorch `Novograd` optimizer that it compares the fused kernel
against (/root/reference/tests/L0/run_optimizers/test_fused_novograd.py:10-114).  That class is executed here from
where it lies (its source text is exec'd, nothing is copied into this repository; the rest of the module needs a CUDA
build of apex and is not imported) on seeded CPU inputs, and the resulting parameters / moments are stored in
tests/golden/ref_novograd.pt.  tests/test_oracle_pins.py regenerates the inputs from the same seeds and pins the
oracle's multi_tensor_novograd restatement against them.

Cases: the reference test's own options (lr 1e-3, betas (0.95, 0), eps 1e-8, no decay, no grad averaging,
test_fused_novograd.py:125-131) and a second one that exercises the norm blend, L2 regularisation inside the moment and
gradient averaging (betas (0.95, 0.98), weight_decay 0.01, grad_averaging True).
"""
import math
from pathlib import Path

import torch
from torch.optim import Optimizer

REF = Path("/root/reference/tests/L0/run_optimizers/test_fused_novograd.py")
OUT = Path(__file__).resolve().parent / "ref_novograd.pt"
SIZES = [4099, 33, 9001]
ITERS = 7
CASES = {
    "reference_test_options": dict(lr=1e-3, betas=(0.95, 0), eps=1e-8, weight_decay=0, grad_averaging=False),
    "blend_decay_averaging": dict(lr=1e-3, betas=(0.95, 0.98), eps=1e-8, weight_decay=0.01, grad_averaging=True),
}


def gen(seed, n, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, generator=g) * scale


def load_reference_class():
    src = REF.read_text()
    a = src.index("class Novograd(Optimizer):")
    b = src.index("class TestFusedNovoGrad")
    ns = {"torch": torch, "Optimizer": Optimizer, "math": math}
    exec(compile(src[a:b], str(REF), "exec"), ns)
    return ns["Novograd"]


def main():
    Novograd = load_reference_class()
    out = {"SIZES": SIZES, "ITERS": ITERS, "cases": {}}
    for name, opts in CASES.items():
        params = [torch.nn.Parameter(gen(300 + i, n, 0.1)) for i, n in enumerate(SIZES)]
        opt = Novograd(params, amsgrad=False, **opts)
        for it in range(ITERS):
            for i, p in enumerate(params):
                p.grad = gen(4000 + 10 * it + i, p.numel(), 0.05)
            opt.step()
        out["cases"][name] = {"opts": opts, "params": [p.detach().clone() for p in params],
                              "exp_avg": [opt.state[p]["exp_avg"].clone() for p in params],
                              "exp_avg_sq": [opt.state[p]["exp_avg_sq"].clone() for p in params]}
    torch.save(out, OUT)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
