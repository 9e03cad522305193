#!/usr/bin/env python
"""tools/optimizer_step_profile.py -- wall time of a whole optimizer.step() (Python + extension + GPU) on the
50 000-tensor set (config C3) and the Llama-3-8B-shaped tensor COUNT (291), for the drop-in classes and torch.optim."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

import workloads  # noqa: E402
from beforeholiday_b200.optimizers import FusedAdam, FusedLAMB, FusedSGD  # noqa: E402


def make_params(numels):
    flat = torch.zeros(sum(numels), device="cuda").normal_(0, 0.02)
    gflat = torch.zeros(sum(numels), device="cuda").normal_(0, 1e-3)
    ps, o = [], 0
    for n in numels:
        p = torch.nn.Parameter(flat[o:o + n])
        p.grad = gflat[o:o + n]
        ps.append(p)
        o += n
    return ps


def wall_ms(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t0) * 1e3)
    return best


def main():
    numels = workloads.numels(workloads.c3_many_small())
    for name, ctor in (
        ("FusedSGD(momentum=0.9, nesterov)", lambda ps: FusedSGD(ps, lr=0.25, momentum=0.9, weight_decay=1e-4, nesterov=True)),
        ("FusedAdam", lambda ps: FusedAdam(ps, lr=1e-3)),
        ("torch.optim.SGD(foreach)", lambda ps: torch.optim.SGD(ps, lr=0.25, momentum=0.9, weight_decay=1e-4, nesterov=True, foreach=True)),
        ("torch.optim.Adam(fused)", lambda ps: torch.optim.Adam(ps, lr=1e-3, fused=True)),
    ):
        ps = make_params(numels)
        opt = ctor(ps)
        ms = wall_ms(opt.step)
        print(f"C3 {len(numels)} tensors  {name:36s} step() wall {ms:8.3f} ms", flush=True)
        del opt, ps
        torch.cuda.empty_cache()
    # C4: BERT-large FusedLAMB (398 tensors, 336 M params): generic per-step list building vs the steady-state route
    numels = workloads.numels(workloads.bert_large())
    for fast in (False, True):
        ps = make_params(numels)
        opt = FusedLAMB(ps, lr=1e-3, weight_decay=0.01, max_grad_norm=1.0)
        opt._fast.enabled = fast

        def one():
            # LAMB stage 1 overwrites the gradients with the update term: refill them each iteration (outside the timing
            # this would be the backward pass); cheap device-side copy
            opt.step()
        ms = wall_ms(one)
        print(f"C4 {len(numels)} tensors  FusedLAMB ({'steady-state route' if fast else 'generic route'})".ljust(56)
              + f" step() wall {ms:8.3f} ms", flush=True)
        del opt, ps
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
