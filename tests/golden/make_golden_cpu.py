#!/usr/bin/env python
"""Generate golden vectors from the REFERENCE's own CPU-capable code, in the build container.

Run here (never on the GPU box: /root/reference does not exist there):

    PYTHONPATH=/root/reference python tests/golden/make_golden_cpu.py

Writes tests/golden/ref_cpu.pt with:
  * scale / axpby + overflow semantics from the reference's Python fallback
    (apex/amp/scaler.py:6-40: scale_check_overflow_python, axpby_check_overflow_python);
  * a 7-iteration RefLAMB trajectory -- the pure-torch LAMB the reference's own test compares FusedLAMB against
    (tests/L0/run_optimizers/test_lamb.py:10-145), imported from that file with a stand-in `amp_C` whose only
    member is a torch.norm-based multi_tensor_l2norm (RefLAMB uses amp_C for nothing else);
  * 7-iteration torch.optim.Adam / AdamW / SGD trajectories with the reference tests' seeds and hyper-parameters
    (tests/L0/run_optimizers/test_fused_optimizer.py:13-17,64-88,278-289) -- the reference's own oracle.

The inputs are regenerated from the recorded seeds by the tests; only outputs are stored.
"""
import importlib.util
import sys
import types
from pathlib import Path

import torch

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent / "ref_cpu.pt"
N = 4099  # odd on purpose: exercises the unaligned tail, like the reference's nelem=278011
ITERS = 7


def gen_inputs(seed, n, dtype=torch.float32, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n, generator=g) * scale).to(dtype)


def main():
    sys.path.insert(0, str(REF))
    # stand-in amp_C: only what RefLAMB touches
    stub = types.ModuleType("amp_C")

    def multi_tensor_l2norm(chunk_size, noop_flag, tensor_lists, per_tensor):
        flat = torch.cat([t.float().reshape(-1) for t in tensor_lists[0]])
        return flat.norm().reshape(1), torch.empty(0)

    stub.multi_tensor_l2norm = multi_tensor_l2norm
    stub.multi_tensor_lamb = None
    for name in ("multi_tensor_scale", "multi_tensor_axpby", "multi_tensor_adam", "multi_tensor_sgd"):
        setattr(stub, name, None)
    sys.modules["amp_C"] = stub

    from apex.amp import scaler as ref_scaler  # noqa: E402  (reference python)

    out = {"N": N, "ITERS": ITERS}

    # ---- scale / axpby python fallbacks ----------------------------------------------------------
    cases = []
    for in_dt in (torch.float16, torch.float32):
        for inject in (None, "inf", "nan"):
            model = gen_inputs(11, N, in_dt, 3.0)
            if inject == "inf":
                model[N // 2] = float("inf")
            if inject == "nan":
                model[-1] = float("nan")
            master = torch.zeros(N, dtype=torch.float32)
            ovf = ref_scaler.scale_check_overflow_python(model, master, 1.0 / 65536.0, check_overflow=True)
            cases.append({"in_dtype": in_dt, "inject": inject, "overflow": bool(ovf), "master": master.clone()})
    out["scale_python"] = cases

    cases = []
    for inject in (None, "inf"):
        model = gen_inputs(12, N, torch.float16, 2.0)
        if inject == "inf":
            model[7] = float("inf")
        stashed = gen_inputs(13, N, torch.float32)
        master = torch.zeros(N, dtype=torch.float32)
        holder = types.SimpleNamespace(data=master, dtype=master.dtype)
        ovf = ref_scaler.axpby_check_overflow_python(model, stashed, holder, 1.0 / 128.0, 1.0, check_overflow=True)
        cases.append({"inject": inject, "overflow": bool(ovf), "master": None if ovf else holder.data.clone()})
    out["axpby_python"] = cases

    # ---- RefLAMB trajectory ------------------------------------------------------------------------
    spec = importlib.util.spec_from_file_location("ref_test_lamb", REF / "tests/L0/run_optimizers/test_lamb.py")
    mod = importlib.util.module_from_spec(spec)
    from apex.multi_tensor_apply import multi_tensor_applier as ref_applier

    spec.loader.exec_module(mod)
    assert ref_applier.available
    lamb = {}
    for wd in (0.0, 0.01):
        sizes = [N, 257, 1]
        params = [torch.nn.Parameter(gen_inputs(100 + i, n).abs() * 0.5 + 0.01) for i, n in enumerate(sizes)]
        opt = mod.RefLAMB(params, lr=5e-4, betas=(0.9, 0.999), eps=1e-6, weight_decay=wd)
        for it in range(ITERS):
            for i, p in enumerate(params):
                p.grad = gen_inputs(1000 + 10 * it + i, p.numel()).reshape(p.shape) * 0.1
            opt.step()
        lamb[wd] = {"sizes": sizes, "params": [p.detach().clone() for p in params],
                    "m": [opt.state[p]["m"].clone() for p in params], "v": [opt.state[p]["v"].clone() for p in params]}
    out["ref_lamb"] = lamb

    # ---- torch.optim trajectories (the reference tests' own oracle) ------------------------------
    traj = {}
    for name, ctor in (
        ("adam_l2", lambda ps: torch.optim.Adam(ps, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False)),
        ("adam_l2_wd", lambda ps: torch.optim.Adam(ps, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.1)),
        ("adamw", lambda ps: torch.optim.AdamW(ps, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.01)),
        ("sgd_mom", lambda ps: torch.optim.SGD(ps, lr=0.25, momentum=0.125)),
        ("sgd_nesterov_wd", lambda ps: torch.optim.SGD(ps, lr=0.1, momentum=0.9, weight_decay=1e-4, nesterov=True)),
    ):
        torch.manual_seed(9876)
        p = torch.nn.Parameter(torch.rand(N))
        opt = ctor([p])
        for it in range(ITERS):
            p.grad = gen_inputs(2000 + it, N).abs() % 1.0  # in [0,1) like torch.rand_like
            opt.step()
        traj[name] = p.detach().clone()
    out["torch_optim"] = traj

    torch.save(out, OUT)
    print("wrote", OUT, OUT.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
