#!/usr/bin/env python
"""tools/ncu_each_kernel.py -- launch every kernel family of the path a couple of times on GPT-2-small / BERT-large /
50k-tensor sized inputs, so that one `ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,
dram__throughput.avg.pct_of_peak_sustained_elapsed` pass yields per-kernel DRAM evidence (profiles/r2_ncu_per_kernel_*;
tools/ncu_table.py joins the CSV with the plan printed here).
Prints, per family, the algorithmic bytes of one launch so the CSV can be compared against them."""
import json
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

import beforeholiday_b200  # noqa: E402
import workloads  # noqa: E402

amp_C = beforeholiday_b200.load_amp_C()
F32, F16, BF16 = torch.float32, torch.float16, torch.bfloat16


def alloc(numels, dt, scale=0.02):
    flat = torch.empty(sum(numels), dtype=dt, device="cuda").normal_(0, scale)
    out, o = [], 0
    for n in numels:
        out.append(flat[o:o + n])
        o += n
    return out


flag = torch.zeros(1, dtype=torch.int32, device="cuda")
plan = []


def run(name, bytes_per_elem, P, fn, reps=2):
    l0 = amp_C.launch_count()
    for _ in range(reps):
        flag.zero_()
        fn()
    torch.cuda.synchronize()
    plan.append({"family": name, "algorithmic_bytes_per_launch": bytes_per_elem * P, "elements": P, "launches": reps,
                 "kernels_per_call": (amp_C.launch_count() - l0) // reps})


# ---- GPT-2-small sized (124 M elements, 148 tensors)
n2 = workloads.numels(workloads.gpt2_small())
P2 = sum(n2)
g16, g32, p32, m, v = alloc(n2, F16, 1.0), alloc(n2, F32), alloc(n2, F32), alloc(n2, F32), [t.abs() for t in alloc(n2, F32)]
p16 = alloc(n2, F16)
run("ScaleOp fp16->fp32", 6, P2, lambda: amp_C.multi_tensor_scale(8192, flag, [g16, g32], 1.0 / 65536))
run("ScaleOp fp32->fp16", 6, P2, lambda: amp_C.multi_tensor_scale(8192, flag, [p32, p16], 1.0))
run("AxpbyOp fp16,fp32->fp32", 10, P2, lambda: amp_C.multi_tensor_axpby(8192, flag, [g16, g32, g32], 1.0 / 65536, 1.0, 0))
run("L2NormOp fp16 (in-launch finalize)", 2, P2, lambda: amp_C.multi_tensor_l2norm(65536, flag, [g16], True))
run("L2NormOp fp32 (in-launch finalize)", 4, P2, lambda: amp_C.multi_tensor_l2norm(65536, flag, [g32], True))
run("L2NormScaleOp fp32->fp32 (in-launch finalize)", 8, P2, lambda: amp_C.multi_tensor_l2norm_scale(65536, flag, [g32, p32], 0.5, False))
run("AdamOp fp32", 28, P2, lambda: amp_C.multi_tensor_adam(8192, flag, [g32, p32, m, v], 5e-4, 0.9, 0.999, 1e-8, 3, 1, 1, 0.01))
nt = len(n2)
T = lambda val, dt=F32: torch.full((nt,), val, dtype=dt, device="cuda")
hp = (T(0.9), T(0.999), T(1, torch.int32), T(1e-8), T(0.01))
run("DistAdamOp fp16 g, fp32 p/m/v, fp16 copy", 28, P2,
    lambda: amp_C.multi_tensor_fused_adam(8192, flag, [p32, m, v, g16, p16], *hp, 5e-4, 65536.0, 3, 1, True))
run("SgdOp fp32 momentum", 20, P2, lambda: amp_C.multi_tensor_sgd(8192, flag, [g32, p32, m], 1e-4, 0.9, 0.0, 0.1, True, False, False, 1.0))
run("AdagradOp fp32", 20, P2, lambda: amp_C.multi_tensor_adagrad(8192, flag, [g32, p32, v], 1e-2, 1e-10, 0, 0.0))
del g16, g32, p32, m, v, p16

# ---- BERT-large sized (336 M elements, 398 tensors): LAMB = stage 1 (+ both per-tensor norms, finalized in-launch) -> stage 2
n4 = workloads.numels(workloads.bert_large())
P4 = sum(n4)
g, p, m, v = alloc(n4, F32, 1e-3), alloc(n4, F32), alloc(n4, F32, 1e-4), [t.abs() for t in alloc(n4, F32, 1e-4)]
gn = torch.tensor([2.0], device="cuda")
run("LambStage1Op (+ in-launch norms) + LambStage2Op fp32 (28 + 12 B)", 40, P4,
    lambda: amp_C.multi_tensor_lamb(8192, flag, [g, p, m, v], 5e-4, 0.9, 0.999, 1e-6, 3, 1, 0.01, 1, 1, gn, 1.0, False))
del g, p, m, v

# ---- 50 000 tensors
n3 = workloads.numels(workloads.c3_many_small())
P3 = sum(n3)
g, w, mom = alloc(n3, F32, 1e-2), alloc(n3, F32), alloc(n3, F32, 1e-3)
run("SgdOp fp32 momentum, 50k tensors", 20, P3, lambda: amp_C.multi_tensor_sgd(8192, flag, [g, w, mom], 1e-4, 0.9, 0.0, 0.25, True, False, False, 1.0))
print(json.dumps(plan), flush=True)
