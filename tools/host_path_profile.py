#!/usr/bin/env python
"""tools/host_path_profile.py -- where the host time of one amp_C call goes at 50 000 tensors (config C3):
list walk + validation (`_flatten_only`) vs the whole call (walk + table lookup + launch)."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

import beforeholiday_b200  # noqa: E402
import workloads  # noqa: E402

amp_C = beforeholiday_b200.load_amp_C()
numels = workloads.numels(workloads.c3_many_small())
P = sum(numels)
flat = [torch.zeros(P, device="cuda") for _ in range(3)]
lists = []
for f in flat:
    o, l = 0, []
    for n in numels:
        l.append(f[o:o + n])
        o += n
    lists.append(l)
flag = torch.zeros(1, dtype=torch.int32, device="cuda")


def host_ms(fn, n=30):
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        fn()
        best = min(best, (time.perf_counter() - t0) * 1e3)
    torch.cuda.synchronize()
    return best


for depth in (2, 3):
    ls = lists[:depth]
    walk = host_ms(lambda: amp_C._flatten_only(ls))
    if depth == 2:
        call = host_ms(lambda: amp_C.multi_tensor_scale(8192, flag, ls, 1.0))
    else:
        call = host_ms(lambda: amp_C.multi_tensor_sgd(8192, flag, ls, 1e-4, 0.9, 0.0, 0.25, True, False, False, 1.0))
    print(f"depth {depth}: {len(numels)} tensors  walk {walk:.3f} ms  whole call {call:.3f} ms  (table+launch {call - walk:.3f} ms)"
          f"  {call * 1e6 / len(numels) / depth:.1f} ns per tensor entry", flush=True)
