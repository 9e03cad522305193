#!/usr/bin/env python
"""tools/bench_dist_lamb.py -- sharded LAMB (DistributedFusedLAMB) step on N B200s of one node, timed on the device,
max over ranks, with a breakdown.  Launch like bench.py's N > 1 leg:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/bench_dist_lamb.py [--workload bert_large|llama3_8b] [--steps 10] [--warmup 3]

Rank 0 prints one JSON line.  (N = 1 works too: no process group needed, plain `python tools/bench_dist_lamb.py`.)
"""
from __future__ import annotations

import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="bert_large", choices=sorted(workloads.WORKLOADS))
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp16"])
    ap.add_argument("--e5m2-allgather", action="store_true", help="parameters travel as one e5m2 byte each (reference option)")
    ap.add_argument("--clip-before-ar", action="store_true", help="clip_after_ar=False (reference option)")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from beforeholiday_b200.contrib.optimizers import DistributedFusedLAMB

    dt = torch.bfloat16 if args.dtype == "bf16" else torch.float16
    shapes = workloads.WORKLOADS[args.workload]()
    numels = workloads.numels(shapes)
    P = sum(numels)
    params = []
    for n in numels:
        t = torch.empty(n, dtype=dt, device=dev)
        t.normal_(0, 0.02)
        params.append(torch.nn.Parameter(t))
    opt = DistributedFusedLAMB(params, lr=1e-4, betas=(0.9, 0.999), eps=1e-6, weight_decay=0.01, max_grad_norm=1.0,
                               dwu_num_blocks=4, dwu_num_chunks=4, e5m2_allgather=args.e5m2_allgather,
                               clip_after_ar=not args.clip_before_ar)
    del params
    opt.set_global_scale(torch.tensor([1024.0], device=dev))
    L = opt.layout
    amp_C = opt._amp_C()

    def fill():
        opt.grads_flat.normal_(0, 1.0)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, steps, prep=None):
        tot = 0.0
        for _ in range(steps):
            if prep:
                prep()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize(dev)
            tot += e0.elapsed_time(e1)
        ms = torch.tensor([tot / steps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    def full():
        opt.complete_reductions()
        opt.step()

    for _ in range(args.warmup):
        fill()
        full()
    assert int(opt._overflow_buf) == 0, "benchmark gradients overflowed"
    l0 = amp_C.launch_count()
    step_ms = timed(full, args.steps, prep=fill)
    launches = (amp_C.launch_count() - l0) / args.steps

    gs, gn = opt._device_scalar(opt.global_scale, "_gs_cache"), opt.L2_grad_norm

    def rs_all():
        for b in range(L.nbuckets):
            opt._reduce_scatter(b)

    def ag_all():
        for b in range(L.nbuckets):
            opt._all_gather(b)

    def stage2_all():
        for b in range(L.nbuckets):
            opt._stage2(b, opt._param_norms, opt._update_norms, gn)

    bd = {
        "reduce_scatter_only": timed(rs_all, 3),
        "grad_norm_only": timed(lambda: opt._shard_sq_norm(), 3),
        "stage1_only": timed(lambda: opt._stage1(gs, gn, 1.0), 3),
        "stage2_only": timed(stage2_all, 3),
        "all_gather_only": timed(ag_all, 3),
    }
    shard = L.state_elems
    if rank == 0:
        print(json.dumps({
            "metric": "sharded_lamb_step_params_per_sec", "value": P / (step_ms * 1e-3), "unit": "params/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_ms, "scaling": "strong", "data": "synthetic",
            "config": {"workload": f"{args.workload}: DistributedFusedLAMB, {args.dtype} grads/params + fp32 p/m/v/u shards "
                                   f"({len(numels)} tensors, {P} params)", "nbuckets": L.nbuckets, "shard_elems": shard,
                       "fragments_rank0": len(opt.frags), "e5m2_allgather": args.e5m2_allgather,
                       "clip_after_ar": not args.clip_before_ar},
            "breakdown_ms": bd, "launches_per_step": launches,
            "kernel_hbm_gbs": {"stage1": 26 * shard / bd["stage1_only"] / 1e6, "stage2": 14 * shard / bd["stage2_only"] / 1e6,
                               "grad_norm": 2 * shard / bd["grad_norm_only"] / 1e6},
        }), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
