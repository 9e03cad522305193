"""bench_dist.py -- the N > 1 leg of bench.py: ZeRO-2 sharded mixed-precision AdamW over N B200s of one node.

Step (timed, all of it): reduce-scatter of the bf16 gradients -> update of this rank's 1/N shard (fp32 p/m/v) ->
all-gather of the updated bf16 parameters.  Total work is fixed (strong scaling); `value` = params of the WHOLE set /
max-over-ranks step time.  The line also breaks the step down (update kernel only, reduce-scatter only, all-gather only)
because the end-to-end step is NVLink-bound: 2 B/param must leave and enter every GPU twice.
"""
from __future__ import annotations

import json
import sys
import time

import torch
import torch.distributed as dist

import workloads
from bench import BYTES_PER_PARAM, HP, METRIC, UNIT, ClockSampler, log, peaks


def _timed(fn, steps, device):
    dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(device)
    dist.barrier()
    t1 = time.time()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), t0, t1


PARITY_SIZES = [1000, 7, 4096, 33333, 1, 129, 300001]


def parity_block(amp_C, rank, world, device):
    """After the timed region: a small sharded run (several buckets, parameter fragments straddling shard boundaries,
    per-parameter weight decay, 3 steps) through every route, compared ON EVERY RANK bit for bit with the single-GPU
    kernel (amp_C.multi_tensor_fused_adam) applied to the rank-order sum of the same gradients.  Gradients are small
    multiples of 2^-10, so every partial sum is exact in bf16 and the order of the reduction cannot matter.  The
    single-GPU kernel is itself held bit-exact to the CPU oracle by tests/ (bench.py may not execute oracle/ outside
    its cpu_baseline leg).  Returns {"nccl": bool, "fused_p2p": bool|None, "fused_multicast": bool|None,
    "fused_hybrid": bool|None, "max_ulp": int}."""
    from beforeholiday_b200.contrib.optimizers import FlatShardedAdam

    sizes, k = PARITY_SIZES, len(PARITY_SIZES)
    wds = [0.1 if i % 2 else 0.0 for i in range(k)]
    hp = dict(lr=1e-2, betas=(0.9, 0.95), eps=1e-8)

    def grads_of(it, r):
        g = torch.Generator().manual_seed(100 * it + r)
        return [(torch.randint(-16, 16, (n,), generator=g).float() / 1024) for n in sizes]

    # single-GPU truth
    torch.manual_seed(0)
    init = [(torch.randn(n) * 0.02).bfloat16() for n in sizes]
    p32 = [x.float().to(device) for x in init]
    m = [torch.zeros(n, device=device) for n in sizes]
    v = [torch.zeros(n, device=device) for n in sizes]
    pc = [torch.zeros(n, dtype=torch.bfloat16, device=device) for n in sizes]
    T = lambda vals, dt: torch.tensor(vals, dtype=dt, device=device)
    hps = (T([hp["betas"][0]] * k, torch.float32), T([hp["betas"][1]] * k, torch.float32), T([1] * k, torch.int32),
           T([hp["eps"]] * k, torch.float32), T(wds, torch.float32))
    flag = torch.zeros(1, dtype=torch.int32, device=device)
    for it in range(3):
        tot = [torch.zeros(n) for n in sizes]
        for r in range(world):
            for t, g in zip(tot, grads_of(it, r)):
                t += g
        gs = [t.bfloat16().to(device) for t in tot]
        amp_C.multi_tensor_fused_adam(8192, flag, [p32, m, v, gs, pc], *hps, hp["lr"], float(world), it + 1, 1, True)
    out = {"max_ulp": 0}
    for route in ("nccl", "fused_p2p", "fused_multicast", "fused_hybrid"):
        try:
            opt = FlatShardedAdam(sizes, device=device, bucket_cap_mb=65536 * 2 / 2 ** 20, per_param_weight_decay=wds,
                                  fused_p2p=(route != "nccl"), **hp)
            if route != "nccl" and not opt.fused_p2p:
                out[route] = None
                continue
            if route in ("fused_multicast", "fused_hybrid") and not opt.has_multicast:
                out[route] = None
                continue
            L = opt.layout
            assert L.nbuckets > 1
            for x, o in zip(init, L.offsets):
                opt.params_flat[o:o + x.numel()].copy_(x)
            opt.init_master_from_params()
            for it in range(3):
                opt.zero_grad()
                for g, n, o in zip(grads_of(it, rank), sizes, L.offsets):
                    opt.grads_flat[o:o + n].copy_(g)
                if route == "nccl":
                    opt.step()
                elif route == "fused_hybrid":
                    opt.step_fused(multicast=True, ld_p2p_of16=5, st_p2p_of16=9)
                else:
                    opt.step_fused(multicast=(route == "fused_multicast"), ld_p2p_of16=0, st_p2p_of16=0)
            torch.cuda.synchronize(device)
            worst = 0
            for want, n, o in zip(pc, sizes, L.offsets):
                a = opt.params_flat[o:o + n].view(torch.int16).to(torch.int32)
                b = want.view(torch.int16).to(torch.int32)
                worst = max(worst, int((a - b).abs().max()))
            ok = torch.tensor([1 if (worst == 0 and int(opt.noop.item()) == 0) else 0], device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            w = torch.tensor([worst], device=device)
            dist.all_reduce(w, op=dist.ReduceOp.MAX)
            out[route] = bool(ok.item())
            out["max_ulp"] = max(out["max_ulp"], int(w.item()))
            del opt
        except Exception as e:  # noqa: BLE001
            out[route] = False
            out.setdefault("errors", {})[route] = repr(e)[:200]
    out["vs"] = ("single-GPU amp_C.multi_tensor_fused_adam on the rank-order sum of the gradients, every rank, all "
                 f"{sum(PARITY_SIZES)} gathered bf16 params, 3 steps, {k} tensors over several buckets")
    return out


def run_distributed(args, rank, world, local_rank, device):
    import beforeholiday_b200
    from beforeholiday_b200.contrib.optimizers import FlatShardedAdam

    amp_C = beforeholiday_b200.load_amp_C()
    shapes = workloads.WORKLOADS[args.workload]()
    numels = workloads.numels(shapes)
    P = sum(numels)
    kw = dict(lr=HP["lr"], betas=(HP["beta1"], HP["beta2"]), eps=HP["eps"], weight_decay=HP["wd"],
              adam_w_mode=True, device=device, bucket_cap_mb=1024.0, grad_dtype=torch.bfloat16, param_dtype=torch.bfloat16)
    fused_err = None
    try:  # symmetric-memory buffers: enables the single-kernel reduce-scatter + update + all-gather route
        opt = FlatShardedAdam(numels, fused_p2p=True, **kw)
    except Exception as e:  # noqa: BLE001
        fused_err = repr(e)[:300]
        opt = FlatShardedAdam(numels, fused_p2p=False, **kw)
    ok = torch.tensor([1 if opt.fused_p2p else 0], device=device)
    dist.all_reduce(ok, op=dist.ReduceOp.MIN)
    fused = bool(ok.item())
    L = opt.layout
    if rank == 0:
        log(f"[bench] {args.workload}: {len(numels)} tensors, {P} params; world={world}, {L.nbuckets} buckets of "
            f"{L.bucket_elems} elems, shard {L.state_elems} elems/rank, {len(opt.frags)} fragments on rank 0")
    CH = 1 << 28
    gen = torch.Generator(device=device).manual_seed(1234)          # same params on every rank
    for o in range(0, L.padded, CH):
        opt.params_flat[o:o + CH].normal_(0, 0.02, generator=gen)
    opt.init_master_from_params()
    gen = torch.Generator(device=device).manual_seed(100 + rank)     # each rank's own gradients
    for o in range(0, L.padded, CH):
        opt.grads_flat[o:o + CH].normal_(0, 1e-3, generator=gen)
    torch.cuda.synchronize(device)

    sampler = ClockSampler(local_rank) if rank == 0 else None
    # NCCL route (reduce_scatter -> update -> all_gather on three streams)
    for _ in range(args.warmup):
        opt.step()
    launches0 = amp_C.launch_count()
    ms_nccl, t0, t1 = _timed(opt.step, args.steps, device)
    launches = amp_C.launch_count() - launches0
    clocks = ClockSampler.summarize(sampler.window(t0, t1)) if sampler else None
    ms, route, ms_fused = ms_nccl, "nccl: reduce_scatter -> update -> all_gather, 3 streams", None
    ms_p2p = ms_mc = ms_hy = hybrid_best = hybrid_search = None
    fused_fn = None
    if fused:
        # fused route: one kernel reads the gradient shard from every peer and writes the parameters to every peer
        for _ in range(args.warmup):
            opt.step_fused()
        launches0 = amp_C.launch_count()
        ms_p2p, t0f, t1f = _timed(lambda: opt.step_fused(multicast=False), args.steps, device)
        n_l = amp_C.launch_count() - launches0
        ms_fused, fused_fn = ms_p2p, (lambda: opt.step_fused(multicast=False))
        fused_name = "fused_p2p: one kernel, ld/st on every peer's buffer over NVLink (+2 barriers)"
        if opt.has_multicast:
            for _ in range(args.warmup):
                opt.step_fused(multicast=True)
            ms_mc, t0m, t1m = _timed(lambda: opt.step_fused(multicast=True), args.steps, device)
            if ms_mc < ms_p2p:
                ms_fused, fused_fn, t0f, t1f = ms_mc, (lambda: opt.step_fused(multicast=True)), t0m, t1m
                fused_name = "fused_multicast: one kernel, multimem.ld_reduce / multimem.st through NVSwitch (+2 barriers)"
            # hybrid: a share of the 32 KB blocks goes peer-to-peer instead of through the switch.  A short search over
            # the split (2 warm-up + 3 steps each), then the best one is timed like the other routes.
            search = {}
            for ld16, st16 in ((2, 2), (0, 2), (0, 4), (0, 6)):
                fn_h = (lambda a=ld16, b=st16: opt.step_fused(multicast=True, ld_p2p_of16=a, st_p2p_of16=b))
                for _ in range(2):
                    fn_h()
                search[(ld16, st16)], _, _ = _timed(fn_h, 3, device)
            bl, bs = min(search, key=search.get)
            hybrid_search = {f"ld {a}/16, st {b}/16": v for (a, b), v in search.items()}
            fn_h = (lambda: opt.step_fused(multicast=True, ld_p2p_of16=bl, st_p2p_of16=bs))
            ms_hy, t0h, t1h = _timed(fn_h, args.steps, device)
            hybrid_best = f"ld {bl}/16, st {bs}/16"
            if ms_hy < ms_fused:
                ms_fused, fused_fn, t0f, t1f = ms_hy, fn_h, t0h, t1h
                fused_name = (f"fused_hybrid: one kernel, {bl}/16 of the gradient loads and {bs}/16 of the parameter stores "
                              "(32 KB blocks) peer-to-peer, the rest multimem.ld_reduce / multimem.st through NVSwitch "
                              "(+2 barriers)")
        if ms_fused < ms_nccl:
            ms, route = ms_fused, fused_name
            launches = n_l
            clocks = ClockSampler.summarize(sampler.window(t0f, t1f)) if sampler else clocks

    # breakdown (each max over ranks)
    ms_update, _, _ = _timed(lambda: [opt._update_bucket(b, float(world)) for b in range(L.nbuckets)], args.steps, device)
    # what a training loop still sees when the drop-in class has reduce-scattered the gradients during backward (hooks):
    # update + all-gather pipeline only
    ms_uag, _, _ = _timed(lambda: (opt.noop.zero_(), opt.update_and_gather(float(world))), max(2, args.steps // 2), device)
    ms_rs, _, _ = _timed(lambda: [opt._reduce_scatter(b) for b in range(L.nbuckets)], max(2, args.steps // 2), device)
    ms_ag, _, _ = _timed(lambda: [opt._all_gather(b) for b in range(L.nbuckets)], max(2, args.steps // 2), device)
    rs_s, ag_s = opt.rs_stream, opt.ag_stream
    opt.rs_stream = None
    ms_serial, _, _ = _timed(opt.step, max(2, args.steps // 2), device)
    opt.rs_stream, opt.ag_stream = rs_s, ag_s

    # e2e: every rank's gradients arrive from pinned host memory each step (H2D inside the timed region), the step's
    # skip flag is read back (D2H).  PCIe-bound: 2 B/param per rank per step.
    e2e = None
    if not args.no_e2e:
        win = min(L.padded, 1 << 30)                       # 2 GiB pinned window per rank, cycled over the flat space
        host = torch.empty(win, dtype=torch.bfloat16, pin_memory=True)
        host.copy_(opt.grads_flat[:win])
        host_flag = torch.zeros(1, dtype=torch.int32, pin_memory=True)

        copy_stream = torch.cuda.Stream(device)
        use_fused = [False]
        bucket_evs = [torch.cuda.Event() for _ in range(L.nbuckets)]

        def e2e_step():
            main = torch.cuda.current_stream(device)
            copy_stream.wait_stream(main)        # the previous step is done with the gradient buffer
            with torch.cuda.stream(copy_stream):
                for b in range(L.nbuckets):
                    lo, hi = L.bucket_range(b)
                    for o in range(lo, hi, win):
                        n = min(win, hi - o)
                        opt.grads_flat[o:o + n].copy_(host[:n], non_blocking=True)
                    bucket_evs[b].record(copy_stream)
            if use_fused[0]:                     # one kernel over everything: it can only start after the last copy
                main.wait_stream(copy_stream)
                fused_fn()
            else:                                # bucket b's reduce-scatter starts as soon as bucket b has landed
                opt.step_count += 1
                opt.noop.zero_()
                for b in range(L.nbuckets):
                    main.wait_event(bucket_evs[b])
                    opt.reduce_gradients([b], side_stream=True)
                opt.update_and_gather(float(world))
            host_flag.copy_(opt.noop, non_blocking=True)
            main.synchronize()

        e2e_step()
        ms_e2e, _, _ = _timed(e2e_step, max(2, min(args.steps, 3)), device)
        e2e_routes = {"nccl_bucket_overlapped": ms_e2e}
        if fused_fn is not None:
            use_fused[0] = True
            e2e_step()
            ms_f, _, _ = _timed(e2e_step, max(2, min(args.steps, 3)), device)
            e2e_routes["fused_after_last_copy"] = ms_f
            ms_e2e = min(ms_e2e, ms_f)
        e2e = {"value": P / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 2 * L.padded * world,
               "d2h_bytes_per_step": 4 * world, "ms_per_step": ms_e2e, "routes_ms": e2e_routes,
               "note": "each rank copies its full bf16 gradient buffer from pinned host memory every step, bucket by "
                       "bucket on a copy stream; on the NCCL route a bucket's reduce-scatter starts as soon as the bucket "
                       "has landed (the fused one-kernel route has to wait for the last copy); the skip flag is read "
                       "back; bytes are summed over ranks"}
    if sampler:
        sampler.stop()
    parity = parity_block(amp_C, rank, world, device)
    if rank == 0:
        peak, peak_src = peaks()
        shard_bytes = BYTES_PER_PARAM * L.state_elems
        achieved = shard_bytes / (ms_update * 1e-3) / 1e9
        nvlink_bytes = 2 * L.padded * (world - 1) / world  # bytes each GPU sends (and receives) per collective
        line = {
            "metric": METRIC, "value": P / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{args.workload}: ZeRO-2 sharded mixed-precision AdamW ({P} params, {len(numels)} tensors): "
                                   f"reduce-scatter bf16 grads -> shard update (fp32 p/m/v) -> all-gather bf16 params",
                       "parallelism": f"optimizer state sharded x{world}; route = {route}",
                       "bytes_per_param": BYTES_PER_PARAM, "l2": "inputs >> L2; no flush needed", "hyper": HP},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": None, "peak_source": peak_src, "kernel": "mt_table_kernel<DistAdamOp<bf16,copy,noop>> (shard update only)",
                         "algorithmic_bytes_per_launch": shard_bytes / max(1, L.nbuckets)},
            "route": route, "fused_p2p_error": fused_err,
            "breakdown_ms": {"step": ms, "step_nccl_overlapped": ms_nccl, "step_fused_p2p": ms_p2p,
                             "step_fused_multicast": ms_mc, "step_fused_hybrid": ms_hy,
                             "hybrid_best_p2p_share": hybrid_best, "hybrid_search_ms": hybrid_search,
                             "step_nccl_serial": ms_serial, "update_only": ms_update,
                             "update_and_gather_after_overlapped_grad_sync": ms_uag,
                             "reduce_scatter_only": ms_rs, "all_gather_only": ms_ag},
            "update_only": {"params_per_s_whole_job": P / (ms_update * 1e-3), "hbm_gbs_per_gpu": achieved},
            "nvlink": {"bytes_sent_per_gpu_per_collective": nvlink_bytes,
                       "reduce_scatter_GBps_per_gpu": nvlink_bytes / (ms_rs * 1e-3) / 1e9,
                       "all_gather_GBps_per_gpu": nvlink_bytes / (ms_ag * 1e-3) / 1e9},
            "clocks": clocks, "gpu_launches": int(launches), "parity": parity,
            "e2e": e2e if e2e is not None else {"value": None, "unit": UNIT, "h2d_bytes_per_step": None,
                                                "d2h_bytes_per_step": None, "note": "skipped (--no-e2e)"},
        }
        print(json.dumps(line), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0
