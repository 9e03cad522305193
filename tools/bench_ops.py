#!/usr/bin/env python
"""tools/bench_ops.py -- per-config table: this library vs the reference's CUDA kernels
(oracle/_ref, compiled unmodified for sm_100a) on the same B200, same buffers, CUDA-event timing.

    python tools/bench_ops.py [--configs c1,c2,c3,c4] [--iters 10]

Prints one JSON object per measurement to stdout.

Measurement infrastructure, not product: like bench.py's `reference_cuda` / `cpu_baseline` legs it loads oracle/_ref (the
reference's kernels) and oracle/mt_oracle (the CPU restatement) as BASELINES timed beside this library; nothing under
beforeholiday_b200/ imports either.
"""
from __future__ import annotations

import argparse
import importlib.machinery
import importlib.util
import json
import sys
import sysconfig
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

import workloads  # noqa: E402


def load_ref(name):
    p = ROOT / "oracle" / "_ref" / (name + sysconfig.get_config_var("EXT_SUFFIX"))
    if not p.exists():
        return None
    loader = importlib.machinery.ExtensionFileLoader(name, str(p))
    mod = importlib.util.module_from_spec(importlib.util.spec_from_loader(name, loader))
    loader.exec_module(mod)
    return mod


def timeit(fn, iters, warm=3, flush=None, burst=8):
    """GPU ms per call (CUDA events) and host wall ms per call (enqueue cost).

    gpu_ms is taken over a BURST of back-to-back calls inside one event pair, so that the host's enqueue latency
    (tens of microseconds: list walk, table lookup, torch allocations) overlaps GPU execution as it does in a training
    loop; timing single calls would charge that latency to sub-100-microsecond kernels.  Working sets here exceed L2
    (>= 250 MB), so consecutive calls do not hit in cache; `flush` is still applied between bursts."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    gpu, host = [], []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()  # > L2: evict between iterations for working sets that would otherwise sit in L2
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        for _ in range(burst):
            fn()
        e1.record()
        host.append((time.perf_counter() - t0) * 1e3 / burst)
        torch.cuda.synchronize()
        gpu.append(e0.elapsed_time(e1) / burst)
    gpu.sort()
    host.sort()
    return gpu[len(gpu) // 2], host[len(host) // 2]


def alloc(shapes, dtype, scale=0.02):
    out = []
    for s in shapes:
        t = torch.empty(s, dtype=dtype, device="cuda")
        t.normal_(0, scale)
        out.append(t)
    return out


def report(config, op, impl, ms, host_ms, nbytes, nparams, launches=None, extra=None):
    d = {"config": config, "op": op, "impl": impl, "gpu_ms": round(ms, 4), "host_ms": round(host_ms, 4),
         "GBps": round(nbytes / ms / 1e6, 1), "Gparam_s": round(nparams / ms / 1e6, 2),
         "pct_of_8TBps": round(nbytes / ms / 1e6 / 80, 1)}
    if launches is not None:
        d["launches"] = launches
    if extra:
        d.update(extra)
    print(json.dumps(d), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c2,c3,c4")
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    import beforeholiday_b200

    mine = beforeholiday_b200.load_amp_C()
    ref = load_ref("amp_C")
    impls = [("b200", mine)] + ([("reference_cuda", ref)] if ref is not None else [])
    flag = torch.zeros(1, dtype=torch.int32, device="cuda")
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    cfgs = args.configs.split(",")

    def launches_of(fn):
        a = mine.launch_count()
        fn()
        return mine.launch_count() - a

    if "c1" in cfgs:
        shapes = workloads.c1_mlp()
        P = workloads.total(shapes)
        g, p, m, v = (alloc(shapes, torch.float32) for _ in range(4))
        v = [t.abs() for t in v]
        for name, mod in impls:
            fn = lambda: mod.multi_tensor_adam(8192, flag, [g, p, m, v], 5e-4, 0.9, 0.999, 1e-8, 3, 0, 1, 0.0)
            ms, host = timeit(fn, args.iters, flush=flush, burst=1)
            report("C1 1M-param MLP fp32", "adam", name, ms, host, 28 * P, P, launches_of(fn) if mod is mine else None)

    if "c2" in cfgs:
        shapes = workloads.gpt2_small()
        P = workloads.total(shapes)
        g16 = alloc(shapes, torch.float16, 1.0)
        g32 = [torch.empty_like(t, dtype=torch.float32) for t in g16]
        p32, m, v = (alloc(shapes, torch.float32) for _ in range(3))
        v = [t.abs() for t in v]
        p16 = [torch.empty_like(t, dtype=torch.float16) for t in p32]
        for name, mod in impls:
            def three_pass():
                mod.multi_tensor_scale(8192, flag, [g16, g32], 1.0 / 65536)
                mod.multi_tensor_adam(8192, flag, [g32, p32, m, v], 5e-4, 0.9, 0.999, 1e-8, 3, 1, 1, 0.01)
                mod.multi_tensor_scale(8192, flag, [p32, p16], 1.0)
            ms, host = timeit(three_pass, args.iters)
            report("C2 GPT-2 124M", "amp 3-pass: unscale+adam fp32+copy-out (40 B/param)", name, ms, host, 40 * P, P,
                   launches_of(three_pass) if mod is mine else None)
            for opname, fn, b in (
                ("scale fp16->fp32", lambda: mod.multi_tensor_scale(8192, flag, [g16, g32], 1.0 / 65536), 6),
                ("adam fp32", lambda: mod.multi_tensor_adam(8192, flag, [g32, p32, m, v], 5e-4, 0.9, 0.999, 1e-8, 3, 1, 1, 0.01), 28),
                ("scale fp32->fp16", lambda: mod.multi_tensor_scale(8192, flag, [p32, p16], 1.0), 6),
                ("axpby fp16+fp32->fp32", lambda: mod.multi_tensor_axpby(8192, flag, [g16, g32, g32], 1.0 / 65536, 1.0, 0), 10),
                ("l2norm fp16", lambda: mod.multi_tensor_l2norm(65536, flag, [g16], False), 2),
                ("l2norm fp32 per-tensor", lambda: mod.multi_tensor_l2norm(65536, flag, [g32], True), 4),
            ):
                ms, host = timeit(fn, args.iters, flush=flush)
                report("C2 GPT-2 124M", opname, name, ms, host, b * P, P)
        # fused route (this library only): inf/norm check on fp16 grads + mixed-precision Adam honouring the flag
        nt = len(shapes)
        hp = [torch.full((nt,), val, dtype=dt, device="cuda") for val, dt in
              ((0.9, torch.float32), (0.999, torch.float32), (1, torch.int32), (1e-8, torch.float32), (0.01, torch.float32))]
        def fused():
            flag.zero_()
            mine.multi_tensor_l2norm(65536, flag, [g16], False)
            mine.multi_tensor_fused_adam(8192, flag, [p32, m, v, g16, p16], *hp, 5e-4, 65536.0, 3, 1, True)
        ms, host = timeit(fused, args.iters)
        report("C2 GPT-2 124M", "fused: inf/norm check + mixed Adam + fp16 copy (30 B/param)", "b200", ms, host, 30 * P, P,
               launches_of(fused))
        del g16, g32, p32, m, v, p16

    if "c3" in cfgs:
        shapes = workloads.c3_many_small()
        P = workloads.total(shapes)
        g, w, mom = (alloc(shapes, torch.float32) for _ in range(3))
        g16 = [torch.empty_like(t, dtype=torch.float16) for t in g]
        for name, mod in impls:
            fn = lambda: mod.multi_tensor_sgd(8192, flag, [g, w, mom], 1e-4, 0.9, 0.0, 0.25, True, False, False, 1.0)
            ms, host = timeit(fn, max(3, args.iters // 2))
            report("C3 50k tensors", "sgd momentum fp32 (20 B/param)", name, ms, host, 20 * P, P,
                   launches_of(fn) if mod is mine else None, {"ns_per_tensor_host": round(host * 1e6 / len(shapes), 1)})
            fn = lambda: mod.multi_tensor_scale(8192, flag, [g16, g], 1.0 / 128)
            ms, host = timeit(fn, max(3, args.iters // 2))
            report("C3 50k tensors", "scale fp16->fp32 (6 B/param)", name, ms, host, 6 * P, P,
                   launches_of(fn) if mod is mine else None, {"ns_per_tensor_host": round(host * 1e6 / len(shapes), 1)})
        # launcher-overhead variant: only the <= 4096-element tensors
        small = [i for i, s in enumerate(shapes) if s[0] <= 4096]
        gs, ws, ms_ = [g[i] for i in small], [w[i] for i in small], [mom[i] for i in small]
        Ps = sum(t.numel() for t in gs)
        for name, mod in impls:
            fn = lambda: mod.multi_tensor_sgd(8192, flag, [gs, ws, ms_], 1e-4, 0.9, 0.0, 0.25, True, False, False, 1.0)
            t, host = timeit(fn, max(3, args.iters // 2))
            report("C3 small-only (%d tensors)" % len(small), "sgd momentum fp32", name, t, host, 20 * Ps, Ps,
                   launches_of(fn) if mod is mine else None, {"ns_per_tensor_host": round(host * 1e6 / len(small), 1)})
        del g, w, mom, g16

    if "c4" in cfgs:
        shapes = workloads.bert_large()
        P = workloads.total(shapes)
        g, p, m, v = (alloc(shapes, torch.float32) for _ in range(4))
        v = [t.abs() for t in v]
        for name, mod in impls:
            def lamb_step():
                gn, _ = mod.multi_tensor_l2norm(65536, flag, [g], False)
                mod.multi_tensor_lamb(8192, flag, [g, p, m, v], 5e-4, 0.9, 0.999, 1e-6, 3, 1, 0.01, 1, 1, gn, 1.0, False)
            ms, host = timeit(lamb_step, args.iters)
            report("C4 BERT-large 336M", "LAMB step incl. grad norm", name, ms, host, (44 if mod is mine else 52) * P, P,
                   launches_of(lamb_step) if mod is mine else None,
                   {"bytes_per_param": 44 if mod is mine else 52})
        # mixed-precision LAMB (fp16 grads + model copy, fp32 masters): the "_mp" kernels
        g16 = [t.half() for t in g]
        p16 = [t.half() for t in p]
        dv = lambda val, dt=torch.float32: torch.tensor([val], dtype=dt, device="cuda")
        lr_d, step_d, mx_d, inv_d, fi_d = dv(5e-4), dv(3, torch.int32), dv(1.0), dv(1.0), dv(0.0)
        def lamb_mp_step():
            gn, _ = mine.multi_tensor_l2norm_mp(65536, flag, [g16], False)
            mine.multi_tensor_lamb_mp(8192, flag, [g16, p, m, v, p16], lr_d, 0.9, 0.999, 1e-6, step_d, 1, 0.01, 1, 1, gn, mx_d,
                                      False, fi_d, inv_d)
        flag.zero_()
        ms, host = timeit(lamb_mp_step, args.iters)
        report("C4 BERT-large 336M", "LAMB-mp step incl. grad norm (fp16 g + copy, fp32 p/m/v)", "b200", ms, host, 38 * P, P,
               launches_of(lamb_mp_step), {"bytes_per_param": 38, "flag": int(flag.item())})
        del g16, p16
        # the CPU step beside it (north star): the oracle's LAMB restatement (OpenMP, all host cores) on the same set
        try:
            import os

            from oracle import mt_oracle as O

            hl = [[t.cpu() for t in l] for l in (g, p, m, v)]
            fl = torch.zeros(1, dtype=torch.int32)
            best = 1e30
            for _ in range(3):
                t0 = time.perf_counter()
                gn, _ = O.multi_tensor_l2norm(65536, fl, [hl[0]], False)
                O.multi_tensor_lamb(8192, fl, hl, 5e-4, 0.9, 0.999, 1e-6, 3, 1, 0.01, 1, 1, gn, 1.0, False)
                best = min(best, (time.perf_counter() - t0) * 1e3)
            report("C4 BERT-large 336M", "LAMB step incl. grad norm", f"cpu_oracle_port x{O.lib().orc_num_threads()} threads "
                   f"({os.cpu_count()} cores)", best, best, 52 * P, P)
        except Exception as e:  # noqa: BLE001
            print(json.dumps({"config": "C4", "cpu_oracle_error": repr(e)[:200]}), flush=True)

    if "c6" in cfgs:
        flag.zero_()   # earlier configs may have raised it (uninitialised fp16 inputs); these kernels honour the flag
        # sharded LAMB kernels on a BERT-large-sized shard: fp16 grads and model copy, fp32 p/m/v/u
        shapes = workloads.bert_large()
        P = workloads.total(shapes)
        nt = len(shapes)
        g16 = alloc(shapes, torch.float16, 1.0)
        p, m, v, u = (alloc(shapes, torch.float32) for _ in range(4))
        v = [t.abs() for t in v]
        p16 = [t.half() for t in p]
        T = lambda val, dt=torch.float32: torch.full((nt,), val, dtype=dt, device="cuda")
        hp = (T(0.9), T(0.999), T(0.1), T(1, torch.int32))
        eps, wd = T(1e-6), T(0.01)
        step, gs, lr = (torch.tensor([v_], dtype=dt, device="cuda") for v_, dt in ((3, torch.int32), (128.0, torch.float32), (1e-4, torch.float32)))
        gn = torch.tensor([4000.0], device="cuda")
        off = torch.arange(nt, device="cuda")
        seg = torch.arange(nt + 1, dtype=torch.int32, device="cuda")
        dl = load_ref("distributed_lamb_cuda")

        def mine_step():
            sq = mine.multi_tensor_lamb_compute_update_term(8192, flag, [g16, p, m, v, u], *hp, step, eps, 1, wd, gs, gn, 1.0,
                                                            seg, torch.float16)
            n = sq.sqrt()
            mine.multi_tensor_lamb_update_weights(8192, flag, [u, p, p16], n[0], n[1], off, lr, wd, gn, False)
        ms, host = timeit(mine_step, args.iters)
        report("C6 sharded LAMB, 336M-param shard", "stage1 (+fused norms) + stage2 (40 B/param)", "b200", ms, host, 40 * P, P,
               launches_of(mine_step))
        fn = lambda: mine.multi_tensor_lamb_compute_update_term(8192, flag, [g16, p, m, v, u], *hp, step, eps, 1, wd, gs, gn, 1.0, seg, torch.float16)
        ms, host = timeit(fn, args.iters)
        report("C6 sharded LAMB, 336M-param shard", "stage1 + fused norms (26 B/param)", "b200", ms, host, 26 * P, P)
        fn = lambda: mine.multi_tensor_lamb_compute_update_term(8192, flag, [g16, p, m, v, u], *hp, step, eps, 1, wd, gs, gn, 1.0)
        ms, host = timeit(fn, args.iters)
        report("C6 sharded LAMB, 336M-param shard", "stage1 (26 B/param)", "b200", ms, host, 26 * P, P)
        n2 = torch.ones(2, nt, device="cuda")
        fn = lambda: mine.multi_tensor_lamb_update_weights(8192, flag, [u, p, p16], n2[0], n2[1], off, lr, wd, gn, False)
        ms, host = timeit(fn, args.iters)
        report("C6 sharded LAMB, 336M-param shard", "stage2 (14 B/param)", "b200", ms, host, 14 * P, P)
        if dl is not None and ref is not None:
            def ref_step():
                _, pn = ref.multi_tensor_l2norm(65536, flag, [p16], True)
                dl.multi_tensor_lamb_compute_update_term(8192, flag, [g16, p, m, v, u], *hp, step, eps, 1, wd, gs, gn, 1.0)
                _, un = ref.multi_tensor_l2norm(65536, flag, [u], True)
                dl.multi_tensor_lamb_update_weights(8192, flag, [u, p, p16], pn, un, off, lr, wd, gn, False)
            ms, host = timeit(ref_step, args.iters)
            report("C6 sharded LAMB, 336M-param shard", "l2norm(p16) + stage1 + l2norm(u) + stage2 (46 B/param)", "reference_cuda",
                   ms, host, 46 * P, P)
            fn = lambda: dl.multi_tensor_lamb_compute_update_term(8192, flag, [g16, p, m, v, u], *hp, step, eps, 1, wd, gs, gn, 1.0)
            ms, host = timeit(fn, args.iters)
            report("C6 sharded LAMB, 336M-param shard", "stage1 (26 B/param)", "reference_cuda", ms, host, 26 * P, P)
            fn = lambda: dl.multi_tensor_lamb_update_weights(8192, flag, [u, p, p16], n2[0], n2[1], off, lr, wd, gn, False)
            ms, host = timeit(fn, args.iters)
            report("C6 sharded LAMB, 336M-param shard", "stage2 (14 B/param)", "reference_cuda", ms, host, 14 * P, P)


if __name__ == "__main__":
    main()
