"""bench_configs.py -- the N = 1 side table of bench.py: BASELINE.json's configs C1-C4, each
timed the same way as the headline C5 step and reported inside the same JSON line under "configs".

Per config: W >= 3 warm-up steps, then rounds of K steps; every step sits between its own pair of CUDA events on the
launching stream, with an L2 flush (outside the event pair) before it, so small working sets (C1 fits L2 many times
over) are measured from HBM like the large ones.  The flush is a 256 MiB write FOLLOWED BY a 256 MiB read of a second
buffer: a write alone leaves L2 full of dirty lines, and the timed pass then pays for writing them back (126 MB of
DRAM writes that are not its own: +3.4 us on the 249 MB norm pass, +15.5 us on the 498 MB one, tools/norm_trace.py
--clean-flush / profiles/r2_norm_trace_flush_modes.txt); after the read pass L2 holds clean lines of a foreign buffer.
Passes under 2 GB also report `ms_after_write_only_flush` for comparison with earlier records.  Rounds repeat until the config has run for ~0.4 s so
that the nvidia-smi sampler sees it; `ms` is the median over rounds of the per-round mean step time.  `gbs` =
bytes_per_param x params / ms with SURVEY 8(d)'s algorithmic bytes; `frac` = gbs / MEASURED_PEAKS hbm_gbs.  Passes that
move less than 2 GB also carry `copy_same_bytes_ms`: a plain `dst.copy_(src)` of the same total bytes under the same
timer -- what this GPU attains at that size, launch latency and ramp included (`frac_of_copy_same_bytes` = that / ms).
`reference_cuda_ms`: the reference's own kernels (oracle/_ref, compiled unmodified for sm_100a) on the same buffers,
same timing -- the kernel to beat; measurement infrastructure only, never on the product path.
"""
from __future__ import annotations

import importlib.machinery
import importlib.util
import sysconfig
import time
from pathlib import Path

import torch

import workloads

ROOT = Path(__file__).resolve().parent


def _load_ref(name):
    p = ROOT / "oracle" / "_ref" / (name + sysconfig.get_config_var("EXT_SUFFIX"))
    if not p.exists():
        return None
    try:
        loader = importlib.machinery.ExtensionFileLoader(name, str(p))
        mod = importlib.util.module_from_spec(importlib.util.spec_from_loader(name, loader))
        loader.exec_module(mod)
        return mod
    except Exception:  # noqa: BLE001
        return None


class _Timer:
    def __init__(self, device, steps, warmup, sampler, summarize):
        self.device, self.steps, self.warmup = device, steps, max(3, warmup)
        self.flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=device)
        self.flush_read = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=device)
        self.sampler, self.summarize = sampler, summarize

    def __call__(self, fn, min_wall_s=0.45, max_rounds=100000, write_only_flush=False):
        """-> (gpu ms per step, host enqueue ms per step, clocks record)."""
        for _ in range(self.warmup):
            fn()
        torch.cuda.synchronize(self.device)
        rounds, host = [], []
        t0 = time.time()
        while True:
            evs = []
            h = 0.0
            for _ in range(self.steps):
                self.flush.zero_()
                if not write_only_flush:
                    self.flush_read.sum()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                s = time.perf_counter()
                fn()
                h += time.perf_counter() - s
                b.record()
                evs.append((a, b))
            torch.cuda.synchronize(self.device)
            rounds.append(sum(a.elapsed_time(b) for a, b in evs) / self.steps)
            host.append(h / self.steps * 1e3)
            if time.time() - t0 >= min_wall_s or len(rounds) >= max_rounds:
                break
        t1 = time.time()
        rounds.sort()
        host.sort()
        clocks = self.summarize(self.sampler.window(t0, t1)) if self.sampler else None
        return rounds[len(rounds) // 2], host[len(host) // 2], clocks, len(rounds)


def _alloc(shapes, dtype, device, scale=0.02):
    out = []
    for s in shapes:
        t = torch.empty(s, dtype=dtype, device=device)
        t.normal_(0, scale)
        out.append(t)
    return out


def run_configs(amp_C, device, steps, warmup, peak, sampler, summarize, log, want=("C1", "C2", "C3", "C4"), with_ref=True):
    ref = _load_ref("amp_C") if with_ref else None
    timer = _Timer(device, steps, warmup, sampler, summarize)
    flag = torch.zeros(1, dtype=torch.int32, device=device)
    out = {}

    def entry(name, op, bpp, P, fn, ref_fn=None, extra=None, ntensors=None):
        l0 = amp_C.launch_count()
        fn()
        launches = amp_C.launch_count() - l0
        ms, host_ms, clocks, rounds = timer(fn)
        gbs = bpp * P / (ms * 1e-3) / 1e9
        d = {"op": op, "params": P, "bytes_per_param": bpp, "ms": ms, "gbs": gbs, "frac": gbs / peak,
             "params_per_s": P / (ms * 1e-3), "launches_per_step": int(launches), "host_enqueue_ms": host_ms,
             "steps": steps, "rounds": rounds, "clocks": clocks, "reference_cuda_ms": None}
        if ntensors:
            d["tensors"] = ntensors
            d["host_ns_per_tensor"] = host_ms * 1e6 / ntensors
        if ref_fn is not None:
            try:
                rms, rhost, _, _ = timer(ref_fn, min_wall_s=0.2)
                d["reference_cuda_ms"] = rms
                d["reference_cuda_host_enqueue_ms"] = rhost
                d["speedup_vs_reference_cuda"] = rms / ms
            except Exception as e:  # noqa: BLE001
                d["reference_cuda_error"] = repr(e)[:160]
        if bpp * P < 2e9:
            # attainable at THIS size: a plain device copy that moves the same number of bytes, timed the same way
            # (MEASURED_PEAKS' copy figure is taken on 2 GiB+; launch latency and ramp weigh more on a sub-GB pass)
            half = max(1, int(bpp * P) // 2)
            src, dst = torch.empty(half, dtype=torch.uint8, device=device), torch.empty(half, dtype=torch.uint8, device=device)
            cms, _, _, _ = timer(lambda: dst.copy_(src), min_wall_s=0.15)
            d["copy_same_bytes_ms"] = cms
            d["frac_of_copy_same_bytes"] = cms / ms
            d["ms_after_write_only_flush"] = timer(fn, min_wall_s=0.15, write_only_flush=True)[0]
            del src, dst
        if extra:
            d.update(extra)
        out[name] = d
        log(f"[bench] {name}: {op}: {ms:.4f} ms, {gbs:.0f} GB/s ({gbs / peak * 100:.1f}% of peak), "
            f"{launches} launches, ref_cuda {d['reference_cuda_ms']}")

    # ---- C1: ~1M-param fp32 MLP, FusedAdam's kernel (28 B/param) -------------------------------------------------------
    if "C1" in want:
        shapes = workloads.c1_mlp()
        P = workloads.total(shapes)
        g, p, m, v = (_alloc(shapes, torch.float32, device) for _ in range(4))
        v = [t.abs_() for t in v]
        mk = lambda mod: (lambda: mod.multi_tensor_adam(8192, flag, [g, p, m, v], 5e-4, 0.9, 0.999, 1e-8, 3, 0, 1, 0.0))
        entry("C1", "multi_tensor_adam fp32, 6 tensors (launch-latency bound: 27.6 MB per step)", 28, P, mk(amp_C),
              mk(ref) if ref else None, ntensors=len(shapes))
        del g, p, m, v

    # ---- C2: GPT-2-small amp step ---------------------------------------------------------------------------------------
    if "C2" in want:
        shapes = workloads.gpt2_small()
        P = workloads.total(shapes)
        nt = len(shapes)
        g16 = _alloc(shapes, torch.float16, device, 1.0)          # true grad x loss scale, finite
        g32 = [torch.empty_like(t, dtype=torch.float32) for t in g16]
        p32, m, v = (_alloc(shapes, torch.float32, device) for _ in range(3))
        v = [t.abs_() for t in v]
        p16 = [torch.empty_like(t, dtype=torch.float16) for t in p32]

        def three_pass(mod):
            def fn():
                mod.multi_tensor_scale(8192, flag, [g16, g32], 1.0 / 65536)
                mod.multi_tensor_adam(8192, flag, [g32, p32, m, v], 5e-4, 0.9, 0.999, 1e-8, 3, 1, 1, 0.01)
                mod.multi_tensor_scale(8192, flag, [p32, p16], 1.0)
            return fn

        entry("C2", "amp step as the reference runs it: unscale fp16->fp32 + Adam fp32 + master->model fp16 copy (3 passes)",
              40, P, three_pass(amp_C), three_pass(ref) if ref else None, ntensors=nt)
        hp = [torch.full((nt,), val, dtype=dt, device=device) for val, dt in
              ((0.9, torch.float32), (0.999, torch.float32), (1, torch.int32), (1e-8, torch.float32), (0.01, torch.float32))]

        def fused():
            flag.zero_()
            amp_C.multi_tensor_l2norm(65536, flag, [g16], False)
            amp_C.multi_tensor_fused_adam(8192, flag, [p32, m, v, g16, p16], *hp, 5e-4, 65536.0, 3, 1, True)

        entry("C2_fused", "fused route: inf/norm check on fp16 grads (2 B) + mixed-precision Adam with fp16 copy-out (28 B)",
              30, P, fused, None, ntensors=nt)
        entry("C2_l2norm_fp16", "multi_tensor_l2norm fp16, chunk 65536 (one launch, in-launch finalize)", 2, P,
              lambda: amp_C.multi_tensor_l2norm(65536, flag, [g16], False),
              (lambda: ref.multi_tensor_l2norm(65536, flag, [g16], False)) if ref else None, ntensors=nt)
        entry("C2_l2norm_fp32", "multi_tensor_l2norm fp32 per-tensor, chunk 65536", 4, P,
              lambda: amp_C.multi_tensor_l2norm(65536, flag, [g32], True),
              (lambda: ref.multi_tensor_l2norm(65536, flag, [g32], True)) if ref else None, ntensors=nt)
        entry("C2_scale_fp16_fp32", "multi_tensor_scale fp16->fp32 (unscale + overflow check)", 6, P,
              lambda: amp_C.multi_tensor_scale(8192, flag, [g16, g32], 1.0 / 65536),
              (lambda: ref.multi_tensor_scale(8192, flag, [g16, g32], 1.0 / 65536)) if ref else None, ntensors=nt)
        # overflow variant: inf at element 0 of tensor 0, nan at the last element of the last tensor -> the flag must
        # come up and the update must leave p/m/v untouched (the step is skipped on the device, no host round trip)
        keep0, keepl = g16[0].view(-1)[0].clone(), g16[-1].view(-1)[-1].clone()
        g16[0].view(-1)[0] = float("inf")
        g16[-1].view(-1)[-1] = float("nan")
        before = [float(x[0].double().sum()) for x in (p32, m, v)]
        fused()
        torch.cuda.synchronize(device)
        skipped_ok = int(flag.item()) == 1 and before == [float(x[0].double().sum()) for x in (p32, m, v)]
        entry("C2_fused_overflow", "fused route with inf/nan gradients: norm pass raises the flag, the update kernel exits "
              "at once (step skipped on device)", 2, P, fused, None, extra={"flag_raised_and_state_untouched": bool(skipped_ok)},
              ntensors=nt)
        g16[0].view(-1)[0] = keep0
        g16[-1].view(-1)[-1] = keepl
        flag.zero_()
        del g16, g32, p32, m, v, p16, hp

    # ---- C3: 50 000 small tensors -- SGD momentum + scale; the reference is launch/host bound here --------------------
    if "C3" in want:
        shapes = workloads.c3_many_small()
        P = workloads.total(shapes)
        nt = len(shapes)
        g, w, mom = (_alloc(shapes, torch.float32, device) for _ in range(3))
        g16 = [torch.empty_like(t, dtype=torch.float16).normal_(0, 1.0) for t in g]
        mk = lambda mod: (lambda: mod.multi_tensor_sgd(8192, flag, [g, w, mom], 1e-4, 0.9, 0.0, 0.25, True, False, False, 1.0))
        entry("C3_sgd", "multi_tensor_sgd momentum+nesterov+wd fp32 over 50 000 tensors", 20, P, mk(amp_C),
              mk(ref) if ref else None, ntensors=nt)
        mk = lambda mod: (lambda: mod.multi_tensor_scale(8192, flag, [g16, g], 1.0 / 128))
        entry("C3_scale", "multi_tensor_scale fp16->fp32 over 50 000 tensors", 6, P, mk(amp_C), mk(ref) if ref else None,
              ntensors=nt)
        del g, w, mom, g16

    # ---- C4: BERT-large FusedLAMB step (grad-norm pass + stage 1 with both per-tensor norms + stage 2) ----------------
    if "C4" in want:
        shapes = workloads.bert_large()
        P = workloads.total(shapes)
        g, p, m, v = (_alloc(shapes, torch.float32, device) for _ in range(4))
        v = [t.abs_() for t in v]

        def lamb(mod):
            def fn():
                gn, _ = mod.multi_tensor_l2norm(65536, flag, [g], False)
                mod.multi_tensor_lamb(8192, flag, [g, p, m, v], 5e-4, 0.9, 0.999, 1e-6, 3, 1, 0.01, 1, 1, gn, 1.0, False)
            return fn

        entry("C4", "LAMB step incl. global grad norm, fp32, ~398 tensors (44 B/param here; the reference moves 52)", 44, P,
              lamb(amp_C), lamb(ref) if ref else None, ntensors=len(shapes))
        del g, p, m, v
    del timer
    torch.cuda.empty_cache()
    return out
