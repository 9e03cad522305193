"""tools/norm_trace.py -- where do the ~13 us go that a norm pass spends beyond bytes / bandwidth?

    python tools/norm_trace.py [--dtype fp16|fp32] > profiles/r2_norm_trace_<dtype>.json

Builds a DEBUG copy of the table cache + norm entry points with -DMT_TRACE (tools/_trace/libmt_trace.so; the product
library carries none of this), runs multi_tensor_l2norm over GPT-2 small's 124 M gradients (chunk 65536, as FusedLAMB
calls it) and reads back per-CTA globaltimer stamps: entry, tile located (table lookup done), tile streamed, tile handed
in (block sum + completion tickets).  Prints the launch's timeline: bytes-in-progress over time (how many CTAs are
streaming in each microsecond), per-phase durations, and head / tail lengths."""
import argparse
import ctypes as C
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch

import workloads
from beforeholiday_b200 import _lib

CSRC = ROOT / "beforeholiday_b200" / "csrc"


def build():
    out = ROOT / "tools" / "_trace"
    out.mkdir(exist_ok=True)
    so = out / "libmt_trace.so"
    srcs = [CSRC / "mt_table.cu", CSRC / "mt_api_norm.cu"]
    deps = srcs + list(CSRC.glob("*.cuh")) + list(CSRC.glob("*.h"))
    if not so.exists() or so.stat().st_mtime < max(p.stat().st_mtime for p in deps):
        subprocess.run(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17",
                        "-lineinfo", "-DMT_TRACE", "-shared", "-cudart", "shared", "-Xcompiler", "-fPIC",
                        f"-I{ROOT / 'include'}", f"-I{CSRC}"] + [str(s) for s in srcs] + ["-o", str(so)], check=True)
    return so


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "fp32"])
    ap.add_argument("--per-tensor", type=int, default=0)
    ap.add_argument("--build-only", action="store_true")
    ap.add_argument("--no-flush", action="store_true", help="leave L2 as the previous launch left it")
    ap.add_argument("--clean-flush", action="store_true",
                    help="after writing the flush buffer, READ a second one: L2 is left full of clean lines, so the "
                         "traced launch does not pay for writing the flush's dirty lines back")
    args = ap.parse_args()
    so = build()
    if args.build_only:
        return
    L = C.CDLL(str(so))
    dev = torch.device("cuda", 0)
    dt = torch.float16 if args.dtype == "fp16" else torch.float32
    numels = workloads.numels(workloads.gpt2_small())
    xs = [torch.randn(n, device=dev, dtype=torch.float32).mul_(1e-3).to(dt) for n in numels]
    lists = _lib.Lists([xs])
    chunk = 65536
    L.mt_norm_scratch_bytes.restype = C.c_size_t
    need = L.mt_norm_scratch_bytes(chunk, lists.ref)
    ntiles = need // 4
    scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    noop = torch.zeros(1, dtype=torch.int32, device=dev)
    ret = torch.zeros(1, dtype=torch.float32, device=dev)
    ret_pt = torch.zeros(len(xs), dtype=torch.float32, device=dev)
    trace = torch.zeros(ntiles * 8, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    flush2 = torch.zeros(64 << 20, dtype=torch.float32, device=dev)
    stream = _lib.stream_ptr(dev)

    def call():
        rc = L.mt_multi_tensor_l2norm(chunk, _lib.vp(noop), lists.ref, args.per_tensor, _lib.vp(ret), _lib.vp(ret_pt),
                                      _lib.vp(scratch), C.c_size_t(need), stream)
        assert rc == 0, rc

    for _ in range(3):
        call()
    assert L.mt_debug_set_trace(_lib.vp(trace)) == 0
    runs = []
    for _ in range(3):
        if not args.no_flush:
            flush.zero_()
            if args.clean_flush:
                flush2.sum()
        trace.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        call()
        e1.record()
        torch.cuda.synchronize()
        t = trace.view(ntiles, 8).cpu().numpy()
        t0 = int(t[:, 0].min())
        entry, located, streamed, done = [(t[:, k] - t0) / 1e3 for k in range(4)]  # us
        end = float(done.max())
        tile_bytes = chunk * xs[0].element_size()
        # CTAs streaming (between "located" and "streamed") in each microsecond
        import numpy as np

        edges = np.arange(0, int(end) + 2)
        active = [int(((located < b + 1) & (streamed > b)).sum()) for b in edges[:-1]]
        resident = [int(((entry < b + 1) & (done > b)).sum()) for b in edges[:-1]]
        tiles_per_tensor = [(n + chunk - 1) // chunk for n in numels]
        first_tile = np.cumsum([0] + tiles_per_tensor)
        slow = []
        for i in np.argsort(done - streamed)[::-1][:6]:
            tensor = int(np.searchsorted(first_tile, i, side="right") - 1)
            slow.append({"tile": int(i), "tensor": tensor, "tiles_of_tensor": tiles_per_tensor[tensor],
                         "hand_in_us": float((done - streamed)[i]), "handed_in_at_us": float(done[i]),
                         "last_done_of_tensor_us": float(done[first_tile[tensor]:first_tile[tensor + 1]].max())})
        runs.append({
            "slowest_hand_ins": slow,
            "event_us": e0.elapsed_time(e1) * 1e3, "kernel_span_us": end, "ntiles": int(ntiles),
            "first_stream_start_us": float(located.min()), "last_stream_end_us": float(streamed.max()),
            "lookup_us_median": float(np.median(located - entry)), "stream_us_median": float(np.median(streamed - located)),
            "stream_us_p10_p90": [float(np.percentile(streamed - located, 10)), float(np.percentile(streamed - located, 90))],
            "hand_in_us_median": float(np.median(done - streamed)), "hand_in_us_max": float((done - streamed).max()),
            "entry_us_of_last_cta": float(entry.max()),
            "ideal_us_at_6550GBps": ntiles * tile_bytes / 6550e3,
            "ctas_streaming_per_us": active, "ctas_resident_per_us": resident,
            "sms_used": int(len(set(t[:, 4].tolist()))),
        })
    print(json.dumps({"dtype": args.dtype, "per_tensor": args.per_tensor, "l2_flushed": not args.no_flush, "clean_flush": args.clean_flush, "bytes": sum(numels) * xs[0].element_size(),
                      "runs": runs}))


if __name__ == "__main__":
    main()
