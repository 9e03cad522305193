"""tools/nvlink_probe.py -- what bounds the fused reduce-scatter + Adam + all-gather kernel on the NVSwitch fabric?

    torchrun --nproc-per-node N tools/nvlink_probe.py [--gb 4] > profiles/r2_nvlink_probe_nN.json

Separates the two link directions: per rank, over its own 1/N shard of two symmetric bf16 buffers
(the bench's gradient / parameter buffers in miniature), times
    mc_ld      multimem.ld_reduce only          (each GPU SENDS (N-1)/N of the buffer, receives 1/N)
    mc_st      multimem.st only                 (each GPU sends 1/N, RECEIVES (N-1)/N)
    mc_both    both in one kernel               (the fused kernel's traffic without the HBM stream and the math)
    p2p_ld / p2p_st / p2p_both                  the same with plain peer loads / stores
and the fused optimizer kernel itself for every hybrid split (ld_p2p_of16, st_p2p_of16) on a --gb sized parameter set.
GB/s figures are per GPU and per direction: bytes that cross this GPU's links in the busier direction / time.
NVLink byte counters (`nvidia-smi nvlink -gt d`) are sampled around the fused-kernel runs when the tool is available.
One JSON object on stdout (rank 0)."""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import torch.distributed as dist


def build_probe():
    src, out = ROOT / "tools" / "nvlink_probe.cu", ROOT / "tools" / "libnvlink_probe.so"
    if not out.exists() or out.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo",
                        "-shared", "-cudart", "shared", "-Xcompiler", "-fPIC", str(src), "-o", str(out)], check=True)
    return C.CDLL(str(out))


def timed(fn, steps, device):
    dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(device)
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=device)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def nvlink_counters(index):
    """Cumulative NVLink payload counters of one GPU in KiB: NVML field values (all links summed) when the driver
    exposes them, else `nvidia-smi nvlink -gt d`.  None when neither answers."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        vals = pynvml.nvmlDeviceGetFieldValues(h, [pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX,
                                                   pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX])
        if all(v.nvmlReturn == 0 for v in vals):
            return {"tx_kib": int(vals[0].value.ullVal), "rx_kib": int(vals[1].value.ullVal), "source": "nvml"}
    except Exception:  # noqa: BLE001
        pass
    try:
        r = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(index)], capture_output=True, text=True, timeout=20)
        tx = rx = 0
        for line in r.stdout.splitlines():
            if "Data Tx" in line:
                tx += int(line.split(":")[-1].split()[0])
            if "Data Rx" in line:
                rx += int(line.split(":")[-1].split()[0])
        if tx or rx:
            return {"tx_kib": tx, "rx_kib": rx, "source": "nvidia-smi"}
        nvlink_counters.raw = (r.stdout + r.stderr)[:400]
    except Exception as e:  # noqa: BLE001
        nvlink_counters.raw = repr(e)[:200]
    return None


nvlink_counters.raw = None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=4.0, help="size of each symmetric bf16 buffer in GB")
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    rank, world, lrank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lrank)
    dev = torch.device("cuda", lrank)
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm_mem

    lib = build_probe() if rank == 0 else None
    dist.barrier()
    lib = lib or build_probe()
    out = {"world": world, "buffer_gb": args.gb, "probe": {}, "fused": {}}

    # ---- raw probes -------------------------------------------------------------------------------------------------------
    n = int(args.gb * 1e9 / 2) // (world * 4096) * (world * 4096)
    g = symm_mem.empty(n, dtype=torch.bfloat16, device=dev)
    p = symm_mem.empty(n, dtype=torch.bfloat16, device=dev)
    g.fill_(1.0)
    p.zero_()
    hg = symm_mem.rendezvous(g, group=dist.group.WORLD)
    hp = symm_mem.rendezvous(p, group=dist.group.WORLD)
    gb = (C.c_void_p * world)(*[int(x) for x in hg.buffer_ptrs])
    pb = (C.c_void_p * world)(*[int(x) for x in hp.buffer_ptrs])
    gmc, pmc = int(getattr(hg, "multicast_ptr", 0) or 0), int(getattr(hp, "multicast_ptr", 0) or 0)
    out["has_multicast"] = bool(gmc and pmc)
    sink = torch.zeros(4, dtype=torch.int32, device=dev)
    shard = n * 2 // world
    stream = torch.cuda.current_stream(dev).cuda_stream
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    for mc in ([1, 0] if out["has_multicast"] else [0]):
        for mode, name in ((1, "ld"), (2, "st"), (3, "both")):
            def fn():
                hg.barrier(channel=0)
                rc = lib.nvlink_probe(world, mc, gb, pb, C.c_void_p(gmc), C.c_void_p(pmc), C.c_size_t(rank * shard),
                                      C.c_size_t(shard), mode, sms * 8, C.c_void_p(sink.data_ptr()), C.c_void_p(stream))
                assert rc == 0, rc
                hp.barrier(channel=1)
            fn()
            ms = timed(fn, args.steps, dev)
            # bytes over this GPU's links in the busier direction.  Peer-to-peer: (N-1)/N of the buffer (the local shard
            # stays local).  Through the switch the requester's own copy travels too (the switch reads / writes every
            # member of the multicast group, this GPU included): the whole buffer.
            link_bytes = n * 2 * (1.0 if mc else (world - 1) / world)
            out["probe"][("mc_" if mc else "p2p_") + name] = {"ms": ms, "GBps_per_dir": link_bytes / ms / 1e6}
    del g, p, hg, hp

    # ---- the fused optimizer kernel under every split ---------------------------------------------------------------------
    from beforeholiday_b200.contrib.optimizers import FlatShardedAdam

    P = int(args.gb * 1e9 / 2)
    numels = [64 * 1024 * 1024] * max(1, P // (64 * 1024 * 1024))
    opt = FlatShardedAdam(numels, lr=1e-4, betas=(0.9, 0.95), weight_decay=0.1, device=dev, fused_p2p=True)
    opt.params_flat.normal_(0, 0.02)
    opt.init_master_from_params()
    opt.grads_flat.normal_(0, 1e-3)
    Pn = sum(numels)
    link_bytes = Pn * 2 * (world - 1) / world
    splits = [(0, 0), (16, 16), (16, 0), (0, 16), (2, 2), (3, 3), (4, 4), (6, 6), (8, 8), (4, 0), (0, 4), (8, 0), (0, 8)]
    if not opt.has_multicast:
        splits = [(16, 16)]
    for ld16, st16 in splits:
        fn = lambda: opt.step_fused(multicast=opt.has_multicast, ld_p2p_of16=ld16, st_p2p_of16=st16)
        for _ in range(2):
            fn()
        c0 = nvlink_counters(lrank) if rank == 0 else None
        ms = timed(fn, args.steps, dev)
        c1 = nvlink_counters(lrank) if rank == 0 else None
        rec = {"ms": ms, "Gparam_s": Pn / ms / 1e6, "ms_scaled_to_llama3_8b": ms * 8030261248 / Pn}
        if c0 and c1:
            rec["nvlink_tx_GB_per_step"] = (c1["tx_kib"] - c0["tx_kib"]) * 1024 / 1e9 / args.steps
            rec["nvlink_rx_GB_per_step"] = (c1["rx_kib"] - c0["rx_kib"]) * 1024 / 1e9 / args.steps
        out["fused"][f"ld{ld16}_st{st16}"] = rec
    out["nvlink_counter_note"] = nvlink_counters.raw
    out["fused_params"] = Pn
    out["fused_link_bytes_ideal_per_dir"] = link_bytes
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
