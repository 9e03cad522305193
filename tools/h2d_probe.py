"""tools/h2d_probe.py -- what bounds the end-to-end leg (gradients from pinned host memory every step) at N GPUs?

    torchrun --nproc-per-node N tools/h2d_probe.py [--gb 2] > profiles/r2_h2d_probe_nN.json

Every rank copies a --gb sized pinned host buffer to its GPU, all ranks at once and then one rank at a time, first with
the process placed wherever the launcher left it and then bound to the CPUs NVML names as local to the GPU
(bind_to_gpu below), with the pinned buffer allocated AFTER the binding so that its pages come
from the local NUMA node.  GB/s per rank; host topology (sockets, NUMA nodes, allowed CPUs) alongside.  Rank 0 prints one
JSON object."""
import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

def _nvml_handle(cuda_index):
    import pynvml

    pynvml.nvmlInit()
    try:  # CUDA_VISIBLE_DEVICES renumbers CUDA devices; NVML does not: go through the UUID
        import torch

        uuid = str(torch.cuda.get_device_properties(cuda_index).uuid)
        return pynvml, pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode() if not uuid.startswith("GPU-") else uuid.encode())
    except Exception:  # noqa: BLE001
        return pynvml, pynvml.nvmlDeviceGetHandleByIndex(cuda_index)


def gpu_local_cpus(cuda_index):
    """Sorted list of the CPUs NVML reports as local to the GPU ([] when NVML does not answer)."""
    try:
        pynvml, h = _nvml_handle(cuda_index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (max(os.cpu_count() or 1, 1) + 63) // 64)
    except Exception:  # noqa: BLE001
        return []
    return [w * 64 + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1]


def bind_to_gpu(cuda_index):
    """Restrict this process to the CPUs local to GPU `cuda_index` (intersected with the CPUs it is allowed to use).
    Returns the list it bound to, or None when there was nothing to do (NVML silent, the local set covers every allowed
    CPU, or the intersection is empty)."""
    if not hasattr(os, "sched_setaffinity"):
        return None
    allowed = os.sched_getaffinity(0)
    local = set(gpu_local_cpus(cuda_index)) & allowed
    if not local or local == allowed:
        return None
    os.sched_setaffinity(0, local)
    return sorted(local)


def h2d_gbs(host, devbuf, steps, dev, together):
    if together:
        dist.barrier()
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        devbuf.copy_(host, non_blocking=True)
    e1.record()
    torch.cuda.synchronize(dev)
    return host.numel() * host.element_size() * steps / e0.elapsed_time(e1) / 1e6


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gb", type=float, default=2.0)
    ap.add_argument("--steps", type=int, default=4)
    args = ap.parse_args()
    rank, world, lrank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lrank)
    dev = torch.device("cuda", lrank)
    dist.init_process_group("nccl", device_id=dev)
    n = int(args.gb * 1e9 / 2)
    devbuf = torch.empty(n, dtype=torch.bfloat16, device=dev)
    out = {"world": world, "gb": args.gb, "modes": {}}
    for mode in ("as_launched", "bound_to_gpu_cpus"):
        if mode == "bound_to_gpu_cpus":
            bound = bind_to_gpu(lrank)
        else:
            bound = None
        host = torch.empty(n, dtype=torch.bfloat16).pin_memory()
        host.fill_(1.0)  # touch every page from this process
        h2d_gbs(host, devbuf, 1, dev, True)
        together = h2d_gbs(host, devbuf, args.steps, dev, True)
        alone = 0.0
        for r in range(world):  # one rank at a time
            dist.barrier()
            if r == rank:
                alone = h2d_gbs(host, devbuf, args.steps, dev, False)
        rec = torch.tensor([together, alone, float(len(os.sched_getaffinity(0)))], device=dev)
        allrec = [torch.zeros_like(rec) for _ in range(world)]
        dist.all_gather(allrec, rec)
        out["modes"][mode] = {"together_GBps": [round(float(x[0]), 1) for x in allrec],
                              "alone_GBps": [round(float(x[1]), 1) for x in allrec],
                              "cpus_in_affinity": [int(x[2]) for x in allrec],
                              "sum_together_GBps": round(sum(float(x[0]) for x in allrec), 1),
                              "rank0_bound_to": bound}
        del host
    if rank == 0:
        def sh(cmd):
            try:
                return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=20).stdout.strip()
            except Exception as e:  # noqa: BLE001
                return repr(e)
        out["lscpu"] = sh("lscpu | grep -i 'numa\\|socket\\|model name\\|^CPU(s)'")
        out["topo"] = sh("nvidia-smi topo -m | head -12")
        out["gpu_local_cpus"] = {str(i): gpu_local_cpus(i) for i in range(world)}
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
