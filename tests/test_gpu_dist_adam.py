"""The reference's DistributedFusedAdam test scenarios (apex/contrib/test/optimizers/test_dist_adam.py:84-387) against
this repo's drop-in class on real GPUs: NCCL process groups of 2 and 4 ranks, the CUDA kernels, the three-stream
pipeline, post-backward hooks that reduce-scatter buckets while backward is still running, GradScaler with device-side
skip, gather-on-root checkpoints; plus the one-kernel NVLink route (fused_p2p) on a bf16 model.  World 1 (a one-rank
NCCL group: the class, its hooks, the scaler protocol and the kernels, without peers) runs on any GPU box; worlds 2 and 4
need `gpurun --gpus 2|4`.  The CPU twin (gloo + oracle) is tests/test_dist_adam_cpu.py."""
import datetime
import os
import socket
import sys
import traceback

import pytest
import torch

pytestmark = pytest.mark.gpu
sys.path.insert(0, os.path.dirname(__file__))
from dist_adam_cases import CASES, matches_pytorch  # noqa: E402

GPU_CASES = dict(CASES)
GPU_CASES["matches_pytorch_bf16_fused_p2p"] = lambda make, rank, device, **kw: matches_pytorch(
    make, rank, device, dtype=torch.bfloat16, param_sync_dtype=torch.bfloat16, grad_sync_dtype=torch.bfloat16,
    overlap_communication=False, fused_p2p=True, rtol=0.15, atol=0.15, **kw)
GPU_CASES["matches_pytorch_bf16_sync"] = lambda make, rank, device, **kw: matches_pytorch(
    make, rank, device, dtype=torch.bfloat16, param_sync_dtype=torch.bfloat16, grad_sync_dtype=torch.bfloat16,
    rtol=0.15, atol=0.15, **kw)
# (bf16 cases: the reference side is torch.optim.AdamW on bf16 parameters WITHOUT fp32 masters, so the two trajectories
# drift by a few bf16 ulps of the layer outputs (2^-4 at |y| ~ 8); these cases check that the 16-bit routes run and
# track, the bit-level parity of the fused kernel is tests/test_gpu_dist.py's and bench_dist.py's parity block's job)


def _worker(rank, world, port, outdir):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    # a case that fails on one rank only leaves the ranks in different collectives: keep the resulting hang short
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev, timeout=datetime.timedelta(seconds=120))
    results = {}
    try:
        from beforeholiday_b200.contrib.optimizers import DistributedFusedAdam

        for name, fn in GPU_CASES.items():
            try:
                fn(DistributedFusedAdam, rank, dev)
                torch.cuda.synchronize()
                ok = torch.ones(1, device=dev)
            except Exception:  # noqa: BLE001
                results[name] = traceback.format_exc()
                with open(os.path.join(outdir, f"rank{rank}.{name}.err"), "w") as f:   # survives a later hang / abort
                    f.write(results[name])
                ok = torch.zeros(1, device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if name not in results:
                results[name] = None if bool(ok.item()) else "failed on another rank"
        torch.save(results, os.path.join(outdir, f"rank{rank}.pt"))
    finally:
        try:
            torch.cuda.synchronize()
            dist.destroy_process_group()
        except Exception:  # noqa: BLE001
            traceback.print_exc()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


_cache = {}


def _results(world, tmp_path_factory):
    if world not in _cache:
        import torch.multiprocessing as mp

        out = tmp_path_factory.mktemp(f"gpu_dist_adam_w{world}")
        ctx = mp.get_context("spawn")
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, str(out))) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(900)
        codes = [p.exitcode for p in procs]
        for p in procs:
            if p.exitcode is None:
                p.kill()
        errs = "".join(f"--- {f.name} ---\n{f.read_text()[-3000:]}\n" for f in sorted(out.glob("rank*.err")))
        assert all(c == 0 for c in codes), f"worker exit codes {codes}\n{errs}"
        _cache[world] = [torch.load(out / f"rank{r}.pt") for r in range(world)]
    return _cache[world]


@pytest.mark.timeout(1200)
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("case", list(GPU_CASES))
def test_dist_adam_reference_scenarios_gpu(case, world, tmp_path_factory):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    res = _results(world, tmp_path_factory)
    for rank, r in enumerate(res):
        assert r[case] is None, f"rank {rank}:\n{r[case]}"
