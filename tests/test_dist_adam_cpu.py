"""The reference's DistributedFusedAdam test scenarios (apex/contrib/test/optimizers/test_dist_adam.py) against this
repo's drop-in class, on CPU: world 2, 3 (an odd size) and 4 (which also runs the 2 x 2 shard / replica grid) over gloo, with the oracle standing in for the CUDA update
(tests/cpu_engine.py).  What is under test is the optimizer's host logic: flat layout, param groups with their own lr,
post-backward hooks and bucket bookkeeping across micro-batches / no_sync, clip_grad_norm, the GradScaler protocol with
device-side skip, gather-on-root checkpointing.  The GPU twin (NCCL, real kernels) is tests/test_gpu_dist_adam.py."""
import os
import socket
import sys
import traceback

import pytest
import torch
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(__file__))
from dist_adam_cases import CASES  # noqa: E402


def _worker(rank, world, port, outdir):
    import torch.distributed as dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from cpu_engine import CpuDistributedFusedAdam

    results = {}
    try:
        for name, fn in CASES.items():
            try:
                fn(CpuDistributedFusedAdam, rank, "cpu")
                ok = torch.ones(1)
            except Exception:  # noqa: BLE001
                results[name] = traceback.format_exc()
                ok = torch.zeros(1)
            # keep the ranks in lockstep: a case that failed on one rank is abandoned by all before the next starts
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if name not in results:
                results[name] = None if bool(ok) else "failed on another rank"
        torch.save(results, os.path.join(outdir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


_cache = {}


def _results(world, tmp_path_factory):
    if world not in _cache:
        out = tmp_path_factory.mktemp(f"dist_adam_w{world}")
        ctx = mp.get_context("spawn")
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, str(out))) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(600)
        codes = [p.exitcode for p in procs]
        for p in procs:
            if p.exitcode is None:
                p.kill()
        assert all(c == 0 for c in codes), f"worker exit codes {codes}"
        _cache[world] = [torch.load(out / f"rank{r}.pt") for r in range(world)]
    return _cache[world]


@pytest.mark.timeout(900)
@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("case", list(CASES))
def test_dist_adam_reference_scenarios_cpu(case, world, tmp_path_factory):
    res = _results(world, tmp_path_factory)
    for rank, r in enumerate(res):
        assert r[case] is None, f"rank {rank}:\n{r[case]}"
