"""tools/sweep_p2p.py -- time the fused reduce-scatter+Adam+all-gather kernel under its tuning knobs.
   torchrun --nproc-per-node N tools/sweep_p2p.py        (one process per GPU; prints on rank 0)"""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import torch.distributed as dist

import workloads
from bench_dist import _timed
from beforeholiday_b200.contrib.optimizers import FlatShardedAdam as DistributedFusedAdam

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
numels = workloads.numels(workloads.llama3_8b())
opt = DistributedFusedAdam(numels, lr=1e-4, betas=(0.9, 0.95), weight_decay=0.1, device=dev, fused_p2p=True)
opt.params_flat.normal_(0, 0.02)
opt.init_master_from_params()
opt.grads_flat.normal_(0, 1e-3)
P = sum(numels)
for mc in ([False, True] if opt.has_multicast else [False]):
    for tune in (0,):   # (round 1 also swept 1/2 groups in flight x 2/3/4 CTAs per SM: no effect, see the log)
        for tile in (0, 4096, 8192, 32768):
            os.environ["MT_TILE_ELEMS"] = str(tile)
            fn = lambda: opt.step_fused(multicast=mc)
            for _ in range(2):
                fn()
            ms, _, _ = _timed(fn, 4, dev)
            if rank == 0:
                print(f"world={world} multicast={mc} tune={tune} tile={tile or 'auto'}: {ms:.2f} ms  {P / ms / 1e6:.1f} Gparam/s", flush=True)
dist.barrier()
dist.destroy_process_group()
