"""Test infrastructure: gloo + oracle stand-ins for the NCCL collectives and the CUDA kernels of the sharded Adam
engine, so that the HOST logic of FlatShardedAdam / DistributedFusedAdam (layout, fragments, bucket bookkeeping, hooks,
GradScaler protocol, checkpointing) runs on CPU at world size > 1.  Nothing here is product code."""
import torch
import torch.distributed as dist

from beforeholiday_b200.contrib.optimizers import DistributedFusedAdam, FlatShardedAdam


def _scalar(x):
    return float(x.reshape(-1)[0]) if isinstance(x, torch.Tensor) else float(x)


class CpuShardedAdam(FlatShardedAdam):
    def _init_comm(self):
        self.rs_group = self.ag_group = self.group
        self.rs_stream = self.ag_stream = None

    def _reduce_scatter(self, b):
        lo, hi = self.layout.bucket_range(b)
        buf = self.grads_flat[lo:hi].float()
        if self.world > 1:
            dist.all_reduce(buf, group=self.group)
        olo, ohi = self.layout.owned_range(b)
        self.grads_flat[olo:ohi].copy_(buf[olo - lo:ohi - lo])  # only the owned sub-shard is defined afterwards

    def _reduce_redundant(self, b):
        if self.redundant_size > 1:
            olo, ohi = self.layout.owned_range(b)
            buf = self.grads_flat[olo:ohi].float()
            dist.all_reduce(buf, group=self.redundant_group)
            self.grads_flat[olo:ohi].copy_(buf)

    def _all_gather(self, b):
        if self.world == 1:
            return
        lo, hi = self.layout.bucket_range(b)
        olo, ohi = self.layout.owned_range(b)
        parts = [torch.empty(self.layout.sub_shard, dtype=torch.float32) for _ in range(self.world)]
        dist.all_gather(parts, self.params_flat[olo:ohi].float().contiguous(), group=self.group)
        self.params_flat[lo:hi].copy_(torch.cat(parts))

    def broadcast_params(self, src_rank_in_group=0):
        if self.world > 1:
            buf = self.params_flat.float()
            dist.broadcast(buf, src=src_rank_in_group, group=self.group)
            self.params_flat.copy_(buf)

    def _launch_update(self, entry, lr, grad_scale, step):
        from oracle import mt_oracle as O

        step = int(step.reshape(-1)[0]) if isinstance(step, torch.Tensor) else int(step)
        O.multi_tensor_fused_adam(8192, self.noop, entry["lists"], *[h.tolist() for h in entry["hp"]], _scalar(lr),
                                  _scalar(grad_scale), step, 1 if self.adam_w_mode else 0, honour_noop=True)

    def local_grad_sq(self, params=None, flag=None):
        gl = self.owned_grad_fragments(params)
        sq = sum(((g.double() ** 2).sum() for g in gl), torch.zeros((), dtype=torch.float64)).float().reshape(1)
        if not bool(torch.isfinite(sq)):
            (self.noop if flag is None else flag).fill_(1)
        return sq

    def cast_out(self, src, dst):
        if int(self.noop) == 0:
            for s, d in zip(src, dst):
                d.copy_(s)


class CpuDistributedFusedAdam(DistributedFusedAdam):
    _engine_cls = CpuShardedAdam
