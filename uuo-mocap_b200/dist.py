"""Frame-range sharding plumbing (torch.distributed: NCCL over NVLink on GPUs, gloo in the CPU tests).

The fitting path shards by contiguous frame ranges; frames are independent
except for (1) the shared betas row -> one small all-reduce per closure, (2) global ``mean``
normalisers -> all-reduced once, (3) the mean-distance correspondence matrix -> one [M,V] all-reduce per
stage, (4) the optional temporal terms -> a 1- or 2-frame halo exchanged with the two neighbours by
P2P send/recv.  Nothing here touches CUDA directly: it works on whatever device the tensors live on.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import torch
import torch.distributed as dist


def frame_range(F: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous range [start, stop) of rank ``rank``; sizes differ by at most one frame."""
    base, rem = divmod(F, world)
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


@dataclass
class Shard:
    rank: int = 0
    world: int = 1
    group: Optional[object] = None

    @classmethod
    def from_env(cls, group=None) -> "Shard":
        if dist.is_available() and dist.is_initialized():
            return cls(dist.get_rank(group), dist.get_world_size(group), group)
        return cls()

    def range(self, F: int) -> Tuple[int, int]:
        return frame_range(F, self.rank, self.world)

    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        """For host-side scalars (float64 CPU tensor): moved to the backend's device when needed."""
        if self.world <= 1:
            return t
        if dist.get_backend(self.group) == "nccl" and not t.is_cuda:
            d = t.cuda()
            dist.all_reduce(d, op=dist.ReduceOp.SUM, group=self.group)
            return d.cpu()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t


def exchange_halo(x: torch.Tensor, n: int, shard: Shard):
    """x [F,D] (this rank's rows).  Returns (prev, next): the last ``n`` rows of the left neighbour and the
    first ``n`` rows of the right neighbour ([n,D] each), or None at the ends of the sequence.
    Each boundary moves n*D floats in each direction (<= 2*207 floats for the pose second difference)."""
    if shard.world <= 1:
        return None, None
    if x.shape[0] < n:
        raise ValueError("a shard must hold at least %d frames for this halo" % n)
    r, w = shard.rank, shard.world
    prev = torch.empty((n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device) if r > 0 else None
    nxt = torch.empty((n,) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device) if r < w - 1 else None
    ops = []
    gr = lambda k: k if shard.group is None else dist.get_global_rank(shard.group, k)
    if r > 0:
        ops.append(dist.P2POp(dist.isend, x[:n].contiguous(), gr(r - 1), group=shard.group))
        ops.append(dist.P2POp(dist.irecv, prev, gr(r - 1), group=shard.group))
    if r < w - 1:
        ops.append(dist.P2POp(dist.isend, x[-n:].contiguous(), gr(r + 1), group=shard.group))
        ops.append(dist.P2POp(dist.irecv, nxt, gr(r + 1), group=shard.group))
    for req in dist.batch_isend_irecv(ops):
        req.wait()
    return prev, nxt
