"""Frame-range sharding plumbing (torch.distributed: NCCL over NVLink on GPUs, gloo in the CPU tests).

The fitting path shards by contiguous frame ranges; frames are independent
except for (1) the shared betas row -> one small all-reduce per closure, (2) global ``mean``
normalisers -> all-reduced once, (3) the mean-distance correspondence matrix -> one [M,V] all-reduce per
stage, (4) the optional temporal terms -> a 1- or 2-frame halo exchanged with the two neighbours by
P2P send/recv.  Two transports: ``torch.distributed`` (NCCL on GPUs, gloo in the CPU tests) for set-up and
the once-per-stage sums, and -- on CUDA shards -- ``PeerComm`` (csrc/peer.cu): the per-iteration sum and halo
written straight into the neighbour's HBM over NVLink by small kernels, so closure + Adam can be replayed from
a CUDA graph on every rank.
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


class PeerComm:
    """One rank's mailbox of the peer-memory transport (uuo_comm_*).  ``connect()`` trades the CUDA IPC handles
    through torch.distributed; ``connect_local`` links communicators created in one process (tests)."""

    def __init__(self, rank: int, world: int, device, timeout_s: Optional[float] = None):
        import ctypes as C
        from . import _lib
        self._lib = _lib
        self.rank, self.world = int(rank), int(world)
        self.device = torch.device(device)
        h = C.c_void_p()
        _lib.check(_lib.lib().uuo_comm_create(self.rank, self.world, self.device.index or 0, C.byref(h)))
        self.handle = h
        if timeout_s is not None:
            _lib.check(_lib.lib().uuo_comm_set_timeout(h, float(timeout_s)))

    def connect(self, group=None) -> "PeerComm":
        import ctypes as C
        L = self._lib.lib()
        nb = L.uuo_comm_handle_bytes()
        buf = (C.c_ubyte * nb)()
        self._lib.check(L.uuo_comm_export(self.handle, buf))
        mine = torch.tensor(list(buf), dtype=torch.uint8, device=self.device)
        every = [torch.empty_like(mine) for _ in range(self.world)]
        dist.all_gather(every, mine, group=group)
        flat = torch.cat(every).cpu().numpy().tobytes()
        self._lib.check(L.uuo_comm_connect(self.handle, flat))
        dist.barrier(group=group)
        return self

    @staticmethod
    def connect_local(comms) -> None:
        import ctypes as C
        arr = (C.c_void_p * len(comms))(*[c.handle for c in comms])
        for c in comms:
            c._lib.check(c._lib.lib().uuo_comm_connect_local(c.handle, arr))

    def all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise self._lib.UuoError("PeerComm.all_reduce_sum_ needs a contiguous fp32 CUDA tensor")
        self._lib.check(self._lib.lib().uuo_comm_all_reduce_sum(self.handle, self._lib.ptr(t), t.numel(),
                                                                  self._lib.stream_ptr(t.device)))
        return t

    def halo_exchange(self, x: torch.Tensor, n: int, prev: Optional[torch.Tensor], nxt: Optional[torch.Tensor],
                      col: int = 0) -> None:
        """x [F,D]; prev / nxt [n,ld]: columns col..col+D of their rows receive the neighbours' rows."""
        F, D = x.shape
        ref = prev if prev is not None else nxt
        ld = ref.shape[1] if ref is not None else D
        self._lib.check(self._lib.lib().uuo_comm_halo_exchange(self.handle, self._lib.ptr(x), F, D, n,
                                                                 self._lib.ptr(prev), self._lib.ptr(nxt), ld, col,
                                                                 self._lib.stream_ptr(x.device)))

    def close(self):
        if self.handle:
            self._lib.lib().uuo_comm_destroy(self.handle)
            self.handle = None


@dataclass
class Shard:
    rank: int = 0
    world: int = 1
    group: Optional[object] = None
    comm: Optional[PeerComm] = None          # peer-memory transport of the per-iteration exchange (CUDA shards)

    @classmethod
    def from_env(cls, group=None, peer_device=None) -> "Shard":
        """``peer_device``: also set up the peer-memory transport on that CUDA device (one process per GPU)."""
        if dist.is_available() and dist.is_initialized():
            s = cls(dist.get_rank(group), dist.get_world_size(group), group)
            if peer_device is not None and s.world > 1:
                s.comm = PeerComm(s.rank, s.world, peer_device).connect(group)
            return s
        return cls()

    @property
    def graph_safe(self) -> bool:
        """True when the per-iteration exchange holds no host-issued collective (one rank, or peer memory)."""
        return self.world <= 1 or self.comm is not None

    def iter_all_reduce_sum_(self, t: torch.Tensor) -> torch.Tensor:
        """The per-iteration small sum (<= 1024 floats): peer memory when set up, else torch.distributed."""
        if self.world > 1 and self.comm is not None and t.is_cuda:
            return self.comm.all_reduce_sum_(t)
        return self.all_reduce_sum_(t)

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


def halo_buffers(n: int, ld: int, shard: Shard, device, dtype=torch.float32):
    """(prev, next) receive buffers [n,ld] of a shard; None at the ends of the sequence."""
    if shard.world <= 1:
        return None, None
    prev = torch.zeros((n, ld), dtype=dtype, device=device) if shard.rank > 0 else None
    nxt = torch.zeros((n, ld), dtype=dtype, device=device) if shard.rank < shard.world - 1 else None
    return prev, nxt


def exchange_halo_into(x: torch.Tensor, n: int, shard: Shard, prev, nxt, col: int = 0) -> None:
    """Fill columns col..col+D of the persistent buffers ``prev`` / ``nxt`` (from ``halo_buffers``) with the
    neighbours' boundary rows of x [F,D]: one peer-memory kernel when the shard has a PeerComm, else P2P send/recv."""
    if shard.world <= 1:
        return
    if shard.comm is not None and x.is_cuda:
        shard.comm.halo_exchange(x, n, prev, nxt, col)
        return
    D = x.shape[1]
    p, q = exchange_halo(x, n, shard)
    if p is not None:
        prev[:, col:col + D].copy_(p)
    if q is not None:
        nxt[:, col:col + D].copy_(q)


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
