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
        self.group = None                     # LocalGroup when one process drives every rank of the world on one device
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
        if self.group is not None:
            return self.group.all_reduce_sum_(self.rank, t)
        self._lib.check(self._lib.lib().uuo_comm_all_reduce_sum(self.handle, self._lib.ptr(t), t.numel(),
                                                                  self._lib.stream_ptr(t.device)))
        return t

    def halo_exchange(self, x: torch.Tensor, n: int, prev: Optional[torch.Tensor], nxt: Optional[torch.Tensor],
                      col: int = 0) -> None:
        """x [F,D]; prev / nxt [n,ld]: columns col..col+D of their rows receive the neighbours' rows."""
        F, D = x.shape
        ref = prev if prev is not None else nxt
        ld = ref.shape[1] if ref is not None else D
        if self.group is not None:
            return self.group.halo_exchange(self.rank, x, n, prev, nxt, ld, col)
        self._lib.check(self._lib.lib().uuo_comm_halo_exchange(self.handle, self._lib.ptr(x), F, D, n,
                                                                 self._lib.ptr(prev), self._lib.ptr(nxt), ld, col,
                                                                 self._lib.stream_ptr(x.device)))

    def close(self):
        if self.handle:
            self._lib.lib().uuo_comm_destroy(self.handle)
            self.handle = None


class LocalGroup:
    """Every rank of a world driven by ONE process on ONE device (the single-GPU parity tests of the sharded closures).
    Each rank runs in its own Python thread (``run``) and issues its kernels on the same CUDA stream; an exchange is a
    host barrier, ONE cooperative launch that plays all ranks (uuo_comm_group_*: block r = rank r, same mailboxes and
    protocol as the per-rank kernels), and a second barrier.  Kernels of different launches never wait on one another."""

    def __init__(self, world: int, device, timeout_s: Optional[float] = 20.0):
        import threading
        self.world = int(world)
        self.device = torch.device(device)
        self.comms = [PeerComm(r, world, device, timeout_s) for r in range(world)]
        PeerComm.connect_local(self.comms)
        for c in self.comms:
            c.group = self
        self._barrier = threading.Barrier(world)
        self._slot = [None] * world

    def shard(self, rank: int) -> "Shard":
        return Shard(rank, self.world, None, self.comms[rank])

    def _collect(self, rank, item, launch):
        self._slot[rank] = item
        self._barrier.wait()
        if rank == 0:
            try:
                launch(list(self._slot))
            except BaseException:
                self._barrier.abort()
                raise
        self._barrier.wait()

    def all_reduce_sum_(self, rank: int, t: torch.Tensor) -> torch.Tensor:
        import ctypes as C
        lib = self.comms[0]._lib

        def launch(ts):
            n = ts[0].numel()
            if any(x.numel() != n for x in ts):
                raise ValueError("every rank must reduce the same number of floats")
            cs = (C.c_void_p * self.world)(*[c.handle for c in self.comms])
            xs = (C.c_void_p * self.world)(*[x.data_ptr() for x in ts])
            lib.check(lib.lib().uuo_comm_group_all_reduce_sum(cs, xs, n, self.world, lib.stream_ptr(self.device)))
        self._collect(rank, t, launch)
        return t

    def halo_exchange(self, rank: int, x, n, prev, nxt, ld, col) -> None:
        import ctypes as C
        lib = self.comms[0]._lib

        def launch(items):
            D = items[0][0].shape[1]
            ptr_of = lambda t: None if t is None else t.data_ptr()
            cs = (C.c_void_p * self.world)(*[c.handle for c in self.comms])
            xs = (C.c_void_p * self.world)(*[it[0].data_ptr() for it in items])
            Fs = (C.c_int * self.world)(*[it[0].shape[0] for it in items])
            ps = (C.c_void_p * self.world)(*[ptr_of(it[1]) for it in items])
            ns = (C.c_void_p * self.world)(*[ptr_of(it[2]) for it in items])
            lib.check(lib.lib().uuo_comm_group_halo_exchange(cs, xs, Fs, D, n, ps, ns, ld, col, self.world,
                                                             lib.stream_ptr(self.device)))
        self._collect(rank, (x, prev, nxt), launch)

    def run(self, fn):
        """fn(rank) on one thread per rank; returns the list of results (re-raises the first failure)."""
        import threading
        out, err = [None] * self.world, [None] * self.world

        def work(r):
            try:
                with torch.cuda.device(self.device):
                    out[r] = fn(r)
            except BaseException as e:      # a failed rank must not leave the others waiting at a barrier
                err[r] = e
                self._barrier.abort()
        ts = [threading.Thread(target=work, args=(r,)) for r in range(self.world)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        real = [e for e in err if e is not None and not isinstance(e, threading.BrokenBarrierError)]
        if real or any(err):
            raise (real[0] if real else [e for e in err if e is not None][0])
        return out

    def close(self):
        for c in self.comms:
            c.close()


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
