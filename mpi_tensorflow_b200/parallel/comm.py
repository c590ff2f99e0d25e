"""MPI-style communicator API and its host-side backends.

The reference's whole communication surface is ``MPI.COMM_WORLD`` with
``Get_rank`` / ``Get_size`` (``mpipy.py:208-210``), six ``Scatter`` calls
(``mpipy.py:236-241``), four ``Gather`` calls per sync (``mpipy.py:121-127``) and a
commented-out ``Barrier`` (``mpipy.py:92-93``).  :class:`Communicator` reproduces
that surface (the capitalised mpi4py-style buffer methods are kept as aliases)
and adds what the reference *meant* to have: ``bcast`` and ``allreduce``
(SURVEY §5.5 "API surface to reproduce").

Backends
--------
``FakeWorld``      N logical ranks in one process (threads, lock-step) — unit tests
``TorchComm``      ``torch.distributed`` gloo (CPU) or nccl (GPU *baseline* only)
``ShmComm``        native shared-memory process group (``csrc/host/shm_comm.cpp``)
``P2PComm``        peer-mapped GPU memory + hand-written kernels (``parallel/p2p.py``)
"""
from __future__ import annotations

import os
import threading
from typing import List, Optional

import numpy as np
import torch


class CommError(RuntimeError):
    pass


class CommTimeout(CommError):
    """A peer failed to arrive at a collective within the timeout."""


class CommAborted(CommError):
    """Another rank raised the job-wide abort flag."""


def _as_tensor(x) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x
    if isinstance(x, np.ndarray):
        return torch.from_numpy(x)
    raise TypeError(f"expected torch.Tensor or numpy.ndarray, got {type(x)}")


class Communicator:
    """Abstract communicator.  All collectives are blocking and in-order; every
    rank must call them in the same sequence (as with MPI)."""

    rank: int
    size: int
    backend: str = "abstract"

    # -- mpi4py-flavoured aliases (mpipy.py:208-210, :121, :236) ----------------
    def Get_rank(self) -> int:
        return self.rank

    def Get_size(self) -> int:
        return self.size

    def Barrier(self) -> None:
        self.barrier()

    def Scatter(self, sendbuf, recvbuf, root: int = 0) -> None:
        self.scatter(None if sendbuf is None else _as_tensor(sendbuf), _as_tensor(recvbuf), root)

    def Gather(self, sendbuf, recvbuf, root: int = 0) -> None:
        self.gather(_as_tensor(sendbuf), None if recvbuf is None else _as_tensor(recvbuf), root)

    def Bcast(self, buf, root: int = 0) -> None:
        self.bcast(_as_tensor(buf), root)

    def Allreduce(self, buf, average: bool = False) -> None:
        self.allreduce(_as_tensor(buf), average)

    # -- to be provided by backends -------------------------------------------
    def barrier(self) -> None:
        raise NotImplementedError

    def bcast(self, t: torch.Tensor, root: int = 0) -> None:
        raise NotImplementedError

    def scatter(self, send: Optional[torch.Tensor], recv: torch.Tensor, root: int = 0) -> None:
        """``send`` (root only) holds ``size`` equal contiguous chunks; chunk r lands in rank r's ``recv``."""
        raise NotImplementedError

    def gather(self, send: torch.Tensor, recv: Optional[torch.Tensor], root: int = 0) -> None:
        raise NotImplementedError

    def allgather(self, send: torch.Tensor, recv: torch.Tensor) -> None:
        raise NotImplementedError

    def allreduce(self, t: torch.Tensor, average: bool = False) -> None:
        """In-place sum (or mean) across ranks; result bit-identical on every rank."""
        raise NotImplementedError

    def abort(self) -> None:
        """Fail-fast: make every rank's pending/next collective raise."""

    def close(self) -> None:
        pass

    # -- conveniences -----------------------------------------------------------
    def bcast_object(self, obj, root: int = 0):
        import pickle
        if self.rank == root:
            payload = np.frombuffer(pickle.dumps(obj), dtype=np.uint8).copy()
            n = torch.tensor([payload.size], dtype=torch.int64)
        else:
            n = torch.zeros(1, dtype=torch.int64)
        self.bcast(n, root)
        buf = torch.from_numpy(payload) if self.rank == root else torch.zeros(int(n), dtype=torch.uint8)
        self.bcast(buf, root)
        return obj if self.rank == root else pickle.loads(buf.numpy().tobytes())

    def allgather_object(self, obj) -> list:
        import pickle
        payload = np.frombuffer(pickle.dumps(obj), dtype=np.uint8).copy()
        sizes = torch.zeros(self.size, dtype=torch.int64)
        self.allgather(torch.tensor([payload.size], dtype=torch.int64), sizes)
        m = int(sizes.max())
        send = torch.zeros(m, dtype=torch.uint8)
        send[:payload.size] = torch.from_numpy(payload)
        recv = torch.zeros(self.size * m, dtype=torch.uint8)
        self.allgather(send, recv)
        return [pickle.loads(recv[r * m:r * m + int(sizes[r])].numpy().tobytes()) for r in range(self.size)]


# ---------------------------------------------------------------------------
# single-process world
# ---------------------------------------------------------------------------
class SelfComm(Communicator):
    """World of one.  Every collective is the identity."""
    backend = "self"

    def __init__(self):
        self.rank, self.size = 0, 1

    def barrier(self):
        pass

    def bcast(self, t, root=0):
        pass

    def scatter(self, send, recv, root=0):
        recv.reshape(-1).copy_(send.reshape(-1)[:recv.numel()])

    def gather(self, send, recv, root=0):
        recv.reshape(-1)[:send.numel()].copy_(send.reshape(-1))

    def allgather(self, send, recv):
        recv.reshape(-1).copy_(send.reshape(-1))

    def allreduce(self, t, average=False):
        pass


# ---------------------------------------------------------------------------
# fake world: N logical ranks as threads of one process
# ---------------------------------------------------------------------------
class _FakeState:
    def __init__(self, n: int, timeout: float):
        self.n = n
        self.timeout = timeout
        self.barrier = threading.Barrier(n)
        self.slots: List[Optional[torch.Tensor]] = [None] * n
        self.aborted = False


class FakeComm(Communicator):
    backend = "fake"

    def __init__(self, state: _FakeState, rank: int):
        self._s = state
        self.rank, self.size = rank, state.n

    def _wait(self):
        if self._s.aborted:
            raise CommAborted("fake world aborted")
        try:
            self._s.barrier.wait(self._s.timeout)
        except threading.BrokenBarrierError:
            if self._s.aborted:
                raise CommAborted("fake world aborted") from None
            raise CommTimeout(f"rank {self.rank}: peer missing at barrier") from None

    def barrier(self):
        self._wait()

    def abort(self):
        self._s.aborted = True
        self._s.barrier.abort()

    def bcast(self, t, root=0):
        if self.rank == root:
            self._s.slots[root] = t
        self._wait()
        if self.rank != root:
            t.copy_(self._s.slots[root].reshape(t.shape))
        self._wait()

    def scatter(self, send, recv, root=0):
        if self.rank == root:
            self._s.slots[root] = send
        self._wait()
        n = recv.numel()
        recv.reshape(-1).copy_(self._s.slots[root].reshape(-1)[self.rank * n:(self.rank + 1) * n])
        self._wait()

    def gather(self, send, recv, root=0):
        self._s.slots[self.rank] = send
        self._wait()
        if self.rank == root:
            n = send.numel()
            for r in range(self.size):
                recv.reshape(-1)[r * n:(r + 1) * n].copy_(self._s.slots[r].reshape(-1))
        self._wait()

    def allgather(self, send, recv):
        self._s.slots[self.rank] = send
        self._wait()
        n = send.numel()
        for r in range(self.size):
            recv.reshape(-1)[r * n:(r + 1) * n].copy_(self._s.slots[r].reshape(-1))
        self._wait()

    def allreduce(self, t, average=False):
        self._s.slots[self.rank] = t.clone()
        self._wait()
        acc = torch.zeros_like(t)
        for r in range(self.size):            # fixed rank order => identical bits on all ranks
            acc += self._s.slots[r]
        if average:
            acc *= 1.0 / self.size
        self._wait()
        t.copy_(acc)


class FakeWorld:
    """``FakeWorld(n).run(fn)`` runs ``fn(comm)`` on n threads, one logical rank
    each, and returns the per-rank results (re-raising the first failure)."""

    def __init__(self, n: int, timeout: float = 60.0):
        self.state = _FakeState(n, timeout)
        self.comms = [FakeComm(self.state, r) for r in range(n)]

    def run(self, fn, *args, **kw):
        results = [None] * self.state.n
        errors: List[Optional[BaseException]] = [None] * self.state.n

        def body(r):
            try:
                results[r] = fn(self.comms[r], *args, **kw)
            except BaseException as e:  # noqa: BLE001
                errors[r] = e
                self.comms[r].abort()

        ts = [threading.Thread(target=body, args=(r,), daemon=True) for r in range(self.state.n)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        primary = [e for e in errors if e is not None and not isinstance(e, CommAborted)]
        if primary:
            raise primary[0]
        if any(errors):
            raise [e for e in errors if e is not None][0]
        return results


# ---------------------------------------------------------------------------
# torch.distributed (gloo on CPU, nccl on GPU as the *baseline* transport)
# ---------------------------------------------------------------------------
class TorchComm(Communicator):
    def __init__(self, backend: str = "gloo", timeout_s: float = 120.0, device: Optional[torch.device] = None):
        import datetime
        import torch.distributed as dist
        self._dist = dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            kw = {}
            if backend == "nccl" and device is not None:
                kw["device_id"] = device
            dist.init_process_group(backend=backend, timeout=datetime.timedelta(seconds=timeout_s), **kw)
        self.backend = dist.get_backend()
        self.rank, self.size = dist.get_rank(), dist.get_world_size()
        self.device = device

    def _wrap(self, fn, *a, **k):
        try:
            return fn(*a, **k)
        except RuntimeError as e:
            msg = str(e)
            if "imeout" in msg or "timed out" in msg:
                raise CommTimeout(msg) from e
            if "Connection closed" in msg or "reset by peer" in msg:
                raise CommAborted(msg) from e
            raise

    def barrier(self):
        self._wrap(self._dist.barrier)

    def bcast(self, t, root=0):
        self._wrap(self._dist.broadcast, t, src=root)

    def scatter(self, send, recv, root=0):
        chunks = None
        if self.rank == root:
            flat = send.reshape(self.size, -1)
            chunks = [flat[r].reshape(recv.shape).contiguous() for r in range(self.size)]
        self._wrap(self._dist.scatter, recv, chunks, src=root)

    def gather(self, send, recv, root=0):
        outs = None
        if self.rank == root:
            outs = [torch.empty_like(send) for _ in range(self.size)]
        self._wrap(self._dist.gather, send.contiguous(), outs, dst=root)
        if self.rank == root:
            n = send.numel()
            for r in range(self.size):
                recv.reshape(-1)[r * n:(r + 1) * n].copy_(outs[r].reshape(-1))

    def allgather(self, send, recv):
        self._wrap(self._dist.all_gather_into_tensor, recv.reshape(-1), send.reshape(-1).contiguous())

    def allreduce(self, t, average=False):
        self._wrap(self._dist.all_reduce, t, op=self._dist.ReduceOp.SUM)
        if average:
            t.mul_(1.0 / self.size)

    def close(self):
        if self._dist.is_initialized():
            self._dist.destroy_process_group()
