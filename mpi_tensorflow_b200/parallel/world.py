"""Process-group bootstrap: the ``MPI_Init`` + ``COMM_WORLD`` of this framework.

``from mpi4py import MPI`` initialises MPI at import (``mpipy.py:5``) and
``mpipy.py:208-210`` reads rank and size.  Here a launcher (``parallel/launch.py``
or ``torchrun``) exports ``RANK`` / ``WORLD_SIZE`` / ``LOCAL_RANK`` /
``MASTER_ADDR`` / ``MASTER_PORT`` and :func:`init` turns them into a
:class:`~.comm.Communicator`.  Unlike the reference (no device pinning at
``mpipy.py:69-71``: every rank lands on GPU 0) each rank is pinned to
``cuda:LOCAL_RANK``.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch

from .comm import Communicator, SelfComm, TorchComm

_WORLD: Optional[Communicator] = None


@dataclass(frozen=True)
class RankInfo:
    rank: int
    world: int
    local_rank: int
    master_addr: str
    master_port: int

    @staticmethod
    def from_env() -> "RankInfo":
        return RankInfo(
            rank=int(os.environ.get("RANK", "0")),
            world=int(os.environ.get("WORLD_SIZE", "1")),
            local_rank=int(os.environ.get("LOCAL_RANK", os.environ.get("RANK", "0"))),
            master_addr=os.environ.get("MASTER_ADDR", "127.0.0.1"),
            master_port=int(os.environ.get("MASTER_PORT", "29500")),
        )


def pick_device(info: RankInfo, want: str = "auto") -> torch.device:
    if want == "cpu" or (want == "auto" and not torch.cuda.is_available()):
        return torch.device("cpu")
    if not torch.cuda.is_available():
        raise RuntimeError("device 'cuda' requested but no GPU is visible")
    dev = torch.device("cuda", info.local_rank % torch.cuda.device_count())
    torch.cuda.set_device(dev)
    return dev


def init(backend: str = "auto", device: str = "auto", timeout_s: float = 120.0) -> Communicator:
    """Create (once) and return the world communicator.

    backend: ``auto`` | ``self`` | ``gloo`` | ``shm`` | ``nccl`` | ``p2p``
      auto -> self (world 1), p2p (GPU, world>1), gloo (CPU, world>1)
    """
    global _WORLD
    if _WORLD is not None:
        return _WORLD
    info = RankInfo.from_env()
    dev = pick_device(info, device)
    if dev.type == "cpu" and info.world > 1 and "OMP_NUM_THREADS" not in os.environ:
        # N ranks on one host: split the cores instead of oversubscribing them N times
        torch.set_num_threads(max(1, (os.cpu_count() or 1) // info.world))
    if backend == "auto":
        backend = "self" if info.world == 1 else ("p2p" if dev.type == "cuda" else "gloo")
    if backend == "self":
        if info.world != 1:
            raise ValueError("backend 'self' needs WORLD_SIZE=1")
        comm: Communicator = SelfComm()
    elif backend == "gloo":
        comm = TorchComm("gloo", timeout_s)
    elif backend == "nccl":
        comm = TorchComm("nccl", timeout_s, device=dev)
    elif backend == "shm":
        from .shm import ShmComm
        comm = ShmComm(info.rank, info.world, timeout_s=timeout_s)
    elif backend == "p2p":
        from .p2p import P2PComm
        comm = P2PComm(info, dev, timeout_s=timeout_s)
    else:
        raise ValueError(f"unknown backend {backend!r}")
    comm.device = dev  # type: ignore[attr-defined]
    _WORLD = comm
    return comm


def world() -> Communicator:
    if _WORLD is None:
        return init()
    return _WORLD


def finalize() -> None:
    global _WORLD
    if _WORLD is not None:
        _WORLD.close()
        _WORLD = None
