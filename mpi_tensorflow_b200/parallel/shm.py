"""Native shared-memory communicator for CPU ranks (``csrc/host/shm_comm.cpp``).

This is the MPI-runtime stand-in: ``mpipy.py:5`` gets its
world from ``MPI_Init``; here rank/size come from the launcher environment and
the transport is one POSIX shm segment.  Timeouts and the abort word turn a
dead peer into :class:`CommTimeout` / :class:`CommAborted`.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch

from ..ops import native_host as nh
from .comm import CommAborted, CommError, CommTimeout, Communicator


def _check(rc: int, what: str):
    if rc == nh.OK:
        return
    if rc == nh.ERR_TIMEOUT:
        raise CommTimeout(f"{what}: a peer did not arrive within the timeout")
    if rc == nh.ERR_ABORTED:
        raise CommAborted(f"{what}: another rank aborted the job")
    raise CommError(f"{what}: native error code {rc}")


class ShmComm(Communicator):
    backend = "shm"

    def __init__(self, rank: int, size: int, job_id: Optional[str] = None,
                 slot_bytes: int = 32 << 20, timeout_s: float = 60.0):
        self.rank, self.size = rank, size
        job = job_id or os.environ.get("MTB_JOB_ID") or os.environ.get("MASTER_PORT", "0")
        self._name = f"/mtb200_{os.getuid()}_{job}".encode()
        self._h = C.c_void_p()
        self._L = nh.lib()
        _check(self._L.mtb_shm_open(self._name, rank, size, slot_bytes, int(timeout_s * 1000),
                                    C.byref(self._h)), "shm_open")

    @staticmethod
    def _ptr(t: torch.Tensor) -> int:
        if t.device.type != "cpu" or not t.is_contiguous():
            raise ValueError("ShmComm needs contiguous CPU tensors")
        return t.data_ptr()

    def set_timeout(self, seconds: float):
        self._L.mtb_shm_set_timeout(self._h, int(seconds * 1000))

    def barrier(self):
        _check(self._L.mtb_shm_barrier(self._h), "barrier")

    def abort(self):
        if self._h:
            self._L.mtb_shm_abort(self._h)

    def bcast(self, t, root=0):
        _check(self._L.mtb_shm_bcast(self._h, self._ptr(t), t.numel() * t.element_size(), root), "bcast")

    def scatter(self, send, recv, root=0):
        sp = self._ptr(send) if (self.rank == root and send is not None) else None
        _check(self._L.mtb_shm_scatter(self._h, sp, self._ptr(recv),
                                       recv.numel() * recv.element_size(), root), "scatter")

    def gather(self, send, recv, root=0):
        rp = self._ptr(recv) if (self.rank == root and recv is not None) else None
        _check(self._L.mtb_shm_gather(self._h, self._ptr(send), rp,
                                      send.numel() * send.element_size(), root), "gather")

    def allgather(self, send, recv):
        _check(self._L.mtb_shm_allgather(self._h, self._ptr(send), self._ptr(recv),
                                         send.numel() * send.element_size()), "allgather")

    def allreduce(self, t, average=False):
        if t.dtype != torch.float32:
            # other dtypes: reduce in fp32 and cast back (exact for ints < 2^24)
            tmp = t.to(torch.float32).contiguous()
            self.allreduce(tmp, average)
            t.copy_(tmp.to(t.dtype))
            return
        _check(self._L.mtb_shm_allreduce_f32(self._h, self._ptr(t), t.numel(), int(average)), "allreduce")

    def heartbeat(self, rank: int) -> int:
        return int(self._L.mtb_shm_heartbeat(self._h, rank))

    def close(self):
        if self._h:
            self._L.mtb_shm_close(self._h)
            self._h = C.c_void_p()
