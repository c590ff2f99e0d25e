"""Data-parallel synchronisation strategies (the reference's only parallelism).

``Cnn.bcast_parameters`` (``mpipy.py:95-153``) is: every 50 steps (``:87``) Gather
the four *weight* tensors to rank 0 (``:121-127``), ``np.mean`` over the rank axis
(``:130-136``) and assign the mean to root's variables only (``:129``, ``:138-152``)
— biases and momentum are never synchronised and non-root ranks keep their own
weights.  SURVEY §7.1 reconciles that with the per-step gradient allreduce the
north-star asks for:

* :class:`GradAllreduce` — primary mode; gradients are averaged every step so
  all replicas stay bit-identical.  On GPUs with the p2p backend the engine does
  this inside the fused kernel and this class is bypassed.
* :class:`WeightAverage` — periodic parameter averaging; by default every rank
  receives the mean, biases included; ``root_only=True, include_bias=False``
  reproduces the reference literally, through the same Gather-to-root path.

Both work on the flat fp32 buffers of ``models.spec`` through any
:class:`~.comm.Communicator`.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import torch

from ..models.spec import ModelSpec
from .comm import Communicator


def _ranges(spec: ModelSpec, include_bias: bool) -> List[Tuple[int, int]]:
    return [(p.offset, p.numel) for p in spec.params if include_bias or not p.is_bias]


class GradAllreduce:
    name = "grad_allreduce"

    def __init__(self, comm: Communicator, spec: ModelSpec):
        self.comm, self.spec = comm, spec

    def reduce_grads(self, flat_grads: torch.Tensor) -> None:
        """In place: g <- mean over ranks."""
        if self.comm.size > 1:
            self.comm.allreduce(flat_grads, average=True)

    def after_step(self, step: int, flat_params: torch.Tensor,
                   flat_slots: Optional[List[torch.Tensor]] = None, engine=None) -> bool:
        return False


class WeightAverage:
    name = "weight_avg"

    def __init__(self, comm: Communicator, spec: ModelSpec, period: int = 50, include_bias: bool = True,
                 include_momentum: bool = False, root_only: bool = False):
        self.comm, self.spec = comm, spec
        self.period, self.include_bias = period, include_bias
        self.include_momentum, self.root_only = include_momentum, root_only

    def reduce_grads(self, flat_grads: torch.Tensor) -> None:
        pass                                   # local SGD between syncs

    def due(self, step: int) -> bool:
        return step > 0 and step % self.period == 0          # mpipy.py:87

    def after_step(self, step: int, flat_params: torch.Tensor,
                   flat_slots: Optional[List[torch.Tensor]] = None, engine=None) -> bool:
        """Returns True when ``flat_params`` was rewritten by this call from OUTSIDE the engine (the caller then
        refreshes derived copies).  With a GPU engine on the peer-memory backend the averaging is one fused kernel
        over NVLink that also rewrites the bf16 shadow (``NativeEngine.average_params``): nothing left to refresh."""
        if not self.due(step) or self.comm.size == 1:
            return False
        if engine is not None and hasattr(engine, "average_params") and getattr(self.comm, "backend", "") == "p2p":
            engine.average_params(_ranges(self.spec, self.include_bias), root_only=self.root_only,
                                  include_momentum=self.include_momentum)
            return False
        bufs = [flat_params] + (list(flat_slots or []) if self.include_momentum else [])
        for buf in bufs:
            if self.root_only:
                self._gather_mean_root(buf)
            elif self.include_bias:
                self.comm.allreduce(buf, average=True)
            else:
                for off, n in _ranges(self.spec, False):
                    seg = buf[off:off + n].contiguous()
                    self.comm.allreduce(seg, average=True)
                    buf[off:off + n].copy_(seg)
        return True

    def _gather_mean_root(self, buf: torch.Tensor) -> None:
        """The literal ``mpipy.py:104-152`` data path: Gather -> mean -> assign on root."""
        c = self.comm
        for off, n in _ranges(self.spec, self.include_bias):
            send = buf[off:off + n].contiguous()
            recv = torch.zeros(c.size * n, dtype=buf.dtype, device=buf.device) if c.rank == 0 else None
            c.gather(send, recv, root=0)                     # mpipy.py:121-127
            if c.rank == 0:                                  # mpipy.py:129
                buf[off:off + n].copy_(recv.reshape(c.size, n).mean(dim=0))   # mpipy.py:130-152


class NoSync:
    name = "none"

    def reduce_grads(self, flat_grads):
        pass

    def after_step(self, step, flat_params, flat_slots=None, engine=None):
        return False


def make_sync(cfg, comm: Communicator, spec: ModelSpec):
    if cfg.sync == "grad_allreduce":
        return GradAllreduce(comm, spec)
    if cfg.sync == "weight_avg":
        return WeightAverage(comm, spec, cfg.sync_period, cfg.sync_include_bias,
                             cfg.sync_include_momentum, cfg.sync_root_only)
    if cfg.sync == "none":
        return NoSync()
    raise ValueError(f"unknown sync mode {cfg.sync!r}")


def check_replicas_equal(comm: Communicator, flat: torch.Tensor) -> bool:
    """True iff every rank holds bit-identical ``flat`` (start-of-run check that
    replaces the reference's reliance on equal seeds, ``mpipy.py:38-52``)."""
    if comm.size == 1:
        return True
    ref = flat.clone()
    comm.bcast(ref, root=0)
    same = torch.tensor([1.0 if torch.equal(ref, flat) else 0.0], device=flat.device)
    comm.allreduce(same, average=False)
    return int(same.item()) == comm.size
