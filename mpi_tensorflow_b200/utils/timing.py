"""Phase timers and NVTX ranges.

The reference's only timer is commented out (``mpipy.py:78``).  GPU phases are
timed with CUDA events on the launching stream (never wall clock); multi-rank
numbers are reduced with max over ranks by the caller (``bench.py``).
"""
from __future__ import annotations

import contextlib
import time
from collections import defaultdict
from typing import Dict, List

import torch


class PhaseTimer:
    """``enabled=False`` (the default of the training loop) makes :meth:`phase` free: no events are created and
    nothing accumulates, so a long run does not pay for timing nobody reads (``TrainConfig.profile_phases``)."""

    def __init__(self, cuda: bool, enabled: bool = True):
        self.enabled = enabled
        self.cuda = cuda and torch.cuda.is_available()
        self._pending: Dict[str, List] = defaultdict(list)
        self.totals_ms: Dict[str, float] = defaultdict(float)
        self.counts: Dict[str, int] = defaultdict(int)

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        if self.cuda:
            torch.cuda.nvtx.range_push(name)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            try:
                yield
            finally:
                b.record()
                torch.cuda.nvtx.range_pop()
                self._pending[name].append((a, b))
        else:
            t = time.perf_counter()
            try:
                yield
            finally:
                self.totals_ms[name] += (time.perf_counter() - t) * 1e3
                self.counts[name] += 1

    def flush(self):
        if self.cuda:
            torch.cuda.synchronize()
            for name, evs in self._pending.items():
                for a, b in evs:
                    self.totals_ms[name] += a.elapsed_time(b)
                    self.counts[name] += 1
            self._pending.clear()

    def summary(self) -> Dict[str, float]:
        self.flush()
        return {k: self.totals_ms[k] / max(1, self.counts[k]) for k in self.totals_ms}
