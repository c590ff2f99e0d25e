"""Rank-prefixed structured logging.

The reference prints a start banner with the rank (``mpipy.py:77``) and
``rank, step, test error`` every 50 steps (``mpipy.py:88``), flushing stdout
(``mpipy.py:90``); loss / lr / predictions are computed and thrown away
(``mpipy.py:89``).  Here every K steps one line carries step, loss, lr, errors,
img/s and sync time, on stdout (rank-prefixed) and optionally as JSONL.
"""
from __future__ import annotations

import json
import sys
import time
from typing import Optional


class RunLogger:
    def __init__(self, rank: int = 0, world: int = 1, jsonl_path: Optional[str] = None,
                 all_ranks: bool = False, stream=None):
        self.rank, self.world = rank, world
        self.enabled = all_ranks or rank == 0
        self.stream = stream or sys.stdout
        self._f = open(jsonl_path, "a") if (jsonl_path and self.enabled) else None
        self.t0 = time.time()

    def banner(self):
        # mpipy.py:77
        print(f"Process ID: {self.rank}  training session starts! (world={self.world})",
              file=self.stream, flush=True)

    def log(self, **rec):
        if not self.enabled:
            return
        rec = {"rank": self.rank, "t": round(time.time() - self.t0, 4), **rec}
        if self._f:
            self._f.write(json.dumps(rec) + "\n")
            self._f.flush()
        parts = []
        for k, v in rec.items():
            if k in ("rank", "t"):
                continue
            parts.append(f"{k}={v:.6g}" if isinstance(v, float) else f"{k}={v}")
        print(f"[{self.rank}] " + " ".join(parts), file=self.stream, flush=True)   # mpipy.py:88,90

    def close(self):
        if self._f:
            self._f.close()
            self._f = None
