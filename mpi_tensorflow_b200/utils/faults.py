"""Fault injection and watchdog.

The reference has no failure handling: a dead rank leaves its peers blocked in
``Gather`` forever (``mpipy.py:121-127``) and the only ``except`` names an
undefined class (``mpipy.py:197``).  This module provides the hook the tests use
to kill or stall a chosen rank at a chosen step (SURVEY §5 "fault-injection
hook") and a host watchdog that aborts the job when a step exceeds its budget.

``MTB_FAULT="rank=1,step=3,kind=exit"``   kinds: exit | sleep | raise
"""
from __future__ import annotations

import os
import threading
import time
from dataclasses import dataclass
from typing import Callable, Optional


@dataclass
class FaultSpec:
    rank: int = -1
    step: int = -1
    kind: str = "exit"
    seconds: float = 3600.0

    @staticmethod
    def from_env(var: str = "MTB_FAULT") -> Optional["FaultSpec"]:
        s = os.environ.get(var)
        if not s:
            return None
        kv = dict(item.split("=", 1) for item in s.split(",") if "=" in item)
        return FaultSpec(int(kv.get("rank", -1)), int(kv.get("step", -1)), kv.get("kind", "exit"),
                         float(kv.get("seconds", 3600.0)))


class InjectedFault(RuntimeError):
    pass


def maybe_inject(spec: Optional[FaultSpec], rank: int, step: int) -> None:
    if spec is None or spec.rank != rank or spec.step != step:
        return
    if spec.kind == "exit":
        os._exit(17)
    if spec.kind == "sleep":
        time.sleep(spec.seconds)
        return
    if spec.kind == "raise":
        raise InjectedFault(f"injected fault on rank {rank} at step {step}")
    raise ValueError(f"unknown fault kind {spec.kind!r}")


class Watchdog:
    """Calls ``on_timeout`` if :meth:`kick` is not called for ``budget_s`` seconds."""

    def __init__(self, budget_s: float, on_timeout: Callable[[], None]):
        self.budget_s, self.on_timeout = budget_s, on_timeout
        self._last = time.monotonic()
        self._stop = threading.Event()
        self._fired = False
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def kick(self):
        self._last = time.monotonic()

    @property
    def fired(self) -> bool:
        return self._fired

    def _run(self):
        while not self._stop.wait(min(0.25, self.budget_s / 4)):
            if time.monotonic() - self._last > self.budget_s:
                self._fired = True
                try:
                    self.on_timeout()
                finally:
                    return

    def stop(self):
        self._stop.set()
