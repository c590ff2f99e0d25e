"""CPU engine: the oracle wrapped behind the engine interface.

Used for the 2-rank CPU plumbing config of ``BASELINE.json`` (SURVEY §7.3 step 2:
"using the oracle as the device") and as the timed CPU baseline B-cpu.  The GPU
engine (``engine_native.py``) exposes the same attributes and methods.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch

from .config import TrainConfig
from .models.oracle import Oracle, dropout_keep_mask
from .models.spec import ModelSpec, flat_size


class CpuEngine:
    device = torch.device("cpu")
    name = "cpu-oracle"

    def __init__(self, spec: ModelSpec, cfg: TrainConfig):
        self.spec, self.cfg = spec, cfg
        # weight decay is applied in apply() on the (possibly averaged) gradient, exactly
        # where the fused device kernel applies it, so the oracle itself runs with wd=0.
        self.oracle = Oracle(spec, seed=cfg.seed, weight_decay=0.0)
        n = flat_size(spec)
        self.flat_params = self.oracle.flatten(self.oracle.params, n)
        self.flat_grads = torch.zeros(n)
        self.slots: List[torch.Tensor] = []
        if cfg.optimizer in ("momentum", "adam"):
            self.slots.append(torch.zeros(n))
        if cfg.optimizer == "adam":
            self.slots.append(torch.zeros(n))
        self.opt_step = 0
        self._decay_mask = torch.zeros(n)
        for p in spec.params:
            if p.decay:
                self._decay_mask[p.offset:p.offset + p.numel] = 1.0
        self.last_aux = {}

    # -- parameters live in flat_params; the oracle sees views -------------------
    def _sync_oracle(self):
        self.oracle.params = self.oracle.unflatten(self.flat_params)

    def compute_grads(self, x, y, step: int) -> float:
        x = torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x).float()
        y = torch.as_tensor(np.asarray(y) if not isinstance(y, torch.Tensor) else y).long()
        self._sync_oracle()
        mask = dropout_keep_mask(self.spec, x.shape[0], self.cfg.seed, step) if self.cfg.dropout else None
        loss, g, aux = self.oracle.loss_and_grads(x, y, mask)
        self.flat_grads.copy_(self.oracle.flatten(g, self.flat_grads.numel()))
        # report the reference's loss (CE + L2 term, mpipy.py:55-58)
        if self.cfg.weight_decay:
            l2 = 0.5 * float(((self.flat_params * self._decay_mask) ** 2).sum())
            loss += self.cfg.weight_decay * l2
        self.last_aux = aux
        return loss

    def apply(self, lr: float) -> None:
        c = self.cfg
        self.opt_step += 1
        g = self.flat_grads
        if c.weight_decay:
            g = g + c.weight_decay * self._decay_mask * self.flat_params        # mpipy.py:57-58
        w = self.flat_params
        if c.optimizer == "sgd":
            w -= lr * g
        elif c.optimizer == "momentum":                                         # mpipy.py:65
            m = self.slots[0]
            m.mul_(c.momentum).add_(g)
            w -= lr * m
        elif c.optimizer == "adam":
            m, v = self.slots
            m.mul_(c.beta1).add_(g, alpha=1 - c.beta1)
            v.mul_(c.beta2).addcmul_(g, g, value=1 - c.beta2)
            t = self.opt_step
            lr_t = lr * (1 - c.beta2 ** t) ** 0.5 / (1 - c.beta1 ** t)
            w -= lr_t * m / (v.sqrt() + c.adam_eps)
        else:
            raise ValueError(c.optimizer)

    def predict(self, x, dropout_step: Optional[int] = None) -> torch.Tensor:
        x = torch.as_tensor(np.asarray(x) if not isinstance(x, torch.Tensor) else x).float()
        self._sync_oracle()
        mask = None
        if dropout_step is not None:                    # reference leaves dropout on in eval
            mask = dropout_keep_mask(self.spec, x.shape[0], self.cfg.seed, dropout_step)
        return torch.softmax(self.oracle.forward(x, mask), dim=1)

    def synchronize(self):
        pass

    # -- checkpoint views ---------------------------------------------------------
    def get_params(self):
        return {p.key: self.flat_params[p.offset:p.offset + p.numel].reshape(p.shape).numpy().copy()
                for p in self.spec.params}

    def get_slots(self):
        names = ["m", "v"]
        return {p.key: {names[i]: s[p.offset:p.offset + p.numel].reshape(p.shape).numpy().copy()
                        for i, s in enumerate(self.slots)} for p in self.spec.params}

    def set_state(self, params, slots, step: int):
        names = ["m", "v"]
        for p in self.spec.params:
            self.flat_params[p.offset:p.offset + p.numel] = torch.from_numpy(
                np.asarray(params[p.key], dtype=np.float32).reshape(-1))
            for i, s in enumerate(self.slots):
                if slots and p.key in slots and names[i] in slots[p.key]:
                    s[p.offset:p.offset + p.numel] = torch.from_numpy(
                        np.asarray(slots[p.key][names[i]], dtype=np.float32).reshape(-1))
        self.opt_step = step
