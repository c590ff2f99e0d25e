"""CPU oracle: a plain fp32 restatement of every op the reference instantiates.

This file is the correctness reference for every CUDA kernel in ``csrc/`` and
the timed CPU baseline (SURVEY §4.2 "Unit — CPU reference model", §6.2 B-cpu).
Forward, backward and the optimizers are written out by hand (im2col + matmul,
explicit pool routing, explicit softmax-CE gradient) — no autograd — so that
each device kernel has a formula to be compared with.  Autograd is used only
by the tests as a second opinion.

Op map:
  O1/O4  conv 5x5 SAME s1 NHWC x HWIO        ``mpipy.py:156,159``
  O2/O5  bias + ReLU                         ``mpipy.py:157,160``
  O3/O6  max-pool 2x2 s2                     ``mpipy.py:158,161``
  O7     flatten (h, w, c)                   ``mpipy.py:162-164``
  O8     FC1 + bias + ReLU                   ``mpipy.py:165``
  O9     dropout(keep)                       ``mpipy.py:166``
  O10    FC2 + bias                          ``mpipy.py:167``
  O11    sparse softmax CE, mean             ``mpipy.py:55-56``
  O12    5e-4 * sum l2_loss(FC params)       ``mpipy.py:57-58``
  O13    softmax heads                       ``mpipy.py:67-68``
  O14    staircase exponential LR decay      ``mpipy.py:60-64``
  O15    momentum apply                      ``mpipy.py:65-66``
"""
from __future__ import annotations

import math
from typing import Dict, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from .spec import KSIZE, ModelSpec
from ..utils import philox

Params = Dict[str, torch.Tensor]


def bf16_round(x: torch.Tensor) -> torch.Tensor:
    return x.to(torch.bfloat16).to(torch.float32)


def init_params(spec: ModelSpec, seed: int = 1) -> Params:
    """Seeded init identical on every rank (``mpipy.py:38-53``: all ranks use
    seed=1 so no initial broadcast is needed).  Stream id = parameter index, so
    each tensor draws an independent Philox stream."""
    out: Params = {}
    for i, p in enumerate(spec.params):
        if p.init == "trunc_normal":
            v = philox.truncated_normal(p.numel, p.init_value, seed, i)
        elif p.init == "zeros":
            v = np.zeros(p.numel, dtype=np.float32)
        else:
            v = np.full(p.numel, p.init_value, dtype=np.float32)
        out[p.key] = torch.from_numpy(v.reshape(p.shape).copy())
    return out


# --------------------------------------------------------------------------
# primitive ops, NHWC / HWIO, fp32
# --------------------------------------------------------------------------
def _im2col(x: torch.Tensor) -> torch.Tensor:
    """[B,H,W,C] -> [B*H*W, 25*C] with column order (kh, kw, c) — the row order
    of an HWIO filter reshaped to [25*C, O]."""
    B, H, W, C = x.shape
    pad = KSIZE // 2
    xp = F.pad(x, (0, 0, pad, pad, pad, pad))
    cols = []
    for kh in range(KSIZE):
        for kw in range(KSIZE):
            cols.append(xp[:, kh:kh + H, kw:kw + W, :])
    return torch.stack(cols, dim=3).reshape(B * H * W, KSIZE * KSIZE * C)


def _col2im(cols: torch.Tensor, shape) -> torch.Tensor:
    B, H, W, C = shape
    pad = KSIZE // 2
    xp = torch.zeros(B, H + 2 * pad, W + 2 * pad, C, dtype=cols.dtype)
    cols = cols.reshape(B, H, W, KSIZE * KSIZE, C)
    t = 0
    for kh in range(KSIZE):
        for kw in range(KSIZE):
            xp[:, kh:kh + H, kw:kw + W, :] += cols[:, :, :, t, :]
            t += 1
    return xp[:, pad:pad + H, pad:pad + W, :]


def conv5x5(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    B, H, W, C = x.shape
    O = w.shape[3]
    return (_im2col(x) @ w.reshape(-1, O)).reshape(B, H, W, O)


def conv5x5_wgrad(x: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    C, O = x.shape[3], dy.shape[3]
    return (_im2col(x).t() @ dy.reshape(-1, O)).reshape(KSIZE, KSIZE, C, O)


def conv5x5_dgrad(dy: torch.Tensor, w: torch.Tensor, xshape) -> torch.Tensor:
    O = w.shape[3]
    return _col2im(dy.reshape(-1, O) @ w.reshape(-1, O).t(), xshape)


def maxpool2x2(x: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Returns (pooled, argmax index 0..3 = 2*dy + dx; first max wins on ties)."""
    B, H, W, C = x.shape
    assert H % 2 == 0 and W % 2 == 0, "oracle pool assumes even extents (28 -> 14 -> 7)"
    v = x.reshape(B, H // 2, 2, W // 2, 2, C).permute(0, 1, 3, 5, 2, 4).reshape(B, H // 2, W // 2, C, 4)
    val, idx = v.max(dim=-1)
    # torch.max returns an arbitrary index on ties; make it the first one
    first = (v == val.unsqueeze(-1)).to(torch.uint8).argmax(dim=-1)
    return val, first.to(torch.uint8)


def maxpool2x2_bwd(dp: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    B, h, w, C = dp.shape
    g = torch.zeros(B, h, w, C, 4, dtype=dp.dtype)
    g.scatter_(-1, idx.long().unsqueeze(-1), dp.unsqueeze(-1))
    return g.reshape(B, h, w, C, 2, 2).permute(0, 1, 4, 2, 5, 3).reshape(B, 2 * h, 2 * w, C)


def softmax_ce(logits: torch.Tensor, labels: torch.Tensor):
    """mean CE, dlogits = (softmax - onehot)/B, probs, error count."""
    B = logits.shape[0]
    m = logits.max(dim=1, keepdim=True).values
    e = torch.exp(logits - m)
    s = e.sum(dim=1, keepdim=True)
    probs = e / s
    ll = (logits - m - torch.log(s)).gather(1, labels.long().unsqueeze(1)).squeeze(1)
    loss = -ll.mean()
    d = probs.clone()
    d[torch.arange(B), labels.long()] -= 1.0
    d /= B
    errors = int((probs.argmax(dim=1) != labels.long()).sum())
    return loss, d, probs, errors


def error_rate(pred, labels) -> float:
    """``100 - 100*mean(argmax(pred,1) == labels)`` percent — semantics of the
    missing ``convolutional.error_rate`` used at ``mpipy.py:86``."""
    pred = np.asarray(pred)
    labels = np.asarray(labels)
    return 100.0 - 100.0 * float(np.mean(np.argmax(pred, 1) == labels))


def learning_rate(step: int, batch: int, decay_rows: int, base: float = 0.01,
                  decay: float = 0.95, staircase: bool = True) -> float:
    """``tf.train.exponential_decay(0.01, iter*batch, shard_rows, 0.95, staircase=True)``
    (``mpipy.py:60-64``).  Evaluated in fp32 like TF does."""
    p = np.float32(step) * np.float32(batch) / np.float32(decay_rows)
    if staircase:
        p = np.floor(p)
    return float(np.float32(base) * np.power(np.float32(decay), p, dtype=np.float32))


# --------------------------------------------------------------------------
# the network
# --------------------------------------------------------------------------
class Oracle:
    """Forward / backward / update of the MNIST CNN on CPU fp32.

    ``emulate_bf16=True`` rounds the weights and every stored activation /
    activation-gradient to bf16 at the same points the device path does, so a
    bf16 kernel can be compared with a tolerance that only has to absorb
    accumulation-order differences.
    """

    def __init__(self, spec: ModelSpec, params: Optional[Params] = None, seed: int = 1,
                 weight_decay: float = 5e-4, emulate_bf16: bool = False):
        self.spec = spec
        self.params = params if params is not None else init_params(spec, seed)
        self.weight_decay = weight_decay
        self.emulate_bf16 = emulate_bf16
        self.slots: Dict[str, Dict[str, torch.Tensor]] = {}
        self.step_count = 0

    def _r(self, x):
        return bf16_round(x) if self.emulate_bf16 else x

    # ---- forward (``mpipy.py:155-167``) -----------------------------------
    def forward(self, x: torch.Tensor, keep_mask: Optional[torch.Tensor] = None, cache: bool = False):
        P = {k: self._r(v) for k, v in self.params.items()}
        spec = self.spec
        x = self._r(x.reshape(-1, spec.image_size, spec.image_size, 1).float())
        a1 = torch.relu(conv5x5(x, P["conv1_weight"]) + P["conv1_bias"])
        p1, i1 = maxpool2x2(a1)
        p1 = self._r(p1)
        a2 = torch.relu(conv5x5(p1, P["conv2_weight"]) + P["conv2_bias"])
        p2, i2 = maxpool2x2(a2)
        p2 = self._r(p2)
        flat = p2.reshape(p2.shape[0], -1)                       # (h, w, c) order
        h = torch.relu(flat @ P["fc1_weight"] + P["fc1_bias"])
        if keep_mask is not None:
            h = h * keep_mask.reshape(h.shape).float() / spec.keep_prob
        h = self._r(h)
        logits = h @ P["fc2_weight"] + P["fc2_bias"]
        if cache:
            self._cache = dict(x=x, p1=p1, i1=i1, p2=p2, i2=i2, flat=flat, h=h,
                               keep=keep_mask, P=P)
        return logits

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        """Softmax probabilities, dropout off (the reference leaves it on in
        eval, ``mpipy.py:166`` via ``:68``; see ``Trainer(eval_dropout=True)``)."""
        return torch.softmax(self.forward(x), dim=1)

    # ---- loss + gradients -------------------------------------------------
    def loss_and_grads(self, x, labels, keep_mask=None):
        spec = self.spec
        logits = self.forward(x, keep_mask, cache=True)
        c = self._cache
        P = c["P"]
        ce, dlogits, probs, errors = softmax_ce(logits, labels)
        l2 = sum(0.5 * (self.params[p.key] ** 2).sum() for p in spec.params if p.decay)
        loss = ce + self.weight_decay * l2
        g: Params = {}
        # FC2 (O20)
        g["fc2_weight"] = c["h"].t() @ dlogits
        g["fc2_bias"] = dlogits.sum(0)
        dh = dlogits @ P["fc2_weight"].t()
        # dropout + ReLU grad (O21, O18); h > 0 already implies kept & relu active
        dh = dh * (c["h"] > 0).float()
        if c["keep"] is not None:
            dh = dh / spec.keep_prob
        dh = self._r(dh)
        g["fc1_weight"] = c["flat"].t() @ dh
        g["fc1_bias"] = dh.sum(0)
        dflat = self._r(dh @ P["fc1_weight"].t())
        dp2 = dflat.reshape(c["p2"].shape)
        # pool2 + relu2 grad (O19, O18)
        da2 = maxpool2x2_bwd(dp2 * (c["p2"] > 0).float(), c["i2"])
        g["conv2_weight"] = conv5x5_wgrad(c["p1"], da2)
        g["conv2_bias"] = da2.sum((0, 1, 2))
        dp1 = self._r(conv5x5_dgrad(da2, P["conv2_weight"], c["p1"].shape))
        da1 = maxpool2x2_bwd(dp1 * (c["p1"] > 0).float(), c["i1"])
        g["conv1_weight"] = conv5x5_wgrad(c["x"], da1)
        g["conv1_bias"] = da1.sum((0, 1, 2))
        # L2 term on FC weights and biases (O12): grad = wd * x
        for p in spec.params:
            if p.decay and self.weight_decay:
                g[p.key] = g[p.key] + self.weight_decay * self.params[p.key]
        return float(loss), g, dict(ce=float(ce), probs=probs, errors=errors, logits=logits)

    # ---- optimizers ---------------------------------------------------------
    def apply(self, grads: Params, lr: float, optimizer: str = "momentum", momentum: float = 0.9,
              beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8):
        """TF-1 update rules.  momentum: ``accum = m*accum + g; w -= lr*accum``
        (``tf.train.MomentumOptimizer``, ``mpipy.py:65``).  adam:
        ``lr_t = lr*sqrt(1-b2^t)/(1-b1^t); w -= lr_t*m/(sqrt(v)+eps)``."""
        self.step_count += 1
        t = self.step_count
        for k, w in self.params.items():
            gk = grads[k].reshape(w.shape)
            s = self.slots.setdefault(k, {})
            if optimizer == "sgd":
                w -= lr * gk
            elif optimizer == "momentum":
                a = s.setdefault("m", torch.zeros_like(w))
                a.mul_(momentum).add_(gk)
                w -= lr * a
            elif optimizer == "adam":
                m = s.setdefault("m", torch.zeros_like(w))
                v = s.setdefault("v", torch.zeros_like(w))
                m.mul_(beta1).add_(gk, alpha=1 - beta1)
                v.mul_(beta2).addcmul_(gk, gk, value=1 - beta2)
                lr_t = lr * math.sqrt(1 - beta2 ** t) / (1 - beta1 ** t)
                w -= lr_t * m / (v.sqrt() + eps)
            else:
                raise ValueError(optimizer)

    def train_step(self, x, labels, lr, keep_mask=None, **opt):
        loss, g, aux = self.loss_and_grads(x, labels, keep_mask)
        self.apply(g, lr, **opt)
        return loss, aux

    # ---- flat-buffer helpers ------------------------------------------------
    def flatten(self, tensors: Params, size: Optional[int] = None) -> torch.Tensor:
        from .spec import flat_size
        out = torch.zeros(size or flat_size(self.spec), dtype=torch.float32)
        for p in self.spec.params:
            out[p.offset:p.offset + p.numel] = tensors[p.key].reshape(-1)
        return out

    def unflatten(self, flat: torch.Tensor) -> Params:
        return {p.key: flat[p.offset:p.offset + p.numel].reshape(p.shape).clone()
                for p in self.spec.params}


def dropout_keep_mask(spec: ModelSpec, batch: int, seed: int, step: int) -> torch.Tensor:
    """The mask the device FC1 epilogue generates for ``step`` (shared Philox stream)."""
    m = philox.dropout_mask(batch * spec.fc, spec.keep_prob, seed, step)
    return torch.from_numpy(m.reshape(batch, spec.fc))
