"""Philox4x32-10 counter-based RNG in NumPy.

The reference draws dropout masks and truncated-normal initial weights from
TensorFlow's Philox generator (``mpipy.py:166`` ``tf.nn.dropout(..., seed=1)``;
``mpipy.py:38-52`` ``tf.truncated_normal(..., seed=1)``).  Bit parity with TF's
stream is neither feasible nor useful; what matters is
that the CPU oracle and the CUDA kernels of THIS framework draw the same
numbers, so that a full train step with dropout enabled can be compared
element by element.  ``csrc/philox.cuh`` implements exactly this function.

Stream convention (shared with the kernels):
  counter = (index // 4, stream_lo, stream_hi, purpose), key = (seed_lo, seed_hi)
  value   = output word ``index % 4``
"""
from __future__ import annotations

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = np.uint32(0x9E3779B9)
_W1 = np.uint32(0xBB67AE85)
_MASK = np.uint64(0xFFFFFFFF)

PURPOSE_DROPOUT = 1
PURPOSE_INIT = 2
PURPOSE_DATA = 3


def philox4x32(c0, c1, c2, c3, k0, k1, rounds: int = 10):
    """Vectorised Philox4x32.  All inputs broadcastable uint32 arrays."""
    c0 = np.asarray(c0, dtype=np.uint32).copy()
    c1 = np.broadcast_to(np.asarray(c1, dtype=np.uint32), c0.shape).copy()
    c2 = np.broadcast_to(np.asarray(c2, dtype=np.uint32), c0.shape).copy()
    c3 = np.broadcast_to(np.asarray(c3, dtype=np.uint32), c0.shape).copy()
    k0 = np.uint32(k0)
    k1 = np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(rounds):
            p0 = _M0 * c0.astype(np.uint64)
            p1 = _M1 * c2.astype(np.uint64)
            hi0 = (p0 >> np.uint64(32)).astype(np.uint32)
            lo0 = (p0 & _MASK).astype(np.uint32)
            hi1 = (p1 >> np.uint64(32)).astype(np.uint32)
            lo1 = (p1 & _MASK).astype(np.uint32)
            n0 = hi1 ^ c1 ^ k0
            n1 = lo1
            n2 = hi0 ^ c3 ^ k1
            n3 = lo0
            c0, c1, c2, c3 = n0, n1, n2, n3
            k0 = np.uint32((int(k0) + int(_W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(_W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


def random_u32(n: int, seed: int, stream: int, purpose: int, offset: int = 0) -> np.ndarray:
    """``n`` uint32 values for element indices ``offset .. offset+n``.  One Philox evaluation yields the four
    words of a block of four consecutive indices, so only the blocks that are touched are evaluated (in chunks,
    to bound the temporaries)."""
    if n <= 0:
        return np.empty(0, dtype=np.uint32)
    b0, b1 = offset // 4, (offset + n + 3) // 4
    out = np.empty((b1 - b0) * 4, dtype=np.uint32)
    chunk = 1 << 20
    for lo in range(b0, b1, chunk):
        hi = min(b1, lo + chunk)
        blk = np.arange(lo, hi, dtype=np.uint64).astype(np.uint32)
        w = philox4x32(blk, np.uint32(stream & 0xFFFFFFFF), np.uint32((stream >> 32) & 0xFFFFFFFF),
                       np.uint32(purpose), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        out[(lo - b0) * 4:(hi - b0) * 4] = np.stack(w, axis=1).reshape(-1)
    skip = offset - b0 * 4
    return out[skip:skip + n]


def uniform01(n: int, seed: int, stream: int, purpose: int, offset: int = 0) -> np.ndarray:
    """fp32 uniforms in [0, 1): ``(u32 >> 8) * 2^-24`` (exact in fp32)."""
    u = random_u32(n, seed, stream, purpose, offset)
    return ((u >> np.uint32(8)).astype(np.float32)) * np.float32(1.0 / 16777216.0)


def dropout_mask(n: int, keep_prob: float, seed: int, step: int) -> np.ndarray:
    """Boolean keep-mask; element kept iff u < keep_prob (TF keeps iff
    ``floor(keep + u) == 1`` — same distribution, ``mpipy.py:166``)."""
    return uniform01(n, seed, step, PURPOSE_DROPOUT) < np.float32(keep_prob)


def truncated_normal(n: int, stddev: float, seed: int, stream: int) -> np.ndarray:
    """Normal(0, stddev) truncated to +-2 stddev by resampling (the semantics of
    ``tf.truncated_normal``, ``mpipy.py:38``).  Attempt ``t`` of element ``i``
    uses Philox counter (i, stream, t, PURPOSE_INIT): words 0,1 feed one
    Box-Muller draw.  At most 16 attempts; P(all fail) < 1e-21."""
    out = np.zeros(n, dtype=np.float32)
    todo = np.arange(n, dtype=np.uint32)
    for t in range(16):
        if todo.size == 0:
            break
        w0, w1, _, _ = philox4x32(todo, np.uint32(stream), np.uint32(t), np.uint32(PURPOSE_INIT),
                                  seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
        u1 = ((w0 >> np.uint32(8)).astype(np.float32) + np.float32(1.0)) * np.float32(1.0 / 16777216.0)  # (0,1]
        u2 = (w1 >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)
        r = np.sqrt(np.float32(-2.0) * np.log(u1)).astype(np.float32)
        z = (r * np.cos(np.float32(2.0 * np.pi) * u2)).astype(np.float32)
        ok = np.abs(z) <= np.float32(2.0)
        out[todo[ok]] = z[ok] * np.float32(stddev)
        todo = todo[~ok]
    return out
