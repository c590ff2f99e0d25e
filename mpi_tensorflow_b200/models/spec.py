"""Model presets and the parameter registry.

The reference hard-codes one network as eight unnamed ``tf.Variable``s
(``mpipy.py:38-53``) and a forward function (``mpipy.py:155-167``).  Here the
network is parametric: a :class:`ModelSpec` fixes channel counts, FC width,
dropout keep-probability and class count, and derives every tensor shape, the
TensorFlow variable names used for checkpoint exchange and the
flat-buffer offsets that the fused allreduce/update kernel works on.

Presets
-------
``tutorial512``   bit-for-bit the shapes of ``mpipy.py:38-53`` (FC 512, keep 0.5)
``official1024``  the tensorflow/models official MNIST CNN that ``README.md:4``
                  points at (FC 1024, dropout rate 0.4 => keep 0.6)
``wide``          conv 128/256 channels, FC 4096 (comm-bound config of
                  ``BASELINE.json``), 52,248,074 parameters
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Tuple

IMAGE_SIZE = 28          # mpipy.py:19
NUM_CLASSES = 10         # mpipy.py:21 (misnamed ``num_channel`` there)
KSIZE = 5                # 5x5 kernels, mpipy.py:38,42


@dataclass(frozen=True)
class ParamInfo:
    """One trainable tensor.

    ``shape`` is the TensorFlow layout (conv: HWIO, dense: [in, out]) which is
    also the layout of the fp32 master copy in the flat parameter buffer.
    """
    key: str                 # short attribute name, e.g. ``fc1_weight`` (mpipy.py:46)
    shape: Tuple[int, ...]
    tf1_name: str            # name under the shipped script's unnamed variables
    keras_name: str          # name in the official-model checkpoint
    init: str                # ``trunc_normal`` | ``zeros`` | ``const``
    init_value: float        # stddev for trunc_normal, value for const
    decay: bool              # L2 term applies (mpipy.py:57-58: FC weights and biases)
    is_bias: bool
    layer: str               # conv1 | conv2 | fc1 | fc2 (bucket id for overlap)
    offset: int = 0          # element offset in the flat buffer (filled by ModelSpec)

    @property
    def numel(self) -> int:
        n = 1
        for s in self.shape:
            n *= s
        return n


@dataclass(frozen=True)
class ModelSpec:
    name: str
    c1: int = 32             # conv1 output channels (mpipy.py:38)
    c2: int = 64             # conv2 output channels (mpipy.py:42)
    fc: int = 512            # hidden width (mpipy.py:46)
    keep_prob: float = 0.5   # mpipy.py:166
    num_classes: int = NUM_CLASSES
    image_size: int = IMAGE_SIZE
    params: Tuple[ParamInfo, ...] = field(default=(), compare=False)

    # ---- derived sizes -------------------------------------------------
    @property
    def pooled1(self) -> int:          # 14
        return (self.image_size + 1) // 2

    @property
    def pooled2(self) -> int:          # 7
        return (self.pooled1 + 1) // 2

    @property
    def flat_dim(self) -> int:         # 3136 = 7*7*64, flatten order (h, w, c): mpipy.py:162-164
        return self.pooled2 * self.pooled2 * self.c2

    @property
    def num_params(self) -> int:
        return sum(p.numel for p in self.params)

    def param(self, key: str) -> ParamInfo:
        for p in self.params:
            if p.key == key:
                return p
        raise KeyError(key)

    @property
    def keys(self) -> List[str]:
        return [p.key for p in self.params]

    def buckets(self) -> List[Tuple[str, int, int]]:
        """(layer, offset, numel) per layer, in REVERSE layer order.

        This is the order gradients become available during backward (fc2 first,
        conv1 last) and therefore the static launch order of the fused
        allreduce buckets (SURVEY §5.5 "Overlap").  Weight and bias of a layer
        are adjacent in the flat buffer so a bucket is one contiguous range.
        """
        out = []
        for layer in ("fc2", "fc1", "conv2", "conv1"):
            ps = [p for p in self.params if p.layer == layer]
            off = min(p.offset for p in ps)
            end = max(p.offset + p.numel for p in ps)
            out.append((layer, off, _align(end, 16) - off))
        return out

    def flops_per_image_fwd(self) -> float:
        s, k = self.image_size, KSIZE
        conv1 = 2.0 * s * s * k * k * 1 * self.c1
        conv2 = 2.0 * self.pooled1 ** 2 * k * k * self.c1 * self.c2
        fc1 = 2.0 * self.flat_dim * self.fc
        fc2 = 2.0 * self.fc * self.num_classes
        return conv1 + conv2 + fc1 + fc2


def _align(n: int, a: int) -> int:
    return (n + a - 1) // a * a


def make_spec(name: str, c1: int, c2: int, fc: int, keep_prob: float,
              num_classes: int = NUM_CLASSES, image_size: int = IMAGE_SIZE) -> ModelSpec:
    p2 = ((image_size + 1) // 2 + 1) // 2
    flat = p2 * p2 * c2
    raw = [
        # key, shape, tf1 name, keras name, init, value, decay, is_bias, layer
        ("conv1_weight", (KSIZE, KSIZE, 1, c1), "Variable", "conv2d/kernel", "trunc_normal", 0.1, False, False, "conv1"),
        ("conv1_bias", (c1,), "Variable_1", "conv2d/bias", "zeros", 0.0, False, True, "conv1"),       # mpipy.py:41
        ("conv2_weight", (KSIZE, KSIZE, c1, c2), "Variable_2", "conv2d_1/kernel", "trunc_normal", 0.1, False, False, "conv2"),
        ("conv2_bias", (c2,), "Variable_3", "conv2d_1/bias", "const", 0.1, False, True, "conv2"),     # mpipy.py:45
        ("fc1_weight", (flat, fc), "Variable_4", "dense/kernel", "trunc_normal", 0.1, True, False, "fc1"),
        ("fc1_bias", (fc,), "Variable_5", "dense/bias", "const", 0.1, True, True, "fc1"),             # mpipy.py:49
        ("fc2_weight", (fc, num_classes), "Variable_6", "dense_1/kernel", "trunc_normal", 0.1, True, False, "fc2"),
        ("fc2_bias", (num_classes,), "Variable_7", "dense_1/bias", "const", 0.1, True, True, "fc2"),  # mpipy.py:53
    ]
    params = []
    off = 0
    for key, shape, tf1, keras, init, val, decay, is_bias, layer in raw:
        # every tensor starts on a 16-element boundary (64 B in fp32, 32 B in bf16)
        # so the fused kernels can use 16-byte vector accesses on every parameter
        off = _align(off, 16)
        pi = ParamInfo(key, tuple(shape), tf1, keras, init, val, decay, is_bias, layer, off)
        params.append(pi)
        off += pi.numel
    return ModelSpec(name=name, c1=c1, c2=c2, fc=fc, keep_prob=keep_prob,
                     num_classes=num_classes, image_size=image_size, params=tuple(params))


PRESETS: Dict[str, ModelSpec] = {
    "tutorial512": make_spec("tutorial512", 32, 64, 512, 0.5),
    "official1024": make_spec("official1024", 32, 64, 1024, 0.6),
    "wide": make_spec("wide", 128, 256, 4096, 0.6),
    # not a reference config: a small network for fast CPU tests of the plumbing
    "tiny": make_spec("tiny", 8, 16, 32, 0.5),
}


def get_spec(name: str) -> ModelSpec:
    try:
        return PRESETS[name]
    except KeyError:
        raise ValueError(f"unknown model preset {name!r}; choose from {sorted(PRESETS)}")


def flat_size(spec: ModelSpec) -> int:
    """Length (elements) of the flat parameter / gradient / slot buffers, padded
    so that any world size up to 16 divides it into 16-byte aligned slices."""
    last = spec.params[-1]
    return _align(last.offset + last.numel, 16 * 16 * 4)
