"""Dataset acquisition, parsing, splitting and sharding.

Replaces ``data_exist_here`` (``mpipy.py:185-199``), the missing
``convolutional.extract_data / extract_labels`` (``mpipy.py:12,215-218``; semantics
in SURVEY §2.2) and the split + shard arithmetic of ``main`` (``mpipy.py:211-241``).

Differences from the reference, all deliberate:
  * only rank 0 touches the filesystem/network; there is a real exception type
    (the reference names an undefined ``DownloadError``, ``mpipy.py:197``);
  * the train split really has ``tr_size`` rows — the reference scatters 50,000
    rows into buffers sized for 55,000 and trains on the zero tail
    (``mpipy.py:221`` vs ``:230``).  ``reference_bug=True`` reproduces that;
  * a synthetic generator exists because the build box has no network.
"""
from __future__ import annotations

import gzip
import os
import struct
from dataclasses import dataclass
from typing import Tuple

import numpy as np

from . import philox

DATA_URL = "https://storage.googleapis.com/cvdf-datasets/mnist/"   # mpipy.py:17
FILES = {
    "train_images": "train-images-idx3-ubyte.gz",
    "train_labels": "train-labels-idx1-ubyte.gz",
    "test_images": "t10k-images-idx3-ubyte.gz",
    "test_labels": "t10k-labels-idx1-ubyte.gz",
}
PIXEL_DEPTH = 255.0


class DownloadError(RuntimeError):
    """Raised when a dataset file is missing and cannot be fetched."""


def maybe_download(fname: str, data_dir: str = "./data", allow_network: bool = True) -> str:
    """Return the path of ``fname`` under ``data_dir``, downloading if absent."""
    os.makedirs(data_dir, exist_ok=True)
    path = os.path.join(data_dir, fname)
    if os.path.exists(path):
        return path
    if not allow_network:
        raise DownloadError(f"{path} not found and network access is disabled")
    import urllib.request
    tmp = path + f".part{os.getpid()}"
    try:
        urllib.request.urlretrieve(DATA_URL + fname, tmp)
        os.replace(tmp, path)                      # atomic: no torn file if ranks race
    except Exception as e:                         # noqa: BLE001
        if os.path.exists(tmp):
            os.remove(tmp)
        raise DownloadError(f"download of {fname} failed: {e}") from e
    return path


def _open(path: str):
    return gzip.open(path, "rb") if path.endswith(".gz") else open(path, "rb")


def extract_data(path: str, num_images: int, image_size: int = 28) -> np.ndarray:
    """IDX3 -> float32 ``[n, 28, 28, 1]`` rescaled to [-0.5, 0.5]:
    ``(x - 127.5) / 255``."""
    with _open(path) as f:
        magic, n, rows, cols = struct.unpack(">IIII", f.read(16))
        if magic != 2051:
            raise ValueError(f"{path}: bad IDX3 magic {magic}")
        if rows != image_size or cols != image_size:
            raise ValueError(f"{path}: {rows}x{cols} images, expected {image_size}")
        if num_images > n:
            raise ValueError(f"{path}: holds {n} images, {num_images} requested")
        buf = f.read(num_images * rows * cols)
    u8 = np.frombuffer(buf, dtype=np.uint8)
    return normalize_u8(u8).reshape(num_images, rows, cols, 1)


def extract_labels(path: str, num_images: int) -> np.ndarray:
    """IDX1 -> int64 ``[n]``."""
    with _open(path) as f:
        magic, n = struct.unpack(">II", f.read(8))
        if magic != 2049:
            raise ValueError(f"{path}: bad IDX1 magic {magic}")
        if num_images > n:
            raise ValueError(f"{path}: holds {n} labels, {num_images} requested")
        buf = f.read(num_images)
    return np.frombuffer(buf, dtype=np.uint8).astype(np.int64)


def normalize_u8(u8: np.ndarray) -> np.ndarray:
    return (u8.astype(np.float32) - np.float32(PIXEL_DEPTH / 2.0)) / np.float32(PIXEL_DEPTH)


def write_idx(images_u8: np.ndarray, labels_u8: np.ndarray, img_path: str, lbl_path: str) -> None:
    """Write an IDX3/IDX1 pair (gz if the name ends in .gz).  Used by tests and
    to export synthetic data in the format the reference downloads."""
    n, r, c = images_u8.shape[:3]
    op = gzip.open if img_path.endswith(".gz") else open
    with op(img_path, "wb") as f:
        f.write(struct.pack(">IIII", 2051, n, r, c))
        f.write(np.ascontiguousarray(images_u8, dtype=np.uint8).tobytes())
    op = gzip.open if lbl_path.endswith(".gz") else open
    with op(lbl_path, "wb") as f:
        f.write(struct.pack(">II", 2049, n))
        f.write(np.ascontiguousarray(labels_u8, dtype=np.uint8).tobytes())


# --------------------------------------------------------------------------
# split arithmetic (mpipy.py:211-222)
# --------------------------------------------------------------------------
@dataclass(frozen=True)
class SplitSizes:
    world: int
    tr_size: int       # 55000//N*N   mpipy.py:211
    ts_size: int       # 10000//N*N   mpipy.py:212
    val_size: int      # 5000//N*N    mpipy.py:213

    @property
    def train_rows_per_rank(self) -> int:
        return self.tr_size // self.world

    @property
    def test_rows_per_rank(self) -> int:
        return self.ts_size // self.world

    @property
    def val_rows_per_rank(self) -> int:
        return self.val_size // self.world

    def steps(self, epochs: int, batch: int) -> int:
        """``iteration * rows // batch`` (``mpipy.py:79``)."""
        return epochs * self.train_rows_per_rank // batch

    def reference_real_rows_per_rank(self) -> int:
        """Rows that carry data under the reference's bug: ``(tr_size - val_size)/N``."""
        return (self.tr_size - self.val_size) // self.world


def split_sizes(world: int, n_train: int = 55000, n_test: int = 10000, n_val: int = 5000) -> SplitSizes:
    return SplitSizes(world, n_train // world * world, n_test // world * world, n_val // world * world)


def split_train_val(images: np.ndarray, labels: np.ndarray, sizes: SplitSizes,
                    reference_bug: bool = False):
    """Root-side split.  The first ``val_size`` rows are validation
    (``mpipy.py:219-220``).  Fixed behaviour: the next ``tr_size`` rows are the
    train set.  ``reference_bug=True``: rows ``[val_size:tr_size)`` followed by a
    zero tail up to ``tr_size`` rows — exactly what each rank ends up holding
    after ``mpipy.py:221,230,236`` (SURVEY §2.1 "Data-split arithmetic")."""
    v, t = sizes.val_size, sizes.tr_size
    val = (images[:v], labels[:v])
    if not reference_bug:
        if images.shape[0] < v + t:
            raise ValueError(f"need {v + t} rows for val+train, have {images.shape[0]}")
        return val, (images[v:v + t], labels[v:v + t])
    # reference: per-rank buffer = tr_size/N rows, first (t-v)/N real, rest zero
    real_i, real_l = images[v:t], labels[v:t]
    n = sizes.world
    per, real_per = t // n, (t - v) // n
    ti = np.zeros((t,) + images.shape[1:], dtype=images.dtype)
    tl = np.zeros((t,), dtype=labels.dtype)
    for r in range(n):
        ti[r * per:r * per + real_per] = real_i[r * real_per:(r + 1) * real_per]
        tl[r * per:r * per + real_per] = real_l[r * real_per:(r + 1) * real_per]
    return val, (ti, tl)


def shard(array: np.ndarray, rank: int, world: int) -> np.ndarray:
    """Contiguous equal shard — what ``comm.Scatter`` delivers (``mpipy.py:236-241``)."""
    per = array.shape[0] // world
    return array[rank * per:(rank + 1) * per]


# --------------------------------------------------------------------------
# synthetic data (BASELINE.json: "synthetic 28x28 data")
# --------------------------------------------------------------------------
def synthetic_mnist(n: int, seed: int = 0, image_size: int = 28, num_classes: int = 10,
                    learnable: bool = True, offset: int = 0) -> Tuple[np.ndarray, np.ndarray]:
    """Deterministic fake MNIST: uint8 images ``[n,28,28]`` and uint8 labels.

    ``learnable=True`` draws each image as a class-specific blob pattern plus
    noise so that training visibly reduces the loss; otherwise pure noise.
    Rows ``offset..offset+n`` of one conceptual infinite dataset, so any rank can
    generate exactly its shard without materialising the rest."""
    npx = image_size * image_size
    lab = (philox.random_u32(n, seed, 0, philox.PURPOSE_DATA, offset) % np.uint32(num_classes)).astype(np.uint8)
    noise = philox.random_u32(n * npx, seed, 1, philox.PURPOSE_DATA, offset * npx)
    img = (noise >> np.uint32(24)).astype(np.float32).reshape(n, image_size, image_size)
    if learnable:
        yy, xx = np.mgrid[0:image_size, 0:image_size].astype(np.float32)
        protos = np.zeros((num_classes, image_size, image_size), dtype=np.float32)
        for c in range(num_classes):
            cy = 6.0 + 16.0 * ((c * 7) % num_classes) / num_classes
            cx = 6.0 + 16.0 * ((c * 3) % num_classes) / num_classes
            protos[c] = 255.0 * np.exp(-((yy - cy) ** 2 + (xx - cx) ** 2) / (2.0 * 3.0 ** 2))
        img = 0.35 * img + 0.65 * protos[lab]
    return np.clip(img, 0, 255).astype(np.uint8), lab


@dataclass
class Shards:
    train_x: np.ndarray
    train_y: np.ndarray
    test_x: np.ndarray
    test_y: np.ndarray
    val_x: np.ndarray
    val_y: np.ndarray


def load_root_dataset(source: str, sizes: SplitSizes, data_dir: str = "./data", seed: int = 0,
                      reference_bug: bool = False, n_total_train: int = 60000):
    """What rank 0 does at ``mpipy.py:214-222``: parse (or synthesise), split."""
    if source == "mnist":
        paths = {k: maybe_download(v, data_dir) for k, v in FILES.items()}
        tr_x = extract_data(paths["train_images"], n_total_train)
        tr_y = extract_labels(paths["train_labels"], n_total_train)
        ts_x = extract_data(paths["test_images"], sizes.ts_size)
        ts_y = extract_labels(paths["test_labels"], sizes.ts_size)
    elif source == "synthetic":
        need = sizes.val_size + sizes.tr_size
        u8, tr_y = synthetic_mnist(need, seed)
        tr_x = normalize_u8(u8).reshape(need, 28, 28, 1)
        tr_y = tr_y.astype(np.int64)
        u8, ts_y = synthetic_mnist(sizes.ts_size, seed, offset=1 << 20)
        ts_x = normalize_u8(u8).reshape(sizes.ts_size, 28, 28, 1)
        ts_y = ts_y.astype(np.int64)
    else:
        raise ValueError(f"unknown data source {source!r}")
    (val_x, val_y), (trn_x, trn_y) = split_train_val(tr_x, tr_y, sizes, reference_bug)
    return Shards(trn_x, trn_y, ts_x, ts_y, val_x, val_y)


class SequentialSampler:
    """The reference's batch window: ``offset = (step*B) % (rows - B)``
    (``mpipy.py:80-82``) — unshuffled, wraps, never emits a ragged batch."""

    def __init__(self, rows: int, batch: int):
        if rows <= batch:
            raise ValueError(f"shard of {rows} rows is not larger than batch {batch}")
        self.rows, self.batch = rows, batch

    def offset(self, step: int) -> int:
        return (step * self.batch) % (self.rows - self.batch)

    def indices(self, step: int) -> np.ndarray:
        o = self.offset(step)
        return np.arange(o, o + self.batch)


class ShuffleSampler:
    """Optional epoch-wise shuffle (not in the reference)."""

    def __init__(self, rows: int, batch: int, seed: int = 0):
        self.rows, self.batch, self.seed = rows, batch, seed
        self.per_epoch = rows // batch

    def indices(self, step: int) -> np.ndarray:
        epoch, i = divmod(step, self.per_epoch)
        perm = np.random.default_rng(self.seed + epoch).permutation(self.rows)
        return perm[i * self.batch:(i + 1) * self.batch]
