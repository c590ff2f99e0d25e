"""Checkpoint export / import with TensorFlow variable names and layouts.

The reference never writes a checkpoint (no ``tf.train.Saver``; SURVEY §5
"Checkpoint / resume"), so "the reference's checkpoint" is defined by variable
creation order and layout, which *are* in the reference (``mpipy.py:38-53,59,65``):

  tf1 names    ``Variable`` .. ``Variable_7``, ``Variable_8`` (iter_), ``<v>/Momentum``
  keras names  ``conv2d/kernel`` .. ``dense_1/bias``, ``global_step``, ``<v>/Adam``,
               ``<v>/Adam_1``, ``beta1_power``, ``beta2_power``  (official model)

Two on-disk formats, same name -> tensor map, TF layouts (HWIO, [in,out]):
  * ``.npz``  — flat archive, loadable with one line of NumPy;
  * TF bundle-V2 — ``<prefix>.index`` (a LevelDB-format table of BundleEntryProto)
    + ``<prefix>.data-00000-of-00001`` (raw little-endian tensors, masked CRC32C),
    the format ``tf.train.Saver`` / ``tf.train.load_checkpoint`` use.  No
    TensorFlow is installed here, so fidelity is established by self round-trip
    and by the format invariants the tests check (magic, CRCs, sorted keys).
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Optional, Tuple

import numpy as np

from ..models.spec import ModelSpec

DT_FLOAT, DT_INT32, DT_INT64 = 1, 3, 9
_DT = {np.dtype("float32"): DT_FLOAT, np.dtype("int32"): DT_INT32, np.dtype("int64"): DT_INT64}
_NP = {v: k for k, v in _DT.items()}
_TABLE_MAGIC = 0xDB4775248B80FB57
_MASK_DELTA = 0xA282EAD8


# ---------------------------------------------------------------------------
# name maps
# ---------------------------------------------------------------------------
def name_map(spec: ModelSpec, style: str = "tf1") -> Dict[str, str]:
    """param key -> checkpoint variable name."""
    if style == "tf1":
        return {p.key: p.tf1_name for p in spec.params}
    if style == "keras":
        return {p.key: p.keras_name for p in spec.params}
    raise ValueError(style)


def slot_suffixes(optimizer: str) -> Dict[str, str]:
    """optimizer slot -> TF name suffix."""
    if optimizer == "momentum":
        return {"m": "/Momentum"}                   # created by mpipy.py:65
    if optimizer == "adam":
        return {"m": "/Adam", "v": "/Adam_1"}
    return {}


def to_named(spec: ModelSpec, params: Dict[str, np.ndarray], slots: Optional[Dict[str, Dict[str, np.ndarray]]],
             step: int, optimizer: str, style: str = "tf1", beta1: float = 0.9, beta2: float = 0.999):
    names = name_map(spec, style)
    out: Dict[str, np.ndarray] = {}
    for p in spec.params:
        out[names[p.key]] = np.ascontiguousarray(np.asarray(params[p.key], dtype=np.float32).reshape(p.shape))
        for slot, suffix in slot_suffixes(optimizer).items():
            if slots and p.key in slots and slot in slots[p.key]:
                out[names[p.key] + suffix] = np.ascontiguousarray(
                    np.asarray(slots[p.key][slot], dtype=np.float32).reshape(p.shape))
    if style == "tf1":
        out["Variable_8"] = np.asarray(step, dtype=np.float32)       # iter_, fp32 scalar, mpipy.py:59
    else:
        out["global_step"] = np.asarray(step, dtype=np.int64)
        if optimizer == "adam":
            out["beta1_power"] = np.asarray(beta1 ** (step + 1), dtype=np.float32)
            out["beta2_power"] = np.asarray(beta2 ** (step + 1), dtype=np.float32)
    return out


def from_named(spec: ModelSpec, named: Dict[str, np.ndarray], optimizer: str, style: str = "tf1",
               dense_channels_first: bool = False):
    """Inverse of :func:`to_named`.  ``dense_channels_first=True`` converts a
    ``dense/kernel`` whose rows are ordered (c,h,w) — the upstream official
    model's GPU default — to this framework's (h,w,c) order."""
    names = name_map(spec, style)
    params, slots = {}, {}
    for p in spec.params:
        n = names[p.key]
        if n not in named:
            raise KeyError(f"checkpoint lacks variable {n!r} ({p.key})")

        def fix(a):
            a = np.asarray(a, dtype=np.float32)
            if a.shape != p.shape:
                raise ValueError(f"{n}: shape {a.shape}, expected {p.shape}")
            if dense_channels_first and p.key == "fc1_weight":
                s, c = spec.pooled2, spec.c2
                a = a.reshape(c, s, s, spec.fc).transpose(1, 2, 0, 3).reshape(p.shape)
            return np.ascontiguousarray(a)

        params[p.key] = fix(named[n])
        for slot, suffix in slot_suffixes(optimizer).items():
            if n + suffix in named:
                slots.setdefault(p.key, {})[slot] = fix(named[n + suffix])
    step_name = "Variable_8" if style == "tf1" else "global_step"
    step = int(np.asarray(named[step_name]).item()) if step_name in named else 0
    return params, slots, step


# ---------------------------------------------------------------------------
# npz
# ---------------------------------------------------------------------------
def save_npz(path: str, named: Dict[str, np.ndarray]) -> None:
    tmp = path + f".tmp{os.getpid()}"
    with open(tmp, "wb") as f:
        np.savez(f, **{k.replace("/", "%2F"): v for k, v in named.items()})
    os.replace(tmp, path)


def load_npz(path: str) -> Dict[str, np.ndarray]:
    with np.load(path) as z:
        return {k.replace("%2F", "/"): z[k] for k in z.files}


# ---------------------------------------------------------------------------
# CRC32C
# ---------------------------------------------------------------------------
def _crc32c(data: bytes) -> int:
    try:
        from ..ops.native_host import crc32c
        return crc32c(data)
    except Exception:  # noqa: BLE001  (pure-python fallback, slow)
        tbl = _py_table()
        c = 0xFFFFFFFF
        for b in data:
            c = tbl[(c ^ b) & 0xFF] ^ (c >> 8)
        return c ^ 0xFFFFFFFF


_PY_TABLE = None


def _py_table():
    global _PY_TABLE
    if _PY_TABLE is None:
        t = []
        for i in range(256):
            c = i
            for _ in range(8):
                c = (c >> 1) ^ 0x82F63B78 if c & 1 else c >> 1
            t.append(c)
        _PY_TABLE = t
    return _PY_TABLE


def masked_crc32c(data: bytes) -> int:
    c = _crc32c(data)
    return ((((c >> 15) | (c << 17)) & 0xFFFFFFFF) + _MASK_DELTA) & 0xFFFFFFFF


# ---------------------------------------------------------------------------
# minimal protobuf wire helpers
# ---------------------------------------------------------------------------
def _varint(n: int) -> bytes:
    n &= (1 << 64) - 1
    out = bytearray()
    while True:
        b = n & 0x7F
        n >>= 7
        if n:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _read_varint(buf: bytes, pos: int) -> Tuple[int, int]:
    shift = val = 0
    while True:
        b = buf[pos]
        pos += 1
        val |= (b & 0x7F) << shift
        if not b & 0x80:
            return val, pos
        shift += 7


def _field(num: int, wire: int, payload: bytes) -> bytes:
    return _varint((num << 3) | wire) + payload


def _parse(buf: bytes) -> Dict[int, list]:
    out: Dict[int, list] = {}
    pos = 0
    while pos < len(buf):
        key, pos = _read_varint(buf, pos)
        num, wire = key >> 3, key & 7
        if wire == 0:
            v, pos = _read_varint(buf, pos)
        elif wire == 2:
            n, pos = _read_varint(buf, pos)
            v = buf[pos:pos + n]
            pos += n
        elif wire == 5:
            v = struct.unpack_from("<I", buf, pos)[0]
            pos += 4
        elif wire == 1:
            v = struct.unpack_from("<Q", buf, pos)[0]
            pos += 8
        else:
            raise ValueError(f"unsupported wire type {wire}")
        out.setdefault(num, []).append(v)
    return out


def _shape_proto(shape) -> bytes:
    # TensorShapeProto { repeated Dim dim = 2 { int64 size = 1 } }
    return b"".join(_field(2, 2, _varint(len(d)) + d) for d in (_field(1, 0, _varint(int(s))) for s in shape))


def _entry_proto(dtype: int, shape, offset: int, size: int, crc: int) -> bytes:
    # BundleEntryProto: dtype=1, shape=2, shard_id=3, offset=4, size=5, crc32c=6 (fixed32)
    sp = _shape_proto(shape)
    out = _field(1, 0, _varint(dtype)) + _field(2, 2, _varint(len(sp)) + sp)
    if offset:
        out += _field(4, 0, _varint(offset))
    out += _field(5, 0, _varint(size)) + _field(6, 5, struct.pack("<I", crc))
    return out


def _header_proto() -> bytes:
    # BundleHeaderProto: num_shards=1, endianness=2 (LITTLE=0, omitted), version=3 {producer=1}
    ver = _field(1, 0, _varint(1))
    return _field(1, 0, _varint(1)) + _field(3, 2, _varint(len(ver)) + ver)


# ---------------------------------------------------------------------------
# LevelDB table (uncompressed)
# ---------------------------------------------------------------------------
def _block(entries, restart_interval: int = 16) -> bytes:
    out = bytearray()
    restarts = []
    last = b""
    for i, (k, v) in enumerate(entries):
        if i % restart_interval == 0:
            restarts.append(len(out))
            shared = 0
        else:
            shared = 0
            m = min(len(last), len(k))
            while shared < m and last[shared] == k[shared]:
                shared += 1
        out += _varint(shared) + _varint(len(k) - shared) + _varint(len(v)) + k[shared:] + v
        last = k
    if not restarts:
        restarts = [0]
    for r in restarts:
        out += struct.pack("<I", r)
    out += struct.pack("<I", len(restarts))
    return bytes(out)


def _emit_block(f, contents: bytes) -> bytes:
    """Write block + 5-byte trailer; return its BlockHandle encoding."""
    off = f.tell()
    f.write(contents)
    trailer_type = b"\x00"                                     # kNoCompression
    f.write(trailer_type + struct.pack("<I", masked_crc32c(contents + trailer_type)))
    return _varint(off) + _varint(len(contents))


def _write_table(path: str, items: Dict[bytes, bytes], block_bytes: int = 4096) -> None:
    keys = sorted(items)
    with open(path, "wb") as f:
        index_entries = []
        cur, cur_size = [], 0
        for k in keys:
            cur.append((k, items[k]))
            cur_size += len(k) + len(items[k]) + 8
            if cur_size >= block_bytes:
                index_entries.append((cur[-1][0], _emit_block(f, _block(cur))))
                cur, cur_size = [], 0
        if cur:
            index_entries.append((cur[-1][0], _emit_block(f, _block(cur))))
        meta_handle = _emit_block(f, _block([]))
        index_handle = _emit_block(f, _block(index_entries, restart_interval=1))
        footer = meta_handle + index_handle
        footer += b"\x00" * (40 - len(footer))
        footer += struct.pack("<Q", _TABLE_MAGIC)
        f.write(footer)


def _read_block(buf: bytes, off: int, size: int, verify: bool = True):
    contents = buf[off:off + size]
    ctype = buf[off + size]
    crc = struct.unpack_from("<I", buf, off + size + 1)[0]
    if ctype != 0:
        raise ValueError("compressed table blocks are not supported")
    if verify and crc != masked_crc32c(contents + bytes([ctype])):
        raise ValueError("table block CRC mismatch")
    n_restarts = struct.unpack_from("<I", contents, len(contents) - 4)[0]
    end = len(contents) - 4 - 4 * n_restarts
    pos, last = 0, b""
    out = []
    while pos < end:
        shared, pos = _read_varint(contents, pos)
        non_shared, pos = _read_varint(contents, pos)
        vlen, pos = _read_varint(contents, pos)
        k = last[:shared] + contents[pos:pos + non_shared]
        pos += non_shared
        out.append((k, contents[pos:pos + vlen]))
        pos += vlen
        last = k
    return out


def _read_table(path: str) -> Dict[bytes, bytes]:
    buf = open(path, "rb").read()
    if len(buf) < 48 or struct.unpack_from("<Q", buf, len(buf) - 8)[0] != _TABLE_MAGIC:
        raise ValueError(f"{path}: not a table file (bad magic)")
    footer = buf[-48:]
    pos = 0
    _, pos = _read_varint(footer, pos)          # metaindex offset
    _, pos = _read_varint(footer, pos)          # metaindex size
    ioff, pos = _read_varint(footer, pos)
    isize, pos = _read_varint(footer, pos)
    out = {}
    for _, handle in _read_block(buf, ioff, isize):
        boff, p = _read_varint(handle, 0)
        bsize, p = _read_varint(handle, p)
        for k, v in _read_block(buf, boff, bsize):
            out[k] = v
    return out


# ---------------------------------------------------------------------------
# TF bundle V2
# ---------------------------------------------------------------------------
def save_tf_bundle(prefix: str, named: Dict[str, np.ndarray]) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(prefix)), exist_ok=True)
    data_path = prefix + ".data-00000-of-00001"
    items = {b"": _header_proto()}
    off = 0
    with open(data_path + ".tmp", "wb") as f:
        for name in sorted(named):
            a = np.asarray(named[name], order="C")      # (ascontiguousarray would turn 0-d into 1-d)
            if a.dtype not in _DT:
                raise TypeError(f"{name}: dtype {a.dtype} not supported in bundle")
            raw = a.astype(a.dtype.newbyteorder("<"), copy=False).tobytes()
            f.write(raw)
            items[name.encode()] = _entry_proto(_DT[a.dtype], a.shape, off, len(raw), masked_crc32c(raw))
            off += len(raw)
    os.replace(data_path + ".tmp", data_path)
    _write_table(prefix + ".index.tmp", items)
    os.replace(prefix + ".index.tmp", prefix + ".index")
    # the ``checkpoint`` state file tf.train.latest_checkpoint reads
    with open(os.path.join(os.path.dirname(os.path.abspath(prefix)), "checkpoint"), "w") as f:
        base = os.path.basename(prefix)
        f.write(f'model_checkpoint_path: "{base}"\nall_model_checkpoint_paths: "{base}"\n')


def load_tf_bundle(prefix: str, verify: bool = True) -> Dict[str, np.ndarray]:
    table = _read_table(prefix + ".index")
    hdr = _parse(table.get(b"", b""))
    shards = hdr.get(1, [1])[0]
    if shards != 1:
        raise ValueError(f"bundle has {shards} shards; only single-shard bundles are supported")
    if hdr.get(2, [0])[0] != 0:
        raise ValueError("big-endian bundles are not supported")
    data = open(prefix + ".data-00000-of-00001", "rb").read()
    out = {}
    for k, v in table.items():
        if k == b"":
            continue
        e = _parse(v)
        dtype = e.get(1, [0])[0]
        if dtype not in _NP:
            continue                               # e.g. strings (save counters); not ours
        shape = []
        if 2 in e:
            for d in _parse(e[2][0]).get(2, []):
                shape.append(_parse(d).get(1, [0])[0])
        off, size = e.get(4, [0])[0], e.get(5, [0])[0]
        raw = data[off:off + size]
        if verify and 6 in e and e[6][0] != masked_crc32c(raw):
            raise ValueError(f"{k.decode()}: tensor CRC mismatch")
        out[k.decode()] = np.frombuffer(raw, dtype=_NP[dtype].newbyteorder("<")).reshape(shape).copy()
    return out


def save(path: str, named: Dict[str, np.ndarray]) -> None:
    if path.endswith(".npz"):
        save_npz(path, named)
    else:
        save_tf_bundle(path, named)


def load(path: str) -> Dict[str, np.ndarray]:
    if path.endswith(".npz"):
        return load_npz(path)
    if path.endswith(".index"):
        path = path[:-len(".index")]
    return load_tf_bundle(path)
