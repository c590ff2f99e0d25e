"""ctypes binding of ``libmtb200_host.so`` (C ABI in ``include/mtb200_host.h``)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

_LIB = None

OK, ERR_ARG, ERR_SYS, ERR_TIMEOUT, ERR_ABORTED, ERR_FORMAT = range(6)


def lib() -> C.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    path = Path(__file__).resolve().parents[1] / "lib" / "libmtb200_host.so"
    if not path.exists():
        from . import build
        build.build_host()
    L = C.CDLL(str(path))
    vp, i32, u64, i64 = C.c_void_p, C.c_int, C.c_uint64, C.c_int64
    L.mtb_shm_open.argtypes = [C.c_char_p, i32, i32, u64, i32, C.POINTER(vp)]
    L.mtb_shm_open.restype = i32
    for name, args in {
        "mtb_shm_close": [vp], "mtb_shm_rank": [vp], "mtb_shm_size": [vp], "mtb_shm_barrier": [vp],
        "mtb_shm_bcast": [vp, vp, u64, i32],
        "mtb_shm_scatter": [vp, vp, vp, u64, i32],
        "mtb_shm_gather": [vp, vp, vp, u64, i32],
        "mtb_shm_allgather": [vp, vp, vp, u64],
        "mtb_shm_allreduce_f32": [vp, vp, u64, i32],
        "mtb_idx_read_images": [C.c_char_p, i64, i64, i64, vp, vp],
        "mtb_idx_read_labels": [C.c_char_p, i64, vp],
    }.items():
        getattr(L, name).argtypes = args
        getattr(L, name).restype = i32
    L.mtb_shm_set_timeout.argtypes = [vp, i32]
    L.mtb_shm_set_timeout.restype = None
    L.mtb_shm_abort.argtypes = [vp]
    L.mtb_shm_abort.restype = None
    L.mtb_shm_heartbeat.argtypes = [vp, i32]
    L.mtb_shm_heartbeat.restype = u64
    L.mtb_normalize_u8.argtypes = [vp, vp, i64]
    L.mtb_normalize_u8.restype = None
    L.mtb_crc32c.argtypes = [vp, u64, C.c_uint32]
    L.mtb_crc32c.restype = C.c_uint32
    _LIB = L
    return L


def crc32c(data: bytes, init: int = 0) -> int:
    buf = C.create_string_buffer(data, len(data))
    return int(lib().mtb_crc32c(C.cast(buf, C.c_void_p), len(data), init))


def read_idx_images(path: str, n: int, rows: int = 28, cols: int = 28):
    import numpy as np
    out = np.empty((n, rows, cols, 1), dtype=np.float32)
    rc = lib().mtb_idx_read_images(path.encode(), n, rows, cols, out.ctypes.data, None)
    if rc != OK:
        raise ValueError(f"mtb_idx_read_images({path}) failed with code {rc}")
    return out


def read_idx_labels(path: str, n: int):
    import numpy as np
    out = np.empty((n,), dtype=np.int64)
    rc = lib().mtb_idx_read_labels(path.encode(), n, out.ctypes.data)
    if rc != OK:
        raise ValueError(f"mtb_idx_read_labels({path}) failed with code {rc}")
    return out
