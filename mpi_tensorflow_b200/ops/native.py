"""ctypes binding of ``libmtb200.so`` (C ABI in ``include/mtb200.h``).

The library is built in-tree by ``ops/build.py`` for sm_100a only.  On a GPU box a
missing or unloadable library is a hard error — there is no PyTorch fallback on
the compute path (the round-end driver records which .so files were loaded).
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Optional

import torch

_LIB: Optional[C.CDLL] = None
LAUNCHES = 0          # kernels launched through this binding (bench.py reports it)

vp, i32, i64, u64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_ulonglong, C.c_float


class FusedArgs(C.Structure):
    _fields_ = [
        ("world", i32), ("rank", i32), ("algo", i32), ("optimizer", i32),
        ("grads", vp * 16), ("params", vp * 16), ("shadow", vp * 16),
        ("mc_grads", vp), ("mc_params", vp), ("mc_shadow", vp),
        ("slot_m", vp), ("slot_v", vp),
        ("flags", vp * 16), ("seq", vp), ("status", vp),
        ("offset", i64), ("count", i64),
        ("scale", f32), ("lr0", f32), ("lr_decay", f32), ("lr_override", f32),
        ("batch", i32), ("decay_rows", i64), ("step_ptr", vp),
        ("momentum", f32), ("beta1", f32), ("beta2", f32), ("eps", f32), ("wd", f32),
        ("n_ranges", i32), ("range_begin", i64 * 8), ("range_end", i64 * 8),
        ("grid", i32), ("block", i32),
        ("n_sharded", i32), ("shard_begin", i64 * 4), ("shard_end", i64 * 4),
        ("advance_step", vp), ("done_counter", vp),
        ("loss_src", vp), ("loss_mail", vp), ("loss_ring", i32),
        ("n_prered", i32), ("prered_begin", i64 * 2), ("prered_end", i64 * 2),
        ("chan", i32), ("no_flags", i32),
        ("push_recv", vp * 16), ("push_stride", i64),
        ("grads_lp", vp * 16), ("mc_grads_lp", vp), ("lp_begin", i64), ("lp_end", i64),
        ("dst_mask", C.c_uint),
        ("l2_blocks", vp), ("l2_slots", vp * 16), ("l2_row", i32), ("l2_rows_n", i32 * 4),
        ("trace", vp), ("err_src", vp), ("aux_mail", vp),
    ]


class GemmRS(C.Structure):
    """mtb_gemm_rs: where the reduce-scatter epilogue of mtb_gemm_bf16_rs adds its tiles."""
    _fields_ = [("world", i32), ("rank", i32), ("base", i64), ("per", i64), ("out", vp * 16), ("flags", i32)]


_SIGS = {
    "mtb_device_info": [i32, vp, vp, vp, vp],
    "mtb_arena_alloc": [C.c_size_t, C.POINTER(vp)],
    "mtb_arena_free": [vp],
    "mtb_ipc_export": [vp, vp],
    "mtb_ipc_import": [vp, C.POINTER(vp)],
    "mtb_ipc_close": [vp],
    "mtb_enable_peer_access": [i32],
    "mtb_gemm_bf16": [vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, vp, i32, f32, u64, vp, i32, i32,
                      vp, vp, vp],
    "mtb_gemm_bf16_rs": [vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, C.POINTER(GemmRS), vp],
    "mtb_gemm_tf32": [vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, vp, i32, f32, u64, vp, i32, i32, vp, vp, vp],
    "mtb_conv5x5_gemm": [i32, vp, vp, vp, i32, i32, i32, i32, vp, i32, i32, vp, vp, vp],
    "mtb_conv1_fwd_f32": [vp, vp, vp, vp, vp, i32, i32, vp],
    "mtb_conv1_wgrad_f32": [vp, vp, vp, vp, vp, vp, i32, i32, vp],
    "mtb_maxpool2x2_fwd_f32": [vp, vp, vp, i32, i32, i32, i32, vp],
    "mtb_unpool_relu_bwd_f32": [vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "mtb_im2col_5x5_f32": [vp, vp, i32, i32, i32, i32, vp],
    "mtb_col2im_5x5_f32": [vp, vp, i32, i32, i32, i32, vp],
    "mtb_fc2_softmax_ce_f32": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "mtb_fc2_dgrad_f32": [vp, vp, vp, vp, i32, i32, i32, f32, vp],
    "mtb_fc2_wgrad_f32": [vp, vp, vp, vp, i32, i32, i32, vp],
    "mtb_colsum_f32": [vp, vp, i64, i32, vp],
    "mtb_fc2_softmax_ce_fast_win_f32": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, i64, vp],
    "mtb_fc2_bwd_fused_f32": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, vp],
    "mtb_init_trunc_normal": [vp, i64, f32, u64, C.c_uint, vp],
    "mtb_fill_f32": [vp, i64, f32, vp],
    "mtb_zero_multi": [vp, vp, i32, vp],
    "mtb_normalize_u8_dev": [vp, vp, i64, vp],
    "mtb_synthetic_batch": [vp, vp, i64, i64, u64, i32, vp],
    "mtb_batch_window": [vp, vp, vp, i64, i32, vp, vp, vp],
    "mtb_cast_f32_to_bf16": [vp, vp, i64, vp],
    "mtb_cast_bf16_to_f32": [vp, vp, i64, vp],
    "mtb_l2_half_sumsq": [vp, i64, vp, vp],
    "mtb_conv1_fwd": [vp, vp, vp, vp, vp, i32, i32, vp],
    "mtb_conv1_wgrad": [vp, vp, vp, vp, vp, vp, i32, i32, vp],
    "mtb_conv1_fwd_tc": [vp, vp, vp, vp, vp, i32, vp],
    "mtb_conv1_wgrad_tc": [vp, vp, vp, vp, vp, vp, i32, vp],
    "mtb_conv1_fwd_tc_win": [vp, vp, vp, vp, vp, i32, vp, i64, vp],
    "mtb_conv1_fwd_tc_win_zero": [vp, vp, vp, vp, vp, i32, vp, i64, vp, vp, i32, vp],
    "mtb_conv1_wgrad_tc_win": [vp, vp, vp, vp, vp, vp, i32, vp, i64, vp],
    "mtb_fc2_softmax_ce_fast_win": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp, i64, vp],
    "mtb_maxpool2x2_fwd": [vp, vp, vp, i32, i32, i32, i32, vp],
    "mtb_unpool_relu_bwd": [vp, vp, vp, vp, i32, i32, i32, i32, vp],
    "mtb_im2col_5x5": [vp, vp, i32, i32, i32, i32, vp],
    "mtb_col2im_5x5": [vp, vp, i32, i32, i32, i32, vp],
    "mtb_fc2_softmax_ce": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "mtb_fc2_softmax_ce_fast": [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, vp],
    "mtb_fc2_bwd_fused": [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, f32, vp],
    "mtb_fc2_dgrad": [vp, vp, vp, vp, i32, i32, i32, f32, vp],
    "mtb_fc2_wgrad": [vp, vp, vp, vp, i32, i32, i32, vp],
    "mtb_colsum_bf16": [vp, vp, i64, i32, vp],
    "mtb_fused_allreduce_update": [C.POINTER(FusedArgs), vp],
    "mtb_p2p_barrier": [C.POINTER(FusedArgs), vp],
    "mtb_advance_step": [vp, vp],
    "mtb_debug_occupy": [i32, i32, i32, i64, vp],
    "mtb_debug_shift_mma": [vp, vp, vp, i32, i32, i32, i32, i32, vp],
    "mtb_debug_shift_mma32": [vp, vp, vp, i32, i32, i32, i32, vp],
    "mtb_conv2_fprop_pool": [vp, vp, vp, vp, vp, i32, vp],
    "mtb_conv2_dgrad": [vp, vp, vp, i32, vp],
    "mtb_conv2_wgrad": [vp, vp, vp, vp, vp, i32, vp],
    "mtb_conv2_dgrad_unpool": [vp, vp, vp, vp, vp, i32, vp],
    "mtb_conv2_wgrad_unpool": [vp, vp, vp, vp, vp, vp, vp, i32, vp],
}


def lib_path() -> Path:
    return Path(__file__).resolve().parents[1] / "lib" / "libmtb200.so"


def lib() -> C.CDLL:
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not path.exists():
        from . import build
        build.build_cuda()
    L = C.CDLL(str(path))
    for name, args in _SIGS.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = i32
    L.mtb_last_error.restype = C.c_char_p
    L.mtb_gemm_workspace_bytes.argtypes = [i32, i32, i32, i32]
    L.mtb_gemm_workspace_bytes.restype = C.c_size_t
    L.mtb_gemm_counter_bytes.argtypes = [i32, i32, i32]
    L.mtb_gemm_counter_bytes.restype = C.c_size_t
    L.mtb_conv2_supported.argtypes = [i32, i32, i32]
    L.mtb_conv2_supported.restype = i32
    L.mtb_conv2_wgrad_workspace_bytes.argtypes = [i32]
    L.mtb_conv2_wgrad_workspace_bytes.restype = C.c_size_t
    L.mtb_fused_max_blocks.restype = i32
    L.mtb_fused_args_size.restype = i32
    if L.mtb_fused_args_size() != C.sizeof(FusedArgs):
        raise ImportError(f"FusedArgs layout drifted from include/mtb200.h: {C.sizeof(FusedArgs)} vs {L.mtb_fused_args_size()} bytes")
    L.mtb_fused_max_l2_blocks.restype = i32
    L.mtb_fused_flag_bytes.restype = i32
    _LIB = L
    return L


class NativeError(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().mtb_last_error()
        raise NativeError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


_KERNELS_PER_CALL = {"mtb_conv2_wgrad": 2, "mtb_conv2_wgrad_unpool": 2}      # main kernel + partial-sum reduction
# (mtb_conv5x5_gemm in wgrad mode with splits > 1 also launches two; `call` counts it below)


def call(name: str, *args) -> None:
    """Call ``name(*args, current_stream)`` and raise on a non-zero return."""
    global LAUNCHES
    LAUNCHES += _KERNELS_PER_CALL.get(name, 1)
    if name == "mtb_conv5x5_gemm" and args[0] == 2 and args[10] > 1:
        LAUNCHES += 1
    check(getattr(lib(), name)(*args, stream_ptr()), name)


def ptr(t) -> Optional[int]:
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


class _RawCuda:
    """Zero-copy view of raw device memory through ``__cuda_array_interface__``."""

    def __init__(self, p: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (p, False), "version": 2}


def tensor_from_ptr(p: int, nbytes: int, device: torch.device) -> torch.Tensor:
    """uint8 tensor aliasing ``nbytes`` at device pointer ``p`` (arena / peer memory)."""
    return torch.as_tensor(_RawCuda(p, nbytes), device=device)


# ---------------------------------------------------------------------------
# typed wrappers
# ---------------------------------------------------------------------------
def gemm_bf16(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, a_mn: bool, b_mn: bool,
              bias: Optional[torch.Tensor] = None, relu: bool = False, keep_prob: float = 1.0, seed: int = 0,
              step_ptr: Optional[torch.Tensor] = None, bn: int = 0, splits: int = 1,
              workspace: Optional[torch.Tensor] = None, counters: Optional[torch.Tensor] = None) -> None:
    """out[M,N] = op(A) op(B).  K-major operand is stored [MN, K]; MN-major is stored [K, MN]."""
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    lda, ldb, ldc = A.stride(0), B.stride(0), out.stride(0)
    call("mtb_gemm_bf16", ptr(A), ptr(B), ptr(out), M, N, K, int(a_mn), int(b_mn), lda, ldb, ldc,
         int(out.dtype == torch.bfloat16), ptr(bias), int(relu), float(keep_prob), seed, ptr(step_ptr), bn, splits,
         ptr(workspace), ptr(counters))


def gemm_bf16_rs(A: torch.Tensor, B: torch.Tensor, M: int, N: int, K: int, a_mn: bool, b_mn: bool, ldc: int, bn: int,
                 rs: GemmRS) -> None:
    """op(A) op(B) added into the owners' gradient slices over peer memory (GEMM -> reduce-scatter)."""
    assert A.dtype == torch.bfloat16 and B.dtype == torch.bfloat16
    call("mtb_gemm_bf16_rs", ptr(A), ptr(B), M, N, K, int(a_mn), int(b_mn), A.stride(0), B.stride(0), ldc, bn,
         C.byref(rs))


def gemm_tf32(A: torch.Tensor, B: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, a_mn: bool, b_mn: bool,
              bias: Optional[torch.Tensor] = None, relu: bool = False, keep_prob: float = 1.0, seed: int = 0,
              step_ptr: Optional[torch.Tensor] = None, bn: int = 0, splits: int = 1,
              workspace: Optional[torch.Tensor] = None, counters: Optional[torch.Tensor] = None) -> None:
    """fp32 operands and output, TF32 tensor-core math; same conventions as :func:`gemm_bf16`."""
    assert A.dtype == torch.float32 and B.dtype == torch.float32 and out.dtype == torch.float32
    call("mtb_gemm_tf32", ptr(A), ptr(B), ptr(out), M, N, K, int(a_mn), int(b_mn), A.stride(0), B.stride(0),
         out.stride(0), ptr(bias), int(relu), float(keep_prob), seed, ptr(step_ptr), bn, splits, ptr(workspace),
         ptr(counters))


def gemm_scratch_sizes(M: int, N: int, bn: int, splits: int):
    L = lib()
    return int(L.mtb_gemm_workspace_bytes(M, N, bn, splits)), int(L.mtb_gemm_counter_bytes(M, N, bn))
