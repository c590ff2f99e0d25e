"""The allocation-free split-K GEMM entry point (``mtb_gemm_bf16_splitk_noalloc``).

Who needs it: code that drives ``libmtb200.so`` through its C ABI from a place where device
memory cannot be requested — a plug-in living inside a process whose allocator has already
taken the whole GPU, a loop that must stay free of allocator calls once started, or a plain C
host with nothing to allocate scratch with.  The regular entry point makes the caller supply
the split-K workspace and tile counters; this one keeps them in a fixed array inside the
library.

The test talks to the library through ctypes only (the same C ABI such a caller would use) and
checks, on two calls in a row (the second call meets the scratch the first one left behind),
that the result is the one ``mtb_gemm_bf16`` gives with caller-provided scratch; and that a
problem too large for the built-in array is refused instead of overrunning it.

Collected last on purpose (``zz``): a run that stops at the first failure has by then run
everything else.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

vp, i32, i64, f32, u64 = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_ulonglong


def _lib():
    from mpi_tensorflow_b200.ops import native
    L = C.CDLL(str(native.lib_path()))
    L.mtb_gemm_bf16_splitk_noalloc.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, vp, i32, i32,
                                               i32, vp]
    L.mtb_gemm_bf16_splitk_noalloc.restype = i32
    L.mtb_gemm_bf16.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, i64, i64, i64, i32, vp, i32, f32, u64, vp, i32,
                                i32, vp, vp, vp]
    L.mtb_gemm_bf16.restype = i32
    L.mtb_gemm_workspace_bytes.argtypes = [i32, i32, i32, i32]
    L.mtb_gemm_workspace_bytes.restype = C.c_size_t
    L.mtb_gemm_counter_bytes.argtypes = [i32, i32, i32]
    L.mtb_gemm_counter_bytes.restype = C.c_size_t
    L.mtb_gemm_noalloc_max_workspace_bytes.restype = C.c_size_t
    L.mtb_last_error.restype = C.c_char_p
    return L


@pytest.mark.parametrize("M,N,K,a_mn,b_mn,bn,splits,relu", [
    (800, 64, 6272, True, True, 64, 16, False),       # conv2 weight gradient as a GEMM (K = 32 images x 196 pixels)
    (300, 200, 1000, False, True, 64, 4, True),       # ragged tiles, bias + ReLU epilogue
])
def test_noalloc_splitk_matches_regular_entry_point(M, N, K, a_mn, b_mn, bn, splits, relu):
    L = _lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    g = torch.Generator(device="cpu").manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g).to(torch.bfloat16).to(dev)
    B = torch.randn(K, N, generator=g).to(torch.bfloat16).to(dev)
    bias = torch.randn(N, generator=g).to(dev) if relu else None
    A_st = A.t().contiguous() if a_mn else A.contiguous()
    B_st = B.contiguous() if b_mn else B.t().contiguous()
    lda, ldb = A_st.stride(0), B_st.stride(0)
    stream = torch.cuda.current_stream().cuda_stream
    bias_p = bias.data_ptr() if bias is not None else None

    # reference: the existing entry point with caller-allocated workspace and counters
    ws = torch.zeros(L.mtb_gemm_workspace_bytes(M, N, bn, splits), dtype=torch.uint8, device=dev)
    cnt = torch.zeros(L.mtb_gemm_counter_bytes(M, N, bn), dtype=torch.uint8, device=dev)
    ref = torch.zeros(M, N, device=dev)
    rc = L.mtb_gemm_bf16(A_st.data_ptr(), B_st.data_ptr(), ref.data_ptr(), M, N, K, int(a_mn), int(b_mn), lda, ldb, N,
                         0, bias_p, int(relu), 1.0, 0, None, bn, splits, ws.data_ptr(), cnt.data_ptr(), stream)
    assert rc == 0, L.mtb_last_error()
    torch.cuda.synchronize()
    expect = A.float() @ B.float()
    if relu:
        expect = torch.relu(expect + bias)
    assert float((ref - expect).norm() / expect.norm()) < 1e-4

    assert ws.numel() <= L.mtb_gemm_noalloc_max_workspace_bytes()
    outs = [torch.full((M, N), float("nan"), device=dev) for _ in range(2)]
    # the first call on a device resolves the scratch array's address (documented start-up step; the CUDA runtime
    # may load the kernel image lazily at that point, which is the driver's business, not an allocation of ours)
    rc = L.mtb_gemm_bf16_splitk_noalloc(A_st.data_ptr(), B_st.data_ptr(), outs[0].data_ptr(), M, N, K, int(a_mn),
                                        int(b_mn), lda, ldb, N, 0, bias_p, int(relu), bn, splits, stream)
    assert rc == 0, L.mtb_last_error()
    outs[0].fill_(float("nan"))
    torch.cuda.synchronize()
    # "allocates nothing" is checked, not assumed: neither the driver's free-memory figure nor the tensor
    # library's own books may move across the two calls (the outputs were allocated above, by the caller)
    free0, _ = torch.cuda.mem_get_info(dev)
    torch_alloc0, torch_res0 = torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)
    for call in range(2):                         # the second call starts from the first one's leftovers
        out = outs[call]
        rc = L.mtb_gemm_bf16_splitk_noalloc(A_st.data_ptr(), B_st.data_ptr(), out.data_ptr(), M, N, K, int(a_mn),
                                            int(b_mn), lda, ldb, N, 0, bias_p, int(relu), bn, splits, stream)
        assert rc == 0, L.mtb_last_error()
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 == free0, f"device memory moved by {free0 - free1} bytes across the allocation-free calls"
    assert (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev)) == (torch_alloc0, torch_res0)
    for call, out in enumerate(outs):
        assert torch.isfinite(out).all(), f"call {call}"
        assert float((out - ref).norm() / ref.norm()) < 1e-5, f"call {call}"


def test_noalloc_rejects_unsupported_tile_width():
    """bn = 256 exists in the regular entry point only; here it is an argument error, not a silent change of tile."""
    L = _lib()
    dev = torch.device("cuda", 0)
    A = torch.zeros(128, 128, dtype=torch.bfloat16, device=dev)
    B = torch.zeros(128, 128, dtype=torch.bfloat16, device=dev)
    out = torch.zeros(128, 128, device=dev)
    rc = L.mtb_gemm_bf16_splitk_noalloc(A.data_ptr(), B.data_ptr(), out.data_ptr(), 128, 128, 128, 0, 0, 128, 128, 128,
                                        0, None, 0, 256, 2, torch.cuda.current_stream().cuda_stream)
    assert rc == 2                                # MTB_E_ARG
    assert b"bn" in L.mtb_last_error()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs a second GPU")
def test_noalloc_on_a_second_device():
    """Every GPU that loads the library has its own copy of the array and its own kernel attributes."""
    L = _lib()
    dev = torch.device("cuda", 1)
    with torch.cuda.device(dev):
        M, N, K = 256, 128, 1024
        g = torch.Generator(device="cpu").manual_seed(7)
        A = torch.randn(M, K, generator=g).to(torch.bfloat16).to(dev)
        B = torch.randn(N, K, generator=g).to(torch.bfloat16).to(dev)
        out = torch.zeros(M, N, device=dev)
        for _ in range(2):
            rc = L.mtb_gemm_bf16_splitk_noalloc(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, 0, 0, K, K, N, 0,
                                                None, 0, 64, 4, torch.cuda.current_stream(dev).cuda_stream)
            assert rc == 0, L.mtb_last_error()
        torch.cuda.synchronize(dev)
        expect = A.float() @ B.float().t()
        assert float((out - expect).norm() / expect.norm()) < 1e-4


def test_noalloc_refuses_shapes_beyond_the_builtin_scratch():
    L = _lib()
    dev = torch.device("cuda", 0)
    M, N, K = 3136, 1024, 1024                   # 200 tiles x 4 splits x 64 KiB partials: far above 4 MiB
    A = torch.zeros(M, K, dtype=torch.bfloat16, device=dev)
    B = torch.zeros(N, K, dtype=torch.bfloat16, device=dev)
    out = torch.zeros(M, N, device=dev)
    rc = L.mtb_gemm_bf16_splitk_noalloc(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, 0, 0, K, K, N, 0, None, 0,
                                        128, 4, torch.cuda.current_stream().cuda_stream)
    assert rc == 4                                # MTB_E_LIMIT
    assert b"exceeds" in L.mtb_last_error()
    torch.cuda.synchronize()
