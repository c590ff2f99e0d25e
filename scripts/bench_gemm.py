"""Time the tcgen05 GEMM engine on the FC shapes against cuBLAS (torch.matmul), L2-warm, CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_tensorflow_b200.ops import native as N
dev = torch.device("cuda", 0)

def timeit(fn, iters=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
shapes = [("fc1 fprop", B, 1024, 3136, False, True, 64), ("fc1 fprop bn128", B, 1024, 3136, False, True, 128),
          ("fc1 dgrad", B, 3136, 1024, False, False, 128), ("fc1 dgrad bn256", B, 3136, 1024, False, False, 256),
          ("fc1 wgrad", 3136, 1024, B, True, True, 128), ("fc1 wgrad bn256", 3136, 1024, B, True, True, 256), ("fc1 fprop bn256", B, 1024, 3136, False, True, 256),
          ("fc1 dgrad bn128x2", B, 3136, 1024, False, False, 129), ("fc1 wgrad bn128x2", 3136, 1024, B, True, True, 129)]
for name, M, Nn, K, a_mn, b_mn, bn in shapes:
    A = torch.randn(M, K, device=dev).to(torch.bfloat16)
    Bm = torch.randn(K, Nn, device=dev).to(torch.bfloat16)
    A_st = A.t().contiguous() if a_mn else A
    B_st = Bm if b_mn else Bm.t().contiguous()
    out = torch.zeros(M, Nn, device=dev, dtype=torch.bfloat16 if "wgrad" not in name else torch.float32)
    g = torch.cuda.CUDAGraph()
    N.gemm_bf16(A_st, B_st, out, M, Nn, K, a_mn, b_mn, bn=bn)
    torch.cuda.synchronize()
    with torch.cuda.graph(g):
        for _ in range(10):
            N.gemm_bf16(A_st, B_st, out, M, Nn, K, a_mn, b_mn, bn=bn)
    us = timeit(lambda: g.replay(), iters=20) / 10
    ref = timeit(lambda: torch.matmul(A, Bm))
    fl = 2.0 * M * Nn * K
    print(f"{name:18s} M={M} N={Nn} K={K}: ours {us:7.1f} us ({fl/us/1e6:6.0f} TFLOP/s)   cuBLAS {ref:7.1f} us ({fl/ref/1e6:6.0f} TFLOP/s)", flush=True)
