"""Explore the smem-descriptor base-offset semantics (see csrc/debug_shift.cu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_tensorflow_b200.ops import native as N

dev = torch.device("cuda", 0)
torch.manual_seed(0)
for mode, rb in ((0, 128), (0, 64), (1, 128)):
    rows = 256 if mode == 0 else 96
    if mode == 0:
        kel = rb // 2
        A = torch.randn(rows, kel).to(torch.bfloat16).to(dev)
        B = torch.randn(64, kel).to(torch.bfloat16).to(dev)
    else:
        A = torch.randn(rows, 128).to(torch.bfloat16).to(dev)     # [K rows][M]
        B = torch.randn(64, 32).to(torch.bfloat16).to(dev)        # [N][K]
    for shift in (0, 1, 2, 3, 5, 8, 9, 18, 37):
        res = []
        for bo in sorted({0, shift & 7, (shift >> 1) & 7, (-shift) & 7}):
            out = torch.zeros(128, 64, device=dev)
            N.call("mtb_debug_shift_mma", A.data_ptr(), B.data_ptr(), out.data_ptr(), mode, rb, rows, shift, bo)
            torch.cuda.synchronize()
            if mode == 0:
                ref = A[shift:shift + 128].float() @ B.float().t()
            else:
                ref = A[shift:shift + 32].float().t() @ B.float().t()
            err = float((out - ref).norm() / ref.norm())
            res.append(f"bo={bo}:{err:.2e}")
        print(f"mode={mode} rb={rb} shift={shift:2d}  " + "  ".join(res), flush=True)
