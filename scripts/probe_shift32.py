import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_tensorflow_b200.ops import native as N
dev = torch.device("cuda", 0)
torch.manual_seed(0)
rows = 300
A = torch.randn(rows, 16).to(torch.bfloat16).to(dev)
for shift in (0, 1, 3, 4, 7, 12, 33):
    B = torch.randn(32, 16).to(torch.bfloat16).to(dev)
    out = torch.zeros(128, 32, device=dev)
    N.call("mtb_debug_shift_mma32", A.data_ptr(), B.data_ptr(), out.data_ptr(), 2, rows, shift, 0)
    torch.cuda.synchronize()
    ref = A[shift:shift + 128].float() @ B.float().t()
    print("mode2 shift", shift, float((out - ref).norm() / ref.norm()))
for shift in (0, 1, 5, 16, 37):
    for ar in (1, 4):
        B = torch.randn(16, 32).to(torch.bfloat16).to(dev)
        out = torch.zeros(128, 32, device=dev)
        N.call("mtb_debug_shift_mma32", A.data_ptr(), B.data_ptr(), out.data_ptr(), 3, rows, shift, ar)
        torch.cuda.synchronize()
        ref = torch.zeros(128, 32)
        Af, Bf = A.float().cpu(), B.float().cpu()
        for a in range(8):
            ref[a * 16:(a + 1) * 16] = Af[shift + a * ar: shift + a * ar + 16].t() @ Bf
        print("mode3 shift", shift, "atom_rows", ar, float((out.cpu() - ref).norm() / ref.norm()))
