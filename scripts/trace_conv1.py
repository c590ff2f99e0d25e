import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_tensorflow_b200.ops import native as N
dev = torch.device("cuda", 0)
B = 1024
x = (torch.rand(B, 784) - 0.5).to(dev); w = (torch.randn(25, 32) * 0.1).to(dev); bias = torch.zeros(32, device=dev)
out = torch.zeros(B, 14, 14, 32, dtype=torch.bfloat16, device=dev); idx = torch.zeros(B, 14, 14, 32, dtype=torch.uint8, device=dev)
for _ in range(3):
    N.call("mtb_conv1_fwd_tc", x.data_ptr(), w.data_ptr(), bias.data_ptr(), out.data_ptr(), idx.data_ptr(), B)
torch.cuda.synchronize()
buf = (C.c_longlong * (64 * 16))()
L = N.lib(); L.mtb_debug_trace_read.argtypes = [C.c_void_p]; L.mtb_debug_trace_read(buf)
t0 = min(v for v in buf if v > 0)
names = ["mma:start", "mma:x_full", "mma:a_empty", "mma:issued", "prod:start", "prod:x_empty", "prod:xs_done", "prod:x5_done", "epi:start", "epi:a_full", "epi:done"]
print("it  " + " ".join(f"{n:>13s}" for n in names))
for it in range(8):
    print(f"{it:2d}  " + " ".join(f"{(buf[it*16+c]-t0) if buf[it*16+c] else -1:13d}" for c in range(11)))
