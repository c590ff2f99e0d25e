import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_tensorflow_b200.ops import native as N
dev = torch.device("cuda", 0)
B = 1024
x = (torch.rand(B, 784) - 0.5).to(dev)
P = torch.relu(torch.randn(B, 14, 14, 32)).to(torch.bfloat16).to(dev); dP = torch.randn(B, 14, 14, 32).to(torch.bfloat16).to(dev)
idx = torch.randint(0, 4, (B, 14, 14, 32), dtype=torch.uint8).to(dev)
dW = torch.zeros(25, 32, device=dev); db = torch.zeros(32, device=dev)
for _ in range(3):
    N.call("mtb_conv1_wgrad_tc", x.data_ptr(), dP.data_ptr(), P.data_ptr(), idx.data_ptr(), dW.data_ptr(), db.data_ptr(), B)
torch.cuda.synchronize()
buf = (C.c_longlong * (64 * 16))()
L = N.lib(); L.mtb_debug_trace_read.argtypes = [C.c_void_p]; L.mtb_debug_trace_read(buf)
rows = range(16, 24)
t0 = min(buf[r*16+c] for r in rows for c in range(9) if buf[r*16+c] > 0)
names = ["mma:start", "mma:x_full", "mma:issued", "-", "prod:start", "prod:x_empty", "prod:xs", "prod:dA", "prod:done"]
print("it  " + " ".join(f"{n:>12s}" for n in names))
for r in rows:
    print(f"{r-16:2d}  " + " ".join(f"{(buf[r*16+c]-t0) if buf[r*16+c] else -1:12d}" for c in range(9)))

print("kernel: " + " ".join(f"{(buf[30*16+c]-t0):d}" for c in range(6)), "(entry, prologue done, pdl done, a_full, atomics done, end)")
