import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_tensorflow_b200.ops import native as N
dev = torch.device("cuda", 0)
B = 1024
x = torch.relu(torch.randn(B, 14, 14, 32)).to(torch.bfloat16).to(dev)
P = torch.relu(torch.randn(B, 7, 7, 64)).to(torch.bfloat16).to(dev); dflat = torch.randn(B, 7, 7, 64).to(torch.bfloat16).to(dev)
idx = torch.randint(0, 4, (B, 7, 7, 64), dtype=torch.uint8).to(dev)
dW = torch.zeros(800, 64, device=dev); db = torch.zeros(64, device=dev)
ws = torch.zeros(N.lib().mtb_conv2_wgrad_workspace_bytes(B), dtype=torch.uint8, device=dev)
for _ in range(3):
    N.call("mtb_conv2_wgrad_unpool", x.data_ptr(), dflat.data_ptr(), P.data_ptr(), idx.data_ptr(), dW.data_ptr(), db.data_ptr(), ws.data_ptr(), B)
torch.cuda.synchronize()
buf = (C.c_longlong * (64 * 8))()
L = N.lib(); L.mtb_debug_trace3_read.argtypes = [C.c_void_p]; L.mtb_debug_trace3_read(buf)
t0 = min(v for v in buf if v > 0)
print("it   mma:start  mma:p_full  mma:issued  bld:start  bld:p_empty  bld:done")
for it in range(22):
    print(f"{it:2d} " + " ".join(f"{(buf[it*8+c]-t0) if buf[it*8+c] else -1:11d}" for c in range(6)))
