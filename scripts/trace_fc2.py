import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mpi_tensorflow_b200.ops import native as N
dev = torch.device("cuda", 0)
B, F = 1024, 1024
h = torch.relu(torch.randn(B, F)).to(torch.bfloat16).to(dev); W = (torch.randn(F, 10) * 0.1).to(dev)
dl = torch.randn(B, 10, device=dev); dh = torch.zeros(B, F, dtype=torch.bfloat16, device=dev)
dW = torch.zeros(F, 10, device=dev); db2 = torch.zeros(10, device=dev); db1 = torch.zeros(F, device=dev)
for _ in range(3):
    N.call("mtb_fc2_bwd_fused", h.data_ptr(), dl.data_ptr(), W.data_ptr(), dh.data_ptr(), dW.data_ptr(), db2.data_ptr(), db1.data_ptr(), B, F, 10, 1.0)
torch.cuda.synchronize()
buf = (C.c_longlong * 64)()
L = N.lib(); L.mtb_debug_trace2_read.argtypes = [C.c_void_p]; L.mtb_debug_trace2_read(buf)
print("fc2_bwd block 0 (cycles): entry->pdl %d, ->loads+sync %d, ->rows %d, ->red stores+sync+atomics %d, ->end %d" % tuple(buf[i+1]-buf[i] for i in range(5)))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50):
    N.call("mtb_fc2_bwd_fused", h.data_ptr(), dl.data_ptr(), W.data_ptr(), dh.data_ptr(), dW.data_ptr(), db2.data_ptr(), db1.data_ptr(), B, F, 10, 1.0)
e1.record(); torch.cuda.synchronize()
print("back-to-back avg us:", e0.elapsed_time(e1) / 50 * 1e3)
