"""Does this box expose NVLS multicast through torch's symmetric memory?  (run under the launcher, 2+ GPUs)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
torch.cuda.set_device(dev)
backend = sys.argv[1] if len(sys.argv) > 1 else "gloo"
dist.init_process_group(backend, rank=rank, world_size=world, device_id=dev if backend == "nccl" else None)
import torch.distributed._symmetric_memory as symm
try:
    t = symm.empty(1 << 20, dtype=torch.float32, device=dev)
    h = symm.rendezvous(t, group=dist.group.WORLD)
    print(rank, "backend", backend, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "multicast_ptr", hex(h.multicast_ptr),
          "signal_pads", len(h.signal_pad_ptrs), flush=True)
    t.fill_(rank + 1)
    torch.cuda.synchronize(); dist.barrier()
    peer = h.get_buffer((rank + 1) % world, (8,), torch.float32)
    print(rank, "peer view", peer[:2].tolist(), flush=True)
except Exception as e:
    import traceback; traceback.print_exc()
    print(rank, "FAILED", type(e).__name__, str(e)[:300], flush=True)
os._exit(0)
