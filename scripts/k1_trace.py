"""Where the bucketed K1 sits in the step (MTB_K1_TRACE=1): %globaltimer stamps written by the hidden launch and by the
closing launch, read after graph replays.  Prints, per rank, microseconds relative to the hidden launch's start:
hidden end, tail start, tail end, and the step period.   torchrun ... scripts/k1_trace.py"""
import json
import os
import sys

os.environ["MTB_K1_TRACE"] = "1"
import torch  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_tensorflow_b200 import TrainConfig, Trainer  # noqa: E402
from mpi_tensorflow_b200.ops import native as N  # noqa: E402
from mpi_tensorflow_b200.parallel import world  # noqa: E402

n = int(os.environ.get("WORLD_SIZE", "1"))
comm = world.init("p2p" if n > 1 else "self", "cuda", timeout_s=120)
dev = comm.device
B = 1024
rows = 16 * B
x = torch.empty(rows, 784, device=dev)
y = torch.empty(rows, dtype=torch.int32, device=dev)
N.call("mtb_synthetic_batch", x.data_ptr(), y.data_ptr(), rows, comm.rank * rows, 1, 1)
tr = Trainer(TrainConfig(model="official1024", batch=B, eval_every=0, log_every=0), comm)
tr.attach_device_dataset(x, y)
for _ in range(20):
    tr.step(read_loss=False)
torch.cuda.synchronize()
recs = []
prev_tail_end = None
for i in range(12):
    tr.step(read_loss=False)
    torch.cuda.synchronize()
    t = tr.engine._trace.cpu().numpy().astype("int64")
    h0, h1, c0, c1, t0, t1 = (int(v) for v in t.reshape(-1))
    rec = {"hidden_us": (h1 - h0) / 1e3, "tail_start_after_hidden_end_us": (t0 - h1) / 1e3,
           "tail_us": (t1 - t0) / 1e3, "hidden_start_to_tail_end_us": (t1 - h0) / 1e3}
    if c0 > h0:                                   # a hidden conv2 bucket ran this step
        rec.update(conv2_bucket_start_after_hidden_start_us=(c0 - h0) / 1e3, conv2_bucket_us=(c1 - c0) / 1e3,
                   tail_start_after_conv2_bucket_end_us=(t0 - c1) / 1e3)
    recs.append(rec)
med = {k: sorted(r[k] for r in recs)[len(recs) // 2] for k in recs[0]}
allm = comm.allgather_object(med) if n > 1 else [med]
if comm.rank == 0:
    for r, m in enumerate(allm):
        print(json.dumps({"rank": r, **{k: round(v, 1) for k, v in m.items()}}), flush=True)
tr.close()
world.finalize()
