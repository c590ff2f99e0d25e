"""Per-phase device times of one training step (NativeEngine.timed_step), bucketed K1 vs the single launch.

    python scripts/phases.py [--batch 1024] [--model official1024] [--dtype bf16]
Under torch.distributed.run it uses the p2p backend and reports the max over ranks.  Eager launches: the host
paces these steps, so read the phases relative to each other, not as the graph-replayed step time.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_tensorflow_b200 import TrainConfig, Trainer  # noqa: E402
from mpi_tensorflow_b200.ops import native as N  # noqa: E402
from mpi_tensorflow_b200.parallel import world  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--model", default="official1024")
    ap.add_argument("--dtype", default="bf16")
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    n = int(os.environ.get("WORLD_SIZE", "1"))
    comm = world.init("p2p" if n > 1 else "self", "cuda", timeout_s=120)
    dev = comm.device
    rows = 8 * a.batch
    x = torch.empty(rows, 784, device=dev)
    y = torch.empty(rows, dtype=torch.int32, device=dev)
    N.call("mtb_synthetic_batch", x.data_ptr(), y.data_ptr(), rows, comm.rank * rows, 1, 1)
    for overlap in (True, False):
        cfg = TrainConfig(model=a.model, batch=a.batch, dtype=a.dtype, eval_every=0, log_every=0, overlap=overlap,
                          cuda_graph=False)
        tr = Trainer(cfg, comm)
        tr.attach_device_dataset(x, y)
        for _ in range(5):
            tr.engine.timed_step()
        acc = {}
        for _ in range(a.reps):
            ph = tr.engine.timed_step()
            for k, v in ph.items():
                acc.setdefault(k, []).append(v)
        med = {k: sorted(v)[len(v) // 2] for k, v in acc.items()}
        if n > 1:
            allm = comm.allgather_object(med)
            med = {k: max(m[k] for m in allm) for k in med}
        if comm.rank == 0:
            print(json.dumps({"overlap": overlap, "n_gpus": n, **{k: round(v, 1) for k, v in med.items()}}), flush=True)
        tr.close()
    world.finalize()


if __name__ == "__main__":
    main()
