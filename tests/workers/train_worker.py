"""Per-rank training body: BASELINE.json config 1 (2 ranks on CPU, full plumbing)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mpi_tensorflow_b200 import TrainConfig, Trainer  # noqa: E402
from mpi_tensorflow_b200.parallel import world  # noqa: E402


def main():
    backend, model, steps = sys.argv[1], sys.argv[2], int(sys.argv[3])
    outdir = sys.argv[4]
    comm = world.init(backend, "cpu", timeout_s=30)
    cfg = TrainConfig(model=model, batch=16, max_steps=steps, device="cpu", backend=backend,
                      eval_every=0, log_every=5, step_timeout_s=20, dropout=False,
                      checkpoint_dir=os.path.join(outdir, "ckpt") if comm.rank == 0 or True else None)
    tr = Trainer(cfg, comm)
    tr.load_data()
    res = tr.fit()
    torch.save(tr.engine.flat_params.clone(), os.path.join(outdir, f"params_{comm.rank}.pt"))
    if comm.rank == 0:
        with open(os.path.join(outdir, "result.json"), "w") as f:
            json.dump({k: v for k, v in res.items() if k != "phases_ms"}, f)
    print(f"TRAIN_OK {comm.rank}", flush=True)
    tr.close()
    world.finalize()


if __name__ == "__main__":
    main()
