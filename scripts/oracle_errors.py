"""Relative error of every gradient of one engine step against the bf16-emulating CPU oracle (what
tests/test_gpu_engine.py::test_one_step_grads_match_bf16_oracle bounds), to set that bound from measurements."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_tensorflow_b200 import TrainConfig, Trainer  # noqa: E402
from mpi_tensorflow_b200.models.oracle import Oracle, dropout_keep_mask  # noqa: E402
from mpi_tensorflow_b200.models.spec import get_spec  # noqa: E402
from mpi_tensorflow_b200.utils import data as D  # noqa: E402

for model, batch, dropout, dtype in [("tutorial512", 64, False, "bf16"), ("official1024", 100, True, "bf16"),
                                     ("official1024", 100, True, "fp32")]:
    for seed in (0, 1, 2):
        cfg = TrainConfig(model=model, batch=batch, eval_every=0, log_every=0, dropout=dropout, cuda_graph=False, dtype=dtype)
        tr = Trainer(cfg, device=torch.device("cuda", 0))
        u8, lab = D.synthetic_mnist(512, seed)
        x = D.normalize_u8(u8).reshape(512, 28, 28, 1)
        sh = D.Shards(x, lab.astype(np.int64), x[:128], lab[:128].astype(np.int64), x[:64], lab[:64].astype(np.int64))
        tr.load_data(sh)
        eng, spec = tr.engine, get_spec(model)
        p0 = {k: torch.from_numpy(v) for k, v in eng.get_params().items()}
        eng.compute_grads(None, None, 0)
        grads = eng.flat_grads.cpu()
        orc = Oracle(spec, params=p0, weight_decay=0.0, emulate_bf16=(dtype == "bf16"))
        mask = dropout_keep_mask(spec, batch, cfg.seed, 0) if dropout else None
        _, rg, _ = orc.loss_and_grads(torch.from_numpy(sh.train_x[:batch]), torch.from_numpy(sh.train_y[:batch]), mask)
        errs = {p.key: float((grads[p.offset:p.offset + p.numel] - rg[p.key].reshape(-1)).norm() / (rg[p.key].norm() + 1e-12))
                for p in spec.params}
        print(model, dtype, "seed", seed, " ".join(f"{k}={v:.2e}" for k, v in errs.items()), flush=True)
        tr.close()
