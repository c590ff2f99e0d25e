"""``python -m mpi_tensorflow_b200 [flags]`` — the ``main()`` of ``mpipy.py:201-247``.

Run under a launcher for more than one rank::

    python -m mpi_tensorflow_b200.parallel.launch -np 8 -m mpi_tensorflow_b200 --model official1024 --batch 1024
    torchrun --nproc-per-node 8 --master-addr 127.0.0.1 -m mpi_tensorflow_b200 ...
"""
from __future__ import annotations

import argparse
import json
import sys

from . import config as C
from .parallel import world
from .trainer import Trainer


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="mpi_tensorflow_b200")
    C.add_args(ap)
    ap.add_argument("--reference-semantics", action="store_true",
                    help="start from TrainConfig.reference(): the shipped script's literal behaviour")
    ns = ap.parse_args(argv)
    cfg = C.from_args(ns)
    if ns.reference_semantics:
        ref = C.TrainConfig.reference()
        for k, v in ref.to_dict().items():
            if getattr(cfg, k) == getattr(C.TrainConfig(), k):
                setattr(cfg, k, v)
    comm = world.init(cfg.backend, cfg.device, cfg.step_timeout_s)
    tr = Trainer(cfg, comm)
    try:
        res = tr.fit()
        if comm.rank == 0:
            print(json.dumps({"result": res}, default=float))
    finally:
        tr.close()
        world.finalize()
    return 0


if __name__ == "__main__":
    sys.exit(main())
