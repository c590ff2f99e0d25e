"""mpi_tensorflow_b200 — B200-native MPI-style data-parallel trainer.

Capabilities of ``youzhenfei1995/mpi-Tensorflow`` (``mpipy.py``), redesigned for
8xB200: rank/world setup, dataset scatter, the MNIST CNN, momentum-SGD with
exponential LR decay, periodic weight averaging *and* per-step gradient
allreduce fused with the optimizer update over NVLink peer memory.
"""
from .config import TrainConfig
from .models.spec import PRESETS, ModelSpec, get_spec
from .parallel.comm import (CommAborted, CommError, CommTimeout, Communicator, FakeWorld,
                            SelfComm)
from .parallel.world import finalize, init, world
from .trainer import Trainer

__version__ = "0.1.0"
__all__ = ["TrainConfig", "Trainer", "init", "finalize", "world", "Communicator", "FakeWorld",
           "SelfComm", "CommError", "CommTimeout", "CommAborted", "ModelSpec", "PRESETS", "get_spec"]
