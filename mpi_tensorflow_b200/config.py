"""Run configuration: one dataclass + an argparse front end.

The reference's "config system" is five module constants (``mpipy.py:17-21``)
plus hyper-parameters hard-coded inline: lr 0.01 / decay 0.95 (``:60-63``),
momentum 0.9 (``:65``), L2 5e-4 (``:57``), dropout keep 0.5 (``:166``), sync period 50
(``:87``), seed 1 (``:38-52``), epochs 2 (``:18``), batch 64 (``:20``).  Every one of
those is a field here, with the reference's value as the default of the
``tutorial512`` preset, and every ``BASELINE.json`` config is expressible.
"""
from __future__ import annotations

import argparse
from dataclasses import asdict, dataclass, fields
from typing import Optional


@dataclass
class TrainConfig:
    # model
    model: str = "tutorial512"          # tutorial512 | official1024 | wide
    dtype: str = "bf16"                 # bf16: bf16 tensor-core operands, fp32 accumulate / master / optimizer.
                                        # fp32: fp32 storage everywhere, TF32 tensor-core GEMMs, fp32 CUDA-core conv1/head.
    seed: int = 1                       # mpipy.py:40
    # data
    data: str = "synthetic"             # synthetic | mnist
    data_dir: str = "./data"
    reference_split_bug: bool = False   # reproduce mpipy.py:221-vs-230 zero tail
    sampler: str = "sequential"         # sequential (mpipy.py:80-82) | shuffle
    # optimisation
    batch: int = 64                     # per-rank, mpipy.py:20
    epochs: int = 2                     # mpipy.py:18 ``iteration``
    max_steps: int = 0                  # >0 overrides epochs
    optimizer: str = "momentum"         # momentum | sgd | adam
    lr: float = 0.01                    # mpipy.py:60
    lr_decay: float = 0.95              # mpipy.py:63
    lr_decay_rows: int = 0              # 0 => per-rank shard rows (mpipy.py:62)
    momentum: float = 0.9               # mpipy.py:65
    beta1: float = 0.9
    beta2: float = 0.999
    adam_eps: float = 1e-8
    weight_decay: float = 5e-4          # mpipy.py:57 (FC weights+biases only)
    dropout: bool = True                # mpipy.py:166
    # parallelism / sync
    sync: str = "grad_allreduce"        # grad_allreduce (every step) | weight_avg (every K) | none
    sync_period: int = 50               # mpipy.py:87
    sync_include_bias: bool = True      # reference: False (bias lines commented, mpipy.py:122-128)
    sync_include_momentum: bool = False
    sync_root_only: bool = False        # True reproduces mpipy.py:129 (mean stays on root)
    comm_algo: str = "auto"             # auto | oneshot | twoshot | nvls | push (closing bucket only) | nccl (baseline)
    fuse_wgrad_rs: bool = False         # FC1 weight gradient: GEMM -> reduce-scatter in one kernel (tiles added into the
                                        # owner rank's slice over NVLink); K1 then skips that range's peer reduction
    comm_dtype: str = "fp32"            # gradient wire format: fp32, or bf16 for the FC1 weight gradient (98 % of the
                                        # bytes; fp32 accumulation in K1 / in the switch, needs dtype="bf16")
    shard_master: bool = True           # two-shot / nvls: the big weights' fp32 master stays with the slice owner,
                                        # only their bf16 shadow is broadcast (gathered on demand for checkpoints)
    backend: str = "auto"               # auto | p2p | nccl | gloo | shm | self
    device: str = "auto"                # auto | cuda | cpu
    overlap: bool = True                # bucketed K1: the fc bucket's allreduce+update (narrow blocks, comm stream) hides
                                        # under the conv backward; a few-KB "push" bucket with one flag round closes the step
    # eval / logging
    eval_every: int = 50                # reference: 1 (mpipy.py:86)
    eval_dropout: bool = False          # reference: True (mpipy.py:166 via :68)
    eval_batch: int = 64                # mpipy.py:170
    log_every: int = 50                 # mpipy.py:87
    log_jsonl: Optional[str] = None
    log_all_ranks: bool = False
    profile_phases: bool = False        # at every log line, run one step eagerly with CUDA events between its phases
                                        # (fwd / bwd / allreduce+update) and add their times and the bus GB/s to the line
    # runtime
    conv_impl: str = "auto"             # auto | implicit (shift-GEMM, no im2col) | im2col (+ GEMM)
    cuda_graph: bool = True
    checkpoint_dir: Optional[str] = None
    checkpoint_every: int = 0
    checkpoint_names: str = "tf1"       # tf1 (Variable_N) | keras (conv2d/kernel ...)
    resume: Optional[str] = None
    step_timeout_s: float = 120.0

    def to_dict(self):
        return asdict(self)

    @staticmethod
    def reference() -> "TrainConfig":
        """The shipped script, literally: weights-only root-only average every 50
        steps, eval every step with dropout on, zero-tail split."""
        return TrainConfig(model="tutorial512", dtype="fp32", sync="weight_avg", sync_period=50,
                           sync_include_bias=False, sync_root_only=True, eval_every=1,
                           eval_dropout=True, reference_split_bug=True)


def add_args(ap: argparse.ArgumentParser) -> None:
    for f in fields(TrainConfig):
        name = "--" + f.name.replace("_", "-")
        if f.type in ("bool", bool):
            ap.add_argument(name, action=argparse.BooleanOptionalAction, default=f.default)
        elif f.type in ("int", int):
            ap.add_argument(name, type=int, default=f.default)
        elif f.type in ("float", float):
            ap.add_argument(name, type=float, default=f.default)
        else:
            ap.add_argument(name, type=str, default=f.default)


def from_args(ns: argparse.Namespace) -> TrainConfig:
    return TrainConfig(**{f.name: getattr(ns, f.name) for f in fields(TrainConfig)})
