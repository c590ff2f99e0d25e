"""Training driver: ``Cnn.__init__`` + ``run_process`` + ``eval_in_batches`` of the
reference (``mpipy.py:23-93``, ``:169-183``) and the data staging of ``main``
(``mpipy.py:201-244``), as a reusable class instead of a constructor that *is*
the program (SURVEY §1 "Control-flow oddity").

Public API a user of the reference switches to::

    from mpi_tensorflow_b200 import TrainConfig, Trainer, init
    comm = init()                       # MPI.COMM_WORLD              mpipy.py:208
    tr = Trainer(TrainConfig(model="official1024", batch=1024), comm)
    tr.load_data()                      # parse + 6x Scatter          mpipy.py:214-241
    tr.fit()                            # run_process                 mpipy.py:76-93
    loss = tr.step(images, labels)      # or drive it yourself, host batches in
"""
from __future__ import annotations

import os
import time
from typing import Optional

import numpy as np
import torch

from .config import TrainConfig
from .models.oracle import error_rate, learning_rate
from .models.spec import get_spec
from .parallel.comm import CommError, Communicator, SelfComm
from .parallel.sync import check_replicas_equal, make_sync
from .utils import checkpoint as ckpt
from .utils import data as D
from .utils.faults import FaultSpec, Watchdog, maybe_inject
from .utils.logging import RunLogger
from .utils.timing import PhaseTimer


def make_engine(cfg: TrainConfig, comm: Communicator, device: torch.device):
    spec = get_spec(cfg.model)
    if device.type == "cuda":
        from .engine_native import NativeEngine
        return NativeEngine(spec, cfg, comm, device, conv_impl=cfg.conv_impl)
    from .engine_cpu import CpuEngine
    return CpuEngine(spec, cfg)


class Trainer:
    def __init__(self, cfg: TrainConfig, comm: Optional[Communicator] = None,
                 device: Optional[torch.device] = None, engine=None):
        self.cfg = cfg
        self.comm = comm or SelfComm()
        if device is None:
            device = getattr(self.comm, "device", None) or torch.device(
                "cuda" if (cfg.device != "cpu" and torch.cuda.is_available()) else "cpu")
        self.device = device
        self.spec = get_spec(cfg.model)
        self.engine = engine or make_engine(cfg, self.comm, device)
        self.sync = make_sync(cfg, self.comm, self.spec)
        self.log = RunLogger(self.comm.rank, self.comm.size, cfg.log_jsonl, cfg.log_all_ranks)
        self.timer = PhaseTimer(device.type == "cuda", enabled=cfg.profile_phases)
        self.global_step = 0
        self.shards: Optional[D.Shards] = None
        self.sampler = None
        self.decay_rows = cfg.lr_decay_rows
        self._fault = FaultSpec.from_env()
        # The GPU engine runs every strategy through its fused, graph-captured step: with the per-step gradient
        # allreduce inside K1 (sync="grad_allreduce"), or as LOCAL steps (sync="weight_avg" between two averagings,
        # sync="none") - the cross-GPU reduction happens only when the strategy asks for it.
        self._fused = getattr(self.engine, "can_fuse", False)
        self._peers = cfg.sync == "grad_allreduce"
        if cfg.resume:
            self.load_checkpoint(cfg.resume)

    # ------------------------------------------------------------------ data
    def load_data(self, shards: Optional[D.Shards] = None) -> D.Shards:
        """Rank 0 parses/synthesises and splits; six scatters hand every rank its
        contiguous shard (``mpipy.py:214-241``).  Labels travel as int64."""
        c, comm = self.cfg, self.comm
        sizes = D.split_sizes(comm.size)
        if shards is None:
            # Parsing or synthesising 60,000 images on rank 0 takes tens of seconds on a slow host; the others wait in
            # the first scatter, so the per-step failure-detection timeout must not apply to this phase.
            if hasattr(comm, "set_timeout"):
                comm.set_timeout(max(600.0, float(c.step_timeout_s)))
            root = D.load_root_dataset(c.data, sizes, c.data_dir, c.seed, c.reference_split_bug) \
                if comm.rank == 0 else None
            per = dict(train=sizes.train_rows_per_rank, test=sizes.test_rows_per_rank,
                       val=sizes.val_rows_per_rank)
            out = {}
            for split in ("train", "test", "val"):
                rx = torch.zeros(per[split], 28, 28, 1, dtype=torch.float32)          # mpipy.py:230
                ry = torch.zeros(per[split], dtype=torch.int64)                        # mpipy.py:231
                sx = torch.from_numpy(np.ascontiguousarray(getattr(root, f"{split}_x"))) if root else None
                sy = torch.from_numpy(np.ascontiguousarray(getattr(root, f"{split}_y"))) if root else None
                comm.scatter(sx, rx, root=0)                                           # mpipy.py:236,238,240
                comm.scatter(sy, ry, root=0)                                           # mpipy.py:237,239,241
                out[split] = (rx.numpy(), ry.numpy())
            shards = D.Shards(out["train"][0], out["train"][1], out["test"][0], out["test"][1],
                              out["val"][0], out["val"][1])
            if hasattr(comm, "set_timeout"):
                comm.set_timeout(float(c.step_timeout_s))
        self.shards = shards
        rows = shards.train_x.shape[0]
        self.sampler = (D.SequentialSampler if c.sampler == "sequential" else D.ShuffleSampler)(rows, c.batch)
        if not self.decay_rows:
            self.decay_rows = rows                                                     # mpipy.py:62
        if hasattr(self.engine, "attach_dataset"):
            self.engine.attach_dataset(shards)
        return shards

    def attach_device_dataset(self, images: torch.Tensor, labels: torch.Tensor) -> None:
        """Use a shard that already lives in device memory (``[rows,784]`` fp32 + int32 labels):
        ``step()`` without arguments then walks it with the reference's sequential window
        (``mpipy.py:80-82``) entirely on the device."""
        self.engine.attach_device_dataset(images, labels)
        self.sampler = D.SequentialSampler(images.shape[0], self.cfg.batch)
        if not self.decay_rows:
            self.decay_rows = images.shape[0]

    # ------------------------------------------------------------------ step
    def lr_at(self, step: int) -> float:
        c = self.cfg
        if c.lr_decay == 1.0 or not self.decay_rows:
            return c.lr
        return learning_rate(step, c.batch, self.decay_rows, c.lr, c.lr_decay)          # mpipy.py:60-64

    def step(self, images=None, labels=None, read_loss: bool = True) -> float:
        """One optimizer step.  With ``images``/``labels`` (host or device
        tensors, NHWC float and int labels) the batch is copied to the device as
        needed; without, the next window of the resident shard is used
        (``mpipy.py:80-85``).  Returns the loss of this step as a float; with
        ``read_loss=False`` the fused GPU path skips the device-to-host read (and
        the host sync it implies) and returns 0.0."""
        step = self.global_step
        maybe_inject(self._fault, self.comm.rank, step)
        lr = self.lr_at(step)
        if images is None:
            if self.sampler is None:
                raise RuntimeError("no data: call load_data() or pass a batch")
            if hasattr(self.engine, "resident_batch") and isinstance(self.sampler, D.SequentialSampler):
                # the GPU engine walks the reference's sequential window over its resident shard on the device
                images, labels = self.engine.resident_batch(self.sampler, step)
            else:
                # any other sampler (sampler="shuffle") picks its rows on the host and feeds them like a caller's batch
                if self.shards is None:
                    raise RuntimeError(f"sampler={self.cfg.sampler!r} needs the host shards (load_data()); a dataset "
                                       f"attached on the device is walked with the sequential window only")
                idx = self.sampler.indices(step)
                images, labels = self.shards.train_x[idx], self.shards.train_y[idx]
        if self._fused:
            with self.timer.phase("step"):
                loss = self.engine.fused_step(images, labels, lr, step, read_loss=read_loss, peers=self._peers)
            if not self._peers:
                with self.timer.phase("weight_sync"):
                    self.sync.after_step(step, self.engine.flat_params, self.engine.slots, engine=self.engine)
        else:
            with self.timer.phase("fwd_bwd"):
                loss = self.engine.compute_grads(images, labels, step)
            with self.timer.phase("allreduce"):
                self.sync.reduce_grads(self.engine.flat_grads)
            with self.timer.phase("update"):
                self.engine.apply(lr)
            with self.timer.phase("weight_sync"):
                if self.sync.after_step(step, self.engine.flat_params, self.engine.slots):
                    if hasattr(self.engine, "params_changed"):
                        self.engine.params_changed()
        self.global_step += 1
        return float(loss)

    def fit_stream(self, images, labels, steps: int):
        """Train ``steps`` steps from a dataset kept in host memory, one host->device batch copy
        and one device->host loss read per step (the reference's feed loop, ``mpipy.py:79-85``),
        with copies overlapped with compute on the GPU engine.  ``images`` [rows,28,28,1] or
        [rows,784], float32 (already normalised) or uint8 (raw pixels as the IDX files hold them:
        a quarter of the bytes per copy, normalised by the first kernel of the step as
        ``extract_data`` does on the host); ``labels`` [rows] integer.  They are pinned here if
        they are not already.  Returns the list of per-step losses."""
        x = torch.as_tensor(images)
        if x.dtype != torch.uint8:                   # uint8 = raw pixels: normalised on the device by the GPU engine
            x = x.to(torch.float32)
        y = torch.as_tensor(labels).to(torch.int32)
        if not self._fused or not hasattr(self.engine, "stream_steps") or self.cfg.sync == "weight_avg":
            B, rows = self.cfg.batch, x.shape[0]
            if x.dtype == torch.uint8:
                x = (x.to(torch.float32) - 127.5) / 255.0                       # extract_data, mpipy.py:215
            out = []
            for i in range(steps):
                o = (self.global_step * B) % (rows - B) if rows > B else 0       # mpipy.py:80-82, global step
                out.append(self.step(x[o:o + B], y[o:o + B]))
            return out
        x = x if x.is_pinned() else x.pin_memory()
        y = y if y.is_pinned() else y.pin_memory()
        if not self.decay_rows:
            self.decay_rows = x.shape[0]
            self.engine.decay_rows = self.engine.decay_rows or x.shape[0]
        losses = self.engine.stream_steps(x, y, steps, peers=self._peers)
        self.global_step += steps
        return losses

    # ------------------------------------------------------------------ eval
    def eval_in_batches(self, data, dropout_step: Optional[int] = None) -> np.ndarray:
        """Softmax predictions ``[size, classes]`` in fixed batches; the ragged tail
        is served by re-running the last ``B`` rows and keeping the needed suffix
        (``mpipy.py:169-183``)."""
        B = self.cfg.eval_batch
        size = data.shape[0]
        if size < B:
            raise ValueError(f"batch size for evals larger than dataset: {size}")          # mpipy.py:171-172
        if hasattr(self.engine, "predict_all"):
            return self.engine.predict_all(data, B, dropout_step)
        out = np.empty((size, self.spec.num_classes), dtype=np.float32)
        for begin in range(0, size, B):
            end = begin + B
            if end <= size:
                out[begin:end] = self.engine.predict(data[begin:end], dropout_step).cpu().numpy()
            else:
                tail = self.engine.predict(data[-B:], dropout_step).cpu().numpy()
                out[begin:] = tail[begin - size:]
        return out

    def evaluate(self, split: str = "test") -> float:
        x = getattr(self.shards, f"{split}_x")
        y = getattr(self.shards, f"{split}_y")
        ds = self.global_step if self.cfg.eval_dropout else None
        return error_rate(self.eval_in_batches(x, ds), y)                                   # mpipy.py:86

    # ------------------------------------------------------------------ loop
    def total_steps(self) -> int:
        c = self.cfg
        if c.max_steps:
            return c.max_steps
        return c.epochs * self.shards.train_y.shape[0] // c.batch                          # mpipy.py:79

    def fit(self, steps: Optional[int] = None) -> dict:
        c = self.cfg
        if self.shards is None:
            self.load_data()
        if not check_replicas_equal(self.comm, self.engine.flat_params):
            raise RuntimeError("replicas differ before training (seeded init broken?)")
        steps = steps or self.total_steps()
        self.log.banner()                                                                   # mpipy.py:77
        wd = Watchdog(c.step_timeout_s, self.comm.abort) if self.comm.size > 1 else None
        t0 = time.time()
        last_loss, test_err = float("nan"), float("nan")
        imgs_at_log, t_at_log = 0, t0
        phases = {}
        try:
            while self.global_step < steps:
                step = self.global_step
                log_now = bool(c.log_every) and step > 0 and step % c.log_every == 0
                if log_now and c.profile_phases and self._fused and hasattr(self.engine, "timed_step"):
                    # this step runs eagerly with events between its phases (every rank takes the same branch)
                    phases = self.engine.timed_step(peers=self._peers)
                    last_loss = self.engine._await_loss(self.engine.opt_step)
                    if not self._peers:
                        self.sync.after_step(step, self.engine.flat_params, self.engine.slots, engine=self.engine)
                    self.global_step += 1
                else:
                    last_loss = self.step()
                if wd:
                    wd.kick()
                if c.eval_every and (step % c.eval_every == 0) and (c.eval_every == 1 or step > 0):
                    test_err = self.evaluate("test")
                if log_now:                                                                # mpipy.py:87
                    self.engine.synchronize()
                    now = time.time()
                    ips = (step * c.batch - imgs_at_log) / max(1e-9, now - t_at_log)
                    imgs_at_log, t_at_log = step * c.batch, now
                    rec = dict(step=step, loss=last_loss, lr=self.lr_at(step), test_error=test_err,
                               img_per_s=ips * self.comm.size)
                    if hasattr(self.engine, "train_errors") and self._fused:               # mpipy.py:89, kept
                        rec["train_error"] = 100.0 * self.engine.train_errors(self.engine.opt_step) / c.batch
                    if phases:
                        rec.update(fwd_us=phases["fwd"], bwd_us=phases.get("bwd", 0.0) + phases.get("bwd_fc", 0.0)
                                   + phases.get("bwd_conv", 0.0), allreduce_us=phases["k1_hidden"] + phases["k1_tail"],
                                   allreduce_exposed_us=phases["k1_tail"], bus_gbs=phases["bus_gbs"])
                    self.log.log(**rec)
                if c.checkpoint_dir and c.checkpoint_every and step > 0 and step % c.checkpoint_every == 0:
                    self.save_checkpoint(os.path.join(c.checkpoint_dir, f"model.ckpt-{step}"))
        except CommError:
            self.log.log(event="comm_failure", step=self.global_step)
            raise
        finally:
            if wd:
                wd.stop()
        self.engine.synchronize()
        dt = time.time() - t0
        final_err = self.evaluate("test") if self.shards is not None else float("nan")
        res = dict(steps=self.global_step, loss=last_loss, test_error=final_err, seconds=dt,
                   img_per_s=self.global_step * c.batch * self.comm.size / max(dt, 1e-9),
                   phases_ms=self.timer.summary())
        self.log.log(event="done", **{k: v for k, v in res.items() if k != "phases_ms"})
        if c.checkpoint_dir:
            self.save_checkpoint(os.path.join(c.checkpoint_dir, f"model.ckpt-{self.global_step}"))
        return res

    # ------------------------------------------------------------------ checkpoint
    def save_checkpoint(self, path: str) -> None:
        """Rank 0 writes; everyone waits.  Names/layouts per SURVEY §5.4."""
        c = self.cfg
        if hasattr(self.engine, "sync_state"):
            self.engine.sync_state()               # collective: masters / slots kept by slice owners are gathered
        if self.comm.rank == 0:
            named = ckpt.to_named(self.spec, self.engine.get_params(), self.engine.get_slots(),
                                  self.global_step, c.optimizer, c.checkpoint_names, c.beta1, c.beta2)
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            ckpt.save(path, named)
        self.comm.barrier()

    def load_checkpoint(self, path: str, dense_channels_first: bool = False) -> None:
        c = self.cfg
        named = ckpt.load(path)
        params, slots, step = ckpt.from_named(self.spec, named, c.optimizer, c.checkpoint_names,
                                              dense_channels_first)
        self.engine.set_state(params, slots, step)
        self.global_step = step

    def close(self):
        self.log.close()
        if hasattr(self.engine, "close"):
            self.engine.close()
