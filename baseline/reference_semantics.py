"""B-faithful: the reference's OWN strategy, timed.

Two arms over the same schedule - the shipped script's, literally (``/root/reference/mpipy.py``):
tutorial CNN (FC512), batch 64, fp32, momentum-SGD with the staircase LR, every rank trains on its own
shard, EVERY step evaluates the whole test shard with dropout on (``mpipy.py:86``), every 50 steps the
four weight tensors are gathered to rank 0, averaged there and assigned on root only (``mpipy.py:87-153``).

  baseline   a literal restatement on the stock libraries: per-step host->device feed of the batch
             (``mpipy.py:83-85``), cuDNN / cuBLAS through PyTorch (TF32 allowed, CUDA-graphed step so the
             session's executor is not charged with Python overhead), eval as ``eval_in_batches`` does it -
             one feed + one fetch per 64 rows (``mpipy.py:174-182``) - and the sync as device->host copies,
             a Gather to rank 0 over the host process group (gloo standing in for MPI), a NumPy mean and
             host->device assigns on root.
  ours       this repository through its public API with ``--reference-semantics``: graph-captured local
             steps, the test shard resident and its evaluation one graph replay + one fetch, the sync one
             in-place averaging kernel over NVLink peer memory (root-only mask).

Both report images/s "train only" (eval switched off, sync kept) and "with eval" (the reference's
actual loop), device-timed, max over ranks.
"""
from __future__ import annotations

import os

import numpy as np
import torch
import torch.distributed as dist
import torch.nn.functional as F

ROWS_TRAIN, ROWS_TEST, BATCH, PERIOD = 55000, 10000, 64, 50


def _region(fn, barrier, reduce_max):
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    return reduce_max(e0.elapsed_time(e1))


def run_baseline(steps: int, warmup: int) -> dict:
    from torch_nccl_trainer import MnistCnn, PRESETS
    n = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    torch.backends.cudnn.benchmark = True
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    if n > 1:
        dist.init_process_group("gloo")                       # host collectives, as the reference's MPI
    torch.manual_seed(1)
    c1, c2, fc, keep = PRESETS["tutorial512"]
    net = MnistCnn(c1, c2, fc, keep).to(dev)
    net.train()                                               # dropout is unconditional in the reference (mpipy.py:166)
    decay = [p for k, p in net.named_parameters() if k.startswith("fc")]
    rest = [p for k, p in net.named_parameters() if not k.startswith("fc")]
    opt = torch.optim.SGD([{"params": decay, "weight_decay": 5e-4}, {"params": rest, "weight_decay": 0.0}],
                          lr=0.01, momentum=0.9, foreach=True)
    rows, trows = ROWS_TRAIN // n, ROWS_TEST // n
    g = torch.Generator().manual_seed(1 + rank)
    hx = (torch.rand(rows, 1, 28, 28, generator=g) - 0.5).pin_memory()
    hy = torch.randint(0, 10, (rows,), generator=g).pin_memory()
    tx = (torch.rand(trows, 1, 28, 28, generator=g) - 0.5).pin_memory()
    ty = torch.randint(0, 10, (trows,), generator=g).numpy()
    x = torch.zeros(BATCH, 1, 28, 28, device=dev)
    y = torch.zeros(BATCH, dtype=torch.int64, device=dev)
    ex = torch.zeros(BATCH, 1, 28, 28, device=dev)
    probs = torch.zeros(BATCH, 10, device=dev)

    def train_body():
        loss = F.cross_entropy(net(x), y)
        opt.zero_grad(set_to_none=False)
        loss.backward()
        opt.step()

    def eval_body():
        with torch.no_grad():
            probs.copy_(F.softmax(net(ex), dim=1))

    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            train_body()
            eval_body()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g_train, g_eval = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    with torch.cuda.graph(g_train):
        train_body()
    with torch.cuda.graph(g_eval):
        eval_body()
    weights = [net.conv1.weight, net.conv2.weight, net.fc1.weight, net.fc2.weight]     # biases are never synced

    def sync():
        for w in weights:                                     # mpipy.py:121-127 .eval() + Gather, one tensor at a time
            send = w.detach().cpu()
            recv = [torch.empty_like(send) for _ in range(n)] if rank == 0 else None
            if n > 1:
                dist.gather(send, recv, dst=0)
            else:
                recv = [send]
            if rank == 0:                                     # mpipy.py:129-152
                mean = np.mean(np.stack([r.numpy() for r in recv]), axis=0)
                with torch.no_grad():
                    w.copy_(torch.from_numpy(mean).to(dev))

    def evaluate():
        out = np.empty((trows, 10), dtype=np.float32)         # mpipy.py:169-183
        for b in range(0, trows, BATCH):
            e = b + BATCH
            if e <= trows:
                ex.copy_(tx[b:e], non_blocking=True)
                g_eval.replay()
                out[b:e] = probs.cpu().numpy()
            else:
                ex.copy_(tx[-BATCH:], non_blocking=True)
                g_eval.replay()
                out[b:] = probs.cpu().numpy()[b - trows:]
        return 100.0 - 100.0 * float(np.mean(np.argmax(out, 1) == ty))

    state = {"step": 0}

    def loop(k, with_eval):
        for _ in range(k):
            st = state["step"]
            o = (st * BATCH) % (rows - BATCH)
            x.copy_(hx[o:o + BATCH], non_blocking=True)       # the per-step feed, mpipy.py:83-85
            y.copy_(hy[o:o + BATCH], non_blocking=True)
            g_train.replay()
            if with_eval:
                evaluate()
            if st > 0 and st % PERIOD == 0:
                sync()
            state["step"] = st + 1

    barrier = (lambda: dist.barrier()) if n > 1 else (lambda: None)

    def rmax(v):
        if n == 1:
            return v
        t = torch.tensor([v])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    loop(warmup, False)
    ms_train = _region(lambda: loop(steps, False), barrier, rmax)
    k_eval = max(4, steps // 10)
    loop(2, True)
    ms_eval = _region(lambda: loop(k_eval, True), barrier, rmax)
    return {"impl": "baseline-literal", "n_gpus": n, "train_only_img_s": n * BATCH * steps / (ms_train / 1e3),
            "train_only_ms_per_step": ms_train / steps, "with_eval_img_s": n * BATCH * k_eval / (ms_eval / 1e3),
            "with_eval_ms_per_step": ms_eval / k_eval, "steps": steps, "eval_steps": k_eval} if rank == 0 else {}


def run_ours(steps: int, warmup: int) -> dict:
    from mpi_tensorflow_b200 import TrainConfig, Trainer
    from mpi_tensorflow_b200.parallel import world
    from mpi_tensorflow_b200.utils import data as D
    n = int(os.environ.get("WORLD_SIZE", "1"))
    comm = world.init("p2p" if n > 1 else "self", "cuda", timeout_s=120)
    cfg = TrainConfig.reference()
    cfg.batch, cfg.eval_every, cfg.log_every, cfg.eval_batch = BATCH, 0, 0, BATCH       # the loop below drives eval itself
    tr = Trainer(cfg, comm)
    rows, trows = ROWS_TRAIN // n, ROWS_TEST // n
    u8, lab = D.synthetic_mnist(rows // 8 + trows, 1, offset=comm.rank * 100000)        # a slice is enough for timing
    xs = D.normalize_u8(u8).reshape(-1, 28, 28, 1)
    ys = lab.astype(np.int64)
    reps = -(-rows // (rows // 8))
    shards = D.Shards(np.tile(xs[:rows // 8], (reps, 1, 1, 1))[:rows], np.tile(ys[:rows // 8], reps)[:rows],
                      xs[rows // 8:], ys[rows // 8:], xs[:64], ys[:64])
    tr.load_data(shards)

    def loop(k, with_eval):
        for _ in range(k):
            tr.step(read_loss=False)                          # local graph step (+ the averaging kernel every 50)
            if with_eval:
                tr.evaluate("test")                           # mpipy.py:86, dropout on (reference config)

    barrier = comm.barrier
    rmax = (lambda v: max(comm.allgather_object(v))) if n > 1 else (lambda v: v)
    loop(warmup, False)
    ms_train = _region(lambda: loop(steps, False), barrier, rmax)
    k_eval = max(4, steps // 10)
    loop(2, True)
    ms_eval = _region(lambda: loop(k_eval, True), barrier, rmax)
    out = {"impl": "ours", "n_gpus": n, "train_only_img_s": n * BATCH * steps / (ms_train / 1e3),
           "train_only_ms_per_step": ms_train / steps, "with_eval_img_s": n * BATCH * k_eval / (ms_eval / 1e3),
           "with_eval_ms_per_step": ms_eval / k_eval, "steps": steps, "eval_steps": k_eval}
    tr.close()
    world.finalize()
    return out if comm.rank == 0 else {}
