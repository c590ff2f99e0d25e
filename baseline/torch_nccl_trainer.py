"""B-lib: the reference-equivalent library build.

The reference itself (``mpipy.py``: TensorFlow 1.x + mpi4py, Python 2) cannot be installed
on this image, so the number to beat is the same training step written against the stock
libraries the reference's own stack would call on this hardware: cuDNN convolutions and
cuBLAS GEMMs through PyTorch, NCCL ``all_reduce`` for the gradient average (what
``bcast_parameters``' Gather + mean stands for, ``mpipy.py:95-153``) and a separate optimizer
step.  It is given every fair advantage: bf16 autocast, channels-last, a flat gradient
buffer reduced with ONE NCCL call, a fused (foreach) momentum update, the dataset resident
on the device, and the whole step captured in a CUDA graph.

None of this package's kernels, engine or communicator is used here.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class MnistCnn(nn.Module):
    """conv5x5x32 - pool - conv5x5x64 - pool - FC - dropout - FC10 (mpipy.py:155-167)."""

    def __init__(self, c1=32, c2=64, fc=1024, keep=0.6):
        super().__init__()
        self.conv1 = nn.Conv2d(1, c1, 5, padding=2)
        self.conv2 = nn.Conv2d(c1, c2, 5, padding=2)
        self.fc1 = nn.Linear(7 * 7 * c2, fc)
        self.fc2 = nn.Linear(fc, 10)
        self.drop = 1.0 - keep

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = x.flatten(1)
        x = F.dropout(F.relu(self.fc1(x)), self.drop, self.training)
        return self.fc2(x)


PRESETS = {"tutorial512": (32, 64, 512, 0.5), "official1024": (32, 64, 1024, 0.6), "wide": (128, 256, 4096, 0.6)}


class Step:
    def __init__(self, model, batch, rows, world, dev, optimizer="momentum", graph=True, seed=1, dtype="bf16"):
        torch.manual_seed(seed)
        self.bf16 = dtype != "fp32"
        c1, c2, fc, keep = PRESETS[model]
        self.net = MnistCnn(c1, c2, fc, keep).to(dev)
        if os.environ.get("BASELINE_CHANNELS_LAST", "1") == "1":
            self.net = self.net.to(memory_format=torch.channels_last)
        self.world, self.B, self.rows, self.dev = world, batch, rows, dev
        params = list(self.net.parameters())
        # one flat gradient buffer -> one NCCL all_reduce per step
        self.flat = torch.zeros(sum(p.numel() for p in params), device=dev)
        off = 0
        for p in params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        decay = [p for n, p in self.net.named_parameters() if n.startswith("fc")]
        rest = [p for n, p in self.net.named_parameters() if not n.startswith("fc")]
        groups = [{"params": decay, "weight_decay": 5e-4}, {"params": rest, "weight_decay": 0.0}]
        if optimizer == "adam":
            self.opt = torch.optim.Adam(groups, lr=1e-4, capturable=True, foreach=True)
        else:
            self.opt = torch.optim.SGD(groups, lr=0.01, momentum=0.9 if optimizer == "momentum" else 0.0,
                                       foreach=True)
        g = torch.Generator(device=dev).manual_seed(seed + dist.get_rank() if world > 1 else seed)
        self.data = (torch.rand(rows, 1, 28, 28, device=dev, generator=g) - 0.5)
        self.labels = torch.randint(0, 10, (rows,), device=dev, generator=g)
        self.step_no = torch.zeros(1, dtype=torch.int64, device=dev)
        self.x = torch.zeros(batch, 1, 28, 28, device=dev).contiguous(memory_format=torch.channels_last)
        self.y = torch.zeros(batch, dtype=torch.int64, device=dev)
        self.loss = torch.zeros((), device=dev)
        self.graph = None
        self.graph_nowin = None
        self.use_graph = graph
        self.idx = torch.arange(batch, device=dev)

    def _window(self):
        # sequential window of the resident shard, offset from a device-side counter (mpipy.py:80-82)
        off = (self.step_no * self.B) % (self.rows - self.B)
        sel = self.idx + off
        self.x.copy_(self.data.index_select(0, sel))
        self.y.copy_(self.labels.index_select(0, sel))
        self.step_no += 1

    def _body(self):
        # bf16 arm: autocast; fp32 arm: fp32 tensors with TF32 tensor-core math allowed (set in bench())
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.bf16):
            logits = self.net(self.x)
        loss = F.cross_entropy(logits.float(), self.y)
        self.flat.zero_()
        loss.backward()
        if self.world > 1:
            dist.all_reduce(self.flat)
            self.flat.mul_(1.0 / self.world)
        self.opt.step()
        self.loss.copy_(loss.detach())

    def warm(self, n=3):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(n):
                self._window()
                self._body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        if self.use_graph:
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._window()
                self._body()

    def step(self, x=None, y=None):
        if x is not None:
            self.x.copy_(x, non_blocking=True)
            self.y.copy_(y, non_blocking=True)
            if self.graph is not None:
                self.graph_nowin.replay()
            else:
                self._body()
            return
        if self.graph is not None:
            self.graph.replay()
        else:
            self._window()
            self._body()

    def capture_host_variant(self):
        if self.use_graph:
            self.graph_nowin = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph_nowin):
                self._body()


def bench(args, metric, ClockSampler, physical_gpu_index) -> dict:
    if os.environ.get("BASELINE_DEBUG"):
        import faulthandler
        import sys
        faulthandler.dump_traceback_later(int(os.environ["BASELINE_DEBUG"]), exit=True, file=sys.stderr)
    n = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    torch.backends.cudnn.benchmark = True
    if args.dtype == "fp32":                      # the same arithmetic our fp32 path uses
        torch.backends.cuda.matmul.allow_tf32 = True
        torch.backends.cudnn.allow_tf32 = True
    if n > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    st = Step(args.model, B, args.rows, n, dev, args.optimizer, graph=not args.no_graph, dtype=args.dtype)
    st.warm(3)
    for _ in range(W):
        st.step()

    def region(fn):
        if n > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        if n > 1:
            dist.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if n > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    sampler = ClockSampler(physical_gpu_index(local))
    with sampler:
        ms = region(lambda: [st.step() for _ in range(K)])
    value = n * B * K / (ms / 1e3)
    e2e = None
    if not args.no_e2e:
        st.capture_host_variant()
        hx = st.data[:16 * B].cpu().pin_memory()
        hy = st.labels[:16 * B].cpu().pin_memory()

        def one(i):
            o = (i % 16) * B
            st.step(hx[o:o + B], hy[o:o + B])
            return float(st.loss.item())

        for i in range(W):
            one(i)
        ms2 = region(lambda: [one(i) for i in range(K)])
        e2e = {"value": n * B * K / (ms2 / 1e3), "unit": "images/s", "ms_per_step": ms2 / K,
               "h2d_bytes_per_step": B * 784 * 4 + B * 8, "d2h_bytes_per_step": 4}
    out = {"metric": metric, "value": value, "unit": "images/s", "n_gpus": n, "steps": K, "warmup": W,
           "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": args.dtype, "data": "synthetic", "impl": "baseline",
           "config": {"model": f"mnist_cnn_{args.model}", "global_batch": n * B, "per_rank_batch": B, "seq_len": 784,
                      "parallelism": f"dp{n}", "optimizer": args.optimizer, "cuda_graph": not args.no_graph,
                      "stack": ("torch cuDNN/cuBLAS bf16 autocast" if args.dtype != "fp32" else "torch cuDNN/cuBLAS fp32 (TF32 allowed)")
                               + " + NCCL all_reduce + foreach optimizer",
                      "loss_last": float(st.loss.item())},
           "clocks": sampler.summary(), "gpu_launches": 0}
    if e2e:
        out["e2e"] = e2e
    # graphs that captured NCCL work keep the communicator busy: drop them and skip the explicit
    # destroy (it blocks forever here); the caller prints the result and exits the process
    st.graph = None
    st.graph_nowin = None
    torch.cuda.synchronize()
    return out if rank == 0 else {}
