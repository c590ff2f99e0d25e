#!/usr/bin/env python
"""Gradient-allreduce sweep (BASELINE.json config 4): the fused P2P kernel vs NCCL, 4 KB - 1 GB.

    python -m mpi_tensorflow_b200.parallel.launch -np 8 bench/allreduce_sweep.py [--max-mb 1024] [--out file.jsonl]

Per message size S (fp32 elements in symmetric, peer-mapped buffers) it times, with CUDA events
after warm-up and as the max over ranks:
  p2p-oneshot / p2p-twoshot   mtb_fused_allreduce_update in plain-allreduce mode (optimizer = none)
  nccl                        torch.distributed.all_reduce on the same data (the reference-equivalent
                              collective; mpipy.py:121-136 is a host Gather + NumPy mean)
and reports bus bandwidth 2(N-1)/N * S / t against the measured NVLink peer-copy figure
(770 GB/s per direction, B200_PROFILING.md).
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_tensorflow_b200.ops import native as N  # noqa: E402
from mpi_tensorflow_b200.parallel import world  # noqa: E402

LINK_GBS = 770.0


def timed(fn, comm, iters):
    comm.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    return max(comm.allgather_object(ms)) if comm.size > 1 else ms


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--max-mb", type=int, default=1024)
    ap.add_argument("--min-bytes", type=int, default=4096, help="first message size of the sweep (x4 steps)")
    ap.add_argument("--out", default=None)
    ap.add_argument("--no-nccl", action="store_true")
    a = ap.parse_args()
    comm = world.init("p2p", "cuda", timeout_s=120)
    dev, n, r = comm.device, comm.size, comm.rank
    max_elems = a.max_mb * (1 << 20) // 4
    src = comm.alloc_symmetric(max_elems * 4)
    dst = comm.alloc_symmetric(max_elems * 4)
    src.typed(torch.float32).normal_()
    nccl_ok = False
    if not a.no_nccl and n > 1:
        try:
            # second process group for the baseline collective; the p2p backend bootstrapped gloo already
            pg = dist.new_group(backend="nccl")
            t = torch.ones(8, device=dev)
            dist.all_reduce(t, group=pg)
            torch.cuda.synchronize()
            nccl_ok = True
        except Exception as e:  # noqa: BLE001
            if r == 0:
                print(f"# nccl unavailable: {e}", flush=True)
    rows = []
    size = max(4096, a.min_bytes)
    while size <= a.max_mb << 20:
        elems = size // 4
        iters = 200 if size <= (1 << 20) else (50 if size <= (64 << 20) else 10)
        rec = {"bytes": size, "n_gpus": n}
        variants = [("p2p_oneshot", 0), ("p2p_twoshot", 1)] + ([("p2p_nvls", 2)] if src.mc_ptr and dst.mc_ptr else [])
        for name, algo in variants:
            if algo == 0 and size > (64 << 20):
                continue
            args = comm.fused_args()
            args.algo, args.optimizer = algo, -1
            if algo == 2:
                args.mc_grads, args.mc_params = src.mc_ptr, dst.mc_ptr
            for k in range(n):
                args.grads[k] = src.ptrs[k]
                args.params[k] = dst.ptrs[k]
            args.offset, args.count = 0, (elems + 4 * n - 1) // (4 * n) * (4 * n)
            args.count = min(args.count, max_elems)
            args.scale = 1.0 / n

            def fn(args=args):
                N.call("mtb_fused_allreduce_update", C.byref(args))

            for _ in range(5):
                fn()
            ms = timed(fn, comm, iters)
            comm.check_status()
            rec[name + "_us"] = ms * 1e3
            rec[name + "_busbw_gbs"] = 2 * (n - 1) / n * size / (ms * 1e-3) / 1e9 if n > 1 else size / (ms * 1e-3) / 1e9
        if nccl_ok:
            buf = src.typed(torch.float32)[:elems].clone()

            def fn2(buf=buf):
                dist.all_reduce(buf, group=pg)

            for _ in range(5):
                fn2()
            ms = timed(fn2, comm, iters)
            rec["nccl_us"] = ms * 1e3
            rec["nccl_busbw_gbs"] = 2 * (n - 1) / n * size / (ms * 1e-3) / 1e9
        times = {k[:-3]: v for k, v in rec.items() if k.startswith("p2p") and k.endswith("_us")}
        best_name = min(times, key=times.get)
        rec["p2p_best"], rec["p2p_best_us"] = best_name, times[best_name]
        if n > 1:
            # Link-time floor of EACH algorithm for its own traffic (bytes over one direction of NVLink at the measured
            # 770 GB/s): one-shot pulls the whole message from N-1 peers; two-shot moves (N-1)/N of it in and as much
            # out again, back to back; with the in-switch reduction a GPU sends its (N-1)/N share of the reads up while
            # the reduced slice comes down, then broadcasts its slice: S/N * (N-1) up twice over ~ S each way.
            floors = {"p2p_oneshot": (n - 1) * size, "p2p_twoshot": 2 * (n - 1) / n * size, "p2p_nvls": size}
            for name, t in times.items():
                fl = floors[name] / (LINK_GBS * 1e9) * 1e6
                rec[name + "_floor_us"] = fl
                rec[name + "_frac_of_floor"] = fl / t
            rec["p2p_best_frac_of_its_floor"] = rec[best_name + "_frac_of_floor"]
        rows.append(rec)
        if r == 0:
            print(json.dumps(rec), flush=True)
        size *= 4
    if r == 0 and a.out:
        with open(a.out, "w") as f:
            for rec in rows:
                f.write(json.dumps(rec) + "\n")
    comm.barrier()
    if nccl_ok:
        os._exit(0)
    world.finalize()


if __name__ == "__main__":
    main()
