"""Per-rank body of the multi-GPU tests (one process per GPU, launched by parallel.launch).

modes:
  algebra   every P2PComm collective on device tensors equals serial NumPy
  fused     the fused allreduce+update kernel: replicas bit-identical, one-shot == two-shot,
            N ranks x batch b == 1 rank x batch N*b (DP equivalence)
  train     Trainer.fit() through the public API, CUDA graph on
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mpi_tensorflow_b200 import TrainConfig, Trainer  # noqa: E402
from mpi_tensorflow_b200.parallel import world  # noqa: E402
from mpi_tensorflow_b200.utils import data as D  # noqa: E402


def shards_for(rank, rows, seed=0):
    u8, lab = D.synthetic_mnist(rows, seed, offset=rank * rows)
    x = D.normalize_u8(u8).reshape(rows, 28, 28, 1)
    y = lab.astype(np.int64)
    return D.Shards(x, y, x[:64], y[:64], x[:64], y[:64])


def algebra(comm):
    r, n, dev = comm.rank, comm.size, comm.device
    rng = np.random.default_rng(99)
    for count in (4, 1000, 100003, (3 << 20) + 1):
        full = rng.standard_normal((n, count)).astype(np.float32)
        t = torch.from_numpy(full[r].copy()).to(dev)
        comm.allreduce(t, average=True)
        ref = full[0].copy()
        for p in range(1, n):
            ref += full[p]
        ref *= np.float32(1.0 / n)
        assert np.allclose(t.cpu().numpy(), ref, rtol=1e-6, atol=1e-6), ("allreduce", count)
        chk = t.clone()
        comm.bcast(chk, root=0)
        assert torch.equal(chk, t), "allreduce result differs across ranks"
        recv = torch.zeros(count, device=dev)
        comm.scatter(torch.from_numpy(full.copy()) if r == 0 else None, recv, root=0)
        assert np.array_equal(recv.cpu().numpy(), full[r]), ("scatter", count)
        out = torch.zeros(n, count, device=dev) if r == 0 else None
        comm.gather(torch.from_numpy(full[r].copy()).to(dev), out, root=0)
        if r == 0:
            assert np.array_equal(out.cpu().numpy(), full), ("gather", count)
        ag = torch.zeros(n * count, device=dev)
        comm.allgather(torch.from_numpy(full[r].copy()).to(dev), ag)
        assert np.array_equal(ag.cpu().numpy().reshape(n, count), full), ("allgather", count)
        b = torch.from_numpy(full[n - 1].copy()).to(dev) if r == n - 1 else torch.zeros(count, device=dev)
        comm.bcast(b, root=n - 1)
        assert np.array_equal(b.cpu().numpy(), full[n - 1]), ("bcast", count)
    for _ in range(50):
        comm.barrier()
    # Flag-protocol stress ("race detection": compute-sanitizer is closed on this pool): 1000 device
    # barriers queued back to back with no host synchronisation between them (sequence numbers only ever grow, a rank
    # may run several rounds ahead of a slow peer), then 100 allreduces of varying small sizes whose exact result shows
    # at once if any rank ever read a peer's buffer from the wrong round.
    for _ in range(1000):
        comm.device_barrier()
    torch.cuda.synchronize()
    comm.check_status()
    sizes = np.random.default_rng(7).integers(1, 5000, size=100)       # the same sequence on every rank
    for it, count in enumerate(sizes):
        t = torch.full((int(count),), float(r + 1 + it), device=dev)
        comm.allreduce(t, average=False)
        assert bool((t == float(n * (n + 1) // 2 + n * it)).all()), ("stress allreduce", it, int(count))
    print(f"GPU_COMM_OK rank {r}/{n}", flush=True)


def fused(comm, outdir):
    r, n, dev = comm.rank, comm.size, comm.device
    b, steps = 32, 6
    finals = {}
    mc = getattr(comm, "has_multicast", False)
    # "+rs": the FC1 weight gradient is reduce-scattered inside its GEMM (tiles added into the owner over NVLink)
    # "+rs": the FC1 weight gradient is reduce-scattered inside its GEMM; "+single": one whole-buffer K1 launch instead
    # of the bucketed overlap; "+bf16": bf16 gradient wire format; "auto" = hidden two-shot / nvls bucket + push tail
    algos = (["oneshot", "twoshot"] + (["nvls"] if mc else []) + ["twoshot+rs"] + (["nvls+rs"] if mc else [])
             + ["auto", "auto+single", "twoshot+single", "auto+bf16"] + (["nvls+bf16"] if mc else []))
    print(f"rank {r}: allocator={'symm' if comm._symm is not None else 'ipc'} multicast={comm.has_multicast}", flush=True)
    slots_of = {}
    for algo in algos:
        opts = algo.split("+")[1:]
        cfg = TrainConfig(model="tutorial512", batch=b, eval_every=0, log_every=0, dropout=False,
                          comm_algo=algo.split("+")[0], fuse_wgrad_rs="rs" in opts, overlap="single" not in opts,
                          comm_dtype="bf16" if "bf16" in opts else "fp32",
                          cuda_graph=(algo != "oneshot"), lr_decay_rows=10 ** 9)
        tr = Trainer(cfg, comm)
        assert tr.engine._rs_mode == ("rs" in opts)
        assert tr.engine._overlap == ("single" not in opts and "rs" not in opts)
        # rank r trains on rows [r*b, (r+1)*b) of every global batch of n*b rows
        rows = steps * b + b
        u8, lab = D.synthetic_mnist(n * rows, 5)
        gx = D.normalize_u8(u8).reshape(n * rows, 28, 28, 1)
        gy = lab.astype(np.int64)
        losses = []
        for s in range(steps):
            lo = s * n * b + r * b
            losses.append(tr.step(torch.from_numpy(gx[lo:lo + b]), torch.from_numpy(gy[lo:lo + b]).int()))
        tr.engine.synchronize()
        comm.check_status()
        tr.engine.sync_state()                   # fp32 masters of the big weights live with their slice owners
        mine = tr.engine.flat_params.clone()
        ref = mine.clone()
        comm.bcast(ref, root=0)
        assert torch.equal(mine, ref), f"{algo}: replicas diverged"
        assert torch.equal(tr.engine.shadow, tr.engine.flat_params.to(torch.bfloat16)), "bf16 shadow stale"
        finals[algo] = mine.cpu()
        slots_of[algo] = tr.engine.slots[0].clone().cpu()       # complete on every rank after sync_state()
        sl = slots_of[algo].to(dev)
        ref_s = sl.clone()
        comm.bcast(ref_s, root=0)
        assert torch.equal(sl, ref_s), f"{algo}: gathered momentum differs across ranks"
        # Ownership is per bucket.  A rank never touches the momentum of a slice it does not own, so that
        # memory still holds its initial zeros: if sync_state() took any part of the FC1 weights' momentum from a rank
        # that does not own it, zeros would show up there (the L2 term makes every true entry non-zero).
        fc1w = tr.spec.param("fc1_weight")
        m_fc1 = slots_of[algo][fc1w.offset:fc1w.offset + fc1w.numel]
        assert float((m_fc1 == 0).float().mean()) < 1e-4, f"{algo}: stale momentum gathered from a non-owner"
        assert all(np.isfinite(losses)) and losses[0] > 1.0, losses      # CE + wd*L2: the L2 part alone is ~3
        if r == 0 and algo == "oneshot":
            # single-process big-batch run on the same rows (no peers: world-1 engine)
            from mpi_tensorflow_b200.parallel.comm import SelfComm
            cfg1 = TrainConfig(model="tutorial512", batch=n * b, eval_every=0, log_every=0, dropout=False,
                               cuda_graph=False, lr_decay_rows=10 ** 9)
            t1 = Trainer(cfg1, SelfComm(), device=dev)
            for s in range(steps):
                lo = s * n * b
                t1.step(torch.from_numpy(gx[lo:lo + n * b]), torch.from_numpy(gy[lo:lo + n * b]).int())
            t1.engine.synchronize()
            big = t1.engine.flat_params.cpu()
            rel = float((big - finals[algo]).norm() / big.norm())
            assert rel < 2e-3, f"DP equivalence: rel diff {rel}"
            t1.close()
        tr.close()
        comm.barrier()
    for algo in algos[1:]:
        rel = float((finals["oneshot"] - finals[algo]).norm() / finals["oneshot"].norm())
        assert rel < (5e-3 if "bf16" in algo else 1e-3), f"oneshot vs {algo}: {rel}"
        # momentum is a sum of six gradients, each carrying the run-to-run noise of the fp32 atomics: percent level
        rel_s = float((slots_of["oneshot"] - slots_of[algo]).norm() / slots_of["oneshot"].norm())
        assert rel_s < 5e-2, f"momentum oneshot vs {algo}: {rel_s}"
    print(f"GPU_FUSED_OK rank {r}/{n}", flush=True)


def train(comm, outdir):
    cfg = TrainConfig(model="official1024", batch=100, max_steps=30, eval_every=10, log_every=10, profile_phases=True,
                      dropout=True, step_timeout_s=60, log_jsonl=os.path.join(outdir, f"log_rank{comm.rank}.jsonl"))
    tr = Trainer(cfg, comm)
    tr.load_data()
    res = tr.fit()
    tr.engine.sync_state()
    mine = tr.engine.flat_params.clone()
    ref = mine.clone()
    comm.bcast(ref, root=0)
    assert torch.equal(mine, ref), "replicas diverged"
    if comm.rank == 0:
        with open(os.path.join(outdir, "result.json"), "w") as f:
            json.dump({k: v for k, v in res.items() if k != "phases_ms"}, f)
        # the log line carries what SURVEY section 5 asks for: train / test error, per-phase times, bus bandwidth
        recs = [json.loads(l) for l in open(os.path.join(outdir, "log_rank0.jsonl"))]
        line = [r for r in recs if r.get("step") == 20][0]
        for k in ("loss", "lr", "train_error", "test_error", "img_per_s", "fwd_us", "bwd_us", "allreduce_us",
                  "allreduce_exposed_us", "bus_gbs"):
            assert k in line, (k, line)
        assert 0.0 <= line["train_error"] <= 100.0 and line["allreduce_us"] > 0 and line["bus_gbs"] > 0, line
    tr.close()
    comm.barrier()
    # the fp32 / TF32 path over the same fused allreduce (full fp32 parameter broadcast, no bf16 shadow)
    cfg32 = TrainConfig(model="official1024", batch=100, dtype="fp32", max_steps=12, eval_every=0, log_every=0,
                        dropout=True, step_timeout_s=60, lr=0.02)
    tr = Trainer(cfg32, comm)
    tr.load_data()
    losses = [tr.step() for _ in range(12)]
    tr.engine.sync_state()
    comm.check_status()
    mine = tr.engine.flat_params.clone()
    ref = mine.clone()
    comm.bcast(ref, root=0)
    assert torch.equal(mine, ref), "fp32 replicas diverged"
    assert np.isfinite(losses).all() and losses[-1] < losses[0], losses
    print(f"GPU_TRAIN_OK rank {comm.rank}/{comm.size}", flush=True)
    tr.close()


def weight_avg(comm, outdir):
    """Reference-style sync on GPUs (mpipy.py:95-153 semantics, fixed so every rank gets the mean):
    local momentum-SGD steps, parameters averaged every 5 steps through the P2P allreduce."""
    cfg = TrainConfig(model="tutorial512", batch=32, max_steps=12, eval_every=0, log_every=0, dropout=False,
                      sync="weight_avg", sync_period=5, sync_include_bias=True)
    tr = Trainer(cfg, comm)
    tr.load_data(shards_for(comm.rank, 256))
    for _ in range(11):                                      # steps 0..10: syncs after steps 5 and 10
        tr.step()
    tr.engine.synchronize()
    mine = tr.engine.flat_params.clone()
    ref = mine.clone()
    comm.bcast(ref, root=0)
    assert torch.equal(mine, ref), "weight_avg: replicas differ right after a sync step"
    tr.step()                                                # one local step on different data: replicas drift apart
    drift = tr.engine.flat_params.clone()
    ref2 = drift.clone()
    comm.bcast(ref2, root=0)
    if comm.rank != 0:
        assert not torch.equal(drift, ref2), "weight_avg: local step did not change the replica"
    tr.close()
    comm.barrier()
    # the averaging kernel itself on known values: all ranks / root only / momentum included
    for root_only, with_m in ((False, True), (True, False)):
        cfg = TrainConfig(model="tutorial512", batch=32, eval_every=0, log_every=0, dropout=False, sync="weight_avg",
                          sync_period=1, sync_include_bias=False, sync_include_momentum=with_m, sync_root_only=root_only)
        tr = Trainer(cfg, comm)
        eng, n, r = tr.engine, comm.size, comm.rank
        eng.flat_params.fill_(float(r + 1))
        eng.slots[0].fill_(float(10 * (r + 1)))
        eng.params_changed()
        torch.cuda.synchronize()
        comm.barrier()
        tr.sync.after_step(1, eng.flat_params, eng.slots, engine=eng)
        torch.cuda.synchronize()
        comm.barrier()
        comm.check_status()
        mean = (n + 1) / 2.0
        for p in tr.spec.params:
            seg = eng.flat_params[p.offset:p.offset + p.numel]
            sh = eng.shadow[p.offset:p.offset + p.numel].float()
            if p.is_bias or (root_only and r != 0):          # biases are not averaged here; off-root keeps its own
                assert bool((seg == r + 1).all()) and bool((sh == r + 1).all()), (p.key, root_only)
            else:
                assert bool((seg == mean).all()), (p.key, root_only, float(seg[0]))
                assert bool((sh == float(torch.tensor(mean).bfloat16())).all()), p.key
            m = eng.slots[0][p.offset:p.offset + p.numel]
            want = 10 * mean if (with_m and not p.is_bias) else 10 * (r + 1)
            assert bool((m == want).all()), (p.key, "momentum", float(m[0]), want)
        tr.close()
        comm.barrier()
    print(f"GPU_WAVG_OK rank {comm.rank}/{comm.size}", flush=True)


def die(comm, outdir):
    """Fault tier on GPUs: rank 1 stalls (a wedged process, not a dead one: the launcher would end the
    job by itself for an exit) and never reaches the collective; rank 0 must come back with a clear
    error (device-side spin limit -> status word -> CommTimeout), not hang with the GPU spinning."""
    import time
    cfg = TrainConfig(model="tutorial512", batch=32, eval_every=0, log_every=0, dropout=False, cuda_graph=False)
    tr = Trainer(cfg, comm)
    tr.load_data(shards_for(comm.rank, 128))
    tr.step()
    if comm.rank == 1:
        time.sleep(150)
        os._exit(17)
    t0 = time.time()
    try:
        tr.step()
        tr.engine.synchronize()
        comm.check_status()
    except Exception as e:  # noqa: BLE001
        print(f"CLEAN_FAILURE {type(e).__name__} after {time.time() - t0:.1f}s", flush=True)
        os._exit(3)
    print("UNEXPECTED_SUCCESS", flush=True)
    os._exit(0)


def main():
    mode, outdir = sys.argv[1], sys.argv[2]
    comm = world.init("p2p", "cuda", timeout_s=60)
    if mode == "algebra":
        algebra(comm)
    elif mode == "fused":
        fused(comm, outdir)
    elif mode == "train":
        train(comm, outdir)
    elif mode == "weight_avg":
        weight_avg(comm, outdir)
    elif mode == "die":
        die(comm, outdir)
    world.finalize()


if __name__ == "__main__":
    main()
