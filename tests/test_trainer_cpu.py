"""Integration tiers on CPU: 2-rank plumbing, DP equivalence,
reference-semantics mode, checkpoint round trip, fault injection."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from mpi_tensorflow_b200 import FakeWorld, TrainConfig, Trainer
from mpi_tensorflow_b200.models.spec import get_spec
from mpi_tensorflow_b200.parallel.comm import CommError
from mpi_tensorflow_b200.utils import checkpoint as ckpt
from mpi_tensorflow_b200.utils import data as D

WORKER = os.path.join(os.path.dirname(__file__), "workers", "train_worker.py")
CPU = torch.device("cpu")


def _shards(n_rows, seed=0, offset=0):
    u8, lab = D.synthetic_mnist(n_rows, seed, offset=offset)
    x = D.normalize_u8(u8).reshape(n_rows, 28, 28, 1)
    y = lab.astype(np.int64)
    return D.Shards(x, y, x[:64], y[:64], x[:64], y[:64])


def _cfg(**kw):
    base = dict(model="tiny", batch=8, device="cpu", eval_every=0, log_every=0, dropout=False,
                weight_decay=5e-4)
    base.update(kw)
    return TrainConfig(**base)


def test_single_rank_loss_decreases_and_eval_tail():
    tr = Trainer(_cfg(batch=32, lr=0.05), device=CPU)
    tr.load_data(_shards(256))
    losses = [tr.step() for _ in range(50)]
    assert losses[-1] < 0.6 * losses[0]
    # ragged tail handling of eval_in_batches (mpipy.py:179-182): 100 rows, batch 64
    x = tr.shards.train_x[:100]
    tr.cfg.eval_batch = 64
    p = tr.eval_in_batches(x)
    full = tr.engine.predict(x).numpy()
    assert p.shape == (100, 10) and np.allclose(p, full, atol=1e-5)
    with pytest.raises(ValueError):
        tr.eval_in_batches(x[:10])


@pytest.mark.parametrize("n", [2, 4])
def test_dp_equivalence_grad_allreduce(n):
    """N ranks x batch b, gradient allreduce == 1 rank x batch N*b on the same rows."""
    b, steps = 4, 4
    big = _shards(n * 64)

    def body(comm):
        r = comm.rank
        tr = Trainer(_cfg(batch=b, optimizer="momentum"), comm, device=CPU)
        tr.decay_rows = 10 ** 9
        for s in range(steps):
            lo = s * n * b + r * b                     # rank r takes slice r of the global batch
            tr.step(big.train_x[lo:lo + b], big.train_y[lo:lo + b])
        return tr.engine.flat_params.clone()

    par = FakeWorld(n).run(body)
    for r in range(1, n):
        assert torch.equal(par[r], par[0])             # replicas bit-identical after every sync
    one = Trainer(_cfg(batch=n * b, optimizer="momentum"), device=CPU)
    one.decay_rows = 10 ** 9
    for s in range(steps):
        lo = s * n * b
        one.step(big.train_x[lo:lo + n * b], big.train_y[lo:lo + n * b])
    assert torch.allclose(par[0], one.engine.flat_params, atol=2e-6, rtol=1e-5)


def test_weight_average_all_ranks_equal_after_sync():
    def body(comm):
        tr = Trainer(_cfg(sync="weight_avg", sync_period=3, optimizer="momentum"), comm, device=CPU)
        tr.load_data(_shards(128, offset=1000 * comm.rank))       # different data per rank
        snaps = []
        for s in range(7):
            tr.step()
            snaps.append(tr.engine.flat_params.clone())
        return snaps

    a, b = FakeWorld(2).run(body)
    assert not torch.equal(a[1], b[1])                 # local SGD between syncs: replicas drift
    assert torch.equal(a[3], b[3])                     # step 3 triggers a sync (step>0 and step%3==0)
    assert not torch.equal(a[4], b[4])
    assert torch.equal(a[6], b[6])


def test_reference_semantics_root_only_weights_only():
    """mpipy.py:121-152 literally: weights (not biases) averaged, result on root only."""
    spec = get_spec("tiny")

    def body(comm):
        tr = Trainer(_cfg(sync="weight_avg", sync_period=2, sync_root_only=True, sync_include_bias=False),
                     comm, device=CPU)
        tr.load_data(_shards(128, offset=1000 * comm.rank))
        tr.step(); tr.step()
        before = tr.engine.flat_params.clone()
        # recompute what step 2 does without the sync to isolate the sync's effect
        tr.step()
        return before, tr.engine.flat_params.clone()

    # serial simulation of the same thing with sync disabled
    def body_nosync(comm):
        tr = Trainer(_cfg(sync="none"), comm, device=CPU)
        tr.load_data(_shards(128, offset=1000 * comm.rank))
        for _ in range(3):
            tr.step()
        return tr.engine.flat_params.clone()

    (b0, a0), (b1, a1) = FakeWorld(2).run(body)
    n0, n1 = FakeWorld(2).run(body_nosync)
    assert torch.equal(a1, n1)                         # non-root keeps its local weights (mpipy.py:129)
    w = spec.param("fc1_weight")
    bia = spec.param("fc1_bias")
    sl = slice(w.offset, w.offset + w.numel)
    assert torch.allclose(a0[sl], 0.5 * (n0[sl] + n1[sl]), atol=1e-7)          # root got the mean
    sb = slice(bia.offset, bia.offset + bia.numel)
    assert torch.equal(a0[sb], n0[sb])                 # biases untouched (mpipy.py:122 commented out)


def test_two_process_cpu_plumbing(tmp_path):
    """BASELINE.json config 1: official MNIST CNN... (tiny preset for CI speed; the
    official1024 preset runs through the same code in test_official_model_two_ranks)."""
    for backend in ("gloo", "shm"):
        out = tmp_path / backend
        out.mkdir()
        cmd = [sys.executable, "-m", "mpi_tensorflow_b200.parallel.launch", "-np", "2", "--timeout", "300",
               "--", WORKER, backend, "tiny", "12", str(out)]
        env = dict(os.environ, PYTHONPATH=os.path.dirname(os.path.dirname(__file__)))
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env, timeout=400)
        assert p.returncode == 0, p.stdout
        assert p.stdout.count("TRAIN_OK") == 2
        a = torch.load(out / "params_0.pt")
        b = torch.load(out / "params_1.pt")
        assert torch.equal(a, b)
        res = json.loads((out / "result.json").read_text())
        assert res["steps"] == 12 and np.isfinite(res["loss"])
        assert (out / "ckpt" / "model.ckpt-12.index").exists()


def test_official_model_two_ranks_fake_world():
    """The named config (official MNIST CNN, synthetic 28x28, 2 ranks, CPU), 2 steps."""
    def body(comm):
        tr = Trainer(TrainConfig(model="official1024", batch=8, device="cpu", eval_every=0, log_every=0,
                                 weight_decay=0.0), comm, device=CPU)
        tr.load_data(_shards(64, offset=100 * comm.rank))
        l0 = tr.step()
        l1 = tr.step()
        return l0, l1, tr.engine.flat_params.clone()

    a, b = FakeWorld(2).run(body)
    assert np.isfinite(a[0]) and np.isfinite(a[1])
    assert torch.equal(a[2], b[2])


@pytest.mark.parametrize("fmt,names,opt", [("npz", "tf1", "momentum"), ("bundle", "tf1", "momentum"),
                                           ("bundle", "keras", "adam")])
def test_checkpoint_roundtrip(tmp_path, fmt, names, opt):
    cfg = _cfg(optimizer=opt, checkpoint_names=names)
    tr = Trainer(cfg, device=CPU)
    tr.load_data(_shards(64))
    for _ in range(3):
        tr.step()
    path = str(tmp_path / ("m.npz" if fmt == "npz" else "model.ckpt-3"))
    tr.save_checkpoint(path)
    named = ckpt.load(path)
    spec = tr.spec
    # name map equals SURVEY §5.4 tables
    if names == "tf1":
        assert set(f"Variable_{i}" if i else "Variable" for i in range(9)) <= set(named)
        assert named["Variable_4"].shape == spec.param("fc1_weight").shape
        assert "Variable/Momentum" in named and named["Variable_8"].dtype == np.float32
    else:
        for k in ("conv2d/kernel", "conv2d_1/bias", "dense/kernel", "dense_1/bias", "global_step",
                  "dense/kernel/Adam", "dense/kernel/Adam_1", "beta1_power"):
            assert k in named, k
        assert named["global_step"].dtype == np.int64 and int(named["global_step"]) == 3
    tr2 = Trainer(cfg, device=CPU)
    tr2.load_checkpoint(path)
    assert tr2.global_step == 3
    assert torch.equal(tr2.engine.flat_params, tr.engine.flat_params)
    for s1, s2 in zip(tr.engine.slots, tr2.engine.slots):
        assert torch.equal(s1, s2)
    tr2.load_data(_shards(64))
    assert tr.step() == pytest.approx(tr2.step(), abs=0)       # resumed run continues bit-exactly


def test_bundle_format_invariants(tmp_path):
    named = {"b/x": np.arange(6, dtype=np.float32).reshape(2, 3), "a": np.asarray(7, dtype=np.int64),
             **{f"v{i:03d}": np.full((40,), i, dtype=np.float32) for i in range(300)}}   # > one 4 KiB block
    prefix = str(tmp_path / "ck")
    ckpt.save_tf_bundle(prefix, named)
    raw = open(prefix + ".index", "rb").read()
    assert raw[-8:] == bytes.fromhex("57fb808b247547db")        # leveldb table magic, little-endian
    got = ckpt.load_tf_bundle(prefix)
    assert set(got) == set(named)
    for k in named:
        assert np.array_equal(got[k], named[k]) and got[k].dtype == named[k].dtype
    # corruption is detected through the masked CRC32C
    data = bytearray(open(prefix + ".data-00000-of-00001", "rb").read())
    data[5] ^= 0xFF
    open(prefix + ".data-00000-of-00001", "wb").write(bytes(data))
    with pytest.raises(ValueError):
        ckpt.load_tf_bundle(prefix)
    assert ckpt.masked_crc32c(b"123456789") == ((((0xE3069283 >> 15) | (0xE3069283 << 17)) & 0xFFFFFFFF) + 0xA282EAD8) & 0xFFFFFFFF


def test_channels_first_dense_import():
    spec = get_spec("tiny")
    rng = np.random.default_rng(0)
    params = {p.key: rng.standard_normal(p.shape).astype(np.float32) for p in spec.params}
    named = ckpt.to_named(spec, params, None, 0, "sgd", "keras")
    s, c = spec.pooled2, spec.c2
    hwc = named["dense/kernel"].reshape(s, s, c, spec.fc)
    named_cf = dict(named)
    named_cf["dense/kernel"] = hwc.transpose(2, 0, 1, 3).reshape(-1, spec.fc)       # (c,h,w) rows
    back, _, _ = ckpt.from_named(spec, named_cf, "sgd", "keras", dense_channels_first=True)
    assert np.array_equal(back["fc1_weight"], params["fc1_weight"])


def test_fault_injection_mid_training(monkeypatch):
    monkeypatch.setenv("MTB_FAULT", "rank=1,step=2,kind=raise")

    def body(comm):
        tr = Trainer(_cfg(), comm, device=CPU)
        tr.load_data(_shards(64))
        for _ in range(5):
            tr.step()

    from mpi_tensorflow_b200.utils.faults import InjectedFault
    with pytest.raises((InjectedFault, CommError)):
        FakeWorld(2, timeout=2.0).run(body)


def test_loss_mailbox_decoding_without_a_gpu():
    """The host side of the loss mailbox (engine_native.NativeEngine._await_loss): one 8-byte word
    carries {step number : fp32 loss}; a stale or missing word must time out, never be returned."""
    import struct
    import types

    from mpi_tensorflow_b200.engine_native import NativeEngine

    ring = NativeEngine._MAIL_RING
    mail = np.zeros(ring, dtype=np.int64)
    eng = types.SimpleNamespace(_mail_np=mail, _MAIL_RING=ring, cfg=types.SimpleNamespace(step_timeout_s=0.2))

    def post(step, loss):
        bits = struct.unpack("<I", struct.pack("<f", loss))[0]
        mail[(step - 1) % ring] = np.array([(step << 32) | bits], dtype=np.uint64).view(np.int64)[0]

    post(5, 1.25)
    assert NativeEngine._await_loss(eng, 5) == 1.25
    post(5 + ring, -3.5)                                       # the slot is reused one ring later
    assert NativeEngine._await_loss(eng, 5 + ring) == -3.5
    with pytest.raises(RuntimeError):
        NativeEngine._await_loss(eng, 5)                       # overwritten: the old step can no longer be read
    with pytest.raises(RuntimeError):
        NativeEngine._await_loss(eng, 6, timeout=0.05)         # never mailed
    post(2 ** 31 + 7, 0.5)                                     # step numbers with the top bit set decode too
    assert NativeEngine._await_loss(eng, 2 ** 31 + 7) == 0.5


def test_crash_then_resume_matches_uninterrupted_run(tmp_path):
    """Failure detection + checkpoint/resume through the command line: a rank dies at step 9,
    the launcher takes the job down, a second launch resumes from the last checkpoint and ends with exactly
    the parameters, momentum and step counter of a run that never crashed."""
    root = os.path.dirname(os.path.dirname(__file__))
    env = dict(os.environ, PYTHONPATH=root)
    env.pop("MTB_FAULT", None)

    def launch(ckdir, extra=(), fault=None):
        cmd = [sys.executable, "-m", "mpi_tensorflow_b200.parallel.launch", "-np", "2", "--timeout", "240", "-m",
               "mpi_tensorflow_b200", "--device", "cpu", "--backend", "shm", "--model", "tiny", "--batch", "16",
               "--max-steps", "12", "--eval-every", "0", "--log-every", "0", "--step-timeout-s", "20",
               "--checkpoint-dir", str(ckdir), "--checkpoint-every", "4", *extra]
        e = dict(env, MTB_FAULT=fault) if fault else env
        return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=e, timeout=300)

    a, b = tmp_path / "a", tmp_path / "b"
    p = launch(a)
    assert p.returncode == 0, p.stdout[-2000:]
    p = launch(b, fault="rank=1,step=9,kind=exit")
    assert p.returncode != 0, "the job must fail when a rank dies"
    assert (b / "model.ckpt-8.index").exists() and not (b / "model.ckpt-12.index").exists(), p.stdout[-2000:]
    p = launch(b, extra=("--resume", str(b / "model.ckpt-8")))
    assert p.returncode == 0, p.stdout[-2000:]
    want, got = ckpt.load(str(a / "model.ckpt-12")), ckpt.load(str(b / "model.ckpt-12"))
    assert set(want) == set(got) and len(want) >= 9
    for k in want:
        assert np.array_equal(want[k], got[k]), k


def test_fit_stream_accepts_raw_uint8_pixels():
    """``fit_stream`` takes the dataset as raw uint8 pixels (what the IDX files hold) as well as host-normalised
    float32: both train identically (the CPU engine normalises on the host, ``(x - 127.5) / 255`` as ``extract_data``
    does for the reference, ``mpipy.py:215``; the GPU engine does it in the first kernel of the step)."""
    u8, lab = D.synthetic_mnist(96, 5)
    x32 = D.normalize_u8(u8).reshape(96, 28, 28, 1)
    cfg = TrainConfig(model="tiny", batch=16, device="cpu", eval_every=0, log_every=0, dropout=False)
    a = Trainer(cfg)
    la = a.fit_stream(x32, lab, 7)
    b = Trainer(cfg)
    lb = b.fit_stream(u8.reshape(96, 28, 28, 1), lab, 7)
    assert a.global_step == b.global_step == 7
    assert np.allclose(la, lb, rtol=0, atol=0)
    # the window walks on from the global step (the reference's offset arithmetic), also across calls
    lc = a.fit_stream(x32, lab, 3) + []
    d = Trainer(cfg)
    ld = d.fit_stream(x32, lab, 10)
    assert np.allclose(la + lc, ld, rtol=0, atol=0)


def test_shuffle_sampler_is_honoured_by_an_engine_with_a_resident_shard():
    """An engine that keeps the shard on the device walks it with the sequential window itself; any other sampler
    must not be silently replaced by that window: its rows are picked on the host and fed as a batch."""
    import types
    seen = []

    class StubEngine:
        can_fuse, name, flat_params, slots = True, "stub", torch.zeros(4), []

        def attach_dataset(self, shards):
            pass

        def resident_batch(self, sampler, step):
            return None, None

        def fused_step(self, images, labels, lr, step, read_loss=True, peers=True):
            seen.append(None if images is None else np.asarray(labels).copy())
            return 0.0

    u8, lab = D.synthetic_mnist(64, 2)
    x = D.normalize_u8(u8).reshape(64, 28, 28, 1)
    sh = D.Shards(x, lab.astype(np.int64), x[:16], lab[:16].astype(np.int64), x[:16], lab[:16].astype(np.int64))
    for sampler in ("sequential", "shuffle"):
        seen.clear()
        tr = Trainer(TrainConfig(model="tiny", batch=16, device="cpu", sampler=sampler, eval_every=0, log_every=0),
                     engine=StubEngine())
        tr.load_data(sh)
        for _ in range(3):
            tr.step()
        if sampler == "sequential":
            assert all(s is None for s in seen)                     # the engine gathers the window itself
        else:
            want = [sh.train_y[tr.sampler.indices(i)] for i in range(3)]
            assert all(np.array_equal(a, b) for a, b in zip(seen, want))
    tr2 = Trainer(TrainConfig(model="tiny", batch=16, device="cpu", sampler="shuffle", eval_every=0, log_every=0),
                  engine=StubEngine())
    tr2.sampler = D.ShuffleSampler(64, 16)
    with pytest.raises(RuntimeError):
        tr2.step()
