"""End-to-end checks of the GPU engine on one B200: one step against the
bf16-emulating CPU oracle, loss going down, CUDA-graph replay == eager, checkpoint
round trip through the device, and the public Trainer API with host batches."""
import numpy as np
import pytest
import torch

from mpi_tensorflow_b200 import TrainConfig, Trainer
from mpi_tensorflow_b200.models.oracle import Oracle, dropout_keep_mask
from mpi_tensorflow_b200.models.spec import get_spec
from mpi_tensorflow_b200.utils import data as D

pytestmark = pytest.mark.gpu


def _shards(n_rows, seed=0):
    u8, lab = D.synthetic_mnist(n_rows, seed)
    x = D.normalize_u8(u8).reshape(n_rows, 28, 28, 1)
    y = lab.astype(np.int64)
    return D.Shards(x, y, x[:128], y[:128], x[:64], y[:64])


def _cfg(**kw):
    base = dict(model="tutorial512", batch=64, eval_every=0, log_every=0, dropout=False, weight_decay=5e-4)
    base.update(kw)
    return TrainConfig(**base)


@pytest.mark.parametrize("model,batch,dropout", [("tutorial512", 64, False), ("official1024", 100, True)])
def test_one_step_grads_match_bf16_oracle(model, batch, dropout):
    cfg = _cfg(model=model, batch=batch, dropout=dropout, cuda_graph=False)
    tr = Trainer(cfg, device=torch.device("cuda", 0))
    sh = _shards(512)
    tr.load_data(sh)
    eng = tr.engine
    spec = get_spec(model)
    p0 = {k: torch.from_numpy(v) for k, v in eng.get_params().items()}
    loss = eng.compute_grads(None, None, 0)
    grads = eng.flat_grads.cpu()
    orc = Oracle(spec, params=p0, weight_decay=0.0, emulate_bf16=True)
    mask = dropout_keep_mask(spec, batch, cfg.seed, 0) if dropout else None
    x = torch.from_numpy(sh.train_x[:batch])
    y = torch.from_numpy(sh.train_y[:batch])
    rloss, rg, aux = orc.loss_and_grads(x, y, mask)
    l2 = sum(0.5 * float((p0[p.key] ** 2).sum()) for p in spec.params if p.decay)
    assert abs(loss - (rloss + cfg.weight_decay * l2)) < 2e-2 * max(1.0, abs(rloss))
    for p in spec.params:
        g = grads[p.offset:p.offset + p.numel]
        r = rg[p.key].reshape(-1)
        rel = float((g - r).norm() / (r.norm() + 1e-12))
        # Bounds set from measurements over several seeds (profiles/r2/engine_vs_oracle_errors.txt: conv1 <= 2.3e-2,
        # conv2 <= 1.4e-2, fc1 <= 8.9e-3, fc2 <= 1.6e-3), about 1.5x the worst seen.  The error grows towards the input
        # because every layer boundary rounds to bf16 once more, and conv1 runs on fp32 inputs / weights on the device
        # while the oracle rounds them to bf16.  A 2 % systematic error in any layer from conv2 up now fails.
        tol = {"conv1": 3.5e-2, "conv2": 2.2e-2, "fc1": 1.4e-2, "fc2": 4e-3}[p.layer]
        assert rel < tol, f"{p.key}: rel err {rel}"


def test_training_reduces_loss_and_graph_matches_eager():
    sh = _shards(1024)
    runs = {}
    for graph in (False, True):
        tr = Trainer(_cfg(batch=64, lr=0.02, cuda_graph=graph, dropout=True), device=torch.device("cuda", 0))
        tr.load_data(sh)
        losses = [tr.step() for _ in range(40)]
        runs[graph] = (losses, tr.engine.flat_params.clone())
        assert losses[-1] < 0.7 * losses[0], losses
        err = tr.evaluate("test")
        assert 0.0 <= err <= 100.0
        tr.close()
    # atomics in the small wgrad kernels reorder fp32 sums and bf16 rounding amplifies that
    # over 40 steps, so "equal" means close
    a, b = runs[False][1], runs[True][1]
    assert float((a - b).norm() / a.norm()) < 2e-2
    assert abs(runs[False][0][-1] - runs[True][0][-1]) < 0.1


def test_public_api_host_batches_and_checkpoint(tmp_path):
    dev = torch.device("cuda", 0)
    cfg = _cfg(model="official1024", batch=100, optimizer="adam", lr=1e-3, lr_decay=1.0, dropout=True)
    tr = Trainer(cfg, device=dev)
    sh = _shards(400)
    x = torch.from_numpy(sh.train_x).pin_memory()
    y = torch.from_numpy(sh.train_y.astype(np.int32)).pin_memory()
    losses = [tr.step(x[(i % 4) * 100:(i % 4 + 1) * 100], y[(i % 4) * 100:(i % 4 + 1) * 100]) for i in range(24)]
    assert np.isfinite(losses).all() and losses[-1] < losses[0]
    path = str(tmp_path / "model.ckpt")
    tr.save_checkpoint(path)
    before = tr.engine.get_params()
    tr2 = Trainer(cfg, device=dev)
    tr2.load_checkpoint(path)
    after = tr2.engine.get_params()
    for k in before:
        assert np.array_equal(before[k], after[k])
    assert tr2.global_step == 24
    p1 = tr.engine.predict(x[:64]).cpu()
    p2 = tr2.engine.predict(x[:64]).cpu()
    assert torch.allclose(p1, p2, atol=1e-6)
    assert torch.allclose(p1.sum(1), torch.ones(64), atol=1e-4)


def test_wide_preset_runs():
    tr = Trainer(_cfg(model="wide", batch=32, optimizer="adam", lr=1e-4, lr_decay=1.0), device=torch.device("cuda", 0))
    tr.load_data(_shards(256))
    losses = [tr.step() for _ in range(6)]
    assert np.isfinite(losses).all()


def test_fit_stream_matches_stepwise_training():
    """Host-streamed training (H2D prefetch into a ring of input buffers, lagged loss reads) gives the same losses as
    calling step(images, labels) synchronously on the same windows."""
    dev = torch.device("cuda", 0)
    sh = _shards(512)
    x = torch.from_numpy(sh.train_x)
    y = torch.from_numpy(sh.train_y.astype(np.int32))
    cfg = _cfg(batch=64, lr=0.02, dropout=True, lr_decay_rows=10 ** 9)
    a = Trainer(cfg, device=dev)
    la = a.fit_stream(x, y, 20)
    b = Trainer(cfg, device=dev)
    win = [(i * 64) % (512 - 64) for i in range(20)]          # the reference's window on the global step, past one wrap
    lb = [b.step(x[o:o + 64], y[o:o + 64]) for o in win]
    assert len(la) == 20 and a.global_step == 20
    # atomics reorder fp32 sums, so the two trajectories drift apart slowly: the first steps must
    # agree closely, later ones only in trend
    assert np.allclose(la[:3], lb[:3], rtol=5e-3, atol=5e-3)
    assert abs(np.mean(la[-5:]) - np.mean(lb[-5:])) < 0.3 * max(np.mean(lb[-5:]), 0.1)
    assert la[-1] < la[0] and lb[-1] < lb[0]


def test_fit_stream_from_raw_uint8_pixels():
    """fit_stream on the dataset's raw uint8 pixels (a quarter of the bytes over PCIe, normalised by the first kernel of
    the step) trains exactly like fit_stream on the host-normalised float32 images."""
    dev = torch.device("cuda", 0)
    u8, lab = D.synthetic_mnist(512, 3)
    x32 = torch.from_numpy(D.normalize_u8(u8).reshape(512, 28, 28, 1))
    xu8 = torch.from_numpy(u8.reshape(512, 28, 28, 1))
    y = torch.from_numpy(lab.astype(np.int32))
    cfg = _cfg(batch=64, lr=0.02, dropout=True, lr_decay_rows=10 ** 9)
    a = Trainer(cfg, device=dev)
    la = a.fit_stream(x32, y, 12)
    b = Trainer(cfg, device=dev)
    lb = b.fit_stream(xu8, y, 12)
    assert np.allclose(la[:3], lb[:3], rtol=5e-3, atol=5e-3), (la[:3], lb[:3])
    assert lb[-1] < lb[0]
    a.close()
    b.close()


def test_loss_mailbox_matches_device_accumulator():
    """The loss the update kernel mails into pinned host memory is the device accumulator's
    value for exactly that step (step number and value travel in one 8-byte store)."""
    tr = Trainer(_cfg(batch=64), device=torch.device("cuda", 0))
    tr.load_data(_shards(256))
    eng = tr.engine
    spec = get_spec("tutorial512")

    def l2_now():
        p = eng.get_params()
        return sum(0.5 * float((p[q.key].astype(np.float64) ** 2).sum()) for q in spec.params if q.decay)

    for i in range(1, 20):                       # more steps than ring slots: slots are reused
        l2 = l2_now()                            # the loss uses the weights BEFORE the update (mpipy.py:55-58)
        mailed = tr.step()
        torch.cuda.synchronize()
        assert eng.opt_step == i and int(eng.dev_step.item()) == i
        assert mailed == pytest.approx(float(eng.loss_acc.item()) + 5e-4 * l2, rel=2e-6)
        assert eng.train_errors(i) == int(eng.err_acc.item())
    with pytest.raises(RuntimeError):
        eng._await_loss(eng.opt_step + 1, timeout=0.2)      # a step that never ran
    # rewinding the step counter (checkpoint restore) must not match losses mailed before the rewind
    params, slots = eng.get_params(), eng.get_slots()
    eng.set_state(params, slots, 3)
    tr.global_step = 3
    for i in range(4, 8):
        l2 = l2_now()
        mailed = tr.step()
        torch.cuda.synchronize()
        assert eng.opt_step == i and mailed == pytest.approx(float(eng.loss_acc.item()) + 5e-4 * l2, rel=2e-6)


def test_fused_step_loss_equals_unfused_loss(monkeypatch):
    """Same API, one meaning: the loss of the fused graph step (mailed by K1) is the loss compute_grads returns for
    the same batch and weights, CE + 5e-4 * L2 as in the reference (mpipy.py:55-58)."""
    monkeypatch.setenv("MTB_OVERLAP_N1", "1")        # bucketed K1 on a lone GPU (the default; the variable can turn it off)
    dev = torch.device("cuda", 0)
    sh = _shards(256)
    x = torch.from_numpy(sh.train_x[:64])
    y = torch.from_numpy(sh.train_y[:64].astype(np.int32))
    for overlap in (True, False):
        tr = Trainer(_cfg(batch=64, overlap=overlap), device=dev)
        eager = tr.engine.compute_grads(x, y, 0)            # leaves the weights untouched
        fused = tr.step(x, y)
        assert fused == pytest.approx(eager, rel=1e-5), (overlap, fused, eager)
        assert fused > 1.0                                   # the L2 term alone is ~3 for this model
        tr.close()


def test_bucketed_overlap_equals_single_launch(monkeypatch):
    """The hidden fc bucket + closing bucket update exactly what the single whole-buffer launch updates."""
    monkeypatch.setenv("MTB_OVERLAP_N1", "1")
    dev = torch.device("cuda", 0)
    sh = _shards(512)
    out = {}
    for overlap in (True, False):
        tr = Trainer(_cfg(batch=64, lr=0.02, overlap=overlap, optimizer="adam", lr_decay=1.0), device=dev)
        tr.load_data(sh)
        losses = [tr.step() for _ in range(5)]
        out[overlap] = (tr.engine.flat_params.clone(), [s.clone() for s in tr.engine.slots], losses)
        assert tr.engine.launches_per_step == (13 if overlap else 12)
        tr.close()
    # the only difference between the two is the order in which fp32 atomics of the backward kernels land
    a, b = out[True], out[False]
    assert float((a[0] - b[0]).norm() / b[0].norm()) < 1e-3
    assert np.allclose(a[2], b[2], rtol=1e-3)


def test_bf16_gradient_wire_format_single_gpu():
    """comm_dtype='bf16': the FC1 weight gradient is produced and consumed as bf16 (2 bytes per element on the
    wire); everything downstream of the load stays fp32, so training follows the fp32-wire run closely."""
    dev = torch.device("cuda", 0)
    sh = _shards(512)
    runs = {}
    for wire in ("fp32", "bf16"):
        tr = Trainer(_cfg(batch=64, lr=0.02, comm_dtype=wire), device=dev)
        tr.load_data(sh)
        runs[wire] = ([tr.step() for _ in range(10)], tr.engine.flat_params.clone())
        tr.close()
    assert runs["bf16"][0][-1] < runs["bf16"][0][0]
    assert float((runs["bf16"][1] - runs["fp32"][1]).norm() / runs["fp32"][1].norm()) < 5e-3
    with pytest.raises(ValueError):
        Trainer(_cfg(dtype="fp32", comm_dtype="bf16"), device=dev)


@pytest.mark.parametrize("size,B", [(256, 64), (200, 64), (130, 100)])
def test_graph_captured_eval_matches_eager_predict(size, B):
    """eval_in_batches over a resident split, whole loop in one CUDA graph (ragged tail included), gives the
    probabilities the eager per-batch predict gives; the second call replays the graph."""
    dev = torch.device("cuda", 0)
    sh = _shards(512)
    tr = Trainer(_cfg(batch=64, eval_batch=B), device=dev)
    tr.load_data(sh)
    for _ in range(3):
        tr.step()
    data = sh.train_x[:size]
    got = tr.eval_in_batches(data)
    assert got.shape == (size, 10) and np.allclose(got.sum(1), 1.0, atol=1e-4)
    ref = np.concatenate([tr.engine.predict(torch.from_numpy(data[i:i + 1].repeat(B, 0))).cpu().numpy()[:1]
                          for i in range(0, size, 37)])
    assert np.allclose(got[::37], ref, atol=1e-5)
    tr.step()                                                # weights move: the replay must see the new ones
    again = tr.eval_in_batches(data)
    assert len(tr.engine._eval_cache) == 1 and not np.array_equal(again, got)
    ref2 = tr.engine.predict(torch.from_numpy(data[:B])).cpu().numpy()
    assert np.allclose(again[:B], ref2, atol=1e-5)
    tr.close()


def test_timed_step_reports_phases(monkeypatch):
    monkeypatch.setenv("MTB_OVERLAP_N1", "1")
    tr = Trainer(_cfg(batch=64), device=torch.device("cuda", 0))
    tr.load_data(_shards(256))
    tr.step()
    ph = tr.engine.timed_step()
    assert tr.engine.opt_step == 2 and int(tr.engine.dev_step.item()) == 2
    for k in ("fwd", "bwd_fc", "bwd_conv", "k1_hidden", "k1_tail", "step"):
        assert ph[k] > 0.0, (k, ph)
    assert ph["step"] >= ph["fwd"] + ph["bwd_fc"] + ph["bwd_conv"]
    tr.close()
