"""CPU reference model tier: hand-written math vs autograd."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from mpi_tensorflow_b200.models import oracle as O
from mpi_tensorflow_b200.models.spec import get_spec, make_spec
from mpi_tensorflow_b200.utils import philox

SMALL = make_spec("tiny", 4, 8, 16, 0.5)


def _torch_forward(P, x, keep=None, keep_prob=0.5):
    t = x.permute(0, 3, 1, 2)
    h = F.relu(F.conv2d(t, P["conv1_weight"].permute(3, 2, 0, 1), P["conv1_bias"], padding=2))
    h = F.max_pool2d(h, 2)
    h = F.relu(F.conv2d(h, P["conv2_weight"].permute(3, 2, 0, 1), P["conv2_bias"], padding=2))
    h = F.max_pool2d(h, 2)
    h = h.permute(0, 2, 3, 1).reshape(x.shape[0], -1)
    h = F.relu(h @ P["fc1_weight"] + P["fc1_bias"])
    if keep is not None:
        h = h * keep.float() / keep_prob
    return h @ P["fc2_weight"] + P["fc2_bias"]


def test_param_counts_match_reference():
    assert get_spec("tutorial512").num_params == 1663370
    assert get_spec("official1024").num_params == 3274634
    assert get_spec("wide").num_params == 52248074
    s = get_spec("tutorial512")
    assert s.param("fc1_weight").shape == (3136, 512) and s.flat_dim == 3136


def test_forward_backward_match_autograd():
    torch.manual_seed(0)
    spec = SMALL
    orc = O.Oracle(spec, seed=1, weight_decay=5e-4)
    x = torch.randn(6, 28, 28, 1) * 0.3
    y = torch.randint(0, 10, (6,))
    keep = O.dropout_keep_mask(spec, 6, 1, 0)
    loss, g, aux = orc.loss_and_grads(x, y, keep)
    P = {k: v.clone().requires_grad_(True) for k, v in orc.params.items()}
    logits = _torch_forward(P, x, keep, spec.keep_prob)
    ref = F.cross_entropy(logits, y) + 5e-4 * sum(0.5 * (P[p.key] ** 2).sum() for p in spec.params if p.decay)
    ref.backward()
    assert abs(loss - float(ref)) < 1e-5
    assert torch.allclose(aux["logits"], logits.detach(), atol=1e-5)
    for k in P:
        assert torch.allclose(g[k], P[k].grad, atol=2e-5, rtol=1e-4), k


def test_pool_first_max_wins_and_bwd_routes():
    x = torch.zeros(1, 2, 2, 1)
    v, i = O.maxpool2x2(x)                       # all ties -> index 0
    assert int(i) == 0
    d = O.maxpool2x2_bwd(torch.ones(1, 1, 1, 1), i)
    assert d.reshape(-1).tolist() == [1, 0, 0, 0]
    x = torch.tensor([[1., 5.], [7., 3.]]).reshape(1, 2, 2, 1)
    v, i = O.maxpool2x2(x)
    assert float(v) == 7 and int(i) == 2         # 2*dy+dx


def test_lr_schedule_staircase():
    # mpipy.py:60-64: 0.01 * 0.95^floor(step*64/rows)
    rows = 55000
    assert O.learning_rate(0, 64, rows) == pytest.approx(0.01)
    assert O.learning_rate(859, 64, rows) == pytest.approx(0.01)          # 54976 < 55000
    assert O.learning_rate(860, 64, rows) == pytest.approx(0.0095)
    assert O.learning_rate(1718, 64, rows) == pytest.approx(0.01 * 0.95 ** 1)
    assert O.learning_rate(1719, 64, rows) == pytest.approx(0.01 * 0.95 ** 2)


def test_optimizers_match_torch():
    spec = SMALL
    for opt in ("sgd", "momentum", "adam"):
        orc = O.Oracle(spec, seed=2, weight_decay=0.0)
        P = {k: v.clone().requires_grad_(True) for k, v in orc.params.items()}
        if opt == "sgd":
            topt = torch.optim.SGD(P.values(), lr=0.05)
        elif opt == "momentum":
            topt = torch.optim.SGD(P.values(), lr=0.05, momentum=0.9)   # same rule for constant lr
        else:
            topt = torch.optim.Adam(P.values(), lr=0.05, eps=1e-8)
        torch.manual_seed(1)
        for _ in range(3):
            # |g| bounded away from 0: TF's epsilon sits outside the bias correction (eps-hat form)
            g = {k: torch.randn_like(v).sign() * (0.5 + torch.rand_like(v)) for k, v in orc.params.items()}
            orc.apply(g, 0.05, optimizer=opt)
            for k in P:
                P[k].grad = g[k].clone()
            topt.step()
        for k in P:
            assert torch.allclose(orc.params[k], P[k].detach(), atol=1e-5, rtol=1e-4), (opt, k)


def test_error_rate():
    pred = np.eye(10, dtype=np.float32)[[1, 2, 3, 4]]
    assert O.error_rate(pred, np.array([1, 2, 0, 4])) == pytest.approx(25.0)


def test_training_reduces_loss_on_synthetic():
    from mpi_tensorflow_b200.utils import data as D
    spec = SMALL
    orc = O.Oracle(spec, seed=1, weight_decay=0.0)
    u8, lab = D.synthetic_mnist(256, seed=0)
    x = torch.from_numpy(D.normalize_u8(u8)).reshape(-1, 28, 28, 1)
    y = torch.from_numpy(lab.astype(np.int64))
    first = last = None
    for s in range(60):
        i = (s * 32) % (256 - 32)
        loss, _ = orc.train_step(x[i:i + 32], y[i:i + 32], 0.05, optimizer="momentum")
        first = first if first is not None else loss
        last = loss
    assert last < 0.5 * first


def test_philox_known_answer_and_streams():
    # Random123 known-answer test for philox4x32-10, counter=0 key=0
    out = philox.philox4x32(np.zeros(1, np.uint32), 0, 0, 0, 0, 0)
    assert [int(o[0]) for o in out] == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    out = philox.philox4x32(np.full(1, 0xffffffff, np.uint32), 0xffffffff, 0xffffffff, 0xffffffff,
                            0xffffffff, 0xffffffff)
    assert [int(o[0]) for o in out] == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    m = philox.dropout_mask(100000, 0.6, 1, 7)
    assert abs(m.mean() - 0.6) < 0.01
    assert not np.array_equal(m, philox.dropout_mask(100000, 0.6, 1, 8))
    z = philox.truncated_normal(200000, 0.1, 1, 0)
    assert np.abs(z).max() <= 0.2 + 1e-6 and abs(z.std() - 0.088) < 0.002   # std of N(0,1)|<2 = 0.8796
    a = philox.uniform01(10, 1, 0, 3, offset=6)
    b = philox.uniform01(16, 1, 0, 3)
    assert np.array_equal(a, b[6:])


def test_emulate_bf16_close_to_fp32():
    spec = SMALL
    x = torch.randn(4, 28, 28, 1) * 0.3
    a = O.Oracle(spec, seed=1).forward(x)
    b = O.Oracle(spec, seed=1, emulate_bf16=True).forward(x)
    assert torch.allclose(a, b, atol=0.05, rtol=0.05)
