"""Per-kernel numerics on a B200 (SURVEY §4.2 "Unit — each sm_100a kernel").

Every test calls one C-ABI entry point of ``libmtb200.so`` and compares it with a plain
PyTorch fp32 restatement of the same op (inputs rounded to bf16 where the kernel
consumes bf16, so the tolerance only absorbs accumulation order and output rounding).
"""
import ctypes as C

import numpy as np
import pytest
import torch

from mpi_tensorflow_b200.models import oracle as O
from mpi_tensorflow_b200.utils import data as D
from mpi_tensorflow_b200.utils import philox

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def N():
    from mpi_tensorflow_b200.ops import native
    native.lib()
    return native


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


def _bf(t):
    return t.to(torch.bfloat16)


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


# ----------------------------------------------------------------------------- GEMM
GEMM_SHAPES = [
    # M, N, K, bn, splits
    (128, 64, 64, 64, 1),
    (128, 128, 256, 128, 1),
    (100, 1024, 3136, 64, 1),        # fc1 fprop, batch 100
    (1024, 1024, 3136, 128, 1),      # fc1 fprop, batch 1024
    (3136, 1024, 100, 128, 1),       # fc1 wgrad, K = batch (ragged, TMA zero fill)
    (100, 3136, 1024, 128, 1),       # fc1 dgrad
    (19600, 64, 800, 64, 1),         # conv2 fprop as GEMM
    (800, 64, 19600, 64, 16),        # conv2 wgrad, split-K
    (19600, 800, 64, 128, 1),        # conv2 dgrad
    (300, 200, 136, 64, 3),          # ragged everything + split-K
    (1024, 3136, 1024, 256, 1),      # wide tile (BN = 256): fc1 dgrad shape
    (3136, 1024, 200, 256, 1),       # wide tile, ragged K
]


@pytest.mark.parametrize("a_mn", [False, True])
@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("shape", GEMM_SHAPES)
def test_gemm_bf16_all_layouts(N, dev, shape, a_mn, b_mn):
    M, Nn, K, bn, splits = shape
    if (a_mn and M % 8) or (b_mn and Nn % 8) or ((not a_mn or not b_mn) and K % 8):
        pytest.skip("row pitch must be a multiple of 16 bytes")
    g = torch.Generator(device="cpu").manual_seed(M * 7 + Nn * 3 + K)
    A = _bf(torch.randn(M, K, generator=g)).to(dev)
    B = _bf(torch.randn(K, Nn, generator=g)).to(dev)
    ref = A.float() @ B.float()
    A_st = A.t().contiguous() if a_mn else A.contiguous()            # MN-major stored [K, M]
    B_st = B.contiguous() if b_mn else B.t().contiguous()            # K-major stored [N, K]
    out = torch.full((M, Nn), float("nan"), device=dev)
    ws = cnt = None
    if splits > 1:
        wsb, cb = N.gemm_scratch_sizes(M, Nn, bn, splits)
        ws = torch.zeros(wsb, dtype=torch.uint8, device=dev)
        cnt = torch.zeros(cb, dtype=torch.uint8, device=dev)
    for _ in range(2):                                               # second call checks the self-cleaning counters
        out.fill_(float("nan"))
        N.gemm_bf16(A_st, B_st, out, M, Nn, K, a_mn, b_mn, bn=bn, splits=splits, workspace=ws, counters=cnt)
        torch.cuda.synchronize()
        assert torch.isfinite(out).all()
        assert _rel(out, ref) < 2e-5 * max(1.0, K / 2048) + 1e-5


def test_gemm_epilogue_bias_relu_bf16_out(N, dev):
    M, Nn, K = 200, 512, 256
    A = _bf(torch.randn(M, K)).to(dev)
    W = _bf(torch.randn(K, Nn) * 0.1).to(dev)
    bias = torch.randn(Nn, device=dev)
    ref = torch.relu(A.float() @ W.float() + bias)
    out = torch.zeros(M, Nn, dtype=torch.bfloat16, device=dev)
    N.gemm_bf16(A, W, out, M, Nn, K, False, True, bias=bias, relu=True, bn=64)
    torch.cuda.synchronize()
    assert _rel(out, ref) < 4e-3
    assert float((out.float() - ref).abs().max()) < 0.05


def test_gemm_dropout_matches_philox_stream(N, dev):
    M, Nn, K, keep, seed, step = 96, 512, 128, 0.6, 7, 5
    A = _bf(torch.rand(M, K) + 0.1).to(dev)
    W = _bf(torch.rand(K, Nn) * 0.1 + 0.01).to(dev)
    step_t = torch.tensor([step], dtype=torch.int64, device=dev)
    out = torch.zeros(M, Nn, device=dev)
    N.gemm_bf16(A, W, out, M, Nn, K, False, True, keep_prob=keep, seed=seed, step_ptr=step_t, bn=64)
    torch.cuda.synchronize()
    mask = torch.from_numpy(philox.dropout_mask(M * Nn, keep, seed, step).reshape(M, Nn)).to(dev)
    ref = (A.float() @ W.float()) * mask / keep
    assert torch.equal(out != 0, mask)
    assert _rel(out, ref) < 1e-5


# ----------------------------------------------------------------------------- conv1
@pytest.mark.parametrize("B,C1,tc", [(3, 32, False), (70, 32, False), (5, 128, False), (1, 32, True), (7, 32, True),
                                       (150, 32, True), (299, 32, True)])
def test_conv1_fwd_and_wgrad(N, dev, B, C1, tc):
    """CUDA-core (any C1) and tensor-core (C1 = 32, bf16 operands) versions against the fp32 oracle."""
    x = (torch.rand(B, 28, 28) - 0.5).to(dev)
    w = (torch.randn(5, 5, 1, C1) * 0.1).to(dev)
    bias = (torch.randn(C1) * 0.1).to(dev)
    out = torch.zeros(B, 14, 14, C1, dtype=torch.bfloat16, device=dev)
    idx = torch.full((B, 14, 14, C1), 9, dtype=torch.uint8, device=dev)
    if tc:
        N.call("mtb_conv1_fwd_tc", x.data_ptr(), w.data_ptr(), bias.data_ptr(), out.data_ptr(), idx.data_ptr(), B)
    else:
        N.call("mtb_conv1_fwd", x.data_ptr(), w.data_ptr(), bias.data_ptr(), out.data_ptr(), idx.data_ptr(), B, C1)
    torch.cuda.synchronize()
    xr = x.cpu().reshape(B, 28, 28, 1)
    if tc:      # the tensor-core kernel multiplies bf16-rounded pixels and weights
        a = torch.relu(O.conv5x5(O.bf16_round(xr), O.bf16_round(w.cpu())) + bias.cpu())
    else:
        a = torch.relu(O.conv5x5(xr, w.cpu()) + bias.cpu())
    p, i = O.maxpool2x2(a)
    assert float((out.cpu().float() - p).abs().max()) < 1e-2 * max(1.0, float(p.abs().max()))
    # argmax may differ only where two window entries tie numerically
    agree = (idx.cpu().long() == i.long()).float().mean()
    assert agree > 0.999
    assert int(idx.max()) <= 3
    # wgrad: gradient of sum(dP * pool(relu(conv))) w.r.t. w and bias
    dP = _bf(torch.randn(B, 14, 14, C1)).to(dev)
    dW = torch.zeros(25, C1, device=dev)
    db = torch.zeros(C1, device=dev)
    if tc:
        N.call("mtb_conv1_wgrad_tc", x.data_ptr(), dP.data_ptr(), out.data_ptr(), idx.data_ptr(), dW.data_ptr(),
               db.data_ptr(), B)
    else:
        N.call("mtb_conv1_wgrad", x.data_ptr(), dP.data_ptr(), out.data_ptr(), idx.data_ptr(), dW.data_ptr(),
               db.data_ptr(), B, C1)
    torch.cuda.synchronize()
    g = dP.cpu().float() * (out.cpu().float() > 0).float()
    da = O.maxpool2x2_bwd(g, idx.cpu().long())
    ref_dw = O.conv5x5_wgrad(O.bf16_round(xr) if tc else xr, da).reshape(25, C1)
    ref_db = da.sum((0, 1, 2))
    assert _rel(dW.cpu(), ref_dw) < 1e-4
    assert _rel(db.cpu(), ref_db) < 1e-4


# ----------------------------------------------------------------------------- pool / im2col
@pytest.mark.parametrize("B,H,Cc", [(4, 14, 64), (3, 28, 32), (2, 14, 256)])
def test_maxpool_and_unpool(N, dev, B, H, Cc):
    a = _bf(torch.randn(B, H, H, Cc)).to(dev)
    out = torch.zeros(B, H // 2, H // 2, Cc, dtype=torch.bfloat16, device=dev)
    idx = torch.zeros(B, H // 2, H // 2, Cc, dtype=torch.uint8, device=dev)
    N.call("mtb_maxpool2x2_fwd", a.data_ptr(), out.data_ptr(), idx.data_ptr(), B, H, H, Cc)
    torch.cuda.synchronize()
    p, i = O.maxpool2x2(a.cpu().float())
    assert torch.equal(out.cpu().float(), p)
    assert torch.equal(idx.cpu().long(), i.long())
    dP = _bf(torch.randn(B, H // 2, H // 2, Cc)).to(dev)
    dA = torch.full((B, H, H, Cc), 7.0, dtype=torch.bfloat16, device=dev)
    N.call("mtb_unpool_relu_bwd", dP.data_ptr(), out.data_ptr(), idx.data_ptr(), dA.data_ptr(), B, H, H, Cc)
    torch.cuda.synchronize()
    ref = O.maxpool2x2_bwd(dP.cpu().float() * (p > 0).float(), i)
    assert torch.equal(dA.cpu().float(), ref)


@pytest.mark.parametrize("B,H,Cc", [(3, 14, 32), (2, 14, 128)])
def test_im2col_col2im(N, dev, B, H, Cc):
    x = _bf(torch.randn(B, H, H, Cc)).to(dev)
    col = torch.zeros(B * H * H, 25 * Cc, dtype=torch.bfloat16, device=dev)
    N.call("mtb_im2col_5x5", x.data_ptr(), col.data_ptr(), B, H, H, Cc)
    torch.cuda.synchronize()
    ref = O._im2col(x.cpu().float())
    assert torch.equal(col.cpu().float().reshape(ref.shape), ref)
    dcol = _bf(torch.randn(B * H * H, 25 * Cc)).to(dev)
    dx = torch.zeros(B, H, H, Cc, dtype=torch.bfloat16, device=dev)
    N.call("mtb_col2im_5x5", dcol.data_ptr(), dx.data_ptr(), B, H, H, Cc)
    torch.cuda.synchronize()
    refdx = O._col2im(dcol.cpu().float().reshape(ref.shape), (B, H, H, Cc))
    assert float((dx.cpu().float() - refdx).abs().max()) < 0.05 * float(refdx.abs().max())
    assert _rel(dx.cpu(), refdx) < 5e-3


# ----------------------------------------------------------------------------- loss
@pytest.mark.parametrize("B,F", [(64, 512), (100, 1024), (7, 4096)])
def test_fc2_softmax_ce_fwd_bwd(N, dev, B, F):
    NC = 10
    h = _bf(torch.relu(torch.randn(B, F))).to(dev)
    W = (torch.randn(F, NC) * 0.1).to(dev)
    bias = (torch.randn(NC) * 0.1).to(dev)
    y = torch.randint(0, NC, (B,), dtype=torch.int32).to(dev)
    logits, probs, dl = (torch.zeros(B, NC, device=dev) for _ in range(3))
    loss = torch.zeros(1, device=dev)
    err = torch.zeros(1, dtype=torch.int32, device=dev)
    N.call("mtb_fc2_softmax_ce", h.data_ptr(), W.data_ptr(), bias.data_ptr(), y.data_ptr(), logits.data_ptr(),
           probs.data_ptr(), dl.data_ptr(), loss.data_ptr(), err.data_ptr(), B, F, NC)
    torch.cuda.synchronize()
    rl = h.cpu().float() @ W.cpu() + bias.cpu()
    ce, rdl, rp, rerr = O.softmax_ce(rl, y.cpu().long())
    assert torch.allclose(logits.cpu(), rl, atol=1e-3, rtol=1e-4)
    assert torch.allclose(probs.cpu(), rp, atol=1e-5, rtol=1e-3)
    assert torch.allclose(dl.cpu(), rdl, atol=1e-6, rtol=1e-3)
    assert abs(float(loss) - float(ce)) < 1e-4 * max(1.0, float(ce))
    assert int(err) == int(rerr)
    # backward
    dh = torch.zeros(B, F, dtype=torch.bfloat16, device=dev)
    N.call("mtb_fc2_dgrad", dl.data_ptr(), W.data_ptr(), h.data_ptr(), dh.data_ptr(), B, F, NC, 2.0)
    dW = torch.zeros(F, NC, device=dev)
    db = torch.zeros(NC, device=dev)
    N.call("mtb_fc2_wgrad", h.data_ptr(), dl.data_ptr(), dW.data_ptr(), db.data_ptr(), B, F, NC)
    cs = torch.zeros(F, device=dev)
    N.call("mtb_colsum_bf16", dh.data_ptr(), cs.data_ptr(), B, F)
    torch.cuda.synchronize()
    rdh = (rdl @ W.cpu().t()) * (h.cpu().float() > 0).float() * 2.0
    assert _rel(dh.cpu(), rdh) < 5e-3
    assert _rel(dW.cpu(), h.cpu().float().t() @ rdl) < 1e-4
    assert _rel(db.cpu(), rdl.sum(0)) < 1e-4
    assert _rel(cs.cpu(), dh.cpu().float().sum(0)) < 1e-4
    # engine variants: same math, shared-memory weights / one fused backward pass
    l2, p2, d2 = (torch.zeros(B, NC, device=dev) for _ in range(3))
    loss2 = torch.zeros(1, device=dev)
    err2 = torch.zeros(1, dtype=torch.int32, device=dev)
    N.call("mtb_fc2_softmax_ce_fast", h.data_ptr(), W.data_ptr(), bias.data_ptr(), y.data_ptr(), l2.data_ptr(),
           p2.data_ptr(), d2.data_ptr(), loss2.data_ptr(), err2.data_ptr(), B, F, NC)
    dh2 = torch.zeros(B, F, dtype=torch.bfloat16, device=dev)
    dW2, db2, db1 = torch.zeros(F, NC, device=dev), torch.zeros(NC, device=dev), torch.zeros(F, device=dev)
    N.call("mtb_fc2_bwd_fused", h.data_ptr(), dl.data_ptr(), W.data_ptr(), dh2.data_ptr(), dW2.data_ptr(),
           db2.data_ptr(), db1.data_ptr(), B, F, NC, 2.0)
    torch.cuda.synchronize()
    assert torch.allclose(l2, logits, atol=1e-4, rtol=1e-4) and torch.allclose(d2, dl, atol=1e-7, rtol=1e-3)
    assert abs(float(loss2) - float(loss)) < 1e-4 and int(err2) == int(err)
    assert _rel(dh2.cpu(), rdh) < 5e-3
    assert _rel(dW2.cpu(), h.cpu().float().t() @ rdl) < 1e-4
    assert _rel(db2.cpu(), rdl.sum(0)) < 1e-4
    assert _rel(db1.cpu(), dh2.cpu().float().sum(0)) < 1e-4


# ----------------------------------------------------------------------------- init / data / misc
def test_trunc_normal_matches_host_philox(N, dev):
    n = 10007
    dst = torch.zeros(n, device=dev)
    N.call("mtb_init_trunc_normal", dst.data_ptr(), n, 0.1, 1, 4)
    torch.cuda.synchronize()
    ref = philox.truncated_normal(n, 0.1, 1, 4)
    assert np.allclose(dst.cpu().numpy(), ref, atol=2e-6)
    assert float(dst.abs().max()) <= 0.2 + 1e-6


def test_synthetic_batch_matches_host(N, dev):
    n, off = 37, 1000
    img = torch.zeros(n, 784, device=dev)
    lab = torch.zeros(n, dtype=torch.int32, device=dev)
    N.call("mtb_synthetic_batch", img.data_ptr(), lab.data_ptr(), n, off, 3, 1)
    torch.cuda.synchronize()
    u8, rl = D.synthetic_mnist(n, 3, offset=off)
    ref = D.normalize_u8(u8).reshape(n, 784)
    assert np.array_equal(lab.cpu().numpy(), rl.astype(np.int32))
    diff = np.abs(img.cpu().numpy() - ref)
    # expf on device vs numpy may move a pixel across a uint8 boundary very rarely
    assert (diff > 1e-6).mean() < 1e-3 and diff.max() <= 1.0 / 255 + 1e-6


def test_batch_window_and_casts_and_l2(N, dev):
    rows, B = 500, 64
    data = torch.randn(rows, 784, device=dev)
    labels = torch.randint(0, 10, (rows,), dtype=torch.int32, device=dev)
    x = torch.zeros(B, 784, device=dev)
    y = torch.zeros(B, dtype=torch.int32, device=dev)
    for step in (0, 3, 9):
        st = torch.tensor([step], dtype=torch.int64, device=dev)
        N.call("mtb_batch_window", data.data_ptr(), labels.data_ptr(), st.data_ptr(), rows, B, x.data_ptr(),
               y.data_ptr())
        off = (step * B) % (rows - B)                                  # mpipy.py:80
        assert torch.equal(x, data[off:off + B]) and torch.equal(y, labels[off:off + B])
    src = torch.randn(1001, device=dev)
    b = torch.zeros(1001, dtype=torch.bfloat16, device=dev)
    N.call("mtb_cast_f32_to_bf16", src.data_ptr(), b.data_ptr(), 1001)
    assert torch.equal(b, src.to(torch.bfloat16))
    f = torch.zeros(1001, device=dev)
    N.call("mtb_cast_bf16_to_f32", b.data_ptr(), f.data_ptr(), 1001)
    assert torch.equal(f, b.float())
    acc = torch.zeros(1, device=dev)
    N.call("mtb_l2_half_sumsq", src.data_ptr(), 1001, acc.data_ptr())
    assert abs(float(acc) - 0.5 * float((src.double() ** 2).sum())) < 1e-3
    u8 = torch.randint(0, 256, (999,), dtype=torch.uint8, device=dev)
    o = torch.zeros(999, device=dev)
    N.call("mtb_normalize_u8_dev", u8.data_ptr(), o.data_ptr(), 999)
    assert torch.allclose(o, (u8.float() - 127.5) / 255.0)


# ----------------------------------------------------------------------------- K1 on one GPU
@pytest.mark.parametrize("opt", ["sgd", "momentum", "adam"])
def test_fused_update_single_gpu_matches_oracle(N, dev, opt):
    n = 4096 + 512
    g = torch.randn(n)
    w = torch.randn(n)
    m0, v0 = torch.randn(n) * 0.1, torch.rand(n) * 0.1
    step = 3
    lr0, decay, batch, rows, wd = 0.01, 0.95, 64, 100, 5e-4
    lo, hi = 1024, 3072                                              # range that gets the L2 term
    W, G = w.clone().to(dev), g.clone().to(dev)
    M, V = m0.clone().to(dev), v0.clone().to(dev)
    S = torch.zeros(n, dtype=torch.bfloat16, device=dev)
    seq = torch.zeros(512, dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    st = torch.tensor([step], dtype=torch.int64, device=dev)
    a = N.FusedArgs()
    a.world, a.rank, a.algo = 1, 0, 0
    a.optimizer = {"sgd": 0, "momentum": 1, "adam": 2}[opt]
    a.grads[0], a.params[0], a.shadow[0] = G.data_ptr(), W.data_ptr(), S.data_ptr()
    a.slot_m, a.slot_v = M.data_ptr(), V.data_ptr()
    a.seq, a.status = seq.data_ptr(), status.data_ptr()
    a.offset, a.count, a.scale = 0, n, 1.0
    a.lr0, a.lr_decay, a.lr_override = lr0, decay, -1.0
    a.batch, a.decay_rows, a.step_ptr = batch, rows, st.data_ptr()
    a.momentum, a.beta1, a.beta2, a.eps, a.wd = 0.9, 0.9, 0.999, 1e-8, wd
    a.n_ranges = 1
    a.range_begin[0], a.range_end[0] = lo, hi
    N.call("mtb_fused_allreduce_update", C.byref(a))
    torch.cuda.synchronize()
    lr = O.learning_rate(step, batch, rows, lr0, decay)
    gg = g.clone()
    gg[lo:hi] += wd * w[lo:hi]
    if opt == "sgd":
        ref = w - lr * gg
    elif opt == "momentum":
        mm = 0.9 * m0 + gg
        ref = w - lr * mm
        assert torch.allclose(M.cpu(), mm, atol=1e-6, rtol=1e-5)
    else:
        t = step + 1
        mm = 0.9 * m0 + 0.1 * gg
        vv = 0.999 * v0 + 0.001 * gg * gg
        lr_t = lr * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
        ref = w - lr_t * mm / (vv.sqrt() + 1e-8)
        assert torch.allclose(V.cpu(), vv, atol=1e-6, rtol=1e-5)
    assert torch.allclose(W.cpu(), ref, atol=2e-6, rtol=2e-5)
    assert torch.equal(S.cpu(), W.cpu().to(torch.bfloat16))
    assert int(status) == 0


# ----------------------------------------------------------------------------- descriptor probe
@pytest.mark.parametrize("mode,rb", [(0, 128), (0, 64), (1, 128)])
def test_smem_descriptor_row_shift(N, dev, mode, rb):
    """A tcgen05 operand may start at any row of a swizzled tile with base-offset 0: the
    property the shift-GEMM convolutions are built on."""
    rows = 256 if mode == 0 else 96
    if mode == 0:
        A = _bf(torch.randn(rows, rb // 2)).to(dev)
        B = _bf(torch.randn(64, rb // 2)).to(dev)
    else:
        A = _bf(torch.randn(rows, 128)).to(dev)
        B = _bf(torch.randn(64, 32)).to(dev)
    for shift in (0, 1, 5, 18, 37, 64):
        out = torch.zeros(128, 64, device=dev)
        N.call("mtb_debug_shift_mma", A.data_ptr(), B.data_ptr(), out.data_ptr(), mode, rb, rows, shift, 0)
        torch.cuda.synchronize()
        ref = (A[shift:shift + 128].float() @ B.float().t()) if mode == 0 else \
            (A[shift:shift + 32].float().t() @ B.float().t())
        assert _rel(out, ref) < 1e-5


# ----------------------------------------------------------------------------- conv2 implicit GEMM
@pytest.mark.parametrize("B", [1, 3, 150, 301])
def test_conv2_fprop_pool_implicit_gemm(N, dev, B):
    x = _bf(torch.relu(torch.randn(B, 14, 14, 32))).to(dev)
    w = _bf(torch.randn(5, 5, 32, 64) * 0.05).to(dev)
    bias = (torch.randn(64) * 0.1).to(dev)
    out = torch.zeros(B, 7, 7, 64, dtype=torch.bfloat16, device=dev)
    idx = torch.full((B, 7, 7, 64), 9, dtype=torch.uint8, device=dev)
    N.call("mtb_conv2_fprop_pool", x.data_ptr(), w.data_ptr(), bias.data_ptr(), out.data_ptr(), idx.data_ptr(), B)
    torch.cuda.synchronize()
    a = torch.relu(O.conv5x5(x.cpu().float(), w.cpu().float()) + bias.cpu())
    a = a.to(torch.bfloat16).float()                      # the kernel pools the bf16-rounded activation
    p, i = O.maxpool2x2(a)
    got = out.cpu().float()
    assert float((got - p).abs().max()) < 0.03 * max(1.0, float(p.abs().max()))
    assert _rel(got, p) < 5e-3
    assert ((idx.cpu().long() == i.long()).float().mean()) > 0.995
    assert int(idx.max()) <= 3


# Implicit-GEMM convolution through the GEMM engine's 4-D TMA producer (any channel counts, bf16 or fp32 / TF32):
# the path of the fp32 configuration and of the widened model.  Reference: the plain fp32 oracle.
_IGEMM_CASES = [(3, 32, 64, True), (100, 32, 64, True), (2, 64, 128, True), (2, 128, 256, False), (37, 64, 64, False)]


def _igemm_scratch(N, dev, cin, cout, splits):
    bn = 256 if cout % 256 == 0 else (128 if cout % 128 == 0 else 64)
    ws, cn = N.gemm_scratch_sizes(25 * cin, cout, bn, splits)
    return (torch.zeros(max(ws, 16), dtype=torch.uint8, device=dev),
            torch.zeros(max(cn, 16), dtype=torch.uint8, device=dev))


@pytest.mark.parametrize("B,cin,cout,tf32", _IGEMM_CASES)
def test_conv5x5_implicit_gemm_fprop_dgrad_wgrad(N, dev, B, cin, cout, tf32):
    dt = torch.float32 if tf32 else torch.bfloat16
    g = torch.Generator().manual_seed(B * 1000 + cin)
    x = torch.relu(torch.randn(B, 14, 14, cin, generator=g)).to(dt)
    w = (torch.randn(5, 5, cin, cout, generator=g) * 0.05).to(dt)
    dy = (torch.randn(B, 14, 14, cout, generator=g) * (torch.rand(B, 14, 14, cout, generator=g) < 0.5)).to(dt)
    bias = torch.randn(cout, generator=g) * 0.1
    xd, wd, dyd, bd = x.to(dev), w.to(dev), dy.to(dev), bias.to(dev)
    tol = 2e-3 if tf32 else 5e-3                  # TF32 operands (10-bit mantissa) / bf16 output rounding
    # fprop + bias + ReLU
    out = torch.full((B, 14, 14, cout), float("nan"), dtype=dt, device=dev)
    N.call("mtb_conv5x5_gemm", 0, xd.data_ptr(), wd.data_ptr(), out.data_ptr(), B, cin, cout, int(tf32), bd.data_ptr(),
           1, 1, None, None)
    torch.cuda.synchronize()
    ref = torch.relu(O.conv5x5(x.float(), w.float()) + bias)
    assert torch.isfinite(out).all()
    assert _rel(out.cpu(), ref) < tol
    # dgrad
    dx = torch.full((B, 14, 14, cin), float("nan"), dtype=dt, device=dev)
    N.call("mtb_conv5x5_gemm", 1, dyd.data_ptr(), wd.data_ptr(), dx.data_ptr(), B, cin, cout, int(tf32), None, 0, 1,
           None, None)
    torch.cuda.synchronize()
    ref = O.conv5x5_dgrad(dy.float(), w.float(), (B, 14, 14, cin))
    assert torch.isfinite(dx).all()
    assert _rel(dx.cpu(), ref) < tol
    # wgrad, split-K over the batch, twice (self-cleaning tile counters)
    splits = max(1, min(B, 5))
    ws, cnt = _igemm_scratch(N, dev, cin, cout, splits)
    dW = torch.full((5, 5, cin, cout), float("nan"), device=dev)
    ref = O.conv5x5_wgrad(x.float(), dy.float()).reshape(5, 5, cin, cout)
    for rep in range(2):
        N.call("mtb_conv5x5_gemm", 2, xd.data_ptr(), dyd.data_ptr(), dW.data_ptr(), B, cin, cout, int(tf32), None, 0,
               splits, ws.data_ptr(), cnt.data_ptr())
        torch.cuda.synchronize()
        assert torch.isfinite(dW).all(), rep
        assert _rel(dW.cpu(), ref) < tol, rep
        dW.fill_(float("nan"))


@pytest.mark.parametrize("B", [1, 5, 149, 300])
def test_conv2_dgrad_implicit_gemm(N, dev, B):
    dy = _bf(torch.randn(B, 14, 14, 64)).to(dev)
    w = _bf(torch.randn(5, 5, 32, 64) * 0.05).to(dev)
    dx = torch.zeros(B, 14, 14, 32, dtype=torch.bfloat16, device=dev)
    N.call("mtb_conv2_dgrad", dy.data_ptr(), w.data_ptr(), dx.data_ptr(), B)
    torch.cuda.synchronize()
    ref = O.conv5x5_dgrad(dy.cpu().float(), w.cpu().float(), (B, 14, 14, 32))
    assert _rel(dx.cpu(), ref) < 5e-3


@pytest.mark.parametrize("B", [1, 4, 148, 333])
def test_conv2_wgrad_implicit_gemm(N, dev, B):
    x = _bf(torch.relu(torch.randn(B, 14, 14, 32))).to(dev)
    dy = _bf(torch.randn(B, 14, 14, 64) * (torch.rand(B, 14, 14, 64) < 0.25)).to(dev)
    dW = torch.full((800, 64), float("nan"), device=dev)
    db = torch.zeros(64, device=dev)
    ws = torch.zeros(N.lib().mtb_conv2_wgrad_workspace_bytes(B), dtype=torch.uint8, device=dev)
    for rep in range(2):
        db.zero_()
        N.call("mtb_conv2_wgrad", x.data_ptr(), dy.data_ptr(), dW.data_ptr(), db.data_ptr(), ws.data_ptr(), B)
        torch.cuda.synchronize()
        ref = O.conv5x5_wgrad(x.cpu().float(), dy.cpu().float()).reshape(800, 64)
        assert _rel(dW.cpu(), ref) < 1e-4
        assert _rel(db.cpu(), dy.cpu().float().sum((0, 1, 2))) < 1e-4


@pytest.mark.parametrize("B", [1, 6, 150, 297])
def test_conv2_backward_with_fused_unpool(N, dev, B):
    """dgrad / wgrad fed with the pooled gradient + argmax give what unpool -> dgrad / wgrad give."""
    x = _bf(torch.relu(torch.randn(B, 14, 14, 32))).to(dev)
    w = _bf(torch.randn(5, 5, 32, 64) * 0.05).to(dev)
    a = _bf(torch.randn(B, 14, 14, 64))                      # pre-pool activation (after ReLU some are 0)
    a = torch.relu(a)
    p, i = O.maxpool2x2(a.float())
    P = _bf(p).to(dev)
    idx = i.to(torch.uint8).to(dev)
    dflat = _bf(torch.randn(B, 7, 7, 64)).to(dev)
    # reference path: separate un-pool kernel, then the plain kernels
    dA = torch.zeros(B, 14, 14, 64, dtype=torch.bfloat16, device=dev)
    N.call("mtb_unpool_relu_bwd", dflat.data_ptr(), P.data_ptr(), idx.data_ptr(), dA.data_ptr(), B, 14, 14, 64)
    dx_ref = torch.zeros(B, 14, 14, 32, dtype=torch.bfloat16, device=dev)
    N.call("mtb_conv2_dgrad", dA.data_ptr(), w.data_ptr(), dx_ref.data_ptr(), B)
    ws = torch.zeros(N.lib().mtb_conv2_wgrad_workspace_bytes(B), dtype=torch.uint8, device=dev)
    dW_ref, db_ref = torch.zeros(800, 64, device=dev), torch.zeros(64, device=dev)
    N.call("mtb_conv2_wgrad", x.data_ptr(), dA.data_ptr(), dW_ref.data_ptr(), db_ref.data_ptr(), ws.data_ptr(), B)
    # fused path
    dx = torch.full((B, 14, 14, 32), 3.0, dtype=torch.bfloat16, device=dev)
    N.call("mtb_conv2_dgrad_unpool", dflat.data_ptr(), P.data_ptr(), idx.data_ptr(), w.data_ptr(), dx.data_ptr(), B)
    dW, db = torch.full((800, 64), float("nan"), device=dev), torch.zeros(64, device=dev)
    N.call("mtb_conv2_wgrad_unpool", x.data_ptr(), dflat.data_ptr(), P.data_ptr(), idx.data_ptr(), dW.data_ptr(),
           db.data_ptr(), ws.data_ptr(), B)
    torch.cuda.synchronize()
    assert torch.equal(dx, dx_ref)
    assert _rel(dW, dW_ref) < 1e-5 and _rel(db, db_ref) < 1e-5
    ref = O.conv5x5_dgrad(dA.cpu().float(), w.cpu().float(), (B, 14, 14, 32))
    assert _rel(dx.cpu(), ref) < 5e-3


# ----------------------------------------------------------------------------- GEMM -> reduce-scatter epilogue
@pytest.mark.parametrize("staged", [True, False])
@pytest.mark.parametrize("bn,world,K", [(128, 3, 100), (256, 2, 64), (128, 8, 1024)])
def test_gemm_reduce_scatter_epilogue(N, dev, bn, world, K, staged):
    """mtb_gemm_bf16_rs adds X^T dY into whichever "rank's" buffer owns each 16-byte vector.  The kernel does not
    care whether a buffer is peer or local memory, so one GPU with `world` local buffers checks the routing."""
    M, Nn, base = 3136, 512, 1024                      # fc1 wgrad of the tutorial model, ragged last M tile
    X = _bf(torch.randn(K, M)).to(dev)                 # MN-major A: stored [K, M]
    dY = _bf(torch.randn(K, Nn)).to(dev)               # MN-major B: stored [K, N]
    total = base + M * Nn + 512
    per = -(-(total // 4) // world) * 4
    bufs = [torch.ones(total, device=dev) for _ in range(world)]
    rs = N.GemmRS()
    rs.world, rs.rank, rs.base, rs.per, rs.flags = world, 1, base, per, 0 if staged else 1
    for r in range(world):
        rs.out[r] = bufs[r].data_ptr()
    for _ in range(2):                                 # it ADDS: two calls leave twice the product
        N.gemm_bf16_rs(X, dY, M, Nn, K, True, True, Nn, bn, rs)
    torch.cuda.synchronize()
    prod = (X.float().t() @ dY.float()).reshape(-1)
    owner = (torch.arange(base, base + M * Nn, device=dev) // 4) // (per // 4)
    for r in range(world):
        got = bufs[r][base:base + M * Nn] - 1.0
        want = torch.where(owner == r, 2.0 * prod, torch.zeros_like(prod))
        assert _rel(got, want) < 1e-5 + 2e-5 * K / 2048, (r, _rel(got, want))
        assert torch.equal(got[owner != r], torch.zeros_like(got[owner != r]))
        assert float(bufs[r][:base].min()) == 1.0 and float(bufs[r][base + M * Nn:].max()) == 1.0
