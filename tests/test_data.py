"""Data tier: IDX parse/normalise, split arithmetic, shards, synthetic."""
import numpy as np
import pytest

from mpi_tensorflow_b200.utils import data as D


def test_split_sizes_table():
    # SURVEY §2.1 per-N table (mpipy.py:211-213, :79)
    exp = {1: (55000, 1718), 2: (27500, 859), 4: (13750, 429), 8: (6875, 214)}
    for n, (rows, steps) in exp.items():
        s = D.split_sizes(n)
        assert s.train_rows_per_rank == rows
        assert s.steps(2, 64) == steps
    s3 = D.split_sizes(3)
    assert (s3.tr_size, s3.ts_size, s3.val_size) == (54999, 9999, 4998)


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
def test_shards_partition_exactly(n):
    sizes = D.split_sizes(n, n_train=550, n_test=100, n_val=50)
    x = np.arange(sizes.val_size + sizes.tr_size, dtype=np.float32).reshape(-1, 1, 1, 1)
    y = np.arange(sizes.val_size + sizes.tr_size, dtype=np.int64)
    (vx, vy), (tx, ty) = D.split_train_val(x, y, sizes)
    assert tx.shape[0] == sizes.tr_size and vx.shape[0] == sizes.val_size
    got = np.concatenate([D.shard(ty, r, n) for r in range(n)])
    assert np.array_equal(got, ty)                      # exact partition, no zero tail
    assert np.array_equal(ty, np.arange(sizes.val_size, sizes.val_size + sizes.tr_size))


def test_reference_bug_mode_zero_tail():
    # mpipy.py:221 vs :230 — each rank: (tr-val)/N real rows then val/N zero rows
    n = 2
    sizes = D.split_sizes(n, n_train=550, n_test=100, n_val=50)
    x = np.ones((600, 2, 2, 1), dtype=np.float32)
    y = np.arange(1, 601, dtype=np.int64)
    _, (tx, ty) = D.split_train_val(x, y, sizes, reference_bug=True)
    per, real = sizes.tr_size // n, (sizes.tr_size - sizes.val_size) // n
    for r in range(n):
        sh = D.shard(ty, r, n)
        assert (sh[:real] != 0).all() and (sh[real:] == 0).all()
        assert D.shard(tx, r, n)[real:].sum() == 0
    assert per - real == sizes.val_size // n


def test_idx_roundtrip_python_and_native(tmp_path):
    u8, lab = D.synthetic_mnist(37, seed=3)
    ip, lp = str(tmp_path / "img.gz"), str(tmp_path / "lbl.gz")
    D.write_idx(u8, lab, ip, lp)
    x = D.extract_data(ip, 30)
    y = D.extract_labels(lp, 30)
    assert x.shape == (30, 28, 28, 1) and x.dtype == np.float32
    assert np.allclose(x.reshape(30, 28, 28), (u8[:30].astype(np.float32) - 127.5) / 255.0)
    assert x.min() >= -0.5 and x.max() <= 0.5
    assert y.dtype == np.int64 and np.array_equal(y, lab[:30])
    from mpi_tensorflow_b200.ops import native_host as nh
    assert np.array_equal(nh.read_idx_images(ip, 30), x)
    assert np.array_equal(nh.read_idx_labels(lp, 30), y)
    with pytest.raises(ValueError):
        D.extract_data(lp, 1)
    with pytest.raises(ValueError):
        nh.read_idx_images(lp, 1)


def test_synthetic_deterministic_and_offset():
    a, la = D.synthetic_mnist(16, seed=5)
    b, lb = D.synthetic_mnist(16, seed=5)
    assert np.array_equal(a, b) and np.array_equal(la, lb)
    c, lc = D.synthetic_mnist(8, seed=5, offset=8)
    assert np.array_equal(c, a[8:]) and np.array_equal(lc, la[8:])
    assert len(np.unique(la)) > 3


def test_sequential_sampler_matches_reference_window():
    s = D.SequentialSampler(1000, 64)
    for step in (0, 1, 14, 15, 100):
        assert s.offset(step) == (step * 64) % (1000 - 64)       # mpipy.py:80
    with pytest.raises(ValueError):
        D.SequentialSampler(64, 64)


def test_missing_file_raises_real_exception(tmp_path):
    with pytest.raises(D.DownloadError):
        D.maybe_download("nope.gz", str(tmp_path), allow_network=False)
