"""Property tests (hypothesis) for the host-side pieces whose contract is arithmetic: the
data split and batch window of ``mpipy.py:211-241, 80-82``, the collective algebra every sync
mode is built on (``mpipy.py:121-136``), the checkpoint containers, the dropout stream and the
loss mailbox word.  CPU only."""
import struct
import types

import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from mpi_tensorflow_b200.parallel.comm import FakeWorld
from mpi_tensorflow_b200.utils import checkpoint as ckpt
from mpi_tensorflow_b200.utils import data as D
from mpi_tensorflow_b200.utils import philox

FAST = settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture])


@FAST
@given(world=st.integers(1, 16), rows=st.integers(1, 5000))
def test_shards_are_a_partition_of_the_rounded_dataset(world, rows):
    """Every row below ``rows // world * world`` lands on exactly one rank, in order (mpipy.py:211-213, 236-241)."""
    a = np.arange(rows)
    parts = [D.shard(a, r, world) for r in range(world)]
    per = rows // world
    assert all(len(p) == per for p in parts)
    assert np.array_equal(np.concatenate(parts) if per else np.zeros(0, int), a[:per * world])


@FAST
@given(world=st.sampled_from([1, 2, 3, 4, 8]))
def test_split_sizes_follow_the_reference_arithmetic(world):
    s = D.split_sizes(world)
    assert s.train_rows_per_rank == 55000 // world and s.test_rows_per_rank == 10000 // world
    assert s.val_rows_per_rank == 5000 // world
    assert s.steps(2, 64) == 2 * (55000 // world) // 64                       # mpipy.py:79
    assert s.reference_real_rows_per_rank() == (55000 // world * world - 5000 // world * world) // world


@FAST
@given(rows=st.integers(2, 100000), batch=st.integers(1, 4096), step=st.integers(0, 10 ** 7))
def test_sequential_window_matches_reference_formula(rows, batch, step):
    if rows <= batch:
        with pytest.raises(ValueError):
            D.SequentialSampler(rows, batch)
        return
    idx = D.SequentialSampler(rows, batch).indices(step)
    off = (step * batch) % (rows - batch)                                     # mpipy.py:80
    assert idx[0] == off and idx[-1] == off + batch - 1 and idx[-1] < rows and len(idx) == batch


@FAST
@given(n=st.integers(1, 6), count=st.integers(1, 2000), seed=st.integers(0, 2 ** 31 - 1), average=st.booleans())
def test_allreduce_scatter_gather_algebra(n, count, seed, average):
    """allreduce == serial NumPy sum (or mean), bit-identical on every rank; scatter o gather == identity;
    sizes that are not a multiple of any vector width included."""
    full = np.random.default_rng(seed).standard_normal((n, count)).astype(np.float32)

    def body(comm):
        r = comm.rank
        t = torch.from_numpy(full[r].copy())
        comm.allreduce(t, average=average)
        mine = torch.zeros(count)
        comm.scatter(torch.from_numpy(full.copy()) if r == 0 else None, mine, 0)
        back = torch.zeros(n, count) if r == 0 else None
        comm.gather(mine, back, 0)
        every = torch.zeros(n, count)
        comm.allgather(mine, every)
        return t, back, every

    res = FakeWorld(n).run(body)
    ref = full.astype(np.float64).sum(0) / (n if average else 1)
    for t, back, every in res:
        assert np.allclose(t.numpy(), ref, rtol=1e-5, atol=1e-5)
        assert torch.equal(t, res[0][0])
        assert np.array_equal(every.numpy(), full)
    assert np.array_equal(res[0][1].numpy(), full)


_shapes = st.lists(st.integers(1, 7), min_size=0, max_size=4).map(tuple)


@FAST
@given(shapes=st.lists(_shapes, min_size=1, max_size=6), seed=st.integers(0, 2 ** 31 - 1),
       fmt=st.sampled_from(["npz", "bundle"]))
def test_checkpoint_containers_roundtrip_bit_exact(tmp_path_factory, shapes, seed, fmt):
    rng = np.random.default_rng(seed)
    named = {}
    for i, shp in enumerate(shapes):
        kind = i % 3
        if kind == 0:
            named[f"Variable_{i}"] = rng.standard_normal(shp).astype(np.float32)
        elif kind == 1:
            named[f"dense_{i}/kernel/Momentum"] = rng.standard_normal(shp).astype(np.float32)
        else:
            named[f"global_step_{i}"] = rng.integers(-2 ** 40, 2 ** 40, size=shp, dtype=np.int64)
    d = tmp_path_factory.mktemp("ck")
    path = str(d / ("c.npz" if fmt == "npz" else "model.ckpt-7"))
    ckpt.save(path, named)
    got = ckpt.load(path)
    assert set(got) == set(named)
    for k, v in named.items():
        assert got[k].dtype == v.dtype and got[k].shape == v.shape and np.array_equal(got[k], v), k


@FAST
@given(data=st.binary(min_size=0, max_size=300))
def test_crc32c_native_and_python_agree(data):
    """The masked crc32c that guards every bundle block: table-driven Python == the C++ routine."""
    assert ckpt._crc32c(data) == _py_crc(data)
    m = ckpt.masked_crc32c(data)
    assert 0 <= m < 2 ** 32


def _py_crc(data: bytes) -> int:
    crc = 0xFFFFFFFF
    for b in data:
        crc ^= b
        for _ in range(8):
            crc = (crc >> 1) ^ (0x82F63B78 if crc & 1 else 0)
    return crc ^ 0xFFFFFFFF


@FAST
@given(keep=st.floats(0.05, 0.95), seed=st.integers(0, 2 ** 40), step=st.integers(0, 10 ** 6))
def test_dropout_stream_is_deterministic_and_unbiased(keep, seed, step):
    n = 20000
    a = philox.dropout_mask(n, keep, seed, step)
    assert np.array_equal(a, philox.dropout_mask(n, keep, seed, step))
    assert abs(a.mean() - keep) < 5 * np.sqrt(keep * (1 - keep) / n) + 1e-3
    assert not np.array_equal(a, philox.dropout_mask(n, keep, seed, step + 1))      # a new mask every step


@FAST
@given(step=st.integers(1, 2 ** 32 - 1), loss=st.floats(width=32, allow_nan=False))
def test_loss_mailbox_word_roundtrip(step, loss):
    from mpi_tensorflow_b200.engine_native import NativeEngine

    ring = NativeEngine._MAIL_RING
    mail = np.zeros(ring, dtype=np.int64)
    eng = types.SimpleNamespace(_mail_np=mail, _MAIL_RING=ring, cfg=types.SimpleNamespace(step_timeout_s=0.05))
    bits = struct.unpack("<I", struct.pack("<f", loss))[0]
    mail[(step - 1) % ring] = np.array([(step << 32) | bits], dtype=np.uint64).view(np.int64)[0]
    got = NativeEngine._await_loss(eng, step)
    assert struct.pack("<f", got) == struct.pack("<f", loss)
    if step + ring < 2 ** 32:
        with pytest.raises(RuntimeError):
            NativeEngine._await_loss(eng, step + ring)          # same slot, later step: not there yet


# ---------------------------------------------------------------------------------------------------------
# Ownership of optimizer state under bucketed, owner-sliced K1 launches: when a buffer is completed
# across ranks, every element must come from exactly one rank, and inside a sliced bucket from the rank whose
# slice the kernel's arithmetic (per = ceil(vectors / world)) says it is.
@settings(max_examples=60, deadline=None)
@given(world=st.integers(1, 8), n4=st.integers(8, 400), cuts=st.lists(st.integers(0, 400), min_size=0, max_size=4),
       lo4=st.integers(0, 200), span4=st.integers(1, 300))
def test_sliced_bucket_ownership_tiles_every_range(world, n4, cuts, lo4, span4):
    from mpi_tensorflow_b200.engine_native import NativeEngine as E
    n = 4 * n4
    pts = sorted({min(4 * c, n) for c in cuts} | {0, n})
    buckets = [(a, b) for a, b in zip(pts[:-1], pts[1:])][::2]          # every other interval is an owner-sliced bucket
    lo = min(4 * lo4, n - 4)
    hi = min(n, lo + 4 * span4)
    cover = np.zeros(n, dtype=np.int32)
    for r in range(world):
        for a, b in E.contribution_pieces(r, world, buckets, lo, hi):
            assert lo <= a < b <= hi
            cover[a:b] += 1
    assert (cover[lo:hi] == 1).all() and cover[:lo].sum() == 0 and cover[hi:].sum() == 0
    for blo, bhi in buckets:                                            # slices of one bucket partition it, in rank order
        prev = blo
        for r in range(world):
            olo, ohi = E.owned_slice(blo, bhi, r, world)
            assert olo == prev or olo == ohi or (olo >= prev and ohi <= bhi)
            if ohi > olo:
                assert olo == prev
                prev = ohi
        assert prev == bhi
