"""Comm algebra tier: every collective equals serial NumPy, for
several world sizes and message sizes, on the fake world, the native shm backend
and gloo; plus the fault tier (row 9): a dead rank gives a clean error, not a hang."""
import io
import os
import subprocess
import sys
import contextlib

import numpy as np
import pytest
import torch

from mpi_tensorflow_b200.parallel.comm import CommAborted, CommTimeout, FakeWorld, SelfComm
from mpi_tensorflow_b200.parallel import launch

WORKER = os.path.join(os.path.dirname(__file__), "workers", "comm_worker.py")


def _run_launch(nproc, args, timeout=120, env=None):
    cmd = [sys.executable, "-m", "mpi_tensorflow_b200.parallel.launch", "-np", str(nproc), "--timeout",
           str(timeout), "--", WORKER] + args
    e = dict(os.environ)
    e["PYTHONPATH"] = os.path.dirname(os.path.dirname(__file__)) + os.pathsep + e.get("PYTHONPATH", "")
    if env:
        e.update(env)
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=e,
                       timeout=timeout + 30)
    return p.returncode, p.stdout


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
def test_fake_world_algebra(n):
    rng = np.random.default_rng(0)
    full = rng.standard_normal((n, 1001)).astype(np.float32)

    def body(comm):
        r = comm.rank
        t = torch.from_numpy(full[r].copy())
        comm.allreduce(t, average=True)
        recv = torch.zeros(1001)
        comm.scatter(torch.from_numpy(full.copy()) if r == 0 else None, recv, 0)
        out = torch.zeros(n, 1001) if r == 0 else None
        comm.gather(torch.from_numpy(full[r].copy()), out, 0)
        b = torch.from_numpy(full[0].copy()) if r == 0 else torch.zeros(1001)
        comm.Bcast(b.numpy(), 0)            # mpi4py-style alias on numpy buffers
        comm.Barrier()
        return t, recv, out, b

    res = FakeWorld(n).run(body)
    ref = full.sum(0) / n
    for r, (t, recv, out, b) in enumerate(res):
        assert np.allclose(t.numpy(), ref, atol=1e-6)
        assert torch.equal(t, res[0][0])                    # bit-identical across ranks
        assert np.array_equal(recv.numpy(), full[r])
        assert np.array_equal(b.numpy(), full[0])
    assert np.array_equal(res[0][2].numpy(), full)


def test_fake_world_dead_rank_times_out_cleanly():
    def body(comm):
        if comm.rank == 1:
            return "skipped"                  # never joins the collective
        comm.allreduce(torch.ones(4))

    with pytest.raises((CommTimeout, CommAborted)):
        FakeWorld(2, timeout=0.5).run(body)


def test_self_comm_identity():
    c = SelfComm()
    t = torch.arange(5.)
    c.allreduce(t, average=True)
    assert t.tolist() == [0, 1, 2, 3, 4]
    assert c.Get_rank() == 0 and c.Get_size() == 1


@pytest.mark.parametrize("backend,n", [("shm", 2), ("shm", 4), ("gloo", 2)])
def test_multiprocess_algebra(backend, n):
    rc, out = _run_launch(n, [backend])
    assert rc == 0, out
    assert out.count("COMM_OK") == n, out


@pytest.mark.parametrize("backend", ["shm", "gloo"])
def test_dead_rank_clean_failure(backend):
    # kill one rank before the allreduce: peers must time out / abort with a clear error
    rc, out = _run_launch(2, [backend, "die"], timeout=90, env={"MTB_TEST_TIMEOUT": "5"})
    assert rc != 0
    assert "UNEXPECTED_SUCCESS" not in out, out


def test_launcher_sets_env_and_prefixes(tmp_path):
    prog = tmp_path / "p.py"
    prog.write_text("import os; print(os.environ['RANK'], os.environ['WORLD_SIZE'], os.environ['LOCAL_RANK'])")
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        rc = launch.launch(3, [str(prog)])
    assert rc == 0
    lines = sorted(l for l in buf.getvalue().splitlines() if l.strip())
    assert lines == ["[0] 0 3 0", "[1] 1 3 1", "[2] 2 3 2"]


def test_launcher_fail_fast(tmp_path):
    prog = tmp_path / "p.py"
    prog.write_text("import os, sys, time\nif os.environ['RANK']=='1': sys.exit(5)\ntime.sleep(60)\n")
    import time
    t0 = time.time()
    with contextlib.redirect_stdout(io.StringIO()):
        rc = launch.launch(2, [str(prog)])
    assert rc == 5 and time.time() - t0 < 20
