"""Per-rank body for multi-process communicator tests (run under parallel.launch)."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from mpi_tensorflow_b200.parallel import world  # noqa: E402


def main():
    backend = sys.argv[1]
    mode = sys.argv[2] if len(sys.argv) > 2 else "algebra"
    comm = world.init(backend, "cpu", timeout_s=float(os.environ.get("MTB_TEST_TIMEOUT", "30")))
    r, n = comm.rank, comm.size
    if mode == "die":
        # rank 1 dies before the collective; the others must fail with a clear error, not hang
        if r == 1:
            os._exit(17)
        t = torch.ones(1000)
        try:
            comm.allreduce(t)
        except Exception as e:  # noqa: BLE001
            print(f"CLEAN_FAILURE {type(e).__name__}", flush=True)
            os._exit(3)
        print("UNEXPECTED_SUCCESS", flush=True)
        return
    rng = np.random.default_rng(1234)
    for count in (1, 7, 1000, 100003, (9 << 20) // 4 + 5):   # last one exceeds one shm slot at 8 MiB
        full = rng.standard_normal((n, count)).astype(np.float32)
        # allreduce sum / avg equal serial numpy and are bit-identical on all ranks
        t = torch.from_numpy(full[r].copy())
        comm.allreduce(t, average=True)
        ref = full[0].copy()
        for p in range(1, n):
            ref += full[p]
        ref *= np.float32(1.0 / n)
        assert np.allclose(t.numpy(), ref, rtol=1e-6, atol=1e-6), ("allreduce", count)
        chk = t.clone()
        comm.bcast(chk, root=0)
        assert torch.equal(chk, t), "allreduce result differs across ranks"
        # scatter
        recv = torch.zeros(count)
        comm.scatter(torch.from_numpy(full.copy()) if r == 0 else None, recv, root=0)
        assert np.array_equal(recv.numpy(), full[r]), ("scatter", count)
        # gather
        out = torch.zeros(n, count) if r == 0 else None
        comm.gather(torch.from_numpy(full[r].copy()), out, root=0)
        if r == 0:
            assert np.array_equal(out.numpy(), full), ("gather", count)
        # allgather
        ag = torch.zeros(n * count)
        comm.allgather(torch.from_numpy(full[r].copy()), ag)
        assert np.array_equal(ag.numpy().reshape(n, count), full), ("allgather", count)
        # bcast from a non-zero root
        b = torch.from_numpy(full[n - 1].copy()) if r == n - 1 else torch.zeros(count)
        comm.bcast(b, root=n - 1)
        assert np.array_equal(b.numpy(), full[n - 1]), ("bcast", count)
    # int64 labels travel through scatter too (mpipy.py:237)
    lab = torch.arange(n * 5, dtype=torch.int64) if r == 0 else None
    got = torch.zeros(5, dtype=torch.int64)
    comm.scatter(lab, got, root=0)
    assert got.tolist() == list(range(r * 5, r * 5 + 5))
    objs = comm.allgather_object({"rank": r})
    assert [o["rank"] for o in objs] == list(range(n))
    assert comm.bcast_object("hello" if r == 0 else None) == "hello"
    comm.barrier()
    print(f"COMM_OK rank {r}/{n}", flush=True)
    world.finalize()


if __name__ == "__main__":
    main()
