"""Multi-GPU tiers: one process per GPU over NVLink peer memory."""
import json
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

WORKER = os.path.join(os.path.dirname(__file__), "workers", "gpu_worker.py")


def _run(nproc, mode, outdir, timeout=420):
    cmd = [sys.executable, "-m", "mpi_tensorflow_b200.parallel.launch", "-np", str(nproc), "--timeout",
           str(timeout), "--", WORKER, mode, str(outdir)]
    e = dict(os.environ)
    e["PYTHONPATH"] = os.path.dirname(os.path.dirname(__file__)) + os.pathsep + e.get("PYTHONPATH", "")
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=e, timeout=timeout + 60)
    return p.returncode, p.stdout


def _world():
    return min(torch.cuda.device_count(), 8)


def test_p2p_collectives_algebra(tmp_path):
    n = _world()
    rc, out = _run(n, "algebra", tmp_path)
    assert rc == 0 and out.count("GPU_COMM_OK") == n, out[-4000:]


def test_fused_allreduce_dp_equivalence(tmp_path):
    n = _world()
    rc, out = _run(n, "fused", tmp_path)
    assert rc == 0 and out.count("GPU_FUSED_OK") == n, out[-4000:]


def test_weight_average_mode_on_gpus(tmp_path):
    n = _world()
    rc, out = _run(n, "weight_avg", tmp_path)
    assert rc == 0 and out.count("GPU_WAVG_OK") == n, out[-4000:]


def test_dead_rank_gives_clean_failure_on_gpus(tmp_path):
    """Only 2 ranks: one stalls, the survivor must report a timeout within the spin limit."""
    rc, out = _run(2, "die", tmp_path, timeout=180)
    assert rc != 0
    assert "UNEXPECTED_SUCCESS" not in out, out[-3000:]
    assert "CLEAN_FAILURE" in out, out[-3000:]


def test_trainer_fit_multi_gpu(tmp_path):
    n = _world()
    rc, out = _run(n, "train", tmp_path)
    assert rc == 0 and out.count("GPU_TRAIN_OK") == n, out[-4000:]
    res = json.load(open(tmp_path / "result.json"))
    assert res["steps"] == 30 and res["loss"] == res["loss"]
