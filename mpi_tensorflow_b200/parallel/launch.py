"""SPMD launcher: ``python -m mpi_tensorflow_b200.parallel.launch -np N prog.py [args]``.

Replaces the ``mpirun -np N python mpipy.py`` the reference implies but does not
ship (``README.md:4``; entry guard ``mpipy.py:246-247``).  One OS process per rank,
``RANK`` / ``WORLD_SIZE`` / ``LOCAL_RANK`` / ``MASTER_ADDR`` / ``MASTER_PORT`` /
``MTB_JOB_ID`` in the environment, stdout/stderr forwarded with a ``[rank]``
prefix.  Fail-fast: when any rank exits non-zero the others are terminated (the
reference would leave them blocked in ``Gather``, ``mpipy.py:121-127``).
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import threading
import time
import uuid
from typing import List


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _pump(stream, prefix: str, sink):
    for line in iter(stream.readline, b""):
        sink.write(prefix + line.decode(errors="replace"))
        sink.flush()
    stream.close()


def launch(nproc: int, argv: List[str], module: bool = False, env_extra=None, prefix: bool = True,
           timeout: float | None = None, grace: float = 5.0) -> int:
    port = free_port()
    job = uuid.uuid4().hex[:12]
    procs: List[subprocess.Popen] = []
    threads = []
    for r in range(nproc):
        env = dict(os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(nproc),
                    "LOCAL_WORLD_SIZE": str(nproc), "MASTER_ADDR": "127.0.0.1",
                    "MASTER_PORT": str(port), "MTB_JOB_ID": job})
        env.setdefault("OMP_NUM_THREADS", str(max(1, (os.cpu_count() or 1) // nproc)))
        if env_extra:
            env.update(env_extra)
        cmd = [sys.executable, "-u"] + (["-m"] if module else []) + list(argv)
        p = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                             start_new_session=True)
        procs.append(p)
        t = threading.Thread(target=_pump, args=(p.stdout, f"[{r}] " if prefix else "", sys.stdout),
                             daemon=True)
        t.start()
        threads.append(t)

    def kill_all(sig=signal.SIGTERM):
        for p in procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, sig)       # exact process groups we started
                except ProcessLookupError:
                    pass

    rc = 0
    t0 = time.time()
    try:
        while True:
            codes = [p.poll() for p in procs]
            bad = [c for c in codes if c not in (None, 0)]
            if bad:
                rc = bad[0]
                time.sleep(0.2)
                kill_all()
                break
            if all(c == 0 for c in codes):
                break
            if timeout is not None and time.time() - t0 > timeout:
                rc = 124
                kill_all()
                break
            time.sleep(0.05)
    except KeyboardInterrupt:
        rc = 130
        kill_all(signal.SIGINT)
    t1 = time.time()
    for p in procs:
        try:
            p.wait(max(0.1, grace - (time.time() - t1)))
        except subprocess.TimeoutExpired:
            try:
                os.killpg(p.pid, signal.SIGKILL)
            except ProcessLookupError:
                pass
            p.wait()
    for t in threads:
        t.join(1.0)
    return rc


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="mpi_tensorflow_b200.parallel.launch",
                                 description="one process per rank, MPI-style")
    ap.add_argument("-np", "--nproc", type=int, default=1)
    ap.add_argument("-m", dest="module", action="store_true", help="run argv[0] as a module")
    ap.add_argument("--no-prefix", action="store_true")
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("prog", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    if not a.prog:
        ap.error("nothing to run")
    prog = a.prog[1:] if a.prog[0] == "--" else a.prog
    return launch(a.nproc, prog, module=a.module, prefix=not a.no_prefix, timeout=a.timeout)


if __name__ == "__main__":
    sys.exit(main())
