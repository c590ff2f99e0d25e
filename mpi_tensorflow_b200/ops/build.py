"""In-tree build of the two native libraries.

  lib/libmtb200_host.so   g++   csrc/host/*.cpp       CPU runtime (shm process group, IDX loader)
  lib/libmtb200.so        nvcc  csrc/*.cu             sm_100a kernels + device runtime, C ABI

The reference has no build system at all (zero native files); its
native code lives in TensorFlow / MPI.  Here everything native is compiled from
this tree, for exactly one target: ``-gencode arch=compute_100a,code=sm_100a``.
The .so files are git-ignored but travel to the GPU box with the snapshot.

Run ``python -m mpi_tensorflow_b200.ops.build`` (add ``--force`` to rebuild).
"""
from __future__ import annotations

import hashlib
import re
import shlex
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path
from typing import List

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
INCLUDE = ROOT / "include"
LIBDIR = Path(__file__).resolve().parents[1] / "lib"
OBJDIR = ROOT / "build" / "obj"

NVCC = os.environ.get("NVCC", shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--use_fast_math",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
    "--expt-relaxed-constexpr",
    *(["-DMTB_TRACE"] if os.environ.get("MTB_TRACE") else []),
    f"-I{INCLUDE}", f"-I{CSRC}",
]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread", f"-I{INCLUDE}"]
CXX_FLAGS += shlex.split(os.environ.get("CXXFLAGS", ""))       # e.g. -fsanitize=address,undefined (scripts/sanitize.sh)


def _digest(paths: List[Path], extra: str) -> str:
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        h.update(str(p.relative_to(ROOT)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def _stamp_ok(stamp: Path, digest: str, target: Path) -> bool:
    return target.exists() and stamp.exists() and stamp.read_text().strip() == digest


def _run(cmd: List[str], log: Path | None = None) -> str:
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if log is not None:
        log.write_text(" ".join(cmd) + "\n" + p.stdout)
    if p.returncode != 0:
        sys.stderr.write(p.stdout)
        raise RuntimeError(f"build failed ({p.returncode}): {' '.join(cmd[:4])} ...")
    return p.stdout


def build_host(force: bool = False, verbose: bool = False) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted((CSRC / "host").glob("*.cpp"))
    hdrs = sorted(INCLUDE.glob("*.h"))
    target = LIBDIR / "libmtb200_host.so"
    stamp = LIBDIR / ".host.stamp"
    d = _digest(srcs + hdrs, " ".join(CXX_FLAGS))
    if not force and _stamp_ok(stamp, d, target):
        return target
    cmd = [CXX, *CXX_FLAGS, "-shared", "-o", str(target), *map(str, srcs), "-lz", "-lrt"]
    out = _run(cmd)
    if verbose:
        print(out)
    stamp.write_text(d)
    return target


def build_cuda(force: bool = False, verbose: bool = False, jobs: int = 8) -> Path:
    LIBDIR.mkdir(parents=True, exist_ok=True)
    OBJDIR.mkdir(parents=True, exist_ok=True)
    srcs = sorted(CSRC.glob("*.cu"))
    hdrs = sorted(CSRC.glob("*.cuh")) + sorted(INCLUDE.glob("*.h"))
    target = LIBDIR / "libmtb200.so"
    if not srcs:
        raise RuntimeError("no CUDA sources under csrc/")
    hdr_digest = _digest(hdrs, " ".join(NVCC_FLAGS))

    def compile_one(src: Path) -> Path:
        obj = OBJDIR / (src.stem + ".o")
        stamp = OBJDIR / (src.stem + ".stamp")
        # a .cu that includes another .cu (gemm_rs_sm100.cu) depends on it too
        deps = [CSRC / m for m in re.findall(r'#include "(\w+\.cu)"', src.read_text())]
        d = _digest([src] + deps, hdr_digest)
        if not force and _stamp_ok(stamp, d, obj):
            return obj
        out = _run([NVCC, *NVCC_FLAGS, "-c", str(src), "-o", str(obj)], log=OBJDIR / (src.stem + ".log"))
        if verbose:
            print(out)
        stamp.write_text(d)
        return obj

    with ThreadPoolExecutor(max_workers=jobs) as ex:
        objs = list(ex.map(compile_one, srcs))
    link_stamp = LIBDIR / ".cuda.stamp"
    d = _digest(srcs + hdrs, " ".join(NVCC_FLAGS))
    if force or not _stamp_ok(link_stamp, d, target):
        _run([NVCC, "-shared", "-o", str(target), *map(str, objs), "-gencode",
              "arch=compute_100a,code=sm_100a", "-lcudart"])
        link_stamp.write_text(d)
    return target


def build_all(force: bool = False, verbose: bool = False):
    return build_host(force, verbose), build_cuda(force, verbose)


if __name__ == "__main__":
    force = "--force" in sys.argv
    verbose = "-v" in sys.argv or "--verbose" in sys.argv
    only_host = "--host" in sys.argv
    if only_host:
        print(build_host(force, verbose))
    else:
        for p in build_all(force, verbose):
            print(p)
