"""Builds libnntoolkitcore_b200.so in-tree: nvcc (sm_100a only) for csrc/*.cu, gcc for host/*.c, one shared library.

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.  `python -m nntoolkitcore_b200.build`
rebuilds whatever is out of date; `--force` rebuilds everything.  cudart is linked statically and the driver API is
never linked (TMA descriptors are obtained through cudaGetDriverEntryPoint), so the library loads on a box without
a GPU driver -- every compute entry point then fails loudly with NNTK_ERR_CUDA.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
ROOT = PKG.parent
INC = ROOT / "include"
OBJ = PKG / "lib" / "obj"
LIB = PKG / "lib" / "libnntoolkitcore_b200.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CC = os.environ.get("CC", "gcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
              f"-I{INC}", f"-I{PKG / 'csrc'}", *os.environ.get("NNTK_NVCC_EXTRA", "").split()]
CC_FLAGS = ["-O2", "-std=gnu11", "-fPIC", "-Wall", "-Wextra", "-Wno-unused-parameter", f"-I{INC}", f"-I{PKG / 'host'}"]


def _newer(src: Path, dst: Path, deps: list[Path]) -> bool:
    if not dst.exists():
        return True
    t = dst.stat().st_mtime
    return any(p.stat().st_mtime > t for p in [src, *deps])


def _run(cmd: list[str]) -> None:
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
        raise RuntimeError(f"build step failed: {cmd[0]} {cmd[-1]}")


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(parents=True, exist_ok=True)
    cu = sorted((PKG / "csrc").glob("*.cu"))
    cs = sorted((PKG / "host").glob("*.c"))
    headers = sorted(INC.rglob("*.h")) + sorted((PKG / "csrc").glob("*.h")) + sorted((PKG / "csrc").glob("*.cuh")) + \
        sorted((PKG / "host").glob("*.h"))
    jobs, objs = [], []
    for s in cu:
        o = OBJ / (s.stem + ".cu.o")
        objs.append(o)
        if force or _newer(s, o, headers):
            jobs.append([NVCC, *NVCC_FLAGS, "-c", str(s), "-o", str(o)])
    for s in cs:
        o = OBJ / (s.stem + ".c.o")
        objs.append(o)
        if force or _newer(s, o, headers):
            jobs.append([CC, *CC_FLAGS, "-c", str(s), "-o", str(o)])
    if verbose:
        for j in jobs:
            print(" ".join(j))
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        list(ex.map(_run, jobs))
    if jobs or not LIB.exists():
        _run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(LIB), *map(str, objs),
              "-cudart", "static", "-lpthread", "-ldl", "-lrt", "-lm"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
