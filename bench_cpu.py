"""CPU side of the benchmark: times the UNMODIFIED reference (oracle/_ref/libnntk_ref.so, the checker built by
oracle/Makefile) on the host cores.  Only bench.py's `cpu_baseline` leg and `--impl reference` arm use this file; nothing
here is on the product path.  Workers import numpy + ctypes only (no torch, no CUDA), one process per core, each with its
own handles (reference handles carry scratch and state and are not shareable, layers/gru.c:39-45)."""
from __future__ import annotations

import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _ref():
    import helpers
    return helpers.reference_lib()


def _audio(i, n):
    import helpers
    return helpers.audio_clip(i % 1024, n)


# ---- per-workload unit functions: process `count` units starting at `first`, return seconds of pure Apply time ------
def _c2_units(first, count, cfg):
    from nntoolkitcore_b200 import layers as L
    ref = _ref()
    s = L.Spectrogram(ref, cfg["nfft"], cfg["window_size"], cfg["noverlap"], cfg["input_size"], 1.0, "hann_window")
    clips = [_audio(first + i, cfg["input_size"]) for i in range(count)]
    out = np.empty(s.out_shape, np.float32)
    t0 = time.perf_counter()
    for c in clips:
        s.apply(c, out)
    dt = time.perf_counter() - t0
    s.close()
    return dt


UNIT_FNS = {"c2": _c2_units}


def _worker(args):
    workload, first, count, cfg = args
    return UNIT_FNS[workload](first, count, cfg)


def run_reference(workload: str, cfg: dict, units: int, procs: int | None = None):
    """Runs `units` independent units of `workload` over `procs` worker processes.  Returns (wall seconds of the slowest
    worker's Apply loop, procs).  Input generation and handle creation are outside the timed loops."""
    from concurrent.futures import ProcessPoolExecutor
    import multiprocessing as mp
    procs = procs or host_cores()
    procs = max(1, min(procs, units))
    per = [units // procs + (1 if i < units % procs else 0) for i in range(procs)]
    jobs, first = [], 0
    for p in per:
        jobs.append((workload, first, p, cfg))
        first += p
    t0 = time.perf_counter()
    with ProcessPoolExecutor(max_workers=procs, mp_context=mp.get_context("spawn")) as ex:
        times = list(ex.map(_worker, jobs))
    wall_all = time.perf_counter() - t0
    return max(times), procs, wall_all
