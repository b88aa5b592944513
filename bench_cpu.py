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


def _f1_units(first, count, cfg):
    """row f1: LogMelSpectrogram (80 bands, 20 Hz .. 8 kHz) on the C2 framing, one 10-s clip per unit"""
    from nntoolkitcore_b200 import layers as L
    ref = _ref()
    s = L.Spectrogram(ref, cfg["nfft"], cfg["window_size"], cfg["noverlap"], cfg["input_size"], 1.0, "hann_window")
    f = L.LogMelSpectrogram(ref, s, 80, 16000, 20.0, 8000.0)
    clips = [_audio(first + i, cfg["input_size"]) for i in range(count)]
    out = np.empty(f.out_shape, np.float32)
    t0 = time.perf_counter()
    for c in clips:
        f.apply(c, out)
    dt = time.perf_counter() - t0
    f.close(); s.close()
    return dt


def _timed(fn, count):
    t0 = time.perf_counter()
    for i in range(count):
        fn(i)
    return time.perf_counter() - t0


def _c1_units(first, count, cfg):
    """C1: single GRU layer in=40 H=128 T=100, one sequence per unit, FRESH handle per sequence (zero state)"""
    import helpers
    from nntoolkitcore_b200 import layers as L
    ref = _ref()
    rng = np.random.default_rng(1 + first)
    xs = rng.standard_normal((count, 100, 40)).astype(np.float32)
    grus = []
    for _ in range(count):
        g = L.Recurrent(ref, "gru", 40, 128, 100); helpers.fill_recurrent(g, 100); grus.append(g)
    return _timed(lambda i: grus[i].apply(xs[i]), count)


def _c3_units(first, count, cfg):
    """C3: 3 x (Conv1d k5 s1 -> BatchNorm -> ReLU), 128 -> 256 -> 256 -> 256, T=1000, one sample per unit"""
    import helpers
    from nntoolkitcore_b200 import layers as L
    ref = _ref()
    rng = np.random.default_rng(3 + first)
    xs = rng.standard_normal((count, 1000, 128)).astype(np.float32)
    T, stack = 1000, []
    for i, (ci, co) in enumerate([(128, 256), (256, 256), (256, 256)]):
        conv = L.Conv1d(ref, ci, co, 5, 1, T); helpers.fill_conv(conv, 300 + i)
        T = conv.out_shape[0]
        bn = L.BatchNorm(ref, co, 1e-3, T); helpers.fill_bn(bn, 310 + i)
        stack.append((conv, bn, L.Activation(ref, L.ACT_RELU, T * co)))

    def one(i):
        v = xs[i]
        for conv, bn, relu in stack:
            v = relu.apply(bn.apply(conv.apply(v)))
    return _timed(one, count)


def _c4_units(first, count, cfg):
    """C4: 2 x LSTM(512 -> 512, v2, sequences), T=500, one sequence per unit, fresh handles per sequence"""
    import helpers
    from nntoolkitcore_b200 import layers as L
    ref = _ref()
    rng = np.random.default_rng(4 + first)
    xs = rng.standard_normal((count, 500, 512)).astype(np.float32)
    pairs = []
    for _ in range(count):
        a = L.Recurrent(ref, "lstm", 512, 512, 500, True, True); helpers.fill_recurrent(a, 400)
        b = L.Recurrent(ref, "lstm", 512, 512, 500, True, True); helpers.fill_recurrent(b, 401)
        pairs.append((a, b))
    return _timed(lambda i: pairs[i][1].apply(pairs[i][0].apply(xs[i])), count)


def _c5_units(first, count, cfg):
    """C5: spectrogram(512, win 400, hop 160) -> BN(257) -> Conv1d(257->256,k5)+ReLU -> 2 x GRU(256) -> TDD(256->12)+softmax"""
    import helpers
    from nntoolkitcore_b200 import layers as L
    ref = _ref()
    n = 160000
    clips = [_audio(first + i, n) for i in range(count)]
    spec = L.Spectrogram(ref, 512, 400, 240, n, 1.0, "hann_window")
    T0, F = spec.out_shape
    bn = L.BatchNorm(ref, F, 1e-3, T0); helpers.fill_bn(bn, 500)
    conv = L.Conv1d(ref, F, 256, 5, 1, T0); helpers.fill_conv(conv, 501)
    T1 = conv.out_shape[0]
    relu = L.Activation(ref, L.ACT_RELU, T1 * 256)
    tdd = L.Dense(ref, 256, 12, L.ACT_SOFTMAX, ts=T1); helpers.fill_dense(tdd, 504)
    grus = []
    for _ in range(count):
        g1 = L.Recurrent(ref, "gru", 256, 256, T1); helpers.fill_recurrent(g1, 502)
        g2 = L.Recurrent(ref, "gru", 256, 256, T1); helpers.fill_recurrent(g2, 503)
        grus.append((g1, g2))
    return _timed(lambda i: tdd.apply(grus[i][1].apply(grus[i][0].apply(relu.apply(conv.apply(bn.apply(spec.apply(clips[i]))))))), count)


UNIT_FNS = {"c1": _c1_units, "c2": _c2_units, "c3": _c3_units, "c4": _c4_units, "c5": _c5_units, "f1": _f1_units}
AUDIO_SEC_PER_UNIT = {"c1": 1.0, "c2": 10.0, "c3": 10.0, "c4": 5.0, "c5": 10.0, "f1": 10.0}


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


def _pool_worker(conn, workload, cfg):
    """persistent worker: imports and library load happen once; every request builds fresh (zero-state) handles for its
    units outside the timed loop, exactly like run_reference's one-shot workers"""
    _ref()
    while True:
        msg = conn.recv()
        if msg is None:
            break
        first, count = msg
        try:
            conn.send(UNIT_FNS[workload](first, count, cfg) if count > 0 else 0.0)
        except Exception as e:      # surfaced by Pool.run
            conn.send(e)


class Pool:
    """`procs` long-lived worker processes (one per host core) for the --impl reference arm and the cpu_baseline leg, so that a
    K-step run does not pay process start-up K times.  run(units) -> seconds of the slowest worker's timed Apply loop."""

    def __init__(self, workload: str, cfg: dict, procs: int | None = None):
        import multiprocessing as mp
        ctx = mp.get_context("spawn")
        self.procs = procs or host_cores()
        self.workers = []
        for _ in range(self.procs):
            a, b = ctx.Pipe()
            pr = ctx.Process(target=_pool_worker, args=(b, workload, cfg), daemon=True)
            pr.start()
            self.workers.append((pr, a))
        self.next_first = 0

    def run(self, units: int) -> float:
        n = min(self.procs, max(1, units))
        per = [units // n + (1 if i < units % n else 0) for i in range(n)]
        for (_, conn), c in zip(self.workers, per):
            conn.send((self.next_first, c))
            self.next_first += c
        times = []
        for (_, conn), _c in zip(self.workers, per):
            t = conn.recv()
            if isinstance(t, Exception):
                raise t
            times.append(t)
        return max(times)

    def close(self):
        for pr, conn in self.workers:
            try:
                conn.send(None)
            except Exception:
                pass
        for pr, _ in self.workers:
            pr.join(timeout=10)
            if pr.is_alive():
                pr.terminate()


if __name__ == "__main__":
    # python bench_cpu.py c1 c3 c4 c5 [--units-per-core K]: one JSON line per config (reference CPU path, all host cores)
    import json
    names = [a for a in sys.argv[1:] if a in UNIT_FNS] or list(UNIT_FNS)
    per_core = {"c1": 64, "c2": 8, "c3": 1, "c4": 1, "c5": 1, "f1": 8}
    for nm in names:
        procs = host_cores()
        units = per_core[nm] * procs
        t, procs, wall = run_reference(nm, dict(nfft=512, window_size=512, noverlap=352, input_size=160000), units, procs)
        print(json.dumps({"config": nm, "impl": "reference-cpu", "cores": procs, "units": units, "slowest_worker_s": t,
                          "audio_s_per_s": units * AUDIO_SEC_PER_UNIT[nm] / t}), flush=True)
