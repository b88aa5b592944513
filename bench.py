#!/usr/bin/env python
"""bench.py -- audio-seconds/second of the hot path on N B200s (contract: the task statement / DESIGN.md sections 5-6).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c2|c3|c4] [--impl ours|reference]
                    [--mode tf32|x3|fp32] [--no-extra] [--no-e2e] [--no-cpu]

Headline workload (default) = BASELINE.json's metric "audio-sec/sec (spectrogram+conv1d+GRU/LSTM fwd)": the C5 chain
    SpectrogramApplyBatch -> BatchNormApplyInferenceBatch -> Conv1d(+ReLU) -> GRU -> GRU -> TimeDistributedDense(+softmax)
over 8 192 synthetic 10-s clips PER GPU (N = 8: 65 536 clips = BASELINE configs[4]), as ONE chained pass per step with every
intermediate resident in HBM.  Reference chain: signal/spectrogram.c:113, layers/batch_norm.c:166, conv_1d.c:149,
gru.c:189, time_distributed_dense.c:52.  The other BASELINE configs ride in the same JSON line as `extra`:
    c2  spectrogram, 4 096 clips per GPU (weak);  c3  Conv1d stack, B = 2 048 per GPU (weak);
    c4  2 x LSTM-512, T = 500, B = 4 096 / N per rank (STRONG scaling, as BASELINE configs[3] asks).
N > 1 is launched by torchrun (one rank per GPU); units shard by batch with NO collective on the data path; NCCL carries
only the barrier and the max-over-ranks of the device time.  Rank 0 prints ONE JSON line.

A "step" = one pass of the workload over one batch of synthetic input already resident in HBM (`value`); `e2e` is the same
chain fed from HOST memory (pinned, and pageable beside it) through the library's public C calls, H2D and D2H inside the
timed region.  `--impl reference` times the unmodified reference (oracle/_ref) on the host cores on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

_OUT = sys.stdout

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

METRIC = "audio-sec/sec"
MODES = {"fp32": 0, "tf32": 1, "x3": 2}
DTYPE = {"fp32": "f32 (FFMA, 1e-5 contract)",
         "tf32": "tf32 conv/dense (f16 operands in the fused front end) + f16-operand recurrent steps on tcgen05, f32 accumulate and state (stated tolerance 1e-3)",
         "x3": "3xTF32 (conv/dense) and 3x split-fp16 (fused GRU) operands on tcgen05, f32 accumulate and state, f32-equivalent (1e-5 contract)"}
CLIP_N, CLIP_S = 160000, 10.0

WORKLOADS = {
    "c5": dict(units=8192, audio_s=10.0, strong=False, l2="per step: 5.2 GB of audio in, 8.3 GB per intermediate >> 126 MB L2 (no flush needed)",
               desc="C5 KWS chain: spectrogram(nfft 512, win 400, hop 160, Hann) -> BN(257) -> Conv1d(257->256,k5)+ReLU -> "
                    "2 x GRU(256) -> TDD(256->12)+softmax, 10-s clips @16 kHz"),
    "c2": dict(units=4096, audio_s=10.0, strong=False, l2="inputs+outputs 6.8 GB per step >> 126 MB L2 (no flush needed)",
               desc="C2 spectrogram magnitude nfft=512 window=512 hop=160 Hann @16kHz, 10-s clips"),
    "c3": dict(units=2048, audio_s=10.0, strong=False, l2="activations >= 1 GB per layer >> 126 MB L2 (no flush needed)",
               desc="C3 Conv1d stack 3 x (k5 s1 + BN + ReLU), 128->256->256->256, T=1000"),
    "c4": dict(units=4096, audio_s=5.0, strong=True, l2="x, h1, h2: 4.2 GB each at B = 4096 >> 126 MB L2 (no flush needed)",
               desc="C4 2 x LSTM(512->512, v2, sequences), T=500, batch 4096 sharded over the GPUs"),
}


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_tflops": d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0)),
                "source": "measured (MEASURED_PEAKS.json; bf16 = sustained)"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1400.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region (B200_PROFILING.md recipe), polled in-process through NVML (the
    library nvidia-smi itself queries) every 10 ms; nvidia-smi child as the fallback.  Only samples between
    mark_start() and mark_end() are summarised."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc, self.h, self.stop = index, [], None, None, False
        self.t0 = self.t1 = None
        self.source = None

    def __enter__(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[self.index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else self.index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nv = pynvml
            self.mx = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            get = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            for _ in range(3):      # the first NVML queries take tens of ms: pay that before the timed region
                pynvml.nvmlDeviceGetClockInfo(self.h, pynvml.NVML_CLOCK_SM)
                get(self.h)
            self.source = "nvml"
            threading.Thread(target=self._poll, daemon=True).start()
        except Exception:
            self.h = None
            try:
                self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                              "--format=csv,noheader,nounits", "-lms", "20"], stdout=subprocess.PIPE, text=True)
                self.source = "nvidia-smi"
                threading.Thread(target=self._read, daemon=True).start()
            except OSError:
                self.proc = None
        return self

    def _poll(self):
        nv = self.nv
        bits = [("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]
        get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self.stop:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                mask = int(get(self.h))
                self.rows.append((time.perf_counter(), sm, self.mx, [n for n, b in bits if mask & b]))
            except Exception:
                pass
            time.sleep(0.01)      # ~100 samples per second: NVML queries take driver locks the launching thread also needs

    def _read(self):
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.proc.stdout:
            c = [v.strip() for v in line.split(",")]
            if len(c) >= 6 and c[0].isdigit() and c[1].isdigit():
                self.rows.append((time.perf_counter(), int(c[0]), int(c[1]),
                                  [n for n, v in zip(names, c[2:6]) if v.lower().startswith("active")]))

    def mark_start(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def __exit__(self, *a):
        self.stop = True
        if self.proc:
            self.proc.terminate()

    def summary(self):
        t0, t1 = self.t0 or 0.0, self.t1 or float("inf")
        rows = [r for r in self.rows if t0 <= r[0] <= t1]
        window = "timed region"
        if not rows:
            rows = [r for r in self.rows if t0 - 0.05 <= r[0] <= t1 + 0.05]
            window = "timed region +-50 ms"
        sm = [r[1] for r in rows]
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(r[2] for r in rows) if rows else None,
                "reasons": sorted({n for r in rows for n in r[3]}), "samples": len(sm), "source": self.source, "window": window}


# ---- synthetic inputs (SURVEY.md section 8d generators) ------------------------------------------------------------------
_POOL = {}


def audio_pool(n_pool):
    import helpers
    if n_pool not in _POOL:
        have = max(_POOL) if _POOL else 0
        base = _POOL[have] if have else np.zeros((0, CLIP_N), np.float32)
        if have >= n_pool:
            return base[:n_pool]
        _POOL[n_pool] = np.concatenate([base, np.stack([helpers.audio_clip(i, CLIP_N) for i in range(have, n_pool)])])
    return _POOL[n_pool]


def tile_pool(torch, pool_dev, units, dev):
    x = torch.empty((units,) + tuple(pool_dev.shape[1:]), dtype=torch.float32, device=dev)
    n_pool = pool_dev.shape[0]
    for i in range(0, units, n_pool):
        k = min(n_pool, units - i)
        x[i:i + k] = pool_dev[:k]
    return x


class Env:
    """what every workload builder needs"""

    def __init__(self, torch, lib, L, dev, stream, rank, world, args):
        self.torch, self.lib, self.L, self.dev, self.stream = torch, lib, L, dev, stream
        self.rank, self.world, self.args = rank, world, args
        self.pk = peaks()
        self.tf32_peak = None

    def launches(self):
        return int(self.lib.NNTKDebugLaunchCount()) if hasattr(self.lib, "NNTKDebugLaunchCount") else None

    def measure_tf32_peak(self):
        """dense TF32 GEMM rate of this GPU, same method as the driver's bf16 figure (library matmul, best of 8)"""
        if self.tf32_peak is not None:
            return self.tf32_peak
        torch = self.torch
        old = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = True
        try:
            n = 8192
            a = torch.randn(n, n, device=self.dev); b = torch.randn(n, n, device=self.dev)
            for _ in range(2):
                a @ b
            best = 1e9
            for _ in range(8):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); a @ b; e1.record(); e1.synchronize()
                best = min(best, e0.elapsed_time(e1))
            self.tf32_peak = 2.0 * n ** 3 / (best * 1e-3) / 1e12
        finally:
            torch.backends.cuda.matmul.allow_tf32 = old
        return self.tf32_peak


# ---- workload builders: each returns {"step": fn, "stages": [(name, fn, work)], ...} -------------------------------------
def build_c5(env, units, with_input=True):
    import helpers
    torch, lib, L, dev = env.torch, env.lib, env.L, env.dev
    x = None
    if with_input:
        pool = torch.from_numpy(audio_pool(min(env.args.pool, units))).to(dev)
        x = tile_pool(torch, pool, units, dev)
        del pool
    spec = L.Spectrogram(lib, 512, 400, 240, CLIP_N, 1.0, "hann_window")
    T0, F = spec.out_shape
    bn = L.BatchNorm(lib, F, 1e-3, T0); helpers.fill_bn(bn, 500)
    conv = L.Conv1d(lib, F, 256, 5, 1, T0); helpers.fill_conv(conv, 501)
    T1 = conv.out_shape[0]
    relu = L.Activation(lib, L.ACT_RELU, T1 * 256)
    g1 = L.Recurrent(lib, "gru", 256, 256, T1); helpers.fill_recurrent(g1, 502)
    g2 = L.Recurrent(lib, "gru", 256, 256, T1); helpers.fill_recurrent(g2, 503)
    tdd = L.Dense(lib, 256, 12, L.ACT_SOFTMAX, ts=T1); helpers.fill_dense(tdd, 504)
    L.freeze(lib, bn, conv, g1, g2, tdd)
    fused_front = hasattr(lib, "SpectrogramBNConv1dApplyBatch") and not env.args.unfused_front
    s = None if fused_front else torch.empty(units, T0, F, device=dev)
    c = torch.empty(units, T1, 256, device=dev)
    h1 = torch.empty_like(c)
    h2 = c                                   # the conv output is dead once GRU1 has run: reuse it
    y = torch.empty(units, T1, 12, device=dev)

    def front():
        if fused_front:
            L.front_end(lib, spec, bn, conv, None, relu, x, c)
        else:
            spec.apply_batch(x, s)
            bn.apply_batch(s, s)
            conv.apply_batch(s, c, act=relu)

    def step():
        front()
        g1.apply_batch(c, h1)
        g2.apply_batch(h1, h2)
        tdd.apply_batch(h2, y)

    spec_bytes = 4.0 * units * (CLIP_N + T0 * F)
    conv_flop = 2.0 * units * T1 * 256 * F * 5
    gru_flop = 2.0 * units * T1 * (256 * 768 + 256 * 768)
    gru_bytes = 4.0 * units * T1 * (256 + 256)
    tdd_bytes = 4.0 * units * T1 * (256 + 12)
    tmode = env.args.mode != "fp32"
    if fused_front:
        stages = [("spectrogram+BN+conv+ReLU (fused front end)", front,
                   dict(bound="tensor" if tmode else "ffma", flop=conv_flop, kind="f16" if env.args.mode == "tf32" else "tf32",
                        bytes=4.0 * units * (CLIP_N + T1 * 256), kernels=2,
                        note="two kernels: K1 writing fp16 rows at the conv's pitch (instruction-bound) + the single-pass fp16 implicit GEMM "
                             "(2-CTA clusters share the weight stream); profiles/r2_launches_c5.csv: 4.0 ms + 5.3 ms at 8 192 clips; "
                             "two empty gated launches of the fp32 fallback ride behind them"))]
    else:
        stages = [("spectrogram", lambda: spec.apply_batch(x, s), dict(bound="hbm", bytes=spec_bytes)),
                  ("batchnorm", lambda: bn.apply_batch(s, s), dict(bound="hbm", bytes=8.0 * units * T0 * F)),
                  ("conv1d+relu", lambda: conv.apply_batch(s, c, act=relu), dict(bound="tensor" if tmode else "ffma", flop=conv_flop, kind="tf32"))]
    stages += [("gru1", lambda: g1.apply_batch(c, h1), dict(bound="tensor" if tmode else "ffma", flop=gru_flop, bytes=gru_bytes, kind="f16")),
               ("gru2", lambda: g2.apply_batch(h1, h2), dict(bound="tensor" if tmode else "ffma", flop=gru_flop, bytes=gru_bytes, kind="f16")),
               ("tdd+softmax", lambda: tdd.apply_batch(h2, y), dict(bound="hbm", bytes=tdd_bytes))]
    return dict(step=step, stages=stages, x=x, y=y, units=units, in_bytes=units * CLIP_N * 4, out_bytes=units * T1 * 12 * 4,
                layers=dict(spec=spec, bn=bn, conv=conv, relu=relu, g1=g1, g2=g2, tdd=tdd), T0=T0, T1=T1, F=F,
                fused_front=fused_front)


def build_c2(env, units):
    torch, lib, L, dev = env.torch, env.lib, env.L, env.dev
    pool = torch.from_numpy(audio_pool(min(env.args.pool, units))).to(dev)
    x = tile_pool(torch, pool, units, dev)
    del pool
    spec = L.Spectrogram(lib, 512, 512, 352, CLIP_N, 1.0, "hann_window")
    T0, F = spec.out_shape
    y = torch.empty(units, T0, F, device=dev)
    step = lambda: spec.apply_batch(x, y)
    return dict(step=step, stages=[("spectrogram", step, dict(bound="hbm", bytes=4.0 * units * (CLIP_N + T0 * F)))],
                x=x, y=y, units=units, in_bytes=units * CLIP_N * 4, out_bytes=units * T0 * F * 4, layers=dict(spec=spec))


def build_c3(env, units):
    import helpers
    torch, lib, L, dev = env.torch, env.lib, env.L, env.dev
    T = 1000
    g = torch.Generator(device="cpu"); g.manual_seed(3)
    pool = torch.randn(min(units, 256), T, 128, generator=g).to(dev)          # N(0,1) features, seed 3 (section 8d)
    x = tile_pool(torch, pool, units, dev)
    del pool
    stack, bufs, flop, t = [], [], 0.0, T
    for i, (ci, co) in enumerate([(128, 256), (256, 256), (256, 256)]):
        conv = L.Conv1d(lib, ci, co, 5, 1, t); helpers.fill_conv(conv, 300 + i)
        t_out = conv.out_shape[0]
        bn = L.BatchNorm(lib, co, 1e-3, t_out); helpers.fill_bn(bn, 310 + i)
        relu = L.Activation(lib, L.ACT_RELU, t_out * co)
        L.freeze(lib, conv, bn)
        stack.append((conv, bn, relu))
        bufs.append(torch.empty(units, t_out, co, device=dev))
        flop += 2.0 * units * t_out * co * ci * 5
        t = t_out

    def step():
        v = x
        for (conv, bn, relu), o in zip(stack, bufs):
            conv.apply_batch(v, o, bn=bn, act=relu)
            v = o
    tmode = env.args.mode != "fp32"
    return dict(step=step, stages=[("conv stack", step, dict(bound="tensor" if tmode else "ffma", flop=flop, kind="tf32"))],
                x=x, y=bufs[-1], units=units, layers=dict(stack=stack), in_bytes=units * T * 128 * 4, out_bytes=bufs[-1].numel() * 4)


def build_c4(env, units):
    import helpers
    torch, lib, L, dev = env.torch, env.lib, env.L, env.dev
    T, H = 500, 512
    g = torch.Generator(device="cpu"); g.manual_seed(4)
    pool = torch.randn(min(units, 128), T, H, generator=g).to(dev)
    x = tile_pool(torch, pool, units, dev)
    del pool
    l1 = L.Recurrent(lib, "lstm", H, H, T, True, True); helpers.fill_recurrent(l1, 400)
    l2 = L.Recurrent(lib, "lstm", H, H, T, True, True); helpers.fill_recurrent(l2, 401)
    L.freeze(lib, l1, l2)
    y1 = torch.empty(units, T, H, device=dev); y2 = torch.empty_like(y1)

    def step():
        l1.apply_batch(x, y1)
        l2.apply_batch(y1, y2)
    flop = 2 * 2.0 * units * T * (H * 4 * H + H * 4 * H)
    tmode = env.args.mode != "fp32"
    return dict(step=step, stages=[("2 x lstm", step, dict(bound="tensor" if tmode else "ffma", flop=flop, kind="f16", bytes=3 * 4.0 * units * T * H))],
                x=x, y=y2, units=units, layers=dict(l1=l1, l2=l2), in_bytes=units * T * H * 4, out_bytes=units * T * H * 4)


BUILDERS = {"c5": build_c5, "c2": build_c2, "c3": build_c3, "c4": build_c4}


def close_layers(obj):
    if isinstance(obj, dict):
        for v in obj.values():
            close_layers(v)
    elif isinstance(obj, (list, tuple)):
        for v in obj:
            close_layers(v)
    elif hasattr(obj, "close"):
        obj.close()


# ---- timing ----------------------------------------------------------------------------------------------------------------
def barrier(env):
    if env.world > 1:
        import torch.distributed as dist
        dist.barrier()
    env.torch.cuda.synchronize(env.dev)


def max_over_ranks(env, v):
    if env.world == 1:
        return float(v)
    import torch.distributed as dist
    t = env.torch.tensor([v], dtype=env.torch.float64, device=env.dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def time_steps(env, step, steps, warmup, clk=None):
    """W untimed steps, then exactly K steps between barrier + synchronize, CUDA events on the library stream"""
    torch = env.torch
    for _ in range(warmup):
        step()
    barrier(env)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
    n0 = env.launches()
    if clk:
        clk.mark_start()
    ev[0].record(env.stream)
    for i in range(steps):
        step()
        ev[i + 1].record(env.stream)
    barrier(env)
    if clk:
        clk.mark_end()
    n1 = env.launches()
    total_ms = max_over_ranks(env, ev[0].elapsed_time(ev[-1]))
    per = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
    return total_ms, per, (n1 - n0) if n0 is not None else None


def time_stages(env, stages, iters=3):
    """per-stage device time (median of `iters`), the stages run back to back in chain order exactly like a step"""
    torch = env.torch
    times = [[] for _ in stages]
    for _ in range(iters):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(stages) + 1)]
        ev[0].record(env.stream)
        for i, (_, fn, _) in enumerate(stages):
            fn()
            ev[i + 1].record(env.stream)
        torch.cuda.synchronize(env.dev)
        for i in range(len(stages)):
            times[i].append(ev[i].elapsed_time(ev[i + 1]))
    return [float(np.median(t)) for t in times]


def stage_report(env, stages, ms):
    out = []
    for (name, _, w), t in zip(stages, ms):
        r = {"stage": name, "ms": t, "bound": w["bound"]}
        if w["bound"] == "hbm":
            r.update(algorithmic_bytes=w["bytes"], achieved=w["bytes"] / (t * 1e-3) / 1e9, unit="GB/s", peak=env.pk["hbm_gbs"])
        elif w["bound"] == "tensor":
            peak = env.pk["bf16_tflops"] if w.get("kind") == "f16" else env.measure_tf32_peak()
            r.update(algorithmic_flop=w["flop"], achieved=w["flop"] / (t * 1e-3) / 1e12, unit="TFLOP/s", peak=peak,
                     peak_kind="bf16 sustained (MEASURED_PEAKS.json)" if w.get("kind") == "f16" else "tf32 dense, library matmul measured in this run")
            if "bytes" in w:
                r["algorithmic_bytes"] = w["bytes"]
        else:
            r.update(algorithmic_flop=w["flop"], achieved=w["flop"] / (t * 1e-3) / 1e12, unit="TFLOP/s", peak=None)
        r["frac"] = (r["achieved"] / r["peak"]) if r.get("peak") else None
        if "kernels" in w:
            r["kernels"] = w["kernels"]; r["note"] = w["note"]
        out.append(r)
    return out


def traffic_for(workload, stage_name, units):
    """(dram bytes per launch, kernel name) of the stage's dominant kernel from the committed ncu capture, when one exists for
    this workload, stage and size"""
    tp = ROOT / "profiles" / "traffic.json"
    if not tp.exists():
        return None, None
    d = json.loads(tp.read_text())
    for v in d.values():
        if isinstance(v, dict) and v.get("workload") == workload and v.get("stage") == stage_name:
            return (v.get("dram_bytes_per_launch") if v.get("units") == units else None), v.get("kernel")
    return None, None


# ---- end to end: host buffers -> chain -> host buffers ----------------------------------------------------------------------
def e2e_c5(env, wl, steps, pinned=True):
    """The same chain fed from HOST memory through the library's public C calls: NNTKCopyToDevice (H2D of a chunk of clips) ->
    the six Apply calls on device pointers -> NNTKCopyToHost (the [994, 12] posteriors), chunks double-buffered over two
    library streams so the copy of chunk i+1 overlaps the compute of chunk i.  Every byte of input crosses PCIe inside the
    timed region."""
    torch, lib, dev = env.torch, env.lib, env.dev
    units, T1 = wl["units"], wl["T1"]
    chunk = min(units, env.args.e2e_chunk)
    xh = torch.empty((units, CLIP_N), dtype=torch.float32, pin_memory=pinned)
    xh.copy_(wl["x"])
    yh = torch.empty((units, T1, 12), dtype=torch.float32, pin_memory=pinned)
    sub = build_c5(env, chunk, with_input=False)         # chunk-sized scratch; the weights come from the same seeds
    xd = [torch.empty((chunk, CLIP_N), dtype=torch.float32, device=dev) for _ in range(2)]
    yd = [torch.empty((chunk, T1, 12), dtype=torch.float32, device=dev) for _ in range(2)]
    s_copy, s_run, s_out = torch.cuda.Stream(dev), torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    Lr = sub["layers"]
    fused = sub["fused_front"]
    L = env.L
    # chunk schedule: full chunks and, with --e2e-tail > 0, a smaller final chunk (the compute of the LAST chunk is the part of the step
    # no copy hides).  Measured: the taper does not pay -- the fused recurrent kernel's time barely depends on the batch (a latency
    # chain of T steps), so an extra chunk costs a whole chain
    bounds, a0 = [], 0
    while units - a0 > chunk:
        bounds.append((a0, chunk)); a0 += chunk
    rest = units - a0
    tail = int(rest * env.args.e2e_tail) // 128 * 128
    if 0 < tail < rest and rest >= 1024:
        bounds += [(a0, rest - tail), (a0 + rest - tail, tail)]
    else:
        bounds.append((a0, rest))
    n_chunks = len(bounds)

    def run_chunk(i, slot):
        a, n = bounds[i]
        b = a + n
        with torch.cuda.stream(s_copy):
            xd[slot][:n].copy_(xh[a:b], non_blocking=True)
            ready = torch.cuda.Event(); ready.record(s_copy)
        s_run.wait_event(ready)
        lib.NNTKSetStream(s_run.cuda_stream)
        xin, yout = xd[slot][:n], yd[slot][:n]
        c, h1 = sub["_c"][:n], sub["_h1"][:n]
        if fused:
            L.front_end(lib, Lr["spec"], Lr["bn"], Lr["conv"], None, Lr["relu"], xin, c)
        else:
            s = sub["_s"][:n]
            Lr["spec"].apply_batch(xin, s); Lr["bn"].apply_batch(s, s); Lr["conv"].apply_batch(s, c, act=Lr["relu"])
        Lr["g1"].apply_batch(c, h1); Lr["g2"].apply_batch(h1, c); Lr["tdd"].apply_batch(c, yout)
        done = torch.cuda.Event(); done.record(s_run)
        s_out.wait_event(done)
        with torch.cuda.stream(s_out):
            yh[a:b].copy_(yout, non_blocking=True)
            out_done = torch.cuda.Event(); out_done.record(s_out)
        return done, out_done

    # chunk-sized scratch (the builder allocated per-chunk buffers of its own; reuse them)
    sub["_c"] = torch.empty(chunk, T1, 256, device=dev); sub["_h1"] = torch.empty_like(sub["_c"])
    sub["_s"] = None if fused else torch.empty(chunk, wl["T0"], wl["F"], device=dev)

    def one_pass():
        pend = [None, None]
        for i in range(n_chunks):
            slot = i & 1
            if pend[slot] is not None:            # slot reuse: its compute and its D2H must be done
                s_copy.wait_event(pend[slot][0]); s_run.wait_event(pend[slot][1])
            pend[slot] = run_chunk(i, slot)
        torch.cuda.synchronize(dev)

    saved = env.stream
    one_pass()                                     # warm-up
    barrier(env)
    t0 = time.perf_counter()
    for _ in range(steps):
        one_pass()
    dt = max_over_ranks(env, time.perf_counter() - t0)
    lib.NNTKSetStream(saved.cuda_stream)
    ok = bool(torch.allclose(yh[:4].to(dev), wl["y"][:4], rtol=0, atol=2e-3))
    close_layers(sub["layers"])
    del sub, xd, yd
    return dict(value=units * CLIP_S * env.world * steps / dt, ms_per_step=1e3 * dt / steps, steps=steps, matches_device_path=ok,
                chunk_clips=chunk, chunk_schedule=[n for _, n in bounds], host_memory="pinned" if pinned else "pageable (malloc)"), xh


def host_ceiling(env, xh):
    """plain pinned cudaMemcpyAsync H2D of this rank's input, all ranks at once: what the box's host side allows"""
    torch = env.torch
    d = torch.empty_like(xh, device=env.dev)
    d.copy_(xh, non_blocking=True); torch.cuda.synchronize(env.dev)
    barrier(env)
    t0 = time.perf_counter()
    d.copy_(xh, non_blocking=True)
    torch.cuda.synchronize(env.dev)
    dt = max_over_ranks(env, time.perf_counter() - t0)
    return xh.numel() * 4 / dt / 1e9


def e2e_hostcall(env, wl, steps):
    """c2: SpectrogramApplyBatch(host in, host out) -- the chunked H2D / compute / D2H pipeline inside the C call"""
    torch = env.torch
    xh = torch.empty(tuple(wl["x"].shape), dtype=torch.float32, pin_memory=True); xh.copy_(wl["x"])
    yh = torch.empty(tuple(wl["y"].shape), dtype=torch.float32, pin_memory=True)
    spec = wl["layers"]["spec"]
    spec.apply_batch(xh.numpy(), yh.numpy())
    barrier(env)
    t0 = time.perf_counter()
    for _ in range(steps):
        spec.apply_batch(xh.numpy(), yh.numpy())
    torch.cuda.synchronize(env.dev)
    dt = max_over_ranks(env, time.perf_counter() - t0)
    return dict(value=wl["units"] * CLIP_S * env.world * steps / dt, unit="audio-s/s", ms_per_step=1e3 * dt / steps, steps=steps,
                h2d_bytes_per_step=wl["in_bytes"], d2h_bytes_per_step=wl["out_bytes"],
                matches_device_path=bool(torch.equal(yh[:8].to(env.dev), wl["y"][:8])),
                api="SpectrogramApplyBatch(host in, host out)")


# ---- one workload, measured ------------------------------------------------------------------------------------------------
def run_workload(env, name, steps, warmup, clk=None, with_e2e=False, with_stages=True):
    w = WORKLOADS[name]
    units = env.args.units if (env.args.units and name == env.args.workload) else w["units"]
    if w["strong"]:
        from nntoolkitcore_b200 import sharding
        lo, hi = sharding.unit_range(units, env.rank, env.world)
        local_units = hi - lo
    else:
        local_units = units
    wl = BUILDERS[name](env, local_units)
    total_ms, per, launches = time_steps(env, wl["step"], steps, warmup, clk)
    total_units = units if w["strong"] else units * env.world
    value = total_units * w["audio_s"] * steps / (total_ms * 1e-3)
    res = dict(workload=name, value=value, unit="audio-s/s", ms_per_step=total_ms / steps, ms_per_step_median=float(np.median(per)), steps_ms=[round(float(v), 3) for v in per],
               steps=steps, warmup=warmup, scaling="strong" if w["strong"] else "weak", units_per_gpu=local_units,
               units_total=total_units, gpu_launches=launches, desc=w["desc"])
    if with_stages:
        ms = time_stages(env, wl["stages"])
        res["stages"] = stage_report(env, wl["stages"], ms)
    e2e = None
    if with_e2e:
        es = max(1, min(steps, env.args.e2e_steps))
        if name == "c5":
            e2e, xh = e2e_c5(env, wl, es, pinned=True)
            e2e.update(unit="audio-s/s", h2d_bytes_per_step=wl["in_bytes"], d2h_bytes_per_step=wl["out_bytes"],
                       api="NNTKCopyToDevice-style H2D of clip chunks -> SpectrogramApplyBatch/BatchNorm/Conv1d/GRU/GRU/TDD on device "
                           "pointers -> D2H of the posteriors; double-buffered over the library stream (NNTKSetStream)")
            e2e["h2d_ceiling_gbs_per_gpu_all_ranks_at_once"] = host_ceiling(env, xh)
            del xh
            if not env.args.no_pageable:
                pg, xh2 = e2e_c5(env, wl, 1, pinned=False)
                e2e["pageable"] = {k: pg[k] for k in ("value", "ms_per_step", "host_memory", "matches_device_path")}
                del xh2
        elif name == "c2":
            e2e = e2e_hostcall(env, wl, es)
    close_layers(wl["layers"])
    del wl
    env.torch.cuda.empty_cache()
    return res, e2e


def pick_roofline(env, res):
    """the dominant KERNEL by CUDA-event share of the step: stages that are one kernel launch compete with their full time, a stage
    made of several kernels with its time per kernel (its split is in the committed ncu launch list)"""
    st = max(res["stages"], key=lambda s: s["ms"] / s.get("kernels", 1))
    share = st["ms"] / sum(s["ms"] for s in res["stages"])
    traffic, kernel = traffic_for(res["workload"], st["stage"], res["units_per_gpu"])
    if st["bound"] == "hbm":
        alg = st["algorithmic_bytes"]
    else:
        alg = st["algorithmic_flop"]
    return {"bound": "hbm" if st["bound"] == "hbm" else "tensor", "stage": st["stage"], "kernel": kernel, "achieved": st["achieved"],
            "peak": st.get("peak"), "unit": st["unit"], "frac": st.get("frac"), "traffic": traffic,
            "peak_source": env.pk["source"] + ("; " + st["peak_kind"] if "peak_kind" in st else ""),
            "algorithmic_work": alg, "launch_ms": st["ms"], "share_of_step": share,
            "traffic_source": "profiles/traffic.json (ncu --set full capture, per launch)" if traffic else None}


# ---- the reference arm -----------------------------------------------------------------------------------------------------
def run_reference_arm(args):
    """--impl reference: the unmodified reference CPU implementation on the host cores, a bounded sample per step"""
    import bench_cpu
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    name = args.workload
    w = WORKLOADS[name]
    procs = bench_cpu.host_cores()
    per_core = {"c5": 2, "c2": 4, "c3": 1, "c4": 1}[name]
    units = max(per_core * procs, 16)
    cfg = dict(nfft=512, window_size=512, noverlap=352, input_size=CLIP_N)
    pool = bench_cpu.Pool(name, cfg, procs)
    try:
        for _ in range(1 if args.warmup else 0):
            pool.run(min(units, procs))
        times = [pool.run(units) for _ in range(args.steps)]
    finally:
        pool.close()
    total = float(sum(times))
    value = units * args.steps * w["audio_s"] / total
    sample = (f"{units} units per step over {procs} worker processes (of the {w['units']}-unit per-GPU workload), {args.steps} steps; "
              f"each worker runs the unmodified reference chain on its own handles")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "strong" if w["strong"] else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, args.gpus, w["units"]),
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": procs, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_OUT, flush=True)


def workload_config(name, n_gpus, units):
    w = WORKLOADS[name]
    per = f"{units} units sharded over the GPUs" if w["strong"] else f"{units} units per GPU"
    return {"workload": f"{w['desc']}; {per}", "units": units, "units_are": "10-s clips" if w["audio_s"] == 10.0 else "5-s sequences",
            "parallelism": f"batch-sharded x{n_gpus}, no collective", "l2": w["l2"]}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import nntoolkitcore_b200 as nn
    from nntoolkitcore_b200 import layers as L

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world == 1 and args.gpus > 1:
        raise SystemExit("launch N>1 through torchrun (one rank per GPU)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    affinity = None
    if world > 1:
        # keep this rank's threads (and the first touch of its pinned buffers) on the GPU's own NUMA node
        try:
            import pynvml
            pynvml.nvmlInit()
            before = len(os.sched_getaffinity(0))
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
            affinity = {"cpus_before": before, "cpus_after": len(os.sched_getaffinity(0))}
        except Exception as e:      # recorded in the line: a silent failure here was suspected of the r1 e2e scaling
            affinity = {"error": repr(e)[:120], "cpus": len(os.sched_getaffinity(0))}
        dist.init_process_group("nccl", device_id=dev)
    lib = nn.lib()      # raises if the CUDA library is not built: there is no fallback
    assert lib.NNTKInit(local) == 0, lib.NNTKLastError()
    assert lib.NNTKSetMathMode(MODES[args.mode]) == 0
    stream = torch.cuda.current_stream(dev)
    lib.NNTKSetStream(stream.cuda_stream)
    env = Env(torch, lib, L, dev, stream, rank, world, args)
    name = args.workload

    # ---- cpu baseline first (rank 0, N == 1 only), before the GPU is busy ---------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import bench_cpu
        procs = bench_cpu.host_cores()
        units = max({"c5": 6, "c2": 16, "c3": 2, "c4": 1}[name] * procs, 16)      # c5: ~20 core-seconds of reference work
        pool = bench_cpu.Pool(name, dict(nfft=512, window_size=512, noverlap=352, input_size=CLIP_N), procs)
        try:
            pool.run(procs)
            t = pool.run(units)
        finally:
            pool.close()
        cpu_baseline = {"value": units * WORKLOADS[name]["audio_s"] / t, "unit": "audio-s/s", "cores": procs, "kind": "reference",
                        "sample": f"{units} of {WORKLOADS[name]['units']} units, unmodified reference (oracle/_ref) over {procs} processes, {t:.2f} s"}

    warmup = max(args.warmup, 3)
    with ClockSampler(local) as clk:
        res, e2e = run_workload(env, name, args.steps, warmup, clk, with_e2e=not args.no_e2e)
    roofline = pick_roofline(env, res)

    # the 1e-5-contract mode beside the headline mode (same workload, fewer steps)
    other = None
    if not args.no_extra and args.mode == "tf32":
        assert lib.NNTKSetMathMode(MODES["x3"]) == 0
        o, _ = run_workload(env, name, max(3, args.steps // 4), 3, with_stages=False)
        other = {"mode": "x3", "dtype": DTYPE["x3"], "value": o["value"], "ms_per_step": o["ms_per_step"]}
        assert lib.NNTKSetMathMode(MODES[args.mode]) == 0

    extra = {}
    if not args.no_extra:
        for ex in ("c2", "c3", "c4"):
            if ex == name:
                continue
            r, e = run_workload(env, ex, max(3, min(args.steps, 10)), 3, with_e2e=(ex == "c2" and not args.no_e2e))
            r["config"] = workload_config(ex, world, WORKLOADS[ex]["units"])
            r["roofline"] = pick_roofline(env, r)
            if e:
                r["e2e"] = e
            extra[ex] = r

    if rank == 0:
        w = WORKLOADS[name]
        line = {
            "metric": METRIC, "value": res["value"], "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": res["ms_per_step"], "ms_per_step_median": res["ms_per_step_median"], "steps_ms": res["steps_ms"], "higher_is_better": True,
            "scaling": res["scaling"], "vs_baseline": None, "dtype": DTYPE[args.mode], "math_mode": args.mode, "data": "synthetic",
            "config": workload_config(name, world, args.units or w["units"]),
            "clocks": clk.summary(), "roofline": roofline, "stages": res.get("stages"), "cpu_baseline": cpu_baseline, "e2e": e2e,
            "gpu_launches": res["gpu_launches"], "fp32_contract_mode": other, "extra": extra, "affinity": affinity,
        }
        print(json.dumps(line), file=_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


def _claim_stdout():
    """Keep the JSON line alone on stdout: libraries that write to fd 1 (NCCL's version banner) go to stderr."""
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c5", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", default="tf32", choices=sorted(MODES))
    ap.add_argument("--units", type=int, default=0, help="override units per GPU of the headline workload (debug)")
    ap.add_argument("--pool", type=int, default=1024, help="distinct synthetic clips (tiled to the batch)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--e2e-chunk", type=int, default=2048,
                    help="clips per H2D/compute chunk of the end-to-end pipeline (the fused recurrent kernel is a latency chain of "
                         "T steps whatever the batch up to ~8k sequences: few large chunks beat many small ones; measured at 8 192 clips: "
                         "1 408 -> 110.3 ms, 2 048 -> 109.7, 2 816 -> 112.7, 4 096 -> 116.0 per step)")
    ap.add_argument("--e2e-tail", type=float, default=0.0,
                    help="fraction of the last chunk split off as a final small chunk (0: no taper; measured at 8 192 clips: 0 -> 109.9 ms, "
                         "0.25 -> 112.9, 0.5 -> 134.1 per step: the extra chunk costs more than the shorter tail saves)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-pageable", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--unfused-front", action="store_true", help="run spectrogram, BN and conv as three calls")
    args = ap.parse_args()
    _claim_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
