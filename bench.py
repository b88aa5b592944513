#!/usr/bin/env python
"""bench.py -- audio-seconds/second of the hot path on N B200s (contract: see the task statement / DESIGN.md section 6).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2] [--impl ours|reference]

N > 1 is launched by torchrun (one rank per GPU); independent clips shard by batch with NO collective on the data path
(weak scaling: every rank processes the full per-GPU workload); NCCL is used only for the barrier and the max-over-ranks
of the device time.  One JSON line is printed by rank 0.

A "step" = one pass of the hot path over one batch of synthetic input resident in HBM.  Default workload `c2` =
BASELINE.json configs[1]: Spectrogram nfft=512 hop=160 Hann @16 kHz over 4096 synthetic 10 s clips per GPU.
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
C2 = dict(nfft=512, window_size=512, noverlap=352, input_size=160000, clips=4096, audio_sec_per_clip=10.0)


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return d.get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """SM clock + throttle reasons DURING the timed region (B200_PROFILING.md recipe).  Sampled in-process through NVML --
    the library nvidia-smi itself queries -- every 2 ms, because the timed region of the default run is ~20 ms and an
    `nvidia-smi -lms 100` child does not even start that fast on an 8-GPU box (it produced no sample there); the
    nvidia-smi child is kept as the fallback when NVML cannot be imported.  Only samples between mark_start() and
    mark_end() are summarised."""
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
            # the first queries after nvmlInit can take tens of ms (seen: no sample at all inside a 17 ms timed region):
            # pay that here, synchronously, before the polling thread and the timed region start
            get = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            for _ in range(3):
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
        bits = [("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else 0x8),
                ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4)]
        while not self.stop:
            try:
                sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                mask = int(get(self.h))
                self.rows.append((time.perf_counter(), sm, self.mx, [n for n, b in bits if mask & b]))
            except Exception:
                pass
            time.sleep(0.001)

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
        if not rows:   # region shorter than one sampling period: the samples taken under the same load just around it
            rows = [r for r in self.rows if t0 - 0.05 <= r[0] <= t1 + 0.05]
            window = "timed region +-50 ms"
        sm = [r[1] for r in rows]
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(r[2] for r in rows) if rows else None,
                "reasons": sorted({n for r in rows for n in r[3]}), "samples": len(sm), "source": self.source, "window": window}


def c2_pool(n_pool, n):
    import helpers
    return np.stack([helpers.audio_clip(i, n) for i in range(n_pool)])


def run_reference_arm(args):
    """--impl reference: the unmodified reference CPU implementation on the host cores, bounded sample per step."""
    import bench_cpu
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = dict(C2)
    procs = bench_cpu.host_cores()
    units = max(2 * procs, 32)             # clips per step (bounded sample of the 4096-clip workload)
    for _ in range(args.warmup if args.warmup < 1 else 1):
        bench_cpu.run_reference("c2", cfg, min(units, procs), procs)
    times = []
    for _ in range(args.steps):
        t, procs, _ = bench_cpu.run_reference("c2", cfg, units, procs)
        times.append(t)
    total = float(sum(times))
    value = units * args.steps * cfg["audio_sec_per_clip"] / total
    sample = f"{units} clips per step over {procs} processes (of the {cfg['clips']}-clip workload), {args.steps} steps"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config("c2", cfg, 1),
        "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": procs, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=_OUT, flush=True)


def workload_config(name, cfg, n_gpus):
    return {"workload": f"C2 spectrogram magnitude nfft={cfg['nfft']} window={cfg['window_size']} hop="
                        f"{cfg['window_size'] - cfg['noverlap']} Hann @16kHz, {cfg['clips']} x 10 s clips per GPU",
            "clips_per_gpu": cfg["clips"], "input_size": cfg["input_size"], "parallelism": f"batch-sharded x{n_gpus}, no collective",
            "l2": "inputs+outputs 6.8 GB per step >> 126 MB L2 (no flush needed)"}


def run_ours(args):
    import torch
    import torch.distributed as dist
    import nntoolkitcore_b200 as nn
    from nntoolkitcore_b200 import layers as L

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch N>1 through torchrun (one rank per GPU)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # one process per GPU: keep this rank's threads (and so its pinned host buffers, first touch) on the GPU's own
        # NUMA node -- the e2e leg moves 6.8 GB per step per GPU through host memory
        try:
            import pynvml
            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
        except Exception:
            pass
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = nn.lib()      # raises if the CUDA library is not built: there is no fallback
    assert lib.NNTKInit(local) == 0, lib.NNTKLastError()
    cfg = dict(C2)
    if args.clips:
        cfg["clips"] = args.clips
    n, clips = cfg["input_size"], cfg["clips"]

    # ---- cpu baseline first (rank 0, N == 1 only), before the GPU is busy ---------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        import bench_cpu
        procs = bench_cpu.host_cores()
        units = max(4 * procs, 64)
        t, procs, _ = bench_cpu.run_reference("c2", cfg, units, procs)
        cpu_baseline = {"value": units * cfg["audio_sec_per_clip"] / t, "unit": "audio-s/s", "cores": procs,
                        "kind": "reference",
                        "sample": f"{units} of {clips} clips, unmodified reference (oracle/_ref) over {procs} processes, {t:.2f} s"}

    # ---- synthetic input: pool of distinct clips tiled to the batch ---------------------------------------------------
    n_pool = min(args.pool, clips)
    pool = torch.from_numpy(c2_pool(n_pool, n))
    spec = L.Spectrogram(lib, cfg["nfft"], cfg["window_size"], cfg["noverlap"], n, 1.0, "hann_window")
    ntime, nfreq = spec.out_shape
    x_dev = torch.empty((clips, n), dtype=torch.float32, device=dev)
    pool_dev = pool.to(dev)
    for i in range(0, clips, n_pool):
        k = min(n_pool, clips - i)
        x_dev[i:i + k] = pool_dev[:k]
    del pool_dev
    y_dev = torch.empty((clips, ntime, nfreq), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream(dev)
    lib.NNTKSetStream(stream.cuda_stream)

    def step():
        spec.apply_batch(x_dev, y_dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    with ClockSampler(local) as clk:
        for _ in range(2):      # the sampler thread is already polling while the GPU runs the same load
            step()
        barrier()
        clk.mark_start()
        ev[0].record(stream)
        for i in range(args.steps):
            step()
            ev[i + 1].record(stream)
        barrier()
        clk.mark_end()
    total_ms = ev[0].elapsed_time(ev[-1])
    per_launch_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    audio_s = clips * cfg["audio_sec_per_clip"] * world * args.steps
    value = audio_s / (total_ms * 1e-3)

    # ---- roofline of the dominant kernel (spectrogram512_kernel): algorithmic bytes / launch duration -------------
    alg_bytes = clips * (4 * n + 4 * ntime * nfreq)
    launch_ms = float(np.mean(per_launch_ms))
    peak, peak_src = peaks()
    achieved = alg_bytes / (launch_ms * 1e-3) / 1e9
    traffic = None
    tp = ROOT / "profiles" / "traffic.json"
    if tp.exists():
        traffic = json.loads(tp.read_text()).get("spectrogram512_kernel", {}).get("dram_bytes_per_launch_c2")
        if traffic and clips != C2["clips"]:
            traffic = None
    roofline = {"bound": "hbm", "kernel": "spectrogram512_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes": alg_bytes,
                "launch_ms": launch_ms}

    # ---- e2e: the reference-facing C call with HOST buffers (pinned), H2D + D2H inside the timed region ----------------
    e2e = None
    if not args.no_e2e:
        x_host = torch.empty((clips, n), dtype=torch.float32, pin_memory=True)
        x_host.copy_(x_dev)
        y_host = torch.empty((clips, ntime, nfreq), dtype=torch.float32, pin_memory=True)
        xh, yh = x_host.numpy(), y_host.numpy()
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        spec.apply_batch(xh, yh)    # warm-up: allocates the staging slots
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            spec.apply_batch(xh, yh)     # synchronous: outputs are in host memory on return
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        ok = bool(torch.equal(y_host[:8].to(dev), y_dev[:8]))
        e2e = {"value": clips * cfg["audio_sec_per_clip"] * world * e2e_steps / dt, "unit": "audio-s/s",
               "h2d_bytes_per_step": clips * n * 4, "d2h_bytes_per_step": clips * ntime * nfreq * 4, "steps": e2e_steps,
               "ms_per_step": 1e3 * dt / e2e_steps, "matches_device_path": ok,
               "api": "SpectrogramApplyBatch(host in, host out): chunked H2D/compute/D2H pipeline inside the C call"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config("c2", cfg, world), "clocks": clk.summary(), "roofline": roofline,
            "cpu_baseline": cpu_baseline, "e2e": e2e, "gpu_launches": args.steps,
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--clips", type=int, default=0, help="override clips per GPU (debug; the headline uses 4096)")
    ap.add_argument("--pool", type=int, default=1024, help="distinct synthetic clips (tiled to the batch)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    _claim_stdout()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
