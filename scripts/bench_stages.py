#!/usr/bin/env python
"""Per-stage timing of the hot path at the BASELINE.json config shapes (C1, C3, C4, C5), CUDA events on the library
stream, device-resident buffers, weights frozen.  One JSON line per stage.  Development tool -- bench.py is the contract.

    python scripts/bench_stages.py [--stage c1|c3|c4|c5|f1|f2|f4|all] [--mode fp32|tf32|x3] [--batch-scale 0.25] [--iters 3]
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402
import nntoolkitcore_b200 as nn  # noqa: E402
from nntoolkitcore_b200 import layers as L  # noqa: E402

dev = torch.device("cuda:0")


def timeit(fn, iters, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return min(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))


def emit(**kw):
    print(json.dumps(kw), flush=True)


def c1(lib, args):
    gru = L.Recurrent(lib, "gru", 40, 128, 100); helpers.fill_recurrent(gru, 100)
    L.freeze(lib, gru)
    x = torch.randn(100, 40, device=dev)
    out = torch.empty(100, 128, device=dev)
    ms = timeit(lambda: gru.apply(x, out), 20, 5)
    emit(stage="C1 GRU in40 H128 T100 B1 (stateful single-sequence call)", ms=ms, us_per_step=ms * 10, audio_s_per_s=1.0 / (ms * 1e-3))


def c3(lib, args):
    B = max(1, int(2048 * args.batch_scale)); T = 1000
    chans = [(128, 256), (256, 256), (256, 256)]
    x = torch.randn(B, T, 128, device=dev)
    total_ms, total_flop = 0.0, 0.0
    for i, (ci, co) in enumerate(chans):
        conv = L.Conv1d(lib, ci, co, 5, 1, T); helpers.fill_conv(conv, 300 + i)
        t_out = conv.out_shape[0]
        bn = L.BatchNorm(lib, co, 1e-3, t_out); helpers.fill_bn(bn, 310 + i)
        relu = L.Activation(lib, L.ACT_RELU, t_out * co)
        L.freeze(lib, conv, bn)
        y = torch.empty(B, t_out, co, device=dev)
        ms = timeit(lambda: conv.apply_batch(x, y, bn=bn, act=relu), args.iters)
        flop = 2.0 * B * t_out * co * ci * 5
        byts = 4.0 * B * (T * ci + t_out * co)
        emit(stage=f"C3 conv{i + 1} {ci}->{co} k5 T{T} B{B} +BN+ReLU", mode=args.mode, ms=ms, tflops=flop / ms / 1e9, gbs=byts / ms / 1e6)
        total_ms += ms; total_flop += flop
        x, T = y, t_out
    emit(stage=f"C3 stack B{B}", mode=args.mode, ms=total_ms, tflops=total_flop / total_ms / 1e9, audio_s_per_s=B * 10.0 / (total_ms * 1e-3))


def c4(lib, args):
    B = max(1, int(4096 * args.batch_scale)); T = 500; H = 512
    x = torch.randn(B, T, H, device=dev)
    total_ms = 0.0
    for i in range(2):
        lstm = L.Recurrent(lib, "lstm", H, H, T, True, True); helpers.fill_recurrent(lstm, 400 + i)
        L.freeze(lib, lstm)
        y = torch.empty(B, T, H, device=dev)
        ms = timeit(lambda: lstm.apply_batch(x, y), args.iters, 1)
        flop = 2.0 * B * T * (H * 4 * H + H * 4 * H)
        emit(stage=f"C4 lstm{i + 1} in{H} H{H} T{T} B{B}", mode=args.mode, ms=ms, tflops=flop / ms / 1e9, us_per_step=ms * 1e3 / T)
        total_ms += ms
        x = y
    emit(stage=f"C4 2xLSTM B{B}", mode=args.mode, ms=total_ms, tflops=2 * 2.0 * B * T * 8 * H * H / total_ms / 1e9,
         audio_s_per_s=B * 5.0 / (total_ms * 1e-3))


def c5(lib, args):
    B = max(1, int(8192 * args.batch_scale / 8)); n = 160000
    pool = torch.from_numpy(np.stack([helpers.audio_clip(i, n) for i in range(min(B, 64))])).to(dev)
    x = pool[torch.arange(B, device=dev) % pool.shape[0]].contiguous()
    spec = L.Spectrogram(lib, 512, 400, 240, n, 1.0, "hann_window")
    T0, F = spec.out_shape
    bn = L.BatchNorm(lib, F, 1e-3, T0); helpers.fill_bn(bn, 500)
    conv = L.Conv1d(lib, F, 256, 5, 1, T0); helpers.fill_conv(conv, 501)
    T1 = conv.out_shape[0]
    relu = L.Activation(lib, L.ACT_RELU, T1 * 256)
    g1 = L.Recurrent(lib, "gru", 256, 256, T1); helpers.fill_recurrent(g1, 502)
    g2 = L.Recurrent(lib, "gru", 256, 256, T1); helpers.fill_recurrent(g2, 503)
    tdd = L.Dense(lib, 256, 12, L.ACT_SOFTMAX, ts=T1); helpers.fill_dense(tdd, 504)
    L.freeze(lib, bn, conv, g1, g2, tdd)
    s = torch.empty(B, T0, F, device=dev); sb = torch.empty_like(s)
    c = torch.empty(B, T1, 256, device=dev); h1 = torch.empty_like(c); h2 = torch.empty_like(c)
    y = torch.empty(B, T1, 12, device=dev)
    stages = [("spectrogram", lambda: spec.apply_batch(x, s)), ("batchnorm", lambda: bn.apply_batch(s, sb)),
              ("conv+relu", lambda: conv.apply_batch(sb, c, act=relu)), ("gru1", lambda: g1.apply_batch(c, h1)),
              ("gru2", lambda: g2.apply_batch(h1, h2)), ("tdd+softmax", lambda: tdd.apply_batch(h2, y))]
    total = 0.0
    for name, fn in stages:
        ms = timeit(fn, args.iters, 1)
        emit(stage=f"C5 {name} B{B}", mode=args.mode, ms=ms)
        total += ms
    emit(stage=f"C5 chain B{B}", mode=args.mode, ms=total, audio_s_per_s=B * 10.0 / (total * 1e-3))


def f1(lib, args):
    """fused log-mel front-end at the C2 shape (4096 x 10 s clips x batch-scale), against the same
    three stages run separately (spectrogram -> MelFilterBankApply -> the log lives only in the fused / chained C call)"""
    B = max(1, int(4096 * args.batch_scale)); n = 160000
    pool = torch.from_numpy(np.stack([helpers.audio_clip(i, n) for i in range(min(B, 64))])).to(dev)
    x = pool[torch.arange(B, device=dev) % pool.shape[0]].contiguous()
    spec = L.Spectrogram(lib, 512, 512, 352, n, 1.0, "hann_window")
    T0, F = spec.out_shape
    for n_mels in (40, 80, 128):
        lm = L.LogMelSpectrogram(lib, spec, n_mels, 16000, 20.0, 8000.0)
        out = torch.empty(B, T0, n_mels, device=dev)
        ms = timeit(lambda: lm.apply_batch(x, out), max(args.iters, 5), 2)
        alg = 4.0 * B * (n + T0 * n_mels)
        emit(stage=f"F1 log-mel fused n_mels={n_mels} B{B}", ms=ms, gbs=alg / ms / 1e6, audio_s_per_s=B * 10.0 / (ms * 1e-3))
        lm.close()
    s = torch.empty(B, T0, F, device=dev)
    bank = L.MelFilterBank(lib, 80, 512, 16000, 20.0, 8000.0)
    m = torch.empty(B, T0, 80, device=dev)
    ms_s = timeit(lambda: spec.apply_batch(x, s), max(args.iters, 5), 2)
    ms_m = timeit(lambda: bank.apply(s, m), max(args.iters, 5), 2)
    emit(stage=f"F1 unfused spectrogram + mel GEMM n_mels=80 B{B} (log not included)", ms=ms_s + ms_m, spectrogram_ms=ms_s, mel_ms=ms_m,
         audio_s_per_s=B * 10.0 / ((ms_s + ms_m) * 1e-3))


def f4(lib, args):
    """row f4: batched 512-point complex DFT, planar planes, 16 algorithmic bytes per point"""
    B = max(1, int((1 << 20) * args.batch_scale))
    for nfft in (512, 256, 100):
        re = torch.randn(B, nfft, device=dev); im = torch.randn(B, nfft, device=dev)
        o_re = torch.empty_like(re); o_im = torch.empty_like(im)
        d = L.DFT(lib, nfft, True)
        ms = timeit(lambda: d.perform(re, im, o_re, o_im), max(args.iters, 5), 2)
        emit(stage=f"F4 DFT nfft={nfft} forward B{B}", ms=ms, gbs=16.0 * B * nfft / ms / 1e6, transforms_per_s=B / (ms * 1e-3))
        d.close()


def f2(lib, args):
    """row f2: bidirectional helpers at the C5 GRU shape [B, 994, 256], 8 bytes per element moved"""
    B = max(1, int(1024 * args.batch_scale)); T, H = 994, 256
    a = torch.randn(B, T, H, device=dev); b = torch.randn(B, T, H, device=dev)
    o1 = torch.empty_like(a); o2 = torch.empty(B, T, 2 * H, device=dev)
    n = a.numel()
    for name, fn, byts in (("reverse", lambda: L.bd_reverse_input(lib, a, o1), 8.0 * n),
                           ("merge_concat", lambda: L.bd_merge_concat(lib, a, b, o2), 16.0 * n),
                           ("merge_sum", lambda: L.bd_merge_sum(lib, a, b, o1), 12.0 * n)):
        ms = timeit(fn, max(args.iters, 5), 2)
        emit(stage=f"F2 bd_{name} [B{B},{T},{H}]", ms=ms, gbs=byts / ms / 1e6)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--stage", default="all")
    ap.add_argument("--mode", default="fp32", choices=["fp32", "tf32", "x3"])
    ap.add_argument("--batch-scale", type=float, default=1.0)
    ap.add_argument("--iters", type=int, default=3)
    args = ap.parse_args()
    lib = nn.lib()
    assert lib.NNTKInit(0) == 0, lib.NNTKLastError()
    lib.NNTKSetStream(torch.cuda.current_stream().cuda_stream)
    assert lib.NNTKSetMathMode({"fp32": 0, "tf32": 1, "x3": 2}[args.mode]) == 0
    for name, fn in (("c1", c1), ("c3", c3), ("c4", c4), ("c5", c5), ("f1", f1), ("f2", f2), ("f4", f4)):
        if args.stage == name or (args.stage == "all" and name[0] == "c"):
            fn(lib, args)


if __name__ == "__main__":
    main()
