#!/usr/bin/env python
"""Small-shape drive of every kernel family with cross-CTA communication or asynchronous staging, for compute-sanitizer
(memcheck / racecheck / synccheck -- one tool per gpurun call, B200_PROFILING.md):

    compute-sanitizer --tool racecheck python scripts/sanitize_cases.py

Covers: the fp32 DSMEM cluster kernel, the TF32 and fp16 tcgen05 cluster kernels (incl. the one-slice remainder launch), the fused
persistent pair kernel (FUSED 1- and 3-pass, XW mode, clusters of 2 / 4, split and 32-row shapes, the speculative pair of launches and
the per-step decision), the per-step tensor-core kernel, the tcgen05 GEMM (one-CTA, multicast clusters, cta_group::2, gated launches),
K1 (dense, pitched, fp16 rows) and the host-buffer staging pipeline with >= 3 chunks.  Results are checked against the oracle port so a
"clean" run also means a correct one."""
import os, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers, port_cases, nntoolkitcore_b200 as nn
from nntoolkitcore_b200 import layers as L

lib = nn.lib(); assert lib.NNTKInit(0) == 0
port = helpers.oracle_port()
dev = torch.device("cuda:0")
rng = np.random.default_rng(0)


def rec(cell, in_size, H, T, B, mode, tol, env=None, seq=True):
    for k, v in (env or {}).items():
        os.environ[k] = v
    assert lib.NNTKSetMathMode(mode) == 0
    x = rng.standard_normal((B, T, in_size)).astype(np.float32)
    r = L.Recurrent(lib, cell, in_size, H, T, return_sequences=seq); helpers.fill_recurrent(r, 77)
    got = r.apply_batch(torch.from_numpy(x).to(dev)); torch.cuda.synchronize()
    path = lib.NNTKDebugLastRecurrentPath().decode()
    pick = sorted({0, B // 2, B - 1})
    w = port_cases.recurrent_weights(cell, in_size, H, 77)
    want = np.stack([port_cases.recurrent(port, cell, in_size, H, T, seq, True, w, x[i]) for i in pick])
    err, l2 = helpers.assert_parity(got[pick].cpu().numpy(), want, tol=tol, what=f"{cell} {in_size}->{H} T{T} B{B} mode {mode}")
    print(f"ok {cell} in={in_size} H={H} T={T} B={B} mode={mode} path={path} err={err:.2e}", flush=True)
    r.close(); lib.NNTKSetMathMode(2)


rec("gru", 24, 64, 6, 5, 0, 1e-5)                       # fp32 DSMEM cluster kernel (2 CTAs)
rec("lstm", 24, 128, 5, 9, 0, 1e-5)                     # fp32 cluster kernel, 4 CTAs
rec("gru", 32, 96, 5, 40, 1, 1e-3)                      # H % 64 != 0: fp16 cluster kernel (3 CTAs)
rec("lstm", 48, 160, 4, 300, 1, 1e-3)                   # fp16 cluster kernel with a remainder launch
rec("gru", 64, 64, 5, 130, 1, 1e-3)                     # fused pair kernel, FUSED 3-pass, cluster of 4
rec("lstm", 64, 128, 4, 70, 1, 1e-3, seq=False)         # fused pair kernel, LSTM, one pair (cluster of 2)
rec("gru", 40, 128, 4, 140, 1, 1e-3)                    # fused pair kernel, XW mode (in % 64 != 0)
rec("lstm", 64, 264, 3, 290, 1, 1e-3)                   # per-step tensor-core kernel (H % 32 != 0)
rec("gru", 32, 64, 4, 300, 2, 1e-5)                     # 3xTF32 per-step kernel
rec("gru", 64, 128, 5, 700, 1, 1e-3)                    # fused pair kernel: split mode + 32-row mode candidates, speculative pair of launches
rec("gru", 64, 128, 5, 700, 1, 1e-3, env={"NNTK_RECURRENT_X_SPECULATE": "0"})       # ... the per-step kernel alone
os.environ["NNTK_RECURRENT_X_SPECULATE"] = "1"
rec("lstm", 64, 128, 4, 600, 1, 1e-3, env={"NNTK_RECURRENT_FUSED_SPLIT": "2"})      # split mode forced: DSMEM exchange of h_t halves
os.environ.pop("NNTK_RECURRENT_FUSED_SPLIT")
rec("lstm", 72, 256, 4, 1500, 1, 1e-3)                  # XW mode behind the adaptive projection (cta_group::2 GEMM, gated fp16 / 3xbf16)
rec("gru", 64, 128, 6, 140, 2, 1e-5)                    # fused pair kernel in the fp32-contract mode (three split-fp16 passes for both parts)

# conv on tcgen05 + the fused front end (pitched K1) + the staging pipeline with host buffers (>= 3 chunks of ~32 MiB)
assert lib.NNTKSetMathMode(1) == 0
n, B = 16000, 3
clips = np.stack([helpers.audio_clip(i, n) for i in range(B)])
spec = L.Spectrogram(lib, 512, 400, 240, n, 1.0, "hann_window"); T0, F = spec.out_shape
bn = L.BatchNorm(lib, F, 1e-3, T0); helpers.fill_bn(bn, 1)
conv = L.Conv1d(lib, F, 64, 5, 1, T0); helpers.fill_conv(conv, 2)
relu = L.Activation(lib, L.ACT_RELU, conv.out_shape[0] * 64)
f = L.front_end(lib, spec, bn, conv, None, relu, torch.from_numpy(clips).to(dev)); torch.cuda.synchronize()
want = port_cases.spectrogram(port, 512, 400, 240, n, "hann_window", None, 1.0, clips)
want = port_cases.batch_norm(port, F, 1e-3, bn.gamma, bn.beta, bn.mean, bn.var, want).reshape(B, T0, F)
want = np.maximum(port_cases.conv(port, F, 64, 5, 1, T0, conv.W, conv.b, want), 0)
helpers.assert_parity(f.cpu().numpy(), want, tol=1e-3, what="fused front end")
print("ok fused front end", flush=True)
lib.NNTKSetMathMode(2)
big = np.stack([helpers.audio_clip(i % 4, 160000) for i in range(40)])          # 40 x (0.64 + 1.02) MB: several staging chunks
s2 = L.Spectrogram(lib, 512, 512, 352, 160000, 1.0, "hann_window")
out = s2.apply_batch(big)
ref1 = port_cases.spectrogram(port, 512, 512, 352, 160000, "hann_window", None, 1.0, big[:4])
helpers.assert_parity(out[:4], ref1, what="staged spectrogram"); assert np.array_equal(out[36:40], out[:4])
print("ok staged host-buffer pipeline", flush=True)
print("all sanitize cases ok")
