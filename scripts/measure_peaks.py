#!/usr/bin/env python
"""Roofline denominators MEASURED_PEAKS.json does not carry: cuBLAS TF32 GEMM throughput (the tensor-pipe ceiling of the
kind::tf32 kernels) and cuBLAS fp32 (FFMA) GEMM throughput (the ceiling of the 1e-5 fp32-mode kernels).  Same method as
the driver's bf16 figure: 8192^3 torch.matmul, best of 10, CUDA events.  One JSON line."""
import json
import torch


def best(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        t.append(a.elapsed_time(b))
    return min(t)


N = 8192
x = torch.randn(N, N, device="cuda"); y = torch.randn(N, N, device="cuda")
out = {}
for name, flag in (("tf32_tflops", True), ("fp32_ffma_tflops", False)):
    torch.backends.cuda.matmul.allow_tf32 = flag
    ms = best(lambda: torch.matmul(x, y))
    out[name] = 2.0 * N ** 3 / ms / 1e9
out["how"] = "torch.matmul fp32 8192^3, allow_tf32 on/off, best of 10, CUDA events"
print(json.dumps(out))
