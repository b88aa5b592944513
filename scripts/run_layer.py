#!/usr/bin/env python
"""Development driver: one recurrent layer (or conv) at a given shape on cuda:0, timed with CUDA events -- the small command
ncu wraps (`--set full -k regex:<kernel>`).   python scripts/run_layer.py gru 256 256 100 8192 [--mode tf32] [--iters 3]"""
import argparse, sys
from pathlib import Path
import numpy as np, torch
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers, nntoolkitcore_b200 as nn
from nntoolkitcore_b200 import layers as L

ap = argparse.ArgumentParser()
ap.add_argument("cell"); ap.add_argument("in_size", type=int); ap.add_argument("H", type=int); ap.add_argument("T", type=int); ap.add_argument("B", type=int)
ap.add_argument("--mode", default="tf32"); ap.add_argument("--iters", type=int, default=3); ap.add_argument("--scale", type=float, default=1.0)
a = ap.parse_args()
lib = nn.lib(); assert lib.NNTKInit(0) == 0
dev = torch.device("cuda:0")
lib.NNTKSetStream(torch.cuda.current_stream().cuda_stream)
assert lib.NNTKSetMathMode({"fp32": 0, "tf32": 1, "x3": 2}[a.mode]) == 0
x = (torch.randn(a.B, a.T, a.in_size, device=dev) * a.scale).contiguous()
r = L.Recurrent(lib, a.cell, a.in_size, a.H, a.T); helpers.fill_recurrent(r, 502); L.freeze(lib, r)
y = torch.empty(a.B, a.T, a.H, device=dev)
r.apply_batch(x, y); torch.cuda.synchronize()
ts = []
for _ in range(a.iters):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r.apply_batch(x, y); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
print(f"{a.cell} in={a.in_size} H={a.H} T={a.T} B={a.B} mode={a.mode}: {np.median(ts):.3f} ms ({np.median(ts) * 1e3 / a.T:.2f} us/step), path {lib.NNTKDebugLastRecurrentPath().decode()}")
