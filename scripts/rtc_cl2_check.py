"""Runs the per-step recurrent kernel (TF32 mode) on fixed inputs and saves the outputs: run once with NNTK_RTC_CL2=0
and once with =1 -- the pair mode does the same arithmetic in the same order, so the two files must be bit-identical."""
import sys, os
import numpy as np
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import nntoolkitcore_b200 as nn
from nntoolkitcore_b200 import layers as L
import helpers

lib = nn.lib()
assert lib.NNTKInit(0) == 0
assert lib.NNTKSetMathMode(1) == 0
outs = {}
for cell, in_size, H, T, B in [("lstm", 64, 512, 12, 1600), ("gru", 40, 512, 9, 1700), ("lstm", 32, 320, 7, 1920), ("gru", 64, 384, 6, 2500)]:
    rng = np.random.default_rng(B)
    xs = rng.standard_normal((B, T, in_size)).astype(np.float32)
    g = L.Recurrent(lib, cell, in_size, H, T, True); helpers.fill_recurrent(g, 7)
    outs[f"{cell}_{H}_{B}"] = g.apply_batch(xs)
    g.close()
np.savez(sys.argv[1], **outs)
print({k: float(np.abs(v).mean()) for k, v in outs.items()})
