#!/usr/bin/env python
"""Development probe: structure of the 3xTF32 (NNTK_MATH_TF32X3) error against the unmodified reference on conv shapes.
Prints rel-L2, the least-squares slope of (got - want) on want (a systematic shrink = truncating accumulate), and the rel-L2
left after removing that slope."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402
import nntoolkitcore_b200 as nn  # noqa: E402
from nntoolkitcore_b200 import layers as L  # noqa: E402

lib = nn.lib(); ref = helpers.reference_lib()
assert lib.NNTKInit(0) == 0
for cin, cout, k, T, B in [(128, 256, 5, 300, 2), (256, 256, 5, 130, 2), (256, 256, 1, 130, 2), (512, 2048, 1, 64, 2), (64, 96, 3, 257, 2), (512, 256, 5, 40, 2), (1024, 128, 7, 30, 1), (2048, 64, 7, 20, 1)]:
    rng = np.random.default_rng(3)
    x = rng.standard_normal((B, T, cin)).astype(np.float32)
    c_ref, c_new = L.Conv1d(ref, cin, cout, k, 1, T), L.Conv1d(lib, cin, cout, k, 1, T)
    helpers.fill_conv(c_ref, 300); helpers.copy_weights(c_new, c_ref)
    want = np.stack([c_ref.apply(v) for v in x]).astype(np.float64)
    for mode in (0, 2):
        lib.NNTKSetMathMode(mode)
        got = c_new.apply_batch(x).astype(np.float64)
        e = got - want
        slope = float((e * want).sum() / (want * want).sum())
        l2 = float(np.linalg.norm(e) / np.linalg.norm(want))
        l2c = float(np.linalg.norm(e - slope * want) / np.linalg.norm(want))
        print(f"K={cin * k:5d} N={cout:5d} mode {mode}: rel-L2 {l2:.3e}  slope {slope:+.3e}  rel-L2 without slope {l2c:.3e}")
    lib.NNTKSetMathMode(0)
