"""Error of the TF32-mode per-step recurrent kernel against the CPU reference at C4's length (T = 500, H = 512)."""
import sys, os
import numpy as np
import torch
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import nntoolkitcore_b200 as nn
from nntoolkitcore_b200 import layers as L
import helpers

lib = nn.lib()
assert lib.NNTKInit(0) == 0
assert lib.NNTKSetMathMode(1) == 0
ref = helpers.reference_lib()
for cell, in_size, H, T, B in [("lstm", 512, 512, 500, 4096), ("gru", 256, 512, 500, 4096), ("gru", 128, 256, 994, 3840), ("lstm", 128, 256, 994, 1024)]:
    rng = np.random.default_rng(B)
    xs = rng.standard_normal((B, T, in_size)).astype(np.float32)
    g = L.Recurrent(lib, cell, in_size, H, T, True, True); helpers.fill_recurrent(g, 7)
    L.freeze(lib, g)
    xd = torch.from_numpy(xs).cuda(); yd = torch.empty(B, T, H, device="cuda")
    g.apply_batch(xd, yd)           # device-resident: the whole batch reaches the per-step kernel
    torch.cuda.synchronize()
    got = yd.cpu().numpy()
    errs = []
    for b in (0, B // 5, B // 3, B // 2, (3 * B) // 4, B - 1):
        r = L.Recurrent(ref, cell, in_size, H, T, True); helpers.copy_weights(r, g)
        want = r.apply(xs[b]); r.close()
        errs.append(float(np.abs(got[b] - want).max() / max(1.0, np.abs(want).max())))
    g.close()
    print(cell, "max abs err (scale 1):", errs, flush=True)
