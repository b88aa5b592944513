"""Generates tests/golden/golden_v1.npz (hot path) and golden_v2.npz from the UNMODIFIED reference (oracle/_ref/libnntk_ref.so, built from
/root/reference by oracle/Makefile).  The reference ships no tests or golden vectors of its own (README.md:25), so these
outputs of the reference itself are the pin.  Run in the build container (needs /root/reference once, to build _ref):

    python tests/golden/make_golden.py

Inputs and weights are regenerated from seeds by tests/golden_cases.py; only the reference OUTPUTS are stored.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers  # noqa: E402
import golden_cases  # noqa: E402

if __name__ == "__main__":
    ref = helpers.reference_lib()
    out = {}
    for name, case in golden_cases.CASES.items():
        out[name] = case(ref)
        print(f"{name:32s} {out[name].shape} mean {out[name].mean():+.6f}")
    np.savez_compressed(Path(__file__).parent / "golden_v1.npz", **out)
    nxt = {}
    for name, case in golden_cases.NEXT_CASES.items():
        nxt[name] = case(ref)
        print(f"{name:32s} {nxt[name].shape} mean {nxt[name].mean():+.6f}")
    np.savez_compressed(Path(__file__).parent / "golden_v2.npz", **nxt)
