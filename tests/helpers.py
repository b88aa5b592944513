"""Test infrastructure: loads the checkers (oracle/_ref reference build and the oracle C port), synthetic inputs and
the parity metric of SURVEY.md (|a-b| <= tol * max(|ref|_inf per row, 1) and rel-L2 <= tol)."""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE = ROOT / "oracle"
REF_SO = ORACLE / "_ref" / "libnntk_ref.so"
PORT_SO = ORACLE / "liboracle_port.so"
FP32_TOL = 1e-5   # north_star: fp32 path within 1e-5 relative
TF32_TOL = 1e-3   # north_star: TF32/bf16 tensor-core path within 1e-3
DEFAULT_MATH_MODE = int(__import__("os").environ.get("NNTK_MATH_MODE", "2"))   # the library's process default (b200_ext.h)

_ref = _port = None


def _make(target):
    subprocess.run(["make", "-C", str(ORACLE), target], check=True, capture_output=True)


def reference_lib():
    global _ref
    if _ref is None:
        if not REF_SO.exists():
            _make("ref")
        if not REF_SO.exists():
            raise RuntimeError("oracle/_ref/libnntk_ref.so missing and /root/reference is not available to build it")
        from nntoolkitcore_b200 import capi
        _ref = capi.bind(str(REF_SO), ext=False)
    return _ref


def oracle_port():
    global _port
    if _port is None:
        if not PORT_SO.exists():
            _make("port")
        _port = C.CDLL(str(PORT_SO))
    return _port


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def parity(got, want, tol, rows_axis=-1):
    """Returns (max scaled abs error, rel-L2).  Scaled error: |a-b| / max(|ref|_inf over the row, 1)."""
    got = np.asarray(got, np.float64)
    want = np.asarray(want, np.float64)
    assert got.shape == want.shape, (got.shape, want.shape)
    assert np.isfinite(got).all(), "non-finite values in the output"
    if want.size == 0:
        return 0.0, 0.0
    scale = np.maximum(np.abs(want).max(axis=rows_axis, keepdims=True), 1.0)
    err = float((np.abs(got - want) / scale).max())
    denom = float(np.linalg.norm(want))
    l2 = float(np.linalg.norm(got - want) / denom) if denom > 0 else float(np.linalg.norm(got - want))
    return err, l2


def assert_parity(got, want, tol=FP32_TOL, what=""):
    err, l2 = parity(got, want, tol)
    assert err <= tol and l2 <= tol, f"{what}: scaled max err {err:.3e}, rel-L2 {l2:.3e} exceed {tol:g}"
    return err, l2


# ---- synthetic data ---------------------------------------------------------------------------
def audio_clip(i, n=160000):
    rng = np.random.Generator(np.random.PCG64(20000 + i))
    f = rng.uniform(100, 4000)
    t = np.arange(n)
    x = 0.1 * rng.standard_normal(n) + 0.3 * np.sin(2 * np.pi * f * t / 16000.0)
    return np.clip(x, -1, 1).astype(np.float32)


def glorot(rng, fan_in, fan_out, shape):
    lim = np.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, shape).astype(np.float32)


def fill_recurrent(layer, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    G = layer.W.shape[1] // layer.H
    layer.W[:] = glorot(rng, layer.in_size, layer.H, layer.W.shape)
    layer.U[:] = glorot(rng, layer.H, layer.H, layer.U.shape)
    layer.b_i[:] = (0.05 * rng.standard_normal(G * layer.H)).astype(np.float32)
    layer.b_h[:] = (0.05 * rng.standard_normal(G * layer.H)).astype(np.float32)


def fill_conv(layer, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    cout, cin, k = layer.W.shape
    layer.W[:] = glorot(rng, cin * k, cout * k, layer.W.shape)
    layer.b[:] = (0.05 * rng.standard_normal(cout)).astype(np.float32)


def fill_dense(layer, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    layer.W[:] = glorot(rng, layer.in_size, layer.out_size, layer.W.shape)
    layer.b[:] = (0.05 * rng.standard_normal(layer.out_size)).astype(np.float32)


def fill_bn(layer, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    C_ = layer.C
    layer.set(rng.uniform(0.5, 1.5, C_).astype(np.float32), (0.1 * rng.standard_normal(C_)).astype(np.float32),
              (0.1 * rng.standard_normal(C_)).astype(np.float32), rng.uniform(0.5, 1.5, C_).astype(np.float32))


def copy_weights(dst, src):
    """copy every numpy weight view of `src` (a layers.* object) into `dst` (same class, other library)"""
    for name in ("W", "U", "b", "b_i", "b_h", "gamma", "beta", "mean", "var"):
        if hasattr(src, name):
            getattr(dst, name)[...] = getattr(src, name)
