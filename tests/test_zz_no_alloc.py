"""GPU tier, collected LAST on purpose: the allocation-free recurrent entry point
nntk_cuda_recurrent_static_scratch (include/nntk_b200_shim.h) against the existing nntk_cuda_recurrent.

The entry point has no counterpart in the reference; it was added for callers that cannot
call cudaMalloc.  Its scratch is a file-scope device array of NNTK_RECURRENT_STATIC_SCRATCH_BYTES (32 MiB) shared by every call,
so the test makes two calls in a row with DIFFERENT shapes and inputs (the second must not see the first one's leftovers), checks
each against the existing entry point run on a caller-owned workspace -- through the C-ABI, same device pointers -- and against
the CPU oracle, and checks that a problem beyond the array is rejected with a clear error instead of running.
"""
import ctypes as C

import numpy as np
import pytest

import helpers
from helpers import assert_parity, FP32_TOL

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_BYTES = 32 << 20          # NNTK_RECURRENT_STATIC_SCRATCH_BYTES
ACT_SIGMOID, ACT_TANH = 2, 3
GATES = {"gru": 3, "lstm": 4}


class RecurrentArgs(C.Structure):        # nntk_recurrent_args, include/nntk_b200_shim.h
    _fields_ = [("cell", C.c_int), ("in_", C.c_int), ("H", C.c_int), ("T", C.c_int), ("batch", C.c_int),
                ("return_sequences", C.c_int), ("v2", C.c_int), ("act", C.c_int * 5), ("act_a", C.c_float * 5),
                ("d_W", C.c_void_p), ("d_U", C.c_void_p), ("d_bi", C.c_void_p), ("d_bh", C.c_void_p), ("d_x", C.c_void_p),
                ("d_h", C.c_void_p), ("d_c", C.c_void_p), ("d_out", C.c_void_p), ("d_work", C.c_void_p), ("work_bytes", C.c_size_t)]


def _bind(lib):
    for name in ("nntk_cuda_recurrent", "nntk_cuda_recurrent_static_scratch"):
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = C.c_int, [C.POINTER(RecurrentArgs)]
    lib.nntk_cuda_recurrent_workspace.restype = C.c_size_t
    lib.nntk_cuda_recurrent_workspace.argtypes = [C.POINTER(RecurrentArgs)]
    lib.nntk_cuda_last_error.restype = C.c_char_p
    return lib


def _args(cell, in_size, H, T, B, rs, dev_w, x, h, c, out):
    a = RecurrentArgs()
    a.cell, a.in_, a.H, a.T, a.batch, a.return_sequences, a.v2 = GATES[cell], in_size, H, T, B, int(rs), 1
    acts = [ACT_SIGMOID, ACT_SIGMOID, ACT_TANH, 0, 0] if cell == "gru" else [ACT_SIGMOID, ACT_SIGMOID, ACT_TANH, ACT_SIGMOID, ACT_TANH]
    for i in range(5):
        a.act[i] = acts[i]
        a.act_a[i] = 1.0
    a.d_W, a.d_U, a.d_bi, a.d_bh = (t.data_ptr() for t in dev_w)
    a.d_x, a.d_h, a.d_out = x.data_ptr(), h.data_ptr(), out.data_ptr()
    a.d_c = c.data_ptr() if c is not None else None
    return a


def _both(lib, port, cell, in_size, H, T, B, rs, seed):
    """one problem through the static-scratch entry point and through the existing one; returns nothing, asserts parity"""
    import torch
    import port_cases
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(seed)
    w = port_cases.recurrent_weights(cell, in_size, H, seed)
    dev_w = [torch.from_numpy(np.ascontiguousarray(v)).to(dev) for v in (w.W, w.U, w.b_i, w.b_h)]
    xh = rng.standard_normal((B, T, in_size)).astype(np.float32)
    h0 = (0.1 * rng.standard_normal((B, H))).astype(np.float32)
    c0 = (0.1 * rng.standard_normal((B, H))).astype(np.float32)
    x = torch.from_numpy(xh).to(dev)
    oshape = (B, T, H) if rs else (B, H)
    res = {}
    for which in ("static", "existing"):
        h = torch.from_numpy(h0).to(dev)
        c = torch.from_numpy(c0).to(dev) if cell == "lstm" else None
        out = torch.full(oshape, float("nan"), device=dev)
        a = _args(cell, in_size, H, T, B, rs, dev_w, x, h, c, out)
        if which == "existing":
            work = torch.empty(int(lib.nntk_cuda_recurrent_workspace(C.byref(a))) // 4 + 64, device=dev)
            a.d_work, a.work_bytes = work.data_ptr(), work.numel() * 4
            rc = lib.nntk_cuda_recurrent(C.byref(a))
        else:
            rc = lib.nntk_cuda_recurrent_static_scratch(C.byref(a))
        assert rc == 0, (which, lib.nntk_cuda_last_error())
        torch.cuda.synchronize()
        res[which] = (out.cpu().numpy(), h.cpu().numpy(), None if c is None else c.cpu().numpy())
    what = f"{cell} in={in_size} H={H} T={T} B={B} rs={rs}"
    assert np.isfinite(res["static"][0]).all(), f"{what}: output not fully written"
    assert_parity(res["static"][0], res["existing"][0], tol=FP32_TOL, what=f"static-scratch vs existing entry point, output, {what}")
    assert_parity(res["static"][1], res["existing"][1], tol=FP32_TOL, what=f"static-scratch vs existing entry point, h_T, {what}")
    if cell == "lstm":
        assert_parity(res["static"][2], res["existing"][2], tol=FP32_TOL, what=f"static-scratch vs existing entry point, c_T, {what}")
    for b in (0, B - 1):                 # and against the CPU oracle (reference arithmetic, gru.c:129-204 / lstm.c:185-268)
        want = port_cases.recurrent(port, cell, in_size, H, T, rs, True, w, xh[b], h=h0[b].copy(), c=c0[b].copy())
        assert_parity(res["static"][0][b], want, tol=FP32_TOL, what=f"static-scratch vs oracle, {what}, sequence {b}")


def test_static_scratch_entry_point_two_calls_in_a_row(cuda, product, port):
    lib = _bind(product)
    mode = product.NNTKSetMathMode(0)    # the existing entry point on its fp32 path too: the two then run the same kernels
    assert mode == 0
    try:
        # call 1: GRU, sequences returned, a shape the persistent fp32 cluster kernel takes; fills 5.3 MB of the array
        _both(lib, port, "gru", 40, 128, 27, 9, True, 600)
        # call 2 right behind it: other cell, other shape, last state only (ping-pong buffers in the scratch), larger footprint
        _both(lib, port, "lstm", 24, 64, 19, 300, False, 601)
        # and a third that runs the per-step FFMA kernels (batch above the cluster kernel's limit) over the same array
        _both(lib, port, "gru", 32, 96, 11, 700, True, 602)
    finally:
        product.NNTKSetMathMode(helpers.DEFAULT_MATH_MODE)


def test_static_scratch_entry_point_in_the_default_math_mode(cuda, product, port):
    """whatever the process math mode, the entry point runs its fp32 path and restores the mode"""
    lib = _bind(product)
    assert product.NNTKSetMathMode(helpers.DEFAULT_MATH_MODE) == 0
    _both(lib, port, "gru", 40, 128, 20, 5, True, 603)
    lib.nntk_cuda_math_mode.restype = C.c_int
    assert lib.nntk_cuda_math_mode() == helpers.DEFAULT_MATH_MODE


def test_static_scratch_entry_point_rejects_what_does_not_fit(cuda, product):
    import torch
    lib = _bind(product)
    dev = torch.device("cuda:0")
    cell, in_size, H, T, B = "gru", 8, 128, 100, 300          # 300 * 100 * 384 * 4 = 46 MB > 32 MiB
    assert (B * T * 3 * H + 6 * B * H) * 4 > STATIC_SCRATCH_BYTES
    dev_w = [torch.zeros(n, device=dev) for n in (in_size * 3 * H, H * 3 * H, 3 * H, 3 * H)]
    x = torch.zeros(B, T, in_size, device=dev); h = torch.zeros(B, H, device=dev)
    out = torch.full((B, T, H), 7.0, device=dev)
    a = _args(cell, in_size, H, T, B, True, dev_w, x, h, None, out)
    rc = lib.nntk_cuda_recurrent_static_scratch(C.byref(a))
    msg = lib.nntk_cuda_last_error().decode()
    assert rc != 0 and "static array" in msg and "NNTK_RECURRENT_STATIC_SCRATCH_BYTES" in msg, (rc, msg)
    torch.cuda.synchronize()
    assert float(out.min()) == 7.0 and float(out.max()) == 7.0, "a rejected call must not touch the output"
