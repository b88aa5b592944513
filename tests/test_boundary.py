"""CPU tier: the drop-in boundary.  The C-ABI library loads without a GPU, exports every symbol the headers declare,
its host-side set-up logic (ConfigCreate arithmetic, window tables, weight block layout) equals the reference's, and
every compute entry point fails loudly -- no CPU fallback -- when no CUDA device exists."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

import helpers
from nntoolkitcore_b200 import capi, layers as L

ROOT = Path(__file__).resolve().parent.parent
DECL = re.compile(r"^[A-Za-z_][\w \*]*?\b(\w+)\s*\(", re.M)


def _declared_functions(header: Path):
    text = re.sub(r"/\*.*?\*/", "", header.read_text(), flags=re.S)
    text = re.sub(r"^\s*#.*$", "", text, flags=re.M)
    text = re.sub(r"typedef[^;{]*\([^;]*;", "", text)          # function-pointer typedefs
    text = re.sub(r"\{[^{}]*\}", "", text)                      # struct / enum bodies
    return {m.group(1) for m in DECL.finditer(text)} - {"defined", "sizeof"}


def test_every_declared_symbol_is_exported(product):
    headers = sorted((ROOT / "include").rglob("*.h"))
    assert len(headers) >= 14
    missing = []
    for h in headers:
        for name in _declared_functions(h):
            if not hasattr(product, name):
                missing.append(f"{h.name}:{name}")
    assert not missing, missing


def test_binding_tables_cover_the_headers(product):
    declared = set()
    for h in (ROOT / "include").rglob("*.h"):
        declared |= _declared_functions(h)
    bound = set(capi.REFERENCE_API) | set(capi.EXT_API) | set(capi.SHIM_SYMBOLS)
    assert declared <= bound, sorted(declared - bound)


def test_no_oracle_or_reference_code_in_the_product(product):
    import subprocess
    import nntoolkitcore_b200 as nn
    syms = subprocess.run(["nm", "-D", "--defined-only", str(nn.LIB_PATH)], capture_output=True, text=True).stdout
    assert "oracle_" not in syms and "kiss_fft" not in syms and "op_mat_mul" not in syms
    deps = subprocess.run(["ldd", str(nn.LIB_PATH)], capture_output=True, text=True).stdout
    assert "oracle" not in deps and "nntk_ref" not in deps


@pytest.mark.parametrize("args", [(512, 512, 352, 160000, 1.0), (512, 400, 240, 160000, 1.0), (256, 200, 13, 999, 0.5),
                                  (100, 100, 0, 1000, 2.0)])
def test_spectrogram_config_matches_reference(product, ref, args):
    a, b = product.SpectrogramConfigCreate(*args), ref.SpectrogramConfigCreate(*args)
    for f, _ in capi.SpectrogramConfig._fields_:
        assert getattr(a, f) == getattr(b, f), f


@pytest.mark.parametrize("args", [(128, 256, 5, 1, 1000), (257, 256, 5, 1, 998), (16, 24, 3, 2, 41), (32, 32, 15, 4, 5511)])
def test_conv_config_matches_reference(product, ref, args):
    a, b = product.Conv1dConfigCreate(*args), ref.Conv1dConfigCreate(*args)
    for f, _ in capi.Conv1dConfig._fields_:
        assert getattr(a, f) == getattr(b, f), f


def test_window_tables_bit_exact(product, ref):
    for name in L.Spectrogram.WINDOWS:
        for size in (2, 400, 512, 777):
            a = np.zeros(size, np.float32); b = np.zeros(size, np.float32)
            getattr(product, name)(helpers.fptr(a), size); getattr(ref, name)(helpers.fptr(b), size)
            assert np.array_equal(a, b), (name, size)


def test_weight_blocks_are_contiguous_like_the_reference(product):
    """W|U|b_i|b_h, W|b and gamma|beta|mean|var are single blocks, so one memcpy loads a layer"""
    g = L.Recurrent(product, "gru", 40, 128, 100)
    base = g.W.ctypes.data
    assert g.U.ctypes.data == base + 4 * 40 * 384
    assert g.b_i.ctypes.data == g.U.ctypes.data + 4 * 128 * 384
    assert g.b_h.ctypes.data == g.b_i.ctypes.data + 4 * 384
    assert not g.W.any() and not g.b_h.any()          # zero-filled like f_malloc
    c = L.Conv1d(product, 8, 16, 3, 1, 20)
    assert c.b.ctypes.data == c.W.ctypes.data + 4 * 8 * 16 * 3
    bn = L.BatchNorm(product, 7, 1e-3, 5)
    assert bn.var.ctypes.data == bn.gamma.ctypes.data + 4 * 21
    for x in (g, c, bn):
        x.close()


def test_lstm_activation_struct_order(product):
    """struct order (candidate, input, forget, output_gate, output) differs from the Create argument order"""
    acts = product.LSTMActivationsCreate(1, 2, 3, 4, 5)
    assert (acts.input_gate_activation, acts.forget_gate_activation, acts.candidate_gate_activation,
            acts.output_gate_activation, acts.output_activation) == (1, 2, 3, 4, 5)
    g = product.GRUActivationsCreate(1, 2, 3)
    assert (g.z_gate_activation, g.h_gate_activation, g.r_gate_activation) == (1, 2, 3)


def test_invalid_configs_are_refused(product):
    assert not product.SpectrogramCreateMagnitude(product.SpectrogramConfigCreate(256, 512, 0, 4096, 1.0))  # window > nfft
    assert not product.Conv1dCreateForInference(product.Conv1dConfigCreate(4, 4, 3, 0, 10))                 # stride 0


def test_apply_fails_loudly_without_a_gpu(product):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    s = L.Spectrogram(product, 512, 512, 352, 16000)
    with pytest.raises(L.NNTKError, match="no CUDA device"):
        s.apply_batch(np.zeros((1, 16000), np.float32))
    g = L.Recurrent(product, "gru", 8, 8, 4)
    with pytest.raises(L.NNTKError):
        g.apply(np.zeros((4, 8), np.float32))
    assert product.NNTKInit(0) == -3
    assert b"no CPU fallback" in product.NNTKLastError()


def test_custom_activation_runs_callers_callback_on_host_buffers(product):
    seen = []

    def cb(impl, src, dst, n):
        seen.append(n)
        C.memmove(dst, src, 4 * n)

    fn = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int)(cb)
    act = product.ActivationFunctionCreate(6, None, None, C.cast(fn, C.c_void_p), None, None)
    x = np.arange(6, dtype=np.float32); y = np.zeros(6, np.float32)
    product.ActivationFunctionApply(act, x.ctypes.data, y.ctypes.data)
    assert seen == [6] and np.array_equal(x, y)
    product.ActivationFunctionDestroy(act)


@pytest.mark.parametrize("args", [(40, 512, 16000, 20.0, 8000.0), (80, 512, 16000, 0.0, 8000.0), (128, 512, 16000, 0.0, 8000.0),
                                  (23, 512, 8000, 100.0, 3800.0), (376, 512, 16000, 0.0, 8000.0), (1, 512, 16000, 0.0, 8000.0)])
def test_mel_work_list_reproduces_the_table(product, args):
    """host logic of the fused log-mel kernel (include/nntk_b200_shim.h, nntk_mel_args): walking the per-lane work list the
    way the kernel does -- accumulate weight * spectrum[slot], emit at flagged entries -- gives the dense table product for
    every band exactly once, and the list is balanced and nearly free of shared-memory bank conflicts"""
    n_mels, n_fft = args[0], args[1]
    nb = n_fft // 2 + 1

    class Head(C.Structure):   # leading fields of the product's private struct (host/internal.h)
        _fields_ = [("cfg", capi.MelFilterBankConfig), ("weights", C.POINTER(C.c_float)), ("entries", C.c_void_p),
                    ("n_iter", C.c_int)]
    m = L.MelFilterBank(product, *args)
    head = C.cast(m.handle, C.POINTER(Head)).contents
    w = np.ctypeslib.as_array(head.weights, shape=(nb, n_mels)).copy()
    n_iter = head.n_iter
    ent = np.ctypeslib.as_array(C.cast(head.entries, C.POINTER(C.c_uint32)), shape=(n_iter, 32, 2)).copy()
    m.close()
    packed, weight = ent[..., 0], ent[..., 1].copy().view(np.float32)
    spec = np.abs(np.random.default_rng(1).standard_normal(nb)).astype(np.float64) + 0.1
    parked = np.zeros(nb + nb // 16 + 16)
    k = np.arange(nb)
    parked[k + k // 16] = spec
    got = np.full(n_mels, np.nan)
    for lane in range(32):
        acc = 0.0
        for it in range(n_iter):
            p = int(packed[it, lane])
            acc += float(weight[it, lane]) * parked[(p & 0xff8) // 8]
            if p & 0x80000000:
                band = (p >> 15) & 0xffff
                assert np.isnan(got[band]), f"band {band} emitted twice"
                got[band] = acc
                acc = 0.0
    assert not np.isnan(got).any(), "a band was never emitted"
    np.testing.assert_allclose(got, spec @ w.astype(np.float64), rtol=1e-12, atol=1e-12)
    work = int(np.maximum((w != 0).sum(0), 1).sum())                                       # an empty band still costs one entry
    assert n_iter <= max(int(np.ceil(work / 32)) + 2, int((w != 0).sum(0).max()))         # longest-first balance
    # 8-byte accesses: a half-warp is conflict-free when its 16 slots differ mod 16
    slots = (packed & 0xff8) // 8
    waves = sum(np.bincount(slots[it, h * 16:(h + 1) * 16] % 16, minlength=16).max() for it in range(n_iter) for h in (0, 1))
    if n_mels <= 128:   # (hundreds of 1-4 bin bands leave the ordering no choice)
        assert waves <= 2.2 * 2 * n_iter, f"{waves} wavefronts for {n_iter} iterations"
