"""GPU tier: robustness properties the advisor asked for: in-place calls, two host threads on two handles, the
tensor-core bias compensations outside the zero-mean regime they were fitted on, silent input through the fused log-mel."""
import threading

import numpy as np
import pytest

import helpers
from helpers import assert_parity, FP32_TOL, TF32_TOL
from nntoolkitcore_b200 import layers as L

pytestmark = pytest.mark.gpu


def _to_dev(dev, a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(dev)


@pytest.mark.parametrize("kind", [L.ACT_SIGMOID, L.ACT_TANH, L.ACT_RELU])
def test_activation_in_place(cuda, product, ref, kind):
    """the reference allows input == output (activation.c:23-25 applies in place through its scratch); so must the kernels"""
    rng = np.random.default_rng(1)
    x = rng.standard_normal(1000).astype(np.float32)
    want = L.Activation(ref, kind, 1000, a=2.0).apply(x.copy())
    a = L.Activation(product, kind, 1000, a=2.0)
    d = _to_dev(cuda, x)
    a.apply(d, d)
    assert_parity(d.cpu().numpy(), want, what="in-place activation (device)")
    h = x.copy()
    a.apply(h, h)
    assert_parity(h, want, what="in-place activation (host)")


@pytest.mark.parametrize("V", [12, 40, 257])
def test_softmax_in_place_and_dense_softmax_wide(cuda, product, ref, V):
    """softmax over vectors wider than a warp, applied in place (Dense applies its softmax on the output buffer itself)"""
    rng = np.random.default_rng(2)
    x = rng.standard_normal((7, V)).astype(np.float32)
    want = L.Activation(ref, L.ACT_SOFTMAX, 7, vector_size=V).apply(x.copy())
    a = L.Activation(product, L.ACT_SOFTMAX, 7, vector_size=V)
    d = _to_dev(cuda, x)
    a.apply(d, d)
    assert_parity(d.cpu().numpy(), want, what=f"in-place softmax V={V}")
    d_ref, d_new = L.Dense(ref, 64, V, L.ACT_SOFTMAX), L.Dense(product, 64, V, L.ACT_SOFTMAX)
    helpers.fill_dense(d_ref, 9); helpers.copy_weights(d_new, d_ref)
    xi = rng.standard_normal((5, 64)).astype(np.float32)
    assert_parity(d_new.apply_batch(xi), np.stack([d_ref.apply(v) for v in xi]), what=f"dense softmax V={V}")


def test_two_host_threads_on_two_handles(cuda, product, ref):
    """distinct handles are independent across threads in the reference; here every public entry takes the library's
    API lock, so a host-buffer Apply (staging pipeline, swaps the stream) and a device-buffer Apply from another thread serialise
    instead of racing on the process-wide stream / math mode / workspace"""
    import torch
    rng = np.random.default_rng(3)
    n = 160000
    clips = np.stack([helpers.audio_clip(i % 3, n) for i in range(24)])
    spec = L.Spectrogram(product, 512, 512, 352, n, 1.0, "hann_window")
    want_spec = np.stack([L.Spectrogram(ref, 512, 512, 352, n, 1.0, "hann_window").apply(clips[i]) for i in range(3)])
    xg = rng.standard_normal((70, 40, 32)).astype(np.float32)
    gru = L.Recurrent(product, "gru", 32, 64, 40); helpers.fill_recurrent(gru, 5)
    gr = L.Recurrent(ref, "gru", 32, 64, 40); helpers.copy_weights(gr, gru)
    want_gru = gr.apply(xg[0])
    xd = _to_dev(cuda, xg)
    errs, outs = [], {}

    def host_side():
        try:
            for _ in range(4):
                outs["spec"] = spec.apply_batch(clips)
        except Exception as e:      # noqa: BLE001
            errs.append(e)

    def device_side():
        try:
            torch.cuda.set_device(0)
            for _ in range(40):
                outs["gru"] = gru.apply_batch(xd)
            torch.cuda.synchronize()
        except Exception as e:      # noqa: BLE001
            errs.append(e)

    ts = [threading.Thread(target=host_side), threading.Thread(target=device_side)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errs, errs
    product.NNTKSynchronize()
    assert_parity(outs["spec"][:3], want_spec, what="spectrogram from the host-buffer thread")
    assert np.array_equal(outs["spec"][21:24], outs["spec"][:3])
    assert_parity(outs["gru"][0].cpu().numpy(), want_gru, what="GRU from the device-buffer thread")


@pytest.mark.parametrize("mode,tol", [(2, FP32_TOL), (1, TF32_TOL)])
@pytest.mark.parametrize("cin,k", [(256, 5), (600, 5)])         # K = 1 280 (3xTF32) and 3 000 (above the 1 536 cap: FFMA in mode 2)
def test_tensor_core_compensation_outside_the_zero_mean_regime(cuda, product, ref, mode, tol, cin, k):
    """The tensor core truncates (operands in TF32 mode, the fp32 accumulate in both); the library compensates the MEAN shrink
    with constants fitted on zero-mean data (gemm_tc.cu).  Non-negative activations (post-ReLU) against constant-sign weights
    make every partial sum grow monotonically -- the worst case for a truncating accumulate"""
    rng = np.random.default_rng(4)
    T, cout, B = 64, 64, 2
    x = np.abs(rng.standard_normal((B, T, cin))).astype(np.float32)
    c_ref, c_new = L.Conv1d(ref, cin, cout, k, 1, T), L.Conv1d(product, cin, cout, k, 1, T)
    helpers.fill_conv(c_ref, 31)
    c_ref.W[:] = np.abs(c_ref.W)
    c_ref.b[:] = 0.25
    helpers.copy_weights(c_new, c_ref)
    want = np.stack([c_ref.apply(v) for v in x])
    assert product.NNTKSetMathMode(mode) == 0
    try:
        got = c_new.apply_batch(x)
    finally:
        product.NNTKSetMathMode(helpers.DEFAULT_MATH_MODE)
    assert_parity(got, want, tol=tol, what=f"conv, non-negative operands, K = {cin * k}, mode {mode}")


def test_fused_log_mel_on_silence_and_near_silence(cuda, product, ref):
    """log(x + 1.5849e-13) at x = 0 is where the MUFU log of the fused epilogue has the least headroom"""
    n = 16000
    clips = np.zeros((3, n), np.float32)
    clips[1] = 1e-7 * np.random.default_rng(5).standard_normal(n).astype(np.float32)
    clips[2, ::400] = 1e-3
    s_ref, s_new = L.Spectrogram(ref, 512, 512, 352, n, 1.0, "hann_window"), L.Spectrogram(product, 512, 512, 352, n, 1.0, "hann_window")
    f_ref, f_new = L.LogMelSpectrogram(ref, s_ref, 40, 16000, 20.0, 8000.0), L.LogMelSpectrogram(product, s_new, 40, 16000, 20.0, 8000.0)
    want = np.stack([f_ref.apply(c) for c in clips])
    got = f_new.apply_batch(clips)
    assert np.isfinite(got).all()
    assert_parity(got, want, what="fused log-mel on silent / near-silent clips")
