"""GPU parity tests: the CUDA path, called through the C-ABI, against the unmodified reference (oracle/_ref) on the
same seeded inputs.  Tolerance: 1e-5 on the metric of helpers.parity (north_star fp32 contract)."""
import numpy as np
import pytest

import helpers
from helpers import assert_parity
from nntoolkitcore_b200 import layers as L

pytestmark = pytest.mark.gpu


def _to_dev(cuda, a):
    import torch
    return torch.from_numpy(a).to(cuda)


# ------------------------------------------------------------------------------------------------ spectrogram (K1)
SPEC_CASES = [
    # nfft, win, noverlap, n, window, psd_fs, norm
    (512, 512, 352, 16000, "hann_window", None, 1.0),      # C2 shape (1 s)
    (512, 400, 240, 16000, "hann_window", None, 1.0),      # C5 shape: window < nfft, zero padded
    (512, 512, 352, 16000, "hamming_window", 16000, 1.0),  # PSD
    (512, 400, 240, 8000, "blackman_window", 16000, 0.5),  # PSD + fft_normalization_factor
    (512, 512, 0, 5120, "ones", None, 2.0),                # no overlap, default window, norm
    (512, 512, 511, 1200, "periodic_hann_window", None, 1.0),  # hop 1
    (512, 300, 17, 4001, "periodic_hamming_window", None, 1.0),  # ragged: odd sizes, unaligned frame starts
    (256, 256, 128, 4096, "hann_window", None, 1.0),       # generic power-of-two path
    (64, 48, 16, 1000, "hann_window", 8000, 1.0),
    (100, 100, 50, 2000, "hann_window", None, 1.0),        # generic non-power-of-two path (direct DFT)
    (1024, 1024, 512, 8192, "hann_window", None, 1.0),
]


@pytest.mark.parametrize("nfft,win,nov,n,window,fs,norm", SPEC_CASES)
def test_spectrogram_matches_reference(cuda, product, ref, nfft, win, nov, n, window, fs, norm):
    clips = np.stack([helpers.audio_clip(i, n) for i in range(3)])
    s_ref = L.Spectrogram(ref, nfft, win, nov, n, norm, window, fs)
    want = np.stack([s_ref.apply(c) for c in clips])
    s = L.Spectrogram(product, nfft, win, nov, n, norm, window, fs)
    assert s.out_shape == s_ref.out_shape
    got_host = s.apply_batch(clips)                       # host buffers through the C-ABI
    got_dev = s.apply_batch(_to_dev(cuda, clips)).cpu().numpy()   # device buffers
    assert_parity(got_host, want, what="spectrogram host path")
    assert np.array_equal(got_host, got_dev)
    one = s.apply(clips[1])                               # the reference's own single-clip call
    assert np.array_equal(one, got_host[1])
    s.close(); s_ref.close()


def test_spectrogram_analytic(cuda, product):
    """impulse at in-frame offset j -> flat spectrum w[j]/sum(w); DC input -> out[0] == 1"""
    n, nfft = 2048, 512
    s = L.Spectrogram(product, nfft, nfft, nfft - 160, n, 1.0, "hann_window")
    w = np.zeros(nfft, np.float32)
    product.hann_window(helpers.fptr(w), nfft)
    x = np.zeros((2, n), np.float32)
    x[0, 100] = 1.0
    x[1, :] = 1.0
    out = s.apply_batch(x)
    assert np.allclose(out[0, 0], w[100] / w.sum(), rtol=1e-5, atol=1e-7)
    assert np.allclose(out[1, :, 0], 1.0, rtol=1e-5)
    s.close()


def test_spectrogram_c2_full_clip_and_tile_edges(cuda, product, ref):
    """one full 10 s clip (997 frames = 31 tiles of 32 + one of 5, odd frame count)"""
    x = helpers.audio_clip(7)
    s_ref = L.Spectrogram(ref, 512, 512, 352, x.size)
    s = L.Spectrogram(product, 512, 512, 352, x.size)
    assert s.out_shape == (997, 257)
    assert_parity(s.apply(x), s_ref.apply(x), what="C2 clip")
    s.close(); s_ref.close()


def test_spectrogram_empty_batch(cuda, product):
    s = L.Spectrogram(product, 512, 512, 352, 16000)
    out = s.apply_batch(np.zeros((0, 16000), np.float32))
    assert out.shape == (0, 97, 257)
    s.close()


# ------------------------------------------------------------------------------------------------ element-wise (K5)
@pytest.mark.parametrize("C_,count", [(256, 50), (257, 33), (3, 7)])
def test_batch_norm_bit_exact(cuda, product, ref, C_, count):
    rng = np.random.default_rng(5)
    x = rng.standard_normal((4, count, C_)).astype(np.float32)
    bn_ref, bn = L.BatchNorm(ref, C_, 1e-3, count), L.BatchNorm(product, C_, 1e-3, count)
    helpers.fill_bn(bn_ref, 11); helpers.copy_weights(bn, bn_ref)
    want = np.stack([bn_ref.apply(v) for v in x])
    got = bn.apply_batch(x)
    # same six fp32 operations in the same order (batch_norm.c:151-162): bit-exact
    assert np.array_equal(got, want)
    assert np.array_equal(bn.apply(x[2]), want[2])
    assert np.array_equal(bn.apply_batch(_to_dev(cuda, x)).cpu().numpy(), want)
    bn.close(); bn_ref.close()


@pytest.mark.parametrize("kind,a", [(L.ACT_SIGMOID, 1.0), (L.ACT_TANH, 1.0), (L.ACT_IDENTITY, 1.0), (L.ACT_RELU, 1.0),
                                    (L.ACT_RELU, 0.25)])
def test_activation_matches_reference(cuda, product, ref, kind, a):
    rng = np.random.default_rng(6)
    x = (3 * rng.standard_normal((5, 1003))).astype(np.float32)
    a_ref, a_new = L.Activation(ref, kind, 1003, a), L.Activation(product, kind, 1003, a)
    want = np.stack([a_ref.apply(v) for v in x])
    got = a_new.apply(x, batch=5)
    assert_parity(got, want, what="activation")
    assert np.array_equal(a_new.apply(x[0]), got[0])
    a_ref.close(); a_new.close()


@pytest.mark.parametrize("V", [12, 40, 1])
def test_softmax_matches_reference(cuda, product, ref, V):
    rng = np.random.default_rng(7)
    x = (2 * rng.standard_normal((6, 9, V))).astype(np.float32)
    a_ref, a_new = L.Activation(ref, L.ACT_SOFTMAX, 9, vector_size=V), L.Activation(product, L.ACT_SOFTMAX, 9, vector_size=V)
    want = np.stack([a_ref.apply(v) for v in x])
    got = a_new.apply(x, batch=6)
    assert_parity(got, want, what="softmax")
    a_ref.close(); a_new.close()
