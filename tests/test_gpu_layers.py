"""GPU parity tests for the GEMM-shaped layers (Conv1d, Dense, TimeDistributedDense) and the recurrent cells
(RNN / GRU / LSTM), through the C-ABI, against the unmodified reference."""
import numpy as np
import pytest

import helpers
from helpers import assert_parity
from nntoolkitcore_b200 import layers as L

pytestmark = pytest.mark.gpu


def _to_dev(cuda, a):
    import torch
    return torch.from_numpy(a).to(cuda)


# ------------------------------------------------------------------------------------------------ Conv1d (K2)
CONV_CASES = [  # cin, cout, k, stride, T, batch
    (128, 256, 5, 1, 100, 3),     # C3 layer-1 shape, short T
    (257, 256, 5, 1, 60, 2),      # C5 conv: unaligned channel count
    (16, 24, 3, 2, 41, 4),        # stride 2, ragged sizes
    (8, 200, 7, 3, 100, 2),       # N not a multiple of the tile
    (5, 3, 1, 1, 9, 1),           # kernel 1
    (32, 32, 15, 4, 196, 2),      # the trigger-word model's conv shape class
]


@pytest.mark.parametrize("cin,cout,k,s,T,B", CONV_CASES)
def test_conv1d_matches_reference(cuda, product, ref, cin, cout, k, s, T, B):
    rng = np.random.default_rng(3)
    x = rng.standard_normal((B, T, cin)).astype(np.float32)
    c_ref, c_new = L.Conv1d(ref, cin, cout, k, s, T), L.Conv1d(product, cin, cout, k, s, T)
    assert c_new.out_shape == c_ref.out_shape
    helpers.fill_conv(c_ref, 300); helpers.copy_weights(c_new, c_ref)
    want = np.stack([c_ref.apply(v) for v in x])
    got = c_new.apply_batch(x)
    assert_parity(got, want, what="conv1d host path")
    assert np.array_equal(c_new.apply_batch(_to_dev(cuda, x)).cpu().numpy(), got)
    assert np.array_equal(c_new.apply(x[0]), got[0])
    c_ref.close(); c_new.close()


def test_conv1d_bn_relu_fused_equals_three_reference_calls(cuda, product, ref):
    cin, cout, k, T, B = 128, 256, 5, 64, 3
    rng = np.random.default_rng(33)
    x = rng.standard_normal((B, T, cin)).astype(np.float32)
    c_ref, c_new = L.Conv1d(ref, cin, cout, k, 1, T), L.Conv1d(product, cin, cout, k, 1, T)
    helpers.fill_conv(c_ref, 301); helpers.copy_weights(c_new, c_ref)
    t_out = c_ref.out_shape[0]
    bn_ref, bn_new = L.BatchNorm(ref, cout, 1e-3, t_out), L.BatchNorm(product, cout, 1e-3, t_out)
    helpers.fill_bn(bn_ref, 302); helpers.copy_weights(bn_new, bn_ref)
    for a in (1.0, 0.5):
        r_ref, r_new = L.Activation(ref, L.ACT_RELU, t_out * cout, a), L.Activation(product, L.ACT_RELU, t_out * cout, a)
        want = np.stack([r_ref.apply(bn_ref.apply(c_ref.apply(v))) for v in x])
        got = c_new.apply_batch(x, bn=bn_new, act=r_new)
        assert_parity(got, want, what="conv+bn+relu fused")
        # and the unfused product calls agree with the fused kernel
        unfused = r_new.apply(bn_new.apply_batch(c_new.apply_batch(x)), batch=B)
        assert_parity(unfused, got, what="fused vs unfused")
        r_ref.close(); r_new.close()


def test_conv1d_delta_kernel_is_shifted_copy(cuda, product):
    cin = cout = 8
    c = L.Conv1d(product, cin, cout, 3, 1, 20)
    c.W[:] = 0
    for ch in range(cin):
        c.W[ch, ch, 1] = 1.0
    c.b[:] = np.arange(cout, dtype=np.float32)
    x = np.random.default_rng(1).standard_normal((1, 20, cin)).astype(np.float32)
    # bit-exact copies are a property of the FFMA kernel (NNTK_MATH_FP32); the tensor-core modes round through their operand split
    assert product.NNTKSetMathMode(0) == 0
    try:
        out = c.apply_batch(x)
    finally:
        product.NNTKSetMathMode(helpers.DEFAULT_MATH_MODE)
    assert np.array_equal(out[0], x[0, 1:19] + c.b)
    c.close()


# ------------------------------------------------------------------------------------------------ Dense / TDD (K3)
@pytest.mark.parametrize("act,a", [(L.ACT_NONE, 1.0), (L.ACT_RELU, 2.0), (L.ACT_TANH, 1.0), (L.ACT_SIGMOID, 1.0),
                                   (L.ACT_SOFTMAX, 1.0)])
def test_dense_matches_reference(cuda, product, ref, act, a):
    rng = np.random.default_rng(8)
    x = rng.standard_normal((7, 256)).astype(np.float32)
    d_ref, d_new = L.Dense(ref, 256, 12, act, a), L.Dense(product, 256, 12, act, a)
    helpers.fill_dense(d_ref, 500); helpers.copy_weights(d_new, d_ref)
    want = np.stack([d_ref.apply(v) for v in x])
    got = d_new.apply_batch(x)
    assert_parity(got, want, what="dense")
    assert np.array_equal(d_new.apply(x[3]), got[3])
    d_ref.close(); d_new.close()


def test_time_distributed_dense_softmax(cuda, product, ref):
    ts, B = 50, 3
    rng = np.random.default_rng(9)
    x = rng.standard_normal((B, ts, 256)).astype(np.float32)
    d_ref = L.Dense(ref, 256, 12, L.ACT_SOFTMAX, ts=ts)
    d_new = L.Dense(product, 256, 12, L.ACT_SOFTMAX, ts=ts)
    helpers.fill_dense(d_ref, 501); helpers.copy_weights(d_new, d_ref)
    want = np.stack([d_ref.apply(v) for v in x])
    got = d_new.apply_batch(x)
    assert_parity(got, want, what="tdd softmax")
    assert np.allclose(got.sum(-1), 1.0, atol=1e-5)
    assert np.array_equal(d_new.apply_batch(_to_dev(cuda, x)).cpu().numpy(), got)
    d_ref.close(); d_new.close()


# ------------------------------------------------------------------------------------------------ recurrent (K3 + K4)
def _ref_batch(ref, cell, in_size, H, T, rs, v2, x, seed, acts=None):
    """batched reference = loop of single-sequence calls on FRESH handles (zero state), SURVEY.md 8c"""
    outs = []
    for v in x:
        r = L.Recurrent(ref, cell, in_size, H, T, rs, v2, acts)
        helpers.fill_recurrent(r, seed)
        outs.append(r.apply(v))
        r.close()
    return np.stack(outs)


REC_CASES = [  # cell, in, H, T, batch, return_sequences, v2
    ("gru", 40, 128, 100, 1, True, True),      # C1
    ("gru", 40, 128, 30, 5, False, True),
    ("gru", 256, 256, 12, 130, True, True),    # C5 layer shape, batch > one tile
    ("lstm", 64, 96, 20, 9, True, True),
    ("lstm", 64, 96, 20, 9, True, False),      # v2 off: b_h ignored
    ("lstm", 512, 512, 6, 3, False, True),     # C4 layer shape
    ("gru", 7, 19, 11, 3, True, True),         # ragged sizes (H not a multiple of 4)
    ("lstm", 5, 33, 9, 2, True, True),
]


@pytest.mark.parametrize("cell,in_size,H,T,B,rs,v2", REC_CASES)
def test_recurrent_batch_matches_reference(cuda, product, ref, cell, in_size, H, T, B, rs, v2):
    rng = np.random.default_rng(4)
    x = rng.standard_normal((B, T, in_size)).astype(np.float32)
    want = _ref_batch(ref, cell, in_size, H, T, rs, v2, x, 400)
    r = L.Recurrent(product, cell, in_size, H, T, rs, v2)
    helpers.fill_recurrent(r, 400)
    got = r.apply_batch(x)
    assert_parity(got, want, what=f"{cell} batch")
    assert np.array_equal(r.apply_batch(_to_dev(cuda, x)).cpu().numpy(), got)
    r.close()


@pytest.mark.parametrize("v2", [True, False])
def test_rnn_matches_reference_last_state(cuda, product, ref, v2):
    # the reference's RNNApplyInference is only well defined without return_sequences
    rng = np.random.default_rng(41)
    x = rng.standard_normal((4, 25, 40)).astype(np.float32)
    want = _ref_batch(ref, "rnn", 40, 64, 25, False, v2, x, 401)
    r = L.Recurrent(product, "rnn", 40, 64, 25, False, v2)
    helpers.fill_recurrent(r, 401)
    assert_parity(r.apply_batch(x), want, what="rnn")
    # return_sequences: row t must equal the final state of a T=t+1 run (the reference's evident intent)
    rs = L.Recurrent(product, "rnn", 40, 64, 25, True, v2)
    helpers.copy_weights(rs, r)
    seq = rs.apply_batch(x)
    assert_parity(seq[:, -1], want, what="rnn sequences last row")
    r.close(); rs.close()


@pytest.mark.parametrize("cell", ["gru", "lstm"])
def test_single_sequence_calls_are_stateful_like_the_reference(cuda, product, ref, cell):
    in_size, H, T = 24, 48, 16
    rng = np.random.default_rng(42)
    x = rng.standard_normal((2 * T, in_size)).astype(np.float32)
    r_ref = L.Recurrent(ref, cell, in_size, H, T)
    r_new = L.Recurrent(product, cell, in_size, H, T)
    helpers.fill_recurrent(r_ref, 402); helpers.copy_weights(r_new, r_ref)
    want = np.concatenate([r_ref.apply(x[:T]), r_ref.apply(x[T:])])     # second call continues from h (and c)
    got = np.concatenate([r_new.apply(x[:T]), r_new.apply(x[T:])])
    assert_parity(got, want, what=f"{cell} stateful")
    # two half calls == one call with 2T steps on a fresh handle
    full = L.Recurrent(product, cell, in_size, H, 2 * T)
    helpers.copy_weights(full, r_ref)
    assert_parity(full.apply(x), got, what=f"{cell} chunked == whole")
    # NNTKResetState brings back the fresh-handle result
    r_new.reset_state()
    assert_parity(r_new.apply(x[:T]), want[:T], what=f"{cell} reset")
    for r in (r_ref, r_new, full):
        r.close()


def test_batched_streaming_state_api(cuda, product):
    in_size, H, T, B = 16, 32, 10, 6
    rng = np.random.default_rng(43)
    x = rng.standard_normal((B, 2 * T, in_size)).astype(np.float32)
    whole = L.Recurrent(product, "lstm", in_size, H, 2 * T)
    helpers.fill_recurrent(whole, 403)
    want = whole.apply_batch(x)
    half = L.Recurrent(product, "lstm", in_size, H, T)
    helpers.copy_weights(half, whole)
    h = np.zeros((B, H), np.float32); c = np.zeros((B, H), np.float32)
    a = half.apply_batch(np.ascontiguousarray(x[:, :T]), h_state=h, c_state=c)
    b = half.apply_batch(np.ascontiguousarray(x[:, T:]), h_state=h, c_state=c)
    assert_parity(np.concatenate([a, b], 1), want, what="streaming state")
    assert_parity(h, want[:, -1], what="final h")
    whole.close(); half.close()


def test_custom_activation_in_device_layer_is_refused(cuda, product):
    import ctypes as C
    cb = C.CFUNCTYPE(None, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int)(lambda *a: None)
    act = product.ActivationFunctionCreate(12, None, None, C.cast(cb, C.c_void_p), None, None)
    cfg = product.DenseConfigCreate(8, 12, act)
    d = product.DenseCreateForInference(cfg)
    x = np.zeros(8, np.float32); y = np.zeros(12, np.float32)
    assert product.DenseApplyInference(d, x.ctypes.data, y.ctypes.data) == -2   # NNTK_ERR_UNSUPPORTED, no CPU fallback
    product.DenseDestroy(d)
