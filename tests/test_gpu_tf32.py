"""GPU tier: the optional tensor-core math mode (NNTK_MATH_TF32: tcgen05 kind::tf32 + TMA + TMEM, csrc/gemm_tc.cu)
against the unmodified reference at the stated 1e-3 tolerance; the default fp32 mode is restored afterwards."""
import numpy as np
import pytest

import helpers
from helpers import assert_parity, TF32_TOL
from nntoolkitcore_b200 import layers as L

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


@pytest.fixture()
def tf32(cuda, product):
    assert product.NNTKSetMathMode(1) == 0
    yield product
    product.NNTKSetMathMode(helpers.DEFAULT_MATH_MODE)


@pytest.mark.parametrize("cin,cout,k,s,T,B", [(128, 256, 5, 1, 300, 3), (64, 96, 3, 2, 257, 2), (256, 256, 5, 1, 130, 2),
                                              (16, 8, 1, 1, 40, 1)])
def test_conv1d_tf32(tf32, ref, cin, cout, k, s, T, B):
    rng = np.random.default_rng(3)
    x = rng.standard_normal((B, T, cin)).astype(np.float32)
    c_ref, c_new = L.Conv1d(ref, cin, cout, k, s, T), L.Conv1d(tf32, cin, cout, k, s, T)
    helpers.fill_conv(c_ref, 300); helpers.copy_weights(c_new, c_ref)
    t_out = c_ref.out_shape[0]
    bn_ref, bn_new = L.BatchNorm(ref, cout, 1e-3, t_out), L.BatchNorm(tf32, cout, 1e-3, t_out)
    helpers.fill_bn(bn_ref, 302); helpers.copy_weights(bn_new, bn_ref)
    r_ref, r_new = L.Activation(ref, L.ACT_RELU, t_out * cout), L.Activation(tf32, L.ACT_RELU, t_out * cout)
    want = np.stack([r_ref.apply(bn_ref.apply(c_ref.apply(v))) for v in x])
    got = c_new.apply_batch(x, bn=bn_new, act=r_new)
    err, l2 = assert_parity(got, want, tol=TF32_TOL, what="conv tf32")
    assert l2 > 1e-7, "suspiciously exact: the tensor-core path did not run"


def test_dense_and_tdd_tf32(tf32, ref):
    rng = np.random.default_rng(9)
    x = rng.standard_normal((3, 200, 256)).astype(np.float32)
    d_ref, d_new = L.Dense(ref, 256, 12, L.ACT_SOFTMAX, ts=200), L.Dense(tf32, 256, 12, L.ACT_SOFTMAX, ts=200)
    helpers.fill_dense(d_ref, 501); helpers.copy_weights(d_new, d_ref)
    want = np.stack([d_ref.apply(v) for v in x])
    assert_parity(d_new.apply_batch(x), want, tol=TF32_TOL, what="tdd tf32")


@pytest.mark.parametrize("cell,in_size,H,T,B", [("gru", 256, 256, 10, 200), ("lstm", 512, 512, 6, 130)])
def test_recurrent_projection_tf32(tf32, ref, cell, in_size, H, T, B):
    """TF32 mode: the hoisted x*W projection runs on the tensor cores with split operands -- 3xbf16, or 3xTF32 where the 2-byte pitch rules do not fit (its single-pass rounding alone would cost
    ~1e-3 over long sequences), the recurrent steps as single-pass TF32 (cluster or per-step kernel, by shape)"""
    rng = np.random.default_rng(4)
    x = rng.standard_normal((B, T, in_size)).astype(np.float32)
    outs = []
    for v in x[:4]:
        r = L.Recurrent(ref, cell, in_size, H, T)
        helpers.fill_recurrent(r, 400)
        outs.append(r.apply(v)); r.close()
    r = L.Recurrent(tf32, cell, in_size, H, T)
    helpers.fill_recurrent(r, 400)
    got = r.apply_batch(x)
    assert_parity(got[:4], np.stack(outs), tol=TF32_TOL, what=f"{cell} tf32 projection")


def test_unaligned_channel_count_is_padded_for_tma(tf32, ref):
    """Cin = 257 (the C5 front-end) rows are not 16-byte aligned: the input is re-laid to a 260-float pitch with zero weights
    for the pad channels and runs on the tensor cores like any other shape"""
    rng = np.random.default_rng(5)
    x = rng.standard_normal((2, 40, 257)).astype(np.float32)
    c_ref, c_new = L.Conv1d(ref, 257, 64, 5, 1, 40), L.Conv1d(tf32, 257, 64, 5, 1, 40)
    helpers.fill_conv(c_ref, 7); helpers.copy_weights(c_new, c_ref)
    err, l2 = assert_parity(c_new.apply_batch(x), np.stack([c_ref.apply(v) for v in x]), tol=TF32_TOL, what="padded Cin")
    assert l2 > 1e-7, "suspiciously exact: the tensor-core path did not run"


@pytest.mark.parametrize("cell,in_size,H,T,B,seq", [("gru", 64, 288, 12, 300, True), ("lstm", 96, 320, 40, 260, True),
                                                    ("rnn", 40, 96, 12, 257, True), ("lstm", 64, 264, 7, 385, False),
                                                    ("gru", 32, 68, 50, 270, False)])
def test_recurrent_step_on_tensor_cores(tf32, port, cell, in_size, H, T, B, seq):
    """batch > NNTK_CLUSTER_MAX_BATCH (256) and H not eligible for the persistent tensor-core cluster kernel (H > 256 or
    H % 32 != 0): every timestep is a tcgen05 GEMM with the cell fused into its epilogue
    (csrc/recurrent_tc.cu); rows from the first, a middle and the last 128-row tile are checked against the oracle port
    (the reference's own RNNApplyInference is undefined for sequences, rnn.c:237-241).  The un-gated tanh RNN is the
    exception: with Glorot U its Jacobian does not contract, so TF32 rounding grows step by step (measured 1.06e-3 at
    T = 4 with the step on tensor cores, 1.2e-3 at T = 12 with only the projection in TF32) -- the library keeps the RNN in
    fp32 in every math mode, which this case checks; the gated cells hold 1e-3 on the tensor cores at T = 500 (below)."""
    import port_cases
    rng = np.random.default_rng(6)
    x = rng.standard_normal((B, T, in_size)).astype(np.float32)
    pick = [0, 1, 127, 128, B // 2, B - 2, B - 1]
    w = port_cases.recurrent_weights(cell, in_size, H, 410)
    outs = [port_cases.recurrent(port, cell, in_size, H, T, seq, True, w, x[i]) for i in pick]
    r = L.Recurrent(tf32, cell, in_size, H, T, return_sequences=seq)
    helpers.fill_recurrent(r, 410)
    got = r.apply_batch(x)
    assert_parity(got[pick], np.stack(outs), tol=TF32_TOL, what=f"{cell} tensor-core step")


def test_lstm_c4_shape_long_sequence_tf32(tf32, port):
    """BASELINE config 4 unit (LSTM in=H=512, T=500) through the tensor-core step path: the TF32 rounding must not
    accumulate over 500 recurrent steps beyond the stated 1e-3"""
    import port_cases
    H, T, B = 512, 500, 260
    rng = np.random.default_rng(8)
    x = rng.standard_normal((B, T, H)).astype(np.float32)
    pick = [0, 129, B - 1]
    w = port_cases.recurrent_weights("lstm", H, H, 400)
    outs = [port_cases.recurrent(port, "lstm", H, H, T, True, True, w, x[i]) for i in pick]
    r = L.Recurrent(tf32, "lstm", H, H, T)
    helpers.fill_recurrent(r, 400)
    got = r.apply_batch(x)
    assert_parity(got[pick], np.stack(outs), tol=TF32_TOL, what="lstm C4 shape, T=500, tensor-core step")


@pytest.mark.parametrize("cell,in_size,H,T,B,seq", [("gru", 64, 256, 12, 300, True), ("lstm", 96, 128, 40, 260, True),
                                                    ("lstm", 64, 192, 7, 385, False), ("gru", 32, 64, 50, 270, False),
                                                    ("gru", 40, 128, 100, 33, True), ("gru", 16, 32, 9, 128, True),
                                                    ("gru", 256, 256, 300, 129, True), ("gru", 48, 512, 20, 150, True),
                                                    ("lstm", 64, 512, 12, 70, False), ("lstm", 32, 384, 9, 200, True)])
def test_persistent_tensor_core_cluster_kernel(tf32, port, cell, in_size, H, T, B, seq):
    """TF32 mode, H % 32 == 0, H <= 256, batch >= 32: ONE launch runs all T steps (csrc/recurrent_cluster_tc.cu) -- a cluster
    of H/32 CTAs per 128 sequences keeps U resident in shared memory and exchanges h_t by DSMEM bulk copies.  Cluster sizes
    1, 2, 4, 6, 8, 12, 16 (H = 512: 64-sequence slices, LSTM one slice per cluster, GRU two); ragged last slice; rows from the first, a middle and the last slice against the oracle port; the C5
    GRU(256) shape over 300 steps checks that TF32 rounding of h does not accumulate beyond the stated 1e-3."""
    import port_cases
    rng = np.random.default_rng(16)
    x = rng.standard_normal((B, T, in_size)).astype(np.float32)
    pick = sorted({0, 1, min(127, B - 1), min(128, B - 1), B // 2, B - 2, B - 1})
    w = port_cases.recurrent_weights(cell, in_size, H, 420)
    outs = [port_cases.recurrent(port, cell, in_size, H, T, seq, True, w, x[i]) for i in pick]
    r = L.Recurrent(tf32, cell, in_size, H, T, return_sequences=seq)
    helpers.fill_recurrent(r, 420)
    got = r.apply_batch(x)
    err, l2 = assert_parity(got[pick], np.stack(outs), tol=TF32_TOL, what=f"{cell} tensor-core cluster kernel")
    assert l2 > 1e-7, "suspiciously exact: the tensor-core path did not run"


def test_cluster_kernel_streaming_state_matches_one_shot(tf32):
    """carried state through the persistent kernel: two half-length calls with caller-owned h / c == one full-length call"""
    rng = np.random.default_rng(17)
    B, T, H = 140, 16, 128
    x = rng.standard_normal((B, 2 * T, 64)).astype(np.float32)
    full = L.Recurrent(tf32, "lstm", 64, H, 2 * T); helpers.fill_recurrent(full, 430)
    half = L.Recurrent(tf32, "lstm", 64, H, T); helpers.copy_weights(half, full)
    want = full.apply_batch(x)
    h = np.zeros((B, H), np.float32); c = np.zeros((B, H), np.float32)
    a = half.apply_batch(np.ascontiguousarray(x[:, :T]), h_state=h, c_state=c)
    b = half.apply_batch(np.ascontiguousarray(x[:, T:]), h_state=h, c_state=c)
    assert_parity(np.concatenate([a, b], 1), want, tol=1e-6, what="cluster kernel streaming state")


# ---- NNTK_MATH_TF32X3: tcgen05 with hi/lo operand split (three passes in one K loop).  Operand error drops to ~2^-22; the
# tensor core's truncating fp32 accumulate (a shrink linear in K) is compensated in the epilogue: the mode holds the fp32
# contract, 1e-5.
X3_TOL = helpers.FP32_TOL
@pytest.fixture()
def x3(cuda, product):
    assert product.NNTKSetMathMode(2) == 0
    yield product
    product.NNTKSetMathMode(helpers.DEFAULT_MATH_MODE)


@pytest.mark.parametrize("cin,cout,k,s,T,B", [(128, 256, 5, 1, 300, 3), (64, 96, 3, 2, 257, 2), (256, 256, 5, 1, 130, 2),
                                              (16, 8, 1, 1, 40, 1), (20, 40, 3, 1, 50, 2), (257, 64, 5, 1, 60, 2),
                                              (512, 256, 5, 1, 40, 2), (1024, 128, 7, 1, 30, 1)])   # K = 2 560 and 7 168: above the 3xTF32 cap (1 536), FFMA
def test_conv1d_3xtf32(x3, ref, cin, cout, k, s, T, B):
    rng = np.random.default_rng(3)
    x = rng.standard_normal((B, T, cin)).astype(np.float32)
    c_ref, c_new = L.Conv1d(ref, cin, cout, k, s, T), L.Conv1d(x3, cin, cout, k, s, T)
    helpers.fill_conv(c_ref, 300); helpers.copy_weights(c_new, c_ref)
    t_out = c_ref.out_shape[0]
    bn_ref, bn_new = L.BatchNorm(ref, cout, 1e-3, t_out), L.BatchNorm(x3, cout, 1e-3, t_out)
    helpers.fill_bn(bn_ref, 302); helpers.copy_weights(bn_new, bn_ref)
    r_ref, r_new = L.Activation(ref, L.ACT_RELU, t_out * cout), L.Activation(x3, L.ACT_RELU, t_out * cout)
    want = np.stack([r_ref.apply(bn_ref.apply(c_ref.apply(v))) for v in x])
    got = c_new.apply_batch(x, bn=bn_new, act=r_new)
    assert_parity(got, want, tol=X3_TOL, what="conv 3xTF32")


def test_dense_tdd_and_recurrent_3xtf32(x3, ref):
    rng = np.random.default_rng(9)
    x = rng.standard_normal((3, 200, 256)).astype(np.float32)
    d_ref, d_new = L.Dense(ref, 256, 12, L.ACT_SOFTMAX, ts=200), L.Dense(x3, 256, 12, L.ACT_SOFTMAX, ts=200)
    helpers.fill_dense(d_ref, 501); helpers.copy_weights(d_new, d_ref)
    assert_parity(d_new.apply_batch(x), np.stack([d_ref.apply(v) for v in x]), tol=X3_TOL, what="tdd 3xTF32")
    # recurrent: projection 3xTF32; at this batch (<= 256 per chunk rule aside) the steps run on FFMA kernels
    xs = rng.standard_normal((300, 20, 64)).astype(np.float32)
    outs = []
    for v in xs[:3]:
        r = L.Recurrent(ref, "gru", 64, 128, 20); helpers.fill_recurrent(r, 440)
        outs.append(r.apply(v)); r.close()
    g = L.Recurrent(x3, "gru", 64, 128, 20); helpers.fill_recurrent(g, 440)
    assert_parity(g.apply_batch(xs)[:3], np.stack(outs), tol=X3_TOL, what="gru 3xTF32 projection + fp32 steps")


@pytest.mark.parametrize("cell,in_size,H,T,B,rs", [("gru", 64, 128, 40, 300, True), ("lstm", 96, 256, 120, 260, True),
                                                   ("lstm", 512, 512, 200, 257, True), ("gru", 40, 264, 60, 290, False),
                                                   ("lstm", 32, 68, 300, 270, True)])
def test_recurrent_steps_3xtf32_hold_the_fp32_contract(x3, ref, cell, in_size, H, T, B, rs):
    """TF32X3 mode, batch > 256: projection AND per-step GEMM run on the tensor cores with hi/lo split operands
    (csrc/recurrent_tc.cu, X3), gate math with expf / tanhf -- checked at the fp32 tolerance, 1e-5"""
    rng = np.random.default_rng(T)
    xs = rng.standard_normal((B, T, in_size)).astype(np.float32)
    g = L.Recurrent(x3, cell, in_size, H, T, rs); helpers.fill_recurrent(g, 470)
    got = g.apply_batch(xs)
    for b in (0, B // 2, B - 1):
        r = L.Recurrent(ref, cell, in_size, H, T, rs); helpers.copy_weights(r, g)
        assert_parity(got[b], r.apply(xs[b]), tol=X3_TOL, what=f"{cell} 3xTF32 steps, sequence {b}")
        r.close()
    g.close()
