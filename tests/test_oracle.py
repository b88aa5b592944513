"""CPU tier: pins the oracle C port (oracle/nntk_oracle.c) against the golden vectors produced by the unmodified
reference (tests/golden/golden_v1.npz, made by tests/golden/make_golden.py), and -- when the reference build is
available -- against the reference itself on further seeded cases."""
from pathlib import Path

import numpy as np
import pytest

import helpers
import port_cases
from helpers import assert_parity

GOLDEN = np.load(Path(__file__).parent / "golden" / "golden_v1.npz")
BIT_EXACT = {"bn_257", "conv_128_256_k5", "conv_stride2"} | {n for n in GOLDEN.files if n.startswith("spec_")}


@pytest.mark.parametrize("name", sorted(GOLDEN.files))
def test_port_matches_golden(port, name):
    got = port_cases.run_case(port, name)
    want = GOLDEN[name]
    if name in BIT_EXACT:
        # same algorithm, same operation order as the reference (kissfft butterflies, scalar conv dots, BN op order)
        assert np.array_equal(got, want), f"{name}: port is not bit-exact, max diff {np.abs(got - want).max():.3e}"
    else:
        # GEMV summation order differs from Eigen's vectorised order: tolerance, not bit-equality
        assert_parity(got, want, what=name)


GOLDEN_NEXT = np.load(Path(__file__).parent / "golden" / "golden_v2.npz")


@pytest.mark.parametrize("name", sorted(GOLDEN_NEXT.files))
def test_port_matches_golden_next_rows(port, name):
    """SURVEY.md section 8(f) rows.  DFT (same kissfft butterflies) and the bidirectional copies are bit-exact; the mel
    products sum in a different order than Eigen's -> tolerance"""
    got = port_cases.run_next_case(port, name)
    if name.startswith(("dft_", "bd_")):
        assert np.array_equal(got, GOLDEN_NEXT[name]), name
    else:
        assert_parity(got, GOLDEN_NEXT[name], what=name)


def test_golden_covers_every_case():
    import golden_cases
    assert sorted(GOLDEN.files) == sorted(golden_cases.CASES)
    assert sorted(GOLDEN_NEXT.files) == sorted(golden_cases.NEXT_CASES)


@pytest.mark.parametrize("args", [(40, 512, 16000, 20.0, 8000.0), (80, 512, 16000, 0.0, 8000.0), (128, 512, 16000, 0.0, 8000.0),
                                  (23, 512, 8000, 100.0, 3800.0), (64, 1024, 22050, 50.0, 11025.0), (10, 100, 16000, 300.0, 7000.0)])
def test_mel_tables_bit_exact(port, ref, product, args):
    """the mel table of the port AND of the product's host set-up code equal the reference's bit for bit
    (reference table read back through MelFilterBankApply on an identity matrix)"""
    from nntoolkitcore_b200 import capi, layers as L
    import ctypes as C
    n_mels, n_fft = args[0], args[1]
    nb = n_fft // 2 + 1
    m = L.MelFilterBank(ref, *args)
    want = m.apply(np.eye(nb, dtype=np.float32))
    m.close()
    assert np.array_equal(port_cases.mel_weights(port, *args), want)

    class Head(C.Structure):   # leading fields of the product's private struct (host/internal.h)
        _fields_ = [("cfg", capi.MelFilterBankConfig), ("weights", C.POINTER(C.c_float))]
    pm = L.MelFilterBank(product, *args)
    got = np.ctypeslib.as_array(C.cast(pm.handle, C.POINTER(Head)).contents.weights, shape=(nb, n_mels)).copy()
    pm.close()
    assert np.array_equal(got, want)


def test_port_windows_bit_exact_vs_reference(port, ref):
    for kind, name in enumerate(["ones", "hann_window", "hamming_window", "periodic_hann_window",
                                 "periodic_hamming_window", "blackman_window"]):
        for size in (1, 2, 400, 512, 1023):
            a = np.zeros(size, np.float32); b = np.zeros(size, np.float32)
            port.oracle_window(kind, helpers.fptr(a), size)
            getattr(ref, name)(helpers.fptr(b), size)
            if size == 1 and "periodic" not in name and name not in ("ones", "blackman_window"):
                continue   # symmetric windows divide by size-1 == 0
            assert np.array_equal(a, b), (name, size)


@pytest.mark.parametrize("nfft", [8, 60, 100, 128, 512, 1000])
def test_port_dft_matches_numpy(port, nfft):
    rng = np.random.default_rng(nfft)
    re, im = rng.standard_normal(nfft).astype(np.float32), rng.standard_normal(nfft).astype(np.float32)
    ore, oim = np.zeros(nfft, np.float32), np.zeros(nfft, np.float32)
    port.oracle_dft(nfft, 0, helpers.fptr(re), helpers.fptr(im), helpers.fptr(ore), helpers.fptr(oim))
    want = np.fft.fft(re.astype(np.float64) + 1j * im)
    assert np.abs((ore + 1j * oim) - want).max() <= 2e-6 * np.abs(want).max()


def test_port_gru_matches_torch(port):
    """independent cross-check: gates z,r,h <-> torch r,z,n, weights transposed"""
    import torch
    i, H, T = 40, 128, 50
    w = port_cases.recurrent_weights("gru", i, H, 7)
    x = np.random.default_rng(8).standard_normal((T, i)).astype(np.float32)
    got = port_cases.recurrent(port, "gru", i, H, T, True, True, w, x)
    gru = torch.nn.GRU(i, H, batch_first=True)
    perm = lambda m: np.concatenate([m[..., H:2 * H], m[..., :H], m[..., 2 * H:]], -1)
    with torch.no_grad():
        gru.weight_ih_l0.copy_(torch.from_numpy(perm(w.W).T.copy()))
        gru.weight_hh_l0.copy_(torch.from_numpy(perm(w.U).T.copy()))
        gru.bias_ih_l0.copy_(torch.from_numpy(perm(w.b_i)))
        gru.bias_hh_l0.copy_(torch.from_numpy(perm(w.b_h)))
        want = gru(torch.from_numpy(x)[None])[0][0].numpy()
    assert_parity(got, want, what="gru vs torch")


def test_port_lstm_matches_torch(port):
    import torch
    i, H, T = 32, 64, 40
    w = port_cases.recurrent_weights("lstm", i, H, 9)
    x = np.random.default_rng(10).standard_normal((T, i)).astype(np.float32)
    got = port_cases.recurrent(port, "lstm", i, H, T, True, True, w, x)
    lstm = torch.nn.LSTM(i, H, batch_first=True)
    with torch.no_grad():
        lstm.weight_ih_l0.copy_(torch.from_numpy(w.W.T.copy())); lstm.weight_hh_l0.copy_(torch.from_numpy(w.U.T.copy()))
        lstm.bias_ih_l0.copy_(torch.from_numpy(w.b_i)); lstm.bias_hh_l0.copy_(torch.from_numpy(w.b_h))
        want = lstm(torch.from_numpy(x)[None])[0][0].numpy()
    assert_parity(got, want, what="lstm vs torch")


def test_port_conv_matches_torch(port):
    import torch
    cin, cout, k, s, T = 16, 24, 3, 2, 41
    rng = np.random.default_rng(11)
    W = rng.standard_normal((cout, cin, k)).astype(np.float32); b = rng.standard_normal(cout).astype(np.float32)
    x = rng.standard_normal((2, T, cin)).astype(np.float32)
    got = port_cases.conv(port, cin, cout, k, s, T, W, b, x)
    want = torch.nn.functional.conv1d(torch.from_numpy(x).transpose(1, 2), torch.from_numpy(W), torch.from_numpy(b), stride=s)
    assert_parity(got, want.transpose(1, 2).numpy(), what="conv vs torch")
