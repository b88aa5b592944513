"""Runs the golden cases through the oracle C port (oracle/nntk_oracle.c) instead of a library with the reference API.
The case table mirrors tests/golden_cases.py: same seeds, same shapes."""
import ctypes as C

import numpy as np

import helpers
from helpers import fptr

WIN = {"ones": 0, "hann_window": 1, "hamming_window": 2, "periodic_hann_window": 3, "periodic_hamming_window": 4,
       "blackman_window": 5}
G = {"rnn": 1, "gru": 3, "lstm": 4}


def spectrogram(port, nfft, win, nov, n, window, fs=None, norm=1.0, x=None):
    w = np.zeros(win, np.float32)
    port.oracle_window(WIN[window], fptr(w), win)
    x = np.ascontiguousarray(x, np.float32).reshape(-1, n)
    step = win - nov
    ntime = (n - nov) // step
    out = np.zeros((x.shape[0], ntime, nfft // 2 + 1), np.float32)
    port.oracle_spectrogram(nfft, win, nov, n, C.c_float(norm), fptr(w), 0 if fs is None else 1, int(fs or 0),
                            C.c_float(0.0), fptr(x), fptr(out), x.shape[0])
    return out


class _Shim:
    pass


def recurrent_weights(cell, in_size, H, seed):
    s = _Shim()
    s.in_size, s.H = in_size, H
    g = G[cell]
    s.W = np.zeros((in_size, g * H), np.float32); s.U = np.zeros((H, g * H), np.float32)
    s.b_i = np.zeros(g * H, np.float32); s.b_h = np.zeros(g * H, np.float32)
    helpers.fill_recurrent(s, seed)
    return s


def recurrent(port, cell, in_size, H, T, rs, v2, w, x, h=None, c=None, acts=None):
    x = np.ascontiguousarray(x, np.float32)
    h = np.zeros(H, np.float32) if h is None else h
    c = np.zeros(H, np.float32) if c is None else c
    out = np.zeros((T, H) if rs else (H,), np.float32)
    if cell == "rnn":
        port.oracle_rnn(in_size, H, T, int(rs), int(v2), (acts or (3,))[0], C.c_float(1.0), fptr(w.W), fptr(w.U),
                        fptr(w.b_i), fptr(w.b_h), fptr(x), fptr(h), fptr(out))
    elif cell == "gru":
        z, hh, r = acts or (2, 3, 2)
        port.oracle_gru(in_size, H, T, int(rs), z, hh, r, fptr(w.W), fptr(w.U), fptr(w.b_i), fptr(w.b_h), fptr(x),
                        fptr(h), fptr(out))
    else:
        i, f, g, o, oo = acts or (2, 2, 3, 2, 3)
        port.oracle_lstm(in_size, H, T, int(rs), int(v2), i, f, g, o, oo, fptr(w.W), fptr(w.U), fptr(w.b_i),
                         fptr(w.b_h), fptr(x), fptr(h), fptr(c), fptr(out))
    return out


def conv(port, cin, cout, k, s, T, W, b, x):
    x = np.ascontiguousarray(x, np.float32).reshape(-1, T, cin)
    t_out = (T - (k - s)) // s
    out = np.zeros((x.shape[0], t_out, cout), np.float32)
    port.oracle_conv1d(cin, cout, k, s, T, fptr(np.ascontiguousarray(W)), fptr(b), fptr(x), fptr(out), x.shape[0])
    return out


def batch_norm(port, C_, eps, gamma, beta, mean, var, x):
    x = np.ascontiguousarray(x, np.float32).reshape(-1, C_)
    out = np.zeros_like(x)
    port.oracle_batch_norm(C_, C.c_float(eps), x.shape[0], fptr(gamma), fptr(beta), fptr(mean), fptr(var), fptr(x), fptr(out))
    return out


def dense(port, i, o, W, b, act, relu_a, x, softmax_vector=None):
    x = np.ascontiguousarray(x, np.float32).reshape(-1, i)
    out = np.zeros((x.shape[0], o), np.float32)
    port.oracle_dense(i, o, fptr(np.ascontiguousarray(W)), fptr(b), 0 if act == 5 else act, C.c_float(relu_a), fptr(x),
                      fptr(out), x.shape[0])
    if act == 5:
        v = softmax_vector or o
        res = np.zeros_like(out)
        port.oracle_activation(5, C.c_float(1.0), v, fptr(out), fptr(res), out.size // v)
        out = res
    return out


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def run_case(port, name):
    """the golden case `name` evaluated by the port"""
    if name.startswith("spec_"):
        args = {
            "spec_c2_1s_hann": (512, 512, 352, 16000, "hann_window", None, 1.0, 0),
            "spec_c5_1s_hann400": (512, 400, 240, 16000, "hann_window", None, 1.0, 1),
            "spec_psd_hamming": (512, 512, 352, 8000, "hamming_window", 16000, 1.0, 2),
            "spec_nfft256": (256, 256, 128, 4096, "hann_window", None, 1.0, 3),
            "spec_nfft100_norm": (100, 100, 50, 2000, "blackman_window", None, 0.5, 4),
        }[name]
        nfft, win, nov, n, window, fs, norm, clip = args
        return spectrogram(port, nfft, win, nov, n, window, fs, norm, helpers.audio_clip(clip, n))[0]
    rec = {"gru_c1": ("gru", 40, 128, 100, True, True, 100), "gru_last_only": ("gru", 40, 128, 30, False, True, 110),
           "lstm_v2": ("lstm", 64, 96, 20, True, True, 120), "lstm_v1": ("lstm", 64, 96, 20, True, False, 130),
           "rnn_v2_last": ("rnn", 40, 64, 25, False, True, 140)}
    if name in rec:
        cell, i, H, T, rs, v2, seed = rec[name]
        w = recurrent_weights(cell, i, H, seed)
        x = _rng(seed + 1).standard_normal((T, i)).astype(np.float32)
        return recurrent(port, cell, i, H, T, rs, v2, w, x)
    if name in ("conv_128_256_k5", "conv_stride2"):
        cin, cout, k, s, T, seed = {"conv_128_256_k5": (128, 256, 5, 1, 40, 150), "conv_stride2": (16, 24, 3, 2, 41, 160)}[name]
        sh = _Shim(); sh.W = np.zeros((cout, cin, k), np.float32); sh.b = np.zeros(cout, np.float32)
        helpers.fill_conv(sh, seed)
        x = _rng(seed + 1).standard_normal((T, cin)).astype(np.float32)
        return conv(port, cin, cout, k, s, T, sh.W, sh.b, x)[0]
    if name == "bn_257":
        sh = _Shim(); sh.C = 257
        store = {}
        sh.set = lambda g, b, m, v: store.update(g=g, b=b, m=m, v=v)
        helpers.fill_bn(sh, 170)
        x = _rng(171).standard_normal((20, 257)).astype(np.float32)
        return batch_norm(port, 257, 1e-3, store["g"], store["b"], store["m"], store["v"], x)
    if name in ("dense_relu", "tdd_softmax"):
        act, ts, seed = {"dense_relu": (4, None, 180), "tdd_softmax": (5, 10, 190)}[name]
        sh = _Shim(); sh.in_size, sh.out_size = 256, 12
        sh.W = np.zeros((256, 12), np.float32); sh.b = np.zeros(12, np.float32)
        helpers.fill_dense(sh, seed)
        x = _rng(seed + 1).standard_normal((256,) if ts is None else (ts, 256)).astype(np.float32)
        out = dense(port, 256, 12, sh.W, sh.b, act, 1.0, x)
        return out[0] if ts is None else out
    raise KeyError(name)


# ---- SURVEY.md section 8(f) rows ---------------------------------------------------------------------------------------
def mel_weights(port, n_mels, n_fft, sr, lo, hi):
    w = np.zeros((n_fft // 2 + 1, n_mels), np.float32)
    port.oracle_mel_weights(n_mels, n_fft, sr, C.c_float(lo), C.c_float(hi), fptr(w))
    return w


def mel_apply(port, n_mels, n_fft, sr, lo, hi, spec):
    spec = np.ascontiguousarray(spec, np.float32)
    w = mel_weights(port, n_mels, n_fft, sr, lo, hi)
    out = np.zeros(spec.shape[:-1] + (n_mels,), np.float32)
    port.oracle_mel_apply(n_mels, n_fft // 2 + 1, fptr(w), fptr(spec), fptr(out), int(np.prod(spec.shape[:-1])))
    return out


def log_mel(port, nfft, win, nov, n, window, n_mels, sr, lo, hi, fs=None, x=None):
    w = np.zeros(win, np.float32)
    port.oracle_window(WIN[window], fptr(w), win)
    x = np.ascontiguousarray(x, np.float32).reshape(-1, n)
    ntime = (n - nov) // (win - nov)
    out = np.zeros((x.shape[0], ntime, n_mels), np.float32)
    port.oracle_log_mel_spectrogram(nfft, win, nov, n, C.c_float(1.0), fptr(w), 0 if fs is None else 1, int(fs or 0),
                                    n_mels, sr, C.c_float(lo), C.c_float(hi), fptr(x), fptr(out), x.shape[0])
    return out


def run_next_case(port, name):
    """the golden_v2 case `name` evaluated by the port (same table as golden_cases.NEXT_CASES)"""
    logmel = {"logmel_40_c2": (512, 512, 352, 16000, "hann_window", 40, 16000, 20.0, 8000.0, None, 0),
              "logmel_80_win400": (512, 400, 240, 16000, "hann_window", 80, 16000, 0.0, 8000.0, None, 1),
              "logmel_128_psd": (512, 512, 352, 8000, "hamming_window", 128, 16000, 0.0, 8000.0, 16000, 2),
              "logmel_nfft256": (256, 256, 128, 4096, "hann_window", 20, 16000, 100.0, 7600.0, None, 3)}
    if name in logmel:
        nfft, win, nov, n, window, n_mels, sr, lo, hi, fs, clip = logmel[name]
        return log_mel(port, nfft, win, nov, n, window, n_mels, sr, lo, hi, fs, helpers.audio_clip(clip, n))[0]
    if name == "mel_23_rows7":
        x = np.abs(_rng(200).standard_normal((7, 257))).astype(np.float32)
        return mel_apply(port, 23, 512, 8000, 100.0, 3800.0, x)
    dft = {"dft_512_fwd": (512, True, 3, 210), "dft_512_inv": (512, False, 2, 211), "dft_128_fwd": (128, True, 2, 212),
           "dft_100_inv": (100, False, 2, 213)}
    if name in dft:
        nfft, forward, batch, seed = dft[name]
        rng = _rng(seed)
        re = rng.standard_normal((batch, nfft)).astype(np.float32)
        im = rng.standard_normal((batch, nfft)).astype(np.float32)
        out = np.zeros((2, batch, nfft), np.float32)
        for b in range(batch):
            port.oracle_dft(nfft, int(not forward), fptr(re[b]), fptr(im[b]), fptr(out[0, b]), fptr(out[1, b]))
        return out
    bd = {"bd_reverse": ("reverse", 3, 7, 5, 220, True), "bd_concat_seq": ("concat", 2, 6, 8, 221, True),
          "bd_concat_last": ("concat", 4, 1, 6, 222, False), "bd_sum": ("sum", 2, 5, 7, 223, True)}
    if name in bd:
        which, batch, T, C_, seed, seq = bd[name]
        rng = _rng(seed)
        shape = (batch, T, C_) if seq else (batch, C_)
        a = rng.standard_normal(shape).astype(np.float32)
        b = rng.standard_normal(shape).astype(np.float32)
        if which == "reverse":
            out = np.zeros_like(a)
            port.oracle_bd_reverse(fptr(a), fptr(out), C_, T, batch)
        elif which == "concat":
            out = np.zeros(shape[:-1] + (2 * C_,), np.float32)
            port.oracle_bd_concat(fptr(a), fptr(b), fptr(out), C_, T if seq else 1, batch)
        else:
            out = np.zeros_like(a)
            port.oracle_bd_sum(fptr(a), fptr(b), fptr(out), C.c_longlong(a.size))
        return out
    raise KeyError(name)
