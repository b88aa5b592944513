"""The golden cases: each entry builds seeded inputs + weights, runs them through the C API of the library it is given
(reference when generating, product when checking on the GPU) and returns the output.  The oracle port is checked against
the same stored outputs through tests/test_oracle.py."""
import numpy as np

import helpers
from nntoolkitcore_b200 import layers as L


def _spec(nfft, win, nov, n, window, fs=None, norm=1.0, clip=0):
    def run(lib):
        s = L.Spectrogram(lib, nfft, win, nov, n, norm, window, fs)
        x = helpers.audio_clip(clip, n)
        out = s.apply(x)
        s.close()
        return out
    return run


def _rec(cell, in_size, H, T, rs, v2, seed):
    def run(lib):
        r = L.Recurrent(lib, cell, in_size, H, T, rs, v2)
        helpers.fill_recurrent(r, seed)
        x = np.random.Generator(np.random.PCG64(seed + 1)).standard_normal((T, in_size)).astype(np.float32)
        out = r.apply(x)
        r.close()
        return out
    return run


def _conv(cin, cout, k, s, T, seed):
    def run(lib):
        c = L.Conv1d(lib, cin, cout, k, s, T)
        helpers.fill_conv(c, seed)
        x = np.random.Generator(np.random.PCG64(seed + 1)).standard_normal((T, cin)).astype(np.float32)
        out = c.apply(x)
        c.close()
        return out
    return run


def _bn(C_, count, seed):
    def run(lib):
        b = L.BatchNorm(lib, C_, 1e-3, count)
        helpers.fill_bn(b, seed)
        x = np.random.Generator(np.random.PCG64(seed + 1)).standard_normal((count, C_)).astype(np.float32)
        out = b.apply(x)
        b.close()
        return out
    return run


def _dense(i, o, act, ts, seed):
    def run(lib):
        d = L.Dense(lib, i, o, act, ts=ts)
        helpers.fill_dense(d, seed)
        shape = (i,) if ts is None else (ts, i)
        x = np.random.Generator(np.random.PCG64(seed + 1)).standard_normal(shape).astype(np.float32)
        out = d.apply(x)
        d.close()
        return out
    return run


def _logmel(nfft, win, nov, n, window, n_mels, sr, lo, hi, fs=None, clip=0):
    def run(lib):
        s = L.Spectrogram(lib, nfft, win, nov, n, 1.0, window, fs)
        f = L.LogMelSpectrogram(lib, s, n_mels, sr, lo, hi)
        out = f.apply(helpers.audio_clip(clip, n))
        f.close(); s.close()
        return out
    return run


def _mel(n_mels, n_fft, sr, lo, hi, rows, seed):
    def run(lib):
        m = L.MelFilterBank(lib, n_mels, n_fft, sr, lo, hi)
        x = np.abs(np.random.Generator(np.random.PCG64(seed)).standard_normal((rows, n_fft // 2 + 1))).astype(np.float32)
        out = m.apply(x)
        m.close()
        return out
    return run


CASES = {
    "spec_c2_1s_hann": _spec(512, 512, 352, 16000, "hann_window"),
    "spec_c5_1s_hann400": _spec(512, 400, 240, 16000, "hann_window", clip=1),
    "spec_psd_hamming": _spec(512, 512, 352, 8000, "hamming_window", fs=16000, clip=2),
    "spec_nfft256": _spec(256, 256, 128, 4096, "hann_window", clip=3),
    "spec_nfft100_norm": _spec(100, 100, 50, 2000, "blackman_window", norm=0.5, clip=4),
    "gru_c1": _rec("gru", 40, 128, 100, True, True, 100),          # BASELINE configs[0]
    "gru_last_only": _rec("gru", 40, 128, 30, False, True, 110),
    "lstm_v2": _rec("lstm", 64, 96, 20, True, True, 120),
    "lstm_v1": _rec("lstm", 64, 96, 20, True, False, 130),
    "rnn_v2_last": _rec("rnn", 40, 64, 25, False, True, 140),
    "conv_128_256_k5": _conv(128, 256, 5, 1, 40, 150),
    "conv_stride2": _conv(16, 24, 3, 2, 41, 160),
    "bn_257": _bn(257, 20, 170),
    "dense_relu": _dense(256, 12, L.ACT_RELU, None, 180),
    "tdd_softmax": _dense(256, 12, L.ACT_SOFTMAX, 10, 190),
}

def _dft(nfft, forward, batch, seed):
    def run(lib):
        rng = np.random.Generator(np.random.PCG64(seed))
        re = rng.standard_normal((batch, nfft)).astype(np.float32)
        im = rng.standard_normal((batch, nfft)).astype(np.float32)
        d = L.DFT(lib, nfft, forward)
        if getattr(lib, "has_ext", False):
            o_re, o_im = d.perform(re, im)
        else:   # the reference transforms one signal per call
            outs = [d.perform(re[b], im[b]) for b in range(batch)]
            o_re, o_im = np.stack([o[0] for o in outs]), np.stack([o[1] for o in outs])
        d.close()
        return np.stack([o_re, o_im])
    return run


def _bd(which, batch, T, C_, seed, seq=True):
    def run(lib):
        rng = np.random.Generator(np.random.PCG64(seed))
        shape = (batch, T, C_) if seq else (batch, C_)
        a = rng.standard_normal(shape).astype(np.float32)
        b = rng.standard_normal(shape).astype(np.float32)
        if which == "reverse":
            return L.bd_reverse_input(lib, a)
        return (L.bd_merge_concat if which == "concat" else L.bd_merge_sum)(lib, a, b)
    return run


# SURVEY.md section 8(f) "next" rows, stored in tests/golden/golden_v2.npz
NEXT_CASES = {
    "logmel_40_c2": _logmel(512, 512, 352, 16000, "hann_window", 40, 16000, 20.0, 8000.0),
    "logmel_80_win400": _logmel(512, 400, 240, 16000, "hann_window", 80, 16000, 0.0, 8000.0, clip=1),
    "logmel_128_psd": _logmel(512, 512, 352, 8000, "hamming_window", 128, 16000, 0.0, 8000.0, fs=16000, clip=2),
    "logmel_nfft256": _logmel(256, 256, 128, 4096, "hann_window", 20, 16000, 100.0, 7600.0, clip=3),
    "mel_23_rows7": _mel(23, 512, 8000, 100.0, 3800.0, 7, 200),
    "dft_512_fwd": _dft(512, True, 3, 210),
    "dft_512_inv": _dft(512, False, 2, 211),
    "dft_128_fwd": _dft(128, True, 2, 212),
    "dft_100_inv": _dft(100, False, 2, 213),
    "bd_reverse": _bd("reverse", 3, 7, 5, 220),
    "bd_concat_seq": _bd("concat", 2, 6, 8, 221),
    "bd_concat_last": _bd("concat", 4, 1, 6, 222, seq=False),
    "bd_sum": _bd("sum", 2, 5, 7, 223),
}
