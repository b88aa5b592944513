"""Thin Python mirror of the C create / get-weights / apply-inference API, one class per reference handle type.

Every class takes the bound library as its first argument, so the same code drives the product library and -- in tests
and the bench's CPU baseline only -- the unmodified reference.  Buffers may be numpy arrays (host path: the C call
stages through device memory and is complete on return) or torch CUDA tensors (device path: enqueued on the library
stream).  No arithmetic happens in this file.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi

ACT_NONE, ACT_IDENTITY, ACT_SIGMOID, ACT_TANH, ACT_RELU, ACT_SOFTMAX = 0, 1, 2, 3, 4, 5


class NNTKError(RuntimeError):
    pass


def _ptr(x):
    if x is None:
        return None
    if isinstance(x, np.ndarray):
        assert x.dtype == np.float32 and x.flags["C_CONTIGUOUS"], "float32 C-contiguous arrays only"
        return x.ctypes.data
    if hasattr(x, "data_ptr"):
        assert x.is_contiguous() and str(x.dtype) == "torch.float32", "float32 contiguous tensors only"
        return x.data_ptr()
    return int(x)


def _view(ptr, n):
    return np.ctypeslib.as_array(ptr, shape=(n,)) if n else np.zeros(0, np.float32)


def _check(lib, rc, who):
    if rc != 0:
        msg = lib.NNTKLastError().decode() if getattr(lib, "has_ext", False) else ""
        raise NNTKError(f"{who} returned {rc}: {msg}")


def _new_out(x, shape):
    if isinstance(x, np.ndarray):
        return np.empty(shape, np.float32)
    import torch
    return torch.empty(shape, dtype=torch.float32, device=x.device)


def make_activation(lib, kind, size, a=1.0, vector_size=1):
    if kind in (None, ACT_NONE):
        return None
    if kind == ACT_IDENTITY:
        return lib.ActivationFunctionCreateIdentity(size)
    if kind == ACT_SIGMOID:
        return lib.ActivationFunctionCreateSigmoid(size)
    if kind == ACT_TANH:
        return lib.ActivationFunctionCreateTanh(size)
    if kind == ACT_RELU:
        return lib.ActivationFunctionCreateReLU(size, a)
    if kind == ACT_SOFTMAX:
        return lib.ActivationFunctionCreateSoftmax(size, vector_size)
    raise ValueError(kind)


class Activation:
    def __init__(self, lib, kind, size, a=1.0, vector_size=1):
        self.lib, self.kind, self.size, self.vector_size = lib, kind, size, vector_size
        self.handle = make_activation(lib, kind, size, a, vector_size)

    def apply(self, x, out=None, batch=None):
        out = _new_out(x, x.shape) if out is None else out
        if batch is None:
            self.lib.ActivationFunctionApply(self.handle, _ptr(x), _ptr(out))
        else:
            _check(self.lib, self.lib.ActivationFunctionApplyBatch(self.handle, _ptr(x), _ptr(out), batch), "ActivationFunctionApplyBatch")
        return out

    def close(self):
        if self.handle:
            self.lib.ActivationFunctionDestroy(self.handle)
            self.handle = None


class Spectrogram:
    WINDOWS = ("ones", "hann_window", "hamming_window", "periodic_hann_window", "periodic_hamming_window",
               "blackman_window")

    def __init__(self, lib, nfft, window_size, noverlap, input_size, norm=1.0, window="hann_window", psd_fs=None):
        self.lib = lib
        self.config = lib.SpectrogramConfigCreate(nfft, window_size, noverlap, input_size, norm)
        self.handle = (lib.SpectrogramCreateMagnitude(self.config) if psd_fs is None
                       else lib.SpectrogramCreatePSD(self.config, int(psd_fs)))
        if not self.handle:
            raise NNTKError("SpectrogramCreate failed")
        if window is not None:
            self.set_window(window)

    def set_window(self, window):
        if isinstance(window, str):
            fn = C.cast(getattr(self.lib, window), C.c_void_p)
        else:   # a Python callable (float*, int): kept alive on self
            self._cb = capi.window_fn(window)
            fn = C.cast(self._cb, C.c_void_p)
        self.lib.SpectrogramSetWindowFunc(self.handle, fn)

    @property
    def out_shape(self):
        return (self.config.ntime_series, self.config.nfreq)

    def apply(self, x, out=None):
        out = _new_out(x, self.out_shape) if out is None else out
        self.lib.SpectrogramApply(self.handle, _ptr(x), _ptr(out))
        return out

    def apply_batch(self, x, out=None):
        n = x.shape[0]
        out = _new_out(x, (n,) + self.out_shape) if out is None else out
        _check(self.lib, self.lib.SpectrogramApplyBatch(self.handle, _ptr(x), _ptr(out), n), "SpectrogramApplyBatch")
        return out

    def close(self):
        if self.handle:
            self.lib.SpectrogramDestroy(self.handle)
            self.handle = None


class MelFilterBank:
    """signal/mel_filterbank.h: [rows, n_fft/2+1] spectra -> [rows, n_mels]"""

    def __init__(self, lib, n_mels, n_fft, sample_rate, lower_hz, upper_hz):
        self.lib, self.n_mels, self.n_bins = lib, n_mels, n_fft // 2 + 1
        self.config = lib.MelFilterBankConfigCreate(n_mels, n_fft, sample_rate, lower_hz, upper_hz)
        self.handle = lib.MelFilterBankCreate(self.config)
        if not self.handle:
            raise NNTKError("MelFilterBankCreate failed")

    def apply(self, spec, out=None):
        rows = int(np.prod(spec.shape[:-1]))
        out = _new_out(spec, tuple(spec.shape[:-1]) + (self.n_mels,)) if out is None else out
        self.lib.MelFilterBankApply(self.handle, _ptr(spec), _ptr(out), rows)
        return out

    def close(self):
        if self.handle:
            self.lib.MelFilterBankDestroy(self.handle)
            self.handle = None


class LogMelSpectrogram:
    """signal/log_mel_spectrogram.h on top of a `Spectrogram` (borrowed: close this object first)"""

    def __init__(self, lib, spectrogram, n_mels, sample_rate, lower_hz, upper_hz, n_fft=None):
        self.lib, self.spectrogram, self.n_mels = lib, spectrogram, n_mels
        self.mel_config = lib.MelFilterBankConfigCreate(n_mels, n_fft or spectrogram.config.nfft, sample_rate, lower_hz, upper_hz)
        self.handle = lib.LogMelSpectrogramCreate(spectrogram.handle, self.mel_config)
        if not self.handle:
            raise NNTKError("LogMelSpectrogramCreate failed")

    @property
    def out_shape(self):
        return (self.spectrogram.config.ntime_series, self.n_mels)

    def apply(self, x, out=None):
        out = _new_out(x, self.out_shape) if out is None else out
        self.lib.LogMelSpectrogramApply(self.handle, _ptr(x), _ptr(out))
        return out

    def apply_batch(self, x, out=None):
        n = x.shape[0]
        out = _new_out(x, (n,) + self.out_shape) if out is None else out
        _check(self.lib, self.lib.LogMelSpectrogramApplyBatch(self.handle, _ptr(x), _ptr(out), n), "LogMelSpectrogramApplyBatch")
        return out

    def close(self):
        if self.handle:
            self.lib.LogMelSpectrogramDestroy(self.handle)
            self.handle = None


class BatchNorm:
    def __init__(self, lib, channels, eps, count):
        self.lib, self.C, self.count = lib, channels, count
        self.handle = lib.BatchNormCreateForInference(lib.BatchNormConfigCreate(channels, eps, count))
        if not self.handle:
            raise NNTKError("BatchNormCreateForInference failed")
        w = lib.BatchNormGetWeights(self.handle).contents
        self.gamma, self.beta = _view(w.gamma, channels), _view(w.beta, channels)
        self.mean, self.var = _view(w.moving_mean, channels), _view(w.moving_variance, channels)

    def set(self, gamma, beta, mean, var):
        self.gamma[:], self.beta[:], self.mean[:], self.var[:] = gamma, beta, mean, var

    def apply(self, x, out=None):
        out = _new_out(x, x.shape) if out is None else out
        _check(self.lib, self.lib.BatchNormApplyInference(self.handle, _ptr(x), _ptr(out)), "BatchNormApplyInference")
        return out

    def apply_batch(self, x, out=None):
        out = _new_out(x, x.shape) if out is None else out
        _check(self.lib, self.lib.BatchNormApplyInferenceBatch(self.handle, _ptr(x), _ptr(out), x.shape[0]), "BatchNormApplyInferenceBatch")
        return out

    def close(self):
        if self.handle:
            self.lib.BatchNormDestroy(self.handle)
            self.handle = None


class Conv1d:
    def __init__(self, lib, cin, cout, k, stride, t_in):
        self.lib = lib
        self.config = lib.Conv1dConfigCreate(cin, cout, k, stride, t_in)
        self.handle = lib.Conv1dCreateForInference(self.config)
        if not self.handle:
            raise NNTKError("Conv1dCreateForInference failed")
        w = lib.Conv1dGetWeights(self.handle).contents
        self.W = _view(w.W, cout * cin * k).reshape(cout, cin, k)
        self.b = _view(w.b, cout)

    @property
    def out_shape(self):
        return (self.config.output_size, self.config.output_feature_channels)

    def apply(self, x, out=None):
        out = _new_out(x, self.out_shape) if out is None else out
        _check(self.lib, self.lib.Conv1dApplyInference(self.handle, _ptr(x), _ptr(out)), "Conv1dApplyInference")
        return out

    def apply_batch(self, x, out=None, bn=None, act=None):
        n = x.shape[0]
        out = _new_out(x, (n,) + self.out_shape) if out is None else out
        if bn is None and act is None:
            rc = self.lib.Conv1dApplyInferenceBatch(self.handle, _ptr(x), _ptr(out), n)
        else:
            rc = self.lib.Conv1dBNActApplyInferenceBatch(self.handle, bn.handle if bn else None,
                                                         act.handle if act else None, _ptr(x), _ptr(out), n)
        _check(self.lib, rc, "Conv1dApplyInferenceBatch")
        return out

    def close(self):
        if self.handle:
            self.lib.Conv1dDestroy(self.handle)
            self.handle = None


class Dense:
    """Dense (ts=None) or TimeDistributedDense (ts=int)."""

    def __init__(self, lib, in_size, out_size, act=ACT_NONE, relu_a=1.0, softmax_vector=None, ts=None):
        self.lib, self.in_size, self.out_size, self.ts = lib, in_size, out_size, ts
        if act == ACT_SOFTMAX:
            v = softmax_vector or out_size
            self.act = make_activation(lib, act, out_size // v, vector_size=v)
        else:
            self.act = make_activation(lib, act, out_size, relu_a)
        dcfg = lib.DenseConfigCreate(in_size, out_size, self.act)
        if ts is None:
            self.handle = lib.DenseCreateForInference(dcfg)
            w = lib.DenseGetWeights(self.handle).contents if self.handle else None
        else:
            self.handle = lib.TimeDistributedDenseCreateForInference(lib.TimeDistributedDenseConfigCreate(ts, dcfg))
            w = lib.TimeDistributedDenseGetWeights(self.handle).contents if self.handle else None
        if not self.handle:
            raise NNTKError("DenseCreateForInference failed")
        self.W = _view(w.W, in_size * out_size).reshape(in_size, out_size)
        self.b = _view(w.b, out_size)

    def _shape(self):
        return (self.out_size,) if self.ts is None else (self.ts, self.out_size)

    def apply(self, x, out=None):
        out = _new_out(x, self._shape()) if out is None else out
        fn = self.lib.DenseApplyInference if self.ts is None else self.lib.TimeDistributedDenseApplyInference
        _check(self.lib, fn(self.handle, _ptr(x), _ptr(out)), "DenseApplyInference")
        return out

    def apply_batch(self, x, out=None):
        n = x.shape[0]
        out = _new_out(x, (n,) + self._shape()) if out is None else out
        fn = self.lib.DenseApplyInferenceBatch if self.ts is None else self.lib.TimeDistributedDenseApplyInferenceBatch
        _check(self.lib, fn(self.handle, _ptr(x), _ptr(out), n), "DenseApplyInferenceBatch")
        return out

    def close(self):
        if self.handle:
            (self.lib.DenseDestroy if self.ts is None else self.lib.TimeDistributedDenseDestroy)(self.handle)
            self.handle = None
        if self.act:
            self.lib.ActivationFunctionDestroy(self.act)
            self.act = None


class Recurrent:
    """RNN / GRU / LSTM.  cell in {"rnn", "gru", "lstm"}.  acts: optional tuple of ACT_* kinds in the C argument order of
    the cell's XActivationsCreate (GRU: z, h, r; LSTM: input, forget, candidate, output_gate, output)."""

    GATES = {"rnn": 1, "gru": 3, "lstm": 4}

    def __init__(self, lib, cell, in_size, hidden, timesteps, return_sequences=True, v2=True, acts=None, relu_a=1.0):
        self.lib, self.cell, self.in_size, self.H, self.T = lib, cell, in_size, hidden, timesteps
        self.return_sequences = return_sequences
        G = self.GATES[cell]
        mk = lambda k: make_activation(lib, k, hidden, relu_a)
        if cell == "rnn":
            self._acts = [mk((acts or (ACT_TANH,))[0])]
            cfg = lib.RNNConfigCreate(in_size, hidden, return_sequences, timesteps, v2, self._acts[0])
            self.handle = lib.RNNCreateForInference(cfg)
            getw = lib.RNNGetWeights
        elif cell == "gru":
            z, h, r = acts or (ACT_SIGMOID, ACT_TANH, ACT_SIGMOID)
            self._acts = [mk(z), mk(h), mk(r)]
            cfg = lib.GRUConfigCreate(in_size, hidden, return_sequences, timesteps, lib.GRUActivationsCreate(*self._acts))
            self.handle = lib.GRUCreateForInference(cfg)
            getw = lib.GRUGetWeights
        else:
            self._acts = [mk(k) for k in (acts or (ACT_SIGMOID, ACT_SIGMOID, ACT_TANH, ACT_SIGMOID, ACT_TANH))]
            cfg = lib.LSTMConfigCreate(in_size, hidden, return_sequences, timesteps, v2, lib.LSTMActivationsCreate(*self._acts))
            self.handle = lib.LSTMCreateForInference(cfg)
            getw = lib.LSTMGetWeights
        if not self.handle:
            raise NNTKError(f"{cell} CreateForInference failed")
        w = getw(self.handle).contents
        self.W = _view(w.W, in_size * G * hidden).reshape(in_size, G * hidden)
        self.U = _view(w.U, hidden * G * hidden).reshape(hidden, G * hidden)
        self.b_i, self.b_h = _view(w.b_i, G * hidden), _view(w.b_h, G * hidden)

    def _fn(self, suffix):
        return getattr(self.lib, {"rnn": "RNN", "gru": "GRU", "lstm": "LSTM"}[self.cell] + suffix)

    def _shape(self):
        return (self.T, self.H) if self.return_sequences else (self.H,)

    def apply(self, x, out=None):
        """single sequence, STATEFUL like the reference"""
        out = _new_out(x, self._shape()) if out is None else out
        _check(self.lib, self._fn("ApplyInference")(self.handle, _ptr(x), _ptr(out)), self.cell + " ApplyInference")
        return out

    def apply_batch(self, x, out=None, h_state=None, c_state=None):
        n = x.shape[0]
        out = _new_out(x, (n,) + self._shape()) if out is None else out
        if h_state is None:
            rc = self._fn("ApplyInferenceBatch")(self.handle, _ptr(x), _ptr(out), n)
        elif self.cell == "lstm":
            rc = self._fn("ApplyInferenceBatchState")(self.handle, _ptr(x), _ptr(out), n, _ptr(h_state), _ptr(c_state))
        else:
            rc = self._fn("ApplyInferenceBatchState")(self.handle, _ptr(x), _ptr(out), n, _ptr(h_state))
        _check(self.lib, rc, self.cell + " ApplyInferenceBatch")
        return out

    def reset_state(self):
        _check(self.lib, self.lib.NNTKResetState(self.handle), "NNTKResetState")

    def close(self):
        if self.handle:
            self._fn("Destroy")(self.handle)
            self.handle = None
        for a in self._acts:
            if a:
                self.lib.ActivationFunctionDestroy(a)
        self._acts = []


def front_end(lib, spec, bn, conv, conv_bn, act, x, out=None):
    """SpectrogramBNConv1dApplyBatch: audio [n, input_size] -> [n, T_out, Cout]; bn / conv_bn / act may be None"""
    n = x.shape[0]
    out = _new_out(x, (n,) + conv.out_shape) if out is None else out
    rc = lib.SpectrogramBNConv1dApplyBatch(spec.handle, bn.handle if bn else None, conv.handle, conv_bn.handle if conv_bn else None,
                                           act.handle if act else None, _ptr(x), _ptr(out), n)
    _check(lib, rc, "SpectrogramBNConv1dApplyBatch")
    return out


def freeze(lib, *layers):
    for l in layers:
        _check(lib, lib.NNTKWeightsFreeze(l.handle), "NNTKWeightsFreeze")


class DFT:
    """signal/dft.h: complex DFT on planar buffers; `perform` takes [nfft] or [batch, nfft] real / imaginary planes"""

    def __init__(self, lib, nfft, forward=True):
        self.lib, self.nfft = lib, nfft
        self.handle = lib.DFTSetupCreate(lib.DFTConfigCreate(nfft, forward, True))
        if not self.handle:
            raise NNTKError("DFTSetupCreate failed")

    def perform(self, re, im, out_re=None, out_im=None):
        out_re = _new_out(re, re.shape) if out_re is None else out_re
        out_im = _new_out(im, im.shape) if out_im is None else out_im
        src = capi.ComplexFloatSplit(_ptr(re), _ptr(im))
        dst = capi.ComplexFloatSplit(_ptr(out_re), _ptr(out_im))
        if re.ndim == 1:
            self.lib.DFTPerform(self.handle, C.byref(src), C.byref(dst))
        else:
            _check(self.lib, self.lib.DFTPerformBatch(self.handle, C.byref(src), C.byref(dst), re.shape[0]), "DFTPerformBatch")
        return out_re, out_im

    def close(self):
        if self.handle:
            self.lib.DFTSetupDestroy(self.handle)
            self.handle = None


# ---- layers/bidirectional.h (inference half): x [batch, T, in], results [batch, rows, out] -------------------------------
def _rcfg(lib, in_size, out_size, return_sequences, timesteps):
    return lib.RecurrentConfigCreate(in_size, out_size, return_sequences, timesteps)


def bd_reverse_input(lib, x, out=None):
    b, t, c = x.shape
    out = _new_out(x, x.shape) if out is None else out
    lib.bd_reverse_input_batch(_ptr(x), _ptr(out), _rcfg(lib, c, c, True, t), b)
    return out


def bd_reverse_backward(lib, y, out=None):
    b, t, c = y.shape
    out = _new_out(y, y.shape) if out is None else out
    lib.bd_reverse_backward_batch(_ptr(y), _ptr(out), _rcfg(lib, c, c, True, t), b)
    return out


def _bd_cfg(lib, r):
    if r.ndim == 3:
        return _rcfg(lib, r.shape[2], r.shape[2], True, r.shape[1]), r.shape[0]
    return _rcfg(lib, r.shape[1], r.shape[1], False, 1), r.shape[0]


def bd_merge_concat(lib, fwd, bwd, out=None):
    cfg, b = _bd_cfg(lib, fwd)
    out = _new_out(fwd, tuple(fwd.shape[:-1]) + (2 * fwd.shape[-1],)) if out is None else out
    scratch = np.zeros(lib.bd_merge_concat_buffer_size(cfg), np.float32)   # the reference needs it; ours ignores it
    lib.bd_merge_concat(_ptr(fwd), _ptr(bwd), _ptr(out), cfg, b, scratch.ctypes.data)
    return out


def bd_merge_sum(lib, fwd, bwd, out=None):
    cfg, b = _bd_cfg(lib, fwd)
    out = _new_out(fwd, fwd.shape) if out is None else out
    lib.bd_merge_sum(_ptr(fwd), _ptr(bwd), _ptr(out), cfg, b)
    return out
