"""ctypes binding of the NNToolkitCore C inference API (include/nntoolkitcore/**.h).

The by-value structs and signatures below are the reference's own ABI, so `bind()` works on
BOTH libraries: libnntoolkitcore_b200.so (the product) and oracle/_ref/libnntk_ref.so (the unmodified reference, used
only by tests / bench as the checker).  The additive b200_ext.h entry points are bound when the library exports them.
"""
from __future__ import annotations

import ctypes as C
from ctypes import POINTER, c_bool, c_float, c_int, c_size_t, c_void_p, c_char_p

FP = POINTER(c_float)
window_fn = C.CFUNCTYPE(None, FP, c_int)


class SpectrogramConfig(C.Structure):
    _fields_ = [("nfft", c_int), ("window_size", c_int), ("noverlap", c_int), ("step", c_int), ("input_size", c_int),
                ("nfreq", c_int), ("ntime_series", c_int), ("fft_normalization_factor", c_float)]


class MelFilterBankConfig(C.Structure):
    _fields_ = [("n_mels", c_int), ("n_fft", c_int), ("sample_rate", c_int), ("lower_hz", c_float), ("upper_hz", c_float)]


class DFTConfig(C.Structure):
    _fields_ = [("nfft", c_int), ("forward", c_bool), ("complex", c_bool)]


class ComplexFloatSplit(C.Structure):
    _fields_ = [("real_p", c_void_p), ("imag_p", c_void_p)]


class BatchNormWeights(C.Structure):
    _fields_ = [("gamma", FP), ("beta", FP), ("moving_mean", FP), ("moving_variance", FP)]


class BatchNormConfig(C.Structure):
    _fields_ = [("feature_channels", c_int), ("epsilon", c_float), ("count", c_int)]


class Conv1dConfig(C.Structure):
    _fields_ = [("input_feature_channels", c_int), ("output_feature_channels", c_int), ("kernel_size", c_int),
                ("stride", c_int), ("input_size", c_int), ("output_size", c_int)]


class DefaultWeights(C.Structure):
    _fields_ = [("W", FP), ("b", FP)]


class DenseConfig(C.Structure):
    _fields_ = [("input_size", c_int), ("output_size", c_int), ("activation", c_void_p)]


class TimeDistributedDenseConfig(C.Structure):
    _fields_ = [("dense", DenseConfig), ("ts", c_int)]


class RecurrentWeights(C.Structure):
    _fields_ = [("W", FP), ("U", FP), ("b_i", FP), ("b_h", FP)]


class RecurrentConfig(C.Structure):
    _fields_ = [("input_feature_channels", c_int), ("output_feature_channels", c_int), ("return_sequences", c_bool),
                ("timesteps", c_int)]


class RNNConfig(C.Structure):
    _fields_ = [("base", RecurrentConfig), ("v2", c_bool), ("activation", c_void_p)]


class GRUActivations(C.Structure):
    _fields_ = [("z_gate_activation", c_void_p), ("h_gate_activation", c_void_p), ("r_gate_activation", c_void_p)]


class GRUConfig(C.Structure):
    _fields_ = [("base", RecurrentConfig), ("activations", GRUActivations)]


class LSTMActivations(C.Structure):
    _fields_ = [("candidate_gate_activation", c_void_p), ("input_gate_activation", c_void_p),
                ("forget_gate_activation", c_void_p), ("output_gate_activation", c_void_p),
                ("output_activation", c_void_p)]


class LSTMConfig(C.Structure):
    _fields_ = [("base", RecurrentConfig), ("v2", c_bool), ("activations", LSTMActivations)]


H = c_void_p  # opaque handles

# name -> (restype, argtypes); the reference's inference API
REFERENCE_API = {
    # signal/window.h
    "ones": (None, [FP, c_int]), "hann_window": (None, [FP, c_int]), "hamming_window": (None, [FP, c_int]),
    "periodic_hann_window": (None, [FP, c_int]), "periodic_hamming_window": (None, [FP, c_int]),
    "blackman_window": (None, [FP, c_int]),
    # signal/spectrogram.h
    "SpectrogramConfigCreate": (SpectrogramConfig, [c_int, c_int, c_int, c_int, c_float]),
    "SpectrogramCreateMagnitude": (H, [SpectrogramConfig]),
    "SpectrogramCreatePSD": (H, [SpectrogramConfig, c_int]),
    "SpectrogramGetConfig": (SpectrogramConfig, [H]),
    "SpectrogramSetWindowFunc": (None, [H, c_void_p]),
    "SpectrogramSetScaleFactor": (None, [H, c_float]),
    "SpectrogramApply": (None, [H, c_void_p, c_void_p]),
    "SpectrogramDestroy": (None, [H]),
    # signal/mel_filterbank.h, signal/log_mel_spectrogram.h
    "MelFilterBankConfigCreate": (MelFilterBankConfig, [c_int, c_int, c_int, c_float, c_float]),
    "MelFilterBankCreate": (H, [MelFilterBankConfig]),
    "MelFilterBankApply": (None, [H, c_void_p, c_void_p, c_int]),
    "MelFilterBankDestroy": (None, [H]),
    "LogMelSpectrogramCreate": (H, [H, MelFilterBankConfig]),
    "LogMelSpectrogramApply": (None, [H, c_void_p, c_void_p]),
    "LogMelSpectrogramDestroy": (None, [H]),
    # signal/dft.h
    "DFTConfigCreate": (DFTConfig, [c_int, c_bool, c_bool]), "DFTSetupCreate": (H, [DFTConfig]),
    "DFTPerform": (None, [H, POINTER(ComplexFloatSplit), POINTER(ComplexFloatSplit)]), "DFTSetupDestroy": (None, [H]),
    "split_complex": (None, [c_void_p, POINTER(ComplexFloatSplit), c_int]),
    "join_complex_split": (None, [POINTER(ComplexFloatSplit), c_void_p, c_int]),
    # layers/bidirectional.h (inference half)
    "bd_reverse_input_batch": (None, [c_void_p, c_void_p, RecurrentConfig, c_int]),
    "bd_reverse_backward_batch": (None, [c_void_p, c_void_p, RecurrentConfig, c_int]),
    "bd_merge_concat_buffer_size": (c_int, [RecurrentConfig]),
    "bd_merge_concat": (None, [c_void_p, c_void_p, c_void_p, RecurrentConfig, c_int, c_void_p]),
    "bd_merge_sum": (None, [c_void_p, c_void_p, c_void_p, RecurrentConfig, c_int]),
    # layers/activation*.h
    "ActivationFunctionCreate": (H, [c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "ActivationFunctionCreateIdentity": (H, [c_int]), "ActivationFunctionCreateSigmoid": (H, [c_int]),
    "ActivationFunctionCreateTanh": (H, [c_int]), "ActivationFunctionCreateReLU": (H, [c_int, c_float]),
    "ActivationFunctionCreateSoftmax": (H, [c_int, c_int]),
    "ActivationFunctionApply": (None, [H, c_void_p, c_void_p]), "ActivationFunctionDestroy": (None, [H]),
    # layers/batch_norm.h
    "BatchNormConfigCreate": (BatchNormConfig, [c_int, c_float, c_int]),
    "BatchNormCreateForInference": (H, [BatchNormConfig]),
    "BatchNormGetWeights": (POINTER(BatchNormWeights), [H]),
    "BatchNormApplyInference": (c_int, [H, c_void_p, c_void_p]), "BatchNormDestroy": (None, [H]),
    # layers/conv_1d.h
    "Conv1dConfigCreate": (Conv1dConfig, [c_int, c_int, c_int, c_int, c_int]),
    "Conv1dCreateForInference": (H, [Conv1dConfig]), "Conv1dGetWeights": (POINTER(DefaultWeights), [H]),
    "Conv1dApplyInference": (c_int, [H, c_void_p, c_void_p]), "Conv1dDestroy": (None, [H]),
    # layers/dense.h, time_distributed_dense.h
    "DenseConfigCreate": (DenseConfig, [c_int, c_int, c_void_p]), "DenseCreateForInference": (H, [DenseConfig]),
    "DenseGetWeights": (POINTER(DefaultWeights), [H]), "DenseApplyInference": (c_int, [H, c_void_p, c_void_p]),
    "DenseDestroy": (None, [H]),
    "TimeDistributedDenseConfigCreate": (TimeDistributedDenseConfig, [c_int, DenseConfig]),
    "TimeDistributedDenseCreateForInference": (H, [TimeDistributedDenseConfig]),
    "TimeDistributedDenseGetWeights": (POINTER(DefaultWeights), [H]),
    "TimeDistributedDenseApplyInference": (c_int, [H, c_void_p, c_void_p]),
    "TimeDistributedDenseDestroy": (None, [H]),
    # layers/recurrent.h, rnn.h, gru.h, lstm.h
    "RecurrentConfigCreate": (RecurrentConfig, [c_int, c_int, c_bool, c_int]),
    "RNNConfigCreate": (RNNConfig, [c_int, c_int, c_bool, c_int, c_bool, c_void_p]),
    "RNNCreateForInference": (H, [RNNConfig]), "RNNGetWeights": (POINTER(RecurrentWeights), [H]),
    "RNNApplyInference": (c_int, [H, c_void_p, c_void_p]), "RNNDestroy": (None, [H]),
    "GRUActivationsCreate": (GRUActivations, [c_void_p, c_void_p, c_void_p]),
    "GRUActivationsCreateDefault": (GRUActivations, [c_int]), "GRUActivationsDestroy": (None, [GRUActivations]),
    "GRUConfigCreate": (GRUConfig, [c_int, c_int, c_bool, c_int, GRUActivations]),
    "GRUCreateForInference": (H, [GRUConfig]), "GRUGetWeights": (POINTER(RecurrentWeights), [H]),
    "GRUApplyInference": (c_int, [H, c_void_p, c_void_p]), "GRUDestroy": (None, [H]),
    "LSTMActivationsCreate": (LSTMActivations, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "LSTMActivationsCreateDefault": (LSTMActivations, [c_int]), "LSTMActivationsDestroy": (None, [LSTMActivations]),
    "LSTMConfigCreate": (LSTMConfig, [c_int, c_int, c_bool, c_int, c_bool, LSTMActivations]),
    "LSTMCreateForInference": (H, [LSTMConfig]), "LSTMGetWeights": (POINTER(RecurrentWeights), [H]),
    "LSTMApplyInference": (c_int, [H, c_void_p, c_void_p]), "LSTMDestroy": (None, [H]),
}

# additive entry points of include/nntoolkitcore/b200_ext.h
EXT_API = {
    "NNTKInit": (c_int, [c_int]), "NNTKSetStream": (c_int, [c_void_p]), "NNTKGetStream": (c_void_p, []),
    "NNTKSynchronize": (c_int, []), "NNTKLastError": (c_char_p, []), "NNTKClearError": (None, []),
    "NNTKDebugLaunchCount": (C.c_longlong, []), "NNTKDebugLastRecurrentPath": (c_char_p, []), "NNTKDebugLastRecurrentInputPasses": (c_int, []),
    "NNTKSetMathMode": (c_int, [c_int]), "NNTKGetMathMode": (c_int, []),
    "NNTKDeviceAlloc": (c_void_p, [c_size_t]), "NNTKDeviceFree": (None, [c_void_p]),
    "NNTKHostAlloc": (c_void_p, [c_size_t]), "NNTKHostFree": (None, [c_void_p]),
    "NNTKCopyToDevice": (c_int, [c_void_p, c_void_p, c_size_t]), "NNTKCopyToHost": (c_int, [c_void_p, c_void_p, c_size_t]),
    "NNTKIsDevicePointer": (c_int, [c_void_p]),
    "NNTKWeightsFreeze": (c_int, [c_void_p]), "NNTKWeightsUnfreeze": (c_int, [c_void_p]),
    "NNTKResetState": (c_int, [c_void_p]),
    "SpectrogramApplyBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "LogMelSpectrogramApplyBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "DFTPerformBatch": (c_int, [H, POINTER(ComplexFloatSplit), POINTER(ComplexFloatSplit), c_int]),
    "BatchNormApplyInferenceBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "ActivationFunctionApplyBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "Conv1dApplyInferenceBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "Conv1dBNActApplyInferenceBatch": (c_int, [H, H, H, c_void_p, c_void_p, c_int]),
    "SpectrogramBNConv1dApplyBatch": (c_int, [H, H, H, H, H, c_void_p, c_void_p, c_int]),
    "DenseApplyInferenceBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "TimeDistributedDenseApplyInferenceBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "RNNApplyInferenceBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "GRUApplyInferenceBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "LSTMApplyInferenceBatch": (c_int, [H, c_void_p, c_void_p, c_int]),
    "RNNApplyInferenceBatchState": (c_int, [H, c_void_p, c_void_p, c_int, c_void_p]),
    "GRUApplyInferenceBatchState": (c_int, [H, c_void_p, c_void_p, c_int, c_void_p]),
    "LSTMApplyInferenceBatchState": (c_int, [H, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
}

# the internal shim ABI of include/nntk_b200_shim.h (exported so the symbol test can see the whole boundary)
SHIM_SYMBOLS = [
    "nntk_cuda_init", "nntk_cuda_ensure_init", "nntk_cuda_set_stream", "nntk_cuda_stream", "nntk_cuda_sync",
    "nntk_cuda_last_error", "nntk_cuda_clear_error", "nntk_cuda_set_error", "nntk_cuda_set_math_mode",
    "nntk_cuda_math_mode", "nntk_cuda_sm_count", "nntk_cuda_count_launch", "nntk_cuda_launch_count",
    "nntk_cuda_api_lock", "nntk_cuda_api_unlock", "nntk_cuda_malloc", "nntk_cuda_free", "nntk_cuda_host_alloc",
    "nntk_cuda_host_free", "nntk_cuda_h2d", "nntk_cuda_d2h", "nntk_cuda_d2d", "nntk_cuda_memset",
    "nntk_cuda_is_device_ptr", "nntk_cuda_event_create", "nntk_cuda_event_record", "nntk_cuda_event_sync",
    "nntk_cuda_event_destroy", "nntk_cuda_workspace", "nntk_cuda_staged_run", "nntk_cuda_spectrogram",
    "nntk_cuda_log_mel_spectrogram", "nntk_cuda_log_offset", "nntk_cuda_dft", "nntk_cuda_copy_strided",
    "nntk_cuda_bd_reverse", "nntk_cuda_bd_concat", "nntk_cuda_bd_sum",
    "nntk_cuda_batch_norm", "nntk_cuda_activation", "nntk_cuda_gemm", "nntk_cuda_pack_conv_weights",
    "nntk_cuda_pack_conv_weights_padded", "nntk_cuda_pad_channels", "nntk_cuda_pack_conv_weights_bn",
    "nntk_cuda_spectrogram_pitched",
    "nntk_cuda_recurrent_workspace", "nntk_cuda_recurrent", "nntk_cuda_last_recurrent_path", "nntk_cuda_last_recurrent_xpasses",
    "nntk_cuda_recurrent_static_scratch", "nntk_cuda_spectrogram_half", "nntk_cuda_spectrogram_gate", "nntk_cuda_front_flag",
    "nntk_cuda_gemm_f16", "nntk_cuda_gemm_tc_gate", "nntk_cuda_gemm_tf32", "nntk_cuda_gemm_row_softmax", "nntk_cuda_last_proj_passes",
]


def bind(path: str, ext: bool | None = None) -> C.CDLL:
    """dlopen `path` and attach restype/argtypes.  ext=None: bind the b200 extensions iff the library has them."""
    lib = C.CDLL(str(path), mode=C.RTLD_LOCAL)
    for name, (res, args) in REFERENCE_API.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    has_ext = hasattr(lib, "NNTKInit")
    if ext is True and not has_ext:
        raise RuntimeError(f"{path} does not export the b200 extensions")
    if has_ext and ext is not False:
        for name, (res, args) in EXT_API.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
    lib.has_ext = has_ext and ext is not False
    return lib
