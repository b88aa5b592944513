/*
 * nntoolkitcore/b200_ext.h -- ADDITIVE entry points of the B200 build.  Nothing here changes a reference symbol.
 *
 * Why they exist: the reference inference API has no batch dimension (one clip / one sequence per call, e.g.
 * layers/gru.c:189-204), while the B200 path is only worth running at batch 10^3..10^5.  The batched entries below
 * copy the layout and state semantics of the reference's OWN batched forward (the forward half of
 * *ApplyTrainingBatch): input [B, T, Cin] -> output [B, T_out, Cout] (or [B, H] without return_sequences), and the
 * recurrent state is ZERO at the start of every sequence (layers/gru.c:259-260, lstm.c:438-439, rnn.c:259-260,
 * conv_1d.c:167-183, dense.c:144-162).  The single-sequence reference calls stay stateful as in the reference.
 *
 * Pointer rule for every Apply in this library: input/output may both be HOST pointers (the call stages through
 * device memory -- chunked and double-buffered for the batched entries -- and is complete on return) or both be
 * DEVICE pointers (the call is enqueued on the library stream and returns immediately; order/sync with
 * NNTKSetStream / NNTKSynchronize).
 */
#ifndef NNTK_B200_EXT_H
#define NNTK_B200_EXT_H

#include <stddef.h>
#include "signal/spectrogram.h"
#include "signal/log_mel_spectrogram.h"
#include "signal/dft.h"
#include "layers/activation_default.h"
#include "layers/batch_norm.h"
#include "layers/conv_1d.h"
#include "layers/dense.h"
#include "layers/time_distributed_dense.h"
#include "layers/rnn.h"
#include "layers/gru.h"
#include "layers/lstm.h"
#include "layers/bidirectional.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---- return codes of every int-returning Apply ---------------------------------------------------------------- */
#define NNTK_OK 0
#define NNTK_ERR_MODE (-1)        /* the reference's only error: inference call on a training handle (never here) */
#define NNTK_ERR_UNSUPPORTED (-2) /* e.g. a custom host-callback activation inside a device layer */
#define NNTK_ERR_CUDA (-3)        /* CUDA runtime / launch failure, or no usable sm_100 device */
#define NNTK_ERR_ARG (-4)         /* NULL handle, mixed host/device pointers, bad sizes */

/* ---- runtime ---------------------------------------------------------------------------------------------------- */
/* Selects the CUDA device for this process and creates the runtime state.  Optional: the first Create/Apply calls
 * NNTKInit(current device).  Fails loudly (NNTK_ERR_CUDA + message) when there is no CUDA device. */
int NNTKInit(int device);
/* The library launches on ONE stream per process (default: the legacy default stream).  `cuda_stream` is a
 * cudaStream_t; pass NULL to go back to the default stream. */
int NNTKSetStream(void *cuda_stream);
void *NNTKGetStream(void);
int NNTKSynchronize(void);
/* Last error text of the calling thread ("" if none).  void-returning reference calls (SpectrogramApply,
 * ActivationFunctionApply) report failures only here and on stderr. */
const char *NNTKLastError(void);
void NNTKClearError(void);
/* Kernel launches enqueued by the library since it was loaded (diagnostics / benchmark accounting). */
long long NNTKDebugLaunchCount(void);
/* Which recurrent kernel family served the most recent RNN/GRU/LSTM Apply (see nntk_b200_shim.h). */
const char *NNTKDebugLastRecurrentPath(void);
int NNTKDebugLastRecurrentInputPasses(void);   /* 1 or 3: fp16 passes the fused recurrent kernel took for x.W (synchronises) */

/* GEMM-shaped stages (Conv1d, Dense/TDD, recurrent projections and steps) can run in three math modes. */
#define NNTK_MATH_FP32 0    /* FFMA on CUDA cores; matches the reference within 1e-5 */
#define NNTK_MATH_TF32 1    /* tcgen05 kind::tf32, operands rounded to 10-bit mantissa; 1e-3 */
/* DEFAULT (since round 2; process-wide override: NNTK_MATH_MODE=0|1|2): a caller that only relinks gets the tensor cores at the
 * reference's own tolerance.  Evidence: the whole fp32-contract GPU suite runs in this mode, and tests/test_gpu_robustness.py pins the
 * worst case of the truncating accumulate (non-negative operands, K up to the cap). */
#define NNTK_MATH_TF32X3 2  /* tcgen05, hi/lo operand split, 3 MMAs per product, accumulate truncation compensated: holds the
                               fp32 contract (1e-5) at ~1/3 of the TF32 rate (recurrent steps too, for batches above
                               256); reduction lengths above 1 536 stay on FFMA (the truncating
                               accumulate of the tensor core: csrc/gemm.cu) */
int NNTKSetMathMode(int mode);
int NNTKGetMathMode(void);

/* Device / pinned-host memory for C callers that do not link the CUDA runtime themselves. */
void *NNTKDeviceAlloc(size_t bytes);
void NNTKDeviceFree(void *ptr);
void *NNTKHostAlloc(size_t bytes); /* pinned */
void NNTKHostFree(void *ptr);
int NNTKCopyToDevice(void *dst_device, const void *src_host, size_t bytes); /* async on the library stream */
int NNTKCopyToHost(void *dst_host, const void *src_device, size_t bytes);   /* async on the library stream */
int NNTKIsDevicePointer(const void *ptr);

/* ---- weights and state ------------------------------------------------------------------------------------------ */
/* `layer` is any layer handle of this library (BatchNorm, Conv1d, Dense, TimeDistributedDense, RNN, GRU, LSTM).
 * The reference has no commit call: weights are whatever is behind XGetWeights() when Apply runs.  By default this
 * build therefore restages the (pinned) host block to the device at EVERY Apply.  Freeze stages it once now and
 * skips the restage until Unfreeze -- use it around hot loops. */
int NNTKWeightsFreeze(void *layer);
int NNTKWeightsUnfreeze(void *layer);
/* Zero the carried recurrent state (h, and c for LSTM) of an RNN/GRU/LSTM handle.  The reference offers no reset
 * (state is zeroed only at Create, layers/gru.c:67). */
int NNTKResetState(void *recurrent_layer);

/* ---- batched forwards ------------------------------------------------------------------------------------------- */
/* [n_clips, input_size] -> [n_clips, ntime_series, nfreq] */
int SpectrogramApplyBatch(Spectrogram filter, const float *input, float *output, int n_clips);
/* [n_clips, input_size] -> [n_clips, ntime_series, n_mels]; one fused kernel for nfft == 512 */
int LogMelSpectrogramApplyBatch(LogMelSpectrogram filter, const float *input, float *output, int n_clips);
/* `batch` nfft-point signals stored back to back in each plane: [batch, nfft] -> [batch, nfft] */
int DFTPerformBatch(DFTSetup setup, const ComplexFloatSplit *input, const ComplexFloatSplit *output, int batch);
/* [batch, count, C] -> same */
int BatchNormApplyInferenceBatch(BatchNorm filter, const float *input, float *output, int batch);
/* `batch` repetitions of the activation's own size */
int ActivationFunctionApplyBatch(ActivationFunction filter, const float *input, float *output, int batch);
/* [batch, T, Cin] -> [batch, T_out, Cout] */
int Conv1dApplyInferenceBatch(Conv1d filter, const float *input, float *output, int batch);
/* Conv1d -> BatchNorm(feature_channels == Cout) -> activation fused in one kernel epilogue; bn and/or act may be
 * NULL.  Equals the three reference calls conv_1d.c:149 -> batch_norm.c:166 -> activation.c:23 in that order. */
int Conv1dBNActApplyInferenceBatch(Conv1d filter, BatchNorm bn, ActivationFunction act, const float *input,
                                   float *output, int batch);
/* Fused front end: [n_clips, input_size] audio -> [n_clips, T_out, Cout].  Equals the reference calls signal/spectrogram.c:113 ->
 * layers/batch_norm.c:166 (input_bn over the nfreq bins; may be NULL) -> layers/conv_1d.c:149 -> output_bn / act (may be NULL) per
 * clip; the spectrogram never takes a separate BatchNorm or re-layout pass (the BatchNorm is folded into the conv weights: results
 * agree with the separate calls to ~2e-7 relative, not bit for bit). */
int SpectrogramBNConv1dApplyBatch(Spectrogram spectrogram, BatchNorm input_bn, Conv1d filter, BatchNorm output_bn,
                                  ActivationFunction act, const float *input, float *output, int n_clips);
/* [batch, in] -> [batch, out] */
int DenseApplyInferenceBatch(Dense filter, const float *input, float *output, int batch);
/* [batch, ts, in] -> [batch, ts, out] */
int TimeDistributedDenseApplyInferenceBatch(TimeDistributedDense filter, const float *input, float *output,
                                            int batch);
/* [batch, T, in] -> [batch, T, H] or [batch, H]; zero initial state per sequence; the handle's own carried state is
 * neither read nor written. */
int RNNApplyInferenceBatch(RNN filter, const float *input, float *output, int batch);
int GRUApplyInferenceBatch(GRU filter, const float *input, float *output, int batch);
int LSTMApplyInferenceBatch(LSTM filter, const float *input, float *output, int batch);
/* Batched STREAMING: like the batched calls, but the initial state is read from, and the final state written to,
 * caller-owned [batch, H] buffers (same host/device kind as input), so many concurrent streams can be fed chunk by
 * chunk -- the batch-scale form of the reference's single-stream statefulness (gru.c:197-201, lstm.c:257-265). */
int RNNApplyInferenceBatchState(RNN filter, const float *input, float *output, int batch, float *h_state);
int GRUApplyInferenceBatchState(GRU filter, const float *input, float *output, int batch, float *h_state);
int LSTMApplyInferenceBatchState(LSTM filter, const float *input, float *output, int batch, float *h_state,
                                 float *c_state);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_EXT_H */
