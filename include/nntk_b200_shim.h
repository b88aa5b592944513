/*
 * nntk_b200_shim.h -- the thin C-ABI between the host C layer code (nntoolkitcore_b200/host/*.c) and the
 * hand-written sm_100a CUDA kernels (nntoolkitcore_b200/csrc/*.cu).
 *
 * This is what replaces the reference's backend for the inference path: where the reference's layers call the
 * 25 op_* primitives of core/ops.h:19-67 (implemented by core/default_ops.cc on Eigen/NEON/libm) and DFTPerform
 * (signal/dft.c:34-47 on kissfft), the host code here calls ONE shim function per fused stage.  A per-op_* GPU
 * port would launch a kernel per 128-float vector add, so the boundary sits at layer-Apply granularity.
 *
 * Rules: plain C types only; every pointer named d_* is a DEVICE pointer; every function enqueues on the library
 * stream (nntk_cuda_stream) and returns 0 or a negative NNTK_ERR_* code with the text in nntk_cuda_last_error().
 * There is no CPU implementation behind any of these: without a CUDA device they fail.
 */
#ifndef NNTK_B200_SHIM_H
#define NNTK_B200_SHIM_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* activation kind tags (also stored in ActivationFunction objects) */
enum {
    NNTK_ACT_NONE = 0,
    NNTK_ACT_IDENTITY = 1,
    NNTK_ACT_SIGMOID = 2,
    NNTK_ACT_TANH = 3,
    NNTK_ACT_RELU = 4,
    NNTK_ACT_SOFTMAX = 5,
    NNTK_ACT_CUSTOM = 100
};

enum { NNTK_CELL_RNN = 1, NNTK_CELL_GRU = 3, NNTK_CELL_LSTM = 4 }; /* value = gate count G */

/* ---- runtime ---------------------------------------------------------------------------------------------------- */
int nntk_cuda_init(int device); /* device < 0: keep the current device */
int nntk_cuda_ensure_init(void);
int nntk_cuda_set_stream(void *stream);
void *nntk_cuda_stream(void);
int nntk_cuda_sync(void);
const char *nntk_cuda_last_error(void);
void nntk_cuda_clear_error(void);
void nntk_cuda_set_error(int code, const char *fmt, ...);
int nntk_cuda_set_math_mode(int mode);
int nntk_cuda_math_mode(void);
int nntk_cuda_sm_count(void);
/* kernel launches enqueued by this library since load (every launch site counts itself; bench.py's `gpu_launches`) */
void nntk_cuda_count_launch(int n);
long long nntk_cuda_launch_count(void);
/* Process-wide recursive API lock.  The runtime state behind the shim (library stream, math mode, shared workspace, packed-
 * operand scratch, the host-buffer staging pipeline) is per process, so every public Apply / Freeze entry of host/*.c holds
 * this lock for the duration of its enqueue: calls from different host threads on different handles serialise instead of
 * racing (the reference keeps all scratch per handle, SURVEY.md section 8b "Threading"). */
void nntk_cuda_api_lock(void);
void nntk_cuda_api_unlock(void);

int nntk_cuda_malloc(void **d_ptr, size_t bytes);
void nntk_cuda_free(void *d_ptr);
int nntk_cuda_host_alloc(void **h_ptr, size_t bytes); /* pinned; falls back to malloc when no device exists so that
                                                         Create/GetWeights work on a CPU-only box (Apply still fails) */
void nntk_cuda_host_free(void *h_ptr);
int nntk_cuda_h2d(void *d_dst, const void *h_src, size_t bytes);
int nntk_cuda_d2h(void *h_dst, const void *d_src, size_t bytes);
int nntk_cuda_d2d(void *d_dst, const void *d_src, size_t bytes);
int nntk_cuda_memset(void *d_dst, int value, size_t bytes);
/* events on the library stream (weight-snapshot reuse, host/common.c) */
int nntk_cuda_event_create(void **ev);
int nntk_cuda_event_record(void *ev);
int nntk_cuda_event_sync(void *ev);
void nntk_cuda_event_destroy(void *ev);
/* 1 = device (or managed) pointer, 0 = host pointer */
int nntk_cuda_is_device_ptr(const void *ptr);

/* Shared grow-only device scratch: asks for `want` bytes, accepts anything >= at_least (capped by NNTK_WORKSPACE_MB,
 * default 16 GiB, and by half the free memory). */
int nntk_cuda_workspace(size_t want, size_t at_least, void **d_ptr, size_t *got);

/* Host-buffer pipeline: runs `launch(ctx, d_in, d_out, first_unit, n_units)` over `n_units` independent units whose
 * inputs/outputs live in HOST memory, in chunks, with H2D / compute / D2H of neighbouring chunks overlapped on three
 * streams and two device slots.  Synchronous: the outputs are in h_out on return. */
typedef int (*nntk_launch_fn)(void *ctx, const float *d_in, float *d_out, long long first_unit, int n_units);
int nntk_cuda_staged_run(const void *h_in, void *h_out, size_t unit_in_bytes, size_t unit_out_bytes,
                         long long n_units, nntk_launch_fn launch, void *ctx);

/* ---- K1: spectrogram ---------------------------------------------------------------------------------------------- */
typedef struct {
    int nfft, window_size, step, input_size, nfreq, ntime;
    float norm_factor; /* fft_normalization_factor (signal/spectrogram.c:130-131) */
    int mode;          /* 0 magnitude (spectrogram.c:29-34), 1 PSD (spectrogram.c:41-47) */
    float scale;       /* sum(w) or fs*sum(w^2), computed on the host like spectrogram.c:36-38,49-57 */
    const float *d_window;  /* [window_size] */
    const float *d_twiddle; /* [nfft][2] = (cos, sin)(-2*pi*i/nfft) rounded from double (kiss_fft.c:357-363) */
} nntk_spectrogram_args;
/* d_in [n_clips, input_size] -> d_out [n_clips, ntime, nfreq] */
int nntk_cuda_spectrogram(const nntk_spectrogram_args *a, const float *d_in, float *d_out, int n_clips);

/* The same with rows of out_pitch >= nfreq floats (pad columns zero): what the fused front end (SpectrogramBNConv1dApplyBatch)
 * feeds the Conv1d's TMA im2col view without a re-layout pass.  Returns 1 (nothing enqueued) when the shape has no pitched kernel. */
int nntk_cuda_spectrogram_pitched(const nntk_spectrogram_args *a, const float *d_in, float *d_out, int n_clips, int out_pitch);
/* The tensor-core front end's variant: rows of out_pitch fp16 values (even, nfreq < out_pitch <= 288; same 10 explicit mantissa
 * bits as a TF32 operand, half the bytes).  *d_ovf is set to 1 when a value does not fit fp16's range: the caller's fp32 path,
 * enqueued behind it and gated on that flag, then produces the result (signal/spectrogram.c:113-135 either way). */
int nntk_cuda_spectrogram_half(const nntk_spectrogram_args *a, const float *d_in, void *d_out_half, int n_clips, int out_pitch, int *d_ovf);
/* gate for the NEXT 512-point spectrogram launches, until cleared with NULL: they return at once unless *d_flag == run_if */
void nntk_cuda_spectrogram_gate(const int *d_flag, int run_if);
/* device address of the front end's overflow flag (one int per device, in the library image); NULL without a device */
int *nntk_cuda_front_flag(void);

/* K1 with the log-mel epilogue (signal/log_mel_spectrogram.c:33-38 in one kernel).
 * The mel table is passed as a per-lane work list: entry (iter, lane) = {packed, weight}; packed = 8 * slot(bin) |
 * band << 15 | last << 31 with slot(k) = k + k / 16 (the kernel's conflict-free parking order of the spectrum).  A lane
 * accumulates weight * spectrum[bin] over consecutive entries and, at an entry flagged `last`, emits its sum as mel band
 * `band` and restarts.  The host balances the ~2 non-zeros per bin over the 32 lanes and orders them against bank conflicts. */
typedef struct {
    int n_mels;
    int n_iter;             /* entries per lane */
    const void *d_entries;  /* [n_iter][32] x 8 bytes */
    float log_offset;       /* 1.5849e-13f, log_mel_spectrogram.c:36 */
} nntk_mel_args;
/* d_in [n_clips, input_size] -> d_out [n_clips, ntime, n_mels].  Returns 1 (nothing enqueued) when the fused kernel does not
 * cover the shape (nfft != 512, more than 384 bands): the caller then runs spectrogram -> GEMM -> nntk_cuda_log_offset. */
int nntk_cuda_log_mel_spectrogram(const nntk_spectrogram_args *a, const nntk_mel_args *m, const float *d_in, float *d_out,
                                  int n_clips);
/* d_out[i] = logf(d_in[i] + offset)   (op_vec_add_sc + op_vec_log, log_mel_spectrogram.c:36-37) */
int nntk_cuda_log_offset(const float *d_in, float *d_out, long long n, float offset);

/* Standalone complex DFT on planar buffers (signal/dft.c:34-47).  d_twiddle is the FORWARD table
 * [nfft][2] of nntk_spectrogram_args; inverse != 0 conjugates it.  [batch, nfft] planes, in-place allowed. */
int nntk_cuda_dft(int nfft, int inverse, const float *d_twiddle, const float *d_in_re, const float *d_in_im, float *d_out_re,
                  float *d_out_im, long long batch);
/* strided element copies behind split_complex / join_complex_split on device pointers: dst[i * dst_stride] = src[i * src_stride] */
int nntk_cuda_copy_strided(float *d_dst, int dst_stride, const float *d_src, int src_stride, long long n);

/* ---- K5: element-wise --------------------------------------------------------------------------------------------- */
/* d_params = gamma|beta|mean|var (4*C floats); rows x C elements; layers/batch_norm.c:140-163 */
int nntk_cuda_batch_norm(const float *d_params, int C, float eps, const float *d_in, float *d_out, long long rows);
/* kind != softmax: n elements.  softmax: n vectors of vector_size; layers/activation_default.c */
int nntk_cuda_activation(int kind, float relu_a, int vector_size, const float *d_in, float *d_out, long long n);

/* ---- K2/K3: (implicit) GEMM + bias + BatchNorm + activation --------------------------------------------------------- */
typedef struct {
    /* A is `batch` slabs of `rows` rows: row r of slab s starts at d_A + s*a_slab_stride + r*lda, K floats long.
     * Conv1d: rows = T_out, lda = stride*Cin, K = kernel*Cin, a_slab_stride = T*Cin (overlapping im2col rows).
     * Dense / projections: rows = M, lda = K, batch = 1. */
    const float *d_A;
    long long a_slab_stride;
    int lda;
    int batch, rows;
    int N, K;
    const float *d_B;    /* [K][N] row-major */
    const float *d_bias; /* [N] or NULL */
    const float *d_bn;   /* gamma|beta|mean|var (4*N) or NULL; applied after the bias */
    float bn_eps;
    int act;             /* NNTK_ACT_* element-wise kinds only (softmax is a second pass by the caller) */
    float act_a;
    float *d_C;          /* row r of slab s at d_C + s*c_slab_stride + r*ldc */
    long long c_slab_stride;
    int ldc;
} nntk_gemm_args;
int nntk_cuda_gemm(const nntk_gemm_args *g);
/* C = softmax over each whole row of (A B + bias), N <= 32, in the tensor-core GEMM's epilogue (layers/dense.c:122-142 followed by
 * activation_default.c:149-167 in one kernel).  0 done, 1 not eligible (nothing enqueued), < 0 error */
int nntk_cuda_gemm_row_softmax(const nntk_gemm_args *g);
/* the tcgen05 kernel directly (split 0: single-pass TF32, 1: 3xTF32): 0 done, 1 shape not eligible (nothing enqueued), < 0 error */
int nntk_cuda_gemm_tf32(const nntk_gemm_args *g, int split);
/* Single-pass fp16-operand tcgen05 GEMM for the tensor-core front end: g->d_A points to fp16 rows (lda / a_slab_stride count
 * halves, multiples of 8), d_B is fp32 and repacked to fp16 here; *d_ovf is set when a weight does not fit fp16 and the GEMM
 * kernel only runs while *d_ovf == 0.  Returns 0 done, 1 not eligible (nothing enqueued), < 0 error.  (layers/conv_1d.c:122-155) */
int nntk_cuda_gemm_f16(const nntk_gemm_args *g, int *d_ovf);
/* gate for the NEXT tensor-core GEMM launches, until cleared with NULL: they return at once unless *d_flag == run_if */
void nntk_cuda_gemm_tc_gate(const int *d_flag, int run_if);
/* conv weights W[Cout][Cin][K] -> B[(k*Cin + ci)][co]   (layers/conv_1d.c:128-140 order -> implicit-GEMM order) */
int nntk_cuda_pack_conv_weights(const float *d_W, float *d_B, int cout, int cin, int k);
/* the same with the channel dimension zero-padded to cin_pad (rows k*cin_pad + ci, ci >= cin are zero), and the matching
 * activation repack [rows][C] -> [rows][Cpad]: give unaligned channel counts (Cin = 257) a 16-byte pitch for TMA */
int nntk_cuda_pack_conv_weights_padded(const float *d_W, float *d_B, int cout, int cin, int cin_pad, int k);
int nntk_cuda_pad_channels(const float *d_in, float *d_out, long long rows, int C, int Cpad);
/* Conv weights with an input BatchNorm folded in (fused front end): conv(BN(x)) = conv'(x) with
 *   B'[(k*cin_pad + ci)][co] = W[co][ci][k] * a[ci],  bias'[co] = b[co] + sum_{ci,k} W[co][ci][k] * (beta[ci] - mean[ci] * a[ci]),
 *   a[ci] = gamma[ci] / sqrt(var[ci] + eps)   (layers/batch_norm.c:151-162 composed with layers/conv_1d.c:128-144).
 * d_bn = gamma|beta|mean|var (4*cin); d_bias_out [cout]. */
int nntk_cuda_pack_conv_weights_bn(const float *d_W, const float *d_b, const float *d_bn, float eps, float *d_B, float *d_bias_out,
                                   int cout, int cin, int cin_pad, int k);

/* bidirectional helpers (layers/bidirectional.c:10-85) */
/* d_out[b][t][:] = d_in[b][T-1-t][:], rows of `vec` floats; not in-place */
int nntk_cuda_bd_reverse(const float *d_in, float *d_out, int vec, int T, long long batch);
/* d_out[row] = d_fwd[row] | d_bwd[row], rows of `cols` floats each */
int nntk_cuda_bd_concat(const float *d_fwd, const float *d_bwd, float *d_out, int cols, long long rows);
int nntk_cuda_bd_sum(const float *d_a, const float *d_b, float *d_out, long long n);

/* ---- K3+K4: recurrent layers ---------------------------------------------------------------------------------------- */
typedef struct {
    int cell;            /* NNTK_CELL_* */
    int in, H, T, batch;
    int return_sequences;
    int v2;              /* RNN/LSTM: add b_h (rnn.c:161-163, lstm.c:207-209); GRU always adds it (gru.c:151) */
    int act[5];          /* RNN: [0]; GRU: z, r, h~; LSTM: i, f, g, o, out */
    float act_a[5];      /* ReLU output scale `a` per slot (activation_default.c:123-129) */
    const float *d_W;    /* [in][G*H] */
    const float *d_U;    /* [H][G*H]  */
    const float *d_bi;   /* [G*H] */
    const float *d_bh;   /* [G*H] */
    const float *d_x;    /* [batch, T, in] */
    float *d_h;          /* [batch, H] state: read as h_0, overwritten with h_T */
    float *d_c;          /* [batch, H] LSTM cell state (in/out), else NULL */
    float *d_out;        /* [batch, T, H] or [batch, H] */
    float *d_work;       /* workspace of nntk_cuda_recurrent_workspace() bytes */
    size_t work_bytes;
} nntk_recurrent_args;
size_t nntk_cuda_recurrent_workspace(const nntk_recurrent_args *r);
/* name of the kernel family that ran the most recent recurrent chunk: "fused_tc", "cluster_h16", "cluster_tf32", "cluster_fp32",
 * "step_tc_f16", "step_tc_x3", "step_fp32" (diagnostics; tests pin the dispatch at the BASELINE shapes with it) */
const char *nntk_cuda_last_recurrent_path(void);
/* input passes (1 or 3) the most recent fused recurrent launch took after its adaptive pre-scan; synchronises (tests only) */
int nntk_cuda_last_recurrent_xpasses(void);
/* operand passes the hoisted projection of the most recent large-batch recurrent layer took in tensor-core mode: 1 (single fp16 pass:
 * max |x_t| max |W_n| under the bound), 3 (3xbf16), 0 (no adaptive projection ran); synchronises (tests only) */
int nntk_cuda_last_proj_passes(void);
int nntk_cuda_recurrent(const nntk_recurrent_args *r);
/* Allocation-free variant of nntk_cuda_recurrent.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  The caller
 * passes weights, inputs, state and outputs as device pointers exactly as for nntk_cuda_recurrent; d_work / work_bytes are
 * ignored: the scratch (hoisted projection + ping-pong buffers, 4 (batch T G H + 6 batch H) bytes) lives in a file-scope device
 * array of NNTK_RECURRENT_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Always the fp32 (FFMA) path,
 * whatever the math mode.  Larger problems are rejected with NNTK_ERR_ARG and a message in nntk_cuda_last_error().  Calls no
 * cudaMalloc / cudaFree / allocating library routine.  All calls share the one array: calls whose kernels overlap on the GPU
 * (different streams) race; calls on one stream are safe, nothing in the array has to be cleared between them. */
#define NNTK_RECURRENT_STATIC_SCRATCH_BYTES (32u << 20)
int nntk_cuda_recurrent_static_scratch(const nntk_recurrent_args *r);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_SHIM_H */
