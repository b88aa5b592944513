/*
 * nntk_oracle.h -- CPU restatement ("port") of the NNToolkitCore inference forward path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is linked into, loaded by, or called from the product
 * library (libnntoolkitcore_b200.so).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it, and there only as the checker.
 *
 * Parity pin: every function here is checked against the UNMODIFIED reference compiled into
 * oracle/_ref/libnntk_ref.so (tests/test_oracle.py) and against the golden vectors that library generated
 * (tests/golden/, made by tests/golden/make_golden.py).  The reference itself ships no tests or golden
 * vectors (README.md:25), so that compiled library is the pin.
 *
 * All arithmetic is fp32 with the reference's operation order unless stated; the only deliberate difference is
 * the GEMV summation order (the reference goes through Eigen's vectorised GEMV, core/default_ops.cc:729-741;
 * here it is a plain k-ascending loop), which is why float parity is tolerance-based (1e-5), not bit-exact.
 */
#ifndef NNTK_ORACLE_H
#define NNTK_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* activation kinds (reference: layers/activation_default.c) */
enum { ORACLE_ACT_NONE = 0, ORACLE_ACT_IDENTITY = 1, ORACLE_ACT_SIGMOID = 2, ORACLE_ACT_TANH = 3,
       ORACLE_ACT_RELU = 4, ORACLE_ACT_SOFTMAX = 5 };

/* window kinds (reference: signal/window.c:27-54) */
enum { ORACLE_WIN_ONES = 0, ORACLE_WIN_HANN = 1, ORACLE_WIN_HAMMING = 2, ORACLE_WIN_PERIODIC_HANN = 3,
       ORACLE_WIN_PERIODIC_HAMMING = 4, ORACLE_WIN_BLACKMAN = 5 };

void oracle_window(int kind, float *w, int size);

/* forward/inverse complex DFT, kissfft algorithm (third_party/kissfft/kiss_fft.c:236-300,337-368) */
void oracle_dft(int nfft, int inverse, const float *in_re, const float *in_im, float *out_re, float *out_im);

/* mode 0 = magnitude (signal/spectrogram.c:29-38), 1 = PSD (:41-57, needs fs). scale_override<=0 -> computed. */
void oracle_spectrogram(int nfft, int window_size, int noverlap, int input_size, float norm_factor,
                        const float *window, int mode, int fs, float scale_override,
                        const float *input, float *output, int n_clips);

/* mel filter bank table [n_fft/2+1][n_mels] (signal/mel_filterbank.c:43-100), its apply (:114-116) and the log-mel chain
 * (signal/log_mel_spectrogram.c:33-38) */
void oracle_mel_weights(int n_mels, int n_fft, int sample_rate, float lower_hz, float upper_hz, float *weights);
void oracle_mel_apply(int n_mels, int n_bins, const float *weights, const float *spec, float *mel, int rows);
void oracle_log_mel_spectrogram(int nfft, int window_size, int noverlap, int input_size, float norm_factor,
                                const float *window, int mode, int fs, int n_mels, int sample_rate, float lower_hz,
                                float upper_hz, const float *input, float *output, int n_clips);

void oracle_activation(int kind, float relu_a, int vector_size, const float *in, float *out, int n);

void oracle_batch_norm(int channels, float eps, int count, const float *gamma, const float *beta,
                       const float *mean, const float *var, const float *in, float *out);

void oracle_conv1d(int cin, int cout, int k, int stride, int t_in, const float *W, const float *b,
                   const float *in, float *out, int batch);

void oracle_dense(int in_size, int out_size, const float *W, const float *b, int act, float relu_a,
                  const float *in, float *out, int rows);

/* Recurrent layers.  h (and c) are [H] state vectors, read as the initial state and overwritten with the final
 * state (the reference keeps them in the handle: layers/gru.c:197-201, lstm.c:257-265, rnn.c:236-241).
 * acts[] hold ORACLE_ACT_* kinds.  Output is [T,H] if return_sequences else [H]. */
void oracle_rnn(int in, int H, int T, int return_sequences, int v2, int act, float relu_a,
                const float *W, const float *U, const float *b_i, const float *b_h,
                const float *x, float *h, float *out);
void oracle_gru(int in, int H, int T, int return_sequences, int act_z, int act_h, int act_r,
                const float *W, const float *U, const float *b_i, const float *b_h,
                const float *x, float *h, float *out);
void oracle_lstm(int in, int H, int T, int return_sequences, int v2,
                 int act_i, int act_f, int act_g, int act_o, int act_out,
                 const float *W, const float *U, const float *b_i, const float *b_h,
                 const float *x, float *h, float *c, float *out);

/* bidirectional helpers (layers/bidirectional.c:10-85) */
void oracle_bd_reverse(const float *in, float *out, int vec, int T, int batch);
void oracle_bd_concat(const float *fwd, const float *bwd, float *out, int cols, int rows, int batch);
void oracle_bd_sum(const float *fwd, const float *bwd, float *out, long long n);

#ifdef __cplusplus
}
#endif
#endif
