/*
 * nntk_oracle.c -- plain-C restatement of the NNToolkitCore inference forward path (see nntk_oracle.h).
 * TEST INFRASTRUCTURE ONLY: never linked into the product library.
 *
 * Every function cites the reference file:line (relative to /root/reference) whose arithmetic it follows.
 * Build: gcc -O2 -ffp-contract=off (no FMA contraction, no fast-math) so each fp32 operation rounds separately,
 * exactly like the reference's scalar x86 build (core/default_ops.cc with NEON==0, :13-15).
 */
#include "nntk_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------------------ windows */

/* signal/window.c:13-17: w[i] = alpha - (1 - alpha) * cos(2*pi*i / D); `alpha` is float, `1 - alpha` is a float
 * subtraction, the product with cos() and the final difference are double, the store rounds to float. */
static void hann_family(float *w, int size, int denom, float alpha) {
    for (int i = 0; i < size; ++i) {
        float one_minus = 1 - alpha;
        w[i] = (float)((double)alpha - (double)one_minus * cos(2 * M_PI * i / denom));
    }
}

void oracle_window(int kind, float *w, int size) {
    switch (kind) {
    case ORACLE_WIN_ONES: /* window.c:27-31 */
        for (int i = 0; i < size; ++i) w[i] = 1.0f;
        break;
    case ORACLE_WIN_HANN: hann_family(w, size, size - 1, 0.5f); break;             /* window.c:45-47 */
    case ORACLE_WIN_HAMMING: hann_family(w, size, size - 1, 0.54f); break;         /* window.c:41-43 */
    case ORACLE_WIN_PERIODIC_HANN: hann_family(w, size, size, 0.5f); break;        /* window.c:37-39 */
    case ORACLE_WIN_PERIODIC_HAMMING: hann_family(w, size, size, 0.54f); break;    /* window.c:33-35 */
    case ORACLE_WIN_BLACKMAN:                                                       /* window.c:49-54 */
        for (int i = 0; i < size; ++i) {
            /* `2.f * M_PI * i / size` is double arithmetic (M_PI is double) narrowed to float on assignment */
            float angle = (float)(2.f * M_PI * i / size);
            w[i] = (float)(.42f - .5 * cos(angle) + .08 * cos(2 * angle));
        }
        break;
    default: break;
    }
}

/* ------------------------------------------------------------------------------------------------ DFT */

typedef struct { float r, i; } cpx;

typedef struct {
    int nfft, inverse;
    int factors[64]; /* p1,m1,p2,m2,...  kiss_fft.c:303-328 */
    cpx *tw;
} dft_plan;

/* kiss_fft.c:307-328: powers of 4 first, then 2, then odd primes. */
static void plan_factor(int n, int *fac) {
    int p = 4;
    double floor_sqrt = floor(sqrt((double)n));
    do {
        while (n % p) {
            switch (p) {
            case 4: p = 2; break;
            case 2: p = 3; break;
            default: p += 2; break;
            }
            if (p > floor_sqrt) p = n;
        }
        n /= p;
        *fac++ = p;
        *fac++ = n;
    } while (n > 1);
}

static dft_plan *plan_create(int nfft, int inverse) {
    dft_plan *pl = (dft_plan *)malloc(sizeof(dft_plan));
    pl->nfft = nfft;
    pl->inverse = inverse;
    pl->tw = (cpx *)malloc(sizeof(cpx) * (size_t)nfft);
    /* kiss_fft.c:357-363 + _kiss_fft_guts.h:136-145: phase in double, cos/sin in double, narrowed to float */
    for (int i = 0; i < nfft; ++i) {
        const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
        double phase = -2 * pi * i / nfft;
        if (inverse) phase *= -1;
        pl->tw[i].r = (float)cos(phase);
        pl->tw[i].i = (float)sin(phase);
    }
    plan_factor(nfft, pl->factors);
    return pl;
}

static void plan_destroy(dft_plan *pl) { free(pl->tw); free(pl); }

/* complex multiply, _kiss_fft_guts.h C_MUL: (a.r*b.r - a.i*b.i, a.r*b.i + a.i*b.r) */
static inline cpx cmul(cpx a, cpx b) {
    cpx m;
    m.r = a.r * b.r - a.i * b.i;
    m.i = a.r * b.i + a.i * b.r;
    return m;
}
static inline cpx cadd(cpx a, cpx b) { cpx m = {a.r + b.r, a.i + b.i}; return m; }
static inline cpx csub(cpx a, cpx b) { cpx m = {a.r - b.r, a.i - b.i}; return m; }

/* kiss_fft.c:15-36 */
static void bfly2(cpx *F, int fstride, const dft_plan *pl, int m) {
    for (int j = 0; j < m; ++j) {
        cpx t = cmul(F[m + j], pl->tw[j * fstride]);
        F[m + j] = csub(F[j], t);
        F[j] = cadd(F[j], t);
    }
}

/* kiss_fft.c:38-84 */
static void bfly4(cpx *F, int fstride, const dft_plan *pl, int m) {
    for (int j = 0; j < m; ++j) {
        cpx s0 = cmul(F[j + m], pl->tw[j * fstride]);
        cpx s1 = cmul(F[j + 2 * m], pl->tw[j * fstride * 2]);
        cpx s2 = cmul(F[j + 3 * m], pl->tw[j * fstride * 3]);
        cpx s5 = csub(F[j], s1);
        F[j] = cadd(F[j], s1);
        cpx s3 = cadd(s0, s2);
        cpx s4 = csub(s0, s2);
        F[j + 2 * m] = csub(F[j], s3);
        F[j] = cadd(F[j], s3);
        if (pl->inverse) {
            F[j + m].r = s5.r - s4.i;      F[j + m].i = s5.i + s4.r;
            F[j + 3 * m].r = s5.r + s4.i;  F[j + 3 * m].i = s5.i - s4.r;
        } else {
            F[j + m].r = s5.r + s4.i;      F[j + m].i = s5.i - s4.r;
            F[j + 3 * m].r = s5.r - s4.i;  F[j + 3 * m].i = s5.i + s4.r;
        }
    }
}

/* kiss_fft.c:86-128 */
static void bfly3(cpx *F, int fstride, const dft_plan *pl, int m) {
    cpx epi3 = pl->tw[fstride * m];
    for (int j = 0; j < m; ++j) {
        cpx s1 = cmul(F[j + m], pl->tw[j * fstride]);
        cpx s2 = cmul(F[j + 2 * m], pl->tw[j * fstride * 2]);
        cpx s3 = cadd(s1, s2);
        cpx s0 = csub(s1, s2);
        F[j + m].r = F[j].r - s3.r * 0.5f;
        F[j + m].i = F[j].i - s3.i * 0.5f;
        s0.r *= epi3.i;
        s0.i *= epi3.i;
        F[j] = cadd(F[j], s3);
        F[j + 2 * m].r = F[j + m].r + s0.i;
        F[j + 2 * m].i = F[j + m].i - s0.r;
        F[j + m].r -= s0.i;
        F[j + m].i += s0.r;
    }
}

/* kiss_fft.c:130-190 */
static void bfly5(cpx *F, int fstride, const dft_plan *pl, int m) {
    cpx ya = pl->tw[fstride * m], yb = pl->tw[fstride * 2 * m];
    for (int u = 0; u < m; ++u) {
        cpx *F0 = F + u, *F1 = F0 + m, *F2 = F0 + 2 * m, *F3 = F0 + 3 * m, *F4 = F0 + 4 * m;
        cpx s0 = *F0;
        cpx s1 = cmul(*F1, pl->tw[u * fstride]);
        cpx s2 = cmul(*F2, pl->tw[2 * u * fstride]);
        cpx s3 = cmul(*F3, pl->tw[3 * u * fstride]);
        cpx s4 = cmul(*F4, pl->tw[4 * u * fstride]);
        cpx s7 = cadd(s1, s4), s10 = csub(s1, s4), s8 = cadd(s2, s3), s9 = csub(s2, s3);
        F0->r += s7.r + s8.r;
        F0->i += s7.i + s8.i;
        cpx s5, s6, s11, s12;
        s5.r = s0.r + s7.r * ya.r + s8.r * yb.r;
        s5.i = s0.i + s7.i * ya.r + s8.i * yb.r;
        s6.r = s10.i * ya.i + s9.i * yb.i;
        s6.i = -(s10.r * ya.i) - s9.r * yb.i;
        *F1 = csub(s5, s6);
        *F4 = cadd(s5, s6);
        s11.r = s0.r + s7.r * yb.r + s8.r * ya.r;
        s11.i = s0.i + s7.i * yb.r + s8.i * ya.r;
        s12.r = -(s10.i * yb.i) + s9.i * ya.i;
        s12.i = s10.r * yb.i - s9.r * ya.i;
        *F2 = cadd(s11, s12);
        *F3 = csub(s11, s12);
    }
}

/* kiss_fft.c:192-233 */
static void bfly_generic(cpx *F, int fstride, const dft_plan *pl, int m, int p) {
    cpx *scratch = (cpx *)malloc(sizeof(cpx) * (size_t)p);
    int N = pl->nfft;
    for (int u = 0; u < m; ++u) {
        int k = u;
        for (int q1 = 0; q1 < p; ++q1) { scratch[q1] = F[k]; k += m; }
        k = u;
        for (int q1 = 0; q1 < p; ++q1) {
            int twidx = 0;
            F[k] = scratch[0];
            for (int q = 1; q < p; ++q) {
                twidx += fstride * k;
                if (twidx >= N) twidx -= N;
                F[k] = cadd(F[k], cmul(scratch[q], pl->tw[twidx]));
            }
            k += m;
        }
    }
    free(scratch);
}

/* kiss_fft.c:236-300 (the non-OpenMP branch): decimation in time, p sub-DFTs of size m then a radix-p pass. */
static void dft_work(cpx *F, const cpx *f, int fstride, const int *factors, const dft_plan *pl) {
    int p = factors[0], m = factors[1];
    if (m == 1) {
        for (int j = 0; j < p; ++j) F[j] = f[(size_t)j * fstride];
    } else {
        for (int j = 0; j < p; ++j) dft_work(F + (size_t)j * m, f + (size_t)j * fstride, fstride * p, factors + 2, pl);
    }
    switch (p) {
    case 2: bfly2(F, fstride, pl, m); break;
    case 3: bfly3(F, fstride, pl, m); break;
    case 4: bfly4(F, fstride, pl, m); break;
    case 5: bfly5(F, fstride, pl, m); break;
    default: bfly_generic(F, fstride, pl, m, p); break;
    }
}

static void plan_run(const dft_plan *pl, const cpx *in, cpx *out) { dft_work(out, in, 1, pl->factors, pl); }

/* signal/dft.c:34-47: split -> interleaved, kiss_fft, interleaved -> split. */
void oracle_dft(int nfft, int inverse, const float *in_re, const float *in_im, float *out_re, float *out_im) {
    dft_plan *pl = plan_create(nfft, inverse);
    cpx *a = (cpx *)malloc(sizeof(cpx) * (size_t)nfft), *b = (cpx *)malloc(sizeof(cpx) * (size_t)nfft);
    for (int i = 0; i < nfft; ++i) { a[i].r = in_re[i]; a[i].i = in_im[i]; }
    plan_run(pl, a, b);
    for (int i = 0; i < nfft; ++i) { out_re[i] = b[i].r; out_im[i] = b[i].i; }
    free(a); free(b);
    plan_destroy(pl);
}

/* ------------------------------------------------------------------------------------------------ spectrogram */

/* signal/spectrogram.c:59-70 (sizes), :113-135 (per frame), :29-38 (magnitude), :41-57 (PSD).
 * core/default_ops.cc:638-644 (op_vec_sum: sequential fp32 sum). */
void oracle_spectrogram(int nfft, int window_size, int noverlap, int input_size, float norm_factor,
                        const float *window, int mode, int fs, float scale_override,
                        const float *input, float *output, int n_clips) {
    int step = window_size - noverlap;
    int nfreq = nfft / 2 + 1;
    int ntime = (input_size - noverlap) / step;
    float scale = 0.0f;
    if (scale_override > 0.0f) {
        scale = scale_override;                       /* SpectrogramSetScaleFactor, spectrogram.c:100-102 */
    } else if (mode == 0) {
        for (int i = 0; i < window_size; ++i) scale += window[i];                 /* spectrogram.c:36-38 */
    } else {
        float acc = 0.0f;
        for (int i = 0; i < window_size; ++i) acc += window[i] * window[i];       /* spectrogram.c:49-57 */
        scale = acc * fs;
    }
    dft_plan *pl = plan_create(nfft, 0);
    cpx *a = (cpx *)malloc(sizeof(cpx) * (size_t)nfft), *b = (cpx *)malloc(sizeof(cpx) * (size_t)nfft);
    for (int c = 0; c < n_clips; ++c) {
        const float *in = input + (size_t)c * input_size;
        float *out = output + (size_t)c * ntime * nfreq;
        for (int t = 0; t < ntime; ++t) {
            for (int i = 0; i < nfft; ++i) { a[i].r = 0.0f; a[i].i = 0.0f; }     /* :120-121 zero pad, imag=0 */
            for (int i = 0; i < window_size; ++i) a[i].r = window[i] * in[t * step + i];   /* :124 */
            plan_run(pl, a, b);                                                   /* :128 */
            if (norm_factor != 1.0f)                                              /* :130-131 */
                for (int i = 0; i < nfft; ++i) { b[i].r *= norm_factor; b[i].i *= norm_factor; }
            float *o = out + (size_t)t * nfreq;
            if (mode == 0) {                                                      /* :29-34 */
                for (int k = 0; k < nfreq; ++k) {
                    float m = b[k].r * b[k].r + b[k].i * b[k].i;                  /* default_ops.cc:666-670 */
                    m = sqrtf(m);
                    o[k] = m / scale;
                }
            } else {                                                              /* :41-47 */
                for (int k = 0; k < nfreq; ++k) o[k] = b[k].r * b[k].r + b[k].i * b[k].i;
                float two_over = 2.0f / scale;
                for (int k = 1; k < nfreq - 1; ++k) o[k] = o[k] * two_over;
                o[0] = o[0] / scale;
                o[nfreq - 1] = o[nfreq - 1] / scale;
            }
        }
    }
    free(a); free(b);
    plan_destroy(pl);
}

/* ------------------------------------------------------------------------------------------------ mel / log-mel */

/* signal/mel_filterbank.c:11-23: HTK mel scale as four fp32 vector passes each way */
static float oracle_hz_to_mel(float hz) { float m = hz / 700.0f; m = m + 1.0f; m = logf(m); return m * 1127.0f; }
static float oracle_mel_to_hz(float mel) { float h = mel / 1127.0f; h = expf(h); h = h + -1.0f; return h * 700.0f; }

/* weights [n_fft/2+1][n_mels]: band edges equally spaced in mel between lower_hz and upper_hz (mel_filterbank.c:43-54),
 * bin frequencies k*sr/n_fft (:57-65), triangle = max(min(falling, rising), 0) evaluated in Hz (:68-93), bin 0 zeroed
 * (:93), stored transposed (:96). */
void oracle_mel_weights(int n_mels, int n_fft, int sample_rate, float lower_hz, float upper_hz, float *weights) {
    int n_bins = n_fft / 2 + 1;
    float *band = (float *)malloc(sizeof(float) * (size_t)(n_mels + 2));
    float e0 = oracle_hz_to_mel(lower_hz), e1 = oracle_hz_to_mel(upper_hz);
    float step = (e1 - e0) / (float)(n_mels + 1);
    for (int i = 0; i < n_mels + 2; ++i) band[i] = oracle_mel_to_hz(e0 + (step * i));
    float fstep = (float)sample_rate / (float)n_fft;
    for (int i = 0; i < n_mels; ++i) {
        float lower = band[i], center = band[i + 1], upper = band[i + 2];
        for (int k = 0; k < n_bins; ++k) {
            float f = fstep * (float)k;
            float rising = (f + -1.0f * lower) / (center - lower);
            float falling = (-f + upper) / (upper - center);
            float w = fmaxf(fminf(falling, rising), 0.0f);
            weights[(size_t)k * n_mels + i] = k == 0 ? 0.0f : w;
        }
    }
    free(band);
}

/* MelFilterBankApply = op_mat_mul(spec, weights, mel, rows, n_mels, n_bins) (mel_filterbank.c:114-116); k-ascending sums
 * here, Eigen's blocked order in the reference -> tolerance, not bit-equality */
void oracle_mel_apply(int n_mels, int n_bins, const float *weights, const float *spec, float *mel, int rows) {
    for (int r = 0; r < rows; ++r)
        for (int m = 0; m < n_mels; ++m) {
            float acc = 0.0f;
            for (int k = 0; k < n_bins; ++k) acc += spec[(size_t)r * n_bins + k] * weights[(size_t)k * n_mels + m];
            mel[(size_t)r * n_mels + m] = acc;
        }
}

/* signal/log_mel_spectrogram.c:33-38: spectrogram -> mel -> + 1.5849e-13 -> logf */
void oracle_log_mel_spectrogram(int nfft, int window_size, int noverlap, int input_size, float norm_factor,
                                const float *window, int mode, int fs, int n_mels, int sample_rate, float lower_hz,
                                float upper_hz, const float *input, float *output, int n_clips) {
    int nfreq = nfft / 2 + 1, ntime = (input_size - noverlap) / (window_size - noverlap);
    float *w = (float *)malloc(sizeof(float) * (size_t)nfreq * n_mels);
    float *spec = (float *)malloc(sizeof(float) * (size_t)ntime * nfreq);
    oracle_mel_weights(n_mels, nfft, sample_rate, lower_hz, upper_hz, w);
    for (int c = 0; c < n_clips; ++c) {
        float *o = output + (size_t)c * ntime * n_mels;
        oracle_spectrogram(nfft, window_size, noverlap, input_size, norm_factor, window, mode, fs, 0.0f,
                           input + (size_t)c * input_size, spec, 1);
        oracle_mel_apply(n_mels, nfreq, w, spec, o, ntime);
        for (int i = 0; i < ntime * n_mels; ++i) { float v = o[i] + 1.5849e-13f; o[i] = logf(v); }
    }
    free(w); free(spec);
}

/* ------------------------------------------------------------------------------------------------ activations */

/* layers/activation_default.c: sigmoid :28-33 (neg, expf, +1, 1/x), tanh :65-67 (tanhf), identity :98-103,
 * ReLU :123-129 (max(x,0) then *a iff a != 1), softmax :149-167 (expf, sequential sum, divide; no max shift).
 * For softmax `n` counts vectors of `vector_size`; otherwise `n` counts elements. */
void oracle_activation(int kind, float relu_a, int vector_size, const float *in, float *out, int n) {
    switch (kind) {
    case ORACLE_ACT_NONE:
    case ORACLE_ACT_IDENTITY:
        if (in != out) memmove(out, in, sizeof(float) * (size_t)n);
        break;
    case ORACLE_ACT_SIGMOID:
        for (int i = 0; i < n; ++i) {
            float t = -in[i];
            t = expf(t);
            t = t + 1.0f;
            out[i] = 1.0f / t;
        }
        break;
    case ORACLE_ACT_TANH:
        for (int i = 0; i < n; ++i) out[i] = tanhf(in[i]);
        break;
    case ORACLE_ACT_RELU:
        for (int i = 0; i < n; ++i) {
            float t = in[i] > 0.0f ? in[i] : 0.0f;   /* op_vec_max_sc, default_ops.cc */
            if (relu_a != 1.0f) t = t * relu_a;
            out[i] = t;
        }
        break;
    case ORACLE_ACT_SOFTMAX:
        for (int v = 0; v < n; ++v) {
            const float *x = in + (size_t)v * vector_size;
            float *y = out + (size_t)v * vector_size;
            float sum = 0.0f;
            for (int j = 0; j < vector_size; ++j) y[j] = expf(x[j]);
            for (int j = 0; j < vector_size; ++j) sum += y[j];
            for (int j = 0; j < vector_size; ++j) y[j] = y[j] / sum;
        }
        break;
    default: break;
    }
}

/* ------------------------------------------------------------------------------------------------ batch norm */

/* layers/batch_norm.c:140-163 (six separately rounded fp32 ops), :166-189 (loop over `count` rows). */
void oracle_batch_norm(int channels, float eps, int count, const float *gamma, const float *beta,
                       const float *mean, const float *var, const float *in, float *out) {
    for (int r = 0; r < count; ++r) {
        const float *x = in + (size_t)r * channels;
        float *y = out + (size_t)r * channels;
        for (int c = 0; c < channels; ++c) {
            float x_mu = x[c] - mean[c];
            float var_eps = var[c] + eps;
            float sqrt_var = sqrtf(var_eps);
            float x_norm = x_mu / sqrt_var;
            float gx = x_norm * gamma[c];
            y[c] = gx + beta[c];
        }
    }
}

/* ------------------------------------------------------------------------------------------------ conv1d */

/* layers/conv_1d.c:122-147: valid strided cross-correlation, channels-last in/out, W[Cout][Cin][K];
 * acc = sum over cin of (k-ascending dot from 0, default_ops.cc:224-231), then + b[co].
 * T_out = (T - (K - s)) / s (conv_1d.c:85).  Batched layout [B,T,Cin] -> [B,T_out,Cout] (conv_1d.c:167-183). */
void oracle_conv1d(int cin, int cout, int k, int stride, int t_in, const float *W, const float *b,
                   const float *in, float *out, int batch) {
    int t_out = (t_in - (k - stride)) / stride;
    for (int n = 0; n < batch; ++n) {
        const float *x = in + (size_t)n * t_in * cin;
        float *y = out + (size_t)n * t_out * cout;
        for (int co = 0; co < cout; ++co) {
            for (int p = 0; p < t_out; ++p) {
                float acc = 0.0f;
                for (int ci = 0; ci < cin; ++ci) {
                    float dot = 0.0f;
                    for (int j = 0; j < k; ++j)
                        dot += x[(size_t)(p * stride + j) * cin + ci] * W[((size_t)co * cin + ci) * k + j];
                    acc += dot;
                }
                acc += b[co];
                y[(size_t)p * cout + co] = acc;
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------------ dense */

/* y = x * W  with W[K][N] row-major, k-ascending accumulation (reference: Eigen GEMV, default_ops.cc:735-736). */
static void gemv_row(const float *x, const float *W, float *y, int N, int K) {
    for (int n = 0; n < N; ++n) y[n] = 0.0f;
    for (int kk = 0; kk < K; ++kk) {
        float xv = x[kk];
        const float *w = W + (size_t)kk * N;
        for (int n = 0; n < N; ++n) y[n] += xv * w[n];
    }
}

/* layers/dense.c:122-142 (z = xW + b, a = act(z)); time_distributed_dense.c:52-58 loops this over rows.
 * For softmax the activation is applied per row over the whole out_size vector. */
void oracle_dense(int in_size, int out_size, const float *W, const float *b, int act, float relu_a,
                  const float *in, float *out, int rows) {
    for (int r = 0; r < rows; ++r) {
        float *y = out + (size_t)r * out_size;
        gemv_row(in + (size_t)r * in_size, W, y, out_size, in_size);
        for (int n = 0; n < out_size; ++n) y[n] = y[n] + b[n];
        if (act == ORACLE_ACT_SOFTMAX) oracle_activation(act, relu_a, out_size, y, y, 1);
        else oracle_activation(act, relu_a, 0, y, y, out_size);
    }
}

/* ------------------------------------------------------------------------------------------------ recurrent */

/* layers/rnn.c:144-166 (cell) with the evident-intent output addressing of rnn.c:259-281 (the inference loop at
 * :237 has an indexing bug for return_sequences). */
void oracle_rnn(int in, int H, int T, int return_sequences, int v2, int act, float relu_a,
                const float *W, const float *U, const float *b_i, const float *b_h,
                const float *x, float *h, float *out) {
    float *xw = (float *)malloc(sizeof(float) * 3 * (size_t)H);
    float *hu = xw + H, *gate = hu + H;
    for (int t = 0; t < T; ++t) {
        gemv_row(x + (size_t)t * in, W, xw, H, in);                 /* rnn.c:157 */
        for (int j = 0; j < H; ++j) xw[j] = xw[j] + b_i[j];         /* :158 */
        gemv_row(h, U, hu, H, H);                                   /* :160 */
        if (v2) for (int j = 0; j < H; ++j) hu[j] = hu[j] + b_h[j]; /* :161-163 */
        for (int j = 0; j < H; ++j) gate[j] = hu[j] + xw[j];        /* :164 */
        float *o = out + (return_sequences ? (size_t)t * H : 0);
        oracle_activation(act, relu_a, H, gate, o, act == ORACLE_ACT_SOFTMAX ? 1 : H);   /* :165 */
        memcpy(h, o, sizeof(float) * (size_t)H);
    }
    free(xw);
}

/* layers/gru.c:129-187 (cell), :189-204 (stateful loop). Gate column order [z, r, h~]. */
void oracle_gru(int in, int H, int T, int return_sequences, int act_z, int act_h, int act_r,
                const float *W, const float *U, const float *b_i, const float *b_h,
                const float *x, float *h, float *out) {
    float *xw = (float *)malloc(sizeof(float) * 12 * (size_t)H);
    float *hu = xw + 3 * H, *Zzr = hu + 3 * H, *z = Zzr + 2 * H, *r = z + H, *Zh = r + H, *ht = Zh + H;
    for (int t = 0; t < T; ++t) {
        gemv_row(x + (size_t)t * in, W, xw, 3 * H, in);                      /* gru.c:144 */
        for (int j = 0; j < 3 * H; ++j) xw[j] = xw[j] + b_i[j];              /* :147 */
        gemv_row(h, U, hu, 3 * H, H);                                        /* :150 */
        for (int j = 0; j < 3 * H; ++j) hu[j] = hu[j] + b_h[j];              /* :151 */
        for (int j = 0; j < 2 * H; ++j) Zzr[j] = xw[j] + hu[j];              /* :156 */
        oracle_activation(act_z, 1.0f, H, Zzr, z, act_z == ORACLE_ACT_SOFTMAX ? 1 : H);       /* :162 */
        oracle_activation(act_r, 1.0f, H, Zzr + H, r, act_r == ORACLE_ACT_SOFTMAX ? 1 : H);   /* :163 */
        for (int j = 0; j < H; ++j) Zh[j] = r[j] * hu[2 * H + j];            /* :169 */
        for (int j = 0; j < H; ++j) Zh[j] = Zh[j] + xw[2 * H + j];           /* :171 */
        oracle_activation(act_h, 1.0f, H, Zh, ht, act_h == ORACLE_ACT_SOFTMAX ? 1 : H);       /* :173 */
        float *o = out + (return_sequences ? (size_t)t * H : 0);
        for (int j = 0; j < H; ++j) {                                        /* :177-186 */
            float mz = -z[j];
            mz = mz + 1.0f;
            mz = mz * ht[j];
            float zh = z[j] * h[j];
            o[j] = mz + zh;
        }
        memcpy(h, o, sizeof(float) * (size_t)H);                             /* :201 */
    }
    free(xw);
}

/* layers/lstm.c:185-239 (cell), :241-268 (stateful loop). Gate column order [i, f, g, o]; b_h only if v2. */
void oracle_lstm(int in, int H, int T, int return_sequences, int v2,
                 int act_i, int act_f, int act_g, int act_o, int act_out,
                 const float *W, const float *U, const float *b_i, const float *b_h,
                 const float *x, float *h, float *c, float *out) {
    float *Z = (float *)malloc(sizeof(float) * 14 * (size_t)H);
    float *uH = Z + 4 * H, *gi = uH + 4 * H, *gf = gi + H, *gg = gf + H, *go = gg + H, *cn = go + H, *ct = cn + H;
    for (int t = 0; t < T; ++t) {
        gemv_row(x + (size_t)t * in, W, Z, 4 * H, in);                        /* lstm.c:201 */
        for (int j = 0; j < 4 * H; ++j) Z[j] = Z[j] + b_i[j];                 /* :203 */
        gemv_row(h, U, uH, 4 * H, H);                                         /* :206 */
        if (v2) for (int j = 0; j < 4 * H; ++j) uH[j] = uH[j] + b_h[j];       /* :207-209 */
        for (int j = 0; j < 4 * H; ++j) Z[j] = Z[j] + uH[j];                  /* :210 */
        oracle_activation(act_i, 1.0f, H, Z, gi, act_i == ORACLE_ACT_SOFTMAX ? 1 : H);           /* :214 */
        oracle_activation(act_f, 1.0f, H, Z + H, gf, act_f == ORACLE_ACT_SOFTMAX ? 1 : H);       /* :218 */
        oracle_activation(act_g, 1.0f, H, Z + 2 * H, gg, act_g == ORACLE_ACT_SOFTMAX ? 1 : H);   /* :222 */
        oracle_activation(act_o, 1.0f, H, Z + 3 * H, go, act_o == ORACLE_ACT_SOFTMAX ? 1 : H);   /* :226 */
        for (int j = 0; j < H; ++j) {                                         /* :229-234 */
            float ig = gi[j] * gg[j];
            float fc = gf[j] * c[j];
            cn[j] = fc + ig;
        }
        oracle_activation(act_out, 1.0f, H, cn, ct, act_out == ORACLE_ACT_SOFTMAX ? 1 : H);      /* :237 */
        float *o = out + (return_sequences ? (size_t)t * H : 0);
        for (int j = 0; j < H; ++j) o[j] = go[j] * ct[j];                     /* :238 */
        memcpy(h, o, sizeof(float) * (size_t)H);                              /* :264 */
        memcpy(c, cn, sizeof(float) * (size_t)H);                             /* :265 */
    }
    free(Z);
}

/* ------------------------------------------------------------------------------------------------ bidirectional helpers */

/* layers/bidirectional.c:10-32: per sequence, row t of the output is row T-1-t of the input */
void oracle_bd_reverse(const float *in, float *out, int vec, int T, int batch) {
    for (int b = 0; b < batch; ++b)
        for (int t = 0; t < T; ++t)
            memcpy(out + ((size_t)b * T + t) * vec, in + ((size_t)b * T + (T - 1 - t)) * vec, sizeof(float) * (size_t)vec);
}

/* layers/bidirectional.c:39-55: the three transposes amount to out[b][r] = fwd[b][r] | bwd[b][r] */
void oracle_bd_concat(const float *fwd, const float *bwd, float *out, int cols, int rows, int batch) {
    for (size_t r = 0; r < (size_t)batch * rows; ++r) {
        memcpy(out + r * 2 * cols, fwd + r * cols, sizeof(float) * (size_t)cols);
        memcpy(out + r * 2 * cols + cols, bwd + r * cols, sizeof(float) * (size_t)cols);
    }
}

/* layers/bidirectional.c:75-85: op_vec_add over batch * rows * cols elements */
void oracle_bd_sum(const float *fwd, const float *bwd, float *out, long long n) {
    for (long long i = 0; i < n; ++i) out[i] = fwd[i] + bwd[i];
}
