/* spectrogram.c -- host side of the Spectrogram API (reference: signal/spectrogram.c).  Set-up work (window table,
 * scale factor, twiddle table) happens on the host exactly as in the reference; SpectrogramApply* forwards to the K1
 * kernel through nntk_cuda_spectrogram.  There is no CPU implementation of the transform here. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "internal.h"

struct SpectrogramStruct {
    SpectrogramConfig config;
    int mode;            /* 0 magnitude, 1 PSD */
    int fs;              /* PSD only */
    float *window;       /* host table, window_size floats */
    float scale_factor;
    float *d_window;     /* device mirrors */
    float *d_twiddle;
    int window_dirty;
};

SpectrogramConfig SpectrogramConfigCreate(int nfft, int window_size, int noverlap, int input_size,
                                          float fft_normalization_factor) {
    SpectrogramConfig c;
    c.nfft = nfft;
    c.window_size = window_size;
    c.noverlap = noverlap;
    c.input_size = input_size;
    c.step = window_size - noverlap;
    c.nfreq = nfft / 2 + 1;
    c.ntime_series = c.step > 0 ? (input_size - noverlap) / c.step : 0;
    c.fft_normalization_factor = fft_normalization_factor;
    return c;
}

/* scale = sum(w) (spectrogram.c:36-38) or fs * sum(w^2) (spectrogram.c:49-57); sequential fp32 sums like
 * op_vec_sum (core/default_ops.cc:638-644) */
static void recompute_scale(Spectrogram s) {
    float acc = 0.0f;
    const int n = s->config.window_size;
    if (s->mode == 0) {
        for (int i = 0; i < n; ++i) acc += s->window[i];
        s->scale_factor = acc;
    } else {
        for (int i = 0; i < n; ++i) { float sq = s->window[i] * s->window[i]; acc += sq; }
        s->scale_factor = acc * s->fs;
    }
}

static Spectrogram spectrogram_new(SpectrogramConfig config, int mode, int fs) {
    /* the reference would overflow a stack VLA here (spectrogram.c:120-124); refuse instead */
    if (config.nfft < 1 || config.window_size < 1 || config.window_size > config.nfft || config.step < 1 ||
        config.ntime_series < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "Spectrogram: invalid config (nfft %d, window %d, step %d)", config.nfft,
                            config.window_size, config.step);
        return NULL;
    }
    Spectrogram s = (Spectrogram)calloc(1, sizeof(*s));
    if (!s) return NULL;
    s->config = config;
    s->mode = mode;
    s->fs = fs;
    s->window = (float *)calloc((size_t)config.window_size, sizeof(float));
    SpectrogramSetWindowFunc(s, ones); /* default window, spectrogram.c:88,96 */
    return s;
}

Spectrogram SpectrogramCreateMagnitude(SpectrogramConfig config) { return spectrogram_new(config, 0, 0); }
Spectrogram SpectrogramCreatePSD(SpectrogramConfig config, int fs) { return spectrogram_new(config, 1, fs); }

void SpectrogramSetWindowFunc(Spectrogram filter, window_fn fn) {
    if (!filter || !fn) return;
    fn(filter->window, filter->config.window_size);
    recompute_scale(filter);
    filter->window_dirty = 1;
}
void SpectrogramSetScaleFactor(Spectrogram filter, float factor) { if (filter) filter->scale_factor = factor; }
SpectrogramConfig SpectrogramGetConfig(Spectrogram filter) { return filter->config; }

void SpectrogramDestroy(Spectrogram filter) {
    if (!filter) return;
    nntk_cuda_free(filter->d_window);
    nntk_cuda_free(filter->d_twiddle);
    free(filter->window);
    free(filter);
}

/* device tables: window (re-uploaded after SetWindowFunc) and the kissfft-style twiddles
 * tw[i] = ((float)cos(phi), (float)sin(phi)), phi = -2*pi*i/nfft in double (kiss_fft.c:357-363) */
static int spectrogram_stage(Spectrogram s) {
    int rc;
    const int nfft = s->config.nfft;
    if (!s->d_twiddle) {
        float *tw = (float *)malloc(sizeof(float) * 2 * (size_t)nfft);
        if (!tw) return NNTK_ERR_ARG;
        const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
        for (int i = 0; i < nfft; ++i) {
            double phase = -2 * pi * i / nfft;
            tw[2 * i] = (float)cos(phase);
            tw[2 * i + 1] = (float)sin(phase);
        }
        void *d = NULL;
        rc = nntk_cuda_malloc(&d, sizeof(float) * 2 * (size_t)nfft);
        if (!rc) rc = nntk_cuda_h2d(d, tw, sizeof(float) * 2 * (size_t)nfft);
        if (!rc) rc = nntk_cuda_sync(); /* tw is pageable and freed below */
        free(tw);
        if (rc) { nntk_cuda_free(d); return rc; }
        s->d_twiddle = (float *)d;
    }
    if (!s->d_window) {
        void *d = NULL;
        if ((rc = nntk_cuda_malloc(&d, sizeof(float) * (size_t)s->config.window_size))) return rc;
        s->d_window = (float *)d;
        s->window_dirty = 1;
    }
    if (s->window_dirty) {
        if ((rc = nntk_cuda_h2d(s->d_window, s->window, sizeof(float) * (size_t)s->config.window_size))) return rc;
        if ((rc = nntk_cuda_sync())) return rc; /* host table is pageable and may be rewritten by SetWindowFunc */
        s->window_dirty = 0;
    }
    return 0;
}

static void spectrogram_args(Spectrogram s, nntk_spectrogram_args *out);

int nntk_spectrogram_prepare(Spectrogram s, nntk_spectrogram_args *a) {
    int rc = spectrogram_stage(s);
    if (rc) return rc;
    spectrogram_args(s, a);
    return 0;
}

static int spectrogram_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    nntk_spectrogram_args a;
    spectrogram_args((Spectrogram)ctx, &a);
    return nntk_cuda_spectrogram(&a, d_in, d_out, n);
}

static void spectrogram_args(Spectrogram s, nntk_spectrogram_args *out) {
    nntk_spectrogram_args a;
    a.nfft = s->config.nfft;
    a.window_size = s->config.window_size;
    a.step = s->config.step;
    a.input_size = s->config.input_size;
    a.nfreq = s->config.nfreq;
    a.ntime = s->config.ntime_series;
    a.norm_factor = s->config.fft_normalization_factor;
    a.mode = s->mode;
    a.scale = s->scale_factor;
    a.d_window = s->d_window;
    a.d_twiddle = s->d_twiddle;
    *out = a;
}

int SpectrogramApplyBatch(Spectrogram filter, const float *input, float *output, int n_clips) {
    NNTK_API_GUARD();
    if (!filter || n_clips < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "SpectrogramApplyBatch: bad arguments");
        return NNTK_ERR_ARG;
    }
    int kind = nntk_ptr_kind(input, output, "SpectrogramApply");
    if (kind < 0) return kind;
    int rc = spectrogram_stage(filter);
    if (rc) return rc;
    if (kind == 1) return spectrogram_launch(filter, input, output, 0, n_clips);
    const SpectrogramConfig *c = &filter->config;
    return nntk_cuda_staged_run(input, output, sizeof(float) * (size_t)c->input_size,
                                sizeof(float) * (size_t)c->ntime_series * c->nfreq, n_clips, spectrogram_launch, filter);
}

/* void like the reference (spectrogram.c:113): failures land in NNTKLastError() and on stderr */
void SpectrogramApply(Spectrogram filter, const float *input, float *output) {
    NNTK_API_GUARD();
    (void)SpectrogramApplyBatch(filter, input, output, 1);
}
