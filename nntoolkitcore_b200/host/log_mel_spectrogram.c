/* log_mel_spectrogram.c -- host side of the LogMelSpectrogram API (reference: signal/log_mel_spectrogram.c;
 * SURVEY.md section 8 row f1).  For nfft == 512 the reference's chain SpectrogramApply -> MelFilterBankApply ->
 * op_vec_add_sc(1.5849e-13) -> op_vec_log (:33-38) is ONE kernel launch; other shapes run the same three stages through a
 * device workspace.  There is no CPU implementation here. */
#include <stdlib.h>
#include "internal.h"

#define NNTK_LOG_MEL_OFFSET 1.5849e-13f /* log_mel_spectrogram.c:36 (the double literal narrows to this float) */

struct LogMelSpectrogramStruct {
    MelFilterBankConfig mel_config;
    Spectrogram spectrogram; /* borrowed, like the reference (:23) */
    MelFilterBank bank;      /* owned (:28, :41) */
};

LogMelSpectrogram LogMelSpectrogramCreate(Spectrogram spectrogram, MelFilterBankConfig mel_config) {
    if (!spectrogram) { nntk_cuda_set_error(NNTK_ERR_ARG, "LogMelSpectrogramCreate: NULL spectrogram"); return NULL; }
    const SpectrogramConfig sc = SpectrogramGetConfig(spectrogram);
    if (mel_config.n_fft / 2 + 1 != sc.nfreq) {
        /* the reference multiplies [ntime, nfreq] by a [n_fft/2+1, n_mels] table regardless and reads out of bounds */
        nntk_cuda_set_error(NNTK_ERR_ARG, "LogMelSpectrogramCreate: mel n_fft %d does not match the spectrogram's nfft %d",
                            mel_config.n_fft, sc.nfft);
        return NULL;
    }
    LogMelSpectrogram f = (LogMelSpectrogram)calloc(1, sizeof(*f));
    if (!f) return NULL;
    f->mel_config = mel_config;
    f->spectrogram = spectrogram;
    f->bank = MelFilterBankCreate(mel_config);
    if (!f->bank) { free(f); return NULL; }
    return f;
}

void LogMelSpectrogramDestroy(LogMelSpectrogram f) {
    if (!f) return;
    MelFilterBankDestroy(f->bank);
    free(f);
}

static int logmel_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    LogMelSpectrogram f = (LogMelSpectrogram)ctx;
    nntk_spectrogram_args a;
    int rc = nntk_spectrogram_prepare(f->spectrogram, &a);
    if (rc) return rc;
    if (f->bank->d_entries) {
        nntk_mel_args m;
        m.n_mels = f->mel_config.n_mels;
        m.n_iter = f->bank->n_iter;
        m.d_entries = f->bank->d_entries;
        m.log_offset = NNTK_LOG_MEL_OFFSET;
        rc = nntk_cuda_log_mel_spectrogram(&a, &m, d_in, d_out, n);
        if (rc != 1) return rc;
    }
    /* unfused: clips in workspace-sized chunks through spectrogram -> mel GEMM -> log */
    const size_t spec_bytes = sizeof(float) * (size_t)a.ntime * a.nfreq;
    const size_t out_floats = (size_t)a.ntime * f->mel_config.n_mels;
    if (spec_bytes == 0 || n == 0) return 0;
    void *ws = NULL;
    size_t got = 0;
    if ((rc = nntk_cuda_workspace(spec_bytes * (size_t)n, spec_bytes, &ws, &got))) return rc;
    const int chunk = (int)(got / spec_bytes < (size_t)n ? got / spec_bytes : (size_t)n);
    for (int c0 = 0; c0 < n; c0 += chunk) {
        const int nc = n - c0 < chunk ? n - c0 : chunk;
        float *o = d_out + (size_t)c0 * out_floats;
        if ((rc = nntk_cuda_spectrogram(&a, d_in + (size_t)c0 * a.input_size, (float *)ws, nc))) return rc;
        if ((rc = nntk_mel_gemm(f->bank, (const float *)ws, o, (long long)nc * a.ntime))) return rc;
        if ((rc = nntk_cuda_log_offset(o, o, (long long)nc * (long long)out_floats, NNTK_LOG_MEL_OFFSET))) return rc;
    }
    return 0;
}

int LogMelSpectrogramApplyBatch(LogMelSpectrogram f, const float *input, float *output, int n_clips) {
    NNTK_API_GUARD();
    if (!f || n_clips < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "LogMelSpectrogramApplyBatch: bad arguments");
        return NNTK_ERR_ARG;
    }
    int kind = nntk_ptr_kind(input, output, "LogMelSpectrogramApply");
    if (kind < 0) return kind;
    int rc = nntk_mel_stage(f->bank);
    if (rc) return rc;
    if (kind == 1) return logmel_launch(f, input, output, 0, n_clips);
    const SpectrogramConfig c = SpectrogramGetConfig(f->spectrogram);
    return nntk_cuda_staged_run(input, output, sizeof(float) * (size_t)c.input_size,
                                sizeof(float) * (size_t)c.ntime_series * f->mel_config.n_mels, n_clips, logmel_launch, f);
}

/* void like the reference (:33): failures land in NNTKLastError() and on stderr */
void LogMelSpectrogramApply(LogMelSpectrogram f, const float *input, float *output) {
    NNTK_API_GUARD();
    (void)LogMelSpectrogramApplyBatch(f, input, output, 1);
}
