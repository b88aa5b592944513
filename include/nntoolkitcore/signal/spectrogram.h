/*
 * nntoolkitcore/signal/spectrogram.h -- STFT front-end (framing, window, real DFT, magnitude / PSD).
 *
 * Drop-in for the reference header signal/spectrogram.h:20-44.  The by-value SpectrogramConfig keeps the
 * reference's field order (it is ABI).  SpectrogramApply() runs the batched shared-memory real-FFT kernel
 * (csrc/spectrogram.cu) instead of the reference's per-frame kissfft loop (signal/spectrogram.c:113-135,
 * signal/dft.c:34-47, third_party/kissfft/kiss_fft.c:236-300).
 *
 * Pointers: `input` / `output` may be host pointers (the call stages them through device memory and is complete
 * on return, like the reference) or device pointers (the call is ordered on the library stream, see
 * nntoolkitcore/b200_ext.h).
 */
#ifndef NNTK_B200_SIGNAL_SPECTROGRAM_H
#define NNTK_B200_SIGNAL_SPECTROGRAM_H

#include <stdbool.h>
#include "window.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Reference: signal/spectrogram.h:20-29.  Derived fields are filled by SpectrogramConfigCreate:
 *   step = window_size - noverlap, nfreq = nfft/2 + 1, ntime_series = (input_size - noverlap) / step. */
typedef struct {
    int nfft;
    int window_size;
    int noverlap;
    int step;
    int input_size;
    int nfreq;
    int ntime_series;
    float fft_normalization_factor;
} SpectrogramConfig;

typedef struct SpectrogramStruct *Spectrogram;

/* signal/spectrogram.c:59-70 */
SpectrogramConfig SpectrogramConfigCreate(int nfft, int window_size, int noverlap, int input_size,
                                          float fft_normalization_factor);

/* out[t][k] = sqrt(re^2 + im^2) / sum(w)                      -- signal/spectrogram.c:29-38,92-98.
 * Returns NULL when the configuration cannot be run (window_size > nfft, step <= 0, nfft < 2). */
Spectrogram SpectrogramCreateMagnitude(SpectrogramConfig config);

/* one-sided power spectral density, scale = fs * sum(w^2)     -- signal/spectrogram.c:41-57,83-90 */
Spectrogram SpectrogramCreatePSD(SpectrogramConfig config, int fs);

SpectrogramConfig SpectrogramGetConfig(Spectrogram filter);                 /* spectrogram.c:109-111 */

/* Evaluates fn on the host, recomputes the scale factor, uploads the table -- spectrogram.c:104-107.
 * The default window after Create is `ones`. */
void SpectrogramSetWindowFunc(Spectrogram filter, window_fn fn);

void SpectrogramSetScaleFactor(Spectrogram filter, float factor);           /* spectrogram.c:100-102 */

/* One clip: input[input_size] -> output[ntime_series][nfreq]               -- spectrogram.c:113-135 */
void SpectrogramApply(Spectrogram filter, const float *input, float *output);

void SpectrogramDestroy(Spectrogram filter);                                /* spectrogram.c:137-140 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_SIGNAL_SPECTROGRAM_H */
