/*
 * nntoolkitcore/signal/log_mel_spectrogram.h -- log(mel(spectrogram(x)) + 1.5849e-13).
 *
 * Drop-in for the reference header signal/log_mel_spectrogram.h:11-18 (which lacks the extern "C" guards added here).
 * The reference runs SpectrogramApply into a [ntime, nfreq] buffer, one op_mat_mul with the mel table, op_vec_add_sc and
 * op_vec_log (signal/log_mel_spectrogram.c:33-38).  Here, for nfft == 512, the whole chain is ONE kernel: the mel sums and
 * the log are an epilogue of the K1 spectrogram kernel (csrc/spectrogram.cu), so the spectrogram never reaches HBM.
 *
 * Ownership as in the reference: the filter borrows `spectrogram` (the caller destroys it, after the filter) and owns
 * its mel filter bank.  The mel config's n_fft must equal the spectrogram's nfft (the reference reads out of bounds
 * otherwise); Create returns NULL on a mismatch.
 */
#ifndef NNTK_B200_SIGNAL_LOG_MEL_SPECTROGRAM_H
#define NNTK_B200_SIGNAL_LOG_MEL_SPECTROGRAM_H

#include "spectrogram.h"
#include "mel_filterbank.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct LogMelSpectrogramStruct *LogMelSpectrogram;

LogMelSpectrogram LogMelSpectrogramCreate(Spectrogram spectrogram, MelFilterBankConfig mel_filter_bank_config);

/* input [input_size] -> output [ntime_series, n_mels] */
void LogMelSpectrogramApply(LogMelSpectrogram filter, const float *input, float *output);

void LogMelSpectrogramDestroy(LogMelSpectrogram filter);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_SIGNAL_LOG_MEL_SPECTROGRAM_H */
