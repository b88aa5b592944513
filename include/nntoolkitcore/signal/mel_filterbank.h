/*
 * nntoolkitcore/signal/mel_filterbank.h -- triangular mel filter bank.
 *
 * Drop-in for the reference header signal/mel_filterbank.h:14-38.  MelFilterBankConfig keeps the reference's field
 * order (by-value ABI).  The weight table [n_fft/2+1][n_mels] is built on the host with the reference's fp32 operation
 * order (signal/mel_filterbank.c:11-100: HTK mel scale 1127*ln(1+f/700), triangles in Hz, bin 0 forced to zero);
 * MelFilterBankApply() is the reference's one op_mat_mul (mel_filterbank.c:114-116) as a device GEMM.
 *
 * Pointers may be host (staged, complete on return) or device (ordered on the library stream), see b200_ext.h.
 */
#ifndef NNTK_B200_SIGNAL_MEL_FILTERBANK_H
#define NNTK_B200_SIGNAL_MEL_FILTERBANK_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int n_mels;
    int n_fft;
    int sample_rate;
    float lower_hz;
    float upper_hz;
} MelFilterBankConfig;

MelFilterBankConfig MelFilterBankConfigCreate(int n_mels, int n_fft, int sample_rate, float lower_hz, float upper_hz);

typedef struct MelFilterBankStruct *MelFilterBank;

MelFilterBank MelFilterBankCreate(MelFilterBankConfig config);

/* spectrogram [timesteps, n_fft/2+1] -> mel_spectrogram [timesteps, n_mels] */
void MelFilterBankApply(MelFilterBank filter_bank, const float *spectrogram, float *mel_spectrogram, int timesteps);

void MelFilterBankDestroy(MelFilterBank filter_bank);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_SIGNAL_MEL_FILTERBANK_H */
