/*
 * nntoolkitcore/signal/dft.h -- complex DFT on split (planar) buffers.
 *
 * Drop-in for the reference header signal/dft.h:15-44.  DFTPerform() transforms ONE nfft-point complex signal like the
 * reference (signal/dft.c:34-47: join -> kiss_fft -> split; forward = exp(-2 pi i k n / N), inverse = exp(+...),
 * neither normalised); DFTPerformBatch (b200_ext.h) transforms `batch` signals stored back to back.  nfft == 512 runs the
 * register-resident radix-8 transform of the spectrogram kernel (csrc/spectrogram_core.h), one warp per signal; other sizes
 * run a shared-memory radix-2 (powers of two) or direct (anything else) transform.  `config.complex` is carried but, as in
 * the reference's kissfft path, not consulted: the transform is always complex-to-complex.
 *
 * Pointers: the four planes may all be host pointers (staged through device memory, complete on return) or all device
 * pointers (ordered on the library stream).  In-place (output planes == input planes) is allowed, as in the reference.
 */
#ifndef NNTK_B200_SIGNAL_DFT_H
#define NNTK_B200_SIGNAL_DFT_H

#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct DFTSetupStruct *DFTSetup;

typedef struct {
    int nfft;
    bool forward;
    bool complex;
} DFTConfig;

typedef struct {
    float real;
    float imag;
} ComplexFloat;

typedef struct {
    float *real_p;
    float *imag_p;
} ComplexFloatSplit;

DFTConfig DFTConfigCreate(int nfft, bool forward, bool complex);

DFTSetup DFTSetupCreate(DFTConfig config);

void DFTPerform(DFTSetup setup, ComplexFloatSplit *input, ComplexFloatSplit *output);

void DFTSetupDestroy(DFTSetup setup);

/* interleaved <-> planar (signal/dft.c:60-84); host or device pointers */
void split_complex(const ComplexFloat *complex, ComplexFloatSplit *split, int size);

void join_complex_split(const ComplexFloatSplit *split, ComplexFloat *complex, int size);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_SIGNAL_DFT_H */
