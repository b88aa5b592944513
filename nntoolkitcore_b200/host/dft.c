/* dft.c -- host side of the DFT API (reference: signal/dft.c).  The twiddle table is built on
 * the host like kiss_fft_alloc (third_party/kissfft/kiss_fft.c:357-363: phase in double, cos/sin narrowed to float); the
 * transform itself only exists on the device (csrc/dft.cu). */
#include <math.h>
#include <stdlib.h>
#include "internal.h"

struct DFTSetupStruct {
    DFTConfig config;
    float *d_twiddle; /* forward table, [nfft][2] */
};

DFTConfig DFTConfigCreate(int nfft, bool forward, bool complex) {
    DFTConfig c;
    c.nfft = nfft;
    c.forward = forward;
    c.complex = complex;
    return c;
}

DFTSetup DFTSetupCreate(DFTConfig config) {
    if (config.nfft < 1) { nntk_cuda_set_error(NNTK_ERR_ARG, "DFTSetupCreate: nfft %d", config.nfft); return NULL; }
    DFTSetup s = (DFTSetup)calloc(1, sizeof(*s));
    if (s) s->config = config;
    return s;
}

void DFTSetupDestroy(DFTSetup setup) {
    if (!setup) return;
    nntk_cuda_free(setup->d_twiddle);
    free(setup);
}

static int dft_stage(DFTSetup s) {
    if (s->d_twiddle) return 0;
    const int nfft = s->config.nfft;
    float *tw = (float *)malloc(sizeof(float) * 2 * (size_t)nfft);
    if (!tw) return NNTK_ERR_ARG;
    const double pi = 3.141592653589793238462643383279502884197169399375105820974944;
    for (int i = 0; i < nfft; ++i) {
        const double phase = -2 * pi * i / nfft;
        tw[2 * i] = (float)cos(phase);
        tw[2 * i + 1] = (float)sin(phase);
    }
    void *d = NULL;
    int rc = nntk_cuda_malloc(&d, sizeof(float) * 2 * (size_t)nfft);
    if (!rc) rc = nntk_cuda_h2d(d, tw, sizeof(float) * 2 * (size_t)nfft);
    if (!rc) rc = nntk_cuda_sync();
    free(tw);
    if (rc) { nntk_cuda_free(d); return rc; }
    s->d_twiddle = (float *)d;
    return 0;
}

int DFTPerformBatch(DFTSetup setup, const ComplexFloatSplit *input, const ComplexFloatSplit *output, int batch) {
    NNTK_API_GUARD();
    if (!setup || !input || !output || batch < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "DFTPerform: bad arguments");
        return NNTK_ERR_ARG;
    }
    int kind = nntk_ptr_kind(input->real_p, output->real_p, "DFTPerform");
    if (kind < 0) return kind;
    int kind_i = nntk_ptr_kind(input->imag_p, output->imag_p, "DFTPerform");
    if (kind_i < 0) return kind_i;
    if (kind != kind_i) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "DFTPerform: the four planes must all be host or all be device pointers");
        return NNTK_ERR_ARG;
    }
    int rc = dft_stage(setup);
    if (rc) return rc;
    const int nfft = setup->config.nfft, inverse = !setup->config.forward;
    if (kind == 1)
        return nntk_cuda_dft(nfft, inverse, setup->d_twiddle, input->real_p, input->imag_p, output->real_p, output->imag_p, batch);
    /* host planes: chunks of signals through the shared device workspace, transformed in place there */
    const size_t sig = sizeof(float) * (size_t)nfft;
    void *ws = NULL;
    size_t got = 0;
    if (batch == 0) return 0;
    if ((rc = nntk_cuda_workspace(2 * sig * (size_t)batch, 2 * sig, &ws, &got))) return rc;
    size_t chunk = got / (2 * sig);
    if (chunk > (size_t)batch) chunk = (size_t)batch;
    float *d_re = (float *)ws, *d_im = d_re + chunk * (size_t)nfft;
    for (size_t b0 = 0; b0 < (size_t)batch; b0 += chunk) {
        const size_t nb = (size_t)batch - b0 < chunk ? (size_t)batch - b0 : chunk;
        if ((rc = nntk_cuda_h2d(d_re, input->real_p + b0 * nfft, nb * sig))) return rc;
        if ((rc = nntk_cuda_h2d(d_im, input->imag_p + b0 * nfft, nb * sig))) return rc;
        if ((rc = nntk_cuda_dft(nfft, inverse, setup->d_twiddle, d_re, d_im, d_re, d_im, (long long)nb))) return rc;
        if ((rc = nntk_cuda_d2h(output->real_p + b0 * nfft, d_re, nb * sig))) return rc;
        if ((rc = nntk_cuda_d2h(output->imag_p + b0 * nfft, d_im, nb * sig))) return rc;
        if ((rc = nntk_cuda_sync())) return rc; /* pageable host planes; the workspace is reused by the next chunk */
    }
    return 0;
}

/* void like the reference (dft.c:34): failures land in NNTKLastError() and on stderr */
void DFTPerform(DFTSetup setup, ComplexFloatSplit *input, ComplexFloatSplit *output) {
    NNTK_API_GUARD();
    (void)DFTPerformBatch(setup, input, output, 1);
}

/* dft.c:60-71 / :80-84: plain element loops on host buffers; strided device copies on device buffers */
void split_complex(const ComplexFloat *complex, ComplexFloatSplit *split, int size) {
    NNTK_API_GUARD();
    if (!complex || !split || size < 0) return;
    if (nntk_cuda_is_device_ptr(complex)) {
        (void)nntk_cuda_copy_strided(split->real_p, 1, &complex->real, 2, size);
        (void)nntk_cuda_copy_strided(split->imag_p, 1, &complex->imag, 2, size);
        return;
    }
    for (int i = 0; i < size; ++i) {
        split->real_p[i] = complex[i].real;
        split->imag_p[i] = complex[i].imag;
    }
}

void join_complex_split(const ComplexFloatSplit *split, ComplexFloat *complex, int size) {
    NNTK_API_GUARD();
    if (!complex || !split || size < 0) return;
    if (nntk_cuda_is_device_ptr(complex)) {
        (void)nntk_cuda_copy_strided(&complex->real, 2, split->real_p, 1, size);
        (void)nntk_cuda_copy_strided(&complex->imag, 2, split->imag_p, 1, size);
        return;
    }
    for (int i = 0; i < size; ++i) {
        complex[i].real = split->real_p[i];
        complex[i].imag = split->imag_p[i];
    }
}
