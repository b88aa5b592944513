/* internal.h -- private definitions shared by the host C layer code (nntoolkitcore_b200/host/*.c).
 * The public handle types are opaque in the reference too (struct bodies live in its .c files), so the layouts here
 * are ours: every layer starts with the same header so that NNTKWeightsFreeze / NNTKResetState can take any handle. */
#ifndef NNTK_HOST_INTERNAL_H
#define NNTK_HOST_INTERNAL_H

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>
#include "nntk_b200_shim.h"
#include "nntoolkitcore/b200_ext.h"

/* Scoped hold of the process-wide recursive API lock (nntk_b200_shim.h): first statement of every public entry that enqueues
 * device work or touches the shared runtime state; released on every return path (gcc/clang cleanup attribute). */
static inline void nntk_api_unlock_cleanup_(int *unused) { (void)unused; nntk_cuda_api_unlock(); }
#define NNTK_API_GUARD() int nntk_api_guard_ __attribute__((cleanup(nntk_api_unlock_cleanup_), unused)) = (nntk_cuda_api_lock(), 0)

#define NNTK_LAYER_MAGIC 0x4e544b4cu /* 'NTKL' */
enum { NNTK_L_BATCHNORM = 1, NNTK_L_CONV1D, NNTK_L_DENSE, NNTK_L_TDD, NNTK_L_RNN, NNTK_L_GRU, NNTK_L_LSTM };

typedef struct nntk_layer_hdr {
    uint32_t magic;
    int kind;
    float *h_block;   /* host weight block the caller writes through XGetWeights (pinned when a device exists) */
    float *d_block;   /* device mirror, same layout */
    float *h_snap;    /* private pinned snapshot the H2D reads from: the caller may rewrite h_block right after an asynchronous
                         (device-pointer) Apply returns, so the copy engine must never read h_block itself */
    void *snap_ev;    /* recorded after the H2D of h_snap: the next restage waits for it before overwriting the snapshot */
    size_t n_floats;
    int frozen;       /* NNTKWeightsFreeze: skip the per-Apply restage */
    int staged;       /* device mirror valid */
    int (*after_stage)(struct nntk_layer_hdr *self); /* device-side repack hook, may be NULL */
} nntk_layer_hdr;

int nntk_layer_init(nntk_layer_hdr *h, int kind, size_t n_floats);
void nntk_layer_release(nntk_layer_hdr *h);
/* makes the device mirror current (H2D on the library stream unless frozen and already staged) */
int nntk_layer_stage(nntk_layer_hdr *h);

/* 0 both host, 1 both device, <0 mixed (error recorded) */
int nntk_ptr_kind(const void *in, const void *out, const char *who);

struct ActivationFunctionStruct {
    int kind;        /* NNTK_ACT_* */
    float a;         /* ReLU output scale */
    int input_size;  /* elements (softmax: number of vectors) */
    int vector_size; /* softmax */
    void *implementer;
    ActivationImplementerDestroy destroy_fn;
    ActivationFunctionImpl function;
    ActivationFunctionDerivative derivative, cached_derivative;
};

/* spectrogram.c: makes the device tables of `s` current and fills the kernel arguments (used by the log-mel front-end) */
int nntk_spectrogram_prepare(Spectrogram s, nntk_spectrogram_args *a);

/* mel_filterbank.c */
struct MelFilterBankStruct {
    MelFilterBankConfig config;
    float *weights;    /* host [n_fft/2+1][n_mels] */
    void *entries;     /* host per-lane work list of the fused kernel, [n_iter][32] x 8 bytes (NULL: not expressible) */
    int n_iter;
    float *d_weights;  /* device mirrors, uploaded at the first Apply */
    void *d_entries;
};
int nntk_mel_stage(MelFilterBank b);
int nntk_mel_gemm(MelFilterBank b, const float *d_in, float *d_out, long long rows);

static inline int nntk_act_kind(ActivationFunction f) { return f ? f->kind : NNTK_ACT_NONE; }
static inline float nntk_act_a(ActivationFunction f) { return f ? f->a : 1.0f; }

#endif
