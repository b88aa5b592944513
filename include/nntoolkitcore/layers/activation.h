/*
 * nntoolkitcore/layers/activation.h -- opaque activation objects.
 * Drop-in for the reference header layers/activation.h:19-30 (inference side).
 *
 * The reference activation is a bag of host callbacks (layers/activation.c:14-21).  Here every object also
 * carries a kind tag: the five built-ins of activation_default.h run as CUDA code (standalone kernels in
 * csrc/elementwise.cu, or fused into the GEMM / recurrent epilogues).  An object made with
 * ActivationFunctionCreate() from user callbacks is "custom": ActivationFunctionApply() on HOST pointers calls the
 * user's function as the reference would, but a custom activation cannot be fused into a device layer -- a
 * Dense/RNN/GRU/LSTM configured with one returns NNTK_ERR_UNSUPPORTED (-2) from Apply; there is no CPU fallback.
 */
#ifndef NNTK_B200_LAYERS_ACTIVATION_H
#define NNTK_B200_LAYERS_ACTIVATION_H

#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

/* (implementer, input, output, size) -- layers/activation.h:19 */
typedef void (*ActivationFunctionImpl)(void *, const float *, float *, int);
/* (implementer, z_or_a, d_out, output, size) -- layers/activation.h:20; accepted and stored, never called. */
typedef void (*ActivationFunctionDerivative)(void *, const float *, const float *, float *, int);
typedef void (*ActivationImplementerDestroy)(void *);

typedef struct ActivationFunctionStruct *ActivationFunction;

/* layers/activation.c:34-45 */
ActivationFunction ActivationFunctionCreate(int size, ActivationImplementerDestroy destroy_fn, void *implementer,
                                            ActivationFunctionImpl function,
                                            ActivationFunctionDerivative derivative,
                                            ActivationFunctionDerivative cached_derivative);

/* layers/activation.c:27-32 */
void ActivationFunctionDestroy(ActivationFunction filter);

/* Applies over the `size` given at creation (softmax: `size` vectors of vector_size) -- activation.c:23-25.
 * In-place (input == output) is allowed. */
void ActivationFunctionApply(ActivationFunction filter, const float *input, float *output);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_ACTIVATION_H */
