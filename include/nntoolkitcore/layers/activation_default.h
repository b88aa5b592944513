/*
 * nntoolkitcore/layers/activation_default.h -- the five built-in activations.
 * Drop-in for the reference header layers/activation_default.h:19-27.  Arithmetic follows
 * layers/activation_default.c exactly (IEEE expf/tanhf/division, no fast-math):
 *   sigmoid  1.0f / (expf(-x) + 1.0f)                                   :28-33
 *   tanh     tanhf(x)                                                   :65-67
 *   identity copy (no-op in place)                                      :98-103
 *   ReLU(a)  fmaxf(x, 0) * a   (a scales the OUTPUT; skipped if a == 1) :123-129
 *   softmax  expf(x_j) / sum_j expf(x_j), NO max subtraction            :149-167  (overflows for x > ~88,
 *            exactly like the reference; keep logits bounded)
 */
#ifndef NNTK_B200_LAYERS_ACTIVATION_DEFAULT_H
#define NNTK_B200_LAYERS_ACTIVATION_DEFAULT_H

#include "activation.h"

#ifdef __cplusplus
extern "C" {
#endif

ActivationFunction ActivationFunctionCreateIdentity(int input_size);
ActivationFunction ActivationFunctionCreateSigmoid(int input_size);
ActivationFunction ActivationFunctionCreateTanh(int input_size);
ActivationFunction ActivationFunctionCreateReLU(int input_size, float a);
/* input_size = number of vectors, vector_size = length of each softmax vector */
ActivationFunction ActivationFunctionCreateSoftmax(int input_size, int vector_size);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_ACTIVATION_DEFAULT_H */
