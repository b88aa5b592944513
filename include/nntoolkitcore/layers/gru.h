/*
 * nntoolkitcore/layers/gru.h -- GRU, Keras reset_after=True / cuDNN formulation.
 * Drop-in for the inference side of the reference header layers/gru.h:26-71 (cell: layers/gru.c:129-187).
 *
 *   xw = x W + b_i            hu = h U + b_h                       (W[in][3H], U[H][3H], columns [z, r, h~])
 *   z = act_z(xw_z + hu_z)    r = act_r(xw_r + hu_r)
 *   h~ = act_h(r * hu_h + xw_h)
 *   h' = ((-z) + 1) * h~ + z * h
 *
 * The input projection x W for all timesteps is hoisted into one GEMM; the recurrence runs in the persistent /
 * per-step recurrent kernels (csrc/recurrent.cu).  The handle is stateful: h is zero at Create and carried
 * across GRUApplyInference calls (gru.c:67,197-201).
 */
#ifndef NNTK_B200_LAYERS_GRU_H
#define NNTK_B200_LAYERS_GRU_H

#include <stdbool.h>
#include "activation.h"
#include "recurrent.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef RecurrentWeights GRUWeights;

/* layers/gru.h:26-30: struct order z, h, r */
typedef struct {
    ActivationFunction z_gate_activation;
    ActivationFunction h_gate_activation;
    ActivationFunction r_gate_activation;
} GRUActivations;

typedef struct {
    RecurrentConfig base;
    GRUActivations activations;
} GRUConfig;

typedef struct GRUStruct *GRU;

/* argument order z, h, r -- gru.c:220-230 */
GRUActivations GRUActivationsCreate(ActivationFunction z_gate_activation, ActivationFunction h_gate_activation,
                                    ActivationFunction r_gate_activation);
GRUActivations GRUActivationsCreateDefault(int size);        /* sigmoid, tanh, sigmoid -- gru.c:206-212 */
void GRUActivationsDestroy(GRUActivations activations);      /* gru.c:214-218 */

GRUConfig GRUConfigCreate(int input_feature_channels, int output_feature_channels, bool return_sequences,
                          int timesteps, GRUActivations activations);                 /* gru.c:12-19 */
GRU GRUCreateForInference(GRUConfig config);                                          /* gru.c:110-114 */
GRUWeights *GRUGetWeights(GRU filter);                                                /* gru.c:47-49 */

/* input [T, in] -> output [T, H] (return_sequences) or [H]; 0 on success             -- gru.c:189-204 */
int GRUApplyInference(GRU filter, const float *input, float *output);

void GRUDestroy(GRU filter);                                                          /* gru.c:116-126 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_GRU_H */
