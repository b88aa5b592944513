/*
 * nntoolkitcore/layers/recurrent.h -- config and weight-block types shared by RNN / GRU / LSTM.
 * Drop-in for the inference side of the reference header layers/recurrent.h:17-56.
 *
 * Weight layout (reference: layers/private/recurrent_private.c:29-36): ONE block W|U|b_i|b_h with
 *   W[in][G*H], U[H][G*H], b_i[G*H], b_h[G*H]   (G = 1 RNN, 3 GRU [z,r,h~], 4 LSTM [i,f,g,o]).
 * The block is pinned host memory the caller fills; the device keeps repacked copies.
 */
#ifndef NNTK_B200_LAYERS_RECURRENT_H
#define NNTK_B200_LAYERS_RECURRENT_H

#include <stdbool.h>
#include "shared.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int w;
    int u;
    int b_i;
    int b_h;
    int sum;
} RecurrentWeightsSize;

typedef struct {
    float *W;
    float *U;
    float *b_i;
    float *b_h;
} RecurrentWeights;

/* layers/recurrent.h:34-39 (field order is ABI) */
typedef struct {
    int input_feature_channels;
    int output_feature_channels;
    bool return_sequences;
    int timesteps;
} RecurrentConfig;

/* layers/recurrent.c:7-19 */
RecurrentConfig RecurrentConfigCreate(int input_feature_channels, int output_feature_channels,
                                      bool return_sequences, int timesteps);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_RECURRENT_H */
