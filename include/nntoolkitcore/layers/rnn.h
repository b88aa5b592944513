/*
 * nntoolkitcore/layers/rnn.h -- simple recurrent layer h_t = act(x_t W + b_i + h_{t-1} U [+ b_h if v2]).
 * Drop-in for the inference side of the reference header layers/rnn.h:19-50 (cell: layers/rnn.c:144-166).
 *
 * Deviation (documented): the reference's RNNApplyInference writes step i to
 * output + i*i*H when return_sequences (rnn.c:237 vs :241).  This build implements the evident intent --
 * step i at output + i*H -- which is what the reference's own batched forward does (rnn.c:259-281).
 * The handle is stateful: h persists across Apply calls (rnn.c:236-241); see NNTKResetState in b200_ext.h.
 */
#ifndef NNTK_B200_LAYERS_RNN_H
#define NNTK_B200_LAYERS_RNN_H

#include "activation.h"
#include "recurrent.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef RecurrentWeights RNNWeights;

typedef struct {
    RecurrentConfig base;
    bool v2;
    ActivationFunction activation;
} RNNConfig;

typedef struct RNNStruct *RNN;

RNNConfig RNNConfigCreate(int input_feature_channels, int output_feature_channels, bool return_sequences,
                          int timesteps, bool v2, ActivationFunction activation);    /* rnn.c:49-62 */
RNN RNNCreateForInference(RNNConfig config);                                         /* rnn.c:111-115 */
RNNWeights *RNNGetWeights(RNN filter);                                               /* rnn.c:95-97 */

/* input [T, in] -> output [T, H] (return_sequences) or [H]; 0 on success            -- rnn.c:223-244 */
int RNNApplyInference(RNN filter, const float *input, float *output);

void RNNDestroy(RNN filter);                                                         /* rnn.c:133-142 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_RNN_H */
