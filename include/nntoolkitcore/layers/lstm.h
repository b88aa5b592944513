/*
 * nntoolkitcore/layers/lstm.h -- LSTM, gate column order [i, f, g, o].
 * Drop-in for the inference side of the reference header layers/lstm.h:24-75 (cell: layers/lstm.c:185-239).
 *
 *   Z = (x W + b_i) + (h U [+ b_h iff v2])                         (W[in][4H], U[H][4H])
 *   i = act_i(Z_i)  f = act_f(Z_f)  g = act_g(Z_g)  o = act_o(Z_o)
 *   c' = f * c + i * g          h' = o * act_out(c')
 *
 * The handle is stateful: h and c are zero at Create and carried across LSTMApplyInference calls
 * (lstm.c:160-161,257-265).
 */
#ifndef NNTK_B200_LAYERS_LSTM_H
#define NNTK_B200_LAYERS_LSTM_H

#include <stdbool.h>
#include "activation.h"
#include "recurrent.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef RecurrentWeights LSTMWeights;

/* layers/lstm.h:24-30: struct order candidate, input, forget, output_gate, output -- NOT the Create order */
typedef struct {
    ActivationFunction candidate_gate_activation;
    ActivationFunction input_gate_activation;
    ActivationFunction forget_gate_activation;
    ActivationFunction output_gate_activation;
    ActivationFunction output_activation;
} LSTMActivations;

/* argument order input, forget, candidate, output_gate, output -- lstm.c:17-25 */
LSTMActivations LSTMActivationsCreate(ActivationFunction input_gate_activation,
                                      ActivationFunction forget_gate_activation,
                                      ActivationFunction candidate_gate_activation,
                                      ActivationFunction output_gate_activation,
                                      ActivationFunction output_activation);
LSTMActivations LSTMActivationsCreateDefault(int size);      /* sigm, sigm, tanh, sigm, tanh -- lstm.c:110-118 */
void LSTMActivationsDestroy(LSTMActivations activations);    /* lstm.c:120-126 */

typedef struct {
    RecurrentConfig base;
    bool v2;
    LSTMActivations activations;
} LSTMConfig;

typedef struct LSTMStruct *LSTM;

LSTMConfig LSTMConfigCreate(int input_feature_channels, int output_feature_channels, bool return_sequences,
                            int timesteps, bool v2, LSTMActivations activations);     /* lstm.c:132-138 */
LSTM LSTMCreateForInference(LSTMConfig config);                                       /* lstm.c:154-158 */
LSTMWeights *LSTMGetWeights(LSTM filter);                                             /* lstm.c:128-130 */

/* input [T, in] -> output [T, H] (return_sequences) or [H]; 0 on success             -- lstm.c:241-268 */
int LSTMApplyInference(LSTM filter, const float *input, float *output);

void LSTMDestroy(LSTM filter);                                                        /* lstm.c:172-183 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_LSTM_H */
