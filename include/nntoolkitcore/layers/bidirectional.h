/*
 * nntoolkitcore/layers/bidirectional.h -- data-movement helpers around a forward + a backward recurrent layer
 * (SURVEY.md section 8 row f2).
 *
 * Drop-in for the inference half of the reference header layers/bidirectional.h:14-45 (the *_gradient / bd_accumulate_d_x
 * entries belong to training and are not part of this library):
 *   bd_reverse_input_batch     x [batch, T, in]   -> time-reversed copy           (bidirectional.c:10-24)
 *   bd_reverse_backward_batch  y [batch, T, out]  -> time-reversed copy           (bidirectional.c:26-32)
 *   bd_merge_concat            [batch, rows, out] x 2 -> [batch, rows, 2 out], forward features first (:39-55)
 *   bd_merge_sum               element-wise sum of the two results                (:75-85)
 * with rows = timesteps when config.return_sequences, else 1.  Each is one pass at 8 bytes per element moved.
 * Pointers: all host (staged through device memory, complete on return) or all device (library stream).
 * `buffer` of bd_merge_concat is the reference's transpose scratch; it is accepted and ignored (may be NULL).
 */
#ifndef NNTK_B200_LAYERS_BIDIRECTIONAL_H
#define NNTK_B200_LAYERS_BIDIRECTIONAL_H

#include "recurrent.h"

#ifdef __cplusplus
extern "C" {
#endif

void bd_reverse_input_batch(const float *input, float *output, RecurrentConfig config, int batch);

void bd_reverse_backward_batch(const float *input, float *output, RecurrentConfig config, int batch);

int bd_merge_concat_buffer_size(RecurrentConfig config);

void bd_merge_concat(const float *forward_result, const float *backward_result, float *output, RecurrentConfig config,
                     int batch, float *buffer);

void bd_merge_sum(const float *forward_result, const float *backward_result, float *output, RecurrentConfig config,
                  int batch);

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_BIDIRECTIONAL_H */
