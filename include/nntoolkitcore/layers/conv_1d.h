/*
 * nntoolkitcore/layers/conv_1d.h -- valid, strided 1-D cross-correlation, channels-last.
 * Drop-in for the inference side of the reference header layers/conv_1d.h:22-55.
 *
 * input [T, Cin] -> output [T_out, Cout], T_out = (T - (K - stride)) / stride (layers/conv_1d.c:85);
 * weights W[Cout][Cin][K] followed by b[Cout] in one block (conv_1d.c:32-36).  The reference transposes the input
 * and runs Cout*T_out*Cin scalar K-tap dots (conv_1d.c:122-147); here it is one implicit GEMM
 * [T_out, K*Cin] x [K*Cin, Cout] whose im2col row is the contiguous channels-last span in[x*stride : x*stride+K, :]
 * (csrc/gemm_*.cu), with bias (+BatchNorm +activation when fused through b200_ext.h) in the epilogue.
 */
#ifndef NNTK_B200_LAYERS_CONV_1D_H
#define NNTK_B200_LAYERS_CONV_1D_H

#include "shared.h"

#ifdef __cplusplus
extern "C" {
#endif

/* layers/conv_1d.h:22-29 (field order is ABI) */
typedef struct {
    int input_feature_channels;
    int output_feature_channels;
    int kernel_size;
    int stride;
    int input_size;
    int output_size;
} Conv1dConfig;

typedef DefaultWeights ConvWeights;
typedef DefaultWeightsSize ConvWeightsSize;
typedef struct Conv1dStruct *Conv1d;

/* conv_1d.c:77-87 */
Conv1dConfig Conv1dConfigCreate(int input_feature_channels, int output_feature_channels, int kernel_size,
                                int stride, int inputSize);
Conv1d Conv1dCreateForInference(Conv1dConfig config);                  /* conv_1d.c:98-102 */
ConvWeights *Conv1dGetWeights(Conv1d filter);                          /* conv_1d.c:73-75 */

/* one sample; returns 0, or <0 on failure                             -- conv_1d.c:149-155 */
int Conv1dApplyInference(Conv1d filter, const float *input, float *output);

void Conv1dDestroy(Conv1d filter);                                     /* conv_1d.c:110-120 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_CONV_1D_H */
