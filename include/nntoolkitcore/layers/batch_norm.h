/*
 * nntoolkitcore/layers/batch_norm.h -- inference batch normalisation over [count, feature_channels].
 * Drop-in for the inference side of the reference header layers/batch_norm.h:19-66.
 *
 * y = ((x - mean) / sqrtf(var + eps)) * gamma + beta, six separately rounded fp32 operations in the reference's
 * order (layers/batch_norm.c:151-162).  One vectorised pass on the device (csrc/elementwise.cu) instead of the
 * reference's six passes per row; also available fused behind Conv1d (b200_ext.h).
 */
#ifndef NNTK_B200_LAYERS_BATCH_NORM_H
#define NNTK_B200_LAYERS_BATCH_NORM_H

#ifdef __cplusplus
extern "C" {
#endif

/* One contiguous block gamma|beta|moving_mean|moving_variance, C floats each (batch_norm.c:78-84). */
typedef struct {
    float *gamma;
    float *beta;
    float *moving_mean;
    float *moving_variance;
} BatchNormWeights;

/* layers/batch_norm.h:38-42 */
typedef struct {
    int feature_channels;
    float epsilon;
    int count;
} BatchNormConfig;

typedef struct BatchNormFilterStruct *BatchNorm;

BatchNormConfig BatchNormConfigCreate(int feature_channels, float epsilon, int count);   /* batch_norm.c:65-71 */
BatchNorm BatchNormCreateForInference(BatchNormConfig config);                           /* batch_norm.c:73-86 */
BatchNormWeights *BatchNormGetWeights(BatchNorm filter);                                 /* batch_norm.c:61-63 */

/* input/output: [count, feature_channels]; returns 0, or <0 on failure      -- batch_norm.c:166-189 */
int BatchNormApplyInference(BatchNorm filter, const float *input, float *output);

void BatchNormDestroy(BatchNorm filter);                                                 /* batch_norm.c:88-95 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_BATCH_NORM_H */
