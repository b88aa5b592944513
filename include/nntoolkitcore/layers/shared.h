/*
 * nntoolkitcore/layers/shared.h -- weight-block types shared by Conv1d / Dense / TimeDistributedDense.
 * Drop-in for the inference-side types of the reference header layers/shared.h:20-29.
 *
 * Layout contract (reference: layers/private/weights_private.c:16-21): W and b live in ONE contiguous block,
 * b == W + sizes.w, so a single memcpy of `sum` floats loads a layer.  In this build the block is pinned,
 * device-mapped host memory: the caller writes it through the pointers at any time after Create, and the
 * library restages it to the device on the next Apply (see NNTKWeightsFreeze in b200_ext.h to skip that).
 */
#ifndef NNTK_B200_LAYERS_SHARED_H
#define NNTK_B200_LAYERS_SHARED_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    int w;   /* floats in W */
    int b;   /* floats in b */
    int sum; /* w + b */
} DefaultWeightsSize;

typedef struct {
    float *W;
    float *b;
} DefaultWeights;

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_SHARED_H */
