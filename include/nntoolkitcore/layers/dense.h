/*
 * nntoolkitcore/layers/dense.h -- y = act(x W + b), W[in][out] row-major (Keras kernel layout).
 * Drop-in for the inference side of the reference header layers/dense.h:27-55.
 * Reference arithmetic: layers/dense.c:122-142 (Eigen GEMV + bias, then the activation object; NULL = none).
 */
#ifndef NNTK_B200_LAYERS_DENSE_H
#define NNTK_B200_LAYERS_DENSE_H

#include "activation.h"
#include "shared.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef DefaultWeights DenseWeights;
typedef DefaultWeightsSize DenseWeightsSize;

/* layers/dense.h:27-31.  The activation object stays owned by the caller (dense.c:80-86 never frees it). */
typedef struct {
    int input_size;
    int output_size;
    ActivationFunction activation;
} DenseConfig;

typedef struct DenseStruct *Dense;

DenseConfig DenseConfigCreate(int input_size, int output_size, ActivationFunction activation); /* dense.c:63-69 */
Dense DenseCreateForInference(DenseConfig config);                                             /* dense.c:73-79 */
DenseWeights *DenseGetWeights(Dense filter);                                                   /* dense.c:59-61 */

/* input [in] -> output [out]; 0 on success, <0 on failure                -- dense.c:135-142 */
int DenseApplyInference(Dense filter, const float *input, float *output);

void DenseDestroy(Dense filter);                                                               /* dense.c:81-87 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_DENSE_H */
