/*
 * nntoolkitcore/layers/time_distributed_dense.h -- one Dense applied to each of `ts` rows.
 * Drop-in for the inference side of the reference header layers/time_distributed_dense.h:19-45.
 * The reference loops DenseApplyInference over rows (time_distributed_dense.c:52-58); here it is one GEMM
 * [ts, in] x [in, out] with the bias/activation epilogue.
 */
#ifndef NNTK_B200_LAYERS_TIME_DISTRIBUTED_DENSE_H
#define NNTK_B200_LAYERS_TIME_DISTRIBUTED_DENSE_H

#include "dense.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    DenseConfig dense;
    int ts;
} TimeDistributedDenseConfig;

typedef struct TimeDistributedDenseStruct *TimeDistributedDense;

/* note the argument order (ts first) -- time_distributed_dense.c:18-23 */
TimeDistributedDenseConfig TimeDistributedDenseConfigCreate(int ts, DenseConfig dense);
TimeDistributedDense TimeDistributedDenseCreateForInference(TimeDistributedDenseConfig config); /* :31-35 */
DenseWeights *TimeDistributedDenseGetWeights(TimeDistributedDense filter);                      /* :44-46 */

/* input [ts, in] -> output [ts, out]                                     -- :52-58 */
int TimeDistributedDenseApplyInference(TimeDistributedDense filter, const float *input, float *output);

void TimeDistributedDenseDestroy(TimeDistributedDense filter);                                  /* :68-71 */

#ifdef __cplusplus
}
#endif
#endif /* NNTK_B200_LAYERS_TIME_DISTRIBUTED_DENSE_H */
