/* batch_norm.c -- BatchNorm inference (reference: layers/batch_norm.c:61-95,140-189).  Weight block layout
 * gamma | beta | moving_mean | moving_variance, one contiguous allocation (batch_norm.c:78-84). */
#include <stdlib.h>
#include "internal.h"

struct BatchNormFilterStruct {
    nntk_layer_hdr hdr; /* must be first */
    BatchNormConfig config;
    BatchNormWeights weights;
};

BatchNormConfig BatchNormConfigCreate(int feature_channels, float epsilon, int count) {
    BatchNormConfig c;
    c.feature_channels = feature_channels;
    c.epsilon = epsilon;
    c.count = count;
    return c;
}

BatchNorm BatchNormCreateForInference(BatchNormConfig config) {
    if (config.feature_channels < 1 || config.count < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "BatchNorm: invalid config");
        return NULL;
    }
    BatchNorm f = (BatchNorm)calloc(1, sizeof(*f));
    if (!f) return NULL;
    const size_t C = (size_t)config.feature_channels;
    if (nntk_layer_init(&f->hdr, NNTK_L_BATCHNORM, 4 * C)) { free(f); return NULL; }
    f->config = config;
    f->weights.gamma = f->hdr.h_block;
    f->weights.beta = f->hdr.h_block + C;
    f->weights.moving_mean = f->hdr.h_block + 2 * C;
    f->weights.moving_variance = f->hdr.h_block + 3 * C;
    return f;
}

BatchNormWeights *BatchNormGetWeights(BatchNorm filter) { return &filter->weights; }

void BatchNormDestroy(BatchNorm filter) {
    if (!filter) return;
    nntk_layer_release(&filter->hdr);
    free(filter);
}

/* used by conv_1d.c for the fused epilogue */
const float *nntk_batch_norm_device_params(BatchNorm f, int *channels, float *eps) {
    if (nntk_layer_stage(&f->hdr)) return NULL;
    *channels = f->config.feature_channels;
    *eps = f->config.epsilon;
    return f->hdr.d_block;
}

static int bn_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    BatchNorm f = (BatchNorm)ctx;
    return nntk_cuda_batch_norm(f->hdr.d_block, f->config.feature_channels, f->config.epsilon, d_in, d_out,
                                (long long)n * f->config.count);
}

int BatchNormApplyInferenceBatch(BatchNorm filter, const float *input, float *output, int batch) {
    NNTK_API_GUARD();
    if (!filter || batch < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "BatchNormApplyInference: bad arguments");
        return NNTK_ERR_ARG;
    }
    int kind = nntk_ptr_kind(input, output, "BatchNormApplyInference");
    if (kind < 0) return kind;
    int rc = nntk_layer_stage(&filter->hdr);
    if (rc) return rc;
    if (kind == 1) return bn_launch(filter, input, output, 0, batch);
    const size_t unit = sizeof(float) * (size_t)filter->config.count * filter->config.feature_channels;
    return nntk_cuda_staged_run(input, output, unit, unit, batch, bn_launch, filter);
}

int BatchNormApplyInference(BatchNorm filter, const float *input, float *output) {
    NNTK_API_GUARD();
    return BatchNormApplyInferenceBatch(filter, input, output, 1);
}
