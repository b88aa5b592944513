/* activation.c -- ActivationFunction objects (reference: layers/activation.c, layers/activation_default.c).
 * The five built-ins carry a kind tag so that layers can fuse them into kernel epilogues; stand-alone Apply runs
 * the K5 kernels.  A user-supplied callback object (ActivationFunctionCreate) is the caller's own host code: Apply
 * on host buffers calls it as the reference does (activation.c:23-25); it cannot run inside a device layer, which
 * then fails with NNTK_ERR_UNSUPPORTED rather than falling back to a CPU path. */
#include <stdlib.h>
#include "internal.h"

static ActivationFunction act_new(int kind, int size, float a, int vector_size) {
    ActivationFunction f = (ActivationFunction)calloc(1, sizeof(*f));
    if (!f) return NULL;
    f->kind = kind;
    f->input_size = size;
    f->a = a;
    f->vector_size = vector_size;
    return f;
}

ActivationFunction ActivationFunctionCreate(int size, ActivationImplementerDestroy destroy_fn, void *implementer,
                                            ActivationFunctionImpl function, ActivationFunctionDerivative derivative,
                                            ActivationFunctionDerivative cached_derivative) {
    ActivationFunction f = act_new(NNTK_ACT_CUSTOM, size, 1.0f, 1);
    if (!f) return NULL;
    f->destroy_fn = destroy_fn;
    f->implementer = implementer;
    f->function = function;
    f->derivative = derivative;
    f->cached_derivative = cached_derivative;
    return f;
}

ActivationFunction ActivationFunctionCreateIdentity(int input_size) { return act_new(NNTK_ACT_IDENTITY, input_size, 1.0f, 1); }
ActivationFunction ActivationFunctionCreateSigmoid(int input_size) { return act_new(NNTK_ACT_SIGMOID, input_size, 1.0f, 1); }
ActivationFunction ActivationFunctionCreateTanh(int input_size) { return act_new(NNTK_ACT_TANH, input_size, 1.0f, 1); }
ActivationFunction ActivationFunctionCreateReLU(int input_size, float a) { return act_new(NNTK_ACT_RELU, input_size, a, 1); }
ActivationFunction ActivationFunctionCreateSoftmax(int input_size, int vector_size) {
    return act_new(NNTK_ACT_SOFTMAX, input_size, 1.0f, vector_size);
}

void ActivationFunctionDestroy(ActivationFunction filter) {
    if (!filter) return;
    if (filter->destroy_fn && filter->implementer) filter->destroy_fn(filter->implementer);
    free(filter);
}

typedef struct { ActivationFunction f; } act_ctx;
static int act_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    ActivationFunction f = ((act_ctx *)ctx)->f;
    return nntk_cuda_activation(f->kind, f->a, f->vector_size, d_in, d_out, (long long)n * f->input_size);
}

int ActivationFunctionApplyBatch(ActivationFunction filter, const float *input, float *output, int batch) {
    NNTK_API_GUARD();
    if (!filter || batch < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "ActivationFunctionApply: bad arguments");
        return NNTK_ERR_ARG;
    }
    if (filter->kind == NNTK_ACT_CUSTOM) {
        if (nntk_cuda_is_device_ptr(input) || nntk_cuda_is_device_ptr(output) || !filter->function) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED, "ActivationFunctionApply: a custom host callback cannot run on device buffers");
            return NNTK_ERR_UNSUPPORTED;
        }
        const size_t unit = (size_t)filter->input_size;
        for (int b = 0; b < batch; ++b)
            filter->function(filter->implementer, input + b * unit, output + b * unit, filter->input_size);
        return 0;
    }
    int kind = nntk_ptr_kind(input, output, "ActivationFunctionApply");
    if (kind < 0) return kind;
    act_ctx ctx = {filter};
    if (kind == 1) return act_launch(&ctx, input, output, 0, batch);
    const size_t unit = sizeof(float) * (size_t)filter->input_size * (filter->kind == NNTK_ACT_SOFTMAX ? filter->vector_size : 1);
    return nntk_cuda_staged_run(input, output, unit, unit, batch, act_launch, &ctx);
}

void ActivationFunctionApply(ActivationFunction filter, const float *input, float *output) {
    NNTK_API_GUARD();
    (void)ActivationFunctionApplyBatch(filter, input, output, 1);
}
