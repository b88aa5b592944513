/* dense.c -- Dense and TimeDistributedDense inference (reference: layers/dense.c:59-91,122-162 and
 * layers/time_distributed_dense.c:18-71).  y = act(x*W + b), W[in][out] row-major, block W | b.  Both layers are one
 * GEMM over all rows ([batch] or [batch*ts]) with bias + activation in the epilogue; softmax runs as a second pass. */
#include <stdlib.h>
#include "internal.h"

struct DenseStruct {
    nntk_layer_hdr hdr; /* must be first */
    DenseConfig config;
    DenseWeights weights;
};

DenseConfig DenseConfigCreate(int input_size, int output_size, ActivationFunction activation) {
    DenseConfig c;
    c.input_size = input_size;
    c.output_size = output_size;
    c.activation = activation;
    return c;
}

static int dense_setup(struct DenseStruct *f, DenseConfig config, int kind) {
    if (config.input_size < 1 || config.output_size < 1) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "Dense: invalid config");
        return NNTK_ERR_ARG;
    }
    const size_t w = (size_t)config.input_size * config.output_size;
    if (nntk_layer_init(&f->hdr, kind, w + (size_t)config.output_size)) return NNTK_ERR_ARG;
    f->config = config;
    f->weights.W = f->hdr.h_block;
    f->weights.b = f->hdr.h_block + w;
    return 0;
}

Dense DenseCreateForInference(DenseConfig config) {
    Dense f = (Dense)calloc(1, sizeof(*f));
    if (!f) return NULL;
    if (dense_setup(f, config, NNTK_L_DENSE)) { free(f); return NULL; }
    return f;
}
DenseWeights *DenseGetWeights(Dense filter) { return &filter->weights; }
void DenseDestroy(Dense filter) {
    if (!filter) return;
    nntk_layer_release(&filter->hdr);
    free(filter);
}

typedef struct { struct DenseStruct *f; int rows_per_unit; } dense_ctx;

/* rows = n units * rows_per_unit; [rows, in] -> [rows, out] */
static int dense_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    dense_ctx *dc = (dense_ctx *)ctx;
    const DenseConfig *c = &dc->f->config;
    const long long rows = (long long)n * dc->rows_per_unit;
    if (rows > 0x7fffffffLL) { nntk_cuda_set_error(NNTK_ERR_ARG, "Dense: too many rows"); return NNTK_ERR_ARG; }
    ActivationFunction act = c->activation;
    const int kind = nntk_act_kind(act);
    if (kind == NNTK_ACT_CUSTOM) {
        nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED, "Dense: a custom host-callback activation cannot run on the device");
        return NNTK_ERR_UNSUPPORTED;
    }
    nntk_gemm_args g;
    g.d_A = d_in; g.a_slab_stride = 0; g.lda = c->input_size; g.batch = 1; g.rows = (int)rows;
    g.N = c->output_size; g.K = c->input_size;
    g.d_B = dc->f->hdr.d_block;
    g.d_bias = dc->f->hdr.d_block + (size_t)c->input_size * c->output_size;
    g.d_bn = NULL; g.bn_eps = 0.0f;
    g.act = (kind == NNTK_ACT_SOFTMAX) ? NNTK_ACT_NONE : kind;
    g.act_a = nntk_act_a(act);
    g.d_C = d_out; g.c_slab_stride = 0; g.ldc = c->output_size;
    int rc;
    if (kind == NNTK_ACT_SOFTMAX && act->vector_size == c->output_size && c->output_size <= 32) {
        /* the whole row is one softmax vector and fits one epilogue thread: fused into the tensor-core GEMM when that is eligible */
        rc = nntk_cuda_gemm_row_softmax(&g);
        if (rc <= 0) return rc;
    }
    rc = nntk_cuda_gemm(&g);
    if (rc || kind != NNTK_ACT_SOFTMAX) return rc;
    if (act->vector_size < 1 || c->output_size % act->vector_size) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "Dense: softmax vector_size %d does not divide output_size %d", act->vector_size, c->output_size);
        return NNTK_ERR_ARG;
    }
    return nntk_cuda_activation(NNTK_ACT_SOFTMAX, 1.0f, act->vector_size, d_out, d_out,
                                rows * (c->output_size / act->vector_size));
}

static int dense_apply(struct DenseStruct *f, int rows_per_unit, const float *input, float *output, int batch, const char *who) {
    if (!f || batch < 0) { nntk_cuda_set_error(NNTK_ERR_ARG, "%s: bad arguments", who); return NNTK_ERR_ARG; }
    int kind = nntk_ptr_kind(input, output, who);
    if (kind < 0) return kind;
    int rc = nntk_layer_stage(&f->hdr);
    if (rc) return rc;
    dense_ctx dc = {f, rows_per_unit};
    if (kind == 1) return dense_launch(&dc, input, output, 0, batch);
    return nntk_cuda_staged_run(input, output, sizeof(float) * (size_t)rows_per_unit * f->config.input_size,
                                sizeof(float) * (size_t)rows_per_unit * f->config.output_size, batch, dense_launch, &dc);
}

int DenseApplyInferenceBatch(Dense filter, const float *input, float *output, int batch) {
    NNTK_API_GUARD();
    return dense_apply(filter, 1, input, output, batch, "DenseApplyInference");
}
int DenseApplyInference(Dense filter, const float *input, float *output) {
    NNTK_API_GUARD();
    return dense_apply(filter, 1, input, output, 1, "DenseApplyInference");
}

/* ---- TimeDistributedDense: the same Dense applied to each of ts rows (time_distributed_dense.c:52-58) ------------ */
struct TimeDistributedDenseStruct {
    struct DenseStruct dense; /* hdr first */
    TimeDistributedDenseConfig config;
};

TimeDistributedDenseConfig TimeDistributedDenseConfigCreate(int ts, DenseConfig dense) {
    TimeDistributedDenseConfig c;
    c.dense = dense;
    c.ts = ts;
    return c;
}
TimeDistributedDense TimeDistributedDenseCreateForInference(TimeDistributedDenseConfig config) {
    if (config.ts < 0) { nntk_cuda_set_error(NNTK_ERR_ARG, "TimeDistributedDense: invalid ts"); return NULL; }
    TimeDistributedDense f = (TimeDistributedDense)calloc(1, sizeof(*f));
    if (!f) return NULL;
    if (dense_setup(&f->dense, config.dense, NNTK_L_TDD)) { free(f); return NULL; }
    f->config = config;
    return f;
}
DenseWeights *TimeDistributedDenseGetWeights(TimeDistributedDense filter) { return &filter->dense.weights; }
void TimeDistributedDenseDestroy(TimeDistributedDense filter) {
    if (!filter) return;
    nntk_layer_release(&filter->dense.hdr);
    free(filter);
}
int TimeDistributedDenseApplyInferenceBatch(TimeDistributedDense filter, const float *input, float *output, int batch) {
    NNTK_API_GUARD();
    if (!filter) { nntk_cuda_set_error(NNTK_ERR_ARG, "TimeDistributedDenseApplyInference: NULL handle"); return NNTK_ERR_ARG; }
    return dense_apply(&filter->dense, filter->config.ts, input, output, batch, "TimeDistributedDenseApplyInference");
}
int TimeDistributedDenseApplyInference(TimeDistributedDense filter, const float *input, float *output) {
    NNTK_API_GUARD();
    return TimeDistributedDenseApplyInferenceBatch(filter, input, output, 1);
}
