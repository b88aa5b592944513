/* bidirectional.c -- host side of the bidirectional helpers (reference: layers/bidirectional.c).
 * Device pointers go straight to the copy kernels; host pointers are staged through the shared device workspace. */
#include "internal.h"

typedef int (*bd_launch)(const float *d_a, const float *d_b, float *d_out, const void *ctx, long long units);

/* `units` independent pieces of a_floats (+ b_floats) input floats and out_floats output floats each */
static void bd_run(const float *a, const float *b, float *out, size_t a_floats, size_t b_floats, size_t out_floats, long long units,
                   bd_launch launch, const void *ctx, const char *who) {
    if (units < 0) { nntk_cuda_set_error(NNTK_ERR_ARG, "%s: negative batch", who); return; }
    int kind = nntk_ptr_kind(a, out, who);
    if (kind < 0) return;
    if (b && nntk_cuda_is_device_ptr(b) != kind) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "%s: all buffers must be host or all device pointers", who);
        return;
    }
    if (units == 0) return;
    if (kind == 1) { (void)launch(a, b, out, ctx, units); return; }
    const size_t per = (a_floats + b_floats + out_floats) * sizeof(float);
    void *ws = NULL;
    size_t got = 0;
    if (per == 0 || nntk_cuda_workspace(per * (size_t)units, per, &ws, &got)) return;
    long long chunk = (long long)(got / per);
    if (chunk > units) chunk = units;
    float *d_a = (float *)ws, *d_b = d_a + (size_t)chunk * a_floats, *d_o = d_b + (size_t)chunk * b_floats;
    for (long long u0 = 0; u0 < units; u0 += chunk) {
        const long long n = units - u0 < chunk ? units - u0 : chunk;
        if (nntk_cuda_h2d(d_a, a + (size_t)u0 * a_floats, (size_t)n * a_floats * sizeof(float))) return;
        if (b && nntk_cuda_h2d(d_b, b + (size_t)u0 * b_floats, (size_t)n * b_floats * sizeof(float))) return;
        if (launch(d_a, b ? d_b : NULL, d_o, ctx, n)) return;
        if (nntk_cuda_d2h(out + (size_t)u0 * out_floats, d_o, (size_t)n * out_floats * sizeof(float))) return;
        if (nntk_cuda_sync()) return;
    }
}

typedef struct { int vec, T; } rev_ctx;
static int rev_launch(const float *d_a, const float *d_b, float *d_out, const void *ctx, long long units) {
    (void)d_b;
    const rev_ctx *c = (const rev_ctx *)ctx;
    return nntk_cuda_bd_reverse(d_a, d_out, c->vec, c->T, units);
}
static void reverse_batch(const float *in, float *out, int vec, int T, int batch, const char *who) {
    if (vec < 1 || T < 0) { nntk_cuda_set_error(NNTK_ERR_ARG, "%s: bad config", who); return; }
    rev_ctx c = {vec, T};
    const size_t n = (size_t)vec * (size_t)T;
    bd_run(in, NULL, out, n, 0, n, batch, rev_launch, &c, who);
}
void bd_reverse_input_batch(const float *input, float *output, RecurrentConfig config, int batch) {
    NNTK_API_GUARD();
    reverse_batch(input, output, config.input_feature_channels, config.timesteps, batch, "bd_reverse_input_batch");
}
void bd_reverse_backward_batch(const float *input, float *output, RecurrentConfig config, int batch) {
    NNTK_API_GUARD();
    reverse_batch(input, output, config.output_feature_channels, config.timesteps, batch, "bd_reverse_backward_batch");
}

static int bd_rows(RecurrentConfig c) { return c.return_sequences ? c.timesteps : 1; }

int bd_merge_concat_buffer_size(RecurrentConfig config) { return 2 * bd_rows(config) * config.output_feature_channels; }

typedef struct { int cols, rows; } cat_ctx;
static int cat_launch(const float *d_a, const float *d_b, float *d_out, const void *ctx, long long units) {
    const cat_ctx *c = (const cat_ctx *)ctx;
    return nntk_cuda_bd_concat(d_a, d_b, d_out, c->cols, units * c->rows);
}
void bd_merge_concat(const float *forward_result, const float *backward_result, float *output, RecurrentConfig config,
                     int batch, float *buffer) {
    NNTK_API_GUARD();
    (void)buffer; /* the reference's transpose scratch (bidirectional.c:50-53) */
    cat_ctx c = {config.output_feature_channels, bd_rows(config)};
    if (c.cols < 1 || c.rows < 0 || !backward_result) { nntk_cuda_set_error(NNTK_ERR_ARG, "bd_merge_concat: bad arguments"); return; }
    const size_t n = (size_t)c.cols * (size_t)c.rows;
    bd_run(forward_result, backward_result, output, n, n, 2 * n, batch, cat_launch, &c, "bd_merge_concat");
}

static int sum_launch(const float *d_a, const float *d_b, float *d_out, const void *ctx, long long units) {
    return nntk_cuda_bd_sum(d_a, d_b, d_out, units * *(const long long *)ctx);
}
void bd_merge_sum(const float *forward_result, const float *backward_result, float *output, RecurrentConfig config, int batch) {
    NNTK_API_GUARD();
    long long n = (long long)config.output_feature_channels * bd_rows(config);
    if (n < 0 || !backward_result) { nntk_cuda_set_error(NNTK_ERR_ARG, "bd_merge_sum: bad arguments"); return; }
    bd_run(forward_result, backward_result, output, (size_t)n, (size_t)n, (size_t)n, batch, sum_launch, &n, "bd_merge_sum");
}
