/* conv_1d.c -- Conv1d inference (reference: layers/conv_1d.c:73-155; batched layout from Conv1dApplyTrainingBatch
 * :167-183).  Valid strided cross-correlation, channels-last in and out, weights W[Cout][Cin][K] | b[Cout] in one block
 * (private/weights_private.c:16-21).  On the device the weights are repacked to the implicit-GEMM operand
 * B[(k*Cin + ci)][co]; BatchNorm and a built-in activation can ride in the same kernel's epilogue. */
#include <stdlib.h>
#include "internal.h"

const float *nntk_batch_norm_device_params(BatchNorm f, int *channels, float *eps); /* batch_norm.c */

struct Conv1dStruct {
    nntk_layer_hdr hdr; /* must be first */
    Conv1dConfig config;
    ConvWeights weights;
    float *d_packed; /* [K*Cin][Cout] */
    float *d_packed_pad; /* [K*Cin_pad][Cout], only when Cin % 4 != 0 (tensor-core modes need 16-byte row pitch) */
    float *d_packed_bn;  /* fused front end: [K*pitch][Cout] with the input BatchNorm folded in | bias' [Cout] */
    int packed_bn_pitch;
};

/* Channel count the tensor-core path lays the input out with.  Cin % 4 == 0 needs no repack.  Otherwise the repack is paid
 * anyway, and a 128-byte row pitch (Cin rounded up to 32) keeps every TMA box row inside one 128-byte line: at Cin = 257 the
 * 1040-byte pitch of "round up to 4" ran the implicit GEMM at 456 TFLOP/s against ~700 with aligned rows.  Taken when it costs
 * at most 15 % more K. */
static int cin_pad(const Conv1dConfig *c) {
    const int cin = c->input_feature_channels, p4 = (cin + 3) & ~3, p32 = (cin + 31) & ~31;
    if (p4 == cin) return cin;
    return (long long)p32 * 100 <= (long long)p4 * 115 ? p32 : p4;
}

Conv1dConfig Conv1dConfigCreate(int input_feature_channels, int output_feature_channels, int kernel_size, int stride,
                                int inputSize) {
    Conv1dConfig c;
    c.input_feature_channels = input_feature_channels;
    c.output_feature_channels = output_feature_channels;
    c.kernel_size = kernel_size;
    c.stride = stride;
    c.input_size = inputSize;
    c.output_size = stride > 0 ? (inputSize - (kernel_size - stride)) / stride : 0; /* conv_1d.c:85 */
    return c;
}

static int conv_repack(nntk_layer_hdr *h) {
    Conv1d f = (Conv1d)h;
    const Conv1dConfig *c = &f->config;
    if (!f->d_packed) {
        void *d = NULL;
        int rc = nntk_cuda_malloc(&d, sizeof(float) * (size_t)c->output_feature_channels * c->input_feature_channels * c->kernel_size);
        if (rc) return rc;
        f->d_packed = (float *)d;
    }
    int rc = nntk_cuda_pack_conv_weights(h->d_block, f->d_packed, c->output_feature_channels, c->input_feature_channels,
                                         c->kernel_size);
    if (rc || cin_pad(c) == c->input_feature_channels) return rc;
    if (!f->d_packed_pad) {
        void *d = NULL;
        rc = nntk_cuda_malloc(&d, sizeof(float) * (size_t)c->output_feature_channels * cin_pad(c) * c->kernel_size);
        if (rc) return rc;
        f->d_packed_pad = (float *)d;
    }
    return nntk_cuda_pack_conv_weights_padded(h->d_block, f->d_packed_pad, c->output_feature_channels,
                                              c->input_feature_channels, cin_pad(c), c->kernel_size);
}

Conv1d Conv1dCreateForInference(Conv1dConfig config) {
    if (config.input_feature_channels < 1 || config.output_feature_channels < 1 || config.kernel_size < 1 ||
        config.stride < 1 || config.output_size < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "Conv1d: invalid config");
        return NULL;
    }
    Conv1d f = (Conv1d)calloc(1, sizeof(*f));
    if (!f) return NULL;
    const size_t w = (size_t)config.output_feature_channels * config.input_feature_channels * config.kernel_size;
    if (nntk_layer_init(&f->hdr, NNTK_L_CONV1D, w + (size_t)config.output_feature_channels)) { free(f); return NULL; }
    f->hdr.after_stage = conv_repack;
    f->config = config;
    f->weights.W = f->hdr.h_block;
    f->weights.b = f->hdr.h_block + w;
    return f;
}

ConvWeights *Conv1dGetWeights(Conv1d filter) { return &filter->weights; }

void Conv1dDestroy(Conv1d filter) {
    if (!filter) return;
    nntk_cuda_free(filter->d_packed);
    nntk_cuda_free(filter->d_packed_pad);
    nntk_cuda_free(filter->d_packed_bn);
    nntk_layer_release(&filter->hdr);
    free(filter);
}

typedef struct { Conv1d f; const float *d_bn; float eps; int act; float act_a; } conv_ctx;

static int conv_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    conv_ctx *cc = (conv_ctx *)ctx;
    const Conv1dConfig *c = &cc->f->config;
    const size_t w = (size_t)c->output_feature_channels * c->input_feature_channels * c->kernel_size;
    nntk_gemm_args g;
    g.d_A = d_in;
    g.a_slab_stride = (long long)c->input_size * c->input_feature_channels;
    g.lda = c->stride * c->input_feature_channels;
    g.batch = n;
    g.rows = c->output_size;
    g.N = c->output_feature_channels;
    g.K = c->kernel_size * c->input_feature_channels;
    g.d_B = cc->f->d_packed;
    g.d_bias = cc->f->hdr.d_block + w;
    g.d_bn = cc->d_bn;
    g.bn_eps = cc->eps;
    g.act = cc->act;
    g.act_a = cc->act_a;
    g.d_C = d_out;
    g.c_slab_stride = (long long)c->output_size * c->output_feature_channels;
    g.ldc = c->output_feature_channels;
    const int cp = cin_pad(c);
    if (cp == c->input_feature_channels || nntk_cuda_math_mode() == NNTK_MATH_FP32 || !cc->f->d_packed_pad)
        return nntk_cuda_gemm(&g);
    /* tensor-core modes, Cin % 4 != 0 (the C5 front-end: 257 frequency bins): TMA needs a 16-byte row pitch, so the input is
     * re-laid [T][Cin] -> [T][Cin_pad] (zeros; the weights carry matching zero rows) in chunks that fit the workspace */
    const size_t per_sample = sizeof(float) * (size_t)c->input_size * cp;
    void *ws = NULL; size_t got = 0;
    int rc = nntk_cuda_workspace(per_sample * (size_t)n, per_sample, &ws, &got);
    if (rc) return rc;
    const int chunk = (int)(got / per_sample < (size_t)n ? got / per_sample : (size_t)n);
    g.a_slab_stride = (long long)c->input_size * cp;
    g.lda = c->stride * cp;
    g.K = c->kernel_size * cp;
    g.d_B = cc->f->d_packed_pad;
    for (int b0 = 0; b0 < n; b0 += chunk) {
        const int nb = b0 + chunk <= n ? chunk : n - b0;
        rc = nntk_cuda_pad_channels(d_in + (size_t)b0 * c->input_size * c->input_feature_channels, (float *)ws,
                                    (long long)nb * c->input_size, c->input_feature_channels, cp);
        if (rc) return rc;
        g.d_A = (const float *)ws;
        g.batch = nb;
        g.d_C = d_out + (size_t)b0 * c->output_size * c->output_feature_channels;
        if ((rc = nntk_cuda_gemm(&g))) return rc;
    }
    return 0;
}

int Conv1dBNActApplyInferenceBatch(Conv1d filter, BatchNorm bn, ActivationFunction act, const float *input,
                                   float *output, int batch) {
    NNTK_API_GUARD();
    if (!filter || batch < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "Conv1dApplyInference: bad arguments");
        return NNTK_ERR_ARG;
    }
    int kind = nntk_ptr_kind(input, output, "Conv1dApplyInference");
    if (kind < 0) return kind;
    conv_ctx cc = {filter, NULL, 0.0f, NNTK_ACT_NONE, 1.0f};
    if (act) {
        if (act->kind == NNTK_ACT_SOFTMAX || act->kind == NNTK_ACT_CUSTOM) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED, "Conv1dBNAct: only identity/sigmoid/tanh/ReLU can be fused");
            return NNTK_ERR_UNSUPPORTED;
        }
        cc.act = act->kind;
        cc.act_a = act->a;
    }
    int rc = nntk_layer_stage(&filter->hdr);
    if (rc) return rc;
    if (bn) {
        int ch = 0;
        cc.d_bn = nntk_batch_norm_device_params(bn, &ch, &cc.eps);
        if (!cc.d_bn) return NNTK_ERR_CUDA;
        if (ch != filter->config.output_feature_channels) {
            nntk_cuda_set_error(NNTK_ERR_ARG, "Conv1dBNAct: BatchNorm has %d channels, Conv1d emits %d", ch,
                                filter->config.output_feature_channels);
            return NNTK_ERR_ARG;
        }
    }
    if (kind == 1) return conv_launch(&cc, input, output, 0, batch);
    const Conv1dConfig *c = &filter->config;
    return nntk_cuda_staged_run(input, output, sizeof(float) * (size_t)c->input_size * c->input_feature_channels,
                                sizeof(float) * (size_t)c->output_size * c->output_feature_channels, batch, conv_launch, &cc);
}

/* ---- fused front end: SpectrogramApply -> BatchNormApplyInference -> Conv1dApplyInference (-> BatchNorm -> activation) ------------
 * One K1 launch writes the magnitude rows at the pitch the Conv1d's implicit GEMM wants (pad columns zero), the input BatchNorm
 * is folded into the conv weights (it is affine per input channel), and the conv epilogue carries the optional output BatchNorm
 * and activation: the [clips, ntime, nfreq] spectrogram is written once and read once, where the three separate calls move it
 * five times (K1 write, BN read + write, channel pad read + write, GEMM read).  Reference chain: signal/spectrogram.c:113-135,
 * layers/batch_norm.c:140-189, layers/conv_1d.c:122-155. */
typedef struct {
    Spectrogram s; Conv1d f; const float *d_in_bn; float in_eps; const float *d_bn; float eps; int act; float act_a; int pitch;
} front_ctx;

static int front_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    front_ctx *fc = (front_ctx *)ctx;
    const Conv1dConfig *c = &fc->f->config;
    nntk_spectrogram_args sa;
    int rc = nntk_spectrogram_prepare(fc->s, &sa);
    if (rc) return rc;
    const int cp = fc->pitch;
    /* Tensor-core mode: the spectrogram rows are written and read as fp16 -- the same 10 explicit mantissa bits as the TF32 operand
     * they replace, half the bytes on both sides and twice the tensor rate (C5 front end 13.0 ms -> see DESIGN.md).  fp16 lacks
     * range, so both fp16 kernels raise a device flag when a value does not fit, and the fp32 / TF32 pair is enqueued right behind
     * them, gated on that flag: it costs two empty launches when nothing overflowed and produces the result when something did
     * (audio in int16 scale, huge BatchNorm gains).  NNTK_FRONT_F16=0 disables the fp16 rows. */
    static int want_f16 = -1;
    if (want_f16 < 0) { const char *e = getenv("NNTK_FRONT_F16"); want_f16 = e ? atoi(e) : 1; }
    int *flag = NULL;
    const int f16 = want_f16 && nntk_cuda_math_mode() == NNTK_MATH_TF32 && (cp & 7) == 0 && cp > c->input_feature_channels && cp <= 288 &&
                    (c->output_feature_channels % 32) == 0 &&
                    c->output_feature_channels > 32 && (flag = nntk_cuda_front_flag()) != NULL;
    const size_t rows_f32 = sizeof(float) * (size_t)c->input_size * cp, rows_f16 = rows_f32 / 2;
    const size_t per_clip = rows_f32 + (f16 ? rows_f16 : 0);
    void *ws = NULL; size_t got = 0;
    if ((rc = nntk_cuda_workspace(per_clip * (size_t)n, per_clip, &ws, &got))) return rc;
    const int chunk = (int)(got / per_clip < (size_t)n ? got / per_clip : (size_t)n);
    void *ws16 = (char *)ws + rows_f32 * (size_t)chunk;       /* [chunk] fp16 rows behind the [chunk] fp32 rows */
    nntk_gemm_args g;
    g.a_slab_stride = (long long)c->input_size * cp;
    g.lda = c->stride * cp;
    g.rows = c->output_size;
    g.N = c->output_feature_channels;
    g.K = c->kernel_size * cp;
    g.d_B = fc->f->d_packed_bn;
    g.d_bias = fc->f->d_packed_bn + (size_t)c->kernel_size * cp * c->output_feature_channels;
    g.d_bn = fc->d_bn;
    g.bn_eps = fc->eps;
    g.act = fc->act;
    g.act_a = fc->act_a;
    g.c_slab_stride = (long long)c->output_size * c->output_feature_channels;
    g.ldc = c->output_feature_channels;
    for (int b0 = 0; b0 < n; b0 += chunk) {
        const int nb = b0 + chunk <= n ? chunk : n - b0;
        g.batch = nb;
        g.d_C = d_out + (size_t)b0 * c->output_size * c->output_feature_channels;
        int gated = 0;
        if (f16) {
            if ((rc = nntk_cuda_memset(flag, 0, sizeof(int)))) return rc;
            rc = nntk_cuda_spectrogram_half(&sa, d_in + (size_t)b0 * sa.input_size, ws16, nb, cp, flag);
            if (rc < 0) return rc;
            if (rc == 0) {
                g.d_A = (const float *)ws16;
                rc = nntk_cuda_gemm_f16(&g, flag);
                if (rc < 0) return rc;
                gated = 1;      /* rc == 1: the fp16 rows were written but the GEMM is not eligible -- the ungated path below */
                if (rc == 1) gated = 0;
            }
        }
        if (gated) { nntk_cuda_spectrogram_gate(flag, 1); nntk_cuda_gemm_tc_gate(flag, 1); }
        rc = nntk_cuda_spectrogram_pitched(&sa, d_in + (size_t)b0 * sa.input_size, (float *)ws, nb, cp);
        if (rc == 0) {
            g.d_A = (const float *)ws;
            if (gated) {
                /* only the (gated) tensor-core kernel may run here: the FFMA kernel nntk_cuda_gemm falls back to knows no gate */
                rc = nntk_cuda_gemm_tf32(&g, 0);
                if (rc == 1) { nntk_cuda_set_error(NNTK_ERR_CUDA, "front end: the TF32 fallback GEMM is not eligible for a shape the fp16 GEMM took"); rc = NNTK_ERR_CUDA; }
            } else rc = nntk_cuda_gemm(&g);
        } else if (rc > 0) rc = NNTK_ERR_UNSUPPORTED;
        if (gated) { nntk_cuda_spectrogram_gate(NULL, 0); nntk_cuda_gemm_tc_gate(NULL, 0); }
        if (rc) return rc;
    }
    return 0;
}

int SpectrogramBNConv1dApplyBatch(Spectrogram spectrogram, BatchNorm input_bn, Conv1d filter, BatchNorm output_bn,
                                  ActivationFunction act, const float *input, float *output, int n_clips) {
    NNTK_API_GUARD();
    if (!spectrogram || !filter || n_clips < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "SpectrogramBNConv1dApplyBatch: bad arguments");
        return NNTK_ERR_ARG;
    }
    int kind = nntk_ptr_kind(input, output, "SpectrogramBNConv1dApplyBatch");
    if (kind < 0) return kind;
    const Conv1dConfig *c = &filter->config;
    const SpectrogramConfig sc = SpectrogramGetConfig(spectrogram);
    if (sc.nfreq != c->input_feature_channels || sc.ntime_series != c->input_size) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "SpectrogramBNConv1dApplyBatch: the spectrogram emits [%d, %d], the Conv1d expects [%d, %d]",
                            sc.ntime_series, sc.nfreq, c->input_size, c->input_feature_channels);
        return NNTK_ERR_ARG;
    }
    front_ctx fc = {spectrogram, filter, NULL, 0.0f, NULL, 0.0f, NNTK_ACT_NONE, 1.0f, 0};
    if (act) {
        if (act->kind == NNTK_ACT_SOFTMAX || act->kind == NNTK_ACT_CUSTOM) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED, "SpectrogramBNConv1dApplyBatch: only identity/sigmoid/tanh/ReLU can be fused");
            return NNTK_ERR_UNSUPPORTED;
        }
        fc.act = act->kind;
        fc.act_a = act->a;
    }
    int rc = nntk_layer_stage(&filter->hdr), ch = 0;
    if (rc) return rc;
    if (input_bn) {
        fc.d_in_bn = nntk_batch_norm_device_params(input_bn, &ch, &fc.in_eps);
        if (!fc.d_in_bn) return NNTK_ERR_CUDA;
        if (ch != c->input_feature_channels) {
            nntk_cuda_set_error(NNTK_ERR_ARG, "SpectrogramBNConv1dApplyBatch: input BatchNorm has %d channels, the spectrogram %d bins", ch, sc.nfreq);
            return NNTK_ERR_ARG;
        }
    }
    if (output_bn) {
        fc.d_bn = nntk_batch_norm_device_params(output_bn, &ch, &fc.eps);
        if (!fc.d_bn) return NNTK_ERR_CUDA;
        if (ch != c->output_feature_channels) {
            nntk_cuda_set_error(NNTK_ERR_ARG, "SpectrogramBNConv1dApplyBatch: output BatchNorm has %d channels, Conv1d emits %d", ch, c->output_feature_channels);
            return NNTK_ERR_ARG;
        }
    }
    /* row pitch: 128-byte aligned rows on the tensor-core paths when that costs <= 15 % more K (cin_pad), 16-byte aligned for FFMA */
    const int cin = c->input_feature_channels;
    fc.pitch = nntk_cuda_math_mode() == NNTK_MATH_FP32 ? ((cin + 3) & ~3) : cin_pad(c);
    /* (fp16 rows at the tighter 264-element pitch -- 21 instead of 23 k-blocks per tile -- measured SLOWER: 5.18 vs 4.82 ms for the
     * C5 conv; rows that are not 128-byte aligned cost the TMA im2col loads more than the 9 % fewer MMAs save) */
    if (sc.nfft != 512 || fc.pitch > 288 || !input_bn) {
        /* shapes without a pitched K1 (or nothing to fold): the three reference-shaped calls through a scratch spectrogram */
        const size_t spec_bytes = sizeof(float) * (size_t)n_clips * sc.ntime_series * sc.nfreq;
        if (kind != 1) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED, "SpectrogramBNConv1dApplyBatch: host buffers need nfft = 512 and an input BatchNorm");
            return NNTK_ERR_UNSUPPORTED;
        }
        void *tmp = NULL;
        if ((rc = nntk_cuda_malloc(&tmp, spec_bytes))) return rc;
        rc = SpectrogramApplyBatch(spectrogram, input, (float *)tmp, n_clips);
        if (!rc && input_bn) rc = BatchNormApplyInferenceBatch(input_bn, (float *)tmp, (float *)tmp, n_clips);
        if (!rc) rc = Conv1dBNActApplyInferenceBatch(filter, output_bn, act, (float *)tmp, output, n_clips);
        if (!rc) rc = nntk_cuda_sync();
        nntk_cuda_free(tmp);
        return rc;
    }
    /* conv weights with the input BatchNorm folded in: rebuilt at every call (both weight blocks may have been rewritten) */
    const size_t wn = (size_t)c->kernel_size * fc.pitch * c->output_feature_channels;
    if (!filter->d_packed_bn || filter->packed_bn_pitch != fc.pitch) {
        nntk_cuda_free(filter->d_packed_bn);
        void *d = NULL;
        filter->d_packed_bn = NULL;
        if ((rc = nntk_cuda_malloc(&d, sizeof(float) * (wn + (size_t)c->output_feature_channels)))) return rc;
        filter->d_packed_bn = (float *)d;
        filter->packed_bn_pitch = fc.pitch;
    }
    const size_t w = (size_t)c->output_feature_channels * cin * c->kernel_size;
    if ((rc = nntk_cuda_pack_conv_weights_bn(filter->hdr.d_block, filter->hdr.d_block + w, fc.d_in_bn, fc.in_eps, filter->d_packed_bn,
                                             filter->d_packed_bn + wn, c->output_feature_channels, cin, fc.pitch, c->kernel_size)))
        return rc;
    if (kind == 1) return front_launch(&fc, input, output, 0, n_clips);
    return nntk_cuda_staged_run(input, output, sizeof(float) * (size_t)sc.input_size,
                                sizeof(float) * (size_t)c->output_size * c->output_feature_channels, n_clips, front_launch, &fc);
}

int Conv1dApplyInferenceBatch(Conv1d filter, const float *input, float *output, int batch) {
    NNTK_API_GUARD();
    return Conv1dBNActApplyInferenceBatch(filter, NULL, NULL, input, output, batch);
}
int Conv1dApplyInference(Conv1d filter, const float *input, float *output) {
    NNTK_API_GUARD();
    return Conv1dBNActApplyInferenceBatch(filter, NULL, NULL, input, output, 1);
}
