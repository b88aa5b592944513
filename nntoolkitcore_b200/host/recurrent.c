/* recurrent.c -- RNN / GRU / LSTM inference (reference: layers/recurrent.c:7-19, private/recurrent_private.c:29-41,
 * layers/rnn.c, layers/gru.c, layers/lstm.c).  One shared implementation: the three cells differ only in the gate
 * count G (1 / 3 / 4), the activation slots and the `v2` bias switch; the device work is nntk_cuda_recurrent.
 *
 * Weight block: W[in][G*H] | U[H][G*H] | b_i[G*H] | b_h[G*H], one allocation, like recurrent_weights_create.
 * State: the single-sequence XApplyInference calls are STATEFUL exactly like the reference -- h (and c) are zeroed at
 * Create only and carried across calls (gru.c:67,197-201; lstm.c:257-265; rnn.c:236-241).  The batched calls start
 * every sequence from zero like the reference's own batched forward (gru.c:259-260, lstm.c:438-439, rnn.c:259-260).
 * RNNApplyInference follows the reference's evident intent (output row t = h_t), not its indexing defect at
 * rnn.c:237-241. */
#include <stdlib.h>
#include <string.h>
#include "internal.h"

RecurrentConfig RecurrentConfigCreate(int input_feature_channels, int output_feature_channels, bool return_sequences,
                                      int timesteps) {
    RecurrentConfig c;
    c.input_feature_channels = input_feature_channels;
    c.output_feature_channels = output_feature_channels;
    c.return_sequences = return_sequences;
    c.timesteps = timesteps;
    return c;
}

typedef struct nntk_recurrent {
    nntk_layer_hdr hdr; /* must be first */
    int cell;           /* NNTK_CELL_* = gate count */
    RecurrentConfig base;
    int v2;
    ActivationFunction acts[5]; /* shim slot order: RNN [0]; GRU z, r, h~; LSTM i, f, g, o, out */
    RecurrentWeights weights;
    float *d_h, *d_c;   /* carried single-sequence state, [H] each */
} nntk_recurrent;

static nntk_recurrent *recurrent_new(int kind, int cell, RecurrentConfig base, int v2) {
    if (base.input_feature_channels < 1 || base.output_feature_channels < 1 || base.timesteps < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "recurrent layer: invalid config");
        return NULL;
    }
    nntk_recurrent *f = (nntk_recurrent *)calloc(1, sizeof(*f));
    if (!f) return NULL;
    const size_t in = (size_t)base.input_feature_channels, H = (size_t)base.output_feature_channels, G = (size_t)cell;
    const size_t w = in * G * H, u = H * G * H, b = G * H;
    if (nntk_layer_init(&f->hdr, kind, w + u + 2 * b)) { free(f); return NULL; }
    f->cell = cell;
    f->base = base;
    f->v2 = v2;
    f->weights.W = f->hdr.h_block;
    f->weights.U = f->hdr.h_block + w;
    f->weights.b_i = f->hdr.h_block + w + u;
    f->weights.b_h = f->hdr.h_block + w + u + b;
    return f;
}

static void recurrent_free(nntk_recurrent *f) {
    if (!f) return;
    nntk_cuda_free(f->d_h);
    nntk_cuda_free(f->d_c);
    nntk_layer_release(&f->hdr);
    free(f);
}

static int ensure_state(nntk_recurrent *f) {
    int rc;
    const size_t bytes = sizeof(float) * (size_t)f->base.output_feature_channels;
    if (!f->d_h) {
        void *d = NULL;
        if ((rc = nntk_cuda_malloc(&d, bytes))) return rc;
        if ((rc = nntk_cuda_memset(d, 0, bytes))) { nntk_cuda_free(d); return rc; }
        f->d_h = (float *)d;
    }
    if (f->cell == NNTK_CELL_LSTM && !f->d_c) {
        void *d = NULL;
        if ((rc = nntk_cuda_malloc(&d, bytes))) return rc;
        if ((rc = nntk_cuda_memset(d, 0, bytes))) { nntk_cuda_free(d); return rc; }
        f->d_c = (float *)d;
    }
    return 0;
}

int NNTKResetState(void *layer) {
    NNTK_API_GUARD();
    nntk_recurrent *f = (nntk_recurrent *)layer;
    if (!f || f->hdr.magic != NNTK_LAYER_MAGIC || f->hdr.kind < NNTK_L_RNN || f->hdr.kind > NNTK_L_LSTM) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "NNTKResetState: not a recurrent layer handle");
        return NNTK_ERR_ARG;
    }
    const size_t bytes = sizeof(float) * (size_t)f->base.output_feature_channels;
    int rc = 0;
    if (f->d_h) rc = nntk_cuda_memset(f->d_h, 0, bytes);
    if (!rc && f->d_c) rc = nntk_cuda_memset(f->d_c, 0, bytes);
    return rc;
}

/* ---- device launch ------------------------------------------------------------------------------------------------ */
typedef struct {
    nntk_recurrent *f;
    int mode;          /* 0: carried handle state (batch 1); 1: zero state per sequence; 2: caller state buffers */
    float *h_state, *c_state;   /* mode 2: [batch, H], host or device like the input */
    int state_on_device;
} rec_ctx;

static int rec_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    rec_ctx *rc_ = (rec_ctx *)ctx;
    nntk_recurrent *f = rc_->f;
    const int H = f->base.output_feature_channels, G = f->cell, T = f->base.timesteps;
    const size_t in = (size_t)f->base.input_feature_channels;
    const size_t state_bytes = sizeof(float) * (size_t)n * H;
    nntk_recurrent_args r;
    memset(&r, 0, sizeof(r));
    r.cell = G; r.in = (int)in; r.H = H; r.T = T; r.batch = n;
    r.return_sequences = f->base.return_sequences ? 1 : 0;
    r.v2 = f->v2;
    for (int i = 0; i < 5; ++i) {
        int k = nntk_act_kind(f->acts[i]);
        if (k == NNTK_ACT_CUSTOM || k == NNTK_ACT_SOFTMAX) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED, "recurrent layer: gate activation %d is not a fusable built-in", i);
            return NNTK_ERR_UNSUPPORTED;
        }
        r.act[i] = k;
        r.act_a[i] = nntk_act_a(f->acts[i]);
    }
    const size_t w = in * G * H, u = (size_t)H * G * H, b = (size_t)G * H;
    r.d_W = f->hdr.d_block; r.d_U = f->hdr.d_block + w; r.d_bi = f->hdr.d_block + w + u; r.d_bh = f->hdr.d_block + w + u + b;
    r.d_x = d_in; r.d_out = d_out;
    /* workspace = [state h | state c | projection + ping-pong]; state slots only for modes 1 and 2 */
    const size_t state_slots = (rc_->mode == 0) ? 0 : (G == NNTK_CELL_LSTM ? 2 : 1);
    const size_t state_total = ((state_slots * state_bytes + 255) / 256) * 256;
    size_t want = nntk_cuda_recurrent_workspace(&r) + state_total;
    /* the smallest workspace nntk_cuda_recurrent accepts: one sequence in its own accounting (projection + 6 H of ping-pong,
     * + 256), derived from the device side so the two can never disagree */
    nntk_recurrent_args one = r;
    one.batch = 1;
    const size_t per_seq = sizeof(float) * ((size_t)T * G * H + 6 * (size_t)H) + 256;
    const size_t min_ws = nntk_cuda_recurrent_workspace(&one) > per_seq ? nntk_cuda_recurrent_workspace(&one) : per_seq;
    void *ws = NULL; size_t got = 0;
    int rc = nntk_cuda_workspace(want, state_total + min_ws, &ws, &got);
    if (rc) return rc;
    float *d_hs = (float *)ws, *d_cs = (float *)((char *)ws + state_bytes);
    r.d_work = (float *)((char *)ws + state_total);
    r.work_bytes = got - state_total;
    if (rc_->mode == 0) {
        r.d_h = f->d_h; r.d_c = f->d_c;
    } else {
        r.d_h = d_hs; r.d_c = (G == NNTK_CELL_LSTM) ? d_cs : NULL;
        if (rc_->mode == 1) {
            if ((rc = nntk_cuda_memset(ws, 0, state_slots * state_bytes))) return rc;
        } else {
            float *hs = rc_->h_state + (size_t)first * H, *cs = rc_->c_state ? rc_->c_state + (size_t)first * H : NULL;
            if (rc_->state_on_device) {
                if ((rc = nntk_cuda_d2d(d_hs, hs, state_bytes))) return rc;
                if (r.d_c && (rc = nntk_cuda_d2d(d_cs, cs, state_bytes))) return rc;
            } else {
                if ((rc = nntk_cuda_h2d(d_hs, hs, state_bytes))) return rc;
                if (r.d_c && (rc = nntk_cuda_h2d(d_cs, cs, state_bytes))) return rc;
            }
        }
    }
    if ((rc = nntk_cuda_recurrent(&r))) return rc;
    if (rc_->mode == 2) {
        float *hs = rc_->h_state + (size_t)first * H, *cs = rc_->c_state ? rc_->c_state + (size_t)first * H : NULL;
        if (rc_->state_on_device) {
            if ((rc = nntk_cuda_d2d(hs, d_hs, state_bytes))) return rc;
            if (r.d_c && (rc = nntk_cuda_d2d(cs, d_cs, state_bytes))) return rc;
        } else {
            if ((rc = nntk_cuda_d2h(hs, d_hs, state_bytes))) return rc;
            if (r.d_c && (rc = nntk_cuda_d2h(cs, d_cs, state_bytes))) return rc;
            if ((rc = nntk_cuda_sync())) return rc;   /* workspace is reused by the next chunk */
        }
    }
    return 0;
}

static int rec_apply(nntk_recurrent *f, int mode, const float *input, float *output, int batch, float *h_state,
                     float *c_state, const char *who) {
    NNTK_API_GUARD();
    if (!f || batch < 0) { nntk_cuda_set_error(NNTK_ERR_ARG, "%s: bad arguments", who); return NNTK_ERR_ARG; }
    int kind = nntk_ptr_kind(input, output, who);
    if (kind < 0) return kind;
    rec_ctx ctx = {f, mode, h_state, c_state, kind};
    if (mode == 2) {
        if (!h_state || (f->cell == NNTK_CELL_LSTM && !c_state) || nntk_cuda_is_device_ptr(h_state) != kind ||
            (c_state && nntk_cuda_is_device_ptr(c_state) != kind)) {
            nntk_cuda_set_error(NNTK_ERR_ARG, "%s: state buffers must be non-NULL and of the same kind (host/device) as the input", who);
            return NNTK_ERR_ARG;
        }
    }
    int rc = nntk_layer_stage(&f->hdr);
    if (rc) return rc;
    if (mode == 0 && (rc = ensure_state(f))) return rc;
    if (kind == 1) return rec_launch(&ctx, input, output, 0, batch);
    const size_t T = (size_t)f->base.timesteps, H = (size_t)f->base.output_feature_channels;
    return nntk_cuda_staged_run(input, output, sizeof(float) * T * f->base.input_feature_channels,
                                sizeof(float) * (f->base.return_sequences ? T * H : H), batch, rec_launch, &ctx);
}

/* ---- RNN (layers/rnn.c) ------------------------------------------------------------------------------------------------ */
RNNConfig RNNConfigCreate(int input_feature_channels, int output_feature_channels, bool return_sequences, int timesteps,
                          bool v2, ActivationFunction activation) {
    RNNConfig c;
    c.base = RecurrentConfigCreate(input_feature_channels, output_feature_channels, return_sequences, timesteps);
    c.v2 = v2;
    c.activation = activation;
    return c;
}
RNN RNNCreateForInference(RNNConfig config) {
    nntk_recurrent *f = recurrent_new(NNTK_L_RNN, NNTK_CELL_RNN, config.base, config.v2 ? 1 : 0);
    if (f) f->acts[0] = config.activation;
    return (RNN)f;
}
RNNWeights *RNNGetWeights(RNN filter) { return &((nntk_recurrent *)filter)->weights; }
void RNNDestroy(RNN filter) { recurrent_free((nntk_recurrent *)filter); }
int RNNApplyInference(RNN filter, const float *input, float *output) {
    return rec_apply((nntk_recurrent *)filter, 0, input, output, 1, NULL, NULL, "RNNApplyInference");
}
int RNNApplyInferenceBatch(RNN filter, const float *input, float *output, int batch) {
    return rec_apply((nntk_recurrent *)filter, 1, input, output, batch, NULL, NULL, "RNNApplyInferenceBatch");
}
int RNNApplyInferenceBatchState(RNN filter, const float *input, float *output, int batch, float *h_state) {
    return rec_apply((nntk_recurrent *)filter, 2, input, output, batch, h_state, NULL, "RNNApplyInferenceBatchState");
}

/* ---- GRU (layers/gru.c) ------------------------------------------------------------------------------------------------ */
GRUActivations GRUActivationsCreate(ActivationFunction z_gate_activation, ActivationFunction h_gate_activation,
                                    ActivationFunction r_gate_activation) {
    GRUActivations a;
    a.z_gate_activation = z_gate_activation;
    a.h_gate_activation = h_gate_activation;
    a.r_gate_activation = r_gate_activation;
    return a;
}
GRUActivations GRUActivationsCreateDefault(int size) {
    return GRUActivationsCreate(ActivationFunctionCreateSigmoid(size), ActivationFunctionCreateTanh(size),
                                ActivationFunctionCreateSigmoid(size));
}
void GRUActivationsDestroy(GRUActivations a) {
    ActivationFunctionDestroy(a.z_gate_activation);
    ActivationFunctionDestroy(a.h_gate_activation);
    ActivationFunctionDestroy(a.r_gate_activation);
}
GRUConfig GRUConfigCreate(int input_feature_channels, int output_feature_channels, bool return_sequences, int timesteps,
                          GRUActivations activations) {
    GRUConfig c;
    c.base = RecurrentConfigCreate(input_feature_channels, output_feature_channels, return_sequences, timesteps);
    c.activations = activations;
    return c;
}
GRU GRUCreateForInference(GRUConfig config) {
    nntk_recurrent *f = recurrent_new(NNTK_L_GRU, NNTK_CELL_GRU, config.base, 1);
    if (f) {
        f->acts[0] = config.activations.z_gate_activation;
        f->acts[1] = config.activations.r_gate_activation;
        f->acts[2] = config.activations.h_gate_activation;
    }
    return (GRU)f;
}
GRUWeights *GRUGetWeights(GRU filter) { return &((nntk_recurrent *)filter)->weights; }
void GRUDestroy(GRU filter) { recurrent_free((nntk_recurrent *)filter); }
int GRUApplyInference(GRU filter, const float *input, float *output) {
    return rec_apply((nntk_recurrent *)filter, 0, input, output, 1, NULL, NULL, "GRUApplyInference");
}
int GRUApplyInferenceBatch(GRU filter, const float *input, float *output, int batch) {
    return rec_apply((nntk_recurrent *)filter, 1, input, output, batch, NULL, NULL, "GRUApplyInferenceBatch");
}
int GRUApplyInferenceBatchState(GRU filter, const float *input, float *output, int batch, float *h_state) {
    return rec_apply((nntk_recurrent *)filter, 2, input, output, batch, h_state, NULL, "GRUApplyInferenceBatchState");
}

/* ---- LSTM (layers/lstm.c) ---------------------------------------------------------------------------------------------- */
LSTMActivations LSTMActivationsCreate(ActivationFunction input_gate_activation, ActivationFunction forget_gate_activation,
                                      ActivationFunction candidate_gate_activation,
                                      ActivationFunction output_gate_activation, ActivationFunction output_activation) {
    LSTMActivations a;
    a.input_gate_activation = input_gate_activation;
    a.forget_gate_activation = forget_gate_activation;
    a.candidate_gate_activation = candidate_gate_activation;
    a.output_gate_activation = output_gate_activation;
    a.output_activation = output_activation;
    return a;
}
LSTMActivations LSTMActivationsCreateDefault(int size) {
    return LSTMActivationsCreate(ActivationFunctionCreateSigmoid(size), ActivationFunctionCreateSigmoid(size),
                                 ActivationFunctionCreateTanh(size), ActivationFunctionCreateSigmoid(size),
                                 ActivationFunctionCreateTanh(size));
}
void LSTMActivationsDestroy(LSTMActivations a) {
    ActivationFunctionDestroy(a.input_gate_activation);
    ActivationFunctionDestroy(a.forget_gate_activation);
    ActivationFunctionDestroy(a.candidate_gate_activation);
    ActivationFunctionDestroy(a.output_gate_activation);
    ActivationFunctionDestroy(a.output_activation);
}
LSTMConfig LSTMConfigCreate(int input_feature_channels, int output_feature_channels, bool return_sequences,
                            int timesteps, bool v2, LSTMActivations activations) {
    LSTMConfig c;
    c.base = RecurrentConfigCreate(input_feature_channels, output_feature_channels, return_sequences, timesteps);
    c.v2 = v2;
    c.activations = activations;
    return c;
}
LSTM LSTMCreateForInference(LSTMConfig config) {
    nntk_recurrent *f = recurrent_new(NNTK_L_LSTM, NNTK_CELL_LSTM, config.base, config.v2 ? 1 : 0);
    if (f) {
        f->acts[0] = config.activations.input_gate_activation;
        f->acts[1] = config.activations.forget_gate_activation;
        f->acts[2] = config.activations.candidate_gate_activation;
        f->acts[3] = config.activations.output_gate_activation;
        f->acts[4] = config.activations.output_activation;
    }
    return (LSTM)f;
}
LSTMWeights *LSTMGetWeights(LSTM filter) { return &((nntk_recurrent *)filter)->weights; }
void LSTMDestroy(LSTM filter) { recurrent_free((nntk_recurrent *)filter); }
int LSTMApplyInference(LSTM filter, const float *input, float *output) {
    return rec_apply((nntk_recurrent *)filter, 0, input, output, 1, NULL, NULL, "LSTMApplyInference");
}
int LSTMApplyInferenceBatch(LSTM filter, const float *input, float *output, int batch) {
    return rec_apply((nntk_recurrent *)filter, 1, input, output, batch, NULL, NULL, "LSTMApplyInferenceBatch");
}
int LSTMApplyInferenceBatchState(LSTM filter, const float *input, float *output, int batch, float *h_state,
                                 float *c_state) {
    return rec_apply((nntk_recurrent *)filter, 2, input, output, batch, h_state, c_state, "LSTMApplyInferenceBatchState");
}
