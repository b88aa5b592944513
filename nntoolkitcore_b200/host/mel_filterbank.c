/* mel_filterbank.c -- host side of the MelFilterBank API (reference: signal/mel_filterbank.c).
 * The weight table is built here with the reference's fp32 operation order; the apply is a device GEMM.  For the fused
 * log-mel kernel the table is also flattened into a per-lane work list (see nntk_mel_args in nntk_b200_shim.h). */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "internal.h"

/* mel_filterbank.c:11-23: four separate fp32 vector passes each way (divide, add, logf / expf, multiply) */
static float hertz_to_mel(float hz) {
    float m = hz / 700.0f;
    m = m + 1.0f;
    m = logf(m);
    return m * 1127.0f;
}
static float mel_to_hertz(float mel) {
    float hz = mel / 1127.0f;
    hz = expf(hz);
    hz = hz + -1.0f;
    return hz * 700.0f;
}

MelFilterBankConfig MelFilterBankConfigCreate(int n_mels, int n_fft, int sample_rate, float lower_hz, float upper_hz) {
    MelFilterBankConfig c;
    c.n_mels = n_mels;
    c.n_fft = n_fft;
    c.sample_rate = sample_rate;
    c.lower_hz = lower_hz;
    c.upper_hz = upper_hz;
    return c;
}

/* weights[k * n_mels + i], k = 0..n_fft/2: mel_filterbank.c:43-100 */
static void build_weights(const MelFilterBankConfig *c, float *weights) {
    const int n_mels = c->n_mels, n_bins = c->n_fft / 2 + 1;
    float *band = (float *)malloc(sizeof(float) * (size_t)(n_mels + 2));
    float *bin_hz = (float *)malloc(sizeof(float) * (size_t)n_bins);
    /* band edges: equally spaced on the mel axis, converted back to Hz (:43-54) */
    const float lo = hertz_to_mel(c->lower_hz), hi = hertz_to_mel(c->upper_hz);
    const float mstep = (hi - lo) / (float)(n_mels + 1);
    for (int i = 0; i < n_mels + 2; ++i) band[i] = mel_to_hertz(lo + (mstep * i));
    /* bin centre frequencies (:57-65) */
    const float fstep = (float)c->sample_rate / (float)c->n_fft;
    for (int k = 0; k < n_bins; ++k) bin_hz[k] = fstep * (float)k;
    for (int i = 0; i < n_mels; ++i) {
        const float lower = band[i], center = band[i + 1], upper = band[i + 2];
        const float neg_lower = -1.0f * lower, rise = center - lower, fall = upper - center;
        for (int k = 0; k < n_bins; ++k) {
            float ls = bin_hz[k] + neg_lower;      /* :79-80 */
            ls = ls / rise;
            float us = -bin_hz[k];                 /* :83-86 */
            us = us + upper;
            us = us / fall;
            float w = fminf(us, ls);               /* :90-91 */
            w = fmaxf(w, 0.0f);
            weights[(size_t)k * n_mels + i] = w;
        }
        weights[i] = 0.0f;                         /* bin 0 carries no weight (:93) */
    }
    free(band);
    free(bin_hz);
}

typedef struct { uint32_t packed; float weight; } mel_entry;

/* Per-lane work list for the fused kernel (layout: nntk_mel_args in nntk_b200_shim.h).
 *  - Bands are placed longest-first on the least loaded lane, so a warp needs max(longest band, ~nnz / 32) iterations per
 *    frame pair instead of the sum of the per-round maxima.
 *  - The kernel parks the spectrum as 8-byte (frame a, frame b) pairs at slot k + k / 16, and one iteration is one 8-byte
 *    shared-memory load per lane: a half-warp is conflict-free when its 16 slots differ mod 16.  The order in which a lane
 *    walks the bins of a band is free, so every iteration each lane takes, among the bins left in its band, the one whose
 *    bank pair is least used by the lanes before it (first version, bins ascending: 6.8 wavefronts per load, ncu). */
static int build_entries(struct MelFilterBankStruct *b) {
    const int n_mels = b->config.n_mels, n_bins = b->config.n_fft / 2 + 1;
    if (n_bins > 480 || n_mels < 1 || n_mels >= (1 << 16)) return 0;
    int *nnz = (int *)calloc((size_t)n_mels, sizeof(int)), *order = (int *)malloc(sizeof(int) * (size_t)n_mels);
    int *lane_of = (int *)malloc(sizeof(int) * (size_t)n_mels);
    for (int i = 0; i < n_mels; ++i) {
        order[i] = i;
        for (int k = 0; k < n_bins; ++k) nnz[i] += b->weights[(size_t)k * n_mels + i] != 0.0f;
        if (nnz[i] == 0) nnz[i] = 1; /* an empty band still emits log(offset) */
    }
    for (int i = 1; i < n_mels; ++i) { /* insertion sort, descending */
        int m = order[i], j = i;
        while (j > 0 && nnz[order[j - 1]] < nnz[m]) { order[j] = order[j - 1]; --j; }
        order[j] = m;
    }
    int load[32] = {0};
    for (int i = 0; i < n_mels; ++i) {
        int best = 0;
        for (int l = 1; l < 32; ++l) if (load[l] < load[best]) best = l;
        lane_of[order[i]] = best;
        load[best] += nnz[order[i]];
    }
    int n_iter = 1;
    for (int l = 0; l < 32; ++l) if (load[l] > n_iter) n_iter = load[l];
    mel_entry *e = (mel_entry *)calloc((size_t)n_iter * 32, sizeof(mel_entry));
    /* per-lane band queues (ascending band index) and the bins still to visit in the current band */
    int *queue = (int *)malloc(sizeof(int) * (size_t)n_mels), q_begin[33], q_pos[32];
    unsigned char *left = (unsigned char *)malloc((size_t)32 * n_bins);
    int n_left[32];
    q_begin[0] = 0;
    for (int l = 0; l < 32; ++l) {
        int n = q_begin[l];
        for (int i = 0; i < n_mels; ++i) if (lane_of[i] == l) queue[n++] = i;
        q_begin[l + 1] = n;
        q_pos[l] = q_begin[l];
        n_left[l] = -1; /* band not opened yet */
    }
    for (int it = 0; it < n_iter; ++it) {
        int used[2][16] = {{0}};
        /* open the next band of every lane that needs one, then serve the lanes with the fewest bins left first (they have
         * the least choice), idle lanes last */
        int cnt[32], seq[32];
        for (int l = 0; l < 32; ++l) {
            seq[l] = l;
            if (q_pos[l] == q_begin[l + 1]) { cnt[l] = 1 << 30; continue; }
            if (n_left[l] < 0) {
                const int band = queue[q_pos[l]];
                unsigned char *lf = left + (size_t)l * n_bins;
                n_left[l] = 0;
                for (int k = 0; k < n_bins; ++k) { lf[k] = b->weights[(size_t)k * n_mels + band] != 0.0f; n_left[l] += lf[k]; }
            }
            cnt[l] = n_left[l];
        }
        for (int i = 1; i < 32; ++i) { /* insertion sort by cnt, stable */
            int l = seq[i], j = i;
            while (j > 0 && cnt[seq[j - 1]] > cnt[l]) { seq[j] = seq[j - 1]; --j; }
            seq[j] = l;
        }
        for (int si = 0; si < 32; ++si) {
            const int l = seq[si];
            mel_entry *dst = &e[(size_t)it * 32 + l];
            int *u = used[l >> 4];
            if (q_pos[l] == q_begin[l + 1]) { /* lane is done: a zero-weight read of the least used bank pair */
                int best = 0;
                for (int k = 1; k < 16; ++k) if (u[k] < u[best]) best = k;
                dst->packed = (uint32_t)best * 8u;
                ++u[best];
                continue;
            }
            const int band = queue[q_pos[l]];
            unsigned char *lf = left + (size_t)l * n_bins;
            int pick = -1;
            for (int k = 0; k < n_bins; ++k)
                if (lf[k] && (pick < 0 || u[(k + (k >> 4)) & 15] < u[(pick + (pick >> 4)) & 15])) pick = k;
            float w = 0.0f;
            if (pick < 0) pick = 0; /* empty band: one zero-weight entry */
            else { w = b->weights[(size_t)pick * n_mels + band]; lf[pick] = 0; --n_left[l]; }
            const int slot = pick + (pick >> 4);
            ++u[slot & 15];
            dst->packed = (uint32_t)slot * 8u;
            dst->weight = w;
            if (n_left[l] <= 0) { dst->packed |= 0x80000000u | ((uint32_t)band << 15); ++q_pos[l]; n_left[l] = -1; }
        }
    }
    free(nnz); free(order); free(lane_of); free(queue); free(left);
    b->entries = e;
    b->n_iter = n_iter;
    return 1;
}

MelFilterBank MelFilterBankCreate(MelFilterBankConfig config) {
    if (config.n_mels < 1 || config.n_fft < 2 || config.n_mels >= (1 << 22)) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "MelFilterBank: invalid config (n_mels %d, n_fft %d)", config.n_mels, config.n_fft);
        return NULL;
    }
    MelFilterBank b = (MelFilterBank)calloc(1, sizeof(*b));
    if (!b) return NULL;
    b->config = config;
    b->weights = (float *)calloc((size_t)(config.n_fft / 2 + 1) * config.n_mels, sizeof(float));
    if (!b->weights) { free(b); return NULL; }
    build_weights(&config, b->weights);
    build_entries(b);
    return b;
}

void MelFilterBankDestroy(MelFilterBank b) {
    if (!b) return;
    nntk_cuda_free(b->d_weights);
    nntk_cuda_free(b->d_entries);
    free(b->entries);
    free(b->weights);
    free(b);
}

/* the table never changes after Create (the reference exposes no accessor), so it is uploaded once */
int nntk_mel_stage(MelFilterBank b) {
    int rc;
    if (!b->d_weights) {
        const size_t bytes = sizeof(float) * (size_t)(b->config.n_fft / 2 + 1) * b->config.n_mels;
        void *d = NULL;
        if ((rc = nntk_cuda_malloc(&d, bytes))) return rc;
        if ((rc = nntk_cuda_h2d(d, b->weights, bytes)) || (rc = nntk_cuda_sync())) { nntk_cuda_free(d); return rc; }
        b->d_weights = (float *)d;
    }
    if (!b->d_entries && b->entries) {
        const size_t bytes = sizeof(mel_entry) * 32 * (size_t)b->n_iter;
        void *d = NULL;
        if ((rc = nntk_cuda_malloc(&d, bytes))) return rc;
        if ((rc = nntk_cuda_h2d(d, b->entries, bytes)) || (rc = nntk_cuda_sync())) { nntk_cuda_free(d); return rc; }
        b->d_entries = d;
    }
    return 0;
}

/* rows x [n_bins] -> rows x [n_mels]: op_mat_mul(spectrogram, weights, mel, timesteps, n_mels, n_bins), :114-116.
 * Always fp32 (the table is tiny; a TF32 pass would cost accuracy for nothing). */
int nntk_mel_gemm(MelFilterBank b, const float *d_in, float *d_out, long long rows) {
    if (rows > 0x7fffffffLL) { nntk_cuda_set_error(NNTK_ERR_ARG, "MelFilterBankApply: too many rows"); return NNTK_ERR_ARG; }
    nntk_gemm_args g;
    memset(&g, 0, sizeof(g));
    g.d_A = d_in; g.lda = b->config.n_fft / 2 + 1; g.batch = 1; g.rows = (int)rows;
    g.N = b->config.n_mels; g.K = b->config.n_fft / 2 + 1;
    g.d_B = b->d_weights;
    g.act = NNTK_ACT_NONE; g.act_a = 1.0f;
    g.d_C = d_out; g.ldc = b->config.n_mels;
    const int mode = nntk_cuda_math_mode();
    if (mode) nntk_cuda_set_math_mode(0);
    int rc = nntk_cuda_gemm(&g);
    if (mode) nntk_cuda_set_math_mode(mode);
    return rc;
}

static int mel_launch(void *ctx, const float *d_in, float *d_out, long long first, int n) {
    (void)first;
    return nntk_mel_gemm((MelFilterBank)ctx, d_in, d_out, n);
}

/* void like the reference: failures land in NNTKLastError() and on stderr */
void MelFilterBankApply(MelFilterBank b, const float *spectrogram, float *mel_spectrogram, int timesteps) {
    NNTK_API_GUARD();
    if (!b || timesteps < 0) { nntk_cuda_set_error(NNTK_ERR_ARG, "MelFilterBankApply: bad arguments"); return; }
    int kind = nntk_ptr_kind(spectrogram, mel_spectrogram, "MelFilterBankApply");
    if (kind < 0 || nntk_mel_stage(b)) return;
    if (kind == 1) { (void)mel_launch(b, spectrogram, mel_spectrogram, 0, timesteps); return; }
    (void)nntk_cuda_staged_run(spectrogram, mel_spectrogram, sizeof(float) * (size_t)(b->config.n_fft / 2 + 1),
                               sizeof(float) * (size_t)b->config.n_mels, timesteps, mel_launch, b);
}
