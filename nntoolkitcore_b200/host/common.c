/* common.c -- layer header helpers, pointer-kind check and the additive runtime entry points of b200_ext.h. */
#include <string.h>
#include "internal.h"

int nntk_layer_init(nntk_layer_hdr *h, int kind, size_t n_floats) {
    memset(h, 0, sizeof(*h));
    h->magic = NNTK_LAYER_MAGIC;
    h->kind = kind;
    h->n_floats = n_floats;
    void *p = NULL;
    /* zero-filled like f_malloc (core/memory.c:10-14); pinned when a device exists, plain malloc on a CPU-only box */
    if (nntk_cuda_host_alloc(&p, n_floats * sizeof(float))) return NNTK_ERR_ARG;
    h->h_block = (float *)p;
    return 0;
}

void nntk_layer_release(nntk_layer_hdr *h) {
    if (!h) return;
    if (h->snap_ev) { nntk_cuda_event_sync(h->snap_ev); nntk_cuda_event_destroy(h->snap_ev); }
    nntk_cuda_host_free(h->h_snap);
    nntk_cuda_host_free(h->h_block);
    nntk_cuda_free(h->d_block);
    h->h_block = h->d_block = h->h_snap = NULL;
    h->snap_ev = NULL;
    h->magic = 0;
}

int nntk_layer_stage(nntk_layer_hdr *h) {
    int rc;
    if (h->frozen && h->staged) return 0;
    if (!h->d_block) {
        void *d = NULL;
        if ((rc = nntk_cuda_malloc(&d, h->n_floats * sizeof(float)))) return rc;
        h->d_block = (float *)d;
    }
    /* snapshot first (synchronous host copy), then the asynchronous H2D reads the snapshot: a caller that rewrites its weights
     * right after a device-pointer Apply returned cannot race with the copy engine.  The previous restage's H2D must have
     * drained before the snapshot is overwritten (one event, normally long complete). */
    if (!h->h_snap) {
        void *p = NULL;
        if (nntk_cuda_host_alloc(&p, h->n_floats * sizeof(float))) return NNTK_ERR_ARG;
        h->h_snap = (float *)p;
        if ((rc = nntk_cuda_event_create(&h->snap_ev))) return rc;
    } else if ((rc = nntk_cuda_event_sync(h->snap_ev))) return rc;
    memcpy(h->h_snap, h->h_block, h->n_floats * sizeof(float));
    if ((rc = nntk_cuda_h2d(h->d_block, h->h_snap, h->n_floats * sizeof(float)))) return rc;
    if ((rc = nntk_cuda_event_record(h->snap_ev))) return rc;
    if (h->after_stage && (rc = h->after_stage(h))) return rc;
    h->staged = 1;
    return 0;
}

int nntk_ptr_kind(const void *in, const void *out, const char *who) {
    if (!in || !out) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "%s: NULL input or output", who);
        return NNTK_ERR_ARG;
    }
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    int a = nntk_cuda_is_device_ptr(in), b = nntk_cuda_is_device_ptr(out);
    if (a != b) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "%s: input and output must both be host or both be device pointers", who);
        return NNTK_ERR_ARG;
    }
    return a;
}

static nntk_layer_hdr *as_layer(void *layer, const char *who) {
    nntk_layer_hdr *h = (nntk_layer_hdr *)layer;
    if (!h || h->magic != NNTK_LAYER_MAGIC) {
        nntk_cuda_set_error(NNTK_ERR_ARG, "%s: not a layer handle of this library", who);
        return NULL;
    }
    return h;
}

int NNTKInit(int device) { return nntk_cuda_init(device); }
int NNTKSetStream(void *s) { NNTK_API_GUARD(); return nntk_cuda_set_stream(s); }
void *NNTKGetStream(void) { return nntk_cuda_stream(); }
int NNTKSynchronize(void) { return nntk_cuda_sync(); }
const char *NNTKLastError(void) { return nntk_cuda_last_error(); }
void NNTKClearError(void) { nntk_cuda_clear_error(); }
long long NNTKDebugLaunchCount(void) { return nntk_cuda_launch_count(); }
const char *NNTKDebugLastRecurrentPath(void) { return nntk_cuda_last_recurrent_path(); }
int NNTKDebugLastRecurrentInputPasses(void) { return nntk_cuda_last_recurrent_xpasses(); }
int NNTKSetMathMode(int mode) { NNTK_API_GUARD(); return nntk_cuda_set_math_mode(mode); }
int NNTKGetMathMode(void) { return nntk_cuda_math_mode(); }
void *NNTKDeviceAlloc(size_t bytes) { void *p = NULL; return nntk_cuda_malloc(&p, bytes) ? NULL : p; }
void NNTKDeviceFree(void *p) { nntk_cuda_free(p); }
void *NNTKHostAlloc(size_t bytes) { void *p = NULL; return nntk_cuda_host_alloc(&p, bytes) ? NULL : p; }
void NNTKHostFree(void *p) { nntk_cuda_host_free(p); }
int NNTKCopyToDevice(void *d, const void *h, size_t n) { NNTK_API_GUARD(); int rc = nntk_cuda_ensure_init(); return rc ? rc : nntk_cuda_h2d(d, h, n); }
int NNTKCopyToHost(void *h, const void *d, size_t n) { NNTK_API_GUARD(); int rc = nntk_cuda_ensure_init(); return rc ? rc : nntk_cuda_d2h(h, d, n); }
int NNTKIsDevicePointer(const void *p) { return nntk_cuda_is_device_ptr(p); }

int NNTKWeightsFreeze(void *layer) {
    NNTK_API_GUARD();
    nntk_layer_hdr *h = as_layer(layer, "NNTKWeightsFreeze");
    if (!h) return NNTK_ERR_ARG;
    h->frozen = 0;
    h->staged = 0;
    int rc = nntk_layer_stage(h);
    if (rc) return rc;
    h->frozen = 1;
    return 0;
}
int NNTKWeightsUnfreeze(void *layer) {
    NNTK_API_GUARD();
    nntk_layer_hdr *h = as_layer(layer, "NNTKWeightsUnfreeze");
    if (!h) return NNTK_ERR_ARG;
    h->frozen = 0;
    return 0;
}
