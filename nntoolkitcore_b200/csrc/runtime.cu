// runtime.cu -- process-wide runtime state of the shim: device, stream, error slot, memory helpers and the
// host-buffer staging pipeline.  No compute lives here.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <atomic>
#include <mutex>
#include "common.cuh"

namespace {
std::recursive_mutex g_mu;          // the API lock (nntk_cuda_api_lock) -- also taken by init / staged_run / workspace
std::atomic<long long> g_launches{0};
int g_device = -1;            // -1: not initialised
int g_sm_count = 0;
cudaStream_t g_stream = 0;    // legacy default stream unless the caller sets one
int g_math_mode = 2;            // NNTK_MATH_TF32X3: the fp32 contract (1e-5) on the tensor cores, FFMA above its K cap (b200_ext.h)
bool g_math_mode_set = false;
thread_local char t_err[512] = {0};
thread_local int t_err_code = 0;

// staging pipeline resources (created on first use)
struct Stage {
    cudaStream_t s_in = nullptr, s_run = nullptr, s_out = nullptr;
    void *d_in[2] = {nullptr, nullptr}, *d_out[2] = {nullptr, nullptr};
    size_t in_cap = 0, out_cap = 0;
    cudaEvent_t in_ready[2], run_done[2], out_done[2], in_free[2];
    bool made = false;
} g_stage;
}  // namespace

extern "C" void nntk_cuda_set_error(int code, const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_err, sizeof(t_err), fmt, ap);
    va_end(ap);
    t_err_code = code;
    fprintf(stderr, "[nntk-b200] error %d: %s\n", code, t_err);
}
extern "C" const char *nntk_cuda_last_error(void) { return t_err; }
extern "C" void nntk_cuda_clear_error(void) { t_err[0] = 0; t_err_code = 0; }

extern "C" int nntk_cuda_init(int device) {
    std::lock_guard<std::recursive_mutex> lk(g_mu);
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "no CUDA device available (%s); this library has no CPU fallback",
                            e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
        return NNTK_ERR_CUDA_;
    }
    if (device < 0) NNTK_CUDA_TRY(cudaGetDevice(&device));
    if (device >= count) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "device %d out of range (%d devices)", device, count);
        return NNTK_ERR_ARG_;
    }
    NNTK_CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    NNTK_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "device %d is sm_%d%d; this library is built for sm_100a only", device,
                            prop.major, prop.minor);
        return NNTK_ERR_CUDA_;
    }
    g_device = device;
    g_sm_count = prop.multiProcessorCount;
    return 0;
}

extern "C" int nntk_cuda_ensure_init(void) {
    if (g_device >= 0) {
        // a host thread other than the initialising one may call in: make the device current for it
        int cur = -1;
        if (cudaGetDevice(&cur) == cudaSuccess && cur != g_device) cudaSetDevice(g_device);
        return 0;
    }
    return nntk_cuda_init(-1);
}

extern "C" void nntk_cuda_count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
extern "C" long long nntk_cuda_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }
extern "C" void nntk_cuda_api_lock(void) { g_mu.lock(); }
extern "C" void nntk_cuda_api_unlock(void) { g_mu.unlock(); }
extern "C" int nntk_cuda_sm_count(void) { return g_sm_count > 0 ? g_sm_count : 148; }
extern "C" int nntk_cuda_set_stream(void *stream) { g_stream = (cudaStream_t)stream; return 0; }
extern "C" void *nntk_cuda_stream(void) { return (void *)g_stream; }
extern "C" int nntk_cuda_sync(void) {
    if (g_device < 0) return 0;
    cudaStream_t s;
    { std::lock_guard<std::recursive_mutex> lk(g_mu); s = g_stream; }   // never the pipeline stream a host-buffer Apply swapped in
    NNTK_CUDA_TRY(cudaStreamSynchronize(s));
    return 0;
}
extern "C" int nntk_cuda_set_math_mode(int mode) {
    if (mode < 0 || mode > 2) { nntk_cuda_set_error(NNTK_ERR_ARG_, "math mode %d unknown", mode); return NNTK_ERR_ARG_; }
    g_math_mode = mode;
    g_math_mode_set = true;
    return 0;
}
// The process-wide default comes from NNTK_MATH_MODE (0 FP32, 1 TF32, 2 TF32X3) until NNTKSetMathMode is called.
extern "C" int nntk_cuda_math_mode(void) {
    static bool env_read = false;
    if (!env_read) {
        env_read = true;
        if (const char *e = getenv("NNTK_MATH_MODE")) { const int m = atoi(e); if (m >= 0 && m <= 2 && !g_math_mode_set) g_math_mode = m; }
    }
    return g_math_mode;
}

extern "C" int nntk_cuda_malloc(void **d_ptr, size_t bytes) {
    *d_ptr = nullptr;
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (bytes == 0) bytes = 16;
    NNTK_CUDA_TRY(cudaMalloc(d_ptr, bytes));
    return 0;
}
extern "C" void nntk_cuda_free(void *d_ptr) { if (d_ptr) cudaFree(d_ptr); }

// Weight blocks are handed to the caller right after Create, possibly on a box with no GPU (the CPU test tier
// creates handles and fills weights).  Pinned memory when a device exists, plain malloc otherwise; a tiny header
// remembers which so the free matches.  Apply still fails loudly without a device.
namespace { struct HostHdr { uint64_t magic; uint64_t pinned; uint64_t pad[6]; }; }
extern "C" int nntk_cuda_host_alloc(void **h_ptr, size_t bytes) {
    *h_ptr = nullptr;
    size_t total = bytes + sizeof(HostHdr);
    void *raw = nullptr;
    uint64_t pinned = 0;
    int count = 0;
    if (cudaGetDeviceCount(&count) == cudaSuccess && count > 0 && nntk_cuda_ensure_init() == 0) {
        if (cudaHostAlloc(&raw, total, cudaHostAllocDefault) == cudaSuccess) pinned = 1; else raw = nullptr;
    }
    (void)cudaGetLastError();
    if (!raw) raw = malloc(total);
    if (!raw) { nntk_cuda_set_error(NNTK_ERR_ARG_, "host allocation of %zu bytes failed", bytes); return NNTK_ERR_ARG_; }
    HostHdr *h = (HostHdr *)raw;
    h->magic = 0x4e4e544b484f5354ull;
    h->pinned = pinned;
    *h_ptr = (void *)(h + 1);
    memset(*h_ptr, 0, bytes);   // f_malloc zero-fills (core/memory.c:10-14)
    return 0;
}
extern "C" void nntk_cuda_host_free(void *h_ptr) {
    if (!h_ptr) return;
    HostHdr *h = ((HostHdr *)h_ptr) - 1;
    if (h->magic != 0x4e4e544b484f5354ull) return;
    h->magic = 0;
    if (h->pinned) cudaFreeHost(h); else free(h);
}

extern "C" int nntk_cuda_h2d(void *d, const void *h, size_t bytes) {
    NNTK_CUDA_TRY(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, g_stream));
    return 0;
}
extern "C" int nntk_cuda_d2h(void *h, const void *d, size_t bytes) {
    NNTK_CUDA_TRY(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, g_stream));
    return 0;
}
extern "C" int nntk_cuda_d2d(void *dst, const void *src, size_t bytes) {
    NNTK_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, g_stream));
    return 0;
}
extern "C" int nntk_cuda_memset(void *d, int v, size_t bytes) {
    NNTK_CUDA_TRY(cudaMemsetAsync(d, v, bytes, g_stream));
    return 0;
}
extern "C" int nntk_cuda_event_create(void **ev) {
    cudaEvent_t e;
    NNTK_CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    *ev = (void *)e;
    return 0;
}
extern "C" int nntk_cuda_event_record(void *ev) { NNTK_CUDA_TRY(cudaEventRecord((cudaEvent_t)ev, g_stream)); return 0; }
extern "C" int nntk_cuda_event_sync(void *ev) { NNTK_CUDA_TRY(cudaEventSynchronize((cudaEvent_t)ev)); return 0; }
extern "C" void nntk_cuda_event_destroy(void *ev) { if (ev) cudaEventDestroy((cudaEvent_t)ev); }
extern "C" int nntk_cuda_is_device_ptr(const void *ptr) {
    if (!ptr || g_device < 0) {
        int count = 0;
        if (!ptr || cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { (void)cudaGetLastError(); return 0; }
    }
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    return (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged) ? 1 : 0;
}

// ---- host-buffer staging pipeline -------------------------------------------------------------------------------
// Two device slots; chunk i: H2D on s_in -> launch on s_run (the library stream is pointed at s_run while the
// callback runs) -> D2H on s_out.  PCIe is full duplex, so H2D(i+1), compute(i) and D2H(i-1) overlap.
static int stage_make(size_t in_bytes, size_t out_bytes) {
    Stage &s = g_stage;
    if (!s.made) {
        NNTK_CUDA_TRY(cudaStreamCreateWithFlags(&s.s_in, cudaStreamNonBlocking));
        NNTK_CUDA_TRY(cudaStreamCreateWithFlags(&s.s_run, cudaStreamNonBlocking));
        NNTK_CUDA_TRY(cudaStreamCreateWithFlags(&s.s_out, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) {
            NNTK_CUDA_TRY(cudaEventCreateWithFlags(&s.in_ready[i], cudaEventDisableTiming));
            NNTK_CUDA_TRY(cudaEventCreateWithFlags(&s.run_done[i], cudaEventDisableTiming));
            NNTK_CUDA_TRY(cudaEventCreateWithFlags(&s.out_done[i], cudaEventDisableTiming));
            NNTK_CUDA_TRY(cudaEventCreateWithFlags(&s.in_free[i], cudaEventDisableTiming));
        }
        s.made = true;
    }
    if (in_bytes > s.in_cap) {
        for (int i = 0; i < 2; ++i) { if (s.d_in[i]) cudaFree(s.d_in[i]); s.d_in[i] = nullptr; }
        for (int i = 0; i < 2; ++i) NNTK_CUDA_TRY(cudaMalloc(&s.d_in[i], in_bytes));
        s.in_cap = in_bytes;
    }
    if (out_bytes > s.out_cap) {
        for (int i = 0; i < 2; ++i) { if (s.d_out[i]) cudaFree(s.d_out[i]); s.d_out[i] = nullptr; }
        for (int i = 0; i < 2; ++i) NNTK_CUDA_TRY(cudaMalloc(&s.d_out[i], out_bytes));
        s.out_cap = out_bytes;
    }
    return 0;
}

extern "C" int nntk_cuda_staged_run(const void *h_in, void *h_out, size_t unit_in_bytes, size_t unit_out_bytes,
                                    long long n_units, nntk_launch_fn launch, void *ctx) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (n_units <= 0) return 0;
    std::lock_guard<std::recursive_mutex> lk(g_mu);
    // chunk so that one slot moves ~256 MiB of (in + out): large enough for PCIe efficiency, small enough to overlap
    const size_t target = (size_t)256 << 20;
    size_t per_unit = unit_in_bytes + unit_out_bytes;
    long long chunk = (long long)(target / (per_unit ? per_unit : 1));
    if (chunk < 1) chunk = 1;
    if (chunk > n_units) chunk = n_units;
    if (n_units > 1 && chunk == n_units && per_unit * (size_t)n_units > ((size_t)32 << 20)) chunk = (n_units + 1) / 2;
    rc = stage_make((size_t)chunk * unit_in_bytes, (size_t)chunk * unit_out_bytes);
    if (rc) return rc;
    Stage &s = g_stage;
    cudaStream_t saved = g_stream;
    // The inputs come from the host, but the layer's WEIGHTS were just staged (H2D + repack kernels) on the caller's stream,
    // and the pipeline streams are non-blocking: without this edge the first chunk's kernels could run before the weights
    // had landed (seen once in ~30 runs of the GPU tests as an all-zero conv output: weights, bias and BatchNorm table still
    // zero).  One event on the caller's stream, waited for by the compute stream.
    {
        cudaError_t e = cudaEventRecord(s.in_free[0], saved);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(s.s_run, s.in_free[0], 0);
        if (e != cudaSuccess) { nntk_cuda_set_error(NNTK_ERR_CUDA_, "staged pipeline: %s", cudaGetErrorString(e)); return NNTK_ERR_CUDA_; }
    }
    int result = 0;
    long long n_chunks = (n_units + chunk - 1) / chunk;
    for (long long i = 0; i < n_chunks && result == 0; ++i) {
        int slot = (int)(i & 1);
        long long first = i * chunk;
        int n = (int)((first + chunk <= n_units) ? chunk : (n_units - first));
        // slot reuse: input buffer free once compute(i-2) finished; output buffer free once D2H(i-2) finished
        if (i >= 2) {
            cudaStreamWaitEvent(s.s_in, s.run_done[slot], 0);
            cudaStreamWaitEvent(s.s_run, s.out_done[slot], 0);
        }
        cudaError_t e = cudaMemcpyAsync(s.d_in[slot], (const char *)h_in + (size_t)first * unit_in_bytes,
                                        (size_t)n * unit_in_bytes, cudaMemcpyHostToDevice, s.s_in);
        if (e != cudaSuccess) { nntk_cuda_set_error(NNTK_ERR_CUDA_, "staged H2D: %s", cudaGetErrorString(e)); result = NNTK_ERR_CUDA_; break; }
        cudaEventRecord(s.in_ready[slot], s.s_in);
        cudaStreamWaitEvent(s.s_run, s.in_ready[slot], 0);
        g_stream = s.s_run;
        result = launch(ctx, (const float *)s.d_in[slot], (float *)s.d_out[slot], first, n);
        g_stream = saved;
        if (result) break;
        cudaEventRecord(s.run_done[slot], s.s_run);
        cudaStreamWaitEvent(s.s_out, s.run_done[slot], 0);
        e = cudaMemcpyAsync((char *)h_out + (size_t)first * unit_out_bytes, s.d_out[slot], (size_t)n * unit_out_bytes,
                            cudaMemcpyDeviceToHost, s.s_out);
        if (e != cudaSuccess) { nntk_cuda_set_error(NNTK_ERR_CUDA_, "staged D2H: %s", cudaGetErrorString(e)); result = NNTK_ERR_CUDA_; break; }
        cudaEventRecord(s.out_done[slot], s.s_out);
    }
    g_stream = saved;
    cudaError_t e1 = cudaStreamSynchronize(s.s_in), e2 = cudaStreamSynchronize(s.s_run), e3 = cudaStreamSynchronize(s.s_out);
    if (result == 0 && (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess)) {
        cudaError_t e = e1 != cudaSuccess ? e1 : (e2 != cudaSuccess ? e2 : e3);
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "staged pipeline: %s", cudaGetErrorString(e));
        result = NNTK_ERR_CUDA_;
    }
    return result;
}

// ---- shared device workspace ------------------------------------------------------------------------------------
// One grow-only scratch block for the whole library (hoisted projections, zero states, packed weights are per handle;
// this is only for per-call scratch).  All work is stream-ordered on the library stream, so sharing is safe.
namespace { void *g_work = nullptr; size_t g_work_bytes = 0; }
extern "C" int nntk_cuda_workspace(size_t want, size_t at_least, void **d_ptr, size_t *got) {
    *d_ptr = nullptr; *got = 0;
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    size_t cap = (size_t)16 << 30;
    if (const char *e = getenv("NNTK_WORKSPACE_MB")) { long long mb = atoll(e); if (mb > 0) cap = (size_t)mb << 20; }
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess) { size_t lim = (free_b + g_work_bytes) / 2; if (cap > lim) cap = lim; }
    size_t need = want < cap ? want : cap;
    if (need < at_least) need = at_least;
    if (need > g_work_bytes) {
        NNTK_CUDA_TRY(cudaDeviceSynchronize());
        if (g_work) cudaFree(g_work);
        g_work = nullptr; g_work_bytes = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_work, need));
        g_work_bytes = need;
    }
    *d_ptr = g_work; *got = g_work_bytes;
    return 0;
}
