// elementwise.cu -- K5: standalone BatchNorm and activation kernels (replace layers/batch_norm.c:140-189 and
// layers/activation_default.c:28-33,65-67,98-103,123-129,149-167 of the reference for the inference path).
// Single pass over the tensor (the reference makes 6 passes for BN, 3-4 per activation): HBM roofline = 8 B/element.
// When a BatchNorm / activation directly follows a Conv1d / Dense it is fused into that kernel's epilogue instead
// (csrc/gemm.cu); these kernels serve the stand-alone calls.
#include "common.cuh"

namespace {

// 16-byte streaming accesses: every element is touched once, keep it out of L1.  Not .nc: in == out is legal.
__device__ __forceinline__ float4 ld_stream(const float4 *p) {
    float4 r;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(float4 *p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// prm = gamma | beta | mean | var.  sd = sqrtf(var + eps) is the same value for every row, so it is evaluated per
// element only in the sense of the reference's op order: x_mu / sd with an IEEE divide (batch_norm.c:151-162).
// VEC: C % 4 == 0 and 16-byte aligned pointers.
template <bool VEC>
// The data pointers of every kernel in this file are NOT __restrict__ and are never read through the non-coherent path: the
// reference allows input == output for BatchNorm and the activations, and the library itself calls them in place (Dense softmax,
// the unfused log-mel tail).
__global__ void __launch_bounds__(256) batch_norm_kernel(const float *__restrict__ prm, int C, float eps,
                                                         const float *in, float *out,
                                                         long long total) {
    const float *gamma = prm, *beta = prm + C, *mean = prm + 2 * C, *var = prm + 3 * C;
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        const long long nvec = total >> 2;
        const int cvec = C >> 2;
        for (; i < nvec; i += stride) {
            const int c = (int)(i % cvec) * 4;
            float4 x = ld_stream(reinterpret_cast<const float4 *>(in) + i);
            const float4 g = __ldg(reinterpret_cast<const float4 *>(gamma + c));
            const float4 b = __ldg(reinterpret_cast<const float4 *>(beta + c));
            const float4 m = __ldg(reinterpret_cast<const float4 *>(mean + c));
            const float4 v = __ldg(reinterpret_cast<const float4 *>(var + c));
            x.x = nntk_bn(x.x, g.x, b.x, m.x, v.x, eps);
            x.y = nntk_bn(x.y, g.y, b.y, m.y, v.y, eps);
            x.z = nntk_bn(x.z, g.z, b.z, m.z, v.z, eps);
            x.w = nntk_bn(x.w, g.w, b.w, m.w, v.w, eps);
            st_stream(reinterpret_cast<float4 *>(out) + i, x);
        }
    } else {
        for (; i < total; i += stride) {
            const int c = (int)(i % C);
            out[i] = nntk_bn(in[i], __ldg(gamma + c), __ldg(beta + c), __ldg(mean + c), __ldg(var + c), eps);
        }
    }
}

template <bool VEC>
__global__ void __launch_bounds__(256) activation_kernel(int kind, float a, const float *in, float *out, long long total) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (VEC) {
        const long long nvec = total >> 2;
        for (long long v = i; v < nvec; v += stride) {
            float4 x = ld_stream(reinterpret_cast<const float4 *>(in) + v);
            x.x = nntk_act(kind, a, x.x);
            x.y = nntk_act(kind, a, x.y);
            x.z = nntk_act(kind, a, x.z);
            x.w = nntk_act(kind, a, x.w);
            st_stream(reinterpret_cast<float4 *>(out) + v, x);
        }
        for (long long s = (nvec << 2) + i; s < total; s += stride) out[s] = nntk_act(kind, a, in[s]);
    } else {
        for (; i < total; i += stride) out[i] = nntk_act(kind, a, in[i]);
    }
}

// softmax without max-subtraction, exactly the reference's definition (activation_default.c:149-167):
// e_j = expf(x_j); y_j = e_j / sum(e).  One warp per vector; the sum is a lane-strided partial + shuffle tree
// (the reference sums sequentially: results agree to rounding, not bit-for-bit).
__global__ void __launch_bounds__(256) softmax_kernel(int V, const float *in, float *out, long long n_vec) {
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long r = (((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5); r < n_vec; r += warps) {
        const float *x = in + r * V;
        float *y = out + r * V;
        float s = 0.0f;
        if (V <= 32) {
            float e = (lane < V) ? expf(x[lane]) : 0.0f;
            s = e;
#pragma unroll
            for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane < V) y[lane] = e / s;
        } else {
            // in-place safe: a lane only ever rereads elements it alone writes (j = lane + 32 i), and the sum is complete
            // before the first write
            for (int j = lane; j < V; j += 32) s += expf(x[j]);
#pragma unroll
            for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            for (int j = lane; j < V; j += 32) { const float e = expf(x[j]); y[j] = e / s; }
        }
    }
}

inline unsigned grid_for(long long work_items, int per_block) {
    // enough CTAs for ~8 resident per SM, grid-stride beyond that
    long long blocks = (work_items + per_block - 1) / per_block;
    long long cap = (long long)nntk_cuda_sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (unsigned)blocks;
}


// Any C (e.g. the 257 frequency bins of C5).  gamma | beta | mean | sd sit in shared memory; sd = sqrtf(var + eps) is
// evaluated once per channel -- the same fp32 value the reference recomputes for every row (batch_norm.c:153-156).
// Lane = consecutive element: a warp instruction moves 128 contiguous bytes AND reads 32 consecutive table entries
// (conflict-free).  (The first version gave each thread a float4: lanes were 4 channels apart, every table read was a
// 4-way bank conflict and the kernel ran at 0.64 of the HBM peak.)  Four independent elements per thread and iteration.
__global__ void __launch_bounds__(256) batch_norm_any_kernel(const float *__restrict__ prm, int C, float eps,
                                                             const float *in, float *out,   // in == out is legal
                                                             long long total) {
    extern __shared__ float tab[];       // [4][C]
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        tab[c] = prm[c]; tab[C + c] = prm[C + c]; tab[2 * C + c] = prm[2 * C + c];
        tab[3 * C + c] = sqrtf(__fadd_rn(prm[3 * C + c], eps));
    }
    __syncthreads();
    const float *gamma = tab, *beta = tab + C, *mean = tab + 2 * C, *sd = tab + 3 * C;
    const long long stride = (long long)gridDim.x * blockDim.x;
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int step = (int)(stride % C);
    int c = (int)(t % C);
    for (; t + 3 * stride < total; t += 4 * stride) {
        float x[4];
        int ce[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            x[e] = in[t + e * stride];
            ce[e] = c;
            c += step; if (c >= C) c -= C;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e)
            out[t + e * stride] = __fadd_rn(__fmul_rn(__fdiv_rn(__fsub_rn(x[e], mean[ce[e]]), sd[ce[e]]), gamma[ce[e]]), beta[ce[e]]);
    }
    for (; t < total; t += stride) {
        out[t] = __fadd_rn(__fmul_rn(__fdiv_rn(__fsub_rn(in[t], mean[c]), sd[c]), gamma[c]), beta[c]);
        c += step; if (c >= C) c -= C;
    }
}
}  // namespace

extern "C" int nntk_cuda_batch_norm(const float *d_params, int C, float eps, const float *d_in, float *d_out,
                                    long long rows) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!d_params || !d_in || !d_out || C < 1 || rows < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "batch_norm: bad arguments");
        return NNTK_ERR_ARG_;
    }
    const long long total = rows * C;
    if (total == 0) return 0;
    const bool vec = (C % 4 == 0) && ((((uintptr_t)d_in) | ((uintptr_t)d_out) | ((uintptr_t)d_params)) & 15) == 0;
    const bool aligned = ((((uintptr_t)d_in) | ((uintptr_t)d_out)) & 15) == 0;
    if (vec)
        batch_norm_kernel<true><<<grid_for(total >> 2, 256), 256, 0, nntk_stream()>>>(d_params, C, eps, d_in, d_out, total);
    else if (aligned && C <= 2048 && total >= 4096)
        batch_norm_any_kernel<<<grid_for(total >> 2, 256), 256, 4 * C * sizeof(float), nntk_stream()>>>(d_params, C, eps, d_in, d_out, total);
    else
        batch_norm_kernel<false><<<grid_for(total, 256), 256, 0, nntk_stream()>>>(d_params, C, eps, d_in, d_out, total);
    NNTK_LAUNCH_CHECK("batch_norm_kernel");
    return 0;
}

extern "C" int nntk_cuda_activation(int kind, float relu_a, int vector_size, const float *d_in, float *d_out,
                                    long long n) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!d_in || !d_out || n < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "activation: bad arguments");
        return NNTK_ERR_ARG_;
    }
    if (n == 0) return 0;
    if (kind == NNTK_ACT_SOFTMAX) {
        if (vector_size < 1) {
            nntk_cuda_set_error(NNTK_ERR_ARG_, "softmax: vector_size %d", vector_size);
            return NNTK_ERR_ARG_;
        }
        softmax_kernel<<<grid_for(n, 8), 256, 0, nntk_stream()>>>(vector_size, d_in, d_out, n);
        NNTK_LAUNCH_CHECK("softmax_kernel");
        return 0;
    }
    if (kind == NNTK_ACT_CUSTOM) {
        nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "activation: custom host callbacks cannot run on the device");
        return NNTK_ERR_UNSUPPORTED_;
    }
    if ((kind == NNTK_ACT_IDENTITY || kind == NNTK_ACT_NONE)) {
        if (d_in != d_out) NNTK_CUDA_TRY(cudaMemcpyAsync(d_out, d_in, (size_t)n * sizeof(float), cudaMemcpyDeviceToDevice, nntk_stream()));
        return 0;
    }
    const bool vec = ((((uintptr_t)d_in) | ((uintptr_t)d_out)) & 15) == 0;
    if (vec)
        activation_kernel<true><<<grid_for((n + 3) >> 2, 256), 256, 0, nntk_stream()>>>(kind, relu_a, d_in, d_out, n);
    else
        activation_kernel<false><<<grid_for(n, 256), 256, 0, nntk_stream()>>>(kind, relu_a, d_in, d_out, n);
    NNTK_LAUNCH_CHECK("activation_kernel");
    return 0;
}

// out = logf(in + offset): the tail of the unfused log-mel chain (signal/log_mel_spectrogram.c:36-37)
namespace {
__global__ void __launch_bounds__(256) log_offset_kernel(const float *in, float *out, long long n, float offset) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = logf(in[i] + offset);
}
}  // namespace

extern "C" int nntk_cuda_log_offset(const float *d_in, float *d_out, long long n, float offset) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!d_in || !d_out || n < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "log_offset: bad arguments");
        return NNTK_ERR_ARG_;
    }
    if (n == 0) return 0;
    log_offset_kernel<<<grid_for(n, 256), 256, 0, nntk_stream()>>>(d_in, d_out, n, offset);
    NNTK_LAUNCH_CHECK("log_offset_kernel");
    return 0;
}

// ---- bidirectional helpers (layers/bidirectional.c:10-85): pure data movement at 8 B/element --
namespace {
// out[b][t][:] = in[b][T - 1 - t][:]; V = float4 when the row length and the pointers allow, else float
template <typename V>
__global__ void __launch_bounds__(256) bd_reverse_kernel(const V *__restrict__ in, V *__restrict__ out, int vec, int T, long long total) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long per_seq = (long long)vec * T;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long b = i / per_seq;
        const int r = (int)(i - b * per_seq), t = r / vec, c = r - t * vec;
        out[i] = in[b * per_seq + (long long)(T - 1 - t) * vec + c];
    }
}
// out[row][0:cols] = fwd[row][:], out[row][cols:2 cols] = bwd[row][:]
template <typename V>
__global__ void __launch_bounds__(256) bd_concat_kernel(const V *__restrict__ fwd, const V *__restrict__ bwd, V *__restrict__ out, int cols,
                                                        long long total) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        const long long row = i / (2 * cols);
        const int c = (int)(i - row * 2 * cols);
        out[i] = c < cols ? fwd[row * cols + c] : bwd[row * cols + (c - cols)];
    }
}
__global__ void __launch_bounds__(256) bd_sum_kernel(const float *a, const float *b, float *out, long long n) {      // out may alias a or b
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) out[i] = a[i] + b[i];
}
inline bool aligned16(const void *p) { return (((uintptr_t)p) & 15) == 0; }
}  // namespace

extern "C" int nntk_cuda_bd_reverse(const float *d_in, float *d_out, int vec, int T, long long batch) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!d_in || !d_out || d_in == d_out || vec < 1 || T < 0 || batch < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "bd_reverse: bad arguments (in-place is not supported)");
        return NNTK_ERR_ARG_;
    }
    const long long total = batch * (long long)T * vec;
    if (total == 0) return 0;
    if (vec % 4 == 0 && aligned16(d_in) && aligned16(d_out))
        bd_reverse_kernel<float4><<<grid_for(total >> 2, 256), 256, 0, nntk_stream()>>>(
            reinterpret_cast<const float4 *>(d_in), reinterpret_cast<float4 *>(d_out), vec >> 2, T, total >> 2);
    else
        bd_reverse_kernel<float><<<grid_for(total, 256), 256, 0, nntk_stream()>>>(d_in, d_out, vec, T, total);
    NNTK_LAUNCH_CHECK("bd_reverse_kernel");
    return 0;
}

extern "C" int nntk_cuda_bd_concat(const float *d_fwd, const float *d_bwd, float *d_out, int cols, long long rows) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!d_fwd || !d_bwd || !d_out || cols < 1 || rows < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "bd_concat: bad arguments");
        return NNTK_ERR_ARG_;
    }
    const long long total = rows * 2 * cols;
    if (total == 0) return 0;
    if (cols % 4 == 0 && aligned16(d_fwd) && aligned16(d_bwd) && aligned16(d_out))
        bd_concat_kernel<float4><<<grid_for(total >> 2, 256), 256, 0, nntk_stream()>>>(
            reinterpret_cast<const float4 *>(d_fwd), reinterpret_cast<const float4 *>(d_bwd), reinterpret_cast<float4 *>(d_out),
            cols >> 2, total >> 2);
    else
        bd_concat_kernel<float><<<grid_for(total, 256), 256, 0, nntk_stream()>>>(d_fwd, d_bwd, d_out, cols, total);
    NNTK_LAUNCH_CHECK("bd_concat_kernel");
    return 0;
}

extern "C" int nntk_cuda_bd_sum(const float *d_a, const float *d_b, float *d_out, long long n) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!d_a || !d_b || !d_out || n < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "bd_sum: bad arguments");
        return NNTK_ERR_ARG_;
    }
    if (n == 0) return 0;
    bd_sum_kernel<<<grid_for(n, 256), 256, 0, nntk_stream()>>>(d_a, d_b, d_out, n);
    NNTK_LAUNCH_CHECK("bd_sum_kernel");
    return 0;
}
