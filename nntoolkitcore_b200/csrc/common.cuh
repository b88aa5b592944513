// common.cuh -- shared helpers for the sm_100a kernels behind include/nntk_b200_shim.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "nntk_b200_shim.h"

#define NNTK_ERR_UNSUPPORTED_ (-2)
#define NNTK_ERR_CUDA_ (-3)
#define NNTK_ERR_ARG_ (-4)

// Records a CUDA failure in the thread's last-error slot and returns NNTK_ERR_CUDA from the enclosing function.
#define NNTK_CUDA_TRY(expr)                                                                            \
    do {                                                                                               \
        cudaError_t e__ = (expr);                                                                      \
        if (e__ != cudaSuccess) {                                                                      \
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,           \
                                cudaGetErrorString(e__));                                              \
            return NNTK_ERR_CUDA_;                                                                     \
        }                                                                                              \
    } while (0)

// After a kernel launch: catches launch-configuration errors (asynchronous faults surface at the next sync) and counts
// the launch (nntk_cuda_launch_count(): bench.py's `gpu_launches`).  NNTK_LAUNCH_CHECK_N: `n` launches behind one check.
#define NNTK_LAUNCH_CHECK(name) NNTK_LAUNCH_CHECK_N(name, 1)
#define NNTK_LAUNCH_CHECK_N(name, n)                                                                   \
    do {                                                                                               \
        nntk_cuda_count_launch(n);                                                                     \
        cudaError_t e__ = cudaGetLastError();                                                          \
        if (e__ != cudaSuccess) {                                                                      \
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "launch %s failed: %s", name, cudaGetErrorString(e__)); \
            return NNTK_ERR_CUDA_;                                                                     \
        }                                                                                              \
    } while (0)

static inline cudaStream_t nntk_stream() { return (cudaStream_t)nntk_cuda_stream(); }

// ---- activations, reference arithmetic (layers/activation_default.c) -------------------------------------------
// IEEE expf / tanhf / division: the build never passes --use_fast_math.
__device__ __forceinline__ float nntk_sigmoid(float x) {
    // activation_default.c:28-33: neg, expf, +1, reciprocal
    float t = expf(-x);
    t = t + 1.0f;
    return 1.0f / t;
}
__device__ __forceinline__ float nntk_act(int kind, float a, float x) {
    switch (kind) {
    case NNTK_ACT_SIGMOID: return nntk_sigmoid(x);
    case NNTK_ACT_TANH: return tanhf(x);                       // activation_default.c:65-67
    case NNTK_ACT_RELU: {                                      // activation_default.c:123-129
        float t = fmaxf(x, 0.0f);
        return (a != 1.0f) ? t * a : t;
    }
    default: return x;                                         // identity / none
    }
}
// layers/batch_norm.c:151-162 -- six separately rounded fp32 operations (__f*_rn blocks FMA contraction).
__device__ __forceinline__ float nntk_bn(float x, float gamma, float beta, float mean, float var, float eps) {
    float x_mu = __fsub_rn(x, mean);
    float sd = sqrtf(__fadd_rn(var, eps));
    float xn = __fdiv_rn(x_mu, sd);
    return __fadd_rn(__fmul_rn(xn, gamma), beta);
}
