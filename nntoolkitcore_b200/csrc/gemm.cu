// gemm.cu -- K2/K3, fp32 path: (implicit) GEMM + bias + BatchNorm + activation in one kernel.
//
// Replaces, for the inference path, the reference's Conv1d triple loop of K-tap dot products (layers/conv_1d.c:122-147),
// the per-row Eigen GEMV of Dense / TimeDistributedDense (layers/dense.c:122-133, time_distributed_dense.c:52-58) and
// the per-timestep x*W GEMV of the recurrent cells (rnn.c:157, gru.c:144, lstm.c:201), plus the separate BatchNorm and
// activation passes that follow a Conv1d (batch_norm.c:166, activation.c:23).
//
// Implicit GEMM: with channels-last activations, im2col row x of a Conv1d is the CONTIGUOUS span
// in[x*stride : x*stride + K, :], so A is never materialised: row r of slab s starts at A + s*slab + r*lda and is
// K*Cin floats long (overlapping rows).  The reference's transpose pass (conv_1d.c:125) disappears.
//
// This file is the NNTK_MATH_FP32 mode: FFMA on the CUDA cores, fp32 operands and accumulators, 1e-5 parity with the
// reference.  128x128x16 CTA tile, 8x8 register micro-tile per thread, register-staged double buffering.
// Roofline: fp32 FFMA peak (148 SMs x 128 lanes x 2 flop x clock).  The tcgen05 kind::tf32 mode lives in
// gemm_tc.cu and shares the argument struct and the epilogue definition.
#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16;
constexpr int AS_LD = BM + 4;   // padded leading dimension of the transposed A tile

struct GemmDev {
    const float *A;
    long long a_slab;
    int lda, rows;
    long long M;        // batch * rows
    int N, K;
    const float *B;
    const float *bias;
    const float *bn;    // gamma|beta|mean|var or null
    float bn_eps;
    int act;
    float act_a;
    float *C;
    long long c_slab;
    int ldc;
    int n_tiles;
};

__device__ __forceinline__ float epilogue(const GemmDev &g, float v, int n) {
    if (g.bias) v += __ldg(g.bias + n);
    if (g.bn) v = nntk_bn(v, __ldg(g.bn + n), __ldg(g.bn + g.N + n), __ldg(g.bn + 2 * g.N + n),
                          __ldg(g.bn + 3 * g.N + n), g.bn_eps);
    return nntk_act(g.act, g.act_a, v);
}

template <bool VEC_A, bool VEC_B>
__global__ void __launch_bounds__(256, 2) gemm_fp32_kernel(GemmDev g) {
    __shared__ __align__(16) float As[2][BK][AS_LD];
    __shared__ __align__(16) float Bs[2][BK][BN];

    const int tid = threadIdx.x;
    const int n_tile = blockIdx.x % g.n_tiles;
    const long long m_tile = blockIdx.x / g.n_tiles;
    const long long m0 = m_tile * BM;
    const int n0 = n_tile * BN;

    // ---- global -> register staging maps ---------------------------------------------------------------------------
    // A tile: 128 rows x 16 k.  thread -> rows (tid>>2) and (tid>>2)+64, k quad (tid&3)*4
    const int a_kq = (tid & 3) * 4;
    const float *a_ptr[2];
    bool a_ok[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const long long gr = m0 + (tid >> 2) + 64 * i;
        a_ok[i] = gr < g.M;
        const long long s = a_ok[i] ? gr / g.rows : 0;
        const long long r = a_ok[i] ? gr - s * g.rows : 0;
        a_ptr[i] = g.A + s * g.a_slab + r * (long long)g.lda;
    }
    // B tile: 16 k x 128 n.  thread -> k rows (tid>>5) and (tid>>5)+8, n quad (tid&31)*4
    const int b_k = tid >> 5, b_n = n0 + (tid & 31) * 4;

    float4 ra[2], rb[2];
    auto load_tiles = [&](int k0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int k = k0 + a_kq;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a_ok[i]) {
                if (VEC_A && k + 3 < g.K) {
                    v = __ldg(reinterpret_cast<const float4 *>(a_ptr[i] + k));
                } else {
                    if (k < g.K) v.x = __ldg(a_ptr[i] + k);
                    if (k + 1 < g.K) v.y = __ldg(a_ptr[i] + k + 1);
                    if (k + 2 < g.K) v.z = __ldg(a_ptr[i] + k + 2);
                    if (k + 3 < g.K) v.w = __ldg(a_ptr[i] + k + 3);
                }
            }
            ra[i] = v;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int k = k0 + b_k + 8 * i;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k < g.K) {
                const float *p = g.B + (long long)k * g.N + b_n;
                if (VEC_B && b_n + 3 < g.N) {
                    v = __ldg(reinterpret_cast<const float4 *>(p));
                } else {
                    if (b_n < g.N) v.x = __ldg(p);
                    if (b_n + 1 < g.N) v.y = __ldg(p + 1);
                    if (b_n + 2 < g.N) v.z = __ldg(p + 2);
                    if (b_n + 3 < g.N) v.w = __ldg(p + 3);
                }
            }
            rb[i] = v;
        }
    };
    auto store_tiles = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int m = (tid >> 2) + 64 * i;
            As[buf][a_kq + 0][m] = ra[i].x;
            As[buf][a_kq + 1][m] = ra[i].y;
            As[buf][a_kq + 2][m] = ra[i].z;
            As[buf][a_kq + 3][m] = ra[i].w;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i)
            *reinterpret_cast<float4 *>(&Bs[buf][b_k + 8 * i][(tid & 31) * 4]) = rb[i];
    };

    // ---- compute maps: thread (ty, tx) owns rows {4ty..4ty+3, 64+4ty..} x cols {4tx..4tx+3, 64+4tx..} ---------------
    const int tx = tid & 15, ty = tid >> 4;
    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    const int k_iters = (g.K + BK - 1) / BK;
    load_tiles(0);
    store_tiles(0);
    __syncthreads();
    for (int it = 0; it < k_iters; ++it) {
        const int buf = it & 1;
        if (it + 1 < k_iters) load_tiles((it + 1) * BK);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][k][4 * ty]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][k][64 + 4 * ty]);
            const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[buf][k][4 * tx]);
            const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[buf][k][64 + 4 * tx]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (it + 1 < k_iters) {
            store_tiles(buf ^ 1);
            __syncthreads();
        }
    }

    // ---- epilogue: bias -> BatchNorm -> activation -> store ----------------------------------------------------------
    const bool vec_c = ((g.ldc & 3) == 0) && ((g.c_slab & 3) == 0) && ((((uintptr_t)g.C) & 15) == 0);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const long long gr = m0 + (i < 4 ? 4 * ty + i : 64 + 4 * ty + (i - 4));
        if (gr >= g.M) continue;
        const long long s = gr / g.rows;
        const long long r = gr - s * g.rows;
        float *crow = g.C + s * g.c_slab + r * (long long)g.ldc;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int n = n0 + 64 * h + 4 * tx;
            if (n >= g.N) continue;
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = (n + j < g.N) ? epilogue(g, acc[i][4 * h + j], n + j) : 0.f;
            if (vec_c && n + 3 < g.N) {
                *reinterpret_cast<float4 *>(crow + n) = make_float4(v[0], v[1], v[2], v[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (n + j < g.N) crow[n + j] = v[j];
            }
        }
    }
}

// [rows][C] -> [rows][Cpad], zero filled: gives rows a 16-byte aligned pitch so TMA can address them (Cin = 257 -> 260)
__global__ void __launch_bounds__(256) pad_channels_kernel(const float *__restrict__ in, float *__restrict__ out, long long rows, int C, int Cpad) {
    // one 16-byte store per thread and iteration (Cpad % 4 == 0, out 16-byte aligned); (row, column) advance incrementally
    const int cp4 = Cpad >> 2;
    const long long total4 = rows * cp4, stride = (long long)gridDim.x * blockDim.x;
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long r = i / cp4;
    int c4 = (int)(i - r * cp4);
    const long long dr = stride / cp4;
    const int dc = (int)(stride - dr * cp4);
    for (; i < total4; i += stride) {
        const int c = 4 * c4;
        const float *src = in + r * C + c;
        float4 v;
        v.x = c < C ? __ldg(src) : 0.0f;
        v.y = c + 1 < C ? __ldg(src + 1) : 0.0f;
        v.z = c + 2 < C ? __ldg(src + 2) : 0.0f;
        v.w = c + 3 < C ? __ldg(src + 3) : 0.0f;
        reinterpret_cast<float4 *>(out)[i] = v;
        r += dr; c4 += dc;
        if (c4 >= cp4) { c4 -= cp4; ++r; }
    }
}

// conv weights W[Cout][Cin][K] -> B[(k*cin_pad + ci)][co], rows ci in [cin_src, cin) zero
__global__ void pack_conv_weights_kernel(const float *__restrict__ W, float *__restrict__ B, int cout, int cin, int ks,
                                         int cin_src) {
    const long long total = (long long)cout * cin * ks;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        // i indexes the destination: i = (k*cin + ci)*cout + co  (coalesced writes)
        const int co = (int)(i % cout);
        const long long rest = i / cout;
        const int ci = (int)(rest % cin);
        const int k = (int)(rest / cin);
        B[i] = ci < cin_src ? __ldg(W + ((long long)co * cin_src + ci) * ks + k) : 0.0f;
    }
}

// conv weights with an input BatchNorm folded in: B[(k*cin + ci)][co] = W[co][ci][k] * gamma[ci] / sqrt(var[ci] + eps)
__global__ void pack_conv_weights_bn_kernel(const float *__restrict__ W, const float *__restrict__ bn, float eps, float *__restrict__ B, int cout,
                                            int cin, int ks, int cin_src) {
    const long long total = (long long)cout * cin * ks;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int co = (int)(i % cout);
        const long long rest = i / cout;
        const int ci = (int)(rest % cin);
        const int k = (int)(rest / cin);
        float v = 0.0f;
        if (ci < cin_src) v = __ldg(W + ((long long)co * cin_src + ci) * ks + k) * (__ldg(bn + ci) / sqrtf(__ldg(bn + 3 * cin_src + ci) + eps));
        B[i] = v;
    }
}
// bias'[co] = b[co] + sum_{ci,k} W[co][ci][k] * (beta[ci] - mean[ci] * a[ci]); one warp per output channel, double accumulation
__global__ void fold_bn_bias_kernel(const float *__restrict__ W, const float *__restrict__ b, const float *__restrict__ bn, float eps,
                                    float *__restrict__ bias_out, int cout, int cin, int ks) {
    const int co = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (co >= cout) return;
    double acc = 0.0;
    for (int ci = lane; ci < cin; ci += 32) {
        const float a = __ldg(bn + ci) / sqrtf(__ldg(bn + 3 * cin + ci) + eps);
        const double shift = (double)__ldg(bn + cin + ci) - (double)__ldg(bn + 2 * cin + ci) * (double)a;
        double w = 0.0;
        for (int k = 0; k < ks; ++k) w += (double)__ldg(W + ((long long)co * cin + ci) * ks + k);
        acc += w * shift;
    }
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) bias_out[co] = (float)((double)__ldg(b + co) + acc);
}

}  // namespace

extern "C" int nntk_cuda_pack_conv_weights_bn(const float *d_W, const float *d_b, const float *d_bn, float eps, float *d_B, float *d_bias_out,
                                              int cout, int cin, int cin_pad, int k) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    const long long total = (long long)cout * cin_pad * k;
    if (total <= 0) return 0;
    long long blocks = (total + 255) / 256;
    if (blocks > 4096) blocks = 4096;
    pack_conv_weights_bn_kernel<<<(unsigned)blocks, 256, 0, nntk_stream()>>>(d_W, d_bn, eps, d_B, cout, cin_pad, k, cin);
    fold_bn_bias_kernel<<<(unsigned)((cout + 7) / 8), 256, 0, nntk_stream()>>>(d_W, d_b, d_bn, eps, d_bias_out, cout, cin, k);
    NNTK_LAUNCH_CHECK_N("pack_conv_weights_bn", 2);
    return 0;
}

extern "C" int nntk_cuda_gemm_tf32(const nntk_gemm_args *g, int split);   // gemm_tc.cu
// 3xTF32 holds 1e-5 only while the tensor core's TRUNCATING fp32 accumulate stays small: its shrink grows linearly with the
// number of accumulation steps and depends on the sign statistics of the partial sums -- the fitted compensation
// (gemm_tc.cu X3_ACC_TRUNC_PER_K, zero-mean data) leaves ~5e-9 * K on monotonically growing sums (non-negative activations x
// constant-sign weights: 1.5e-5 at K = 3 000, tests/test_gpu_robustness.py).  1 536 keeps that worst case under 1e-5; longer
// reductions run the FFMA kernel.
#define NNTK_X3_MAX_K 1536

// ---- skinny GEMM: N <= 32 outputs per row (the 12-wide TimeDistributedDense of the KWS chain) in exact fp32 ------------------
// HBM-bound by construction (K floats read per N written), so the fp32-contract modes need no tensor core here: 3xTF32 spent a
// whole extra pass splitting A (C5 TDD: 4.3 ms split + 3.1 ms GEMM for 8.7 GB of algorithmic traffic).  Eight lanes share a row
// (each 16-byte chunk g + 8 i of it), W sits transposed in shared memory ([N][K]: the eight lanes read 128 contiguous bytes, the
// four row groups of a warp the same ones) and every chunk read serves four rows; 3 shuffles per output close the sums; bias,
// activation or the row softmax (activation_default.c:149-167: exp, sum, divide -- the sum taken over the eight lanes' shares)
// follow in registers, each lane finishing the outputs g, g + 8, ... of its rows.
namespace {
// a group of eight lanes works on RPG consecutive rows at once: every W chunk read from shared memory serves all of them
constexpr int SK_RPG = 4;
template <int NMAX, bool SOFTMAX>
__global__ void __launch_bounds__(256) gemm_skinny_kernel(const float *A, long long a_slab, int lda, int rows, long long M, int N, int K,
                                                          const float *__restrict__ B, const float *__restrict__ bias, int act, float act_a,
                                                          float *C, long long c_slab, int ldc) {
    extern __shared__ __align__(16) float s_wt[];            // [N][K]
    for (int i = threadIdx.x; i < N * K; i += blockDim.x) { const int n = i / K, k = i - n * K; s_wt[i] = __ldg(B + (long long)k * N + n); }
    __syncthreads();
    const int lane = threadIdx.x & 31, g = lane & 7;
    const long long groups = ((long long)gridDim.x * blockDim.x) >> 3;
    const int k4 = K >> 2;
    const long long m_pad = (M + 4 * SK_RPG - 1) / (4 * SK_RPG) * (4 * SK_RPG);      // all four groups of a warp run the same trip count (shuffles)
    for (long long r0 = (((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 3) * SK_RPG; r0 < m_pad; r0 += groups * SK_RPG) {
        const float4 *x[SK_RPG];
        bool ok[SK_RPG];
#pragma unroll
        for (int q = 0; q < SK_RPG; ++q) {
            const long long r = r0 + q;
            ok[q] = r < M;
            const long long slab = ok[q] ? r / rows : 0;
            const int row = ok[q] ? (int)(r - slab * rows) : 0;
            x[q] = reinterpret_cast<const float4 *>(A + slab * a_slab + (long long)row * lda);
        }
        float acc[SK_RPG][NMAX];
#pragma unroll
        for (int q = 0; q < SK_RPG; ++q)
#pragma unroll
            for (int n = 0; n < NMAX; ++n) acc[q][n] = 0.0f;
        for (int i = g; i < k4; i += 8) {
            float4 v[SK_RPG];
#pragma unroll
            for (int q = 0; q < SK_RPG; ++q) v[q] = ok[q] ? x[q][i] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int n = 0; n < NMAX; ++n) {
                if (n < N) {
                    const float4 w = *reinterpret_cast<const float4 *>(s_wt + n * K + 4 * i);
#pragma unroll
                    for (int q = 0; q < SK_RPG; ++q) acc[q][n] = fmaf(v[q].x, w.x, fmaf(v[q].y, w.y, fmaf(v[q].z, w.z, fmaf(v[q].w, w.w, acc[q][n]))));
                }
            }
        }
#pragma unroll
        for (int q = 0; q < SK_RPG; ++q) {
            // lane g ends up owning outputs g and g + 8 (.. + 24) of row q
            float mine[NMAX / 8];
#pragma unroll
            for (int n = 0; n < NMAX; ++n) {
                if (n < N) {
                    float s = acc[q][n];
                    s += __shfl_xor_sync(0xffffffffu, s, 1);
                    s += __shfl_xor_sync(0xffffffffu, s, 2);
                    s += __shfl_xor_sync(0xffffffffu, s, 4);
                    if ((n & 7) == g) mine[n >> 3] = s + (bias ? __ldg(bias + n) : 0.0f);
                }
            }
            float sum = 0.0f;
#pragma unroll
            for (int jj = 0; jj < NMAX / 8; ++jj) {
                const int n = g + 8 * jj;
                if (n < N) {
                    if (SOFTMAX) { mine[jj] = expf(mine[jj]); sum += mine[jj]; }
                    else mine[jj] = nntk_act(act, act_a, mine[jj]);
                }
            }
            if (SOFTMAX) {
                sum += __shfl_xor_sync(0xffffffffu, sum, 1);
                sum += __shfl_xor_sync(0xffffffffu, sum, 2);
                sum += __shfl_xor_sync(0xffffffffu, sum, 4);
            }
            if (ok[q]) {
                const long long r = r0 + q;
                const long long slab = r / rows;
                float *c = C + slab * c_slab + (long long)(r - slab * rows) * ldc;
#pragma unroll
                for (int jj = 0; jj < NMAX / 8; ++jj) {
                    const int n = g + 8 * jj;
                    if (n < N) c[n] = SOFTMAX ? mine[jj] / sum : mine[jj];
                }
            }
        }
    }
}
}  // namespace

// 0 done, 1 not eligible
static int gemm_skinny(const nntk_gemm_args *a, bool softmax) {
    static int on = -1;
    if (on < 0) { const char *e = getenv("NNTK_GEMM_SKINNY"); on = e ? atoi(e) : 1; }
    const int N = a->N, K = a->K;
    if (!on || N > 32 || (K & 3) || a->d_bn || (size_t)N * K * sizeof(float) > 96 * 1024 || (a->lda & 3) || (a->a_slab_stride & 3) ||
        (((uintptr_t)a->d_A) & 15) || (!softmax && (a->act == NNTK_ACT_SOFTMAX || a->act == NNTK_ACT_CUSTOM))) return 1;
    const long long M = (long long)a->batch * a->rows;
    if (M == 0) return 0;
    const size_t smem = (size_t)N * K * sizeof(float);
    long long blocks = ((M + SK_RPG - 1) / SK_RPG * 8 + 255) / 256;
    const long long cap = (long long)nntk_cuda_sm_count() * 8;
    if (blocks > cap) blocks = cap;
    cudaStream_t st = nntk_stream();
#define SKINNY_LAUNCH(NM, SM) do { \
        static bool attr = false; \
        if (!attr) { NNTK_CUDA_TRY(cudaFuncSetAttribute(gemm_skinny_kernel<NM, SM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024)); attr = true; } \
        gemm_skinny_kernel<NM, SM><<<(unsigned)blocks, 256, smem, st>>>(a->d_A, a->a_slab_stride, a->lda, a->rows, M, N, K, a->d_B, a->d_bias, a->act, \
                                                                       a->act_a, a->d_C, a->c_slab_stride, a->ldc); } while (0)
    if (N <= 16) { if (softmax) SKINNY_LAUNCH(16, true); else SKINNY_LAUNCH(16, false); }
    else { if (softmax) SKINNY_LAUNCH(32, true); else SKINNY_LAUNCH(32, false); }
#undef SKINNY_LAUNCH
    NNTK_LAUNCH_CHECK("gemm_skinny_kernel");
    return 0;
}

extern "C" void nntk_cuda_gemm_tc_row_softmax(int on);      // gemm_tc.cu
// C = softmax_rows(A B + bias) for N <= 32 on the tensor-core kernel, the softmax over each whole output row taken in the GEMM
// epilogue (TimeDistributedDense + softmax of the KWS chain: saves the second pass over the result).  Same expf / divide as the
// stand-alone softmax kernel.  Returns 0 done, 1 not eligible (FFMA mode, N > 32, K beyond the X3 cap, shape rules), < 0 error.
extern "C" int nntk_cuda_gemm_row_softmax(const nntk_gemm_args *a) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!a || !a->d_A || !a->d_B || !a->d_C || a->batch < 1 || a->rows < 1 || a->N < 1 || a->N > 32 || a->K < 1 || a->d_bn ||
        (a->act != NNTK_ACT_NONE && a->act != NNTK_ACT_IDENTITY)) return 1;
    const int mode = nntk_cuda_math_mode();
    if (mode != 1) return gemm_skinny(a, true);          // the fp32-contract modes: exact fp32, HBM-bound
    nntk_cuda_gemm_tc_row_softmax(1);
    rc = nntk_cuda_gemm_tf32(a, mode == 2);
    nntk_cuda_gemm_tc_row_softmax(0);
    return rc;
}

extern "C" int nntk_cuda_gemm(const nntk_gemm_args *a) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!a || !a->d_A || !a->d_B || !a->d_C || a->batch < 0 || a->rows < 0 || a->N < 1 || a->K < 1) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "gemm: bad arguments");
        return NNTK_ERR_ARG_;
    }
    if (a->act == NNTK_ACT_SOFTMAX || a->act == NNTK_ACT_CUSTOM) {
        nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "gemm: activation kind %d cannot be fused", a->act);
        return NNTK_ERR_UNSUPPORTED_;
    }
    const long long M = (long long)a->batch * a->rows;
    if (M == 0) return 0;
    // TF32X3 promises the fp32 contract (1e-5).  The part of the tensor core's truncating accumulate that the epilogue cannot
    // compensate grows linearly with the reduction length (measured rel-L2 1.5e-6 at K = 1 280, 3.0e-6 at 2 560, 8.0e-6 at
    // 7 168): beyond NNTK_X3_MAX_K the mode runs the FFMA kernel.  Every BASELINE shape has K <= 1 285.
    const int mode = nntk_cuda_math_mode();
    if (mode != 1 && a->N <= 32) {       // narrow outputs in the fp32-contract modes: the exact, HBM-bound skinny kernel
        rc = gemm_skinny(a, false);
        if (rc != 1) return rc;
    }
    if (mode == 1 || (mode == 2 && a->K <= NNTK_X3_MAX_K)) {
        rc = nntk_cuda_gemm_tf32(a, mode == 2);
        if (rc != 1) return rc;   // 1 = shape not eligible for the tensor-core kernel: use the fp32 kernel
    }
    GemmDev g;
    g.A = a->d_A; g.a_slab = a->a_slab_stride; g.lda = a->lda; g.rows = a->rows; g.M = M;
    g.N = a->N; g.K = a->K; g.B = a->d_B; g.bias = a->d_bias; g.bn = a->d_bn; g.bn_eps = a->bn_eps;
    g.act = a->act; g.act_a = a->act_a; g.C = a->d_C; g.c_slab = a->c_slab_stride; g.ldc = a->ldc;
    g.n_tiles = (a->N + BN - 1) / BN;
    const long long m_tiles = (M + BM - 1) / BM;
    const long long grid = m_tiles * g.n_tiles;
    if (grid > 0x7fffffffLL) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "gemm: problem too large for one launch");
        return NNTK_ERR_ARG_;
    }
    const bool vec_a = ((a->lda & 3) == 0) && ((a->a_slab_stride & 3) == 0) && ((((uintptr_t)a->d_A) & 15) == 0);
    const bool vec_b = ((a->N & 3) == 0) && ((((uintptr_t)a->d_B) & 15) == 0);
    cudaStream_t st = nntk_stream();
    if (vec_a && vec_b) gemm_fp32_kernel<true, true><<<(unsigned)grid, 256, 0, st>>>(g);
    else if (vec_a) gemm_fp32_kernel<true, false><<<(unsigned)grid, 256, 0, st>>>(g);
    else if (vec_b) gemm_fp32_kernel<false, true><<<(unsigned)grid, 256, 0, st>>>(g);
    else gemm_fp32_kernel<false, false><<<(unsigned)grid, 256, 0, st>>>(g);
    NNTK_LAUNCH_CHECK("gemm_fp32_kernel");
    return 0;
}

extern "C" int nntk_cuda_pack_conv_weights_padded(const float *d_W, float *d_B, int cout, int cin, int cin_pad, int k) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    const long long total = (long long)cout * cin_pad * k;
    if (total <= 0) return 0;
    long long blocks = (total + 255) / 256;
    if (blocks > 4096) blocks = 4096;
    pack_conv_weights_kernel<<<(unsigned)blocks, 256, 0, nntk_stream()>>>(d_W, d_B, cout, cin_pad, k, cin);
    NNTK_LAUNCH_CHECK("pack_conv_weights_kernel");
    return 0;
}
extern "C" int nntk_cuda_pack_conv_weights(const float *d_W, float *d_B, int cout, int cin, int k) {
    return nntk_cuda_pack_conv_weights_padded(d_W, d_B, cout, cin, cin, k);
}
extern "C" int nntk_cuda_pad_channels(const float *d_in, float *d_out, long long rows, int C, int Cpad) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    const long long total = rows * Cpad;
    if (total <= 0) return 0;
    if ((Cpad & 3) || C > Cpad || (((uintptr_t)d_out) & 15)) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "pad_channels: Cpad %d must be a multiple of 4 >= C %d, output 16-byte aligned", Cpad, C);
        return NNTK_ERR_ARG_;
    }
    long long blocks = (total / 4 + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    pad_channels_kernel<<<(unsigned)blocks, 256, 0, nntk_stream()>>>(d_in, d_out, rows, C, Cpad);
    NNTK_LAUNCH_CHECK("pad_channels_kernel");
    return 0;
}
