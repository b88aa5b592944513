// recurrent.cu -- K3 + K4: RNN / GRU / LSTM inference forward over [batch, T, in] (replaces layers/rnn.c:144-166,223-244,
// layers/gru.c:129-204 and layers/lstm.c:185-268 of the reference for the inference path).
//
// The reference runs, per sequence and per timestep, two GEMVs (x_t*W and h*U) plus ~15 element-wise passes over H floats.
// Here:
//   1. the input projection is hoisted out of the time loop: xw[b,t,:] = x[b,t,:]*W + b_i for ALL timesteps in one GEMM
//      (csrc/gemm.cu / gemm_tc.cu), M = batch*T rows;
//   2. each timestep is ONE kernel: the recurrent GEMM h_{t-1}[B,H] x U[H,G*H] with the gate nonlinearities, the state
//      update and the return_sequences store fused into its epilogue.  A thread owns all G gate columns of its hidden
//      units, so the gate math never leaves registers.  h_{t-1} is read from the previous output row (or a ping-pong
//      buffer), h_t is written to the next: no extra state copies.
// Gate order and formulas follow the reference exactly: GRU [z, r, h~] with reset-after (gru.c:140-186), LSTM [i, f, g, o]
// (lstm.c:199-238), RNN (rnn.c:156-165).
#include <cooperative_groups.h>
#include <stdlib.h>
#include "common.cuh"

namespace {

constexpr int SM_ROWS = 128, SU = 32, SK = 16;
constexpr int SA_LD = SM_ROWS + 4;

struct StepDev {
    int B, H;
    const float *h_prev; long long ld_prev;   // [B, H] view with row stride
    float *h_next; long long ld_next;
    float *c;                                  // [B, H] (LSTM) in/out
    const float *U;                            // [H][G*H]
    const float *bh;                           // [G*H] or null (not added)
    const float *xw; long long ld_xw;          // row b: xw + b*ld_xw, G*H floats (b_i already added)
    int act[5];
    float act_a[5];
    int u_tiles;
};

template <int G>
__global__ void __launch_bounds__(256, 2) recurrent_step_kernel(StepDev p) {
    __shared__ __align__(16) float As[2][SK][SA_LD];
    __shared__ __align__(16) float Us[2][SK][G][SU];
    const int tid = threadIdx.x;
    const int u_tile = blockIdx.x % p.u_tiles;
    const int m_tile = blockIdx.x / p.u_tiles;
    const int m0 = m_tile * SM_ROWS, u0 = u_tile * SU;
    const int GH = G * p.H;
    const bool vec_a = ((p.ld_prev & 3) == 0) && ((((uintptr_t)p.h_prev) & 15) == 0);
    const bool vec_u = ((p.H & 3) == 0) && ((((uintptr_t)p.U) & 15) == 0);

    const int a_kq = (tid & 3) * 4;
    const float *a_ptr[2];
    bool a_ok[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int row = m0 + (tid >> 2) + 64 * i;
        a_ok[i] = row < p.B;
        a_ptr[i] = p.h_prev + (long long)(a_ok[i] ? row : 0) * p.ld_prev;
    }
    // U tile: SK x G x 8 quads; quad q -> k = q / (8G), g = (q / 8) % G, u4 = (q % 8) * 4
    constexpr int UQ = SK * G * 8;
    constexpr int UQ_PER_THREAD = (UQ + 255) / 256;

    float4 ra[2], ru[UQ_PER_THREAD];
    auto load_tiles = [&](int k0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int k = k0 + a_kq;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a_ok[i]) {
                if (vec_a && k + 3 < p.H) v = *reinterpret_cast<const float4 *>(a_ptr[i] + k);
                else {
                    if (k < p.H) v.x = a_ptr[i][k];
                    if (k + 1 < p.H) v.y = a_ptr[i][k + 1];
                    if (k + 2 < p.H) v.z = a_ptr[i][k + 2];
                    if (k + 3 < p.H) v.w = a_ptr[i][k + 3];
                }
            }
            ra[i] = v;
        }
#pragma unroll
        for (int j = 0; j < UQ_PER_THREAD; ++j) {
            const int q = tid + 256 * j;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (q < UQ) {
                const int k = k0 + q / (8 * G), gg = (q >> 3) % G, u = u0 + (q & 7) * 4;
                if (k < p.H) {
                    const float *src = p.U + (long long)k * GH + gg * p.H + u;
                    if (vec_u && u + 3 < p.H) v = __ldg(reinterpret_cast<const float4 *>(src));
                    else {
                        if (u < p.H) v.x = __ldg(src);
                        if (u + 1 < p.H) v.y = __ldg(src + 1);
                        if (u + 2 < p.H) v.z = __ldg(src + 2);
                        if (u + 3 < p.H) v.w = __ldg(src + 3);
                    }
                }
            }
            ru[j] = v;
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
        for (int j = 0; j < UQ_PER_THREAD; ++j) {
            const int q = tid + 256 * j;
            if (q < UQ) *reinterpret_cast<float4 *>(&Us[buf][q / (8 * G)][(q >> 3) % G][(q & 7) * 4]) = ru[j];
        }
    };

    const int ux = tid & 15, ty = tid >> 4;
    float acc[8][G][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int gg = 0; gg < G; ++gg) acc[i][gg][0] = acc[i][gg][1] = 0.f;

    const int k_iters = (p.H + SK - 1) / SK;
    load_tiles(0);
    store_tiles(0);
    __syncthreads();
    for (int it = 0; it < k_iters; ++it) {
        const int buf = it & 1;
        if (it + 1 < k_iters) load_tiles((it + 1) * SK);
#pragma unroll
        for (int k = 0; k < SK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[buf][k][4 * ty]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[buf][k][64 + 4 * ty]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
#pragma unroll
            for (int gg = 0; gg < G; ++gg) {
                const float2 b = *reinterpret_cast<const float2 *>(&Us[buf][k][gg][2 * ux]);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    acc[i][gg][0] = fmaf(a[i], b.x, acc[i][gg][0]);
                    acc[i][gg][1] = fmaf(a[i], b.y, acc[i][gg][1]);
                }
            }
        }
        if (it + 1 < k_iters) {
            store_tiles(buf ^ 1);
            __syncthreads();
        }
    }

    // ---- fused gate epilogue ------------------------------------------------------------------------------------------
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = m0 + (i < 4 ? 4 * ty + i : 64 + 4 * ty + (i - 4));
        if (row >= p.B) continue;
        const float *xw = p.xw + (long long)row * p.ld_xw;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
            const int u = u0 + 2 * ux + e;
            if (u >= p.H) continue;
            float hu[G], xv[G];
#pragma unroll
            for (int gg = 0; gg < G; ++gg) {
                hu[gg] = acc[i][gg][e];
                if (p.bh) hu[gg] = __fadd_rn(hu[gg], __ldg(p.bh + gg * p.H + u));
                xv[gg] = xw[gg * p.H + u];
            }
            float hn;
            if (G == 3) {   // gru.c:156-186
                const float hp = p.h_prev[(long long)row * p.ld_prev + u];
                const float z = nntk_act(p.act[0], p.act_a[0], __fadd_rn(xv[0], hu[0]));
                const float r = nntk_act(p.act[1], p.act_a[1], __fadd_rn(xv[1], hu[1]));
                const float hh = nntk_act(p.act[2], p.act_a[2], __fadd_rn(__fmul_rn(r, hu[2]), xv[2]));
                const float one_minus_z = __fadd_rn(-z, 1.0f);
                hn = __fadd_rn(__fmul_rn(one_minus_z, hh), __fmul_rn(z, hp));
            } else if (G == 4) {   // lstm.c:199-238
                float *cp = p.c + (long long)row * p.H + u;
                const float ig = nntk_act(p.act[0], p.act_a[0], __fadd_rn(xv[0], hu[0]));
                const float fg = nntk_act(p.act[1], p.act_a[1], __fadd_rn(xv[1], hu[1]));
                const float gg_ = nntk_act(p.act[2], p.act_a[2], __fadd_rn(xv[2], hu[2]));
                const float og = nntk_act(p.act[3], p.act_a[3], __fadd_rn(xv[3], hu[3]));
                const float cn = __fadd_rn(__fmul_rn(fg, *cp), __fmul_rn(ig, gg_));
                *cp = cn;
                hn = __fmul_rn(og, nntk_act(p.act[4], p.act_a[4], cn));
            } else {   // rnn.c:156-165
                hn = nntk_act(p.act[0], p.act_a[0], __fadd_rn(hu[0], xv[0]));
            }
            p.h_next[(long long)row * p.ld_next + u] = hn;
        }
    }
}


// ---- K4 persistent cluster kernel (small / medium batch) ---------------------------------------------------------------
// One thread-block CLUSTER of C = H/32 CTAs owns BS sequences for ALL T timesteps (one launch per layer instead of T):
//   * CTA j keeps the recurrent-weight columns of hidden units [32 j, 32 j + 32) -- all G gates -- resident in shared
//     memory for the whole sequence (U slice: H x 32 G floats, row stride padded by 4 floats so a quarter-warp's 16-byte
//     loads of 8 consecutive k rows fall into distinct banks);
//   * h_{t-1} (BS x H) lives in a double-buffered shared-memory vector in EVERY CTA of the cluster;
//   * per step: warp w computes the 4 G pre-activation columns of its 4 hidden units; the 32 lanes split K = H
//     (k = lane + 32 i), FFMA into registers, then warp-shuffle (xor butterfly) reductions; lane 4 b + u applies the gate
//     nonlinearities / state update of unit u of sequence b (reference op order, see recurrent_step_kernel), stores
//     h_t to HBM (return_sequences) and into the h buffer of ALL C CTAs through distributed shared memory;
//   * one cluster barrier (arrive.release / wait.acquire) per timestep publishes h_t.
// The input projection xw = x W + b_i is hoisted (same GEMM as the per-step path).  fp32 FFMA throughout: 1e-5 parity.
constexpr int CL_THREADS = 256, CL_WARPS = 8, CL_UPW = 4, CL_UNITS = CL_WARPS * CL_UPW;   // 32 hidden units per CTA

struct ClusterDev {
    int B, H, T, C;                 // C = cluster size = H / 32
    int return_sequences;
    const float *U, *bh;            // [H][G*H], [G*H] or null
    const float *xw;                // [B][T][G*H]
    float *h_state, *c_state;       // [B][H] in/out
    float *out;                     // [B][T][H] or [B][H]
    int act[5];
    float act_a[5];
};

template <int G, int BS>
__global__ void __launch_bounds__(CL_THREADS, 1) recurrent_cluster_kernel(ClusterDev p) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    constexpr int NCOLW = G * CL_UPW;          // columns per warp, ordered [gate][unit]
    constexpr int NCOL = NCOLW * CL_WARPS;     // columns per CTA
    constexpr int LDU = NCOL + 4;
    extern __shared__ __align__(16) float smem[];
    float *Us = smem;                                  // [H][LDU]
    float *hb = smem + (size_t)p.H * LDU;              // [2][BS][H]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rank = (int)cluster.block_rank();
    const int group = blockIdx.x / p.C;                // batch group of this cluster
    const int b0 = group * BS;
    const int nb = min(BS, p.B - b0);
    const int H = p.H, GH = G * H;
    const int unit0 = rank * CL_UNITS;

    // U slice -> shared memory: column (w, g, u) <- U[k][g*H + unit0 + 4w + u]
    for (int idx = tid; idx < H * NCOL; idx += CL_THREADS) {
        const int k = idx / NCOL, c = idx - k * NCOL;
        const int w = c / NCOLW, r = c - w * NCOLW, g = r / CL_UPW, u = r - g * CL_UPW;
        Us[k * LDU + c] = __ldg(p.U + (size_t)k * GH + g * H + unit0 + CL_UPW * w + u);
    }
    for (int idx = tid; idx < BS * H; idx += CL_THREADS) {
        const int b = idx / H, k = idx - b * H;
        hb[idx] = b < nb ? p.h_state[(size_t)(b0 + b) * H + k] : 0.0f;
    }
    // gate lane: lane = 4 b + u  (b < BS)
    const int gb = lane >> 2, gu = lane & 3;
    const bool gate_lane = gb < nb && lane < 4 * BS;
    const int unit = unit0 + CL_UPW * warp + gu;
    float bh[G], cstate = 0.0f;
#pragma unroll
    for (int g = 0; g < G; ++g) bh[g] = (gate_lane && p.bh) ? __ldg(p.bh + g * H + unit) : 0.0f;
    if (G == 4 && gate_lane) cstate = p.c_state[(size_t)(b0 + gb) * H + unit];
    float *remote[16];
#pragma unroll
    for (int r = 0; r < 16; ++r) remote[r] = r < p.C ? cluster.map_shared_rank(hb, r) : nullptr;
    cluster.sync();   // every CTA's buffers are initialised before anybody writes remotely

    const float *urow = Us + lane * LDU + warp * NCOLW;
    const float *xw_row = p.xw + (size_t)(b0 + (gate_lane ? gb : 0)) * p.T * GH + unit;
    float hn = 0.0f;
    for (int t = 0; t < p.T; ++t) {
        const float *hcur = hb + (t & 1) * BS * H;
        const int nxt = ((t + 1) & 1) * BS * H;
        float xv[G];
#pragma unroll
        for (int g = 0; g < G; ++g) xv[g] = gate_lane ? __ldg(xw_row + (size_t)t * GH + g * H) : 0.0f;
        float acc[BS][NCOLW];
#pragma unroll
        for (int b = 0; b < BS; ++b)
#pragma unroll
            for (int c = 0; c < NCOLW; ++c) acc[b][c] = 0.0f;
#pragma unroll 2
        for (int k0 = 0; k0 < H; k0 += 32) {
            float hv[BS];
#pragma unroll
            for (int b = 0; b < BS; ++b) hv[b] = hcur[b * H + k0 + lane];
            const float *ur = urow + k0 * LDU;
#pragma unroll
            for (int c4 = 0; c4 < NCOLW / 4; ++c4) {
                const float4 u = *reinterpret_cast<const float4 *>(ur + 4 * c4);
#pragma unroll
                for (int b = 0; b < BS; ++b) {
                    acc[b][4 * c4 + 0] = fmaf(hv[b], u.x, acc[b][4 * c4 + 0]);
                    acc[b][4 * c4 + 1] = fmaf(hv[b], u.y, acc[b][4 * c4 + 1]);
                    acc[b][4 * c4 + 2] = fmaf(hv[b], u.z, acc[b][4 * c4 + 2]);
                    acc[b][4 * c4 + 3] = fmaf(hv[b], u.w, acc[b][4 * c4 + 3]);
                }
            }
        }
        // warp-shuffle reductions: every lane ends with every column sum
#pragma unroll
        for (int b = 0; b < BS; ++b)
#pragma unroll
            for (int c = 0; c < NCOLW; ++c) {
                float v = acc[b][c];
                v += __shfl_xor_sync(0xffffffffu, v, 16);
                v += __shfl_xor_sync(0xffffffffu, v, 8);
                v += __shfl_xor_sync(0xffffffffu, v, 4);
                v += __shfl_xor_sync(0xffffffffu, v, 2);
                v += __shfl_xor_sync(0xffffffffu, v, 1);
                acc[b][c] = v;
            }
        // this lane's (b, u) pre-activations
        float hu[G];
#pragma unroll
        for (int g = 0; g < G; ++g) hu[g] = 0.0f;
#pragma unroll
        for (int b = 0; b < BS; ++b)
#pragma unroll
            for (int u = 0; u < CL_UPW; ++u)
                if (lane == 4 * b + u) {
#pragma unroll
                    for (int g = 0; g < G; ++g) hu[g] = acc[b][g * CL_UPW + u];
                }
        if (gate_lane) {
#pragma unroll
            for (int g = 0; g < G; ++g)
                if (p.bh) hu[g] = __fadd_rn(hu[g], bh[g]);
            if (G == 3) {   // gru.c:156-186
                const float hp = hcur[gb * H + unit];
                const float z = nntk_act(p.act[0], p.act_a[0], __fadd_rn(xv[0], hu[0]));
                const float r = nntk_act(p.act[1], p.act_a[1], __fadd_rn(xv[1], hu[1]));
                const float hh = nntk_act(p.act[2], p.act_a[2], __fadd_rn(__fmul_rn(r, hu[2]), xv[2]));
                hn = __fadd_rn(__fmul_rn(__fadd_rn(-z, 1.0f), hh), __fmul_rn(z, hp));
            } else if (G == 4) {   // lstm.c:199-238
                const float ig = nntk_act(p.act[0], p.act_a[0], __fadd_rn(xv[0], hu[0]));
                const float fg = nntk_act(p.act[1], p.act_a[1], __fadd_rn(xv[1], hu[1]));
                const float gg = nntk_act(p.act[2], p.act_a[2], __fadd_rn(xv[2], hu[2]));
                const float og = nntk_act(p.act[3], p.act_a[3], __fadd_rn(xv[3], hu[3]));
                cstate = __fadd_rn(__fmul_rn(fg, cstate), __fmul_rn(ig, gg));
                hn = __fmul_rn(og, nntk_act(p.act[4], p.act_a[4], cstate));
            } else {   // rnn.c:156-165
                hn = nntk_act(p.act[0], p.act_a[0], __fadd_rn(hu[0], xv[0]));
            }
            if (p.return_sequences) p.out[((size_t)(b0 + gb) * p.T + t) * H + unit] = hn;
            const int off = nxt + gb * H + unit;
#pragma unroll
            for (int r = 0; r < 16; ++r)
                if (r < p.C) remote[r][off] = hn;
        }
        cluster.sync();   // h_t is visible in every CTA; everybody is done reading h_{t-1}
    }
    if (gate_lane && p.T > 0) {
        p.h_state[(size_t)(b0 + gb) * H + unit] = hn;
        if (G == 4) p.c_state[(size_t)(b0 + gb) * H + unit] = cstate;
        if (!p.return_sequences) p.out[(size_t)(b0 + gb) * H + unit] = hn;
    }
}

}  // namespace

extern "C" int nntk_cuda_recurrent_fused_plan(const nntk_recurrent_args *r, int nb);          // recurrent_fused.cu
extern "C" size_t nntk_cuda_recurrent_workspace(const nntk_recurrent_args *r) {
    if (!r) return 0;
    // tensor-core mode, fused persistent kernel with the projection computed in-kernel: only the minimum the fallback paths need
    if (nntk_cuda_math_mode() == 1 && r->cell != 1 && nntk_cuda_recurrent_fused_plan(r, r->batch) == 1)
        return (size_t)32 * r->T * r->cell * r->H * sizeof(float) + (size_t)6 * 32 * r->H * sizeof(float) + 256;
    const size_t G = (size_t)r->cell;
    size_t xw = (size_t)((r->batch + 31) / 32 * 32) * r->T * G * r->H * sizeof(float);   // 32-sequence blocks
    size_t pp = (size_t)6 * r->batch * r->H * sizeof(float);   // exact h ping-pong + operand ping-pong (3xTF32: hi and lo parts)
    return xw + pp + 256;
}

template <int G, int BS>
static int launch_cluster(const ClusterDev &p, int groups, size_t smem, cudaStream_t st) {
    auto kern = recurrent_cluster_kernel<G, BS>;
    static size_t attr_smem = 0;
    static bool attr_np = false;
    if (smem > attr_smem) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_smem = smem;
    }
    if (p.C > 8 && !attr_np) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
        attr_np = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(groups * p.C));
    cfg.blockDim = dim3(CL_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    NNTK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, p));
    nntk_cuda_count_launch(1);
    return 0;
}

// Eligibility of the persistent cluster kernel: H a multiple of 32, cluster of H/32 <= 16 CTAs, the U slice plus the h
// buffers within 227 KB of shared memory, and a batch small enough that one launch beats T GEMM-shaped step kernels
// (NNTK_CLUSTER_MAX_BATCH, default 256).
static int recurrent_cluster_try(const nntk_recurrent_args *r, int nb, const float *xw, float *h, float *c, float *out,
                                 int *used) {
    *used = 0;
    static int max_batch = -1;
    if (max_batch < 0) {
        const char *e = getenv("NNTK_CLUSTER_MAX_BATCH");
        max_batch = e ? atoi(e) : 256;
    }
    const int G = r->cell, H = r->H;
    if (nb > max_batch || (H & 31) || H / 32 > 16 || r->T < 1) return 0;
    const int BS = nb >= 4 ? 4 : (nb >= 2 ? 2 : 1);
    const size_t smem = ((size_t)H * (G * CL_UNITS + 4) + (size_t)2 * BS * H) * sizeof(float);
    if (smem > 227 * 1024) return 0;
    ClusterDev p;
    p.B = nb; p.H = H; p.T = r->T; p.C = H / 32; p.return_sequences = r->return_sequences;
    p.U = r->d_U;
    p.bh = ((G == 3) || r->v2) ? r->d_bh : nullptr;
    p.xw = xw; p.h_state = h; p.c_state = c; p.out = out;
    for (int i = 0; i < 5; ++i) { p.act[i] = r->act[i]; p.act_a[i] = r->act_a[i]; }
    const int groups = (nb + BS - 1) / BS;
    cudaStream_t st = nntk_stream();
    int rc;
#define NNTK_CL(GG, BB) rc = launch_cluster<GG, BB>(p, groups, smem, st)
    if (G == 1) { if (BS == 4) NNTK_CL(1, 4); else if (BS == 2) NNTK_CL(1, 2); else NNTK_CL(1, 1); }
    else if (G == 3) { if (BS == 4) NNTK_CL(3, 4); else if (BS == 2) NNTK_CL(3, 2); else NNTK_CL(3, 1); }
    else { if (BS == 4) NNTK_CL(4, 4); else if (BS == 2) NNTK_CL(4, 2); else NNTK_CL(4, 1); }
#undef NNTK_CL
    if (rc) return rc;
    *used = 1;
    return 0;
}

extern "C" int nntk_cuda_recurrent_tc(const nntk_recurrent_args *r, int nb, const float *xw, int xw_blocked, float *h, float *c,
                                      float *out, float *pp, int x3, int *used);   // recurrent_tc.cu

extern "C" int nntk_cuda_recurrent_cluster_tc(const nntk_recurrent_args *r, int nb, const float *xw, float *h, float *c,
                                              float *out, int *used);   // recurrent_cluster_tc.cu

extern "C" int nntk_cuda_recurrent_cluster_h16(const nntk_recurrent_args *r, int nb, const float *xw, int xw_blocked, float *h,
                                               float *c, float *out, int *used);  // recurrent_cluster_h16.cu
extern "C" int nntk_cuda_recurrent_cluster_h16_plan(const nntk_recurrent_args *r, int nb);
extern "C" int nntk_cuda_gemm_tf32_ex(const nntk_gemm_args *a, int split, int c_blocked);   // gemm_tc.cu
extern "C" int nntk_cuda_gemm_proj_adaptive(const nntk_gemm_args *a, float thr);             // gemm_tc.cu
extern "C" int nntk_cuda_recurrent_fused_plan(const nntk_recurrent_args *r, int nb);          // recurrent_fused.cu
extern "C" int nntk_cuda_recurrent_fused(const nntk_recurrent_args *r, int nb, const float *x, const float *xw, float *h, float *c,
                                         float *out, int *used);

// which kernel family ran the most recent recurrent chunk (tests pin the config-shape dispatch with it)
static const char *g_last_path = "none";
extern "C" const char *nntk_cuda_last_recurrent_path(void) { return g_last_path; }

static int recurrent_chunk(const nntk_recurrent_args *r, int b0, int nb, float *work, size_t work_bytes) {
    const int G = r->cell, H = r->H, T = r->T, GH = G * H;
    const float *x = r->d_x + (long long)b0 * T * r->in;
    float *h = r->d_h + (long long)b0 * H;
    float *c = r->d_c ? r->d_c + (long long)b0 * H : nullptr;
    float *out = r->d_out + (long long)b0 * (r->return_sequences ? (long long)T * H : H);
    float *xw = work;
    float *pp = work + (size_t)nb * T * GH;
    // 1. hoisted input projection: xw = x*W + b_i over all timesteps
    nntk_gemm_args g = {};
    g.d_A = x; g.a_slab_stride = 0; g.lda = r->in; g.batch = 1; g.rows = nb * T; g.N = GH; g.K = r->in;
    g.d_B = r->d_W; g.d_bias = r->d_bi; g.d_bn = nullptr; g.act = NNTK_ACT_NONE; g.act_a = 1.f;
    g.d_C = xw; g.c_slab_stride = 0; g.ldc = GH;
    if ((long long)nb * T > 0x7fffffffLL) return NNTK_ERR_ARG_;
    // The un-gated RNN amplifies operand rounding step by step (its Jacobian ~ U does not contract): measured 1.2e-3 at
    // T = 12 with only the PROJECTION in TF32.  It has no BASELINE config, so it simply stays fp32 in every math mode.
    const int mode = nntk_cuda_math_mode();
    // TF32 mode runs the projection with split operands (3xbf16 below, 3xTF32 where that does not fit): emulation of the C5 GRU over 300 steps showed that the
    // single-pass TF32 rounding of x and W in the PROJECTION alone costs 1.2e-3 (each operand ~7e-4), the recurrent TF32
    // rounding only 3.5e-4 -- and the projection is the cheap part.  The tensor-core STEP kernels stay single-pass TF32.
    const int proj_mode = (G == 1) ? 0 : (mode == 1 ? 2 : mode);
    // 0. tensor-core mode, default activations, H % 64 == 0: ONE persistent kernel per layer (recurrent_fused.cu).  Plan 1 also
    //    computes the input projection inside that kernel -- no projection buffer, no GEMM launch; plan 2 reads the blocked
    //    hoisted projection below.
    //    The fp32-contract mode has the same kernel for the GRU with every product in three split-fp16 passes ("fused_x3").
    static int fused_x3 = -1;
    if (fused_x3 < 0) { const char *e = getenv("NNTK_RECURRENT_FUSED_X3"); fused_x3 = e ? atoi(e) : 1; }
    const bool x3_fused = mode == 2 && G == 3 && fused_x3;
    int fplan = ((mode == 1 && G != 1) || x3_fused) ? nntk_cuda_recurrent_fused_plan(r, nb) : 0;
    if (x3_fused && fplan != 1) fplan = 0;
    if (fplan == 1) {
        int used = 0;
        int frc = nntk_cuda_recurrent_fused(r, nb, x, nullptr, h, c, out, &used);
        if (frc) return frc;
        if (used) { g_last_path = x3_fused ? "fused_x3" : "fused_tc"; return 0; }
    }
    if (x3_fused) fplan = 0;
    static int min_batch = -1, use_h16 = -1;
    // never below 32: the blocked projection layout rounds the chunk up to 32 sequences and the chunker only reserves that
    // rounding for chunks of at least 32 (nntk_cuda_recurrent below)
    if (min_batch < 0) { const char *e = getenv("NNTK_CLUSTER_TC_MIN_BATCH"); min_batch = e ? atoi(e) : 32; if (min_batch < 32) min_batch = 32; }
    if (use_h16 < 0) { const char *e = getenv("NNTK_CLUSTER_H16"); use_h16 = e ? atoi(e) : 1; }
    int rc;
    // When the fp16 persistent cluster kernel will take this chunk, the projection is computed timestep-major and written in
    // the 32-sequence-blocked layout that kernel reads with fully coalesced row-per-lane accesses (gemm_tc.cu, c_blocked).
    int xw_blocked = 0;
    const bool h16 = mode == 1 && G != 1 && use_h16 && nb >= min_batch && nntk_cuda_recurrent_cluster_h16_plan(r, nb);
    // The per-step tensor-core kernel profits even more: with the plain [B, T, G*H] layout one step touches B pieces 4 MB
    // apart (C4: 4096 different pages per step -- no operand-traffic change moved its 52 us/step, this is what bound it).
    // It is the planned path when neither cluster kernel can hold the shape (H > 256 or H % 32 != 0) and the batch is large.
    const bool step_tc = mode == 1 && G != 1 && !h16 && nb > 256 && (H > 256 || (H & 31)) && (H & 3) == 0 && (GH & 31) == 0;
    // TF32X3 mode (fp32 contract on the tensor cores): large batches run the same per-step kernel with hi/lo split operands
    const bool step_x3 = mode == 2 && G != 1 && nb > 256 && (H & 3) == 0 && (GH & 31) == 0 && H <= 1536 && r->in <= 1536;
    // the blocked layout rounds the chunk up to 32 sequences: only when that (plus the ping-pong buffers) fits this workspace
    const bool blocked_fits = ((size_t)((nb + 31) / 32 * 32) * T * GH + (size_t)6 * nb * H) * sizeof(float) <= work_bytes;
    if ((h16 || step_tc || step_x3 || fplan == 2) && blocked_fits && (((uintptr_t)xw) & 15) == 0) {
        nntk_gemm_args gb = g;
        gb.batch = T; gb.rows = nb; gb.a_slab_stride = r->in; gb.lda = T * r->in;      // slab = timestep, row = sequence
        // 3xbf16 first (bf16 hi/lo split: ~2^-16 operand error, twice the tensor rate and half the operand bytes of 3xTF32);
        // shapes its 2-byte pitch rules exclude fall back to 3xTF32
        static int proj_bf16 = -1;
        if (proj_bf16 < 0) { const char *e = getenv("NNTK_PROJ_BF16X3"); proj_bf16 = e ? atoi(e) : 1; }
        // ... and a single fp16 pass instead when the inputs are small enough for it (decided on the device: gemm_tc.cu)
        static float adapt_thr = -1.0f;
        if (adapt_thr < 0.0f) { const char *e = getenv("NNTK_RECURRENT_X_ADAPT_THR"); adapt_thr = e ? (float)atof(e) : 6.5f; }
        rc = (proj_bf16 && !step_x3 && mode == 1) ? nntk_cuda_gemm_proj_adaptive(&gb, adapt_thr) : 1;
        if (rc == 1) rc = (proj_bf16 && !step_x3) ? nntk_cuda_gemm_tf32_ex(&gb, 2, 1) : 1;
        if (rc == 1) rc = nntk_cuda_gemm_tf32_ex(&gb, 1, 1);
        if (rc < 0) return rc;
        xw_blocked = rc == 0;
        if (xw_blocked) pp = work + (size_t)((nb + 31) / 32 * 32) * T * GH;      // the blocked xw is rounded up to 32 sequences
    }
    if (!xw_blocked) {
        if (proj_mode != mode) nntk_cuda_set_math_mode(proj_mode);
        rc = nntk_cuda_gemm(&g);
        if (proj_mode != mode) nntk_cuda_set_math_mode(mode);
        if (rc) return rc;
    }
    if (fplan == 2 && xw_blocked) {
        int used = 0;
        rc = nntk_cuda_recurrent_fused(r, nb, x, xw, h, c, out, &used);
        if (rc) return rc;
        if (used) { g_last_path = "fused_tc_xw"; return 0; }
    }
    // 2a'. TF32 mode, H <= 256, batch >= NNTK_CLUSTER_TC_MIN_BATCH (default 32): ONE persistent tensor-core cluster kernel
    //      for all T steps (recurrent_cluster_tc.cu: U resident in shared memory, DSMEM bulk-copy h exchange)
    if (mode == 1 && G != 1) {
        if (h16) {   // fp16-operand, two-slice variant first (default activations only)
            int used = 0;
            rc = nntk_cuda_recurrent_cluster_h16(r, nb, xw, xw_blocked, h, c, out, &used);
            if (rc) return rc;
            if (used) { g_last_path = "cluster_h16"; return 0; }
            if (xw_blocked) {
                nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent: the planned fp16 cluster kernel did not run");
                return NNTK_ERR_CUDA_;
            }
        }
        if (nb >= min_batch && !xw_blocked) {
            int used = 0;
            rc = nntk_cuda_recurrent_cluster_tc(r, nb, xw, h, c, out, &used);
            if (rc) return rc;
            if (used) { g_last_path = "cluster_tf32"; return 0; }
        }
    }
    // 2a. small / medium batch: ONE persistent cluster kernel for all T steps (U resident in shared memory, DSMEM exchange)
    if (!xw_blocked) {
        int used = 0;
        rc = recurrent_cluster_try(r, nb, xw, h, c, out, &used);
        if (rc) return rc;
        if (used) { g_last_path = "cluster_fp32"; return 0; }
    }
    cudaStream_t st = nntk_stream();
    // 2b. large batch, TF32 mode: T tensor-core step kernels with the cell fused into the GEMM epilogue (recurrent_tc.cu)
    if ((mode == 1 || step_x3) && G != 1) {
        int used = 0;
        rc = nntk_cuda_recurrent_tc(r, nb, xw, xw_blocked, h, c, out, pp, step_x3 ? 1 : 0, &used);
        if (rc) return rc;
        if (used) { g_last_path = step_x3 ? "step_tc_x3" : "step_tc_f16"; goto carry_state; }
    }
    if (xw_blocked) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent: the planned tensor-core kernel did not run on the blocked projection");
        return NNTK_ERR_CUDA_;
    }
    {
    // 2c. large batch, fp32 mode: T fused FFMA step kernels
    StepDev p;
    p.B = nb; p.H = H; p.c = c; p.U = r->d_U;
    const bool add_bh = (G == 3) || r->v2;   // gru.c:151 always; rnn.c:161-163 / lstm.c:207-209 only if v2
    p.bh = add_bh ? r->d_bh : nullptr;
    p.ld_xw = (long long)T * GH;
    for (int i = 0; i < 5; ++i) p.act[i] = r->act[i];
    for (int i = 0; i < 5; ++i) p.act_a[i] = r->act_a[i];
    p.u_tiles = (H + SU - 1) / SU;
    const int m_tiles = (nb + SM_ROWS - 1) / SM_ROWS;
    const unsigned grid = (unsigned)(m_tiles * p.u_tiles);
    for (int t = 0; t < T; ++t) {
        if (t == 0) { p.h_prev = h; p.ld_prev = H; }
        else if (r->return_sequences) { p.h_prev = out + (long long)(t - 1) * H; p.ld_prev = (long long)T * H; }
        else { p.h_prev = pp + (size_t)((t - 1) & 1) * nb * H; p.ld_prev = H; }
        if (r->return_sequences) { p.h_next = out + (long long)t * H; p.ld_next = (long long)T * H; }
        else { p.h_next = pp + (size_t)(t & 1) * nb * H; p.ld_next = H; }
        p.xw = xw + (long long)t * GH;
        if (G == 1) recurrent_step_kernel<1><<<grid, 256, 0, st>>>(p);
        else if (G == 3) recurrent_step_kernel<3><<<grid, 256, 0, st>>>(p);
        else recurrent_step_kernel<4><<<grid, 256, 0, st>>>(p);
    }
    NNTK_LAUNCH_CHECK_N("recurrent_step_kernel", T);
    g_last_path = "step_fp32";
    }
carry_state:
    // 3. carry the final state back (reference: f_copy(filter->h, h_t), gru.c:201 / lstm.c:264-265 / rnn.c:241)
    if (T > 0) {
        if (r->return_sequences) {
            NNTK_CUDA_TRY(cudaMemcpy2DAsync(h, (size_t)H * sizeof(float), out + (long long)(T - 1) * H,
                                            (size_t)T * H * sizeof(float), (size_t)H * sizeof(float), nb,
                                            cudaMemcpyDeviceToDevice, st));
        } else {
            const float *last = pp + (size_t)((T - 1) & 1) * nb * H;
            NNTK_CUDA_TRY(cudaMemcpyAsync(out, last, (size_t)nb * H * sizeof(float), cudaMemcpyDeviceToDevice, st));
            NNTK_CUDA_TRY(cudaMemcpyAsync(h, last, (size_t)nb * H * sizeof(float), cudaMemcpyDeviceToDevice, st));
        }
    }
    return 0;
}

extern "C" int nntk_cuda_recurrent(const nntk_recurrent_args *r) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!r || (r->cell != 1 && r->cell != 3 && r->cell != 4) || r->in < 1 || r->H < 1 || r->T < 0 || r->batch < 0 ||
        !r->d_W || !r->d_U || !r->d_bi || !r->d_bh || !r->d_x || !r->d_h || !r->d_out || !r->d_work ||
        (r->cell == 4 && !r->d_c)) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "recurrent: bad arguments");
        return NNTK_ERR_ARG_;
    }
    for (int i = 0; i < 5; ++i)
        if (r->act[i] == NNTK_ACT_SOFTMAX || r->act[i] == NNTK_ACT_CUSTOM) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "recurrent: gate activation kind %d cannot run fused on the device", r->act[i]);
            return NNTK_ERR_UNSUPPORTED_;
        }
    if (r->batch == 0 || r->T == 0) return 0;
    // The fused persistent kernel (tensor-core mode) computes the projection itself: no workspace, so no batch chunking either
    {
        static int fused_x3 = -1;
        if (fused_x3 < 0) { const char *e = getenv("NNTK_RECURRENT_FUSED_X3"); fused_x3 = e ? atoi(e) : 1; }
        const int mm = nntk_cuda_math_mode();
        const bool x3_fused = mm == 2 && r->cell == 3 && fused_x3;
        if (((mm == 1 && r->cell != 1) || x3_fused) && nntk_cuda_recurrent_fused_plan(r, r->batch) == 1) {
            int used = 0;
            rc = nntk_cuda_recurrent_fused(r, r->batch, r->d_x, nullptr, r->d_h, r->d_c, r->d_out, &used);
            if (rc) return rc;
            if (used) { g_last_path = x3_fused ? "fused_x3" : "fused_tc"; return 0; }
        }
    }
    // batch chunks that fit the workspace
    const size_t per_seq = ((size_t)r->T * r->cell * r->H + 6 * (size_t)r->H) * sizeof(float);
    size_t avail = r->work_bytes > 256 ? r->work_bytes - 256 : 0;
    long long chunk = (long long)(avail / per_seq);
    if (chunk < 1) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "recurrent: workspace of %zu bytes holds no sequence (%zu needed each)", r->work_bytes, per_seq);
        return NNTK_ERR_ARG_;
    }
    if (chunk >= r->batch) {
        chunk = r->batch;
        // the blocked projection layout rounds the chunk up to 32 sequences: make sure that still fits
        const size_t need = ((size_t)((chunk + 31) / 32 * 32) * r->T * r->cell * r->H + 6 * (size_t)chunk * r->H) * sizeof(float);
        if (need > avail) chunk = chunk / 32 * 32;
    } else {
        chunk = chunk / 32 * 32;
    }
    if (chunk < 1) chunk = 1;      // tiny workspace: plain layout paths only (nb < 32 never plans the blocked layout)
    for (int b0 = 0; b0 < r->batch; b0 += (int)chunk) {
        const int nb = (int)((b0 + chunk <= r->batch) ? chunk : (r->batch - b0));
        rc = recurrent_chunk(r, b0, nb, r->d_work, avail);
        if (rc) return rc;
    }
    return 0;
}

// ---- allocation-free variant -----------------------------------------------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (code
// under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same layer forward as nntk_cuda_recurrent on its
// fp32 path -- hoisted projection GEMM (FFMA), then the persistent fp32 cluster kernel or the per-step FFMA kernels -- with the
// scratch (projection xw[batch, T, G*H] + 6 [batch, H] ping-pong buffers) taken from ONE file-scope device array instead of the
// caller's workspace.  Every GPU that loads the library holds the array (NNTK_RECURRENT_STATIC_SCRATCH_BYTES) until exit.
// Nothing in it has to start at zero: the projection is fully written before it is read and every ping-pong buffer is written
// by step t before step t + 1 reads it, so the previous call's contents are simply overwritten.  All calls share the array:
// two calls whose kernels overlap on the GPU (different streams, or two host threads each with its own stream) race.
static __device__ __align__(256) float g_recurrent_static_scratch[NNTK_RECURRENT_STATIC_SCRATCH_BYTES / sizeof(float)];

extern "C" int nntk_cuda_recurrent_static_scratch(const nntk_recurrent_args *r) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!r || (r->cell != 1 && r->cell != 3 && r->cell != 4) || r->in < 1 || r->H < 1 || r->T < 0 || r->batch < 0 ||
        !r->d_W || !r->d_U || !r->d_bi || !r->d_bh || !r->d_x || !r->d_h || !r->d_out || (r->cell == 4 && !r->d_c)) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "recurrent_static_scratch: bad arguments");
        return NNTK_ERR_ARG_;
    }
    for (int i = 0; i < 5; ++i)
        if (r->act[i] == NNTK_ACT_SOFTMAX || r->act[i] == NNTK_ACT_CUSTOM) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "recurrent_static_scratch: gate activation kind %d cannot run fused on the device", r->act[i]);
            return NNTK_ERR_UNSUPPORTED_;
        }
    if (r->batch == 0 || r->T == 0) return 0;
    // the whole batch in one chunk (nothing here may grow a buffer): projection + ping-pong must fit the array
    const double need_d = ((double)r->batch * r->T * r->cell * r->H + 6.0 * r->batch * r->H) * sizeof(float);
    if (need_d > (double)NNTK_RECURRENT_STATIC_SCRATCH_BYTES) {
        nntk_cuda_set_error(NNTK_ERR_ARG_,
                            "recurrent_static_scratch: batch %d x T %d x %d gates x H %d needs %.0f bytes of scratch, the static array holds %llu "
                            "(NNTK_RECURRENT_STATIC_SCRATCH_BYTES); use nntk_cuda_recurrent with a caller-owned workspace",
                            r->batch, r->T, r->cell, r->H, need_d, (unsigned long long)NNTK_RECURRENT_STATIC_SCRATCH_BYTES);
        return NNTK_ERR_ARG_;
    }
    nntk_cuda_api_lock();
    // In host code the array's name is only a stand-in: its device address is looked up once per device, with that device current
    // (nntk_cuda_ensure_init above made it so), and that pointer is what the kernels get.
    static float *scratch_of_device[64] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        nntk_cuda_api_unlock();
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_static_scratch: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return NNTK_ERR_CUDA_;
    }
    if (!scratch_of_device[dev]) {
        float *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_recurrent_static_scratch);
        if (e != cudaSuccess || !ptr) {
            nntk_cuda_api_unlock();
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_static_scratch: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return NNTK_ERR_CUDA_;
        }
        scratch_of_device[dev] = ptr;
    }
    // the fp32 path of the dispatcher is the one that neither allocates nor repacks operands into library-owned buffers
    const int mode = nntk_cuda_math_mode();
    if (mode != 0) nntk_cuda_set_math_mode(0);
    rc = recurrent_chunk(r, 0, r->batch, scratch_of_device[dev], (size_t)NNTK_RECURRENT_STATIC_SCRATCH_BYTES);
    if (mode != 0) nntk_cuda_set_math_mode(mode);
    nntk_cuda_api_unlock();
    return rc;
}
