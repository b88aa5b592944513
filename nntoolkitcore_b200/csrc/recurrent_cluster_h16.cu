// recurrent_cluster_h16.cu -- K4, persistent tensor-core recurrence, fp16-operand / two-slice variant (NNTK_MATH_TF32 mode,
// default gate activations, H <= 512).  Same structure as recurrent_cluster_tc.cu (cluster of C = H/32 CTAs, U slice resident
// in shared memory, h exchanged by DSMEM bulk copies, a_full / a_free mbarriers) with two changes that ncu asked for:
//   * operands are fp16 (tcgen05.mma kind::f16, fp32 accumulate).  With the default activations h is bounded by 1, so fp16's
//     10-bit mantissa is EXACTLY the TF32 operand precision (2^-11 relative) -- but the A tile, the U slice and above all the
//     per-step DSMEM exchange (which bounded the TF32 kernel: 45 % of the step was the wait for copies + MMA) are half the
//     bytes.  k-block = 32 units = 64-byte rows, SWIZZLE_64B, 8 KB per 128 sequences.
//   * the halved A tile leaves room for TWO independent 128-sequence slices per cluster.  The step of one slice is a strict
//     chain (MMA -> cell math -> exchange -> MMA ...); with two slices the epilogue warps do the cell math of one slice while
//     the other slice's copies and MMA are in flight, so the exchange latency is hidden instead of added.
// Not eligible (caller falls back to the TF32 cluster kernel, then to the per-step kernels): custom gate activations (h may
// leave the fp16 range), H > 512 or H % 32 != 0.  H = 512 (16-CTA non-portable clusters) runs with 64-sequence slices:
// LSTM-512 (BASELINE config 4) one slice per cluster, GRU-512 two.  fp16 overflow of a caller-supplied |h_0| > 65504 saturates to inf like any fp16 cast.

#include <cuda_fp16.h>
#include "tc_common.cuh"

#ifndef HT_MUFU_TANH
#define HT_MUFU_TANH 1
#endif
#if HT_MUFU_TANH
#define HT_ACT(kind, x) act_mufu(kind, x)
#else
#define HT_ACT(kind, x) act_fast(kind, 1.0f, x)
#endif

namespace {

constexpr int HT_THREADS = 320;              // warp 0: a_free signalling, warp 1: MMA issuer, warps 2..9: epilogue
// Template parameters of the kernel: BM = sequences per slice (128, or 64 when the A tile must shrink: H = 512), SL = slices
// per cluster (2, or 1 when shared memory holds only one A tile: LSTM-512).  One k-block of A = BM rows x 32 halves.

struct HtParams {
    int B, H, T, C;
    int return_sequences;
    const float *bh;                // [G*H] or null (not added)
    const float *xw;                // [B][T][G*H], b_i already added -- or, xw_blocked: [T][B/32][G*H/4][32][4] (gemm_tc.cu)
    int xw_blocked, row_blocks;
    float *h_state, *c_state;       // [B][H] in/out
    float *out;                     // [B][T][H] or [B][H]
};

__device__ __forceinline__ uint32_t ht_cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void ht_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t ht_map(uint32_t saddr, uint32_t rank) {
    uint32_t d;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(saddr), "r"(rank));
    return d;
}
__device__ __forceinline__ void ht_bulk_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t mbar_cluster) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(mbar_cluster) : "memory");
}
__device__ __forceinline__ void ht_mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// tcgen05.commit that arrives on the mbarrier at the SAME offset in every CTA of `mask` once the MMAs issued so far complete:
// the hardware tells all peers "my A tile has been read" -- no software hop between MMA completion and the a_free arrivals
__device__ __forceinline__ void ht_commit_multicast(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
// K-major, SWIZZLE_64B shared-memory matrix descriptor: rows of 64 bytes, 8-row groups 512 bytes apart
__device__ __forceinline__ uint64_t ht_desc(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(512 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)4 << 61;                        // SWIZZLE_64B
    return d;
}
__device__ __forceinline__ uint32_t ht_pack2(float a, float b) {
    __half2 v = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
}

template <int G, int BM, int SL>
__global__ void __launch_bounds__(HT_THREADS, 1)
recurrent_cluster_h16_kernel(const __grid_constant__ CUtensorMap map_u, HtParams p) {
    constexpr int N = 32 * G;
    constexpr int HT_BM = BM, HT_SLICES = SL, HT_KB_BYTES = BM * 64;
    constexpr uint32_t ACC_STRIDE = 128;                   // TMEM columns between the two slices' accumulators
    constexpr uint32_t TMEM_COLS = SL == 2 ? 256 : 128;
    constexpr int S_SIG = NNTK_ACT_SIGMOID, S_TANH = NNTK_ACT_TANH;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int C = p.C, H = p.H, T = p.T, GH = G * H;
    const int a_slice_bytes = C * HT_KB_BYTES;
    uint8_t *sA = smem;                                              // [2 slices][C k-blocks][128 rows][32 halves]
    uint8_t *sB = smem + (size_t)HT_SLICES * a_slice_bytes;          // [C k-blocks][N rows][32 halves]
    uint64_t *bars = reinterpret_cast<uint64_t *>(sB + (size_t)C * N * 64);
    uint64_t *b_full = bars, *a_full = bars + 1, *mma_done = bars + 3, *a_free = bars + 5, *h_ready = bars + 7;   // [2] each
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 9);
    float *s_bh = reinterpret_cast<float *>(bars + 10);              // [G][32]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = ht_cluster_rank();
    const int row0 = (blockIdx.x / C) * (HT_SLICES * HT_BM);
    const uint32_t tx_bytes = (uint32_t)((C - 1) * HT_KB_BYTES);

    if (tid == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_u) : "memory");
        mbar_init(b_full, 1);
        for (int s = 0; s < HT_SLICES; ++s) {
            mbar_init(a_full + s, 2); mbar_init(mma_done + s, 1); mbar_init(a_free + s, (uint32_t)C); mbar_init(h_ready + s, 8);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(b_full, (uint32_t)(C * N * 64));
        for (int kb = 0; kb < C; ++kb) tma_load_2d(sB + (size_t)kb * N * 64, &map_u, b_full, kb * 32, (int)rank * N);
        for (int s = 0; s < HT_SLICES; ++s) mbar_expect_tx(a_full + s, tx_bytes);       // phase 0: the exchange of h_0
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid >= 64 && tid < 64 + N) {
        const int i = tid - 64, g = i >> 5, u = i & 31;
        s_bh[i] = p.bh ? __ldg(p.bh + g * H + (int)rank * 32 + u) : 0.0f;
    }
    tc_fence_before();
    __syncthreads();
    ht_cluster_sync();           // every CTA's barriers exist before anybody arrives on / copies into a peer
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== exchange engine: lane = peer.  For every exchange e and, in the order the epilogue produces them, every slice s:
        //   h_ready[s] (the 8 epilogue warps have written the new h slice into my k-block)  -> push it to CTA `peer`
        // This warp has no global-memory traffic of its own, so the generic->async proxy fence it executes before the copy
        // is cheap; in the epilogue threads the same fence waited for their outstanding xw loads (ncu: 12 % of the step).
        // (The a_free arrivals are not software at all: the MMA issuer's multicast tcgen05.commit delivers them.)
        const int peer = lane;
        if (peer < C) {
            for (int e = 0; e < T; ++e) {
#pragma unroll
                for (int s = 0; s < HT_SLICES; ++s) {
                    const uint32_t my_kb = smem_u32(sA) + s * a_slice_bytes + rank * HT_KB_BYTES;
                    mbar_wait(h_ready + s, (uint32_t)(e & 1));
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    if (peer == (int)rank) mbar_arrive(a_full + s);
                    else ht_bulk_copy(ht_map(my_kb, (uint32_t)peer), my_kb, HT_KB_BYTES, ht_map(smem_u32(a_full + s), (uint32_t)peer));
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: alternates between the two slices =====
        if (lane == 0) {
            // D = f32, A = B = f16, both K-major; N >> 3 at bit 17, M >> 4 at bit 24
            constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(HT_BM >> 4) << 24);
            const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
            mbar_wait(b_full, 0);
            for (int t = 0; t < T; ++t) {
#pragma unroll
                for (int s = 0; s < HT_SLICES; ++s) {
                    mbar_wait(a_full + s, (uint32_t)(t & 1));
                    if (t + 1 < T) mbar_expect_tx(a_full + s, tx_bytes);     // arm the next phase
                    tc_fence_after();
                    const uint32_t tmem_d = tmem_base + s * ACC_STRIDE;
                    for (int kb = 0; kb < C; ++kb) {
                        const uint64_t da = ht_desc(a0 + s * a_slice_bytes + kb * HT_KB_BYTES), db = ht_desc(b0 + kb * N * 64);
#pragma unroll
                        for (int k = 0; k < 2; ++k) ht_mma_f16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                    }
                    tc_commit(mma_done + s);
                    if (t + 1 < T) ht_commit_multicast(a_free + s, (uint16_t)((1u << C) - 1));
                }
            }
        }
    } else {
        // ===== epilogue: thread = row r of BOTH slices x 16 hidden units =====
        // BM = 128: TMEM lane = row.  BM = 64: row m sits in TMEM lane (m % 16) + 32 (m / 16), so each warp's lanes 0..15 own
        // rows 16 q .. 16 q + 15 and lanes 16..31 idle (they still take part in the warp-wide TMEM loads).
        const int q = warp & 3, half = (warp - 2) >> 2;
        const bool active = (BM == 128) || lane < 16;
        const int r = (BM == 128) ? q * 32 + lane : q * 16 + (lane & 15);
        const int ul = half * 16;                          // first unit inside the CTA's 32
        const int unit0 = (int)rank * 32 + ul;             // first hidden unit (global index)
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ul;
        // own A k-block: this thread's two 16-byte chunks of row r (SWIZZLE_64B: chunk index ^ ((row >> 1) & 3))
        const int sw = (r >> 1) & 3;
        uint8_t *stg = reinterpret_cast<uint8_t *>(s_bh + G * 32) + (warp - 2) * 2048;   // output staging, 16-byte aligned
        bool row_ok[HT_SLICES];
        long long rr[HT_SLICES];
        float h[HT_SLICES][16], c[HT_SLICES][G == 4 ? 16 : 1];
#pragma unroll
        for (int s = 0; s < HT_SLICES; ++s) {
            const int row = row0 + s * HT_BM + r;
            row_ok[s] = active && row < p.B;
            rr[s] = row_ok[s] ? row : 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 hv = row_ok[s] ? *reinterpret_cast<const float4 *>(p.h_state + rr[s] * H + unit0 + 4 * j)
                                            : make_float4(0.f, 0.f, 0.f, 0.f);
                h[s][4 * j] = hv.x; h[s][4 * j + 1] = hv.y; h[s][4 * j + 2] = hv.z; h[s][4 * j + 3] = hv.w;
                if (G == 4) {
                    const float4 cv = row_ok[s] ? *reinterpret_cast<const float4 *>(p.c_state + rr[s] * H + unit0 + 4 * j)
                                                : make_float4(0.f, 0.f, 0.f, 0.f);
                    c[s][4 * j] = cv.x; c[s][4 * j + 1] = cv.y; c[s][4 * j + 2] = cv.z; c[s][4 * j + 3] = cv.w;
                }
            }
        }
        float4 xr[G][4];           // ONE buffer: loaded for the other slice while this slice's exchange is in progress
        auto load_xw = [&](int s, int t) {
            if (p.xw_blocked) {
                // lanes = 32 consecutive sequences -> every 16-byte load of the warp covers 512 contiguous bytes
                const float4 *src = reinterpret_cast<const float4 *>(p.xw) +
                                    (((long long)t * p.row_blocks + (rr[s] >> 5)) * (GH >> 2) + (unit0 >> 2)) * 32 + (rr[s] & 31);
#pragma unroll
                for (int g = 0; g < G; ++g)
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        xr[g][j] = row_ok[s] ? __ldg(src + ((long long)g * (H >> 2) + j) * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
                return;
            }
            const float *src = p.xw + (rr[s] * T + t) * (long long)GH + unit0;
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    xr[g][j] = row_ok[s] ? __ldg(reinterpret_cast<const float4 *>(src + g * H) + j) : make_float4(0.f, 0.f, 0.f, 0.f);
        };
        // publishes the fp16-rounded h registers of slice s as k-block `rank` of every CTA's A tile; e = exchange index
        auto exchange = [&](int s, int e) {
            uint8_t *kb = sA + (size_t)s * a_slice_bytes + (size_t)rank * HT_KB_BYTES;
            mbar_wait(a_free + s, (uint32_t)((e & 1) ^ 1));  // e = 0 passes at once; e = t+1 waits for every MMA of step t
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                uint4 v;
                v.x = ht_pack2(h[s][8 * j + 0], h[s][8 * j + 1]); v.y = ht_pack2(h[s][8 * j + 2], h[s][8 * j + 3]);
                v.z = ht_pack2(h[s][8 * j + 4], h[s][8 * j + 5]); v.w = ht_pack2(h[s][8 * j + 6], h[s][8 * j + 7]);
                if (active) *reinterpret_cast<uint4 *>(kb + r * 64 + (((half * 2 + j) ^ sw) << 4)) = v;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(h_ready + s);         // release: the exchange warp fences and pushes
        };
#pragma unroll
        for (int s = 0; s < HT_SLICES; ++s) exchange(s, 0);
        load_xw(0, 0);
        for (int t = 0; t < T; ++t) {
#pragma unroll
            for (int s = 0; s < HT_SLICES; ++s) {
                mbar_wait(mma_done + s, (uint32_t)(t & 1));
                tc_fence_after();
                float acc[G][16];
#pragma unroll
                for (int g = 0; g < G; ++g) tc_ld_32x16(taddr + s * ACC_STRIDE + g * 32, acc[g]);
#pragma unroll
                for (int u = 0; u < 16; ++u) {
                    float hu[G], x[G];
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        hu[g] = acc[g][u] + s_bh[g * 32 + ul + u];
                        x[g] = (&xr[g][u >> 2].x)[u & 3];
                    }
                    if (G == 3) {          // gru.c:156-186
                        const float z = HT_ACT(S_SIG, x[0] + hu[0]);
                        const float rg = HT_ACT(S_SIG, x[1] + hu[1]);
                        const float hh = HT_ACT(S_TANH, fmaf(rg, hu[2], x[2]));
                        h[s][u] = fmaf(z, h[s][u] - hh, hh);               // (1 - z) hh + z h_{t-1}
                    } else {               // lstm.c:199-238
                        const float ig = HT_ACT(S_SIG, x[0] + hu[0]);
                        const float fg = HT_ACT(S_SIG, x[1] + hu[1]);
                        const float gg = HT_ACT(S_TANH, x[2] + hu[2]);
                        const float og = HT_ACT(S_SIG, x[3] + hu[3]);
                        c[s][u] = fmaf(fg, c[s][u], ig * gg);
                        h[s][u] = og * HT_ACT(S_TANH, c[s][u]);
                    }
                }
                // exchange first (the other CTAs are waiting for it), then put the xw of the NEXT slice-step in flight (xr is dead
                // now); issuing the 12 loads BEFORE the barrier try_wait / shared stores of the exchange queued those behind
                // them in the LSU (ncu: 18 % of the step on an a_free wait that never spun)
                if (t + 1 < T) exchange(s, t + 1);
                if (s + 1 < HT_SLICES) load_xw(s + 1, t);
                else if (t + 1 < T) load_xw(0, t + 1);
                if (p.return_sequences) {
                    // h_t rows are 64-byte pieces of [B, T, H] rows: transposed through a 2 KB per-warp tile so that one store
                    // instruction covers 8 sequences x 64 contiguous bytes instead of 32 sequences x 16 bytes
                    constexpr int WROWS = (BM == 128) ? 32 : 16;
                    if (active) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            *reinterpret_cast<float4 *>(stg + (lane & 31) * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) =
                                make_float4(h[s][4 * j], h[s][4 * j + 1], h[s][4 * j + 2], h[s][4 * j + 3]);
                    }
                    __syncwarp();
                    const long long wrow0 = (long long)row0 + s * HT_BM + q * WROWS;      // first sequence of this warp
#pragma unroll
                    for (int i = 0; i < WROWS / 8; ++i) {
                        const int rw = (lane >> 2) + 8 * i, c4 = lane & 3;
                        const float4 w = *reinterpret_cast<const float4 *>(stg + rw * 64 + ((c4 ^ ((rw >> 1) & 3)) << 4));
                        if (wrow0 + rw < p.B)
                            *reinterpret_cast<float4 *>(p.out + ((wrow0 + rw) * T + t) * H + unit0 + c4 * 4) = w;
                    }
                    __syncwarp();
                }
            }
        }
#pragma unroll
        for (int s = 0; s < HT_SLICES; ++s) {
            if (!row_ok[s]) continue;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 hv = make_float4(h[s][4 * j], h[s][4 * j + 1], h[s][4 * j + 2], h[s][4 * j + 3]);
                *reinterpret_cast<float4 *>(p.h_state + rr[s] * H + unit0 + 4 * j) = hv;
                if (!p.return_sequences) *reinterpret_cast<float4 *>(p.out + rr[s] * H + unit0 + 4 * j) = hv;
                if (G == 4)
                    *reinterpret_cast<float4 *>(p.c_state + rr[s] * H + unit0 + 4 * j) =
                        make_float4(c[s][4 * j], c[s][4 * j + 1], c[s][4 * j + 2], c[s][4 * j + 3]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    ht_cluster_sync();           // nobody leaves while a peer could still address this CTA's shared memory
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

// Bt[(rank*G + g)*32 + u][k] = half(U[k][g*H + 32 rank + u])
__global__ void pack_cluster_weights_h16_kernel(const float *__restrict__ U, __half *__restrict__ Bt, int H, int G) {
    const long long total = (long long)G * H * H;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / H), k = (int)(i - (long long)n * H);
        const int rank = n / (G * 32), rem = n - rank * G * 32, g = rem >> 5, u = rem & 31;
        Bt[i] = __float2half_rn(__ldg(U + (long long)k * G * H + g * H + rank * 32 + u));
    }
}

__half *g_ht_scratch = nullptr;
size_t g_ht_halves = 0;

template <int G, int BM, int SL>
int launch_ht(const CUtensorMap &mu, const HtParams &p, int clusters, size_t smem, cudaStream_t st, bool dry, int *max_out = nullptr) {
    auto kern = recurrent_cluster_h16_kernel<G, BM, SL>;
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
    cfg.gridDim = dim3((unsigned)(clusters * p.C));
    cfg.blockDim = dim3(HT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int max_clusters = 0;
    if (cudaOccupancyMaxActiveClusters(&max_clusters, kern, &cfg) != cudaSuccess || max_clusters < 1) {
        (void)cudaGetLastError();
        return 1;     // this cluster shape cannot be scheduled on this device: not eligible, the caller falls back
    }
    if (max_out) *max_out = max_clusters;
    if (dry) return 0;
    NNTK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, mu, p));
    nntk_cuda_count_launch(1);
    return 0;
}

}  // namespace

// static eligibility + configuration: the largest of (128 sequences x 2 slices), (64 x 2), (64 x 1) that fits 227 KB
static int ht_config(const nntk_recurrent_args *r, int nb, int *BM_, int *SL_, size_t *smem_) {
    const int G = r->cell, H = r->H, T = r->T;
    if ((G != 3 && G != 4) || (H & 31) || H > 512 || T < 1 || nb < 1) return 0;
    const int SG = NNTK_ACT_SIGMOID, TH = NNTK_ACT_TANH;
    const bool def = (G == 3) ? (r->act[0] == SG && r->act[1] == SG && r->act[2] == TH)
                              : (r->act[0] == SG && r->act[1] == SG && r->act[2] == TH && r->act[3] == SG && r->act[4] == TH);
    if (!def) return 0;
    const int C = H / 32;
    if (C > 16) return 0;
    int BM = 0, SL = 0;
    size_t smem = 0;
    int cand[4][2] = {{128, 2}, {64, 2}, {64, 1}, {0, 0}};
    static int force_bm = -1, force_sl = 0;
    if (force_bm < 0) {
        force_bm = 0;
        if (const char *e = getenv("NNTK_CLUSTER_H16_FORCE")) { if (sscanf(e, "%d,%d", &force_bm, &force_sl) != 2) force_bm = 0; }
    }
    if (force_bm) { cand[0][0] = force_bm; cand[0][1] = force_sl; cand[1][0] = cand[2][0] = 0; }
    for (int i = 0; i < 3 && !BM; ++i) {
        if (!cand[i][0]) continue;
        const size_t need_smem = (size_t)cand[i][1] * C * cand[i][0] * 64 + (size_t)C * G * 32 * 64 + 96 + (size_t)G * 32 * 4 + 8 * 2048 + 1024;
        if (need_smem <= 227 * 1024) { BM = cand[i][0]; SL = cand[i][1]; smem = need_smem; }
    }
    if (!BM) return 0;
    // 64-sequence slices on 16-CTA clusters (H = 512) keep only ~7 x 64 sequences in flight: measured on C4 they beat the
    // per-step tensor-core kernel up to ~1.5k sequences (B = 512: 19.2 vs 34.9 ms, B = 2048: 55.3 vs 52.6 ms) -- which is
    // the 8-GPU share of C4 (512 per GPU) -- and lose beyond.
    if (BM == 64) {
        static int max64 = -1;
        if (max64 < 0) { const char *e = getenv("NNTK_CLUSTER_H16_MAX_BATCH_64"); max64 = e ? atoi(e) : 1536; }
        if (nb > max64) return 0;
    }
    *BM_ = BM; *SL_ = SL; *smem_ = smem;
    return 1;
}

#define NNTK_HT_DISPATCH(GG, DRY) (BM == 128 ? (SL == 2 ? launch_ht<GG, 128, 2>(mu, p, clusters, smem, st, DRY) : launch_ht<GG, 128, 1>(mu, p, clusters, smem, st, DRY)) \
                                  : SL == 2 ? launch_ht<GG, 64, 2>(mu, p, clusters, smem, st, DRY) : launch_ht<GG, 64, 1>(mu, p, clusters, smem, st, DRY))

// 1 when nntk_cuda_recurrent_cluster_h16 WILL run this chunk (so the caller may lay the projection out for it), else 0
extern "C" int nntk_cuda_recurrent_cluster_h16_plan(const nntk_recurrent_args *r, int nb) {
    int BM, SL; size_t smem;
    if (!ht_config(r, nb, &BM, &SL, &smem) || !get_encode()) return 0;
    CUtensorMap mu = {};
    HtParams p = {};
    p.C = r->H / 32;
    const int clusters = 1;
    cudaStream_t st = nntk_stream();
    const int rc = (r->cell == 3) ? NNTK_HT_DISPATCH(3, true) : NNTK_HT_DISPATCH(4, true);
    return rc == 0;
}

// Runs all T steps of one batch chunk in ONE launch.  *used = 0 when the shape / activations are not eligible.
// xw_blocked: the projection was written in the blocked recurrent layout of nntk_cuda_gemm_tf32_ex.
extern "C" int nntk_cuda_recurrent_cluster_h16(const nntk_recurrent_args *r, int nb, const float *xw, int xw_blocked, float *h,
                                               float *c, float *out, int *used) {
    *used = 0;
    const int G = r->cell, H = r->H, T = r->T;
    int BM, SL; size_t smem;
    if (!ht_config(r, nb, &BM, &SL, &smem)) return 0;
    const int C = H / 32;
    if ((((uintptr_t)h) | ((uintptr_t)out) | ((uintptr_t)xw) | ((uintptr_t)c)) & 15) return 0;
    EncodeTiledFn encode = get_encode();
    if (!encode) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_cluster_h16: cuTensorMapEncodeTiled is not available from this driver");
        return NNTK_ERR_CUDA_;
    }
    cudaStream_t st = nntk_stream();
    const size_t need = (size_t)G * H * H;
    if (need > g_ht_halves) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_ht_scratch) cudaFree(g_ht_scratch);
        g_ht_scratch = nullptr; g_ht_halves = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_ht_scratch, need * sizeof(__half)));
        g_ht_halves = need;
    }
    {
        const long long total = (long long)need;
        const int blocks = (int)((total + 255) / 256 < 1024 ? (total + 255) / 256 : 1024);
        pack_cluster_weights_h16_kernel<<<blocks, 256, 0, st>>>(r->d_U, g_ht_scratch, H, G);
        NNTK_LAUNCH_CHECK("pack_cluster_weights_h16_kernel");
    }
    CUtensorMap mu;
    {
        cuuint64_t dims[2] = {(cuuint64_t)H, (cuuint64_t)G * H};
        cuuint64_t strides[1] = {(cuuint64_t)H * 2};
        cuuint32_t box[2] = {32, (cuuint32_t)(G * 32)};
        cuuint32_t estr[2] = {1, 1};
        CUresult cr = encode(&mu, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void *)g_ht_scratch, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) {
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_cluster_h16: tensor map for U rejected (CUresult %d)", (int)cr);
            return NNTK_ERR_CUDA_;
        }
    }
    HtParams p;
    p.B = nb; p.H = H; p.T = T; p.C = C; p.return_sequences = r->return_sequences;
    p.bh = ((G == 3) || r->v2) ? r->d_bh : nullptr;
    p.xw = xw; p.h_state = h; p.c_state = c; p.out = out;
    p.xw_blocked = xw_blocked; p.row_blocks = (nb + 31) / 32;
    // One launch covers `rows` sequences starting at row0 (a multiple of 32) with slices of bm x sl sequences per cluster.
    auto run = [&](int bm, int sl, int row0, int rows, bool dry, int *max_out) -> int {
        HtParams q = p;
        q.B = rows;
        q.h_state = h + (size_t)row0 * H;
        q.c_state = c ? c + (size_t)row0 * H : nullptr;
        q.out = out + (size_t)row0 * (r->return_sequences ? (size_t)T * H : (size_t)H);
        q.xw = xw_blocked ? xw + (size_t)(row0 / 32) * G * H * 32 : xw + (size_t)row0 * T * G * H;
        const size_t sm = (size_t)sl * C * bm * 64 + (size_t)C * G * 32 * 64 + 96 + (size_t)G * 32 * 4 + 8 * 2048 + 1024;
        const int cl = (rows + sl * bm - 1) / (sl * bm);
        if (bm == 128 && sl == 2) return G == 3 ? launch_ht<3, 128, 2>(mu, q, cl, sm, st, dry, max_out) : launch_ht<4, 128, 2>(mu, q, cl, sm, st, dry, max_out);
        if (bm == 128) return G == 3 ? launch_ht<3, 128, 1>(mu, q, cl, sm, st, dry, max_out) : launch_ht<4, 128, 1>(mu, q, cl, sm, st, dry, max_out);
        if (sl == 2) return G == 3 ? launch_ht<3, 64, 2>(mu, q, cl, sm, st, dry, max_out) : launch_ht<4, 64, 2>(mu, q, cl, sm, st, dry, max_out);
        return G == 3 ? launch_ht<3, 64, 1>(mu, q, cl, sm, st, dry, max_out) : launch_ht<4, 64, 1>(mu, q, cl, sm, st, dry, max_out);
    };
    int rc;
    int main_rows = nb, tail_rows = 0;
    static int split_tail = -1;
    if (split_tail < 0) { const char *e = getenv("NNTK_CLUSTER_H16_SPLIT_TAIL"); split_tail = e ? atoi(e) : 1; }
    if (BM == 128 && SL == 2 && split_tail) {
        // Waves: only `maxc` clusters are resident, and a wave costs the same whether it is full or not (the step is a latency
        // chain).  Two-slice clusters (256 sequences, 6.9 us per step at GRU-256) are the densest; one-slice clusters
        // (128 sequences, 3.8 us) finish a small remainder in ~0.55 of a wave.  Measured at GRU-256, T = 994: 512 sequences
        // 7.6 -> 4.5 ms, 1 920 sequences 9.7 -> 6.7 ms.
        int maxc = 0;
        rc = run(128, 2, 0, nb, true, &maxc);
        if (rc == 1) return 0;
        if (rc) return rc;
        const int wave = maxc * 256;
        const int rem = nb % wave;
        if (rem) {
            const double two = (double)((rem + 255) / 256 + maxc - 1) / maxc >= 1.0 ? (double)(((rem + 255) / 256 + maxc - 1) / maxc) : 1.0;
            const double one = 0.55 * (double)(((rem + 127) / 128 + maxc - 1) / maxc);
            if (one < two) { tail_rows = rem; main_rows = nb - rem; }
        }
    }
    if (main_rows > 0) {
        rc = run(BM, SL, 0, main_rows, false, nullptr);
        if (rc == 1) return 0;
        if (rc) return rc;
    }
    if (tail_rows > 0) {
        rc = run(128, 1, main_rows, tail_rows, false, nullptr);
        if (rc == 1) {
            if (main_rows > 0) { nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_cluster_h16: the one-slice tail launch is not schedulable"); return NNTK_ERR_CUDA_; }
            return 0;
        }
        if (rc) return rc;
    }
    *used = 1;
    return 0;
}
