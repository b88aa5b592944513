// recurrent_tc.cu -- K4, tensor-core path (NNTK_MATH_TF32, large batch): one timestep of RNN / GRU / LSTM as a tcgen05
// GEMM  h_{t-1}[B, H] x U[H, G*H]  whose epilogue IS the cell: it reads the accumulator from TMEM, adds b_h and the hoisted
// input projection xw_t, applies the gate nonlinearities and the state update in the reference's op order
// (gru.c:156-186, lstm.c:199-238, rnn.c:156-165) and writes h_t (and c_t) -- no pre-activation ever reaches HBM.
//
//   * U is re-laid ONCE per call: column (tile j, gate g, unit u) <- U[:, g*H + 64 j + u], K-major, rounded to nearest
//     TF32.  An N tile therefore holds ALL gates of its 64 hidden units and the gate math stays inside one thread.
//   * A = h_{t-1}: the epilogue of step t-1 writes h twice -- exact fp32 to the [B, T, H] output (or an exact ping-pong
//     buffer without return_sequences) and ROUNDED TO NEAREST TF32 to a [2, B, H] ping-pong buffer that TMA reads as the A
//     operand of step t (one 3-D tensor map {H, 2, B}; the buffer index is a TMA coordinate).  Both operands are then
//     unbiased TF32; only step 0 reads the caller's raw state (truncated by the tensor core, bias-compensated).
//   * warp 0 = TMA producer, warp 1 = MMA issuer (M = 128, N = 64 G, K = 8 per instruction), warps 2..9 = two epilogue
//     warpgroups (each thread owns one batch row and 32 of the tile's 64 units); two TMEM accumulator buffers let the
//     epilogue of tile i overlap the MMAs of tile i+1.  Persistent over the (B/128) x (H/64) tiles of the step.
// Operands are TF32 (A truncated by the tensor core and bias-compensated, see tc_common.cuh), accumulation and all gate
// math fp32: 1e-3 parity class.  The X3 variant (NNTK_MATH_TF32X3: hi/lo split operands, three passes, expf / tanhf gate
// math) holds the fp32 contract, 1e-5; the plain FP32 mode never comes here.
#include <stdlib.h>
#include <cuda_fp16.h>
#include "tc_common.cuh"

namespace {

constexpr int R_BM = 128, R_BK = 32, R_UT = 64;     // rows, k per stage, hidden units per N tile
// Epilogue warps (plus the TMA warp and the MMA warp): EW / 4 per TMEM lane quarter, each owning 64 / (EW / 4) units of the tile.
// The fp16 / TF32 variants run 8 (32 units per thread, 8-unit chunks: more warps did not help them, 54.1 -> 55.9 ms on C4);
// the X3 variant's expf / tanhf / IEEE-divide gate math is several times longer, and 16 warps (16 units per thread, 4-unit
// chunks to fit 96 registers) took C4 from 159 to 110 ms.
#ifndef RTC_CH
#define RTC_CH 8    // hidden units per epilogue chunk of the TF32/fp16 step kernel
#endif
#ifndef RTC_STG
#define RTC_STG 0   // 1: stage h_t / c_t through shared memory so every store instruction writes 64-byte row pieces.
                    // Measured on C4 (B = 4096): 56.5 ms against 52.7-54.4 ms with the direct 16-byte stores -- the stores'
                    // cost (NNTK_RTC_DBG=4/8/16: 2.6 + 1.2 + 1.2 ms per layer) is not a coalescing problem.  Off.
#endif
#ifndef RTC_MUFU_TANH
#define RTC_MUFU_TANH 1   // fp16-operand variant: MUFU.TANH gate activations (tc_common.cuh act_mufu); C4 51.7 -> 50.1 ms
#endif
#ifndef RTC_PD
#define RTC_PD 1    // chunks whose xw / state loads are in flight ahead of the gate math.  CH=4 with PD=1, 2 or 3 all time
                    // within 0.5 % of CH=8/PD=1 on C4 and C5 (scripts/rtc_variants.sh): load latency is not the limiter
#endif
__host__ __device__ constexpr int rtc_epi_warps(bool x3) { return x3 ? 16 : 8; }
__host__ __device__ constexpr int rtc_threads(bool x3) { return 64 + 32 * rtc_epi_warps(x3); }

struct RtcParams {
    int B, H, n_tiles, m_tiles, k_blocks, total_tiles;
    int kb_pass;                       // X3: k-blocks per pass (k_blocks = 3 * kb_pass)
    long long part_stride;             // X3: floats between the hi and the lo copy of h in the operand ping-pong
    int a_t;                           // TMA coordinate of h_{t-1} along the time axis of map_a
    const float *bh;                   // permuted like the columns of Bt, or null
    const float *xw; long long ld_xw;  // row b: xw + b*ld_xw, original column order g*H + unit
    const float *xw_next; long long pf_lines;  // next step's blocked projection slab (128-byte lines) to pull into L2, or null
    int xw_blocked, row_blocks;        // xw_blocked: xw points at this step's [B/32][G*H/4][32][4] block (gemm_tc.cu c_blocked)
    const float *h_prev; long long ld_prev;
    float *h_next; long long ld_next;
    float *hr_next;                    // [B, H]: h_t rounded to nearest TF32 = the A operand of step t+1
    int dbg;                           // development switches (NNTK_RTC_DBG): 1 no loads, 2 no stores, 4 no h_t store, 8 no operand-copy store, 16 no c store
    float comp;                        // 1 + truncation-bias compensation when A is raw fp32 (step 0), else 1
    float *c;                          // [B, H] LSTM
    int act[5];
    float act_a[5];
};

template <int G, int STAGES>
struct RtcSmem {
    static constexpr int BN = G * R_UT;
    static constexpr int A_BYTES = R_BM * R_BK * 4;
    static constexpr int B_BYTES = BN * R_BK * 4;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
    // store staging of the non-X3 epilogue: per warp a [32 rows][16 units] fp32 tile of h_t and one of c_t (2 KB each)
    static constexpr int STG_OFFSET = BAR_OFFSET + (((2 * STAGES + 4) * 8 + 16 + 127) & ~127);
    static constexpr int STG_WARP_BYTES = 4096;
    static constexpr int TOTAL = STG_OFFSET + (RTC_STG ? 8 * STG_WARP_BYTES : 0) + 1024;
    static constexpr int ACC_STRIDE = BN <= 64 ? 64 : 256;       // TMEM columns between the two accumulators
    static constexpr int TMEM_COLS = 2 * ACC_STRIDE;
};

__device__ __forceinline__ void tc_ld_32x8(uint32_t taddr, float (&r)[8]) {
    uint32_t u[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __uint_as_float(u[i]);
}

// TMA tile load delivered to the same stage offset (and signalled on the same barrier offset) in both CTAs of a pair
__device__ __forceinline__ void tma_load_2d_mc2(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"((uint16_t)3) : "memory");
}
// tcgen05.commit arriving on the barrier at this offset in both CTAs of a pair
__device__ __forceinline__ void tc_commit_mc2(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

__device__ __forceinline__ void tc_ld_chunk(uint32_t taddr, float (&r)[8]) { tc_ld_32x8(taddr, r); }
__device__ __forceinline__ void tc_ld_chunk(uint32_t taddr, float (&r)[4]) {
    uint32_t u[4];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 4; ++i) r[i] = __uint_as_float(u[i]);
}

// default gate activations compiled in when DEF (the run-time kind switch per call tripled the epilogue's instruction count;
// ncu r1: the per-step kernel ran at 0.53 IPC on ~5000 instructions per thread per tile -- the epilogue, not TMA or MMA)
// X3 (the fp32-contract mode): expf / tanhf / IEEE division like the FFMA kernels (common.cuh), never the MUFU forms.
template <bool DEF, int KIND, bool X3>
__device__ __forceinline__ float rt_act(const RtcParams &p, int slot, float x) {
    if (X3) return nntk_act(p.act[slot], p.act_a[slot], x);
#if RTC_MUFU_TANH
    if (DEF) return act_mufu(KIND, x);
#endif
    return DEF ? act_fast(KIND, 1.0f, x) : act_fast(p.act[slot], p.act_a[slot], x);
}

// H16 (implies the default activations): operands are fp16 (kind::f16; a 128-byte stage row holds 64 k-elements instead of 32): with the default activations h is
// bounded by 1, so fp16's 10-bit mantissa is exactly the TF32 operand precision while U and h cost half the L2 traffic -- the
// per-step kernel re-reads U once per 128-row tile and was bound by exactly that (C4, B = 4096: 196 MB of operands per step).
// X3 (NNTK_MATH_TF32X3, never together with H16): 3xTF32 -- h is kept as tf32 hi + lo parts (h - hi is exact), U as
// [lo | hi | hi] along K, three passes hi*Ulo, lo*Uhi, hi*Uhi in one K loop (small passes first); the accumulate truncation is
// compensated through p.comp like in gemm_tc.cu.  Operand error ~2^-22: the recurrence holds the fp32 contract (1e-5).
// CL2 (H16 only): CTAs run as clusters of two on vertically adjacent tiles (same hidden units, rows m and m+1).  Each CTA loads
// its own h tile and HALF of the shared U slice, multicast into both CTAs' stage buffers: the U traffic from L2 halves
// (256 -> 128 KB per tile).  A stage is free once BOTH MMA issuers have committed it (multicast tcgen05.commit, count 2).
template <int G, int STAGES, bool H16, bool X3, bool CL2 = false>
__global__ void __launch_bounds__(rtc_threads(X3), 1)
recurrent_tc_step_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, RtcParams p) {
    using SL = RtcSmem<G, STAGES>;
    constexpr int BN = SL::BN;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + SL::BAR_OFFSET);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tmem_full = empty_bar + STAGES;     // [2]
    uint64_t *tmem_empty = tmem_full + 2;         // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + 2);

    constexpr int R_EPI_WARPS = rtc_epi_warps(X3), R_EPI_UNITS = R_UT / (R_EPI_WARPS / 4);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");     // the next step may queue behind this one right away
    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], CL2 ? 2 : 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], R_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(SL::TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    if (CL2) asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // work item w -> tile: alone, w is the tile; in a pair, w names (n_t, pair of m tiles) and the cluster rank picks the row tile
    uint32_t crank = 0;
    if (CL2) asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
    const int w_first = CL2 ? blockIdx.x >> 1 : blockIdx.x, w_step = CL2 ? gridDim.x >> 1 : gridDim.x;
    const int w_total = CL2 ? p.n_tiles * ((p.m_tiles + 1) >> 1) : p.total_tiles;
    asm volatile("griddepcontrol.wait;" ::: "memory");                  // everything the previous step wrote is visible from here

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = w_first; tile < w_total; tile += w_step) {
                const int n_t = tile % p.n_tiles, m_t = CL2 ? 2 * (tile / p.n_tiles) + (int)crank : tile / p.n_tiles;
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t *sa = smem + stage * SL::STAGE_BYTES;
                    mbar_expect_tx(&full_bar[stage], SL::STAGE_BYTES);
                    const int pass = X3 ? kb / p.kb_pass : 0, akb = kb - pass * p.kb_pass;
                    tma_load_3d(sa, &map_a, &full_bar[stage], akb * (H16 ? 2 * R_BK : R_BK), X3 ? 2 * p.a_t + (pass == 1 ? 1 : 0) : p.a_t, m_t * R_BM);
                    if (CL2) tma_load_2d_mc2(sa + SL::A_BYTES + crank * (SL::B_BYTES / 2), &map_b, &full_bar[stage], kb * 2 * R_BK, n_t * BN + (int)crank * (BN / 2));
                    else tma_load_2d(sa + SL::A_BYTES, &map_b, &full_bar[stage], kb * (H16 ? 2 * R_BK : R_BK), n_t * BN);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            constexpr uint32_t idesc = (1u << 4) | (H16 ? 0u : ((2u << 7) | (2u << 10))) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(R_BM >> 4) << 24);
            int stage = 0; uint32_t phase = 0, it = 0;
            for (int tile = w_first; tile < w_total; tile += w_step, ++it) {
                const uint32_t buf = it & 1, use = it >> 1;
                mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + buf * SL::ACC_STRIDE;
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * SL::STAGE_BYTES);
                    const uint64_t da = make_desc(sa), db = make_desc(sa + SL::A_BYTES);
#pragma unroll
                    for (int k = 0; k < R_BK / 8; ++k) {
                        if (H16) tc_mma_f16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        else tc_mma_tf32(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                    }
                    if (CL2) tc_commit_mc2(&empty_bar[stage]);     // the peer's producer multicasts into this stage as well
                    else tc_commit(&empty_bar[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                tc_commit(&tmem_full[buf]);
            }
        }
    } else {
        // ===== epilogue: thread = one batch row x 32 hidden units, in four chunks of 8 units.  Everything a chunk needs
        // from HBM (xw, c, h_{t-1}) is independent of the accumulator, so it is loaded one chunk AHEAD -- the first chunk
        // even before the tile's MMAs have finished -- and the load latency hides behind the previous chunk's gate math.
        const int q = warp & 3;                       // TMEM lane quarter this warp may read
        const int half = (warp - 2) >> 2;             // which R_EPI_UNITS units of the tile's 64
        const int row_in_tile = q * 32 + lane;
        const float comp = p.comp;
        constexpr int CH = X3 ? 4 : RTC_CH, NCH = R_EPI_UNITS / CH, NV = CH / 4;
        constexpr int PD = X3 ? 1 : RTC_PD, NS = PD + 1;        // chunks prefetched ahead; ring slots
        constexpr bool STG = !X3 && CH == 8 && RTC_STG;          // stores go through the per-warp staging tile
        float4 *stg = reinterpret_cast<float4 *>(smem + SL::STG_OFFSET + (STG ? (warp - 2) * SL::STG_WARP_BYTES : 0));
        struct Pre { float4 xv[G][NV]; float4 st[NV]; };    // st: c (LSTM) or h_{t-1} (GRU)
        // optional (NNTK_RTC_PF, off by default -- see the host loop): pull the NEXT step's projection slab into L2
        if (p.xw_next) {
            constexpr int N_ET = R_EPI_WARPS * 32;
            for (long long i = (long long)blockIdx.x * N_ET + (threadIdx.x - 64); i < p.pf_lines; i += (long long)gridDim.x * N_ET)
                asm volatile("prefetch.global.L2 [%0];" ::"l"(p.xw_next + i * 32));
        }
        uint32_t it = 0;
        for (int tile = w_first; tile < w_total; tile += w_step, ++it) {
            const int n_t = tile % p.n_tiles, m_t = CL2 ? 2 * (tile / p.n_tiles) + (int)crank : tile / p.n_tiles;
            const uint32_t buf = it & 1, use = it >> 1;
            const int row = m_t * R_BM + row_in_tile;
            const bool row_ok = row < p.B;
            const long long rr = row_ok ? row : 0;
            const int ubase = n_t * R_UT + half * R_EPI_UNITS; // first hidden unit of this thread
            const float *xw = p.xw + rr * p.ld_xw + ubase;
            // blocked layout: float4 slot of (row, column n) = ((row / 32) * (G*H/4) + n / 4) * 32 + row % 32
            const float4 *xwb = reinterpret_cast<const float4 *>(p.xw) + ((rr >> 5) * (long long)((G * p.H) >> 2) + (ubase >> 2)) * 32 + (rr & 31);
            const float *stp = (G == 4) ? p.c + rr * p.H + ubase : p.h_prev + rr * p.ld_prev + ubase;
            float *hn = p.h_next + rr * p.ld_next + ubase;
            float *hr = p.hr_next + rr * p.H + ubase;
            float *cw = (G == 4) ? p.c + rr * p.H + ubase : nullptr;
            const float *bh = p.bh ? p.bh + n_t * BN + half * R_EPI_UNITS : nullptr;
            auto prefetch = [&](int ch, Pre &P) {
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int u = ch * CH + 4 * v;
                    const bool ok = row_ok && ubase + u < p.H && !(p.dbg & 1);
#pragma unroll
                    for (int g = 0; g < G; ++g)
                        P.xv[g][v] = !ok ? make_float4(0.f, 0.f, 0.f, 0.f)
                                         : p.xw_blocked ? __ldg(xwb + ((long long)g * (p.H >> 2) + (u >> 2)) * 32)
                                                        : *reinterpret_cast<const float4 *>(xw + (long long)g * p.H + u);
                    P.st[v] = (ok && G != 1) ? *reinterpret_cast<const float4 *>(stp + u) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
            };
            Pre ring[NS];
#pragma unroll
            for (int i = 0; i < PD; ++i)
                if (i < NCH) prefetch(i, ring[i]);
            mbar_wait(&tmem_full[buf], use & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + buf * SL::ACC_STRIDE + ((uint32_t)(q * 32) << 16) + half * R_EPI_UNITS;
#pragma unroll
            for (int ch = 0; ch < NCH; ++ch) {
                if (ch + PD < NCH) prefetch(ch + PD, ring[(ch + PD) % NS]);
                const Pre &cur = ring[ch % NS];
                float acc[G][CH];
#pragma unroll
                for (int g = 0; g < G; ++g) tc_ld_chunk(taddr + g * R_UT + ch * CH, acc[g]);
#pragma unroll
                for (int v = 0; v < NV; ++v) {
                    const int u = ch * CH + 4 * v;
                    if (!(row_ok && ubase + u < p.H)) continue;          // H is a multiple of 4 (no warp-level sync below this line inside the v loop)
                    float4 bv[G];
#pragma unroll
                    for (int g = 0; g < G; ++g)
                        bv[g] = bh ? __ldg(reinterpret_cast<const float4 *>(bh + g * R_UT + u)) : make_float4(0.f, 0.f, 0.f, 0.f);
                    float out[4], cn[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float hu[G], x[G];
#pragma unroll
                        for (int g = 0; g < G; ++g) {
                            hu[g] = fmaf(acc[g][4 * v + e], comp, (&bv[g].x)[e]);
                            x[g] = (&cur.xv[g][v].x)[e];
                        }
                        const float sv = (&cur.st[v].x)[e];
                        if (G == 3) {
                            const float z = rt_act<H16, NNTK_ACT_SIGMOID, X3>(p, 0, x[0] + hu[0]);
                            const float r = rt_act<H16, NNTK_ACT_SIGMOID, X3>(p, 1, x[1] + hu[1]);
                            const float hh = rt_act<H16, NNTK_ACT_TANH, X3>(p, 2, fmaf(r, hu[2], x[2]));
                            out[e] = fmaf(z, sv - hh, hh);                   // (1 - z) hh + z h_{t-1}
                        } else if (G == 4) {
                            const float ig = rt_act<H16, NNTK_ACT_SIGMOID, X3>(p, 0, x[0] + hu[0]);
                            const float fg = rt_act<H16, NNTK_ACT_SIGMOID, X3>(p, 1, x[1] + hu[1]);
                            const float gg = rt_act<H16, NNTK_ACT_TANH, X3>(p, 2, x[2] + hu[2]);
                            const float og = rt_act<H16, NNTK_ACT_SIGMOID, X3>(p, 3, x[3] + hu[3]);
                            cn[e] = fmaf(fg, sv, ig * gg);
                            out[e] = og * rt_act<H16, NNTK_ACT_TANH, X3>(p, 4, cn[e]);
                        } else {
                            out[e] = act_fast(p.act[0], p.act_a[0], hu[0] + x[0]);
                        }
                    }
                    if (p.dbg & 2) { if (out[0] == 123.456f) hn[u] = out[1] + out[2] + out[3] + cn[0] + cn[1] + cn[2] + cn[3]; continue; }
                    if (STG) {
                        // park the 4 units in this warp's staging tile; the flush below writes whole 64-byte row pieces
                        const int jj = (u >> 2) & 3, slot = lane * 4 + (jj ^ ((lane >> 1) & 3));
                        stg[slot] = make_float4(out[0], out[1], out[2], out[3]);
                        if (G == 4) stg[128 + slot] = make_float4(cn[0], cn[1], cn[2], cn[3]);
                        continue;
                    }
                    if (!(p.dbg & 4)) *reinterpret_cast<float4 *>(hn + u) = make_float4(out[0], out[1], out[2], out[3]);
                    if (p.dbg & 8) {
                    } else if (H16) {
                        __half2 lo = __floats2half2_rn(out[0], out[1]), hi = __floats2half2_rn(out[2], out[3]);
                        uint2 pk = make_uint2(*reinterpret_cast<uint32_t *>(&lo), *reinterpret_cast<uint32_t *>(&hi));
                        *reinterpret_cast<uint2 *>(reinterpret_cast<__half *>(p.hr_next) + rr * p.H + ubase + u) = pk;
                    } else {
                        const float4 hi = make_float4(round_tf32(out[0]), round_tf32(out[1]), round_tf32(out[2]), round_tf32(out[3]));
                        *reinterpret_cast<float4 *>(hr + u) = hi;
                        if (X3)
                            *reinterpret_cast<float4 *>(hr + p.part_stride + u) =
                                make_float4(round_tf32(out[0] - hi.x), round_tf32(out[1] - hi.y), round_tf32(out[2] - hi.z), round_tf32(out[3] - hi.w));
                    }
                    if (G == 4 && !(p.dbg & 16)) *reinterpret_cast<float4 *>(cw + u) = make_float4(cn[0], cn[1], cn[2], cn[3]);
                }
                if (STG && (ch & 1) && !(p.dbg & 2)) {
                    // 16 units of 32 rows are parked: lane -> (row it*8 + lane/4, 16-byte piece lane%4), so one store
                    // instruction covers 8 rows x 64 contiguous bytes instead of 32 rows x 16 bytes
                    __syncwarp();
#pragma unroll
                    for (int it = 0; it < 4; ++it) {
                        const int r = it * 8 + (lane >> 2), j = lane & 3;
                        const long long grow = (long long)m_t * R_BM + q * 32 + r;
                        const int uu = ubase + (ch >> 1) * 16 + 4 * j;
                        if (grow < p.B && uu < p.H) {
                            const int slot = r * 4 + (j ^ ((r >> 1) & 3));
                            const float4 hv = stg[slot];
                            if (!(p.dbg & 4)) *reinterpret_cast<float4 *>(p.h_next + grow * p.ld_next + uu) = hv;
                            if (p.dbg & 8) {
                            } else if (H16) {
                                __half2 lo = __floats2half2_rn(hv.x, hv.y), hi = __floats2half2_rn(hv.z, hv.w);
                                *reinterpret_cast<uint2 *>(reinterpret_cast<__half *>(p.hr_next) + grow * p.H + uu) =
                                    make_uint2(*reinterpret_cast<uint32_t *>(&lo), *reinterpret_cast<uint32_t *>(&hi));
                            } else {
                                *reinterpret_cast<float4 *>(p.hr_next + grow * p.H + uu) =
                                    make_float4(round_tf32(hv.x), round_tf32(hv.y), round_tf32(hv.z), round_tf32(hv.w));
                            }
                            if (G == 4 && !(p.dbg & 16)) *reinterpret_cast<float4 *>(p.c + grow * p.H + uu) = stg[128 + slot];
                        }
                    }
                    __syncwarp();
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[buf]);
        }
    }
    tc_fence_before();
    __syncthreads();
    // neither CTA of a pair may leave while the other can still multicast into its stages or arrive on its barriers
    if (CL2) asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(SL::TMEM_COLS) : "memory");
    }
}

// Bt[(j*G + g)*64 + u][k] = tf32(U[k][g*H + 64 j + u]); padded units and k in [H, ldbt) are zero.  bhp likewise.
__global__ void pack_recurrent_weights_kernel(const float *__restrict__ U, const float *__restrict__ bh, float *__restrict__ Bt,
                                              float *__restrict__ bhp, int H, int G, int ldbt, int n_rows) {
    const long long total = (long long)n_rows * ldbt;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / ldbt), k = (int)(i - (long long)n * ldbt);
        const int j = n / (G * R_UT), r = n - j * G * R_UT, g = r / R_UT, u = r - g * R_UT;
        const int unit = j * R_UT + u;
        const bool ok = unit < H && k < H;
        Bt[i] = ok ? round_tf32(__ldg(U + (long long)k * G * H + g * H + unit)) : 0.0f;
        if (k == 0 && bhp) bhp[n] = (unit < H && bh) ? bh[g * H + unit] : 0.0f;
    }
}

// X3: row n holds [lo | hi | hi], each kpad wide
__global__ void pack_recurrent_weights_x3_kernel(const float *__restrict__ U, const float *__restrict__ bh, float *__restrict__ Bt,
                                                 float *__restrict__ bhp, int H, int G, int kpad, int n_rows) {
    const long long total = (long long)n_rows * kpad;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / kpad), k = (int)(i - (long long)n * kpad);
        const int j = n / (G * R_UT), r = n - j * G * R_UT, g = r / R_UT, u = r - g * R_UT;
        const int unit = j * R_UT + u;
        const float v = (unit < H && k < H) ? __ldg(U + (long long)k * G * H + g * H + unit) : 0.0f;
        const float hi = round_tf32(v);
        float *row = Bt + (long long)n * 3 * kpad;
        row[k] = round_tf32(v - hi); row[kpad + k] = hi; row[2 * kpad + k] = hi;
        if (k == 0 && bhp) bhp[n] = (unit < H && bh) ? bh[g * H + unit] : 0.0f;
    }
}
// X3: the caller's h_0 split into the operand ping-pong slot that step 0 reads
__global__ void split_state_kernel(const float *__restrict__ in, float *__restrict__ hi, float *__restrict__ lo, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const float v = in[i], h = round_tf32(v);
        hi[i] = h; lo[i] = round_tf32(v - h);
    }
}

__global__ void pack_recurrent_weights_h16_kernel(const float *__restrict__ U, const float *__restrict__ bh, __half *__restrict__ Bt,
                                                  float *__restrict__ bhp, int H, int G, int ldbt, int n_rows) {
    const long long total = (long long)n_rows * ldbt;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / ldbt), k = (int)(i - (long long)n * ldbt);
        const int j = n / (G * R_UT), r = n - j * G * R_UT, g = r / R_UT, u = r - g * R_UT;
        const int unit = j * R_UT + u;
        const bool ok = unit < H && k < H;
        Bt[i] = __float2half_rn(ok ? __ldg(U + (long long)k * G * H + g * H + unit) : 0.0f);
        if (k == 0 && bhp) bhp[n] = (unit < H && bh) ? bh[g * H + unit] : 0.0f;
    }
}
__global__ void to_half_kernel(const float *__restrict__ in, __half *__restrict__ out, long long n) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = __float2half_rn(in[i]);
}

float *g_rtc_scratch = nullptr;
size_t g_rtc_floats = 0;

}  // namespace

template <int G, int STAGES, bool H16, bool X3 = false, bool CL2 = false>
static int launch_rtc(const CUtensorMap &ma, const CUtensorMap &mb, const RtcParams &p, cudaStream_t st) {
    using SL = RtcSmem<G, STAGES>;
    static bool attr_done = false;
    static int max_pairs = 0;
    if (!attr_done) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(recurrent_tc_step_kernel<G, STAGES, H16, X3, CL2>, cudaFuncAttributeMaxDynamicSharedMemorySize, SL::TOTAL));
        if (CL2) {
            cudaLaunchConfig_t q = {};
            q.gridDim = dim3(2 * (nntk_cuda_sm_count() / 2)); q.blockDim = dim3(rtc_threads(X3)); q.dynamicSmemBytes = SL::TOTAL;
            cudaLaunchAttribute ca[1];
            ca[0].id = cudaLaunchAttributeClusterDimension; ca[0].val.clusterDim.x = 2; ca[0].val.clusterDim.y = 1; ca[0].val.clusterDim.z = 1;
            q.attrs = ca; q.numAttrs = 1;
            NNTK_CUDA_TRY(cudaOccupancyMaxActiveClusters(&max_pairs, recurrent_tc_step_kernel<G, STAGES, H16, X3, CL2>, &q));
            if (max_pairs < 1) { nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_tc: no 2-CTA cluster fits"); return NNTK_ERR_CUDA_; }
        }
        attr_done = true;
    }
    const int pairs = p.n_tiles * ((p.m_tiles + 1) / 2);
    const int grid = CL2 ? 2 * (pairs < max_pairs ? pairs : max_pairs)
                         : (p.total_tiles < nntk_cuda_sm_count() ? p.total_tiles : nntk_cuda_sm_count());
    // programmatic dependent launch: step t+1's CTAs may start (barrier init, TMEM allocation, descriptor prefetch) on SMs
    // that step t has already left; they wait for all of step t in griddepcontrol.wait before touching h, c or xw
    static int pdl = -1;
    if (pdl < 0) { const char *e = getenv("NNTK_RTC_PDL"); pdl = e ? atoi(e) : 1; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(rtc_threads(X3)); cfg.dynamicSmemBytes = SL::TOTAL; cfg.stream = st;
    cudaLaunchAttribute at[2];
    int na = 0;
    if (CL2) { at[na].id = cudaLaunchAttributeClusterDimension; at[na].val.clusterDim.x = 2; at[na].val.clusterDim.y = 1; at[na].val.clusterDim.z = 1; ++na; }
    if (pdl) { at[na].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[na].val.programmaticStreamSerializationAllowed = 1; ++na; }
    cfg.attrs = at; cfg.numAttrs = na;
    NNTK_CUDA_TRY(cudaLaunchKernelEx(&cfg, recurrent_tc_step_kernel<G, STAGES, H16, X3, CL2>, ma, mb, p));
    return 0;
}

static int encode_h_map(EncodeTiledFn encode, CUtensorMap *m, const void *base, int H, int n_t, long long t_stride,
                        int rows, long long row_stride, bool h16 = false) {
    const int es = h16 ? 2 : 4;
    cuuint64_t dims[3] = {(cuuint64_t)H, (cuuint64_t)n_t, (cuuint64_t)rows};
    cuuint64_t strides[2] = {(cuuint64_t)t_stride * es, (cuuint64_t)row_stride * es};
    cuuint32_t box[3] = {(cuuint32_t)(h16 ? 2 * R_BK : R_BK), 1, R_BM};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = encode(m, h16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, (void *)base, dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_tc: tensor map for h rejected (CUresult %d): H %d rows %d", (int)r, H, rows);
        return NNTK_ERR_CUDA_;
    }
    return 0;
}

// Runs all T steps of one batch chunk on the tensor cores.  *used = 0 when the shape is not eligible (caller falls back to
// the fp32 step kernels).  xw: hoisted projection [nb, T, G*H]; h/c: state [nb, H]; pp: [4, nb, H] (exact ping-pong, then
// the TF32-rounded ping-pong).
extern "C" int nntk_cuda_recurrent_tc(const nntk_recurrent_args *r, int nb, const float *xw, int xw_blocked, float *h, float *c,
                                      float *out, float *pp, int x3, int *used) {
    *used = 0;
    const int G = r->cell, H = r->H, T = r->T;
    if ((H & 3) || T < 1 || nb < 1) return 0;
    if (G == 1) return 0;   // the un-gated RNN stays fp32 in every math mode (see recurrent.cu)
    if ((((uintptr_t)h) | ((uintptr_t)out) | ((uintptr_t)xw) | ((uintptr_t)pp) | ((uintptr_t)c)) & 15) return 0;
    EncodeTiledFn encode = get_encode();
    if (!encode) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_tc: cuTensorMapEncodeTiled is not available from this driver");
        return NNTK_ERR_CUDA_;
    }
    cudaStream_t st = nntk_stream();
    const int n_tiles = (H + R_UT - 1) / R_UT;
    const int n_rows = n_tiles * G * R_UT;
    const int kpad = (H + R_BK - 1) / R_BK * R_BK;
    const int ldbt = x3 ? 3 * kpad : H;   // multiple of 4
    const size_t need = (size_t)n_rows * ldbt + n_rows + 64;
    if (need > g_rtc_floats) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_rtc_scratch) cudaFree(g_rtc_scratch);
        g_rtc_scratch = nullptr; g_rtc_floats = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_rtc_scratch, need * sizeof(float)));
        g_rtc_floats = need;
    }
    float *Bt = g_rtc_scratch;
    float *bhp = g_rtc_scratch + (((size_t)n_rows * ldbt + 3) & ~(size_t)3);
    const bool add_bh = (G == 3) || r->v2;
    // fp16 operands when h is bounded (default activations) and rows of halves stay 16-byte aligned
    static int want_h16 = -1;
    if (want_h16 < 0) { const char *e = getenv("NNTK_RTC_H16"); want_h16 = e ? atoi(e) : 1; }
    const int SG = NNTK_ACT_SIGMOID, TH = NNTK_ACT_TANH;
    const bool def = (G == 3) ? (r->act[0] == SG && r->act[1] == SG && r->act[2] == TH)
                              : (r->act[0] == SG && r->act[1] == SG && r->act[2] == TH && r->act[3] == SG && r->act[4] == TH);
    const bool h16 = !x3 && want_h16 && def && (H % 8 == 0);
    const int es = h16 ? 2 : 4;
    {
        long long total = (long long)n_rows * ldbt;
        int blocks = (int)((total + 255) / 256 < 2048 ? (total + 255) / 256 : 2048);
        if (x3) pack_recurrent_weights_x3_kernel<<<blocks, 256, 0, st>>>(r->d_U, add_bh ? r->d_bh : nullptr, Bt, bhp, H, G, kpad, n_rows);
        else if (h16) pack_recurrent_weights_h16_kernel<<<blocks, 256, 0, st>>>(r->d_U, add_bh ? r->d_bh : nullptr, reinterpret_cast<__half *>(Bt), bhp, H, G, ldbt, n_rows);
        else pack_recurrent_weights_kernel<<<blocks, 256, 0, st>>>(r->d_U, add_bh ? r->d_bh : nullptr, Bt, bhp, H, G, ldbt, n_rows);
        NNTK_LAUNCH_CHECK("pack_recurrent_weights_kernel");
    }
    CUtensorMap m_state, m_hr, mb;
    int rc;
    // NNTK_RTC_CL2: 2-CTA clusters with the U slice multicast (fp16 operands, at least two row tiles)
    static int want_cl2 = -1;
    if (want_cl2 < 0) { const char *e = getenv("NNTK_RTC_CL2"); want_cl2 = e ? atoi(e) : 0; }
    const bool cl2 = want_cl2 && h16 && !x3 && G != 1 && nb > R_BM;
    float *hr = pp + (size_t)2 * nb * H;   // rounded ping-pong [2][nb][H] behind the exact one (fp16 mode: halves, same region)
    if (h16) {
        // step 0 reads h_0 like every other step: converted into slot 1 of the ping-pong ((t - 1) & 1 for t = 0)
        const long long n = (long long)nb * H;
        to_half_kernel<<<(unsigned)((n + 255) / 256 < 2048 ? (n + 255) / 256 : 2048), 256, 0, st>>>(h, reinterpret_cast<__half *>(hr) + n, n);
        NNTK_LAUNCH_CHECK("to_half_kernel");
    } else if (x3) {
        // operand ping-pong [2 steps][hi, lo][nb][H]; step 0 reads slot pair 1 like every odd step
        const long long n = (long long)nb * H;
        split_state_kernel<<<(unsigned)((n + 255) / 256 < 2048 ? (n + 255) / 256 : 2048), 256, 0, st>>>(h, hr + 2 * n, hr + 3 * n, n);
        NNTK_LAUNCH_CHECK("split_state_kernel");
    } else if ((rc = encode_h_map(encode, &m_state, h, H, 1, H, nb, H))) return rc;
    if ((rc = encode_h_map(encode, &m_hr, hr, H, x3 ? 4 : 2, (long long)nb * H, nb, H, h16))) return rc;
    {
        cuuint64_t dims[2] = {(cuuint64_t)ldbt, (cuuint64_t)n_rows};
        cuuint64_t strides[1] = {(cuuint64_t)ldbt * es};
        // pair mode: each CTA of a 2-CTA cluster loads half of the U slice (and multicasts it), so the box is half as tall
        cuuint32_t box[2] = {(cuuint32_t)(h16 ? 2 * R_BK : R_BK), (cuuint32_t)(cl2 ? G * R_UT / 2 : G * R_UT)};
        cuuint32_t estr[2] = {1, 1};
        CUresult cr = encode(&mb, h16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)Bt, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) {
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_tc: tensor map for U rejected (CUresult %d)", (int)cr);
            return NNTK_ERR_CUDA_;
        }
    }
    RtcParams p;
    p.B = nb; p.H = H; p.n_tiles = n_tiles; p.m_tiles = (nb + R_BM - 1) / R_BM;
    p.kb_pass = h16 ? (H + 2 * R_BK - 1) / (2 * R_BK) : (H + R_BK - 1) / R_BK;
    p.k_blocks = x3 ? 3 * p.kb_pass : p.kb_pass;
    p.part_stride = (long long)nb * H;
    p.total_tiles = p.n_tiles * p.m_tiles;
    p.bh = add_bh ? bhp : nullptr;
    p.ld_xw = xw_blocked ? 0 : (long long)T * G * H;
    p.xw_blocked = xw_blocked; p.row_blocks = (nb + 31) / 32;
    { const char *e = getenv("NNTK_RTC_DBG"); p.dbg = e ? atoi(e) : 0; }
    p.c = c;
    for (int i = 0; i < 5; ++i) { p.act[i] = r->act[i]; p.act_a[i] = r->act_a[i]; }
    static int pf = -1;
    // NNTK_RTC_PF=1: each step pulls the next step's projection slab into L2 while its first MMAs run.  Measured on C4:
    // B = 4096 50.6-51.4 -> 53.6-54.4 ms (the 33 MB slab competes with this step's own reads), B = 2048 28.0 -> 27.7 ms.  Off.
    if (pf < 0) { const char *e = getenv("NNTK_RTC_PF"); pf = e ? atoi(e) : 0; }
    for (int t = 0; t < T; ++t) {
        const CUtensorMap *ma;
        const float x3_comp = 1.0f + 2.06e-9f * (float)H;   // truncating accumulate, see gemm_tc.cu (X3_ACC_TRUNC_PER_K)
        if (t == 0 && x3) { ma = &m_hr; p.a_t = 1; p.h_prev = h; p.ld_prev = H; p.comp = x3_comp; }
        else if (t == 0 && h16) { ma = &m_hr; p.a_t = 1; p.h_prev = h; p.ld_prev = H; p.comp = 1.0f; }
        else if (t == 0) { ma = &m_state; p.a_t = 0; p.h_prev = h; p.ld_prev = H; p.comp = 1.0f + TF32_TRUNC_BIAS; }
        else {
            ma = &m_hr; p.a_t = (t - 1) & 1; p.comp = x3 ? x3_comp : 1.0f;
            if (r->return_sequences) { p.h_prev = out + (long long)(t - 1) * H; p.ld_prev = (long long)T * H; }
            else { p.h_prev = pp + (size_t)((t - 1) & 1) * nb * H; p.ld_prev = H; }
        }
        if (r->return_sequences) { p.h_next = out + (long long)t * H; p.ld_next = (long long)T * H; }
        else { p.h_next = pp + (size_t)(t & 1) * nb * H; p.ld_next = H; }
        p.hr_next = h16 ? reinterpret_cast<float *>(reinterpret_cast<__half *>(hr) + (size_t)(t & 1) * nb * H)
                        : hr + (size_t)(t & 1) * (x3 ? 2 : 1) * nb * H;
        p.xw = xw_blocked ? xw + (long long)t * p.row_blocks * 32 * G * H : xw + (long long)t * G * H;
        p.xw_next = (xw_blocked && pf && t + 1 < T) ? p.xw + (long long)p.row_blocks * 32 * G * H : nullptr;
        p.pf_lines = (long long)p.row_blocks * 32 * G * H / 32;
        if (G == 1) rc = launch_rtc<1, 6, false>(*ma, mb, p, st);
        else if (G == 3) rc = x3 ? launch_rtc<3, 4, false, true>(*ma, mb, p, st) : cl2 ? launch_rtc<3, 4, true, false, true>(*ma, mb, p, st) : h16 ? launch_rtc<3, 4, true>(*ma, mb, p, st) : launch_rtc<3, 4, false>(*ma, mb, p, st);
        else rc = x3 ? launch_rtc<4, 4, false, true>(*ma, mb, p, st) : cl2 ? launch_rtc<4, 4, true, false, true>(*ma, mb, p, st) : h16 ? launch_rtc<4, 4, true>(*ma, mb, p, st) : launch_rtc<4, 4, false>(*ma, mb, p, st);
        if (rc) return rc;
    }
    NNTK_LAUNCH_CHECK_N("recurrent_tc_step_kernel", T);
    *used = 1;
    return 0;
}
