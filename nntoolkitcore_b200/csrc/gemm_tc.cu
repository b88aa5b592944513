// gemm_tc.cu -- K2/K3, tensor-core path (NNTK_MATH_TF32): (implicit) GEMM on tcgen05 with TMA-staged tiles and TMEM
// accumulators, bias + BatchNorm + activation fused in the epilogue.  Same contract as csrc/gemm.cu (nntk_gemm_args);
// operands are rounded to TF32 by the tensor core, accumulation is fp32, so parity with the reference is 1e-3, not 1e-5.
// Split-operand variants of the same kernel: 3xTF32 (NNTK_MATH_TF32X3: hi/lo tf32 parts, three passes in one K loop, the
// truncating accumulate compensated -- holds 1e-5) and 3xbf16 (kind::f16, the hoisted recurrent projection of the TF32 mode).
//
// Structure (persistent, warp specialised, one CTA per SM):
//   warp 0      TMA producer: A tile [128 rows x 32 k] and B tile [BN x 32 k] per stage, SWIZZLE_128B, mbarrier tx-count
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma.kind::tf32 (M=128, N=BN, K=8) x4 per stage into one of two
//               TMEM accumulator buffers; tcgen05.commit releases the smem stage / publishes the accumulator
//   warps 2-9   epilogue: two warps per TMEM lane quarter, each draining half of the tile's columns in 32-column chunks:
//               tcgen05.ld, v*scale[n]+shift[n] (bias and BatchNorm folded per channel; read as warp-broadcast 16-byte
//               loads issued under the TMEM read), activation, 16-byte stores; overlaps the next tile's main loop through
//               the second TMEM buffer.  (ncu, r1: with 4 warps and per-element scalar scale/shift loads + bound checks the
//               MMA issuer spent its time waiting on tmem_empty -- the epilogue, not TMA or the tensor pipe, set the pace.)
// Conv1d needs no im2col: with channels-last input the im2col row x is the contiguous span in[x*stride : x*stride+K, :],
// so A is a 3-D TMA tensor {K*Cin, T_out, batch} with OVERLAPPING rows (row stride = stride*Cin floats).
// Weights are re-laid K-major (Bt[N][K]) per call -- they may legally change between Applies.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include "tc_common.cuh"

namespace {

constexpr int BM = 128, BK = 32;          // BK floats = 128 bytes = one SWIZZLE_128B atom
constexpr int UMMA_K = 8;                 // tf32: 32 bytes per MMA along K
constexpr int NUM_THREADS = 320;         // TMA warp, MMA warp, 8 epilogue warps
constexpr int EPI_WARPS = 8;

struct TcParams {
    int N, n_tiles, m_tiles_per_slab, k_blocks;   // k_blocks: total over all passes (3x in split mode)
    int kb_pass;                                  // k-blocks per pass = ceil(K / 32)
    int c_blocked;                                // 0: C row-major.  1: recurrent layout [slab][row/32][N/4][row%32][4] (see nntk_cuda_gemm_tf32_ex)
    int row_blocks;                               // c_blocked: ceil(rows / 32)
    long long total_tiles;
    int rows;
    float *C;
    long long c_slab;
    int ldc;
    const float *scale, *shift;   // folded epilogue: y = act(acc * scale[n] + shift[n])
    int act;
    float act_a;
    int half_fmt;                 // BF16 kernels: operand format of kind::f16 -- 1 bf16 (3xbf16 split), 0 fp16 (single pass, fp16 front end)
    int cs;                       // cluster size (1, 2 or 4; needs n_tiles == 1): the CTAs of a cluster walk their tiles in lockstep and
                                  // each fetches 1 / cs of every B stage, multicast to all -- the weight stream leaves L2 once per cluster
    const int *gate;              // when set the kernel runs only if *gate == gate_run_if (fp16 front end and its fp32 fallback)
    int gate_run_if;
};

template <int BN, int STAGES, bool CG2 = false>
struct SmemLayout {
    static constexpr int A_BYTES = BM * BK * 4;
    static constexpr int B_BYTES = (CG2 ? BN / 2 : BN) * BK * 4;     // CG2: this CTA's half of the pair's B tile
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES;
    static constexpr int STG_OFFSET = BAR_OFFSET + 256;                         // epilogue staging: 8 warps x 2 KB
    static constexpr int TOTAL = STG_OFFSET + 8 * 2048 + 1024;                  // + alignment slack
};

__device__ __forceinline__ uint32_t tc_cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void tc_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// 2-D tile load multicast to the CTAs of `mask`: same shared-memory offset and same mbarrier offset in each of them
__device__ __forceinline__ void tma_load_2d_mc(void *dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
// commit arriving on the barrier at this offset in every CTA of `mask`
__device__ __forceinline__ void tc_commit_mc(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}

// ---- cta_group::2 (CG2 kernels): a CTA pair works on one 256-row tile; each CTA stages its own 128 rows of A and HALF of the B tile
__device__ __forceinline__ uint32_t tc_mapa(uint32_t a, uint32_t rank) {
    uint32_t d; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(rank)); return d;
}
// loads executed by BOTH CTAs; the bytes are counted on the barrier at cluster address bar_cl (the leader's)
__device__ __forceinline__ void tma_load_4d_cg2(void *dst, const CUtensorMap *map, uint32_t bar_cl, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(bar_cl), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_2d_cg2(void *dst, const CUtensorMap *map, uint32_t bar_cl, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(bar_cl), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_mma_cg2(bool f16, uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    if (f16) asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
    else asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tc_commit_cg2(uint64_t *bar) {       // arrives on the barrier at this offset in both CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_arrive_cl(uint32_t a) { asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(a) : "memory"); }

// BF16 = true: the operands are bf16 (kind::f16, K = 16 per MMA, 64 elements per 128-byte k-block) -- the 3xbf16 split mode
// of the hoisted recurrent projection.  Byte-wise the stages, swizzle and descriptors are the same as for tf32.
// CG2 = true: tcgen05.mma.cta_group::2, M = 256 per CTA pair (launched as clusters of 2): each CTA stages its own 128 rows of A and
// half of the B tile, so per MMA an SM reads 8 KB of operands instead of 12 and TMA writes 32 KB per k-block instead of 48 --
// shared-memory bandwidth (reads + TMA writes ~ 190 B/cycle wanted, 128 there) was what held the one-CTA kernel at ~70 % tensor
// activity.  The leader CTA issues; accumulator rows 128 r .. 128 r + 127 live in CTA r's TMEM; each CTA runs its own epilogue.
template <int BN, int STAGES, bool FAST, bool BF16, bool CG2 = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, TcParams p) {
    using SL = SmemLayout<BN, STAGES, CG2>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *full_bar = reinterpret_cast<uint64_t *>(smem + SL::BAR_OFFSET);
    uint64_t *empty_bar = full_bar + STAGES;
    uint64_t *tmem_full = empty_bar + STAGES;     // [2]
    uint64_t *tmem_empty = tmem_full + 2;         // [2]
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (p.gate && __ldg(p.gate) != p.gate_run_if) return;      // the whole grid takes the same branch, before any barrier / TMEM state exists
    constexpr uint32_t TMEM_COLS = 2 * BN;        // two accumulator buffers (256 or 512 columns: powers of two)
    constexpr int BKE = BF16 ? 2 * BK : BK;       // elements per 128-byte k-block

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], CG2 ? 1u : (uint32_t)p.cs); }
        for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full[b], 1); mbar_init(&tmem_empty[b], CG2 ? 2 * EPI_WARPS : EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if (CG2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CG2 || p.cs > 1) tc_cluster_sync();            // every CTA's barriers exist before a peer's multicast or commit can reach them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // Clusters walk in lockstep: every CTA runs the same number of tiles; surplus iterations redo the last tile without storing it.
    // CG2: the walkers are CTA pairs, total_tiles counts 256-row pair tiles and CTA r of the pair owns 128-row tile 2 m + r of the slab
    const uint32_t pr = CG2 ? (tc_cluster_rank() & 1u) : 0u;
    const long long walkers = CG2 ? (gridDim.x >> 1) : gridDim.x, me = CG2 ? (blockIdx.x >> 1) : blockIdx.x;
    const long long n_iter = (p.total_tiles + walkers - 1) / walkers;
    const uint16_t mc_mask = (uint16_t)((1u << p.cs) - 1u);

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            const uint32_t crank = p.cs > 1 ? tc_cluster_rank() : 0u;
            const int b_rows = BN / p.cs;               // this CTA's slice of every B stage
            for (long long i = 0; i < n_iter; ++i) {
                long long tile = me + i * walkers;
                if (tile >= p.total_tiles) tile = p.total_tiles - 1;
                const int n_t = (int)(tile % p.n_tiles);
                const long long rest = tile / p.n_tiles;
                const int m_t = (CG2 ? 2 : 1) * (int)(rest % p.m_tiles_per_slab) + (int)pr;
                const int slab = (int)(rest / p.m_tiles_per_slab);
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t *sa = smem + stage * SL::STAGE_BYTES;
                    if (CG2) {       // both CTAs' bytes are counted on the LEADER's barrier
                        const uint32_t full_cl = tc_mapa(smem_u32(&full_bar[stage]), tc_cluster_rank() & ~1u);
                        if (pr == 0) mbar_expect_tx(&full_bar[stage], 2 * SL::STAGE_BYTES);
                        const int pass = kb / p.kb_pass, akb = kb - pass * p.kb_pass;
                        tma_load_4d_cg2(sa, &map_a, full_cl, akb * BKE, m_t * BM, slab, pass == 1 ? 1 : 0);
                        tma_load_2d_cg2(sa + SL::A_BYTES, &map_b, full_cl, kb * BKE, n_t * BN + (int)pr * (BN / 2));
                        if (++stage == STAGES) { stage = 0; phase ^= 1; }
                        continue;
                    }
                    mbar_expect_tx(&full_bar[stage], SL::STAGE_BYTES);
                    // split (3xTF32) mode: passes hi*Wlo, lo*Whi, hi*Whi -- B' = [Wlo; Whi; Whi] along K, A = hi, lo, hi.  The
                    // two small passes come FIRST: the tensor core's fp32 accumulate truncates (measured: error grows
                    // with the number of accumulation steps at full magnitude), so they run while the sum is still tiny.
                    const int pass = kb / p.kb_pass, akb = kb - pass * p.kb_pass;
                    tma_load_4d(sa, &map_a, &full_bar[stage], akb * BKE, m_t * BM, slab, pass == 1 ? 1 : 0);
                    if (p.cs == 1) tma_load_2d(sa + SL::A_BYTES, &map_b, &full_bar[stage], kb * BKE, n_t * BN);
                    else tma_load_2d_mc(sa + SL::A_BYTES + crank * b_rows * 128, &map_b, &full_bar[stage], kb * BKE, n_t * BN + (int)crank * b_rows, mc_mask);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (CG2: the pair's leader only) =====
        if (lane == 0 && pr == 0) {
            // instruction descriptor: D=f32, A=B=tf32, both K-major, N>>3 at bit 17, M>>4 at bit 24
            // (bf16: a/b format 1, kind::f16)
            const uint32_t fmt = BF16 ? (uint32_t)p.half_fmt : 2u;
            const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((CG2 ? 2 * BM : BM) >> 4) << 24);
            int stage = 0; uint32_t phase = 0;
            uint32_t it = 0;
            for (long long i = 0; i < n_iter; ++i, ++it) {
                const uint32_t buf = it & 1, use = it >> 1;
                mbar_wait(&tmem_empty[buf], (use & 1) ^ 1);     // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + buf * BN;
                for (int kb = 0; kb < p.k_blocks; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(smem + stage * SL::STAGE_BYTES);
                    const uint64_t da = make_desc(sa), db = make_desc(sa + SL::A_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        // advance 32 bytes along K inside the 128-byte swizzle atom: +2 in the (addr >> 4) field
                        if (CG2) tc_mma_cg2(BF16, tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        else if (BF16) tc_mma_f16(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                        else tc_mma_tf32(tmem_d, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                    }
                    if (CG2) tc_commit_cg2(&empty_bar[stage]);     // frees the stage in both CTAs of the pair
                    else if (p.cs == 1) tc_commit(&empty_bar[stage]);   // smem stage free once these MMAs have read it
                    else tc_commit_mc(&empty_bar[stage], mc_mask); // ... in every CTA of the cluster: each of them writes part of it
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                if (CG2) tc_commit_cg2(&tmem_full[buf]);         // accumulator complete: each CTA's epilogue drains its 128 rows
                else tc_commit(&tmem_full[buf]);
            }
        }
    } else {
        // ===== epilogue warps (2..9): TMEM lane quarter = warp % 4, column half = (warp - 2) / 4 =====
        const int q = warp & 3;
        const int half = (warp - 2) >> 2;
        const int row_in_tile = q * 32 + lane;
        uint8_t *stg = smem + SL::STG_OFFSET + (warp - 2) * 2048;      // [32 rows][64 bytes], xor-swizzled 16-byte slots
        uint32_t it = 0;
        const bool vec_ok = ((p.ldc & 3) == 0) && ((p.c_slab & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0);
        const uint32_t tmem_empty_leader = CG2 ? tc_mapa(smem_u32(&tmem_empty[0]), tc_cluster_rank() & ~1u) : 0u;
        for (long long i = 0; i < n_iter; ++i, ++it) {
            const long long tile = me + i * walkers;
            const uint32_t buf = it & 1, use = it >> 1;
            if (tile >= p.total_tiles) {             // lockstep filler: nothing to store
                mbar_wait(&tmem_full[buf], use & 1);
                tc_fence_after();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) { if (CG2) tc_arrive_cl(tmem_empty_leader + 8u * buf); else mbar_arrive(&tmem_empty[buf]); }
                continue;
            }
            const int n_t = (int)(tile % p.n_tiles);
            const long long rest = tile / p.n_tiles;
            const int m_t = (CG2 ? 2 : 1) * (int)(rest % p.m_tiles_per_slab) + (int)pr;
            const long long slab = rest / p.m_tiles_per_slab;
            const int row = m_t * BM + row_in_tile;
            const bool row_ok = row < p.rows;
            float *crow = p.C + slab * p.c_slab + (long long)row * p.ldc;
            float *cwarp = p.C + slab * p.c_slab + (long long)(m_t * BM + q * 32) * p.ldc;   // first row of this warp's 32
            const uint32_t taddr = tmem_base + buf * BN + ((uint32_t)(q * 32) << 16);
            mbar_wait(&tmem_full[buf], use & 1);
            tc_fence_after();
            // BN = 32: one 32-column chunk per tile -- the two warps of a lane quarter take alternate tiles (= alternate accumulator
            // buffers), so all eight epilogue warps work (drained by column-half 0 alone, the epilogue bounded the 12-wide TDD)
            constexpr int HALF_COLS = BN >= 64 ? BN / 2 : BN;
            const int c_begin = BN >= 64 ? half * HALF_COLS : (((int)(it & 1) == half) ? 0 : BN);
#pragma unroll 1
            for (int c0 = c_begin; c0 < (BN >= 64 ? (half + 1) * HALF_COLS : BN); c0 += 32) {
                const int n0 = n_t * BN + c0;
                if (n0 >= p.N) break;                       // warp-uniform
                uint32_t r[32];
                if (FAST) {
                    // N % 32 == 0 and 16-byte aligned rows: the only predicate left is the row bound.  scale / shift come as
                    // warp-broadcast 16-byte loads issued BEFORE the TMEM read is waited for.
                    tc_ld_32x32_nowait(taddr + c0, r);
                    float4 sc[8], sh[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        sc[j] = __ldg(reinterpret_cast<const float4 *>(p.scale + n0) + j);
                        sh[j] = __ldg(reinterpret_cast<const float4 *>(p.shift + n0) + j);
                    }
                    tc_ld_wait();
                    float4 vv[8];
                    if (p.act == NNTK_ACT_RELU) {
                        const float a = p.act_a;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float4 v;
                            v.x = fmaxf(fmaf(__uint_as_float(r[4 * j + 0]), sc[j].x, sh[j].x), 0.0f) * a;
                            v.y = fmaxf(fmaf(__uint_as_float(r[4 * j + 1]), sc[j].y, sh[j].y), 0.0f) * a;
                            v.z = fmaxf(fmaf(__uint_as_float(r[4 * j + 2]), sc[j].z, sh[j].z), 0.0f) * a;
                            v.w = fmaxf(fmaf(__uint_as_float(r[4 * j + 3]), sc[j].w, sh[j].w), 0.0f) * a;
                            vv[j] = v;
                        }
                    } else if (p.act == NNTK_ACT_SIGMOID || p.act == NNTK_ACT_TANH) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float4 v;
                            v.x = nntk_act(p.act, 1.0f, fmaf(__uint_as_float(r[4 * j + 0]), sc[j].x, sh[j].x));
                            v.y = nntk_act(p.act, 1.0f, fmaf(__uint_as_float(r[4 * j + 1]), sc[j].y, sh[j].y));
                            v.z = nntk_act(p.act, 1.0f, fmaf(__uint_as_float(r[4 * j + 2]), sc[j].z, sh[j].z));
                            v.w = nntk_act(p.act, 1.0f, fmaf(__uint_as_float(r[4 * j + 3]), sc[j].w, sh[j].w));
                            vv[j] = v;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float4 v;
                            v.x = fmaf(__uint_as_float(r[4 * j + 0]), sc[j].x, sh[j].x);
                            v.y = fmaf(__uint_as_float(r[4 * j + 1]), sc[j].y, sh[j].y);
                            v.z = fmaf(__uint_as_float(r[4 * j + 2]), sc[j].z, sh[j].z);
                            v.w = fmaf(__uint_as_float(r[4 * j + 3]), sc[j].w, sh[j].w);
                            vv[j] = v;
                        }
                    }
                    if (p.c_blocked) {
                        // recurrent layout: lanes (consecutive rows) are contiguous 16-byte slots -> 512-byte stores, no staging
                        float4 *dstb = reinterpret_cast<float4 *>(p.C) +
                                       (((long long)slab * p.row_blocks + (m_t * (BM / 32) + q)) * (p.N >> 2) + (n0 >> 2)) * 32 + lane;
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            if (row_ok) dstb[j * 32] = vv[j];
                        continue;
                    }
                    // Store through a per-warp 2 KB staging tile so that one store instruction covers 8 rows x 64 contiguous
                    // bytes (8 wavefronts) instead of 32 rows x 16 bytes (32 wavefronts): the row-per-lane TMEM layout made the
                    // LSU, not the tensor pipe, the limiter of short-K GEMMs (ncu r1: 42 % LSU wavefronts at K = 1280).
#pragma unroll
                    for (int hh = 0; hh < 2; ++hh) {
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            *reinterpret_cast<float4 *>(stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) = vv[hh * 4 + j];
                        __syncwarp();
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const int rw = (lane >> 2) + 8 * i, c4 = lane & 3;
                            const float4 w = *reinterpret_cast<const float4 *>(stg + rw * 64 + ((c4 ^ ((rw >> 1) & 3)) << 4));
                            if (m_t * BM + q * 32 + rw < p.rows)
                                *reinterpret_cast<float4 *>(cwarp + (long long)rw * p.ldc + n0 + hh * 16 + c4 * 4) = w;
                        }
                        __syncwarp();
                    }
                } else {
                    tc_ld_32x32(taddr + c0, r);
                    if (BN == 32 && p.act == NNTK_ACT_SOFTMAX) {
                        // the whole output row (N <= 32) sits in this thread: softmax over it here instead of a second pass over
                        // the result (activation_default.c:149-167: e_j = exp(x_j), y_j = e_j / sum, summed in index order, no max shift)
                        if (row_ok) {
                            float e[32], sum = 0.0f;
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const int n = n0 + j;
                                e[j] = 0.0f;
                                if (n < p.N) { e[j] = expf(fmaf(__uint_as_float(r[j]), __ldg(p.scale + n), __ldg(p.shift + n))); sum += e[j]; }
                            }
#pragma unroll
                            for (int j = 0; j < 32; j += 4) {
                                const int n = n0 + j;
                                const float v[4] = {__fdividef(e[j], sum), __fdividef(e[j + 1], sum), __fdividef(e[j + 2], sum), __fdividef(e[j + 3], sum)};
                                if (vec_ok && n + 3 < p.N) *reinterpret_cast<float4 *>(crow + n) = make_float4(v[0], v[1], v[2], v[3]);
                                else
                                    for (int q2 = 0; q2 < 4; ++q2) if (n + q2 < p.N) crow[n + q2] = v[q2];
                            }
                        }
                    } else if (row_ok) {
#pragma unroll
                        for (int j = 0; j < 32; j += 4) {
                            float v[4];
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const int n = n0 + j + e;
                                float x = __uint_as_float(r[j + e]);
                                if (n < p.N) x = nntk_act(p.act, p.act_a, fmaf(x, __ldg(p.scale + n), __ldg(p.shift + n)));
                                v[e] = x;
                            }
                            const int n = n0 + j;
                            if (vec_ok && n + 3 < p.N) *reinterpret_cast<float4 *>(crow + n) = make_float4(v[0], v[1], v[2], v[3]);
                            else
                                for (int e = 0; e < 4; ++e) if (n + e < p.N) crow[n + e] = v[e];
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) { if (CG2) tc_arrive_cl(tmem_empty_leader + 8u * buf); else mbar_arrive(&tmem_empty[buf]); }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (CG2 || p.cs > 1) tc_cluster_sync();     // nobody leaves while a peer's multicast, commit or MMA could still touch this CTA
    if (warp == 1) {
        tc_fence_after();
        if (CG2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

// Bt[n][k] = tf32(B[k][n]), leading dimension ldbt (multiple of 4); tail columns k in [K, kpad) zeroed.  Split mode
// (parts == 3): row n holds [lo | hi | hi], each kpad wide, hi = rn_tf32(B), lo = rn_tf32(B - hi).
__global__ void transpose_weights_kernel(const float *__restrict__ B, float *__restrict__ Bt, int K, int N, int kpad, int ldbt,
                                         int parts) {
    __shared__ float tile[32][33];
    const int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int k = k0 + i, n = n0 + threadIdx.x;
        tile[i][threadIdx.x] = (k < K && n < N) ? B[(long long)k * N + n] : 0.0f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int n = n0 + i, k = k0 + threadIdx.x;
        if (n < N && k < kpad) {
            const float v = tile[threadIdx.x][i], hi = round_tf32(v);
            float *row = Bt + (long long)n * ldbt;
            if (parts == 3) { row[k] = round_tf32(v - hi); row[kpad + k] = hi; row[2 * kpad + k] = hi; }
            else row[k] = hi;
        }
    }
}

// split mode: hi = rn_tf32(x), lo = rn_tf32(x - hi) over the whole extent of A (x - hi is exact in fp32)
__global__ void split_tf32_kernel(const float *__restrict__ x, float *__restrict__ hi, float *__restrict__ lo, long long n) {
    const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(x) + i);
        float4 h, l;
        h.x = round_tf32(v.x); h.y = round_tf32(v.y); h.z = round_tf32(v.z); h.w = round_tf32(v.w);
        l.x = round_tf32(v.x - h.x); l.y = round_tf32(v.y - h.y); l.z = round_tf32(v.z - h.z); l.w = round_tf32(v.w - h.w);
        reinterpret_cast<float4 *>(hi)[i] = h;
        reinterpret_cast<float4 *>(lo)[i] = l;
    }
    for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float v = x[i], h = round_tf32(v);
        hi[i] = h; lo[i] = round_tf32(v - h);
    }
}

// 3xbf16 split: hi = rn_bf16(x), lo = rn_bf16(x - hi) (x - hi is exact in fp32; what is dropped is below 2^-17 |x|).
__device__ __forceinline__ void bf16_split(float v, __nv_bfloat16 &h, __nv_bfloat16 &l) {
    h = __float2bfloat16_rn(v);
    l = __float2bfloat16_rn(v - __bfloat162float(h));
}
__global__ void split_bf16_kernel(const float *__restrict__ x, __nv_bfloat16 *__restrict__ hi, __nv_bfloat16 *__restrict__ lo, long long n,
                                  const int *gate, int gate_run_if) {
    if (gate && __ldg(gate) != gate_run_if) return;
    const long long n4 = n >> 2, stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 v = __ldg(reinterpret_cast<const float4 *>(x) + i);
        __nv_bfloat16 h[4], l[4];
        bf16_split(v.x, h[0], l[0]); bf16_split(v.y, h[1], l[1]); bf16_split(v.z, h[2], l[2]); bf16_split(v.w, h[3], l[3]);
        reinterpret_cast<uint2 *>(hi)[i] = *reinterpret_cast<const uint2 *>(h);
        reinterpret_cast<uint2 *>(lo)[i] = *reinterpret_cast<const uint2 *>(l);
    }
    for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) bf16_split(x[i], hi[i], lo[i]);
}
// Bt[n] = [lo | hi | hi] of B[:, n] in bf16, each part kpad (multiple of 64) wide, zero tail
__global__ void transpose_weights_bf16_kernel(const float *__restrict__ B, __nv_bfloat16 *__restrict__ Bt, int K, int N, int kpad, int ldbt) {
    __shared__ float tile[32][33];
    const int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int k = k0 + i, n = n0 + threadIdx.x;
        tile[i][threadIdx.x] = (k < K && n < N) ? B[(long long)k * N + n] : 0.0f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int n = n0 + i, k = k0 + threadIdx.x;
        if (n < N && k < kpad) {
            __nv_bfloat16 h, l;
            bf16_split(tile[threadIdx.x][i], h, l);
            __nv_bfloat16 *row = Bt + (long long)n * ldbt;
            row[k] = l; row[kpad + k] = h; row[2 * kpad + k] = h;
        }
    }
}

// Bt[n][k] = fp16(B[k][n]), zero tail up to kpad; *ovf = 1 when a weight does not fit fp16 (single-pass fp16 front end)
__global__ void transpose_weights_f16_kernel(const float *__restrict__ B, __half *__restrict__ Bt, int K, int N, int kpad, int *ovf) {
    __shared__ float tile[32][33];
    const int k0 = blockIdx.x * 32, n0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int k = k0 + i, n = n0 + threadIdx.x;
        tile[i][threadIdx.x] = (k < K && n < N) ? B[(long long)k * N + n] : 0.0f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int n = n0 + i, k = k0 + threadIdx.x;
        if (n < N && k < kpad) {
            const float v = tile[threadIdx.x][i];
            if (!(fabsf(v) <= 65504.0f)) *ovf = 1;
            Bt[(long long)n * kpad + k] = __float2half_rn(v);
        }
    }
}

// scale[n], shift[n] such that act(acc*scale + shift) == act(BN(acc + bias)) up to rounding; the scale also carries the
// compensation of the A-operand truncation bias (see TF32_TRUNC_BIAS).
__global__ void fold_epilogue_kernel(const float *bias, const float *bn, float eps, int N, float *scale, float *shift, float comp) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float b = bias ? bias[n] : 0.0f;
    if (bn) {
        const float s = bn[n] / sqrtf(bn[3 * N + n] + eps);
        scale[n] = s * comp;
        shift[n] = (b - bn[2 * N + n]) * s + bn[N + n];
    } else {
        scale[n] = comp;
        shift[n] = b;
    }
}

constexpr float X3_ACC_TRUNC_PER_K = 2.06e-9f;

float *g_scratch = nullptr;      // Bt | scale | shift
size_t g_scratch_floats = 0;
float *g_split = nullptr;        // split mode: hi | lo copies of A
size_t g_split_floats = 0;
int g_tc_cluster = 1;            // cluster size requested for the NEXT launch (its B tensor map must have been encoded with BN / cs rows)
int g_tc_row_softmax = 0;        // the next launch fuses a softmax over the whole output row (N <= 32) into its epilogue
const int *g_tc_gate = nullptr;  // applied to every tensor-core GEMM launch while set (nntk_cuda_gemm_tc_gate)
int g_tc_gate_run_if = 0;

}  // namespace

template <int BN, int STAGES, bool FAST, bool BF16 = false>
static int launch_tc(const CUtensorMap &ma, const CUtensorMap &mb, TcParams &p, cudaStream_t st) {
    using SL = SmemLayout<BN, STAGES>;
    static bool attr_done = false;
    if (!attr_done) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(gemm_tf32_kernel<BN, STAGES, FAST, BF16>, cudaFuncAttributeMaxDynamicSharedMemorySize, SL::TOTAL));
        attr_done = true;
    }
    long long grid = p.total_tiles < nntk_cuda_sm_count() ? p.total_tiles : nntk_cuda_sm_count();
    p.gate = g_tc_gate; p.gate_run_if = g_tc_gate_run_if;
    p.cs = g_tc_cluster;
    g_tc_cluster = 1;
    if (p.cs > 1 && (p.n_tiles != 1 || grid < p.cs || (BN / p.cs) % 8)) {        // the caller encoded B for slices: it must have checked
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "gemm_tc: cluster of %d requested for a launch that cannot take it", p.cs);
        return NNTK_ERR_CUDA_;
    }
    if (p.cs > 1) {
        grid -= grid % p.cs;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(NUM_THREADS); cfg.dynamicSmemBytes = SL::TOTAL; cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)p.cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        NNTK_CUDA_TRY(cudaLaunchKernelEx(&cfg, gemm_tf32_kernel<BN, STAGES, FAST, BF16>, ma, mb, p));
        nntk_cuda_count_launch(1);
        return 0;
    }
    gemm_tf32_kernel<BN, STAGES, FAST, BF16><<<(unsigned)grid, NUM_THREADS, SL::TOTAL, st>>>(ma, mb, p);
    NNTK_LAUNCH_CHECK("gemm_tf32_kernel");
    return 0;
}

// cta_group::2 launch: clusters of 2, p.total_tiles / p.m_tiles_per_slab count 256-row PAIR tiles, the B map's box is BN / 2 rows
template <bool BF16>
static int launch_tc_cg2(const CUtensorMap &ma, const CUtensorMap &mb, TcParams &p, cudaStream_t st) {
    constexpr int BN = 256, STAGES = 6;
    using SL = SmemLayout<BN, STAGES, true>;
    auto kern = gemm_tf32_kernel<BN, STAGES, true, BF16, true>;
    static bool attr_done = false;
    if (!attr_done) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SL::TOTAL));
        attr_done = true;
    }
    long long pairs = nntk_cuda_sm_count() / 2;
    if (p.total_tiles < pairs) pairs = p.total_tiles;
    p.gate = g_tc_gate; p.gate_run_if = g_tc_gate_run_if;
    p.cs = 1;
    g_tc_cluster = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * pairs)); cfg.blockDim = dim3(NUM_THREADS); cfg.dynamicSmemBytes = SL::TOTAL; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    NNTK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, ma, mb, p));
    nntk_cuda_count_launch(1);
    return 0;
}
// whether a launch takes the cta_group::2 kernel: 256-wide n tiles, the vectorised epilogue, and enough 256-row tiles to fill the pairs
static bool want_cg2(int BN, bool fast, int rows, int batch, int n_tiles) {
    static int on = -1;
    if (on < 0) { const char *e = getenv("NNTK_GEMM_CG2"); on = e ? atoi(e) : 1; }
    const long long pair_tiles = (long long)((rows + 2 * BM - 1) / (2 * BM)) * batch * n_tiles;
    return on && BN == 256 && fast && pair_tiles >= nntk_cuda_sm_count();
}

// returns 0 done, 1 not eligible (caller runs the fp32 kernel), <0 error.  split != 0: 3xTF32 (hi/lo operand split, three
// passes folded into ONE K loop of length 3K -- the kernel is unchanged), fp32-class accuracy at a third of the TF32 rate.
extern "C" int nntk_cuda_gemm_tf32_ex(const nntk_gemm_args *a, int split, int c_blocked);
extern "C" int nntk_cuda_gemm_tf32(const nntk_gemm_args *a, int split) { return nntk_cuda_gemm_tf32_ex(a, split, 0); }

// c_blocked = 1 (internal, used by the hoisted recurrent projection): slabs are TIMESTEPS and rows are SEQUENCES
// (a_slab_stride = in, lda = T*in -- interleaved slabs are fine for TMA), and C is written as
//   C[slab][row / 32][N / 4][row % 32][4 floats]
// so that a warp whose lanes are 32 consecutive sequences -- the TMEM row-per-lane layout here AND in the persistent recurrent
// kernel that reads it back -- touches 512 contiguous bytes per 16-byte access instead of 32 different lines.  Needs N % 32 == 0.
extern "C" int nntk_cuda_gemm_tf32_ex(const nntk_gemm_args *a, int split, int c_blocked) {
    if (c_blocked && (a->N % 32 != 0 || (((uintptr_t)a->d_C) & 15))) return 1;
    if ((a->lda & 3) || (a->a_slab_stride & 3) || (((uintptr_t)a->d_A) & 15) || a->rows < 1 || a->batch < 1) return 1;
    EncodeTiledFn encode = get_encode();
    if (!encode) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "gemm_tf32: cuTensorMapEncodeTiled is not available from this driver");
        return NNTK_ERR_CUDA_;
    }
    cudaStream_t st = nntk_stream();
    const int K = a->K, N = a->N;
    // split: 0 single TF32 pass, 1 3xTF32 (fp32 operands split into tf32 hi/lo), 2 3xbf16 (bf16 hi/lo at twice the tensor
    // rate and half the operand bytes; ~2^-16 relative operand error, used by the recurrent projection of the TF32 mode)
    const bool bf = split == 2;
    const bool row_softmax = g_tc_row_softmax && N <= 32 && !c_blocked && split != 2;
    const bool fast = !row_softmax && (c_blocked || ((N % 32 == 0) && (a->ldc % 4 == 0) && (a->c_slab_stride % 4 == 0) && ((((uintptr_t)a->d_C) & 15) == 0)));
    if (g_tc_row_softmax && !row_softmax) return 1;
    if (bf && (!fast || (a->lda & 7) || (a->a_slab_stride & 7))) return 1;
    const int BKE = bf ? 2 * BK : BK;           // elements per k-block
    const size_t esz = bf ? 2 : 4;
    const int parts = split ? 3 : 1;
    const int kpad = split ? ((K + BKE - 1) / BKE) * BKE : ((K + 3) & ~3);
    const int ldbt = parts * kpad;
    const size_t need = (size_t)N * ldbt + 2 * (size_t)N + 64;
    if (need > g_scratch_floats) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_scratch) cudaFree(g_scratch);
        g_scratch = nullptr; g_scratch_floats = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_scratch, need * sizeof(float)));
        g_scratch_floats = need;
    }
    float *Bt = g_scratch;
    float *scale = g_scratch + (((size_t)N * ldbt + 3) & ~(size_t)3);
    float *shift = scale + N;
    dim3 tb(32, 8), tg((kpad + 31) / 32, (N + 31) / 32);
    if (bf) transpose_weights_bf16_kernel<<<tg, tb, 0, st>>>(a->d_B, reinterpret_cast<__nv_bfloat16 *>(Bt), K, N, kpad, ldbt);
    else transpose_weights_kernel<<<tg, tb, 0, st>>>(a->d_B, Bt, K, N, kpad, ldbt, parts);
    // 3xTF32: the operands are exact to ~2^-22, what is left is the tensor core's TRUNCATING fp32 accumulate -- a systematic
    // shrink of every output, linear in the number of accumulation steps (measured on conv shapes K = 192 .. 1280 against the
    // reference: least-squares slope -(2.05..2.16)e-9 * K, residual rel-L2 <= 1.6e-6 once removed).  Compensated here.
    const float comp = split == 1 ? 1.0f + X3_ACC_TRUNC_PER_K * (float)K : (split ? 1.0f : 1.0f + TF32_TRUNC_BIAS);
    fold_epilogue_kernel<<<(N + 127) / 128, 128, 0, st>>>(a->d_bias, a->d_bn, a->bn_eps, N, scale, shift, comp);
    NNTK_LAUNCH_CHECK_N("gemm_tf32 prologue", 2);

    // slab stride of A as the tensor map sees it (elements); in split mode the lo copy sits lo_off elements behind the hi copy
    const long long slab_floats = a->batch > 1 ? a->a_slab_stride : (long long)a->rows * a->lda;
    const void *A = a->d_A;
    long long lo_off = 8;
    if (split) {
        const long long extent = (long long)(a->batch - 1) * slab_floats + (long long)(a->rows - 1) * a->lda + K;
        lo_off = (extent + 63) & ~63LL;
        const size_t want = ((size_t)(lo_off + extent + 64) * esz + 3) / 4;
        if (want > g_split_floats) {
            NNTK_CUDA_TRY(cudaStreamSynchronize(st));
            if (g_split) cudaFree(g_split);
            g_split = nullptr; g_split_floats = 0;
            NNTK_CUDA_TRY(cudaMalloc(&g_split, want * sizeof(float)));
            g_split_floats = want;
        }
        long long blocks = (extent / 4 + 255) / 256;
        if (blocks > 148 * 16) blocks = 148 * 16;
        if (blocks < 1) blocks = 1;
        if (bf) {
            __nv_bfloat16 *hi = reinterpret_cast<__nv_bfloat16 *>(g_split);
            split_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(a->d_A, hi, hi + lo_off, extent, g_tc_gate, g_tc_gate_run_if);
        } else {
            split_tf32_kernel<<<(unsigned)blocks, 256, 0, st>>>(a->d_A, g_split, g_split + lo_off, extent);
        }
        NNTK_LAUNCH_CHECK("split kernel");
        A = g_split;
    }

    // narrow outputs (the 12-wide TDD of C5): a 32-column tile keeps the B stage at 4 KB instead of 16 KB -- with BN = 128 the
    // zero-padded weight tile was as many bytes per stage as the activations and halved the HBM-bound kernel's input rate
    const int BN = N > 128 ? 256 : (N > 32 || bf) ? 128 : 32;
    const bool cg2 = want_cg2(BN, fast, a->rows, a->batch, (N + BN - 1) / BN);
    const CUtensorMapDataType dt = bf ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    CUtensorMap ma, mb;
    {
        // {K, rows, slabs, parts}: parts = hi / lo copies of A in split mode (lo_off elements apart)
        cuuint64_t dims[4] = {(cuuint64_t)K, (cuuint64_t)a->rows, (cuuint64_t)a->batch, (cuuint64_t)(split ? 2 : 1)};
        cuuint64_t slab_bytes = (cuuint64_t)slab_floats * esz;
        if (slab_bytes == 0) slab_bytes = 16;
        cuuint64_t strides[3] = {(cuuint64_t)a->lda * esz, slab_bytes, (cuuint64_t)lo_off * esz};
        cuuint32_t box[4] = {(cuuint32_t)BKE, BM, 1, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        CUresult r = encode(&ma, dt, 4, const_cast<void *>(A), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "gemm_tf32: tensor map for A rejected (CUresult %d): K %d rows %d batch %d lda %d", (int)r, K, a->rows, a->batch, a->lda);
            return NNTK_ERR_CUDA_;
        }
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)ldbt, (cuuint64_t)N};
        cuuint64_t strides[1] = {(cuuint64_t)ldbt * esz};
        cuuint32_t box[2] = {(cuuint32_t)BKE, (cuuint32_t)(cg2 ? BN / 2 : BN)};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode(&mb, dt, 2, (void *)Bt, dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "gemm_tf32: tensor map for B rejected (CUresult %d)", (int)r);
            return NNTK_ERR_CUDA_;
        }
    }
    TcParams p;
    p.N = N;
    p.n_tiles = (N + BN - 1) / BN;
    p.m_tiles_per_slab = (a->rows + BM - 1) / BM;
    p.kb_pass = (K + BKE - 1) / BKE;
    p.k_blocks = parts * p.kb_pass;
    p.c_blocked = c_blocked; p.row_blocks = (a->rows + 31) / 32;
    p.total_tiles = (long long)p.n_tiles * p.m_tiles_per_slab * a->batch;
    p.rows = a->rows;
    p.C = a->d_C; p.c_slab = a->c_slab_stride; p.ldc = a->ldc;
    p.scale = scale; p.shift = shift; p.act = a->act; p.act_a = a->act_a;
    p.half_fmt = 1;
    if (row_softmax) p.act = NNTK_ACT_SOFTMAX;
    if (cg2) {
        p.m_tiles_per_slab = (a->rows + 2 * BM - 1) / (2 * BM);
        p.total_tiles = (long long)p.n_tiles * p.m_tiles_per_slab * a->batch;
        return bf ? launch_tc_cg2<true>(ma, mb, p, st) : launch_tc_cg2<false>(ma, mb, p, st);
    }
    // FAST epilogue: whole 32-column chunks and 16-byte aligned output rows (every BASELINE shape except the 12-wide TDD)
    if (bf) return BN == 256 ? launch_tc<256, 4, true, true>(ma, mb, p, st) : launch_tc<128, 6, true, true>(ma, mb, p, st);
    if (BN == 32) return fast ? launch_tc<32, 8, true>(ma, mb, p, st) : launch_tc<32, 8, false>(ma, mb, p, st);
    if (BN == 256) return fast ? launch_tc<256, 4, true>(ma, mb, p, st) : launch_tc<256, 4, false>(ma, mb, p, st);
    return fast ? launch_tc<128, 6, true>(ma, mb, p, st) : launch_tc<128, 6, false>(ma, mb, p, st);
}

// Gate for the NEXT tensor-core GEMM launches (until cleared with a null flag): the kernels return at once unless *flag == run_if.
// Set and cleared under the API lock by the fp16 front end (host/conv_1d.c) around its fp32 fallback.
extern "C" void nntk_cuda_gemm_tc_row_softmax(int on) { g_tc_row_softmax = on; }
extern "C" void nntk_cuda_gemm_tc_gate(const int *d_flag, int run_if) { g_tc_gate = d_flag; g_tc_gate_run_if = run_if; }

// Single-pass fp16 GEMM of the tensor-core front end: A = fp16 rows produced by nntk_cuda_spectrogram_half (a->d_A points to
// halves; lda and a_slab_stride count halves), B = fp32 [K][N] repacked to fp16 here, fp32 accumulate, the usual folded epilogue.
// fp16 and TF32 operands carry the same 10 explicit mantissa bits, at twice the tensor rate and half the operand bytes; what
// fp16 lacks is range, so *d_ovf is set when a weight does not fit and the kernel itself only runs while *d_ovf == 0 (the
// spectrogram kernel sets the same flag) -- the caller enqueues its TF32 path behind it, gated the other way.
// returns 0 done, 1 not eligible, < 0 error.
static int gemm_f16_ex(const nntk_gemm_args *a, int *d_ovf, const int *d_gate, int gate_run_if, int c_blocked);
static float *g_f16_pre_launch_stats = nullptr;      // adaptive projection (below): its decide kernel runs between the weight repack and the GEMM
static float g_f16_pre_launch_thr = 0.0f;
static void proj_decide_launch(float *stats, float thr, cudaStream_t st);
extern "C" int nntk_cuda_gemm_f16(const nntk_gemm_args *a, int *d_ovf) { return gemm_f16_ex(a, d_ovf, d_ovf, 0, 0); }
// d_ovf: raised when a weight does not fit fp16; (d_gate, gate_run_if): the GEMM kernel runs iff *d_gate == gate_run_if;
// c_blocked: the recurrent projection's output layout (see nntk_cuda_gemm_tf32_ex)
static int gemm_f16_ex(const nntk_gemm_args *a, int *d_ovf, const int *d_gate, int gate_run_if, int c_blocked) {
    if (!a || !d_ovf || (a->lda & 7) || (a->a_slab_stride & 7) || (((uintptr_t)a->d_A) & 15) || a->rows < 1 || a->batch < 1) return 1;
    const int K = a->K, N = a->N;
    if ((N % 32) || (!c_blocked && ((a->ldc % 4) || (a->c_slab_stride % 4))) || (((uintptr_t)a->d_C) & 15) || N <= 32) return 1;
    EncodeTiledFn encode = get_encode();
    if (!encode) return 1;
    cudaStream_t st = nntk_stream();
    const int BKE = 2 * BK;
    const int kpad = ((K + BKE - 1) / BKE) * BKE;
    const size_t need = ((size_t)N * kpad + 1) / 2 + 2 * (size_t)N + 64;
    if (need > g_scratch_floats) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_scratch) cudaFree(g_scratch);
        g_scratch = nullptr; g_scratch_floats = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_scratch, need * sizeof(float)));
        g_scratch_floats = need;
    }
    __half *Bt = reinterpret_cast<__half *>(g_scratch);
    float *scale = g_scratch + ((((size_t)N * kpad + 1) / 2 + 3) & ~(size_t)3);
    float *shift = scale + N;
    dim3 tb(32, 8), tg((kpad + 31) / 32, (N + 31) / 32);
    transpose_weights_f16_kernel<<<tg, tb, 0, st>>>(a->d_B, Bt, K, N, kpad, d_ovf);
    if (g_f16_pre_launch_stats) proj_decide_launch(g_f16_pre_launch_stats, g_f16_pre_launch_thr, st);      // adaptive projection: flag from the stats
    fold_epilogue_kernel<<<(N + 127) / 128, 128, 0, st>>>(a->d_bias, a->d_bn, a->bn_eps, N, scale, shift, 1.0f);
    NNTK_LAUNCH_CHECK_N("gemm_f16 prologue", 2);
    const int BN = N > 128 ? 256 : 128;
    const long long slab = a->batch > 1 ? a->a_slab_stride : (long long)a->rows * a->lda;
    // Every tile streams the whole weight matrix (N <= BN: one n-tile).  Measured on the C5 conv (ncu, 8 192 clips): 148 CTAs each
    // pulling their own copy: 5.65 ms, 87.6 GB through L2, tensor pipe 69.5 % active; clusters of 2 sharing each B stage by
    // multicast: 5.31 ms, 65.7 GB, 73 %; clusters of 4: 8.7 ms (53 GB, 37 %: four CTAs in lockstep on a 4-stage ring stall each
    // other).  What is left is shared-memory bandwidth: per 128-cycle MMA 12 KB of operand reads + 12 KB of TMA writes.
    static int want_cs = -1;
    if (want_cs < 0) { const char *e = getenv("NNTK_GEMM_CLUSTER"); want_cs = e ? atoi(e) : 2; if (want_cs != 1 && want_cs != 2 && want_cs != 4) want_cs = 2; }
    const long long tiles_total = (long long)((a->rows + BM - 1) / BM) * a->batch;
    int cs = (N <= BN && tiles_total >= 2LL * nntk_cuda_sm_count()) ? want_cs : 1;
    while (cs > 1 && ((BN / cs) % 8 || nntk_cuda_sm_count() % cs)) cs >>= 1;
    const bool cg2 = want_cg2(BN, true, a->rows, a->batch, (N + BN - 1) / BN);
    if (cg2) cs = 2;          // the B map's box is half a tile either way
    CUtensorMap ma, mb;
    {
        cuuint64_t dims[4] = {(cuuint64_t)K, (cuuint64_t)a->rows, (cuuint64_t)a->batch, 1};
        cuuint64_t slab_bytes = (cuuint64_t)slab * 2;
        if (slab_bytes == 0) slab_bytes = 16;
        cuuint64_t strides[3] = {(cuuint64_t)a->lda * 2, slab_bytes, 16};
        cuuint32_t box[4] = {(cuuint32_t)BKE, BM, 1, 1};
        cuuint32_t estr[4] = {1, 1, 1, 1};
        CUresult r = encode(&ma, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<float *>(a->d_A), dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "gemm_f16: tensor map for A rejected (CUresult %d): K %d rows %d batch %d lda %d", (int)r, K, a->rows, a->batch, a->lda);
            return NNTK_ERR_CUDA_;
        }
    }
    {
        cuuint64_t dims[2] = {(cuuint64_t)kpad, (cuuint64_t)N};
        cuuint64_t strides[1] = {(cuuint64_t)kpad * 2};
        cuuint32_t box[2] = {(cuuint32_t)BKE, (cuuint32_t)(BN / cs)};
        cuuint32_t estr[2] = {1, 1};
        CUresult r = encode(&mb, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, (void *)Bt, dims, strides, box, estr,
                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "gemm_f16: tensor map for B rejected (CUresult %d)", (int)r);
            return NNTK_ERR_CUDA_;
        }
    }
    TcParams p;
    p.N = N;
    p.n_tiles = (N + BN - 1) / BN;
    p.m_tiles_per_slab = (a->rows + BM - 1) / BM;
    p.kb_pass = (K + BKE - 1) / BKE;
    p.k_blocks = p.kb_pass;
    p.c_blocked = c_blocked; p.row_blocks = (a->rows + 31) / 32;
    p.total_tiles = (long long)p.n_tiles * p.m_tiles_per_slab * a->batch;
    p.rows = a->rows;
    p.C = a->d_C; p.c_slab = a->c_slab_stride; p.ldc = a->ldc;
    p.scale = scale; p.shift = shift; p.act = a->act; p.act_a = a->act_a;
    p.half_fmt = 0;
    // runs only while nothing overflowed; an outer gate (there is none in the front end at this point) would be overridden
    const int *saved = g_tc_gate; const int saved_if = g_tc_gate_run_if;
    g_tc_gate = d_gate; g_tc_gate_run_if = gate_run_if;
    g_tc_cluster = cs;
    if (cg2) {
        p.m_tiles_per_slab = (a->rows + 2 * BM - 1) / (2 * BM);
        p.total_tiles = (long long)p.n_tiles * p.m_tiles_per_slab * a->batch;
    }
    const int rc = cg2 ? launch_tc_cg2<true>(ma, mb, p, st) : BN == 256 ? launch_tc<256, 4, true, true>(ma, mb, p, st) : launch_tc<128, 6, true, true>(ma, mb, p, st);
    g_tc_gate = saved; g_tc_gate_run_if = saved_if;
    return rc;
}

// ---- the hoisted recurrent projection with adaptive operand passes ---------------------------------------------------------
// xw = x W + b for all timesteps is 3xbf16 (three tensor passes + an operand split pass) because single-pass operand rounding costs
// ~1.2e-3 after a few hundred steps on N(0,1) inputs.  That error is linear in max |x_t| max |W_n| (rf_xpasses_note in
// recurrent_fused.cu): when the product is <= thr the SAME single fp16 pass the fused kernel takes is enough -- e.g. the second
// layer of a stack, whose input is the bounded h of the first.  All on the device, nothing synchronises:
//   sampled row norms (every 16th row) + column norms of W -> if the sample already fails, everything fp16 below is skipped;
//   fp32 -> fp16 conversion of x with the FULL row norms (and the fp16 range check) on the way -> decide kernel -> flag;
//   single-pass fp16 GEMM gated on flag == 1;  operand split + 3xbf16 GEMM gated on flag == 0.
namespace {
__device__ float g_proj_stats[8];     // [0] sampled max |x_t|^2, [1] max |W_n|^2, [2] full max |x_t|^2 (Inf: out of fp16 range), [3] int flag, [4] int W overflow
__global__ void __launch_bounds__(256) proj_rownorm_kernel(const float *__restrict__ x, long long rows, int K, int stride, float *stat) {
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    float best = 0.0f;
    for (long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r * stride < rows; r += warps) {
        const float4 *p = reinterpret_cast<const float4 *>(x + r * stride * K);
        float s = 0.0f;
        for (int i = lane; i < (K >> 2); i += 32) { const float4 v = __ldg(p + i); s = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, s)))); }
#pragma unroll
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        best = fmaxf(best, s);
    }
    if (lane == 0) atomicMax(reinterpret_cast<int *>(stat), __float_as_int(best));
}
__global__ void proj_colnorm_kernel(const float *__restrict__ W, int K, int N, float *stat) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    float s = 0.0f;
    if (n < N) for (int k = 0; k < K; ++k) { const float w = __ldg(W + (long long)k * N + n); s = fmaf(w, w, s); }
#pragma unroll
    for (int o = 16; o; o >>= 1) s = fmaxf(s, __shfl_xor_sync(0xffffffffu, s, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int *>(stat), __float_as_int(s));
}
// y = fp16(x) row by row, the true max |x_t|^2 on the way; runs only while the sampled bound holds
__global__ void __launch_bounds__(256) proj_cvt_f16_kernel(const float *__restrict__ x, __half *__restrict__ y, long long rows, int K, float thr, float *stats) {
    if (!(sqrtf(stats[0] * stats[1]) <= thr)) return;
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    float best = 0.0f;
    for (long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows; r += warps) {
        const float4 *p = reinterpret_cast<const float4 *>(x + r * K);
        uint2 *q = reinterpret_cast<uint2 *>(y + r * K);
        float s = 0.0f, big = 0.0f;
        for (int i = lane; i < (K >> 2); i += 32) {
            const float4 v = __ldcs(p + i);
            s = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, s))));
            big = fmaxf(big, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
            const __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
            q[i] = make_uint2(*reinterpret_cast<const uint32_t *>(&a), *reinterpret_cast<const uint32_t *>(&b));
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) { s += __shfl_xor_sync(0xffffffffu, s, o); big = fmaxf(big, __shfl_xor_sync(0xffffffffu, big, o)); }
        if (!(big <= 65504.0f)) s = __int_as_float(0x7f800000);
        best = fmaxf(best, s);
    }
    if (lane == 0) atomicMax(reinterpret_cast<int *>(stats + 2), __float_as_int(best));
}
__global__ void proj_decide_kernel(float *stats, float thr) {
    const bool ok = sqrtf(stats[0] * stats[1]) <= thr && sqrtf(stats[2] * stats[1]) <= thr && reinterpret_cast<int *>(stats)[4] == 0;
    reinterpret_cast<int *>(stats)[3] = ok ? 1 : 0;
}
}  // namespace

// 0 done (xw written, blocked layout), 1 not eligible (nothing enqueued), < 0 error
extern "C" int nntk_cuda_gemm_proj_adaptive(const nntk_gemm_args *a, float thr) {
    static int on = -1;
    if (on < 0) { const char *e = getenv("NNTK_PROJ_ADAPTIVE"); on = e ? atoi(e) : 1; }
    const int K = a->K, N = a->N;
    // the blocked view (slab = timestep, row = sequence) of a CONTIGUOUS [rows_total, K] input
    if (!on || (K & 7) || a->a_slab_stride != K || (long long)a->lda != (long long)a->batch * K || (((uintptr_t)a->d_A) & 15) || (N % 32) || N <= 32) return 1;
    static float *stats_of_device[64] = {nullptr};
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 1;
    if (!stats_of_device[dev]) {
        float *ptr = nullptr;
        if (cudaGetSymbolAddress((void **)&ptr, g_proj_stats) != cudaSuccess) { (void)cudaGetLastError(); return 1; }
        stats_of_device[dev] = ptr;
    }
    float *stats = stats_of_device[dev];
    cudaStream_t st = nntk_stream();
    const long long rows = (long long)a->batch * a->rows;
    const size_t want = ((size_t)rows * K * 2 + 3) / 4 + 64;
    if (want > g_split_floats) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_split) cudaFree(g_split);
        g_split = nullptr; g_split_floats = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_split, want * sizeof(float)));
        g_split_floats = want;
    }
    NNTK_CUDA_TRY(cudaMemsetAsync(stats, 0, 8 * sizeof(float), st));
    proj_rownorm_kernel<<<148 * 2, 256, 0, st>>>(a->d_A, rows, K, 16, stats);
    proj_colnorm_kernel<<<(N + 255) / 256, 256, 0, st>>>(a->d_B, K, N, stats + 1);
    proj_cvt_f16_kernel<<<148 * 8, 256, 0, st>>>(a->d_A, reinterpret_cast<__half *>(g_split), rows, K, thr, stats);
    NNTK_LAUNCH_CHECK_N("proj_rownorm_kernel", 3);
    // the fp16 GEMM's weight repack raises stats[4]; the decide kernel must see it: repack first (inside gemm_f16_ex), so the GEMM is
    // gated on a flag the decide kernel writes BETWEEN the repack and the GEMM -- enqueue the decide kernel from a hook
    nntk_gemm_args h = *a;
    h.d_A = reinterpret_cast<const float *>(g_split);
    g_f16_pre_launch_stats = stats; g_f16_pre_launch_thr = thr;
    int rc = gemm_f16_ex(&h, reinterpret_cast<int *>(stats) + 4, reinterpret_cast<const int *>(stats) + 3, 1, 1);
    g_f16_pre_launch_stats = nullptr;
    if (rc) return rc;
    // everything else when the flag is 0
    const int *saved = g_tc_gate; const int saved_if = g_tc_gate_run_if;
    g_tc_gate = reinterpret_cast<const int *>(stats) + 3; g_tc_gate_run_if = 0;
    rc = nntk_cuda_gemm_tf32_ex(a, 2, 1);
    g_tc_gate = saved; g_tc_gate_run_if = saved_if;
    if (rc == 1) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "proj_adaptive: the 3xbf16 GEMM is not eligible for a shape the fp16 GEMM took");
        return NNTK_ERR_CUDA_;
    }
    return rc;
}

static void proj_decide_launch(float *stats, float thr, cudaStream_t st) {
    proj_decide_kernel<<<1, 1, 0, st>>>(stats, thr);
    nntk_cuda_count_launch(1);
}

// operand passes the most recent adaptive projection took (1: single fp16 pass, 3: 3xbf16; 0: none ran).  Synchronises: tests only.
extern "C" int nntk_cuda_last_proj_passes(void) {
    float *ptr = nullptr;
    if (cudaGetSymbolAddress((void **)&ptr, g_proj_stats) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
    float v[8] = {0};
    if (cudaStreamSynchronize(nntk_stream()) != cudaSuccess || cudaMemcpy(v, ptr, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    if (v[1] == 0.0f) return 0;
    return reinterpret_cast<int *>(v)[3] == 1 ? 1 : 3;
}
