// recurrent_fused.cu -- K3 + K4 in ONE persistent kernel per GRU / LSTM layer (tensor-core mode, default gate activations).
// Replaces, for the shapes it covers, the hoisted-projection GEMM + persistent cluster kernel / per-step kernels of the
// NNTK_MATH_TF32 mode (reference loops: layers/gru.c:129-204, layers/lstm.c:185-268).
//
// Work decomposition.  A CTA PAIR (2-CTA cluster, tcgen05.mma.cta_group::2, M = 128) owns 128 sequences -- 64 per CTA -- for
// ALL T timesteps.  h never leaves the pair: there is no exchange between CTAs beyond the operand sharing of the 2-CTA MMA.
//   * B operands (W for the input part, U for the recurrent part; fp16, repacked once per call so that a tile holds all gates
//     of 64 hidden units) are STREAMED from L2 through a TMA ring, each CTA loading the rows of its 32 units of the tile; the
//     pair's MMA reads both halves.  Nothing is resident: H = 512 works like H = 256.
//   * A operands live in shared memory for the whole kernel: the fp16 image of h_{t-1} [64 rows x H] (rewritten by the
//     epilogue once per step, after the step's last MMA) and -- FUSED mode -- the fp16 image of x_t [64 rows x in], produced by
//     converter warps straight from the caller's fp32 rows: x_{t+1} waits in their REGISTERS while step t runs (the global
//     loads are a whole step old when the tile frees up), so there is no staging buffer and no projection buffer at all:
//     per layer HBM sees x once and h once -- the algorithmic traffic (r1: ~4.5x that).
//     Input passes (RfParams::xp): 3 = x and W split into fp16 hi + lo, products hi*Wlo, lo*Whi, hi*Whi (x . W exact to ~2^-22,
//     like the 3xbf16 projection it replaces; the default), 1 = hi*Whi only (fp16-rounded x and W: measured 1.2e-3 max error on
//     N(0,1) inputs at in = 256, T = 300 -- too coarse for the mode's 1e-3 in general, fine for bounded activations).
//     XW mode (H = 512 / in > 256, where the x tile does not fit) instead reads the hoisted projection in the blocked layout of
//     gemm_tc.cu.
//   * Accumulators in TMEM: one 128-column buffer per 64-unit tile (cta_group::2 layout: lanes 0-63 hold the columns computed
//     from CTA 0's B rows = units 0-31 of the tile, lanes 64-127 units 32-63; scripts/probe_cg2.cu pins this).  GRU columns
//     [x.Wh | z | r | h.Uh]: the input part writes the first three groups (N = 192), the recurrent part accumulates onto the last
//     three (N = 192; the h.Uh group is re-zeroed by the epilogue), so r * (h.Uh + b_hh) + x.Wh keeps its two terms apart
//     (gru.c:169-171).  LSTM columns [i | f | g | o], both parts N = 256.  N >= 192 matters: a pair MMA costs max(57, N/4)
//     cycles (probe), so 32-unit tiles would run at 42 %.
//   * The fp32 recurrent state (h for GRU, c for LSTM) also lives in TMEM (H/2 columns, lane = row x unit half): it is read
//     and written only by the thread that owns (row, units); 64 x H floats would not fit the register file.
//   * Input-part MMAs of the NEXT tile are issued one tile ahead of the recurrent part, and x_{t+1} . W does not depend on h:
//     across the step boundary the tensor pipe works on it while the last tile's cell math finishes.
// Warp roles: 0-7 epilogue (TMEM lane quarter = warp % 4; thread = one row x 16 units of the tile), 8-15 x converters (FUSED
// only), then the four control warps: ring producer (TMA), MMA issuer (leader CTA, one thread), idle, TMEM allocation.  The
// control warps sit LAST on purpose: the warp scheduler favours the highest warp id on a sub-partition, and with the issuer as
// warp 1 (first version) it shared its scheduler with four polling / computing warps and needed 130-170 cycles per MMA
// (the instruction itself: <= 57).  setmaxnreg moves registers from the control and converter warps to the epilogue warps.
// Precision: fp16 operands (round to nearest), fp32 accumulate, fp32 state, MUFU.TANH gates: the 1e-3 class of the tensor-core
// mode; tests/test_gpu_tf32.py and tests/test_gpu_config_shapes.py pin it against the reference at the config shapes.
#include <cuda_fp16.h>
#include <stdlib.h>
#include <vector>
#include "tc_common.cuh"

namespace {

constexpr int RF_ROWS = 64;                  // sequences per CTA (pair: 128)
constexpr int RF_UT = 64;                    // hidden units per tile (pair), 32 per CTA
// FUSED mode epilogue warps: 8 (16 hidden units of a tile per thread; 20 warps per CTA) or 16 (8 units per thread; 28 warps, 80
// registers per epilogue thread).  Measured on the C5 GRU layer (8 192 x 994 steps): 8 warps 7.5 ms, 16 warps 8.25 ms (the
// register diet spills ~360 bytes per thread and the converters load one k-block at a time) -- so 8 it is; -DRF_FUSED_EPW=16 keeps
// the experiment buildable.
#ifndef RF_FUSED_EPW
#define RF_FUSED_EPW 8
#endif
constexpr int RF_THREADS_FUSED = RF_FUSED_EPW == 16 ? 896 : 640, RF_THREADS_XW = 640;
constexpr int RF_MAX_RING = 8;

struct RfParams {
    int B, H, in, T;
    int rows;                        // real sequences per CTA: 64, or 32 (half-filled M tile, twice the pairs: for batches that would leave SMs idle)
    int split;                       // pairs per slice of sequences: 1, or 2 = the hidden units of a slice are split across the two pairs of a
                                     // 4-CTA cluster (each streams half of the weights, owns half of the tiles; h_t halves are exchanged through
                                     // distributed shared memory once per step)
    int return_sequences;
    int ring_stages;
    int xp;                          // FUSED: input passes the launch is PLANNED for (shared memory, packed operands), 1 or 3
    const float *stats;              // adaptive input passes: {max_rows |x_t|^2, max_cols |W_n|^2, passes taken (written by the kernel)}
    float adapt_thr;                 //   3 planned passes drop to 1 when sqrt(stats[0] * stats[1]) <= adapt_thr (see rf_xpasses_note)
    int per_step;                    // 1: the decision is taken per pair and per STEP inside the kernel (no pre-scan of x): the converter warps publish
                                     //    max |x_t|^2 of their rows with the x tile, the issuer skips the two correction passes' MMAs for that step
    int cs;                          // cluster size: 2, 4 or 8 CTAs = 1, 2 or 4 pairs sharing every B stage by TMA multicast
    int xw_blocked_rows;             // XW mode: row_blocks of the blocked projection
    int x3;                          // X3 kernels (fp32-contract mode): h carried as fp16 hi + lo A tiles, three passes for the recurrent part too
    int lookahead;                   // FUSED: tiles the input parts run ahead of the recurrent parts (1 or 2; NNTK_RECURRENT_FUSED_LOOKAHEAD)
    int verify;                      // one-pass plan launched on the strength of a SAMPLED pre-scan: the converters still publish every step's row
                                     //    norms and the issuer raises stats[3] when one breaks the bound (the three-pass launch behind then runs)
    int sel;                         // 0 always run; 1 run iff the sampled bound holds; 2 run iff it does not, or launch 1 raised stats[3]
    const float *h_init, *c_init;    // initial state read by the prologue (a copy when two launches share h_state / c_state), else = h_state / c_state
    const float *x;                  // FUSED: [B][T][in]
    const float *bias_i, *bias_h;    // [G*H] each; bias_h may be null (not added)
    const float *xw;                 // XW mode: blocked projection [T][row_blocks][G*H/4][32][4] (b_i already added)
    float *h_state, *c_state;        // [B][H] in/out
    float *out;                      // [B][T][H] or [B][H]
    long long *dbg;                  // development: per-CTA wait-cycle counters [grid][16] (NNTK_RF_DBG=1), else null
    int skip;                        // development (RF_DEBUG builds, NNTK_RF_SKIP): 1 no cell math / TMEM traffic, 2 no h_t stores,
                                     // 4 no x conversion, 8 no MMAs (results are garbage: timing experiments only)
};
// cycles spent in a wait, accumulated per role (only when p.dbg): 0 mma:ring_full 1 mma:acc_empty 2 mma:h_full 3 mma:x_full
// 4 epi:acc_full 5 epi:total 6 conv:convert 7 conv:x_empty 8 prod:ring_empty 10 mma:total 11 epi:store_wait
// Compiled in only with -DRF_DEBUG=1 (NNTK_NVCC_EXTRA): the counters cost registers the control warps do not have.
#ifndef RF_DEBUG
#define RF_DEBUG 0
#endif
#if RF_DEBUG
#define RF_T0() const long long t0__ = p.dbg ? clock64() : 0
#define RF_T1(i) do { if (p.dbg) dbgc[i] += clock64() - t0__; } while (0)
#define RF_DBG(stmt) do { if (p.dbg) { stmt; } } while (0)
#define RF_SKIP(bit) (p.skip & (bit))
#else
#define RF_SKIP(bit) false
#define RF_T0() do { } while (0)
#define RF_T1(i) do { } while (0)
#define RF_DBG(stmt) do { } while (0)
#endif

__device__ __forceinline__ uint32_t rf_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void rf_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t rf_mapa(uint32_t a, uint32_t rank) {
    uint32_t d; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(rank)); return d;
}
// Arrive on the barrier at cluster address `a` (the leader CTA's copy).  Default (CTA-scope) semantics on both sides, like
// CUTLASS's ClusterBarrier: what crosses the pair is either TMEM traffic (ordered by tcgen05.fence + tcgen05.wait) or shared
// memory read by the async proxy (ordered by fence.proxy.async before the arrive).  Cluster-scope release / acquire cost an
// ERRBAR resp. a CCTL.IVALL per signal (ncu, first version: membar 3.1 of 16 stall cycles per issue).
__device__ __forceinline__ void rf_arrive_cl(uint32_t a) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(a) : "memory");
}
// cluster-scope acquire / release pair for GENERIC data that crosses the pair (the converters' row norms): only the converter warps
// and one issuer wait per step pay for it
__device__ __forceinline__ void rf_mbar_wait_cl(uint64_t *bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWLC:\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t@p bra WDC;\n\tbra WLC;\n\tWDC:\n\t}"
                 ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void rf_arrive_cl_release(uint32_t a) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(a) : "memory");
}
__device__ __forceinline__ bool rf_elect() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// shared-memory accesses by 32-bit shared address: through generic pointers the compiler emitted LD.E / ST.E (address-space
// resolution per access) in the epilogue's dependent chain
__device__ __forceinline__ float4 rf_lds128(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void rf_sts128(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ void rf_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// TMA 2-D load, multicast to the CTAs of `mask` (same shared-memory offset in each); with .cta_group::2 the bytes are counted on
// the barrier at offset `bar` (peer bit cleared) in the PAIR LEADER of every destination CTA -- CUTLASS's SM100_TMA_2SM_LOAD_MULTICAST
__device__ __forceinline__ void rf_tma_2d_cg2_mc(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, uint16_t mask) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
                 ::"r"(dst), "l"(map), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1), "h"(mask) : "memory");
}
__device__ __forceinline__ void rf_tma_store_3d(const CUtensorMap *map, const void *src, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
                 ::"l"(map), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void rf_mma(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// commit arriving on the barrier at this offset in every CTA of `mask` once all MMAs issued so far have completed
__device__ __forceinline__ void rf_commit(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ void rf_ld8(uint32_t taddr, float (&r)[8]) {
    uint32_t u[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]) : "r"(taddr));
#pragma unroll
    for (int i = 0; i < 8; ++i) r[i] = __uint_as_float(u[i]);
}
__device__ __forceinline__ void rf_st8(uint32_t taddr, const float (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                 ::"r"(taddr), "r"(__float_as_uint(r[0])), "r"(__float_as_uint(r[1])), "r"(__float_as_uint(r[2])), "r"(__float_as_uint(r[3])),
                   "r"(__float_as_uint(r[4])), "r"(__float_as_uint(r[5])), "r"(__float_as_uint(r[6])), "r"(__float_as_uint(r[7])) : "memory");
}
__device__ __forceinline__ void rf_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void rf_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t rf_pack_sat(float a, float b) {      // two floats -> packed fp16, round to nearest, finite
    uint32_t r;
    asm("{\n\t.reg .f16 lo, hi;\n\tcvt.rn.satfinite.f16.f32 lo, %1;\n\tcvt.rn.satfinite.f16.f32 hi, %2;\n\tmov.b32 %0, {lo, hi};\n\t}"
        : "=r"(r) : "f"(a), "f"(b));
    return r;
}
// hi = fp16(v) pairs; lo = fp16(v - hi) pairs (v - hi is exact in fp32)
__device__ __forceinline__ void rf_split2(float a, float b, uint32_t &hi, uint32_t &lo) {
    hi = rf_pack_sat(a, b);
    const __half2 h = *reinterpret_cast<const __half2 *>(&hi);
    const float2 f = __half22float2(h);
    lo = rf_pack_sat(a - f.x, b - f.y);
}

// shared-memory plan (host and device agree through this one struct)
template <int G, bool FUSED>
struct RfPlan {
    int a_h, a_l, x_hi, x_lo, out_stg, bias, ring, bars, total;
    int sub_bytes, stage_bytes;
    __host__ __device__ RfPlan(int H, int in, int xp, int ring_stages, bool dbx = false, bool x3 = false) {
        sub_bytes = 32 * G * 128;                            // one k-block of B: [32 G rows of this CTA][64 halves]
        stage_bytes = 2 * sub_bytes;                         // a ring stage holds two k-blocks
        int o = 0;
        a_h = o; o += RF_ROWS * H * 2;                       // [H/64 k-blocks][64 rows][128 B]
        a_l = o; o += x3 ? RF_ROWS * H * 2 : 0;              // X3 kernels: the fp16 image of h - fp16(h)
        x_hi = o; o += FUSED ? RF_ROWS * in * 2 * (dbx ? 2 : 1) : 0;   // [in/64][64 rows][128 B]; DBX kernels: two of them (steps t and t + 1)
        x_lo = o; o += (FUSED && xp == 3) ? RF_ROWS * in * 2 : 0;
        out_stg = o; o += 8 * 2048;                          // per epilogue warp: [32 rows][16 floats] (SWIZZLE_64B) or 16 x [32 rows][8 floats]
        ring = o; o += ring_stages * stage_bytes;
        bias = o; o += 4 * H * 4;
        bars = o; o += 1024;
        total = o + 1024;                                    // + alignment slack
    }
};

// barrier block layout (uint64 slots)
constexpr int RFB_RING_FULL = 0, RFB_RING_EMPTY = RF_MAX_RING, RFB_ACC_FULL = 2 * RF_MAX_RING, RFB_ACC_EMPTY = RFB_ACC_FULL + 4,
              RFB_H_FULL = RFB_ACC_EMPTY + 4, RFB_X_FULL = RFB_H_FULL + 1 /* hi image */, RFB_XL_FULL = RFB_X_FULL + 1 /* lo image, per k-block */,
              RFB_X_EMPTY = RFB_XL_FULL + 4, RFB_A_FREE = RFB_X_EMPTY + 1 /* split mode: both pairs' MMAs of the step have read A_h */,
              RFB_XF2 = RFB_A_FREE + 1 /* DBX kernels: x tile of parity b full */, RFB_XE2 = RFB_XF2 + 2 /* ... free again */,
              RFB_END = RFB_XE2 + 2;

// NCMAX: compile-time bound of the tiles per step (H / 64 <= NCMAX): sizes the register-resident pending-h array.
// DBX (FUSED, one-pass plan only): 16 epilogue warps (8 units of a tile per thread, 96 registers) + 8 slim converter warps that write
// the fp16 image of x_t straight into one of TWO x tiles (no register-resident copy: 40 registers) + 4 control warps = 896 threads.
// X3 (GRU, FUSED): the fp32-contract variant.  Every product is split-fp16 in three passes -- the input part as before, and the
// recurrent part too: h_t is carried as TWO A tiles, fp16(h) and fp16(h - fp16(h)), rebuilt at the end of every step from the exact
// fp32 state in TMEM, against U_lo and U_hi -- and the gates use expf / tanhf instead of MUFU.TANH.  Operands are exact to ~2^-22,
// accumulation and state are fp32: the 1e-5 contract of NNTK_MATH_TF32X3, in ONE launch per layer instead of one per step.
template <int G, bool FUSED, int NCMAX, bool DBX = false, bool X3 = false>
__global__ void __launch_bounds__(FUSED ? (DBX ? 896 : RF_THREADS_FUSED) : RF_THREADS_XW, 1)
recurrent_fused_kernel(const __grid_constant__ CUtensorMap map_w, const __grid_constant__ CUtensorMap map_wl,
                       const __grid_constant__ CUtensorMap map_u, const __grid_constant__ CUtensorMap map_ul,
                       const __grid_constant__ CUtensorMap map_out, RfParams p) {
    constexpr int NW = 64 * G;                              // N of one MMA (LSTM 256, GRU 192)
    constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(NW >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);   // f32 <- f16 x f16, K-major
    constexpr int ROWS_B = 32 * G;                          // B rows per CTA per k-block
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    // Input passes actually taken.  The two correction passes (x_hi . W_lo, x_lo . W_hi) exist because fp16-rounded x and W put
    // ~2^-12 * sqrt(2/3) * |x_t| * |W_n| / ... of noise into every pre-activation of every step; when a pre-scan of this call's x
    // and W bounds |x_t| |W_n| by adapt_thr that noise stays ~4x under the mode's 1e-3 and the passes are skipped (rf_xpasses_note).
    // Every role of every CTA evaluates the same two numbers, so the whole grid takes the same branch.
    if (p.sel) {         // speculative pair of launches (see nntk_cuda_recurrent_fused): the whole grid takes the same branch, before any state exists
        const bool ok = sqrtf(__ldg(p.stats) * __ldg(p.stats + 1)) <= p.adapt_thr;
        const int viol = *reinterpret_cast<const volatile int *>(p.stats + 3);
        if (p.sel == 1 ? !ok : (ok && !viol)) return;
    }
    const bool PS = FUSED && p.per_step && p.xp == 3 && p.stats;      // per-step decision: everything is staged for three passes
    const bool PUB = FUSED && (PS || (p.verify && p.stats));          // the converters publish max |x_t|^2 per step
    const int XP = PS ? 3 : ((p.xp == 3 && p.stats && sqrtf(__ldg(p.stats) * __ldg(p.stats + 1)) <= p.adapt_thr) ? 1 : p.xp);
    const int H = p.H, IN = p.in, T = p.T;
    const int S = p.split;                                  // pairs per slice (1 or 2)
    const int NCT = H / RF_UT, NC = NCT / S;                // tiles per step: of the layer / of this pair
    const int KBH = H / 64, KBX = FUSED ? IN / 64 : 0;      // k-blocks of the recurrent / input part
    const int NB = (512 - H / S / 2) / 128;                 // accumulator buffers (3 when the pair owns <= 256 units, 2 for 512)
    const uint32_t STATE_COL = (uint32_t)(NB * 128);
    const int LA = (p.lookahead >= 2 && NB >= 3) ? 2 : 1;   // input parts issued LA tiles ahead of the recurrent parts
    const RfPlan<G, FUSED> pl(H, IN, p.xp, p.ring_stages, DBX, X3);
    uint8_t *sAl = smem + pl.a_l;
    uint8_t *sAh = smem + pl.a_h, *sXh = smem + pl.x_hi, *sXl = smem + pl.x_lo, *sOut = smem + pl.out_stg, *sRing = smem + pl.ring;
    float *sBias = reinterpret_cast<float *>(smem + pl.bias);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + pl.bars);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + RFB_END);
    float *xs = reinterpret_cast<float *>(smem + pl.bars + 512);     // [2][16] max |x_t|^2 per converter warp of the pair (leader CTA's copy is read)

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);          // warp-uniform for the compiler (role dispatch, uniform registers)
    // FUSED: 8 epilogue warps (16 units of a tile per thread) + 8 converter warps.  XW: 16 epilogue warps (8 units per thread) -- its
    // epilogue also adds the hoisted projection and, with 8 warps, took 4 700 cycles per LSTM-512 tile against 2 064 of MMA.
    constexpr int EPW = FUSED ? (DBX ? 16 : RF_FUSED_EPW) : 16, UPT = 128 / EPW, NE = UPT / 8;
    constexpr int CONV_W0 = FUSED ? EPW : 8, CTRL_W0 = FUSED ? EPW + 8 : 16;     // first converter / control warp
    const uint32_t crank = __shfl_sync(0xffffffffu, rf_rank(), 0);   // rank in the cluster (warp-uniform for the compiler); pairs are (0,1), (2,3), ...
    const uint32_t rank = crank & 1u, leader = crank & ~1u; // rank inside the pair, cluster rank of the pair's leader
    // unsplit: the cs / 2 pairs of a cluster work on different slices and share every B stage by multicast (npairs, pair_i).
    // split: the cluster is ONE slice, pair s = crank / 2 owns tiles [s NC, (s + 1) NC) and streams only their weights.
    const int npairs = S == 2 ? 1 : p.cs >> 1, pair_i = S == 2 ? 0 : (int)(crank >> 1);
    const int tile0 = S == 2 ? (int)(crank >> 1) * NC : 0;
    const uint16_t pair_mask = (uint16_t)(3u << leader), all_mask = S == 2 ? pair_mask : (uint16_t)((1u << p.cs) - 1u);
    const int slice = S == 2 ? (int)(blockIdx.x >> 2) : (int)(blockIdx.x >> 1);
    const int row0 = (slice * 2 + (int)rank) * p.rows;      // first sequence of this CTA
    const int EPA = p.rows == RF_ROWS ? EPW : EPW / 2;      // epilogue warps with real rows (32-row mode: lane quarters 0 and 2 only)
    // tile counters are 32-bit and advanced incrementally everywhere: 64-bit divisions in the issuer's loop cost ~13 000 cycles per
    // step in the first version (more than all of the step's MMAs)

    if (tid == 0 && blockIdx.x == 0 && p.stats && !PUB) const_cast<float *>(p.stats)[2] = (float)XP;
    if (tid == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_w) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_wl) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_u) : "memory");
        if (X3) asm volatile("prefetch.tensormap [%0];" ::"l"(&map_ul) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_out) : "memory");
        for (int s = 0; s < p.ring_stages; ++s) { mbar_init(bars + RFB_RING_FULL + s, 1); mbar_init(bars + RFB_RING_EMPTY + s, (uint32_t)npairs); }
        for (int b = 0; b < 4; ++b) { mbar_init(bars + RFB_ACC_FULL + b, 1); mbar_init(bars + RFB_ACC_EMPTY + b, 2 * EPA); }
        mbar_init(bars + RFB_H_FULL, 2 * EPA * S);
        mbar_init(bars + RFB_A_FREE, (uint32_t)S);
        mbar_init(bars + RFB_X_FULL, 16);
        for (int b = 0; b < 4; ++b) mbar_init(bars + RFB_XL_FULL + b, 16);
        mbar_init(bars + RFB_X_EMPTY, 1);
        for (int b = 0; b < 2; ++b) { mbar_init(bars + RFB_XF2 + b, 16); mbar_init(bars + RFB_XE2 + b, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == CTRL_W0 + 3) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    // bias table [4][H]: GRU {b_i z + b_h z, b_i r + b_h r, b_i h, b_h h}; LSTM {b_i g + b_h g} per gate.  XW mode: b_i is in xw.
    for (int i = tid; i < 4 * H; i += (int)blockDim.x) {
        const int a = i / H, u = i - a * H;
        const float *bi = FUSED ? p.bias_i : nullptr;
        float v;
        if (G == 3) {
            if (a < 2) v = (bi ? __ldg(bi + a * H + u) : 0.f) + (p.bias_h ? __ldg(p.bias_h + a * H + u) : 0.f);
            else if (a == 2) v = bi ? __ldg(bi + 2 * H + u) : 0.f;
            else v = p.bias_h ? __ldg(p.bias_h + 2 * H + u) : 0.f;
        } else {
            v = (bi ? __ldg(bi + a * H + u) : 0.f) + (p.bias_h ? __ldg(p.bias_h + a * H + u) : 0.f);
        }
        sBias[i] = v;
    }
    if (p.rows < RF_ROWS) {
        // 32-row mode: rows 32..63 of every A tile never receive data; give the MMA zeros instead of stale shared memory
        for (int i = tid; i < (pl.out_stg >> 4); i += (int)blockDim.x) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
        rf_fence_async();
    }
    tc_fence_before();
    __syncthreads();
    rf_cluster_sync();               // both CTAs' barriers exist before anybody arrives remotely
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;
#if RF_DEBUG
    long long dbgc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    const long long t_start = p.dbg ? clock64() : 0;
#endif

    if (warp >= CTRL_W0) {
        if (FUSED && !DBX && RF_FUSED_EPW == 8) asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
        else asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == CTRL_W0) {
            // ===== ring producer: this CTA's half of every B stage (two k-blocks), in the order the MMA issuer consumes them =====
            // Like the MMA issuer the WHOLE warp runs the loop converged on warp-uniform values and only the TMA / expect_tx
            // instructions are predicated on elect.sync: issued from one divergent lane every UTMALDG sat in an ELECT / 4 x
            // R2UR.BROADCAST / BRA loop (measured with the kernel's counters: ~180 cycles per TMA, 5 700 of a step's 16 200 cycles,
            // and the step time did not move when MMAs, cell math or stores were switched off: the ring was the floor).
            int stage = 0; uint32_t phase = 0;
            // Every B stage is identical for all pairs of the cluster: this CTA fetches 1 / npairs of the rows of its half and
            // multicasts them to the CTAs of the same half in every pair (L2 -> SM traffic / npairs: with 128 CTAs each pulling its own
            // copy the stream ran at the per-SM share of the L2 bandwidth, ~43 B/cycle, and bounded the step).  A slot is free once ALL
            // pairs' MMAs have read it (ring_empty counts one multicast commit per pair).
            const uint32_t full_l = smem_u32(bars + RFB_RING_FULL), ring_a = smem_u32(sRing);
            const int slice_rows = ROWS_B / npairs;
            uint16_t half_mask = 0;
            for (int i = 0; i < npairs; ++i) half_mask |= (uint16_t)(1u << (2 * i + (int)rank));
            if (S == 2) half_mask = (uint16_t)(1u << crank);
            auto put = [&](const CUtensorMap *map, int kb0, int nkb, int tile) {
                { RF_T0(); mbar_wait(bars + RFB_RING_EMPTY + stage, phase ^ 1); RF_T1(8); }
                RF_T0();
                if (rf_elect()) {
                    if (rank == 0) mbar_expect_tx(bars + RFB_RING_FULL + stage, 2u * (uint32_t)(nkb * pl.sub_bytes));
                    for (int i = 0; i < nkb; ++i)
                        rf_tma_2d_cg2_mc(ring_a + (uint32_t)(stage * pl.stage_bytes + i * pl.sub_bytes + pair_i * slice_rows * 128), map,
                                         full_l + 8u * (uint32_t)stage, (kb0 + i) * 64, (tile * 2 + (int)rank) * ROWS_B + pair_i * slice_rows, half_mask);
                }
                __syncwarp();
                RF_T1(9);
                if (++stage == p.ring_stages) { stage = 0; phase ^= 1; }
            };
            auto put_x = [&](int tile) {
                if (XP == 3) for (int kb = 0; kb < KBX; kb += 2) put(&map_wl, kb, min(2, KBX - kb), tile);
                for (int kb = 0; kb < KBX; kb += 2) put(&map_w, kb, min(2, KBX - kb), tile);
            };
            // issue order (the MMA issuer walks the same one): X(0) .. X(LA - 1), then per tile g: H(g), X(g + LA)
            const int total = T * NC;
            int xc = 0, xg = 0;          // tile (within the step) and linear index of the next input part
            if (FUSED) for (; xg < LA && xg < total; ++xg) { put_x(tile0 + xc); if (++xc == NC) xc = 0; }
            for (int t = 0; t < T; ++t)
                for (int c = 0; c < NC; ++c) {
                    if (X3) for (int kb = 0; kb < KBH; kb += 2) put(&map_ul, kb, min(2, KBH - kb), tile0 + c);      // h_hi . U_lo first
                    for (int kb = 0; kb < KBH; kb += 2) put(&map_u, kb, min(2, KBH - kb), tile0 + c);
                    if (FUSED && xg < total) { put_x(tile0 + xc); if (++xc == NC) xc = 0; ++xg; }
                }
            RF_DBG(if (lane == 0) { p.dbg[blockIdx.x * 16 + 8] = dbgc[8]; p.dbg[blockIdx.x * 16 + 9] = dbgc[9]; });
        } else if (warp == CTRL_W0 + 1 && rank == 0) {
            // ===== MMA issuer (leader CTA): static order  X(0), then per tile g: H(g), X(g + 1) =====
            // The WHOLE warp runs this code converged on warp-uniform values, and only the tcgen05 instructions are predicated on
            // elect.sync (the CUTLASS way): descriptors then live in uniform registers.  Issued from one divergent thread the compiler
            // wrapped every MMA in an ELECT / 5 x R2UR.BROADCAST / BRA loop -- ncu: ~80 cycles of issue overhead per 57-cycle MMA.
            int stage = 0; uint32_t phase = 0;
            const uint32_t tmem = __shfl_sync(0xffffffffu, *tmem_slot, 0);
            const uint32_t ring_a = smem_u32(sRing), ah_a = smem_u32(sAh), xh_a = smem_u32(sXh), xl_a = smem_u32(sXl);
            // one ring stage = up to two k-blocks of B; a0 (and a1 when given: the lo image first, then the hi image) are the A tiles
            auto stage_mmas = [&](uint32_t d, uint32_t a0, uint32_t a1, int kb0, int nkb, bool first_acc) {
                { RF_T0(); mbar_wait(bars + RFB_RING_FULL + stage, phase); RF_T1(0); }
                tc_fence_after();
                { RF_T0();
                if (rf_elect()) {
                for (int i = 0; i < nkb; ++i) {
                    const uint64_t db = make_desc(ring_a + (uint32_t)(stage * pl.stage_bytes + i * pl.sub_bytes));
                    if (RF_SKIP(8)) continue;
                    if (a1) {
                        const uint64_t dl = make_desc(a1 + (uint32_t)((kb0 + i) * RF_ROWS * 128));
#pragma unroll
                        for (int k = 0; k < 4; ++k) rf_mma(d, dl + 2 * k, db + 2 * k, IDESC, 1u);
                    }
                    const uint64_t da = make_desc(a0 + (uint32_t)((kb0 + i) * RF_ROWS * 128));
#pragma unroll
                    for (int k = 0; k < 4; ++k) rf_mma(d, da + 2 * k, db + 2 * k, IDESC, (first_acc && kb0 + i == 0 && k == 0) ? 0u : 1u);
                }
                rf_commit(bars + RFB_RING_EMPTY + stage, all_mask);
                }
                __syncwarp();
                RF_T1(13); }
                if (++stage == p.ring_stages) { stage = 0; phase ^= 1; }
            };
            // a ring stage whose MMAs this pair does not need (per-step decision: one pass): consume it and hand it back
            auto stage_skip = [&]() {
                { RF_T0(); mbar_wait(bars + RFB_RING_FULL + stage, phase); RF_T1(0); }
                if (rf_elect()) rf_commit(bars + RFB_RING_EMPTY + stage, all_mask);
                __syncwarp();
                if (++stage == p.ring_stages) { stage = 0; phase ^= 1; }
            };
            const float wstat = PUB ? __ldg(p.stats + 1) : 0.0f;
            int xp_t = XP, any3 = 0;     // passes of the step whose input parts are being issued
            uint32_t xh_t = xh_a;        // its x tile (DBX: alternates)
            // (t, c) = step and tile; buf = accumulator buffer of the tile, use = how often that buffer has been used before
            auto x_part = [&](int t, int c, int buf, int use) {
                if (use > 0) { RF_T0(); mbar_wait(bars + RFB_ACC_EMPTY + buf, (uint32_t)((use - 1) & 1)); RF_T1(1); }
                tc_fence_after();
                const uint32_t d = tmem + (uint32_t)(buf * 128);
                if (c == 0) {            // the hi image of x_t (and, per-step mode, the row norms the converters published with it)
                    RF_T0();
                    if (DBX) {           // x tile of parity t & 1, used for the (t >> 1)-th time
                        if (PUB) rf_mbar_wait_cl(bars + RFB_XF2 + (t & 1), (uint32_t)((t >> 1) & 1)); else mbar_wait(bars + RFB_XF2 + (t & 1), (uint32_t)((t >> 1) & 1));
                        xh_t = xh_a + (uint32_t)((t & 1) * RF_ROWS * IN * 2);
                    } else
                    if (PUB) rf_mbar_wait_cl(bars + RFB_X_FULL, (uint32_t)(t & 1)); else mbar_wait(bars + RFB_X_FULL, (uint32_t)(t & 1));
                    RF_T1(3);
                    if (PUB) {
                        float mx = 0.0f;
#pragma unroll
                        for (int i = 0; i < 16; ++i) mx = fmaxf(mx, xs[(t & 1) * 16 + i]);
                        const bool small = sqrtf(mx * wstat) <= p.adapt_thr;       // Inf stays, NaN rows were dropped by fmaxf
                        if (PS) { xp_t = small ? 1 : 3; any3 |= xp_t == 3; }
                        else if (!small && lane == 0) *reinterpret_cast<volatile int *>(const_cast<float *>(p.stats) + 3) = 1;   // one-pass plan broken
                    }
                }
                if (PS && xp_t == 1) {   // everything was staged for three passes; this step takes one
                    for (int kb = 0; kb < KBX; kb += 2) stage_skip();
                    for (int kb = 0; kb < KBX; kb += 2) stage_mmas(d, xh_a, 0u, kb, min(2, KBX - kb), true);
                    // the lo image still arrives every step: keep the barrier phases in step (late: it is long there)
                    if (c == NC - 1) for (int i = 0; i < KBX; ++i) mbar_wait(bars + RFB_XL_FULL + i, (uint32_t)(t & 1));
                } else
                if (XP == 3) {           // small terms first: x_hi . W_lo, then x_lo . W_hi + x_hi . W_hi off the same W_hi stage
                    for (int kb = 0; kb < KBX; kb += 2) stage_mmas(d, xh_a, 0u, kb, min(2, KBX - kb), true);
                    for (int kb = 0; kb < KBX; kb += 2) {
                        if (c == 0) for (int i = kb; i < min(kb + 2, KBX); ++i) { RF_T0(); mbar_wait(bars + RFB_XL_FULL + i, (uint32_t)(t & 1)); RF_T1(3); }
                        stage_mmas(d, xh_a, xl_a, kb, min(2, KBX - kb), false);
                    }
                } else {
                    for (int kb = 0; kb < KBX; kb += 2) stage_mmas(d, DBX ? xh_t : xh_a, 0u, kb, min(2, KBX - kb), true);
                }
                if (c == NC - 1) { if (rf_elect()) rf_commit(bars + (DBX ? RFB_XE2 + (t & 1) : RFB_X_EMPTY), pair_mask); __syncwarp(); }     // the x tile may be rewritten once these MMAs are done
            };
            auto h_part = [&](int t, int c, int buf, int use) {
                if (c == 0) { RF_T0(); mbar_wait(bars + RFB_H_FULL, (uint32_t)(t & 1)); RF_T1(2); }
                if (!FUSED && use > 0) { RF_T0(); mbar_wait(bars + RFB_ACC_EMPTY + buf, (uint32_t)((use - 1) & 1)); RF_T1(1); }
                tc_fence_after();
                // The recurrent part ALWAYS accumulates: onto the input part (FUSED) or onto the hoisted projection the epilogue stored
                // into the buffer (XW).  GRU: it lands one 32-column group to the right ([z | r | h.Uh]); x.Wh stays in group 0.
                const uint32_t d = tmem + (uint32_t)(buf * 128) + (G == 3 ? 32u : 0u);
                if (X3) {                // small terms first: h_hi . U_lo, then h_lo . U_hi + h_hi . U_hi off the same U_hi stage
                    for (int kb = 0; kb < KBH; kb += 2) stage_mmas(d, ah_a, 0u, kb, min(2, KBH - kb), false);
                    for (int kb = 0; kb < KBH; kb += 2) stage_mmas(d, ah_a, smem_u32(sAl), kb, min(2, KBH - kb), false);
                } else
                for (int kb = 0; kb < KBH; kb += 2) stage_mmas(d, ah_a, 0u, kb, min(2, KBH - kb), false);
                if (rf_elect()) {
                    rf_commit(bars + RFB_ACC_FULL + buf, pair_mask);
                    if (S == 2 && c == NC - 1) rf_commit(bars + RFB_A_FREE, (uint16_t)0xF);    // this pair is done reading h_{t-1}
                }
                __syncwarp();
            };
            // the tile after (t, c) in issue order and its buffer bookkeeping
            int buf = 0, use = 0;                 // of the tile whose recurrent part is issued next
            // The input parts run LA tiles ahead of the recurrent parts.  LA = 2 (NNTK_RECURRENT_FUSED_LOOKAHEAD=2, three accumulator
            // buffers) puts X(t+1, 0) AND X(t+1, 1) into the tensor pipe while the last tile's epilogue closes step t -- measured
            // SLOWER on the C5 GRU layer (8.11 vs 7.75 ms), so the default stays 1.
            const int total = T * NC;
            int xt = 0, xc = 0, xbuf = 0, xuse = 0, xg = 0;      // step, tile, buffer and use count of the next input part
            auto x_next = [&]() {
                x_part(xt, xc, xbuf, xuse);
                if (++xc == NC) { xc = 0; ++xt; }
                if (++xbuf == NB) { xbuf = 0; ++xuse; }
                ++xg;
            };
            if (FUSED) while (xg < LA && xg < total) x_next();
            for (int t = 0; t < T; ++t)
                for (int c = 0; c < NC; ++c) {
                    h_part(t, c, buf, use);
                    if (FUSED && xg < total) x_next();
                    if (++buf == NB) { buf = 0; ++use; }
                }
            if (PUB && lane == 0) atomicMax(reinterpret_cast<int *>(const_cast<float *>(p.stats) + 2), __float_as_int(any3 ? 3.0f : 1.0f));
            RF_DBG(if (lane == 0) { for (int i = 0; i < 4; ++i) p.dbg[blockIdx.x * 16 + i] = dbgc[i]; p.dbg[blockIdx.x * 16 + 10] = clock64() - t_start;
                   p.dbg[blockIdx.x * 16 + 12] = dbgc[12]; p.dbg[blockIdx.x * 16 + 13] = dbgc[13]; });
        }
    } else if (FUSED && warp >= CONV_W0) {
        // ===== x converters (FUSED only): fp32 rows of x_t from global -> fp16 hi (and lo) k-block tiles =====
        // item (row, 16-byte chunk of the fp16 row = 8 floats): a warp covers 4 rows x 256 contiguous bytes per load instruction and
        // 512 contiguous (swizzled) bytes per store.  x_{t+1} is loaded right after x_t has been converted and waits in registers.
        if (FUSED && DBX) {
            // slim converters: x_t goes global -> registers (one k-block at a time) -> fp16 -> the x tile of parity t & 1, up to two
            // steps ahead of the MMAs; nothing is kept in registers across steps
            asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
            const int ct = tid - CONV_W0 * 32;
            const uint32_t xf_leader = rf_mapa(smem_u32(bars + RFB_XF2), leader);
            const uint32_t xs_leader = rf_mapa(smem_u32(xs), leader) + 4u * (uint32_t)((int)rank * 8 + (ct >> 5));
            const int nit = p.rows == RF_ROWS ? 2 : 1;
            const int r0 = ct >> 3, ch = ct & 7;                 // item (row r0 + 32 it, 16-byte chunk ch of the fp16 row)
            for (int t = 0; t < T; ++t) {
                const int b = t & 1;
                if (t >= 2) mbar_wait(bars + RFB_XE2 + b, (uint32_t)(((t - 2) >> 1) & 1));
                uint8_t *tile = sXh + (size_t)b * (RF_ROWS * IN * 2);
                float ss0 = 0.0f, ss1 = 0.0f;
                for (int kb = 0; kb < KBX; ++kb) {
                    float4 a[2][2];
#pragma unroll
                    for (int it = 0; it < 2; ++it) {
                        const int r = r0 + 32 * it, row = row0 + r;
                        if (it < nit && row < p.B) {
                            const float4 *src = reinterpret_cast<const float4 *>(p.x + ((size_t)row * T + t) * IN + kb * 64 + ch * 8);
                            a[it][0] = __ldg(src); a[it][1] = __ldg(src + 1);
                        } else a[it][0] = a[it][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int it = 0; it < 2; ++it) {
                        const float4 u = a[it][0], w = a[it][1];
                        const float q = fmaf(u.x, u.x, fmaf(u.y, u.y, fmaf(u.z, u.z, fmaf(u.w, u.w, fmaf(w.x, w.x, fmaf(w.y, w.y, fmaf(w.z, w.z, w.w * w.w)))))));
                        if (it == 0) ss0 += q; else ss1 += q;
                        const int r = r0 + 32 * it;
                        if (it < nit)
                            *reinterpret_cast<uint4 *>(tile + (size_t)kb * (RF_ROWS * 128) + r * 128 + ((ch ^ (r & 7)) << 4)) =
                                make_uint4(rf_pack_sat(u.x, u.y), rf_pack_sat(u.z, u.w), rf_pack_sat(w.x, w.y), rf_pack_sat(w.z, w.w));
                    }
                }
                if (PUB) {
#pragma unroll
                    for (int o = 1; o < 8; o <<= 1) { ss0 += __shfl_xor_sync(0xffffffffu, ss0, o); ss1 += __shfl_xor_sync(0xffffffffu, ss1, o); }
                    float m2 = fmaxf(ss0, ss1);
                    m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, 8));
                    m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, 16));
                    if (lane == 0) asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(xs_leader + 64u * (uint32_t)b), "f"(m2) : "memory");
                }
                rf_fence_async();
                __syncwarp();
                if (lane == 0) { if (PUB) rf_arrive_cl_release(xf_leader + 8u * (uint32_t)b); else rf_arrive_cl(xf_leader + 8u * (uint32_t)b); }
            }
        } else
        if (FUSED) {
            if (RF_FUSED_EPW == 8) asm volatile("setmaxnreg.dec.sync.aligned.u32 80;");
            else asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
            const int ct = tid - CONV_W0 * 32;
            const uint32_t xfull0 = rf_mapa(smem_u32(bars + RFB_X_FULL), leader), xlfull0 = rf_mapa(smem_u32(bars + RFB_XL_FULL), leader);
            // item (row, 16-byte chunk of the fp16 row = 8 floats): a warp covers 4 rows x 256 contiguous bytes per load instruction and
            // 512 contiguous (swizzled) bytes per store.  The fp16 HI image of x_{t+1} is built right after x_t has been handed over and
            // waits in registers (32 for in = 256): when the tile frees up it is stored at once and the input part can start.  The LO
            // image (3-pass mode) is recomputed then from a second read of the same rows -- L2 hits, one step old -- while the
            // x_hi . W_lo pass runs.  (Holding raw fp32 rows, 64 registers, spilled; loading k-blocks late stalled the issuer 2 900
            // cycles per step.)
            uint4 hreg[4][2];
            const int nit = p.rows == RF_ROWS ? 2 : 1;           // 32-row mode: items 256.. are rows 32..63 (stay zero)
            auto src_of = [&](int t, int kb, int it) -> const float4 * {
                const int idx = it * 256 + ct, r = idx >> 3, ch = idx & 7, row = row0 + r;
                return (row < p.B && r < p.rows) ? reinterpret_cast<const float4 *>(p.x + ((size_t)row * T + t) * IN + kb * 64 + ch * 8) : nullptr;
            };
            auto tile_off = [&](int kb, int it) -> size_t {
                const int idx = it * 256 + ct, r = idx >> 3, ch = idx & 7;
                return (size_t)kb * (RF_ROWS * 128) + r * 128 + ((ch ^ (r & 7)) << 4);
            };
            // per-step mode: this warp's max |x_t|^2 over its rows, published to the leader CTA with the x tile
            const uint32_t xs_leader = rf_mapa(smem_u32(xs), leader) + 4u * (uint32_t)((int)rank * 8 + (ct >> 5));
            auto preload = [&](int t) {
                float ss[2] = {0.0f, 0.0f};     // sum of squares of this thread's chunk of its two rows, over the k-blocks
                constexpr int KK = RF_FUSED_EPW == 8 ? 2 : 1;   // k-blocks of loads in flight at a time (64 registers per converter thread: one)
#pragma unroll
                for (int k2 = 0; k2 < 4; k2 += KK) {
                    float4 a[KK][2][2];
#pragma unroll
                    for (int kk = 0; kk < KK; ++kk)
#pragma unroll
                        for (int it = 0; it < 2; ++it) {
                            const float4 *src = (k2 + kk < KBX && it < nit) ? src_of(t, k2 + kk, it) : nullptr;
                            if (src) { a[kk][it][0] = __ldg(src); a[kk][it][1] = __ldg(src + 1); }
                            else a[kk][it][0] = a[kk][it][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                        }
#pragma unroll
                    for (int kk = 0; kk < KK; ++kk)
#pragma unroll
                        for (int it = 0; it < 2; ++it) {
                            const float4 u = a[kk][it][0], w = a[kk][it][1];
                            if (PUB) ss[it] = fmaf(u.x, u.x, fmaf(u.y, u.y, fmaf(u.z, u.z, fmaf(u.w, u.w, fmaf(w.x, w.x, fmaf(w.y, w.y, fmaf(w.z, w.z, fmaf(w.w, w.w, ss[it]))))))));
                            hreg[k2 + kk][it] = make_uint4(rf_pack_sat(u.x, u.y), rf_pack_sat(u.z, u.w), rf_pack_sat(w.x, w.y), rf_pack_sat(w.z, w.w));
                        }
                }
                if (PUB) {
                    // the 8 threads of a row are 8 consecutive lanes; then the max over the warp's 4 + 4 rows.  The slot of parity t was
                    // last read by the issuer for step t - 2, before it released the x tile this warp waited for since (X_EMPTY)
#pragma unroll
                    for (int o = 1; o < 8; o <<= 1) { ss[0] += __shfl_xor_sync(0xffffffffu, ss[0], o); ss[1] += __shfl_xor_sync(0xffffffffu, ss[1], o); }
                    float m2 = fmaxf(ss[0], ss[1]);
                    m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, 8));
                    m2 = fmaxf(m2, __shfl_xor_sync(0xffffffffu, m2, 16));
                    if (lane == 0) asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(xs_leader + 64u * (uint32_t)(t & 1)), "f"(m2) : "memory");
                }
            };
            preload(0);
            for (int t = 0; t < T; ++t) {
                if (t >= 1) { RF_T0(); mbar_wait(bars + RFB_X_EMPTY, (uint32_t)((t - 1) & 1)); RF_T1(7); }
                RF_T0();
                if (!RF_SKIP(4)) {
#pragma unroll
                    for (int kb = 0; kb < 4; ++kb) {
                        if (kb >= KBX) break;
#pragma unroll
                        for (int it = 0; it < 2; ++it) if (it < nit) *reinterpret_cast<uint4 *>(sXh + tile_off(kb, it)) = hreg[kb][it];
                    }
                }
                rf_fence_async();
                __syncwarp();
                if (lane == 0) { if (PUB) rf_arrive_cl_release(xfull0); else rf_arrive_cl(xfull0); }
                if (XP == 3) {
#pragma unroll
                    for (int kb = 0; kb < 4; ++kb) {
                        if (kb >= KBX) break;
                        if (RF_SKIP(16)) { if (lane == 0) rf_arrive_cl(xlfull0 + 8u * (uint32_t)kb); continue; }   // timing experiment: no lo image
                        float4 a[2][2];
#pragma unroll
                        for (int it = 0; it < 2; ++it) {
                            const float4 *src = it < nit ? src_of(t, kb, it) : nullptr;
                            if (src) { a[it][0] = __ldg(src); a[it][1] = __ldg(src + 1); } else a[it][0] = a[it][1] = make_float4(0.f, 0.f, 0.f, 0.f);
                        }
#pragma unroll
                        for (int it = 0; it < 2; ++it) {
                            const uint4 h = hreg[kb][it];
                            const float2 h0 = __half22float2(*reinterpret_cast<const __half2 *>(&h.x)), h1 = __half22float2(*reinterpret_cast<const __half2 *>(&h.y)),
                                         h2 = __half22float2(*reinterpret_cast<const __half2 *>(&h.z)), h3 = __half22float2(*reinterpret_cast<const __half2 *>(&h.w));
                            const float4 u = a[it][0], w = a[it][1];
                            const uint4 lo = make_uint4(rf_pack_sat(u.x - h0.x, u.y - h0.y), rf_pack_sat(u.z - h1.x, u.w - h1.y),
                                                        rf_pack_sat(w.x - h2.x, w.y - h2.y), rf_pack_sat(w.z - h3.x, w.w - h3.y));
                            if (!RF_SKIP(4) && it < nit) *reinterpret_cast<uint4 *>(sXl + tile_off(kb, it)) = lo;
                        }
                        rf_fence_async();
                        __syncwarp();
                        if (lane == 0) rf_arrive_cl(xlfull0 + 8u * (uint32_t)kb);
                    }
                }
                RF_T1(6);
                if (t + 1 < T) preload(t + 1);
            }
            RF_DBG(if (ct == 0) { p.dbg[blockIdx.x * 16 + 6] = dbgc[6]; p.dbg[blockIdx.x * 16 + 7] = dbgc[7]; });
        }
    } else {
        // ===== epilogue (warps 0..7): thread = row m of this CTA x 16 hidden units of every tile =====
        if (FUSED && DBX) asm volatile("setmaxnreg.inc.sync.aligned.u32 96;");
        else if (FUSED && RF_FUSED_EPW == 8) asm volatile("setmaxnreg.inc.sync.aligned.u32 128;");
        else if (FUSED) asm volatile("setmaxnreg.inc.sync.aligned.u32 80;");
        else asm volatile("setmaxnreg.inc.sync.aligned.u32 104;");
        const int ew = warp, q = warp & 3, j = ew >> 2;
        const int m = (q & 1) * 32 + lane, hf = q >> 1;
        const int row = row0 + m;
        const bool row_ok = row < p.B;
        const uint32_t tl = tmem + ((uint32_t)(q * 32) << 16);           // this warp's TMEM lanes
        const uint32_t acc_empty0 = rf_mapa(smem_u32(bars + RFB_ACC_EMPTY), leader), hfull0 = rf_mapa(smem_u32(bars + RFB_H_FULL), leader);
        const uint32_t hfull1 = rf_mapa(smem_u32(bars + RFB_H_FULL), leader ^ 2u);       // split mode: the other pair's leader
        const uint32_t sAh_peer = rf_mapa(smem_u32(sAh), crank ^ 2u);                    // split mode: the same rows' A tile in the other pair
        const int sw = m & 7;
        const uint32_t sBias_a = smem_u32(sBias), sAh_a = smem_u32(sAh), sOut_a = smem_u32(sOut);
        if (p.rows == RF_ROWS || !(q & 1)) {       // 32-row mode: lane quarters 1 and 3 hold no sequences
        // ---- prologue: state into TMEM, fp16 h_0 into the A tile, zero the h.Uh groups
        for (int c = 0; c < NCT; ++c) {                  // the A tile holds ALL of h; TMEM only the state of this pair's tiles
            const bool own = c >= tile0 && c < tile0 + NC;
#pragma unroll
            for (int e = 0; e < NE; ++e) {
                const int u0 = c * 64 + hf * 32 + j * UPT + e * 8;
                float hv[8], sv[8];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const float4 a = row_ok ? *reinterpret_cast<const float4 *>(p.h_init + (size_t)row * H + u0 + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
                    hv[4 * i] = a.x; hv[4 * i + 1] = a.y; hv[4 * i + 2] = a.z; hv[4 * i + 3] = a.w;
                    if (G == 4 && own) {
                        const float4 b = row_ok ? *reinterpret_cast<const float4 *>(p.c_init + (size_t)row * H + u0 + 4 * i) : make_float4(0.f, 0.f, 0.f, 0.f);
                        sv[4 * i] = b.x; sv[4 * i + 1] = b.y; sv[4 * i + 2] = b.z; sv[4 * i + 3] = b.w;
                    }
                }
                if (own) {
                    if (G == 4) rf_st8(tl + STATE_COL + (uint32_t)((c - tile0) * 32 + j * UPT + e * 8), sv);
                    else rf_st8(tl + STATE_COL + (uint32_t)((c - tile0) * 32 + j * UPT + e * 8), hv);
                }
                uint4 v;
                v.x = rf_pack_sat(hv[0], hv[1]); v.y = rf_pack_sat(hv[2], hv[3]); v.z = rf_pack_sat(hv[4], hv[5]); v.w = rf_pack_sat(hv[6], hv[7]);
                const int ch = (hf * 32 + j * UPT) / 8 + e;
                *reinterpret_cast<uint4 *>(sAh + (size_t)c * (RF_ROWS * 128) + m * 128 + ((ch ^ sw) << 4)) = v;
                if (X3) {
                    uint4 hi4, lo4;
                    rf_split2(hv[0], hv[1], hi4.x, lo4.x); rf_split2(hv[2], hv[3], hi4.y, lo4.y);
                    rf_split2(hv[4], hv[5], hi4.z, lo4.z); rf_split2(hv[6], hv[7], hi4.w, lo4.w);
                    *reinterpret_cast<uint4 *>(sAl + (size_t)c * (RF_ROWS * 128) + m * 128 + ((ch ^ sw) << 4)) = lo4;
                }
            }
        }
        // XW mode: the hoisted projection of a tile is STORED INTO ITS ACCUMULATOR BUFFER before the tile's MMAs run (which then
        // accumulate onto it): the loads are issued a whole tile earlier, while the previous occupant of the buffer is being
        // processed, so their latency never sits between the MMA and the cell math (loaded and waited for per tile they made the
        // LSTM-512 epilogue 4 700 cycles against 2 064 of MMA).  Column groups as in FUSED mode: GRU [x.Wh | z | r | h.Uh = 0],
        // LSTM [i | f | g | o].
        auto xw_src = [&](int t2, int c2) -> const float4 * {
            return reinterpret_cast<const float4 *>(p.xw) +
                   (((long long)t2 * p.xw_blocked_rows + (row >> 5)) * (G * H >> 2) + (((tile0 + c2) * 64 + hf * 32 + j * UPT) >> 2)) * 32 + (row & 31);
        };
        auto xw_load = [&](float4 (&xn)[G][UPT / 4], int t2, int c2) {
            const float4 *src = xw_src(t2, c2);
#pragma unroll
            for (int gg = 0; gg < G; ++gg)
#pragma unroll
                for (int i = 0; i < UPT / 4; ++i)
                    xn[gg][i] = (row_ok && t2 < T) ? __ldg(src + ((long long)gg * (H >> 2) + i) * 32) : make_float4(0.f, 0.f, 0.f, 0.f);
        };
        auto xw_store = [&](const float4 (&xn)[G][UPT / 4], int b) {       // projection gate gg -> column group: GRU z, r, h -> 1, 2, 0
#pragma unroll
            for (int gg = 0; gg < G; ++gg) {
                const int grp = G == 3 ? (gg == 2 ? 0 : gg + 1) : gg;
#pragma unroll
                for (int e = 0; e < NE; ++e) {
                    const float v8[8] = {xn[gg][2 * e].x, xn[gg][2 * e].y, xn[gg][2 * e].z, xn[gg][2 * e].w,
                                         xn[gg][2 * e + 1].x, xn[gg][2 * e + 1].y, xn[gg][2 * e + 1].z, xn[gg][2 * e + 1].w};
                    rf_st8(tl + (uint32_t)(b * 128 + grp * 32 + j * UPT + e * 8), v8);
                }
            }
        };
        if (G == 3) {
            const float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            for (int b = 0; b < NB; ++b)
#pragma unroll
                for (int e = 0; e < NE; ++e) rf_st8(tl + (uint32_t)(b * 128 + 96 + j * UPT + e * 8), z8);
        }
        if (!FUSED) {
            for (int b = 0; b < NB; ++b) {          // tiles 0 .. NB-1 in issue order
                float4 xn[G][UPT / 4];
                xw_load(xn, b / NC, b % NC);
                xw_store(xn, b);
            }
        }
        rf_st_wait();
        rf_fence_async();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) { rf_arrive_cl(hfull0); if (S == 2) rf_arrive_cl(hfull1); }
        uint32_t hpend[NCMAX][UPT / 2]; // fp16 h_t of the step's tiles x UPT units, written to the A tile once the step's MMAs are done
        int buf = -1, use = 0;          // accumulator buffer of the current tile and how often it has been used before
        for (int t = 0; t < T; ++t) {
#pragma unroll
            for (int c = 0; c < NCMAX; ++c) {
                if (c >= NC) break;
                if (++buf == NB) { buf = 0; ++use; }
                const uint32_t acc = tl + (uint32_t)(buf * 128 + j * UPT);
                float4 xn[G][UPT / 4];   // XW mode: the hoisted projection of the NEXT occupant of this buffer (tile g + NB)
                if (!FUSED) {
                    int c2 = c + NB, t2 = t;
                    while (c2 >= NC) { c2 -= NC; ++t2; }
                    xw_load(xn, t2, c2);
                }
                { RF_T0(); mbar_wait(bars + RFB_ACC_FULL + buf, (uint32_t)(use & 1)); RF_T1(4); }
                tc_fence_after();
                float hn[UPT];
#pragma unroll
                for (int i = 0; i < UPT; ++i) hn[i] = 0.f;
#pragma unroll
                for (int e = 0; e < NE; ++e) {
                    if (RF_SKIP(1)) break;
                    const int ub = (tile0 + c) * 64 + hf * 32 + j * UPT + e * 8;  // first of these 8 hidden units
                    float a0[8], a1[8], a2[8], a3[8], st[8];
                    rf_ld8(acc + (uint32_t)(e * 8), a0);
                    rf_ld8(acc + (uint32_t)(32 + e * 8), a1);
                    rf_ld8(acc + (uint32_t)(64 + e * 8), a2);
                    rf_ld8(acc + (uint32_t)(96 + e * 8), a3);
                    rf_ld8(tl + STATE_COL + (uint32_t)(c * 32 + j * UPT + e * 8), st);
                    rf_ld_wait();
                    float b0[8], b1[8], b2[8], b3[8];
#pragma unroll
                    for (int i = 0; i < 2; ++i) {
                        const float4 v0 = rf_lds128(sBias_a + 4u * (uint32_t)(0 * H + ub + 4 * i)), v1 = rf_lds128(sBias_a + 4u * (uint32_t)(1 * H + ub + 4 * i)),
                                     v2 = rf_lds128(sBias_a + 4u * (uint32_t)(2 * H + ub + 4 * i)), v3 = rf_lds128(sBias_a + 4u * (uint32_t)(3 * H + ub + 4 * i));
                        b0[4 * i] = v0.x; b0[4 * i + 1] = v0.y; b0[4 * i + 2] = v0.z; b0[4 * i + 3] = v0.w;
                        b1[4 * i] = v1.x; b1[4 * i + 1] = v1.y; b1[4 * i + 2] = v1.z; b1[4 * i + 3] = v1.w;
                        b2[4 * i] = v2.x; b2[4 * i + 1] = v2.y; b2[4 * i + 2] = v2.z; b2[4 * i + 3] = v2.w;
                        b3[4 * i] = v3.x; b3[4 * i + 1] = v3.y; b3[4 * i + 2] = v3.z; b3[4 * i + 3] = v3.w;
                    }
                    float sn[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        if (G == 3) {          // gru.c:156-186; columns [x.Wh | z | r | h.Uh]
                            const float xh = a0[u] + b2[u], pz = a1[u] + b0[u], pr = a2[u] + b1[u], hu = a3[u] + b3[u];
                            const float z = X3 ? nntk_act(NNTK_ACT_SIGMOID, 1.0f, pz) : act_mufu(NNTK_ACT_SIGMOID, pz);
                            const float r = X3 ? nntk_act(NNTK_ACT_SIGMOID, 1.0f, pr) : act_mufu(NNTK_ACT_SIGMOID, pr);
                            const float hh = X3 ? nntk_act(NNTK_ACT_TANH, 1.0f, fmaf(r, hu, xh)) : act_mufu(NNTK_ACT_TANH, fmaf(r, hu, xh));
                            sn[u] = fmaf(z, st[u] - hh, hh);               // (1 - z) hh + z h_{t-1}
                            hn[e * 8 + u] = sn[u];
                        } else {               // lstm.c:199-238; columns [i | f | g | o]
                            const float pi = a0[u] + b0[u], pf = a1[u] + b1[u], pg = a2[u] + b2[u], po = a3[u] + b3[u];
                            const float ig = act_mufu(NNTK_ACT_SIGMOID, pi), fg = act_mufu(NNTK_ACT_SIGMOID, pf);
                            const float gg = act_mufu(NNTK_ACT_TANH, pg), og = act_mufu(NNTK_ACT_SIGMOID, po);
                            sn[u] = fmaf(fg, st[u], ig * gg);
                            hn[e * 8 + u] = og * act_mufu(NNTK_ACT_TANH, sn[u]);
                        }
                    }
                    rf_st8(tl + STATE_COL + (uint32_t)(c * 32 + j * UPT + e * 8), sn);
                    if (G == 3) { const float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}; rf_st8(acc + (uint32_t)(96 + e * 8), z8); }
                }
                if (!FUSED) xw_store(xn, buf);
                rf_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) rf_arrive_cl(acc_empty0 + 8u * (uint32_t)buf);       // the accumulator buffer may be overwritten
#pragma unroll
                for (int i = 0; i < UPT / 2; ++i) if (!X3) hpend[c][i] = rf_pack_sat(hn[2 * i], hn[2 * i + 1]);
                if (p.return_sequences && !RF_SKIP(2)) {
                    // this warp's [32 rows][UPT units] piece of h_t -> its own staging tile (UPT = 16: 64-byte rows, SWIZZLE_64B; UPT = 8:
                    // 32-byte rows, unswizzled) -> one TMA store per warp: no CTA-wide barrier in the epilogue (first version: two
                    // bar.sync + a store wait per tile)
                    uint8_t *o = sOut + ew * (32 * UPT * 4);
                    const uint32_t o_a = sOut_a + (uint32_t)(ew * (32 * UPT * 4));
                    { RF_T0(); if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); RF_T1(11); }    // the previous store has read it
                    __syncwarp();
#pragma unroll
                    for (int i = 0; i < UPT / 4; ++i)
                        rf_sts128(o_a + (uint32_t)(lane * (UPT * 4) + ((UPT == 16 ? (i ^ ((lane >> 1) & 3)) : i) << 4)),
                                  __float_as_uint(hn[4 * i]), __float_as_uint(hn[4 * i + 1]), __float_as_uint(hn[4 * i + 2]), __float_as_uint(hn[4 * i + 3]));
                    rf_fence_async();
                    __syncwarp();
                    if (lane == 0) {
                        rf_tma_store_3d(&map_out, o, (tile0 + c) * 64 + hf * 32 + j * UPT, t, row0 + (q & 1) * 32);
                        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                    }
                }
                if (c == NC - 1) {
                    // every recurrent-part MMA of this step has completed (they precede this tile's commit): h_t may now
                    // replace h_{t-1} in the A tile.  The last step's h goes to the caller's state instead.
                    if (t + 1 < T) {
                        // split mode: the other pair's MMAs of this step read ITS copy of h_{t-1}, which this warp is about to overwrite too
                        if (S == 2) mbar_wait(bars + RFB_A_FREE, (uint32_t)(t & 1));
#pragma unroll
                        for (int cc = 0; cc < NCMAX; ++cc) {
                            if (cc >= NC) break;
#pragma unroll
                            for (int e = 0; e < NE; ++e) {
                                const int ch = (hf * 32 + j * UPT) / 8 + e;
                                const uint32_t off = (uint32_t)((tile0 + cc) * (RF_ROWS * 128) + m * 128 + ((ch ^ sw) << 4));
                                if (X3) {
                                    // the exact fp32 h_t of tile cc sits in this thread's state columns (written above, st waited for):
                                    // both A tiles are rebuilt from it -- no register-resident pending copy in this variant
                                    float hx[8];
                                    rf_ld8(tl + STATE_COL + (uint32_t)(cc * 32 + j * UPT + e * 8), hx);
                                    rf_ld_wait();
                                    uint4 hi4, lo4;
                                    rf_split2(hx[0], hx[1], hi4.x, lo4.x); rf_split2(hx[2], hx[3], hi4.y, lo4.y);
                                    rf_split2(hx[4], hx[5], hi4.z, lo4.z); rf_split2(hx[6], hx[7], hi4.w, lo4.w);
                                    rf_sts128(sAh_a + off, hi4.x, hi4.y, hi4.z, hi4.w);
                                    rf_sts128(smem_u32(sAl) + off, lo4.x, lo4.y, lo4.z, lo4.w);
                                    continue;
                                }
                                rf_sts128(sAh_a + off, hpend[cc][4 * e], hpend[cc][4 * e + 1], hpend[cc][4 * e + 2], hpend[cc][4 * e + 3]);
                                if (S == 2)
                                    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sAh_peer + off), "r"(hpend[cc][4 * e]), "r"(hpend[cc][4 * e + 1]),
                                                 "r"(hpend[cc][4 * e + 2]), "r"(hpend[cc][4 * e + 3]) : "memory");
                            }
                        }
                        if (S == 2) asm volatile("fence.proxy.async;" ::: "memory"); else rf_fence_async();
                        // the remote stores must have landed in the other pair's shared memory before its leader's barrier can complete
                        if (S == 2) asm volatile("fence.acq_rel.cluster;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) { rf_arrive_cl(hfull0); if (S == 2) rf_arrive_cl(hfull1); }
                    }
                }
                if (t == T - 1 && row_ok) {
                    // final state (gru.c:201, lstm.c:264-265): exact fp32 h_T (and c_T from TMEM below)
                    float *hs = p.h_state + (size_t)row * H + (tile0 + c) * 64 + hf * 32 + j * UPT;
#pragma unroll
                    for (int i = 0; i < UPT / 4; ++i) {
                        const float4 v = make_float4(hn[4 * i], hn[4 * i + 1], hn[4 * i + 2], hn[4 * i + 3]);
                        *reinterpret_cast<float4 *>(hs + 4 * i) = v;
                        if (!p.return_sequences) *reinterpret_cast<float4 *>(p.out + (size_t)row * H + (tile0 + c) * 64 + hf * 32 + j * UPT + 4 * i) = v;
                    }
                }
            }
        }
        if (G == 4) {
            tc_fence_after();
            for (int c = 0; c < NC; ++c)
#pragma unroll
                for (int e = 0; e < NE; ++e) {
                    float sv[8];
                    rf_ld8(tl + STATE_COL + (uint32_t)(c * 32 + j * UPT + e * 8), sv);
                    rf_ld_wait();
                    if (row_ok) {
                        float *cs = p.c_state + (size_t)row * H + (tile0 + c) * 64 + hf * 32 + j * UPT + e * 8;
                        *reinterpret_cast<float4 *>(cs) = make_float4(sv[0], sv[1], sv[2], sv[3]);
                        *reinterpret_cast<float4 *>(cs + 4) = make_float4(sv[4], sv[5], sv[6], sv[7]);
                    }
                }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");      // this warp's last stores have landed
        }
        RF_DBG(if (ew == 0 && lane == 0) { p.dbg[blockIdx.x * 16 + 4] = dbgc[4]; p.dbg[blockIdx.x * 16 + 5] = clock64() - t_start; p.dbg[blockIdx.x * 16 + 11] = dbgc[11]; });
    }
    tc_fence_before();
    __syncthreads();
    rf_cluster_sync();               // nobody leaves while the peer's MMA / TMA could still touch this CTA
    if (warp == CTRL_W0 + 3) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}

// B operand repack: Wp[(tile * 2 + half) * 32 G + gi * 32 + ul][k] = half(W[k][gate(gi) * H + tile * 64 + half * 32 + ul]);
// gate order `perm`: recurrent part GRU {z, r, h}, input part GRU {h, z, r} (column groups [x.Wh | z | r]), LSTM {i, f, g, o}.
// Wl (optional): the fp16 image of the rounding residual W - half(W), for the 3-pass input part.
__global__ void rf_pack_kernel(const float *__restrict__ W, __half *__restrict__ Wp, __half *__restrict__ Wl, int K, int H, int G, int p0, int p1,
                               int p2, int p3) {
    const long long total = (long long)G * H * K;
    const int perm[4] = {p0, p1, p2, p3};
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / K), k = (int)(i - (long long)n * K);
        const int th = n / (32 * G), rem = n - th * 32 * G, gi = rem >> 5, ul = rem & 31;
        const int tile = th >> 1, half = th & 1;
        const float w = __ldg(W + (long long)k * G * H + perm[gi] * H + tile * 64 + half * 32 + ul);
        const __half hi = __float2half_rn(w);
        Wp[i] = hi;
        if (Wl) Wl[i] = __float2half_rn(w - __half2float(hi));
    }
}

// rf_xpasses_note -- adaptive input passes.  Calibration (tests, emulation, scripts): with ONE fp16 pass for x . W
//   GRU 256 -> 256, N(0,1) inputs (max |x_t| max |W_n| = 18), 300 steps: 1.2e-3 max error   -> three passes needed
//   C5 GRU layers (1.7 and 2.3: post-ReLU features, h in [-1, 1]), 994 steps: 1.3e-4 / 2.1e-4
//   C4 second LSTM-512 layer (5.6: its input is the first layer's h), 500 steps, whole stack vs the reference: 1.6e-4 scaled max
//   error, 2.9e-4 rel-L2 (three passes: 0.7e-4 / 1.5e-4)
// The error is linear in that product, so a bound of 6.5 keeps the single-pass noise at ~4e-4, inside the mode's 1e-3 together
// with the recurrent part's own fp16 rounding (<= 3.6e-4).  NNTK_RECURRENT_X_ADAPT_THR overrides it.
__global__ void __launch_bounds__(256) rf_xnorm_kernel(const float *__restrict__ x, long long rows, int in, float *stats, int stride) {
    const int lane = threadIdx.x & 31;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    const int n4 = in >> 2;
    float best = 0.0f;
    // four rows per warp iteration: their loads are independent and all in flight before the first shuffle (one row at a time the
    // load -> 5 shuffles -> next load chain ran at 4.2 TB/s)
    for (long long r = (((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5) * 4; r * stride < rows; r += warps * 4) {
        float s[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        for (int i = lane; i < n4; i += 32) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if ((r + k) * stride < rows) {          // stride > 1: a sample (every stride-th row); the launch that trusts it verifies every step
                    const float4 v = __ldcs(reinterpret_cast<const float4 *>(x + (r + k) * stride * in) + i);
                    s[k] = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, s[k]))));
                }
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
#pragma unroll
            for (int k = 0; k < 4; ++k) s[k] += __shfl_xor_sync(0xffffffffu, s[k], o);
        }
        best = fmaxf(best, fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3])));          // NaN / Inf inputs: fmaxf drops NaN, Inf stays -> 3 passes
    }
    if (lane == 0) atomicMax(reinterpret_cast<int *>(stats), __float_as_int(best));      // non-negative floats order like ints
}
__global__ void rf_wnorm_kernel(const float *__restrict__ W, int in, int cols, float *stats) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    float s = 0.0f;
    if (n < cols) for (int k = 0; k < in; ++k) { const float w = __ldg(W + (long long)k * cols + n); s = fmaf(w, w, s); }
#pragma unroll
    for (int o = 16; o; o >>= 1) s = fmaxf(s, __shfl_xor_sync(0xffffffffu, s, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<int *>(stats + 1), __float_as_int(s));
}
float *g_rf_stats = nullptr;

__half *g_rf_scratch = nullptr;
size_t g_rf_halves = 0;

template <int G, bool FUSED>
int rf_ring_stages(int H, int in, int xp, bool dbx = false, bool x3 = false) {
    int stages = RF_MAX_RING;
    for (; stages >= 2; --stages) { RfPlan<G, FUSED> pl(H, in, xp, stages, dbx, x3); if (pl.total <= 227 * 1024) break; }
    return stages;
}
CUtensorMap g_rf_map_ul;         // X3 launches: the tensor map of U_lo (set by the caller of rf_launch; a copy of the U map otherwise)

// dry: only report whether a cluster of p.cs CTAs of this kernel can be scheduled (1 = no)
template <int G, bool FUSED, int NCMAX, bool DBX = false, bool X3 = false>
int rf_launch(const CUtensorMap &mw, const CUtensorMap &mwl, const CUtensorMap &mu, const CUtensorMap &mo, RfParams &p, int pairs, cudaStream_t st, int *dry) {
    auto kern = recurrent_fused_kernel<G, FUSED, NCMAX, DBX, X3>;
    const int stages = rf_ring_stages<G, FUSED>(p.H, p.in, p.xp, DBX, X3);
    if (stages < 3) return 1;
    p.ring_stages = stages;
    const RfPlan<G, FUSED> pl(p.H, p.in, p.xp, stages, DBX, X3);
    static size_t attr_smem = 0;
    if ((size_t)pl.total > attr_smem) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total));
        attr_smem = pl.total;
    }
    const int npairs = p.cs / 2, clusters = (pairs + npairs - 1) / npairs;       // surplus pairs run on out-of-range rows
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(clusters * p.cs));
    cfg.blockDim = dim3(FUSED ? (DBX ? 896 : RF_THREADS_FUSED) : RF_THREADS_XW);
    cfg.dynamicSmemBytes = pl.total;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (dry) {           // *dry = pairs this cluster size keeps resident at once (0: cannot be scheduled)
        int maxc = 0;
        if (cudaOccupancyMaxActiveClusters(&maxc, kern, &cfg) != cudaSuccess || maxc < 1) { (void)cudaGetLastError(); *dry = 0; return 1; }
        *dry = (clusters < maxc ? clusters : maxc) * npairs;
        return 0;
    }
    NNTK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, mw, mwl, mu, g_rf_map_ul, mo, p));
    nntk_cuda_count_launch(1);
    return 0;
}

int rf_xpasses() {
    static int xp = 0;
    if (!xp) { const char *e = getenv("NNTK_RECURRENT_X_PASSES"); xp = (e && atoi(e) == 1) ? 1 : 3; }
    return xp;
}

}  // namespace

// 0: not eligible; 1: FUSED (no projection needed); 2: XW mode (needs the blocked hoisted projection)
extern "C" int nntk_cuda_recurrent_fused_plan(const nntk_recurrent_args *r, int nb) {
    static int enabled = -1, min_batch = 0;
    if (enabled < 0) {
        const char *e = getenv("NNTK_RECURRENT_FUSED"); enabled = e ? atoi(e) : 1;
        e = getenv("NNTK_RECURRENT_FUSED_MIN_BATCH"); min_batch = e ? atoi(e) : 65;
    }
    const int G = r->cell, H = r->H, T = r->T;
    if (!enabled || (G != 3 && G != 4) || (H & 63) || H < 64 || H > 512 || T < 1 || nb < min_batch) return 0;
    const int SG = NNTK_ACT_SIGMOID, TH = NNTK_ACT_TANH;
    const bool def = (G == 3) ? (r->act[0] == SG && r->act[1] == SG && r->act[2] == TH)
                              : (r->act[0] == SG && r->act[1] == SG && r->act[2] == TH && r->act[3] == SG && r->act[4] == TH);
    if (!def || !get_encode()) return 0;
    if ((r->in & 63) == 0 && r->in <= 256 && (((uintptr_t)r->d_x) & 15) == 0) {
        const int stages = G == 3 ? rf_ring_stages<3, true>(H, r->in, rf_xpasses()) : rf_ring_stages<4, true>(H, r->in, rf_xpasses());
        if (stages >= 3) return 1;
    }
    return 2;
}

// input passes the most recent FUSED launch actually took (1 or 3; 0: none ran).  Synchronises the stream: diagnostics / tests only.
extern "C" int nntk_cuda_last_recurrent_xpasses(void) {
    if (!g_rf_stats) return 0;
    float v[4] = {0, 0, 0, 0};
    if (cudaStreamSynchronize(nntk_stream()) != cudaSuccess || cudaMemcpy(v, g_rf_stats, sizeof(v), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (int)v[2];
}

float *g_rf_state = nullptr;     // copies of the initial state for the speculative pair of launches
size_t g_rf_state_floats = 0;

// One launch.  spec: 0 plain; 1 / 2 = first (one-pass plan, verified per step, runs iff the sampled bound holds) / second (three-pass
// staging with the per-step decision, runs iff the bound failed or the first launch raised the flag) launch of the speculative pair;
// the caller has prepared g_rf_stats for them.  h_init / c_init: where the prologue reads the initial state.
static int rf_fused_run(const nntk_recurrent_args *r, int nb, const float *x, const float *xw, float *h, float *c, float *out, int *used,
                        int spec, const float *h_init, const float *c_init, int x3 = 0) {
    *used = 0;
    const int G = r->cell, H = r->H, T = r->T, IN = r->in;
    const int plan = nntk_cuda_recurrent_fused_plan(r, nb);
    if (!plan || (plan == 2 && !xw)) return 0;
    const bool fused = plan == 1 && !xw;
    if (x3 && !(fused && G == 3)) return 0;          // the fp32-contract variant exists for the FUSED GRU only
    if ((((uintptr_t)h) | ((uintptr_t)out) | ((uintptr_t)x) | ((uintptr_t)c) | ((uintptr_t)xw)) & 15) return 0;
    EncodeTiledFn encode = get_encode();
    cudaStream_t st = nntk_stream();
    const int xp = fused ? ((spec == 1 && !x3) ? 1 : (x3 ? 3 : rf_xpasses())) : 1;
    const size_t un = (size_t)G * H * H * (x3 ? 2 : 1), wn = fused ? (size_t)G * H * IN : 0, wln = (fused && xp == 3) ? wn : 0;
    if (un + wn + wln > g_rf_halves) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_rf_scratch) cudaFree(g_rf_scratch);
        g_rf_scratch = nullptr; g_rf_halves = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_rf_scratch, (un + wn + wln) * sizeof(__half)));
        g_rf_halves = un + wn + wln;
    }
    __half *Up = g_rf_scratch, *Wp = g_rf_scratch + un, *Wl = wln ? g_rf_scratch + un + wn : nullptr;
    {
        const int blocks = (int)((un + 255) / 256 < 2048 ? (un + 255) / 256 : 2048);
        rf_pack_kernel<<<blocks, 256, 0, st>>>(r->d_U, Up, x3 ? Up + un / 2 : nullptr, H, H, G, 0, 1, 2, 3);
        NNTK_LAUNCH_CHECK("rf_pack_kernel");
        if (fused) {
            const int bw = (int)((wn + 255) / 256 < 2048 ? (wn + 255) / 256 : 2048);
            if (G == 3) rf_pack_kernel<<<bw, 256, 0, st>>>(r->d_W, Wp, Wl, IN, H, G, 2, 0, 1, 0);
            else rf_pack_kernel<<<bw, 256, 0, st>>>(r->d_W, Wp, Wl, IN, H, G, 0, 1, 2, 3);
            NNTK_LAUNCH_CHECK("rf_pack_kernel");
        }
    }
    // cluster size: as many pairs as share a B stream (at most 4), shrunk while the device cannot schedule such a cluster
    int pairs = (nb + 127) / 128;
    static int max_cs = -1;
    if (max_cs < 0) { const char *e = getenv("NNTK_RECURRENT_FUSED_CLUSTER"); max_cs = e ? atoi(e) : 8; if (max_cs != 2 && max_cs != 4 && max_cs != 8) max_cs = 8; }
    int cs = 2;
    CUtensorMap mw, mwl, mu, mo;
    RfParams p = {};
    auto enc2 = [&](CUtensorMap *m, void *base, int K) -> int {
        cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)G * H};
        cuuint64_t strides[1] = {(cuuint64_t)K * 2};
        cuuint32_t box[2] = {64, (cuuint32_t)(32 * G / (p.split == 2 ? 1 : cs / 2))};       // one pair's slice of a CTA's rows
        cuuint32_t es[2] = {1, 1};
        CUresult cr = encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) { nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_fused: weight tensor map rejected (CUresult %d)", (int)cr); return NNTK_ERR_CUDA_; }
        return 0;
    };
    int rc;
    p.H = H; p.in = IN; p.xp = xp; p.rows = RF_ROWS; p.split = 1; p.x3 = x3;
    bool small = H <= 256;
    // NNTK_RECURRENT_FUSED_DBX=1: one-pass FUSED launches take the DBX structure (16 epilogue warps at 96 registers, slim converters,
    // double-buffered x tile).  Measured on the C5 GRU layer: 8.15 ms against 7.85 ms for the 8-warp epilogue -- twice the epilogue
    // warps do not shorten the step, so the step is NOT bound by the epilogue's own speed: per step the tensor pipe is busy
    // ~7.3 k cycles (128 MMAs of >= 57 cycles) and idles through the recurrence's chain (last recurrent MMA -> last tile's
    // epilogue -> h_full -> first recurrent MMA of the next step).  Kept buildable, off by default.
    static int want_dbx = -1;
    if (want_dbx < 0) { const char *e = getenv("NNTK_RECURRENT_FUSED_DBX"); want_dbx = e ? atoi(e) : 0; }
    const bool dbx = want_dbx && fused && xp == 1 && !x3 &&
                     (G == 3 ? rf_ring_stages<3, true>(H, IN, 1, true) : rf_ring_stages<4, true>(H, IN, 1, true)) >= 3;
#define RF_DISPATCH(DRY) \
    (x3 ? (small ? rf_launch<3, true, 4, false, true>(mw, mwl, mu, mo, p, pairs, st, DRY) : rf_launch<3, true, 8, false, true>(mw, mwl, mu, mo, p, pairs, st, DRY)) : \
     dbx ? (G == 3 ? (small ? rf_launch<3, true, 4, true>(mw, mwl, mu, mo, p, pairs, st, DRY) : rf_launch<3, true, 8, true>(mw, mwl, mu, mo, p, pairs, st, DRY)) \
                   : (small ? rf_launch<4, true, 4, true>(mw, mwl, mu, mo, p, pairs, st, DRY) : rf_launch<4, true, 8, true>(mw, mwl, mu, mo, p, pairs, st, DRY))) : \
     G == 3 ? (fused ? (small ? rf_launch<3, true, 4>(mw, mwl, mu, mo, p, pairs, st, DRY) : rf_launch<3, true, 8>(mw, mwl, mu, mo, p, pairs, st, DRY)) \
                     : (small ? rf_launch<3, false, 4>(mw, mwl, mu, mo, p, pairs, st, DRY) : rf_launch<3, false, 8>(mw, mwl, mu, mo, p, pairs, st, DRY))) \
            : (fused ? (small ? rf_launch<4, true, 4>(mw, mwl, mu, mo, p, pairs, st, DRY) : rf_launch<4, true, 8>(mw, mwl, mu, mo, p, pairs, st, DRY)) \
                     : (small ? rf_launch<4, false, 4>(mw, mwl, mu, mo, p, pairs, st, DRY) : rf_launch<4, false, 8>(mw, mwl, mu, mo, p, pairs, st, DRY))))
    // The step is a latency chain, so a second wave doubles the time: take the cluster size that keeps the most pairs resident at
    // once (ties: the larger cluster, it shares the B stream more widely).  Measured at 64 pairs: clusters of 8 run in two waves
    // (43.7 ms per C5 GRU layer), clusters of 4 in one (12.2 ms).
    int best_cs = 0, best_resident = 0;
    auto pick = [&](int np) -> int {
        best_cs = 0; best_resident = 0;
        int c0 = np >= 4 ? 8 : np >= 2 ? 4 : 2;
        if (c0 > max_cs) c0 = max_cs;
        if (p.split == 2) c0 = 4;        // a split slice IS a 4-CTA cluster
        for (int c = c0; c >= (p.split == 2 ? 4 : 2); c /= 2) {
            int resident = 0;
            p.cs = c;
            const int pairs = np;        // RF_DISPATCH reads `pairs`
            const int rc2 = RF_DISPATCH(&resident);
            if (rc2 < 0) return rc2;
            if (resident > best_resident) { best_resident = resident; best_cs = c; }
        }
        return 0;
    };
    // A batch whose 128-sequence pairs would leave most SMs idle is spread wider, as long as everything stays resident in ONE wave
    // (the step is a latency chain: a second wave doubles the time).  In order of preference:
    //   split 2 + 32 rows : 4 x the pairs        split 2 : hidden units of a slice across two pairs (each streams half of the
    //   weights and owns half of the tiles; h_t halves exchanged through DSMEM)        32 rows : half-filled M tile
    // NNTK_RECURRENT_FUSED_ROWS = 64 / 32 and NNTK_RECURRENT_FUSED_SPLIT = 1 / 2 force a shape.
    static int force_rows = -1, force_split = -1;
    if (force_rows < 0) { const char *e = getenv("NNTK_RECURRENT_FUSED_ROWS"); force_rows = e ? atoi(e) : 0; }
    if (force_split < 0) { const char *e = getenv("NNTK_RECURRENT_FUSED_SPLIT"); force_split = e ? atoi(e) : 0; }
    const bool can_split = (H % 128) == 0 && !x3;
    bool chosen = false;
    for (int opt = 0; opt < 4 && !chosen; ++opt) {
        const int o_split = opt < 2 ? 2 : 1, o_rows = (opt & 1) ? 64 : 32;      // (2,32) (2,64) (1,32) (1,64)
        if (o_split == 2 && !can_split) continue;
        if ((force_rows == 32 || force_rows == 64) && o_rows != force_rows) continue;
        if ((force_split == 1 || force_split == 2) && o_split != force_split) continue;
        const int np = ((nb + 2 * o_rows - 1) / (2 * o_rows)) * o_split;
        p.split = o_split; p.rows = o_rows; small = H / o_split <= 256;
        if ((rc = pick(np)) < 0) return rc;
        const bool last = (o_split == 1 && o_rows == 64) || (force_rows && force_split) || (force_split == o_split && force_rows == 0 && o_rows == 64) ||
                          (force_rows == o_rows && force_split == 0 && o_split == 1);
        if (best_cs && (best_resident >= np || last)) { pairs = np; chosen = true; }
    }
    if (!chosen) { p.split = 1; p.rows = RF_ROWS; small = H <= 256; if ((rc = pick(pairs)) < 0) return rc; }
    if (!best_cs) return 0;          // not even a pair fits: not eligible
    cs = best_cs;
    p.cs = cs;
    if ((rc = enc2(&mu, Up, H))) return rc;
    if (x3) { if ((rc = enc2(&g_rf_map_ul, Up + un / 2, H))) return rc; } else g_rf_map_ul = mu;
    if (fused) { if ((rc = enc2(&mw, Wp, IN))) return rc; } else mw = mu;
    if (Wl) { if ((rc = enc2(&mwl, Wl, IN))) return rc; } else mwl = mu;
    if (r->return_sequences) {
        // per-warp pieces of h_t: FUSED [32 rows][16 floats] (SWIZZLE_64B), XW [32 rows][8 floats] (unswizzled)
        cuuint64_t dims[3] = {(cuuint64_t)H, (cuuint64_t)T, (cuuint64_t)nb};
        cuuint64_t strides[2] = {(cuuint64_t)H * 4, (cuuint64_t)T * H * 4};
        const bool wide = fused && RF_FUSED_EPW == 8 && !dbx;
        cuuint32_t box[3] = {(cuuint32_t)(wide ? 16 : 8), 1, 32};
        cuuint32_t es[3] = {1, 1, 1};
        CUresult cr = encode(&mo, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, out, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             wide ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) { nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_fused: output tensor map rejected (CUresult %d)", (int)cr); return NNTK_ERR_CUDA_; }
    } else mo = mu;
    p.B = nb; p.T = T; p.return_sequences = r->return_sequences;
    p.x = x; p.bias_i = r->d_bi;
    p.bias_h = ((G == 3) || r->v2) ? r->d_bh : nullptr;
    p.xw = xw; p.xw_blocked_rows = (nb + 31) / 32;
    static int adaptive = -1;
    static float adapt_thr = 6.5f;
    if (adaptive < 0) {
        const char *e = getenv("NNTK_RECURRENT_X_ADAPTIVE"); adaptive = e ? atoi(e) : 1;
        if ((e = getenv("NNTK_RECURRENT_X_ADAPT_THR"))) adapt_thr = (float)atof(e);
    }
    if (!g_rf_stats) NNTK_CUDA_TRY(cudaMalloc(&g_rf_stats, 4 * sizeof(float)));
    if (!spec) NNTK_CUDA_TRY(cudaMemsetAsync(g_rf_stats, 0, 4 * sizeof(float), st));
    p.stats = g_rf_stats; p.adapt_thr = adapt_thr;
    static int per_step = -1;
    if (per_step < 0) { const char *e = getenv("NNTK_RECURRENT_X_PERSTEP"); per_step = e ? atoi(e) : 1; }
    p.per_step = 0;
    p.sel = spec; p.verify = spec == 1;
    { static int la = -1; if (la < 0) { const char *e = getenv("NNTK_RECURRENT_FUSED_LOOKAHEAD"); la = e ? atoi(e) : 1; } p.lookahead = la; }
    p.h_init = h_init ? h_init : h; p.c_init = c_init ? c_init : c;
    if (x3) {
        p.stats = nullptr; p.adapt_thr = -1.0f; p.sel = 0; p.verify = 0;          // always all three passes, nothing adaptive
    } else if (spec) {
        p.per_step = spec == 2;           // stats prepared by the caller (sampled row norms, W norms, cleared flag)
    } else if (fused && xp == 3 && adaptive && per_step) {
        // the decision moves into the kernel: per pair and per step, from the row norms its converter warps see anyway
        rf_wnorm_kernel<<<(G * H + 255) / 256, 256, 0, st>>>(r->d_W, IN, G * H, g_rf_stats);
        NNTK_LAUNCH_CHECK("rf_wnorm_kernel");
        p.per_step = 1;
    } else if (fused && xp == 3 && adaptive && (IN & 3) == 0) {
        const long long rows = (long long)nb * T;
        rf_xnorm_kernel<<<148 * 8, 256, 0, st>>>(x, rows, IN, g_rf_stats, 1);
        rf_wnorm_kernel<<<(G * H + 255) / 256, 256, 0, st>>>(r->d_W, IN, G * H, g_rf_stats);
        NNTK_LAUNCH_CHECK_N("rf_xnorm_kernel", 2);
    } else if (fused && xp == 3) {
        p.adapt_thr = -1.0f;          // never satisfied: all three passes
    }
    p.h_state = h; p.c_state = c; p.out = out;
    static int want_dbg = -1;
    if (want_dbg < 0) { const char *e = getenv("NNTK_RF_DBG"); want_dbg = e ? atoi(e) : 0; }
    long long *d_dbg = nullptr;
    if (want_dbg) { NNTK_CUDA_TRY(cudaMalloc(&d_dbg, (size_t)(pairs + 4) * 2 * 16 * 8)); NNTK_CUDA_TRY(cudaMemsetAsync(d_dbg, 0, (size_t)(pairs + 4) * 2 * 16 * 8, st)); }
    p.dbg = d_dbg;
    { const char *e = getenv("NNTK_RF_SKIP"); p.skip = e ? atoi(e) : 0; }
    rc = RF_DISPATCH(nullptr);
#undef RF_DISPATCH
    if (d_dbg) {
        static const char *names[14] = {"mma:ring_full", "mma:acc_empty", "mma:h_full", "mma:x_full", "epi:acc_full", "epi:total", "conv:convert", "conv:x_empty",
                                        "prod:ring_empty", "prod:tma_issue", "mma:total", "epi:store_wait", "mma:commit", "mma:issue"};
        std::vector<long long> hst((size_t)pairs * 2 * 16);
        cudaStreamSynchronize(st);
        cudaMemcpy(hst.data(), d_dbg, hst.size() * 8, cudaMemcpyDeviceToHost);
        cudaFree(d_dbg);
        for (int i = 0; i < 14; ++i) {
            double a0 = 0, a1 = 0;
            for (int b = 0; b < pairs; ++b) { a0 += (double)hst[(size_t)(2 * b) * 16 + i]; a1 += (double)hst[(size_t)(2 * b + 1) * 16 + i]; }
            fprintf(stderr, "[rf dbg] %-16s leader %10.0f  peer %10.0f cycles per step (avg over %d pairs; xp %d, ring %d, cluster %d)\n", names[i], a0 / pairs / T, a1 / pairs / T, pairs, p.xp, p.ring_stages, p.cs);
        }
    }
    if (rc == 1) return 0;
    if (rc) return rc;
    *used = 1;
    return 0;
}

// Runs all T steps of one batch chunk.  xw: the blocked hoisted projection (XW mode) or null (FUSED mode).
// FUSED mode, adaptive input passes: a SPECULATIVE PAIR of launches.  Staging everything for three passes and skipping the
// correction passes' MMAs per step (the per-step kernel) still streams W_lo through the ring: measured 8.5 ms per C5 GRU layer
// against 7.4 ms for a kernel planned for one pass.  So: (1) a sampled pre-scan (every 16th row of x: 0.1 ms instead of 1.2)
// gives an optimistic bound; (2) the ONE-PASS kernel runs iff that bound holds, checks every step's true row norms (its
// converters see them anyway) and raises a flag when one breaks the bound; (3) the three-pass / per-step kernel is enqueued
// right behind and runs iff the sampled bound failed or the flag was raised, from a copy of the initial state -- an empty launch
// in the normal case.  All decisions are taken on the device; nothing synchronises.
extern "C" int nntk_cuda_recurrent_fused(const nntk_recurrent_args *r, int nb, const float *x, const float *xw, float *h, float *c,
                                         float *out, int *used) {
    static int speculate = -1, adaptive = -1, per_step = -1;
    if (speculate < 0) {
        const char *e = getenv("NNTK_RECURRENT_X_SPECULATE"); speculate = e ? atoi(e) : 1;
        e = getenv("NNTK_RECURRENT_X_ADAPTIVE"); adaptive = e ? atoi(e) : 1;
        e = getenv("NNTK_RECURRENT_X_PERSTEP"); per_step = e ? atoi(e) : 1;
    }
    const int G = r->cell, H = r->H, T = r->T, IN = r->in;
    if (nntk_cuda_math_mode() == 2) return xw ? 0 : rf_fused_run(r, nb, x, nullptr, h, c, out, used, 0, nullptr, nullptr, 1);
    const bool fused = !xw && nntk_cuda_recurrent_fused_plan(r, nb) == 1;
    if (!(fused && speculate && adaptive && per_step && rf_xpasses() == 3 && (IN & 3) == 0 && (G != 4 || c)))
        return rf_fused_run(r, nb, x, xw, h, c, out, used, 0, nullptr, nullptr);
    cudaStream_t st = nntk_stream();
    const size_t sn = (size_t)nb * H, need = sn * (G == 4 ? 2 : 1);
    if (need > g_rf_state_floats) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_rf_state) cudaFree(g_rf_state);
        g_rf_state = nullptr; g_rf_state_floats = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_rf_state, need * sizeof(float)));
        g_rf_state_floats = need;
    }
    if (!g_rf_stats) NNTK_CUDA_TRY(cudaMalloc(&g_rf_stats, 4 * sizeof(float)));
    NNTK_CUDA_TRY(cudaMemsetAsync(g_rf_stats, 0, 4 * sizeof(float), st));
    NNTK_CUDA_TRY(cudaMemcpyAsync(g_rf_state, h, sn * sizeof(float), cudaMemcpyDeviceToDevice, st));
    if (G == 4) NNTK_CUDA_TRY(cudaMemcpyAsync(g_rf_state + sn, c, sn * sizeof(float), cudaMemcpyDeviceToDevice, st));
    const long long rows = (long long)nb * T;
    rf_xnorm_kernel<<<148 * 4, 256, 0, st>>>(x, rows, IN, g_rf_stats, 16);
    rf_wnorm_kernel<<<(G * H + 255) / 256, 256, 0, st>>>(r->d_W, IN, G * H, g_rf_stats);
    NNTK_LAUNCH_CHECK_N("rf_xnorm_kernel", 2);
    int used1 = 0;
    int rc = rf_fused_run(r, nb, x, nullptr, h, c, out, &used1, 1, g_rf_state, G == 4 ? g_rf_state + sn : nullptr);
    if (rc) return rc;
    if (!used1) return rf_fused_run(r, nb, x, xw, h, c, out, used, 0, nullptr, nullptr);     // nothing was launched: the plain path
    return rf_fused_run(r, nb, x, nullptr, h, c, out, used, 2, g_rf_state, G == 4 ? g_rf_state + sn : nullptr);
}
