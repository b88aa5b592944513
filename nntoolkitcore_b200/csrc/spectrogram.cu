// spectrogram.cu -- K1: batched STFT front-end (replaces signal/spectrogram.c:113-135 + signal/dft.c:34-47 +
// third_party/kissfft/kiss_fft.c:236-300 of the reference for the inference path).
//
// Fast path (nfft == 512): a PERSISTENT grid (2 CTAs per SM) walks tiles of F = 32 consecutive frames of one clip.
//   * The clip span a tile needs, (F-1)*step + window_size samples, is read from HBM once with 16-byte cp.async
//     (LDGSTS) into one of two shared-memory buffers; the copy of tile i+1 is in flight while tile i is transformed.
//     The frame overlap (hop 160 vs window 512 => every sample belongs to 3.2 frames) is served from shared memory.
//   * One warp transforms TWO real frames as one 512-point complex FFT held in registers (csrc/spectrogram_core.h):
//     window (per-lane registers) multiplied on load, three radix-8 passes, two shared-memory exchanges whose addresses
//     are all "per-lane base + immediate", the two frames separated IN REGISTERS (no third exchange), magnitude / PSD
//     scaling applied there.
//   * Each lane writes its bins straight to HBM: per store instruction the warp covers 32 consecutive floats of an
//     output row (128 bytes), so no output staging buffer and no second pass over shared memory is needed
//     (rows are 257 floats, never 16-byte aligned -- L2 merges the partial sectors of neighbouring stores).
//
// Generic path (any other nfft): one CTA per frame, shared-memory radix-2 FFT when nfft is a power of two, direct
// O(nfft^2) DFT from the twiddle table otherwise.  Correct for every nfft the reference accepts, not tuned.
//
// HBM roofline: algorithmic bytes per clip = 4*input_size + 4*ntime*nfreq.
#include <cuda_fp16.h>
#include "common.cuh"
#define NNTK_SPEC_PACKED 1   // complex adds / subtracts of the butterflies as sm_100 FADD2 / FFMA2 (spectrogram_core.h)
#include "spectrogram_core.h"

namespace {

constexpr int SPEC_THREADS = 256, SPEC_WARPS = SPEC_THREADS / 32, SPEC_FW = 4;   // frames per warp tile

struct SpecParams {
    int window_size, step, input_size, ntime, nfreq;
    int frames_per_tile, tiles_per_clip;
    long long total_tiles;
    int n_clips;
    int in_cap;         // floats per input buffer
    float mag_mul;      // magnitude: |norm| * 0.5 / scale            (the split pass yields 2|X|)
    float psd_mid;      // PSD bins 1..nfreq-2: norm^2 * 0.25 * 2/scale (the split pass yields 4|X|^2)
    float psd_edge;     // PSD bins 0, nfreq-1: norm^2 * 0.25 / scale
    int mode;           // 0 magnitude, 1 PSD (generic kernel only; the 512 kernel is templated on it)
    int in_vec_ok;      // input base pointer 16-byte aligned
    int out_pitch;      // floats between output rows (nfreq, or more: the pad columns are written as zeros -- fused front end)
    int n_mels, mel_iter;   // log-mel epilogue (MEL kernels only): bands, work-list entries per lane
    float log_offset;
    int half_out;       // 1: rows of out_pitch fp16 values (the tensor-core front end's conv operand) instead of floats
    int *ovf;           // half_out: set to 1 when a value does not fit fp16 (|v| > 65504 or Inf) -- the caller's fp32 fallback then runs
    const int *gate;    // when set the kernel runs only if *gate == gate_run_if (the fp32 fallback of the fp16 front end)
    int gate_run_if;
};

__device__ __forceinline__ float sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async4(void *smem, const void *gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ uint32_t smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "SPEC_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra SPEC_DONE;\n\t"
        "bra SPEC_WAIT;\n\t"
        "SPEC_DONE:\n\t}" ::"r"(smem_addr(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem, const void *gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_addr(smem)), "l"(gmem), "r"(bytes), "r"(smem_addr(bar)) : "memory");
}

// A tile = frames [f0, f0 + nf) of one clip.  Geometry is advanced incrementally (no divisions in the loop).
struct TileGeom {
    int clip, t;   // clip index, tile index inside the clip
};
__device__ __forceinline__ void tile_advance(const SpecParams &p, TileGeom &g, int dc, int dt) {
    g.clip += dc;
    g.t += dt;
    if (g.t >= p.tiles_per_clip) { g.t -= p.tiles_per_clip; ++g.clip; }
}
// Clip span of a tile -> the warp's shared-memory buffer (smem index s <-> global gi0 - ia + s), asynchronously.
//   16-byte aligned input: lane 0 issues ONE bulk copy (TMA, completion on `bar`) for the whole 16-byte vectors and copies
//   the <= 3 trailing floats itself before arming the barrier (its arrive releases those stores).
//   otherwise: the 32 lanes issue 4-byte cp.async (completion by wait_group + __syncwarp).
__device__ __forceinline__ void tile_load_async(const SpecParams &p, const TileGeom &g, const float *__restrict__ in,
                                                float *s_buf, uint64_t *bar, int lane) {
    const int f0 = g.t * p.frames_per_tile;
    const int nf = min(p.frames_per_tile, p.ntime - f0);
    const long long gi0 = (long long)g.clip * p.input_size + (long long)f0 * p.step;
    const int span = (nf - 1) * p.step + p.window_size;
    if (p.in_vec_ok) {
        if (lane == 0) {
            const int ia = (int)(gi0 & 3);
            const float *gbase = in + (gi0 - ia);
            const int total = ia + span, nvec = total >> 2;
            for (int s = 4 * nvec; s < total; ++s) s_buf[s] = __ldg(gbase + s);
            mbar_arrive_expect_tx(bar, 16u * nvec);
            bulk_g2s(s_buf, gbase, 16u * nvec, bar);
        }
    } else {
        const float *gbase = in + gi0;
        for (int s = lane; s < span; s += 32) cp_async4(s_buf + s, gbase + s);
    }
}

// Every WARP is an independent pipeline: it owns a double-buffered clip-span buffer, two mbarriers and an exchange
// buffer, walks tiles (gw, gw + n_warps, ...) of SPEC_FW frames and never meets a CTA-wide barrier, so the warps of an SM
// drift apart and one warp's shared-memory exchange overlaps another's butterflies.
// Shared memory layout: [SPEC_WARPS][2][in_cap] floats | SPEC_WARPS x 656 float2 | [SPEC_WARPS][2] mbarriers.
// MEL = true (row f1, signal/log_mel_spectrogram.c:33-38): instead of storing the 257 bins, the warp parks the two frames'
// spectra in its (idle) exchange buffer as (a, b) pairs, every lane walks its share of the mel table's non-zeros
// (`mel`: [mel_iter][32] entries {bin | last << 9 | band << 10, weight}, balanced on the host), emits
// log(sum + log_offset) per band, and the warp stores 2 x n_mels contiguous floats.  The spectrogram never reaches HBM.
// HOP5 = step / 32 when the hop is 160 (the BASELINE framing), else 0: lane L reads x[L + 32 j], so with a hop of 5 * 32 the
// second frame's element j IS the first frame's element j + 5 -- 21 shared-memory loads per frame pair instead of 32 (the
// kernel is LSU-wavefront bound: 180 wavefronts per pair, ncu r1).
// WIN: 0 = the window size is a run-time value; otherwise it is this compile-time constant (512 = the full transform; 400 = the
// BASELINE KWS framing).  With a run-time window every element of every frame pays a compare + select (ncu on the C5 framing:
// FSEL 11 % + ISETP 8 % of all instructions); as a constant all but one of the 16 slots resolve at compile time.
template <int WIN, int MODE, bool MEL, int HOP5>
__global__ void __launch_bounds__(SPEC_THREADS, 2)
spectrogram512_kernel(SpecParams p, const float *__restrict__ window, const float2 *__restrict__ tw512,
                      const float *__restrict__ in, float *__restrict__ out, const uint2 *__restrict__ mel) {
    extern __shared__ __align__(16) float smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float *s_in = smem + warp * 2 * p.in_cap;
    float2 *s_x = reinterpret_cast<float2 *>(smem + SPEC_WARPS * 2 * p.in_cap);
    uint64_t *full = reinterpret_cast<uint64_t *>(s_x + SPEC_WARPS * NNTK_SPEC_SMEM_F2) + 2 * warp;

    const int n_warps = gridDim.x * SPEC_WARPS, gw = blockIdx.x * SPEC_WARPS + warp;
    const int dc = n_warps / p.tiles_per_clip, dt = n_warps - dc * p.tiles_per_clip;
    TileGeom g;
    g.clip = gw / p.tiles_per_clip;
    g.t = gw - g.clip * p.tiles_per_clip;
    if (g.clip >= p.n_clips) return;
    if (p.gate && __ldg(p.gate) != p.gate_run_if) return;
    if (lane == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    tile_load_async(p, g, in, s_in, &full[0], lane);
    cp_async_commit();

    // per-lane constants (registers), fetched while the first tile is in flight.  The output scale is folded into the
    // window: |FFT(m w x)| = m |FFT(w x)| (magnitude), |FFT(sqrt(m) w x)|^2 = m |FFT(w x)|^2 (PSD).
    // (A negative scale -- only reachable through SpectrogramSetScaleFactor -- flips the sign afterwards.)
    const int WS = WIN ? WIN : p.window_size;
    SpecLaneConst lc;
    spec_lane_const(lane, tw512, lc);
    const float wscale = MODE == 0 ? fabsf(p.mag_mul) : sqrtf(fabsf(p.psd_mid));
    const bool neg = (MODE == 0 ? p.mag_mul : p.psd_mid) < 0.0f;
    float w[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const int n = lane + 32 * j;
        w[j] = (n < WS) ? __ldg(window + n) * wscale : 0.0f;
    }
    float2 *sx = s_x + warp * NNTK_SPEC_SMEM_F2;
    float2 *x1_st = sx + spec_x1_st_base(lane);
    const float2 *x1_ld = sx + spec_x1_ld_base(lane);
    float2 *x2_st = sx + spec_x2_st_base(lane);
    const float2 *x2_ld0 = sx + spec_x2_ld_base(lane, 0);
    const float2 *x2_ld1 = sx + spec_x2_ld_base(lane, 1);
    const int c1 = spec_lane_c(lane, 0), c2 = spec_lane_c(lane, 1);
    const bool l0 = lane == 0;
    // fp16 rows: a lane holds bins 64 j + c1 and 64 j + c2 -- never two neighbours.  For every pair of slots (A, B) = (j, j + 1) of
    // the same half s one shuffle with the lane that holds the neighbouring bin leaves (A[i], A[i+1]) in the lane of an even local
    // index i and (B[i-1], B[i]) in the lane of an odd one: one 4-byte store per lane, the warp writes 64 + 64 contiguous bytes.
    const int hi0 = c1, hi1 = l0 ? 0 : 32 - c1;                     // local index of this lane's bin inside the 32-bin group, s = 0 / 1
    const int hp0 = spec_lane_c1(hi0 ^ 1);                          // lane holding local index i ^ 1 (spec_lane_c1 is an involution)
    const int hp1 = (hi1 ^ 1) == 0 ? 0 : spec_lane_c1(32 - (hi1 ^ 1));
    float hmax = 0.0f;                                              // largest |value| this lane stored as fp16
    // PSD: bins 0 and nfreq-1 are not doubled (spectrogram.c:41-47): psd_edge / psd_mid = 0.5
    const float edge = 0.5f;

    for (int it = 0;; ++it) {
        __syncwarp();   // every lane is done reading the other buffer
        TileGeom gn = g;
        tile_advance(p, gn, dc, dt);
        const bool more = gn.clip < p.n_clips;
        if (more) tile_load_async(p, gn, in, s_in + ((it + 1) & 1) * p.in_cap, &full[(it + 1) & 1], lane);
        if (p.in_vec_ok) {
            mbar_wait(&full[it & 1], (it >> 1) & 1);
        } else {
            cp_async_commit();
            asm volatile("cp.async.wait_group 1;" ::: "memory");
            __syncwarp();
        }

        const int f0 = g.t * p.frames_per_tile;
        const int nf = min(p.frames_per_tile, p.ntime - f0);
        const long long gi0 = (long long)g.clip * p.input_size + (long long)f0 * p.step;
        const int ia = p.in_vec_ok ? (int)(gi0 & 3) : 0;
        const float *buf = s_in + (it & 1) * p.in_cap + ia + lane;
        const int pitch = p.out_pitch;
        float *orow = out + ((long long)g.clip * p.ntime + f0) * (long long)pitch;
#pragma unroll 1
        for (int fa = 0; fa < nf; fa += 2) {
            const bool has_b = fa + 1 < nf;
            const float *xa = buf + fa * p.step;
            const float *xb = has_b ? xa + p.step : xa;
            float2 v[16];
            if (HOP5 > 0) {
                float r[16 + HOP5];
#pragma unroll
                for (int k = 0; k < 16 + HOP5; ++k) {
                    const bool as_a = k < 16 && (lane + 32 * k < WS);
                    const bool as_b = k >= HOP5 && has_b && (lane + 32 * (k - HOP5) < WS);
                    r[k] = (as_a || as_b) ? xa[32 * k] : 0.0f;
                }
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const bool in_win = lane + 32 * j < WS;
                    const float a = in_win ? r[j] : 0.0f;
                    const float b = has_b ? (in_win ? r[j + HOP5] : 0.0f) : a;
                    v[j] = make_float2(w[j] * a, w[j] * b);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    float a = 0.0f, b = 0.0f;
                    if (lane + 32 * j < WS) {
                        a = xa[32 * j];
                        b = xb[32 * j];
                    }
                    v[j] = make_float2(w[j] * a, w[j] * b);
                }
            }
            spec_pass1(v, lc);
            spec_xchg1_store(v, x1_st);
            __syncwarp();
            spec_xchg1_load(v, x1_ld);
            __syncwarp();
            spec_pass2(v, lc);
            spec_xchg2_store(v, x2_st);
            __syncwarp();
            spec_xchg2_load(v, x2_ld0, x2_ld1);
            __syncwarp();
            spec_pass3(v);
            // split the packed pair in registers and store: slot j -> bin 64 j + c1, slot 4 + j -> bin 64 j + c2
            float pa[9], pb[9];
            spec_split_regs(lane, v, pa, pb);
            float *ra = orow + (long long)fa * pitch;
            float *ra1 = ra + c1, *ra2 = ra + c2;
            if (MODE == 0) {
#pragma unroll
                for (int slot = 0; slot < 9; ++slot) { pa[slot] = sqrt_approx(pa[slot]); pb[slot] = sqrt_approx(pb[slot]); }
            } else {
                pa[8] *= edge; pb[8] *= edge;
                if (l0) { pa[0] *= edge; pb[0] *= edge; }
            }
            if (neg) {
#pragma unroll
                for (int slot = 0; slot < 9; ++slot) { pa[slot] = -pa[slot]; pb[slot] = -pb[slot]; }
            }
            if (MEL) {
                // the exchange buffer is idle here (its last readers passed the __syncwarp after spec_xchg2_load)
                // parked spectra: bin k at slot k + k / 16 as (frame a, frame b) -- with the skew the lanes' bins c + 64 j hit
                // 16 different bank pairs per half-warp (c1 is the identity up to the swaps 1 <-> 16, 9 <-> 24)
                float2 *mg = sx;
                float2 *so = sx + 280;                                  // [n_mels] raw band sums (a, b)
#pragma unroll
                for (int slot = 0; slot < 8; ++slot) {
                    const int k = 64 * (slot & 3) + (slot < 4 ? c1 : c2);
                    mg[k + (k >> 4)] = make_float2(pa[slot], pb[slot]);
                }
                if (l0) mg[256 + 16] = make_float2(pa[8], pb[8]);
                __syncwarp();
                float acc_a = 0.0f, acc_b = 0.0f;
                const uint2 *e = mel + lane;
                const char *mgb = reinterpret_cast<const char *>(mg);
                char *sob = reinterpret_cast<char *>(so);
                // The walk only accumulates and parks raw band sums (a flagged entry is one predicated store): the lanes'
                // bands end at different iterations, so a logf inside the loop would run -- for the whole warp -- on nearly
                // every iteration (first version: 31 iterations cost 2.3x the FFT itself).  The log is taken in the store loop.
#pragma unroll 4
                for (int i = 0; i < p.mel_iter; ++i) {
                    const uint2 en = __ldg(e + 32 * i);
                    const float2 m = *reinterpret_cast<const float2 *>(mgb + (en.x & 0xff8u));
                    acc_a = fmaf(__uint_as_float(en.y), m.x, acc_a);
                    acc_b = fmaf(__uint_as_float(en.y), m.y, acc_b);
                    if ((int)en.x < 0) {
                        *reinterpret_cast<float2 *>(sob + ((en.x >> 12) & 0x7fff8u)) = make_float2(acc_a, acc_b);
                        acc_a = acc_b = 0.0f;
                    }
                }
                __syncwarp();
                float *dst = out + ((long long)g.clip * p.ntime + f0 + fa) * p.n_mels;
                for (int i = lane; i < p.n_mels; i += 32) {
                    const float2 v = so[i];
                    // MUFU log2 (2 instructions; absolute error < 6e-6 on |log| <= 30, inside the 1e-5 contract) -- logf's
                    // ~22-instruction expansion was 200 of the epilogue's ~550 instructions per frame pair
                    dst[i] = __logf(v.x + p.log_offset);
                    if (has_b) dst[p.n_mels + i] = __logf(v.y + p.log_offset);
                }
                __syncwarp();   // before the next pair's exchange-1 stores overwrite mg / so
                continue;
            }
            if (p.half_out) {
                __half *ha = reinterpret_cast<__half *>(out) + ((long long)g.clip * p.ntime + f0 + fa) * (long long)pitch;
#pragma unroll
                for (int s2 = 0; s2 < 2; ++s2) {
                    const int li = s2 ? hi1 : hi0, lp = s2 ? hp1 : hp0;
                    const bool ev = !(li & 1);
#pragma unroll
                    for (int j = 0; j < 4; j += 2) {
                        const int A = 4 * s2 + j, B = A + 1;
                        const float ga = __shfl_sync(0xffffffffu, ev ? pa[B] : pa[A], lp);
                        const float gb = __shfl_sync(0xffffffffu, ev ? pb[B] : pb[A], lp);
                        const float a0 = ev ? pa[A] : ga, a1 = ev ? ga : pa[B];       // two neighbouring bins of frame a
                        const float b0 = ev ? pb[A] : gb, b1 = ev ? gb : pb[B];       // ... of frame b
                        hmax = fmaxf(hmax, fmaxf(fmaxf(fabsf(a0), fabsf(a1)), fmaxf(fabsf(b0), fabsf(b1))));
                        __half *dst = ha + 64 * (ev ? j : j + 1) + 32 * s2 + (li & ~1);
                        *reinterpret_cast<__half2 *>(dst) = __floats2half2_rn(a0, a1);
                        if (has_b) *reinterpret_cast<__half2 *>(dst + pitch) = __floats2half2_rn(b0, b1);
                    }
                }
                if (256 + 2 * lane < pitch) {        // bin 256 and the zero pad
                    if (l0) hmax = fmaxf(hmax, fmaxf(fabsf(pa[8]), fabsf(pb[8])));
                    *reinterpret_cast<__half2 *>(ha + 256 + 2 * lane) = __floats2half2_rn(l0 ? pa[8] : 0.0f, 0.0f);
                    if (has_b) *reinterpret_cast<__half2 *>(ha + pitch + 256 + 2 * lane) = __floats2half2_rn(l0 ? pb[8] : 0.0f, 0.0f);
                }
                continue;
            }
#pragma unroll
            for (int slot = 0; slot < 8; ++slot) {
                float *dst = (slot < 4 ? ra1 : ra2) + 64 * (slot & 3);
                dst[0] = pa[slot];
                if (has_b) dst[pitch] = pb[slot];
            }
            if (pitch == 257) {
                if (l0) {   // bin 256
                    ra[256] = pa[8];
                    if (has_b) ra[256 + 257] = pb[8];
                }
            } else if (256 + lane < pitch) {   // bin 256 and the zero pad columns up to the pitch (<= 288) in one warp store
                ra[256 + lane] = l0 ? pa[8] : 0.0f;
                if (has_b) ra[256 + lane + pitch] = l0 ? pb[8] : 0.0f;
            }
        }
        if (!more) break;
        g = gn;
    }
    if (p.half_out && p.ovf && !(hmax <= 65504.0f)) *p.ovf = 1;
}

// ---- generic nfft -------------------------------------------------------------------------------------------------
// One CTA per frame.  s_a / s_b: ping-pong complex buffers of nfft points.
__global__ void __launch_bounds__(128)
spectrogram_generic_kernel(SpecParams p, int nfft, int log2n, const float *__restrict__ window,
                           const float2 *__restrict__ tw, const float *__restrict__ in, float *__restrict__ out) {
    extern __shared__ __align__(16) float smem[];
    float2 *s_a = reinterpret_cast<float2 *>(smem);
    float2 *s_b = s_a + nfft;
    const int tid = threadIdx.x;
    const long long frame = blockIdx.x;
    const long long clip = frame / p.ntime;
    const int t = (int)(frame - clip * p.ntime);
    const float *x = in + clip * (long long)p.input_size + (long long)t * p.step;
    float *o = out + frame * (long long)p.nfreq;

    if (log2n >= 0) {
        // bit-reversed load, then log2n in-place radix-2 decimation-in-time stages
        for (int n = tid; n < nfft; n += blockDim.x) {
            float val = (n < p.window_size) ? __ldg(window + n) * __ldg(x + n) : 0.0f;
            int r = (log2n > 0) ? (int)(__brev((unsigned)n) >> (32 - log2n)) : 0;
            s_a[r] = make_float2(val, 0.0f);
        }
        __syncthreads();
        for (int s = 0; s < log2n; ++s) {
            const int half = 1 << s;
            const int tstride = nfft >> (s + 1);
            for (int b = tid; b < (nfft >> 1); b += blockDim.x) {
                const int j = b & (half - 1);
                const int i0 = ((b >> s) << (s + 1)) + j;
                const int i1 = i0 + half;
                float2 w = tw[j * tstride];
                float2 u = s_a[i0], y = c_mul(s_a[i1], w);
                s_a[i0] = c_add(u, y);
                s_a[i1] = c_sub(u, y);
            }
            __syncthreads();
        }
    } else {
        for (int n = tid; n < nfft; n += blockDim.x) {
            float val = (n < p.window_size) ? __ldg(window + n) * __ldg(x + n) : 0.0f;
            s_b[n] = make_float2(val, 0.0f);
        }
        __syncthreads();
        for (int k = tid; k < p.nfreq; k += blockDim.x) {
            float re = 0.0f, im = 0.0f;
            int idx = 0;
            for (int n = 0; n < nfft; ++n) {
                float2 w = tw[idx];
                float xv = s_b[n].x;
                re = fmaf(xv, w.x, re);
                im = fmaf(xv, w.y, im);
                idx += k;
                if (idx >= nfft) idx -= nfft;
            }
            s_a[k] = make_float2(re, im);
        }
        __syncthreads();
    }
    for (int k = tid; k < p.nfreq; k += blockDim.x) {
        float2 z = s_a[k];
        float pw = z.x * z.x + z.y * z.y;   // |X|^2; the scale constants below assume the 4x of the packed split
        float r;
        if (p.mode == 0) r = sqrtf(pw) * (2.0f * p.mag_mul);
        else r = pw * 4.0f * ((k == 0 || k == p.nfreq - 1) ? p.psd_edge : p.psd_mid);
        o[k] = r;
    }
}

}  // namespace

static int spectrogram_run(const nntk_spectrogram_args *a, const nntk_mel_args *m, const float *d_in, float *d_out, int n_clips, int out_pitch = 0,
                           int half_out = 0, int *d_ovf = nullptr);

// Gate for the NEXT 512-point launches (until cleared with a null flag): the kernel returns at once unless *flag == run_if.  Used
// by the fp16 front end (host/conv_1d.c): its fp32 fallback is enqueued behind the fp16 kernels and only does work when they
// reported a value outside fp16's range.  Set and cleared under the API lock by the one caller.
static const int *g_spec_gate = nullptr;
static int g_spec_gate_run_if = 0;
extern "C" void nntk_cuda_spectrogram_gate(const int *d_flag, int run_if) { g_spec_gate = d_flag; g_spec_gate_run_if = run_if; }

// The front end's overflow flag: one int per device, in the library's own image (no allocation to manage); host code needs its
// device address, looked up once per device.
static __device__ int g_front_flag;
extern "C" int *nntk_cuda_front_flag(void) {
    static int *addr[64] = {nullptr};
    int dev = -1;
    if (nntk_cuda_ensure_init() || cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    if (!addr[dev]) {
        int *ptr = nullptr;
        if (cudaGetSymbolAddress((void **)&ptr, g_front_flag) != cudaSuccess) { (void)cudaGetLastError(); return nullptr; }
        addr[dev] = ptr;
    }
    return addr[dev];
}

// fp16 rows of `out_pitch` halves (even, nfreq <= out_pitch <= 288): bins as cvt.rn fp16, pad columns zero.  *d_ovf is set to 1
// when some value does not fit fp16 (the caller then runs its fp32 path; the fp16 rows are garbage in that case).  Returns 1
// (nothing enqueued) for shapes without the 512-point kernel.
extern "C" int nntk_cuda_spectrogram_half(const nntk_spectrogram_args *a, const float *d_in, void *d_out_half, int n_clips, int out_pitch, int *d_ovf) {
    if (!a || a->nfft != NNTK_SPEC_N || out_pitch < a->nfreq + 1 || out_pitch > 288 || (out_pitch & 1) || !d_ovf) return 1;
    return spectrogram_run(a, nullptr, d_in, reinterpret_cast<float *>(d_out_half), n_clips, out_pitch, 1, d_ovf);
}

extern "C" int nntk_cuda_spectrogram(const nntk_spectrogram_args *a, const float *d_in, float *d_out, int n_clips) {
    return spectrogram_run(a, nullptr, d_in, d_out, n_clips);
}

// Rows of `out_pitch` >= nfreq floats, the pad columns written as zeros (fused front end: a 16- or 128-byte aligned row pitch
// is what the TMA im2col view of the following Conv1d needs).  Returns 1 (nothing enqueued) when only the dense layout exists
// for this shape (nfft != 512 or a pitch beyond 288).
extern "C" int nntk_cuda_spectrogram_pitched(const nntk_spectrogram_args *a, const float *d_in, float *d_out, int n_clips, int out_pitch) {
    if (!a || a->nfft != NNTK_SPEC_N || out_pitch < a->nfreq || out_pitch > 288) return 1;
    return spectrogram_run(a, nullptr, d_in, d_out, n_clips, out_pitch);
}

extern "C" int nntk_cuda_log_mel_spectrogram(const nntk_spectrogram_args *a, const nntk_mel_args *m, const float *d_in,
                                             float *d_out, int n_clips) {
    if (!m || !m->d_entries || m->n_mels < 1 || m->n_iter < 1) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "log_mel_spectrogram: bad mel arguments");
        return NNTK_ERR_ARG_;
    }
    // the staged band results of a frame pair share the warp's exchange buffer with the parked spectra
    if (!a || a->nfft != NNTK_SPEC_N || m->n_mels > NNTK_SPEC_SMEM_F2 - 280 || m->n_iter > 4096) return 1;
    return spectrogram_run(a, m, d_in, d_out, n_clips);
}

static int spectrogram_run(const nntk_spectrogram_args *a, const nntk_mel_args *m, const float *d_in, float *d_out, int n_clips, int out_pitch,
                           int half_out, int *d_ovf) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!a || !d_in || !d_out || n_clips < 0 || a->nfft < 1 || a->window_size < 1 || a->window_size > a->nfft ||
        a->step < 1 || a->nfreq != a->nfft / 2 + 1 || !a->d_window || !a->d_twiddle) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "spectrogram: bad arguments (nfft %d, window %d, step %d)",
                            a ? a->nfft : -1, a ? a->window_size : -1, a ? a->step : -1);
        return NNTK_ERR_ARG_;
    }
    if (n_clips == 0 || a->ntime <= 0) return 0;
    if ((long long)(a->ntime - 1) * a->step + a->window_size > a->input_size) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "spectrogram: %d frames do not fit input_size %d", a->ntime, a->input_size);
        return NNTK_ERR_ARG_;
    }
    SpecParams p;
    p.window_size = a->window_size;
    p.step = a->step;
    p.input_size = a->input_size;
    p.ntime = a->ntime;
    p.nfreq = a->nfreq;
    p.mode = a->mode;
    const float an = fabsf(a->norm_factor);
    p.mag_mul = an * 0.5f / a->scale;
    p.psd_mid = an * an * 0.25f * (2.0f / a->scale);
    p.psd_edge = an * an * 0.25f / a->scale;
    p.in_vec_ok = (((uintptr_t)d_in) & 15) == 0;
    p.out_pitch = out_pitch > 0 ? out_pitch : a->nfreq;
    p.n_mels = m ? m->n_mels : 0;
    p.mel_iter = m ? m->n_iter : 0;
    p.log_offset = m ? m->log_offset : 0.0f;
    p.half_out = half_out; p.ovf = d_ovf;
    p.gate = (a->nfft == NNTK_SPEC_N) ? g_spec_gate : nullptr; p.gate_run_if = g_spec_gate_run_if;
    cudaStream_t st = nntk_stream();

    if (a->nfft == NNTK_SPEC_N) {
        // frames per warp tile: 4, or 2 when the hop is so large that two CTAs no longer fit an SM
        const size_t xchg = (size_t)SPEC_WARPS * NNTK_SPEC_SMEM_F2 * sizeof(float2);
        int F = SPEC_FW;
        size_t bytes = 0;
        int in_cap = 0;
        for (;; F >>= 1) {
            in_cap = (((F - 1) * a->step + a->window_size + 3 + 3) / 4) * 4;
            bytes = (size_t)SPEC_WARPS * 2 * in_cap * sizeof(float) + xchg + SPEC_WARPS * 2 * sizeof(uint64_t);
            if (bytes <= 112 * 1024 || F == 2) break;
        }
        if (bytes > 227 * 1024) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "spectrogram: step %d too large for the tiled kernel", a->step);
            return NNTK_ERR_UNSUPPORTED_;
        }
        p.frames_per_tile = F;
        p.tiles_per_clip = (a->ntime + F - 1) / F;
        p.total_tiles = (long long)n_clips * p.tiles_per_clip;
        p.in_cap = in_cap;
        p.n_clips = n_clips;
        const bool fullwin = a->window_size == NNTK_SPEC_N;
        typedef void (*SpecKernel)(SpecParams, const float *, const float2 *, const float *, float *, const uint2 *);
        // [hop 160][mel][mode][window class: 0 run-time, 1 full (512), 2 the 400-sample window of the BASELINE framing (hop 160 only)]
#define SPEC_K3(MODE, MEL, HOP) spectrogram512_kernel<0, MODE, MEL, HOP>, spectrogram512_kernel<512, MODE, MEL, HOP>, \
                                spectrogram512_kernel<(HOP ? 400 : 0), MODE, MEL, HOP>
        static const SpecKernel kernels[24] = {
            SPEC_K3(0, false, 0), SPEC_K3(1, false, 0), SPEC_K3(0, true, 0), SPEC_K3(1, true, 0),
            SPEC_K3(0, false, 5), SPEC_K3(1, false, 5), SPEC_K3(0, true, 5), SPEC_K3(1, true, 5)};
#undef SPEC_K3
        static size_t attr_bytes[24] = {0};
        const bool hop160 = a->step == 160;
        const int wc = fullwin ? 1 : (hop160 && a->window_size == 400) ? 2 : 0;
        const int ki = ((hop160 ? 4 : 0) + (m ? 2 : 0) + (a->mode ? 1 : 0)) * 3 + wc;
        const SpecKernel kern = kernels[ki];
        if (bytes > attr_bytes[ki]) {
            NNTK_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
            attr_bytes[ki] = bytes;
        }
        // persistent: two CTAs per SM; warp gw walks tiles gw, gw + n_warps, ...
        if (p.total_tiles > 0x7fffffffLL) {
            nntk_cuda_set_error(NNTK_ERR_ARG_, "spectrogram: %d clips exceed one launch", n_clips);
            return NNTK_ERR_ARG_;
        }
        long long grid = 2LL * nntk_cuda_sm_count();
        const long long need = (p.total_tiles + SPEC_WARPS - 1) / SPEC_WARPS;
        if (grid > need) grid = need;
        kern<<<(unsigned)grid, SPEC_THREADS, bytes, st>>>(p, a->d_window, reinterpret_cast<const float2 *>(a->d_twiddle),
                                                          d_in, d_out, m ? reinterpret_cast<const uint2 *>(m->d_entries) : nullptr);
        NNTK_LAUNCH_CHECK("spectrogram512_kernel");
        return 0;
    }
    if (m) return 1;   // only the 512-point kernel has the mel epilogue
    int log2n = -1;
    if ((a->nfft & (a->nfft - 1)) == 0) {
        log2n = 0;
        while ((1 << log2n) < a->nfft) ++log2n;
    }
    size_t bytes = (size_t)2 * a->nfft * sizeof(float2);
    if (bytes > 200 * 1024) {
        nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "spectrogram: nfft %d too large", a->nfft);
        return NNTK_ERR_UNSUPPORTED_;
    }
    p.frames_per_tile = 1;
    p.tiles_per_clip = a->ntime;
    p.total_tiles = (long long)n_clips * a->ntime;
    p.in_cap = 0;
    p.n_clips = n_clips;
    static size_t attr_bytes_g = 48 * 1024;
    if (bytes > attr_bytes_g) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(spectrogram_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)bytes));
        attr_bytes_g = bytes;
    }
    long long grid = (long long)n_clips * a->ntime;
    if (grid > 0x7fffffffLL) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "spectrogram: too many frames for one launch");
        return NNTK_ERR_ARG_;
    }
    spectrogram_generic_kernel<<<(unsigned)grid, 128, bytes, st>>>(p, a->nfft, log2n, a->d_window,
                                                                  reinterpret_cast<const float2 *>(a->d_twiddle), d_in,
                                                                  d_out);
    NNTK_LAUNCH_CHECK("spectrogram_generic_kernel");
    return 0;
}
