// spectrogram_core.h -- the per-warp arithmetic of the K1 spectrogram kernel (nfft = 512), written once and used by
//   * csrc/spectrogram.cu                 (the sm_100a kernel; one warp = two frames), and
//   * tests/emu/spectrogram_emu.cpp       (a CPU emulation of the 32 lanes, used by the CPU test tier to check the
//                                           index maps / butterflies against the oracle, and to count shared-memory
//                                           bank conflicts of every exchange, before any GPU time is spent).
//
// Algorithm.  The reference runs a 512-point COMPLEX kissfft on each real frame with imag = 0
// (signal/spectrogram.c:120-128, signal/dft.c:34-47).  Here two consecutive real frames a, b are packed as one complex
// signal z = a + i*b, transformed by one 512-point complex FFT (three radix-8 decimation-in-frequency passes,
// 16 points per lane, data in registers, TWO shared-memory exchanges), and separated with the conjugate symmetry
//   A[k] = (Z[k] + conj(Z[N-k])) / 2,   B[k] = (Z[k] - conj(Z[N-k])) / (2i).
// Twiddles come from the host table built exactly like kiss_fft.c:357-363 (double cos/sin rounded to float).
//
// Index maps (N = 512, lane L = 0..31; shared-memory indices are in float2 units):
//   load      v[j]            = z[L + 32 j]                               j = 0..15
//   pass 1    butterfly s=0,1 : inputs v[s + 2m] = z[n' + 64 m], n' = L + 32 s;  y_q[n'] = (sum_m . W8^{mq}) W512^{n' q}
//   xchg 1    E1[q*64 + 2L + s]: the lane's two values of a q are ADJACENT -> 8 x 16-byte stores; the consumer's inputs
//             n'' + 8m and n'' + 8(m+4) are adjacent too -> 8 x 16-byte loads.  No padding, conflict-free.
//   pass 2    butterfly b = L + 32 s, q = b / 8, n'' = b % 8: inputs y_q[n'' + 8 m];
//             u_{q'}[n''] = (sum_m . W8^{m q'}) W64^{n'' q'}
//   xchg 2    E2[82 q + 10 n'' + q']: the lane's q' run is contiguous -> 8 x 16-byte stores (n'' stride of 80 bytes
//             spreads a quarter-warp over all 16-byte slots of a 128-byte line); the consumer reads 16 x 8 bytes
//             (q stride 82 = 2 mod 16 spreads a half-warp over the sixteen 8-byte bank pairs).  Padding only, so every
//             address is one per-lane base plus a compile-time immediate.
//   pass 3    butterfly c = q + 8 q' (lane L owns c1 = pi(L) and c2 = 64 - c1): Z[64 k'' + c] = sum_{n''} E2[c][n''] W8^{n'' k''}
//   split     in registers: bins 64 k'' + c and their mirrors 64 (7 - k'') + (64 - c) are in the same lane.
#pragma once

#ifdef __CUDACC__
#define NNTK_HD __host__ __device__ __forceinline__
#else
#define NNTK_HD inline
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { float2 r; r.x = x; r.y = y; return r; }
#endif

#define NNTK_SPEC_N 512
#define NNTK_SPEC_E2_Q 82    /* exchange-2 stride of q   (float2 units) */
#define NNTK_SPEC_E2_N 10    /* exchange-2 stride of n'' (float2 units) */
#define NNTK_SPEC_SMEM_F2 656 /* float2 slots per warp: exchange 1 uses 512, exchange 2 uses 82*7 + 10*7 + 8 = 652 */

#if defined(__CUDA_ARCH__) && defined(NNTK_SPEC_PACKED)
// sm_100 packed fp32: one FADD2 / FFMA2 issue slot per complex add / subtract (same IEEE results as the scalar forms)
NNTK_HD float2 c_add(float2 a, float2 b) { return __fadd2_rn(a, b); }
NNTK_HD float2 c_sub(float2 a, float2 b) { return __ffma2_rn(b, make_float2(-1.0f, -1.0f), a); }
#else
NNTK_HD float2 c_add(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
NNTK_HD float2 c_sub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
#endif
// (a.x + i a.y)(w.x + i w.y), same operand pairing as kissfft's C_MUL (_kiss_fft_guts.h)
NNTK_HD float2 c_mul(float2 a, float2 w) { return make_float2(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x); }
NNTK_HD float2 c_mul_mi(float2 a) { return make_float2(a.y, -a.x); }  // a * (-i)

// Two adjacent float2 slots moved as one 16-byte access on the device (idx must be even).  The emulator hooks
// NNTK_SPEC_TRACE to record addresses and widths for the bank-conflict count.
#ifndef NNTK_SPEC_TRACE
#define NNTK_SPEC_TRACE(kind, ptr, width)
#endif
NNTK_HD void sm_st2(float2 *sm, int idx, float2 a, float2 b) {
    NNTK_SPEC_TRACE(0, sm + idx, 16);
#ifdef __CUDA_ARCH__
    *reinterpret_cast<float4 *>(sm + idx) = make_float4(a.x, a.y, b.x, b.y);
#else
    sm[idx] = a; sm[idx + 1] = b;
#endif
}
NNTK_HD void sm_ld2(const float2 *sm, int idx, float2 &a, float2 &b) {
    NNTK_SPEC_TRACE(1, sm + idx, 16);
#ifdef __CUDA_ARCH__
    const float4 t = *reinterpret_cast<const float4 *>(sm + idx);
    a = make_float2(t.x, t.y); b = make_float2(t.z, t.w);
#else
    a = sm[idx]; b = sm[idx + 1];
#endif
}
NNTK_HD float2 sm_ld1(const float2 *sm, int idx) {
    NNTK_SPEC_TRACE(1, sm + idx, 8);
    return sm[idx];
}

// 8-point forward DFT, natural-order in and out: x[q] <- sum_m x[m] * exp(-2 pi i m q / 8).
// The sqrt(1/2) factors of W8^1 and W8^3 are folded into the last stage as FMAs.
NNTK_HD void radix8(float2 *x) {
    const float R = 0.70710678118654752440f;
    float2 a0 = c_add(x[0], x[4]), b0 = c_sub(x[0], x[4]);
    float2 a1 = c_add(x[1], x[5]), b1 = c_sub(x[1], x[5]);
    float2 a2 = c_add(x[2], x[6]), b2 = c_sub(x[2], x[6]);
    float2 a3 = c_add(x[3], x[7]), b3 = c_sub(x[3], x[7]);
    // even outputs: 4-point DFT of a
    float2 e0 = c_add(a0, a2), e1 = c_sub(a0, a2), o0 = c_add(a1, a3), o1 = c_mul_mi(c_sub(a1, a3));
    x[0] = c_add(e0, o0);
    x[4] = c_sub(e0, o0);
    x[2] = c_add(e1, o1);
    x[6] = c_sub(e1, o1);
    // odd outputs: 4-point DFT of (b0, b1 W8, b2 W8^2, b3 W8^3) with b1 W8 = R p1, b3 W8^3 = R p3
    b2 = c_mul_mi(b2);
    float2 p1 = make_float2(b1.x + b1.y, b1.y - b1.x);
    float2 p3 = make_float2(b3.y - b3.x, -(b3.x + b3.y));
    float2 f0 = c_add(b0, b2), f1 = c_sub(b0, b2);
    float2 g0 = c_add(p1, p3), g1 = c_mul_mi(c_sub(p1, p3));
#if defined(__CUDA_ARCH__) && defined(NNTK_SPEC_PACKED)
    const float2 RR = make_float2(R, R), NR = make_float2(-R, -R);
    x[1] = __ffma2_rn(RR, g0, f0);
    x[5] = __ffma2_rn(NR, g0, f0);
    x[3] = __ffma2_rn(RR, g1, f1);
    x[7] = __ffma2_rn(NR, g1, f1);
#else
    x[1] = make_float2(fmaf(R, g0.x, f0.x), fmaf(R, g0.y, f0.y));
    x[5] = make_float2(fmaf(-R, g0.x, f0.x), fmaf(-R, g0.y, f0.y));
    x[3] = make_float2(fmaf(R, g1.x, f1.x), fmaf(R, g1.y, f1.y));
    x[7] = make_float2(fmaf(-R, g1.x, f1.x), fmaf(-R, g1.y, f1.y));
#endif
}

// Per-lane constants, loaded once per CTA lifetime.
struct SpecLaneConst {
    float2 tw1[2][7];  // W512^{(L + 32 s) q}, q = 1..7
    float2 tw2[7];     // W64^{(L % 8) q'} = W512^{8 (L % 8) q'}, q' = 1..7
};

NNTK_HD void spec_lane_const(int lane, const float2 *tw512, SpecLaneConst &c) {
    for (int s = 0; s < 2; ++s)
        for (int q = 1; q < 8; ++q) c.tw1[s][q - 1] = tw512[((lane + 32 * s) * q) & 511];
    for (int q = 1; q < 8; ++q) c.tw2[q - 1] = tw512[(8 * (lane & 7) * q) & 511];
}

// pass 1 on v[16] (in place: butterfly s reads/writes v[s + 2m])
NNTK_HD void spec_pass1(float2 *v, const SpecLaneConst &c) {
    for (int s = 0; s < 2; ++s) {
        float2 x[8];
        for (int m = 0; m < 8; ++m) x[m] = v[s + 2 * m];
        radix8(x);
        v[s] = x[0];
        for (int q = 1; q < 8; ++q) v[s + 2 * q] = c_mul(x[q], c.tw1[s][q - 1]);
    }
}
// Exchange 1.  `sm_st` = sm + 2*lane and `sm_ld` = sm + (lane >> 3)*64 + 2*(lane & 7) are the per-lane bases; every
// index below is a compile-time constant.
NNTK_HD int spec_x1_st_base(int lane) { return 2 * lane; }
NNTK_HD int spec_x1_ld_base(int lane) { return (lane >> 3) * 64 + 2 * (lane & 7); }
NNTK_HD void spec_xchg1_store(const float2 *v, float2 *sm_st) {
    for (int q = 0; q < 8; ++q) sm_st2(sm_st, q * 64, v[2 * q], v[2 * q + 1]);
}
NNTK_HD void spec_xchg1_load(float2 *v, const float2 *sm_ld) {
    for (int s = 0; s < 2; ++s)
        for (int m = 0; m < 4; ++m) sm_ld2(sm_ld, 256 * s + 16 * m, v[8 * s + m], v[8 * s + m + 4]);
}
// pass 2 on v[8 s + m]
NNTK_HD void spec_pass2(float2 *v, const SpecLaneConst &c) {
    for (int s = 0; s < 2; ++s) {
        float2 *x = v + 8 * s;
        radix8(x);
        for (int q = 1; q < 8; ++q) x[q] = c_mul(x[q], c.tw2[q - 1]);
    }
}
// Exchange 2: element (q, q', n'') at E2[82 q + 10 n'' + q'].
NNTK_HD int spec_x2_st_base(int lane) { return NNTK_SPEC_E2_Q * (lane >> 3) + NNTK_SPEC_E2_N * (lane & 7); }
// Last-pass butterfly assignment: lane L owns c1 = pi(L) and c2 = 64 - c1 (c1 = 0: c2 = 32).  Butterfly c emits
// Z[64 k'' + c], and the conjugate-symmetry partner of bin 64 k'' + c is 64 (7 - k'') + (64 - c): with this assignment
// BOTH bins of every pair sit in the same lane, so the two packed real frames are separated in registers and the
// third exchange of the textbook scheme disappears.  pi is the identity except for the swaps 1 <-> 16 and 9 <-> 24:
// with the identity, the sixteen c2 values of a half-warp straddle three q' rows of exchange 2 and two of them share an
// 8-byte bank pair (one extra wavefront per load); the two swaps make c1 AND c2 conflict-free in both half-warps
// (found by exhaustive search over the 2^16 bank-preserving assignments; checked by tests/test_spectrogram_emu.py).
NNTK_HD int spec_lane_c1(int lane) {
    return lane == 1 ? 16 : lane == 16 ? 1 : lane == 9 ? 24 : lane == 24 ? 9 : lane;
}
NNTK_HD int spec_lane_c(int lane, int s) {
    const int c1 = spec_lane_c1(lane);
    return s == 0 ? c1 : (c1 == 0 ? 32 : 64 - c1);
}
NNTK_HD int spec_x2_ld_base(int lane, int s) {
    const int c = spec_lane_c(lane, s);
    return NNTK_SPEC_E2_Q * (c & 7) + (c >> 3);
}
NNTK_HD void spec_xchg2_store(const float2 *v, float2 *sm_st) {
    for (int s = 0; s < 2; ++s)
        for (int p = 0; p < 4; ++p)
            sm_st2(sm_st, 4 * NNTK_SPEC_E2_Q * s + 2 * p, v[8 * s + 2 * p], v[8 * s + 2 * p + 1]);
}
NNTK_HD void spec_xchg2_load(float2 *v, const float2 *sm_ld0, const float2 *sm_ld1_) {
    for (int n2 = 0; n2 < 8; ++n2) v[n2] = sm_ld1(sm_ld0, NNTK_SPEC_E2_N * n2);
    for (int n2 = 0; n2 < 8; ++n2) v[8 + n2] = sm_ld1(sm_ld1_, NNTK_SPEC_E2_N * n2);
}
NNTK_HD void spec_pass3(float2 *v) {   // v[k''] = Z[64 k'' + c1], v[8 + k''] = Z[64 k'' + c2]
    radix8(v);
    radix8(v + 8);
}
// pa = 4|A[k]|^2, pb = 4|B[k]|^2 from z = Z[k], m = Z[N-k]:  2A = z + conj(m),  2B = (z - conj(m)) / i
NNTK_HD void spec_pair_power(float2 z, float2 m, float &pa, float &pb) {
#if defined(__CUDA_ARCH__) && defined(NNTK_SPEC_PACKED)
    // (ar, br) = z + m and (bi, ai) = z - m are complex add / subtract: 5 issue slots instead of 8, same roundings
    const float2 s = c_add(z, m), d = c_sub(z, m), t = __fmul2_rn(d, d);
    pa = fmaf(s.x, s.x, t.y);
    pb = fmaf(s.y, s.y, t.x);
    return;
#endif
    float ar = z.x + m.x, ai = z.y - m.y;
    float br = z.y + m.y, bi = z.x - m.x;
    pa = fmaf(ar, ar, ai * ai);
    pb = fmaf(br, br, bi * bi);
}
// In-register split.  Slot j = 0..3: bin 64 j + c1;  slot 4 + j: bin 64 j + c2;  slot 8 (lane 0 only): bin 256.
NNTK_HD int spec_slot_bin(int lane, int slot) {
    if (slot == 8) return 256;
    return 64 * (slot & 3) + spec_lane_c(lane, slot >> 2);
}
NNTK_HD void spec_split_regs(int lane, const float2 *v, float *pa, float *pb) {
    const bool l0 = lane == 0;
    for (int j = 0; j < 4; ++j) {
        // partner of 64 j + c1: lane 0 (c1 = 0): Z[64 (8 - j) mod 512] = v[(8 - j) & 7]; else Z[64 (7 - j) + c2] = v[15 - j]
        float2 m1 = l0 ? v[(8 - j) & 7] : v[15 - j];
        spec_pair_power(v[j], m1, pa[j], pb[j]);
        // partner of 64 j + c2: lane 0 (c2 = 32): v[15 - j]; else Z[64 (7 - j) + c1] = v[7 - j]
        float2 m2 = l0 ? v[15 - j] : v[7 - j];
        spec_pair_power(v[8 + j], m2, pa[4 + j], pb[4 + j]);
    }
    spec_pair_power(v[4], v[4], pa[8], pb[8]);   // bin 256 (meaningful on lane 0 only)
}
