// spectrogram_core.h -- the per-warp arithmetic of the K1 spectrogram kernel (nfft = 512), written once and used by
//   * csrc/spectrogram.cu                 (the sm_100a kernel; one warp = two frames), and
//   * tests/emu/spectrogram_emu.cpp       (a CPU emulation of the 32 lanes, used by the CPU test tier to check the
//                                           index maps / butterflies against the oracle before any GPU time is spent).
//
// Algorithm.  The reference runs a 512-point COMPLEX kissfft on each real frame with imag = 0
// (signal/spectrogram.c:120-128, signal/dft.c:34-47).  Here two consecutive real frames a, b are packed as one complex
// signal z = a + i*b, transformed by one 512-point complex FFT (three radix-8 decimation-in-frequency passes,
// 16 points per lane, data in registers, TWO shared-memory exchanges), and separated with the conjugate symmetry
//   A[k] = (Z[k] + conj(Z[N-k])) / 2,   B[k] = (Z[k] - conj(Z[N-k])) / (2i).
// Twiddles come from the host table built exactly like kiss_fft.c:357-363 (double cos/sin rounded to float).
//
// Index maps (N = 512, lane L = 0..31):
//   load      v[j]            = z[L + 32 j]                               j = 0..15
//   pass 1    butterfly s=0,1 : inputs v[s + 2m] = z[n' + 64 m], n' = L + 32 s;  y_q[n'] = (sum_m . W8^{mq}) W512^{n' q}
//   xchg 1    E1[q*72 + n']   (float2 units; 72 = 64 + 8 pad -> conflict-free LDS.64/STS.64)
//   pass 2    butterfly b = L + 32 s, q = b / 8, n'' = b % 8: inputs E1[q*72 + n'' + 8 m];
//             u_{q'}[n''] = (sum_m . W8^{m q'}) W64^{n'' q'}
//   xchg 2    E2[c*8 + (n'' ^ ((c>>1)&7))], c = 8 q' + q   (XOR-swizzled, no padding)
//   pass 3    butterfly c (lane L owns c = L and c = 64 - L): Z[64 k'' + c] = sum_{n''} E2[c][n''] W8^{n'' k''}
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
#define NNTK_SPEC_SMEM_F2 576 /* float2 slots per warp: exchange 1 uses 8*72, exchange 2 uses 512 */

NNTK_HD float2 c_add(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
NNTK_HD float2 c_sub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
// (a.x + i a.y)(w.x + i w.y), same operand pairing as kissfft's C_MUL (_kiss_fft_guts.h)
NNTK_HD float2 c_mul(float2 a, float2 w) { return make_float2(a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x); }
NNTK_HD float2 c_mul_mi(float2 a) { return make_float2(a.y, -a.x); }  // a * (-i)

// 8-point forward DFT, natural-order in and out: x[q] <- sum_m x[m] * exp(-2 pi i m q / 8).
NNTK_HD void radix8(float2 *x) {
    const float R = 0.70710678118654752440f;
    float2 a0 = c_add(x[0], x[4]), b0 = c_sub(x[0], x[4]);
    float2 a1 = c_add(x[1], x[5]), b1 = c_sub(x[1], x[5]);
    float2 a2 = c_add(x[2], x[6]), b2 = c_sub(x[2], x[6]);
    float2 a3 = c_add(x[3], x[7]), b3 = c_sub(x[3], x[7]);
    // b_m *= W8^m
    b1 = make_float2((b1.x + b1.y) * R, (b1.y - b1.x) * R);
    b2 = c_mul_mi(b2);
    b3 = make_float2((b3.y - b3.x) * R, -(b3.x + b3.y) * R);
    // even outputs: 4-point DFT of a
    float2 e0 = c_add(a0, a2), e1 = c_sub(a0, a2), o0 = c_add(a1, a3), o1 = c_mul_mi(c_sub(a1, a3));
    x[0] = c_add(e0, o0);
    x[4] = c_sub(e0, o0);
    x[2] = c_add(e1, o1);
    x[6] = c_sub(e1, o1);
    // odd outputs: 4-point DFT of the twiddled b
    float2 f0 = c_add(b0, b2), f1 = c_sub(b0, b2), g0 = c_add(b1, b3), g1 = c_mul_mi(c_sub(b1, b3));
    x[1] = c_add(f0, g0);
    x[5] = c_sub(f0, g0);
    x[3] = c_add(f1, g1);
    x[7] = c_sub(f1, g1);
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
NNTK_HD void spec_xchg1_store(int lane, const float2 *v, float2 *sm) {
    for (int s = 0; s < 2; ++s)
        for (int q = 0; q < 8; ++q) sm[q * 72 + lane + 32 * s] = v[s + 2 * q];
}
NNTK_HD void spec_xchg1_load(int lane, float2 *v, const float2 *sm) {
    for (int s = 0; s < 2; ++s) {
        int q = (lane >> 3) + 4 * s, n2 = lane & 7;
        for (int m = 0; m < 8; ++m) v[8 * s + m] = sm[q * 72 + n2 + 8 * m];
    }
}
// pass 2 on v[8 s + m]
NNTK_HD void spec_pass2(float2 *v, const SpecLaneConst &c) {
    for (int s = 0; s < 2; ++s) {
        float2 *x = v + 8 * s;
        radix8(x);
        for (int q = 1; q < 8; ++q) x[q] = c_mul(x[q], c.tw2[q - 1]);
    }
}
// Exchange 2 layout: the 8 inputs of last-pass butterfly c (c = 8 q' + q = low 6 bits of the output bin) live at
// E2[c*8 + (n'' ^ ((c >> 1) & 7))] -- 512 float2, no padding; the XOR makes both the store (lanes vary q, n'') and the
// mirrored load below (lanes vary c) conflict-free for 64-bit accesses.
NNTK_HD int spec_e2_index(int c, int n2) { return c * 8 + (n2 ^ ((c >> 1) & 7)); }
NNTK_HD void spec_xchg2_store(int lane, const float2 *v, float2 *sm) {
    for (int s = 0; s < 2; ++s) {
        int q = (lane >> 3) + 4 * s, n2 = lane & 7;
        for (int qp = 0; qp < 8; ++qp) sm[spec_e2_index(8 * qp + q, n2)] = v[8 * s + qp];
    }
}
// Last-pass butterfly assignment: lane L owns c1 = L and c2 = 64 - L (lane 0: c = 0 and 32).  Butterfly c emits
// Z[64 k'' + c], and the conjugate-symmetry partner of bin 64 k'' + c is 64 (7 - k'') + (64 - c): with this assignment
// BOTH bins of every pair sit in the same lane, so the two packed real frames are separated in registers and the
// third exchange of the textbook scheme disappears.
NNTK_HD int spec_lane_c(int lane, int s) { return s == 0 ? lane : (lane == 0 ? 32 : 64 - lane); }
NNTK_HD void spec_xchg2_load(int lane, float2 *v, const float2 *sm) {
    for (int s = 0; s < 2; ++s) {
        int c = spec_lane_c(lane, s);
        for (int n2 = 0; n2 < 8; ++n2) v[8 * s + n2] = sm[spec_e2_index(c, n2)];
    }
}
NNTK_HD void spec_pass3(float2 *v) {   // v[k''] = Z[64 k'' + c1], v[8 + k''] = Z[64 k'' + c2]
    radix8(v);
    radix8(v + 8);
}
// pa = 4|A[k]|^2, pb = 4|B[k]|^2 from z = Z[k], m = Z[N-k]:  2A = z + conj(m),  2B = (z - conj(m)) / i
NNTK_HD void spec_pair_power(float2 z, float2 m, float &pa, float &pb) {
    float ar = z.x + m.x, ai = z.y - m.y;
    float br = z.y + m.y, bi = z.x - m.x;
    pa = ar * ar + ai * ai;
    pb = br * br + bi * bi;
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
