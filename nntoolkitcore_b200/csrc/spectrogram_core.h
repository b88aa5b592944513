// spectrogram_core.h -- the per-warp arithmetic of the K1 spectrogram kernel (nfft = 512), written once and used by
//   * csrc/spectrogram.cu                 (the sm_100a kernel; one warp = two frames), and
//   * tests/emu/spectrogram_emu.cpp       (a CPU emulation of the 32 lanes, used by the CPU test tier to check the
//                                           index maps / butterflies against the oracle before any GPU time is spent).
//
// Algorithm.  The reference runs a 512-point COMPLEX kissfft on each real frame with imag = 0
// (signal/spectrogram.c:120-128, signal/dft.c:34-47).  Here two consecutive real frames a, b are packed as one complex
// signal z = a + i*b, transformed by one 512-point complex FFT (three radix-8 decimation-in-frequency passes,
// 16 points per lane, data in registers, three shared-memory exchanges), and separated with the conjugate symmetry
//   A[k] = (Z[k] + conj(Z[N-k])) / 2,   B[k] = (Z[k] - conj(Z[N-k])) / (2i).
// Twiddles come from the host table built exactly like kiss_fft.c:357-363 (double cos/sin rounded to float).
//
// Index maps (N = 512, lane L = 0..31):
//   load      v[j]            = z[L + 32 j]                               j = 0..15
//   pass 1    butterfly s=0,1 : inputs v[s + 2m] = z[n' + 64 m], n' = L + 32 s;  y_q[n'] = (sum_m . W8^{mq}) W512^{n' q}
//   xchg 1    E1[q*72 + n']   (float2 units; 72 = 64 + 8 pad -> conflict-free LDS.64/STS.64)
//   pass 2    butterfly b = L + 32 s, q = b / 8, n'' = b % 8: inputs E1[q*72 + n'' + 8 m];
//             u_{q'}[n''] = (sum_m . W8^{m q'}) W64^{n'' q'}
//   xchg 2    E2[(8 q + q') * 9 + n'']   (row stride 9)
//   pass 3    butterfly b = L + 32 s = 8 q + q': inputs E2[b*9 + n''];  Z[64 k'' + 8 q' + q] = sum_{n''} . W8^{n'' k''}
//   xchg 3    ZS[pad(k)] with pad(k) = k + 2 (k >> 4), natural order
//   split     k = L + 32 j (j = 0..7) and k = 256 on lane 0.
#pragma once

#ifdef __CUDACC__
#define NNTK_HD __host__ __device__ __forceinline__
#else
#define NNTK_HD inline
struct float2 { float x, y; };
static inline float2 make_float2(float x, float y) { float2 r; r.x = x; r.y = y; return r; }
#endif

#define NNTK_SPEC_N 512
#define NNTK_SPEC_SMEM_F2 576 /* float2 slots per warp for every exchange: 8*72 = 64*9 = 512 + 64 */

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
NNTK_HD void spec_xchg2_store(int lane, const float2 *v, float2 *sm) {
    for (int s = 0; s < 2; ++s) {
        int q = (lane >> 3) + 4 * s, n2 = lane & 7;
        for (int qp = 0; qp < 8; ++qp) sm[(8 * q + qp) * 9 + n2] = v[8 * s + qp];
    }
}
NNTK_HD void spec_xchg2_load(int lane, float2 *v, const float2 *sm) {
    for (int s = 0; s < 2; ++s)
        for (int n2 = 0; n2 < 8; ++n2) v[8 * s + n2] = sm[(lane + 32 * s) * 9 + n2];
}
NNTK_HD void spec_pass3(float2 *v) {
    radix8(v);
    radix8(v + 8);
}
NNTK_HD int spec_pad(int k) { return k + 2 * (k >> 4); }
NNTK_HD void spec_xchg3_store(int lane, const float2 *v, float2 *sm) {
    for (int s = 0; s < 2; ++s) {
        int q = (lane >> 3) + 4 * s, qp = lane & 7;
        for (int k2 = 0; k2 < 8; ++k2) sm[spec_pad(64 * k2 + 8 * qp + q)] = v[8 * s + k2];
    }
}
// Squared magnitudes (times 4) of the two packed real frames at bin k: pa = 4|A[k]|^2, pb = 4|B[k]|^2.
NNTK_HD void spec_split_power(const float2 *sm, int k, float &pa, float &pb) {
    float2 z = sm[spec_pad(k)];
    float2 n = sm[spec_pad((NNTK_SPEC_N - k) & (NNTK_SPEC_N - 1))];
    float ar = z.x + n.x, ai = z.y - n.y;  // 2 A[k]
    float br = z.y + n.y, bi = z.x - n.x;  // 2 B[k] = (br, -bi)
    pa = ar * ar + ai * ai;
    pb = br * br + bi * bi;
}
