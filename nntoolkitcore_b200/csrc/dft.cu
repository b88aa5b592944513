// dft.cu -- standalone batched complex DFT on planar (split) buffers: SURVEY.md section 8 row f4, replacing
// signal/dft.c:34-47 + third_party/kissfft/kiss_fft.c of the reference (forward exp(-2 pi i kn/N), inverse exp(+...),
// neither normalised).
//
// nfft == 512: one warp per signal, the register-resident three-pass radix-8 transform of the spectrogram kernel
//   (csrc/spectrogram_core.h: 16 points per lane, two shared-memory exchanges); loads are 32 consecutive floats per
//   instruction, and the last pass leaves Z[64 k + c] in lane order c = pi(lane) / 64 - pi(lane), so the stores are
//   32 consecutive floats as well.  The inverse runs the same code on the conjugated input and conjugates the result.
// other sizes: one CTA per signal in shared memory -- radix-2 for powers of two, direct O(N^2) sums otherwise (correct for
//   every nfft the reference accepts, not tuned).
// HBM roofline: 16 bytes per complex point (read re, im; write re, im).
#include "common.cuh"
#define NNTK_SPEC_PACKED 1   // complex adds / subtracts of the butterflies as sm_100 FADD2 / FFMA2 (spectrogram_core.h)
#include "spectrogram_core.h"

namespace {

constexpr int DFT_WARPS = 8;

template <bool INV>
__global__ void __launch_bounds__(DFT_WARPS * 32)
dft512_kernel(const float *re, const float *im, float *ore, float *oim,   // planes may alias (in-place): no __restrict__
              const float2 *__restrict__ tw512, long long batch) {
    __shared__ __align__(16) float2 s_x[DFT_WARPS * NNTK_SPEC_SMEM_F2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    SpecLaneConst lc;
    spec_lane_const(lane, tw512, lc);
    float2 *sx = s_x + warp * NNTK_SPEC_SMEM_F2;
    float2 *x1_st = sx + spec_x1_st_base(lane);
    const float2 *x1_ld = sx + spec_x1_ld_base(lane);
    float2 *x2_st = sx + spec_x2_st_base(lane);
    const float2 *x2_ld0 = sx + spec_x2_ld_base(lane, 0);
    const float2 *x2_ld1 = sx + spec_x2_ld_base(lane, 1);
    const int c1 = spec_lane_c(lane, 0), c2 = spec_lane_c(lane, 1);
    const long long n_warps = (long long)gridDim.x * DFT_WARPS;
    for (long long t = (long long)blockIdx.x * DFT_WARPS + warp; t < batch; t += n_warps) {
        const float *r = re + t * NNTK_SPEC_N + lane, *i = im + t * NNTK_SPEC_N + lane;
        float2 v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float a = r[32 * j], b = i[32 * j];
            v[j] = make_float2(a, INV ? -b : b);
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
        spec_pass3(v);          // v[k] = Z[64 k + c1], v[8 + k] = Z[64 k + c2]
        float *orow = ore + t * NNTK_SPEC_N, *irow = oim + t * NNTK_SPEC_N;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            orow[64 * k + c1] = v[k].x;
            irow[64 * k + c1] = INV ? -v[k].y : v[k].y;
            orow[64 * k + c2] = v[8 + k].x;
            irow[64 * k + c2] = INV ? -v[8 + k].y : v[8 + k].y;
        }
    }
}

// One CTA per signal.  s_a / s_b: complex buffers of nfft points.
__global__ void __launch_bounds__(128)
dft_generic_kernel(int nfft, int log2n, int inverse, const float2 *__restrict__ tw, const float *re, const float *im,
                   float *ore, float *oim) {
    extern __shared__ __align__(16) float smem[];
    float2 *s_a = reinterpret_cast<float2 *>(smem);
    float2 *s_b = s_a + nfft;
    const int tid = threadIdx.x;
    const long long base = (long long)blockIdx.x * nfft;
    const float sgn = inverse ? -1.0f : 1.0f;   // conjugates the forward twiddles
    if (log2n >= 0) {
        for (int n = tid; n < nfft; n += blockDim.x) {
            const int r = (log2n > 0) ? (int)(__brev((unsigned)n) >> (32 - log2n)) : 0;
            s_a[r] = make_float2(re[base + n], im[base + n]);
        }
        __syncthreads();
        for (int s = 0; s < log2n; ++s) {
            const int half = 1 << s, tstride = nfft >> (s + 1);
            for (int b = tid; b < (nfft >> 1); b += blockDim.x) {
                const int j = b & (half - 1);
                const int i0 = ((b >> s) << (s + 1)) + j, i1 = i0 + half;
                float2 w = tw[j * tstride];
                w.y *= sgn;
                const float2 u = s_a[i0], y = c_mul(s_a[i1], w);
                s_a[i0] = c_add(u, y);
                s_a[i1] = c_sub(u, y);
            }
            __syncthreads();
        }
    } else {
        for (int n = tid; n < nfft; n += blockDim.x) s_b[n] = make_float2(re[base + n], im[base + n]);
        __syncthreads();
        for (int k = tid; k < nfft; k += blockDim.x) {
            float2 acc = make_float2(0.0f, 0.0f);
            int idx = 0;
            for (int n = 0; n < nfft; ++n) {
                float2 w = tw[idx];
                w.y *= sgn;
                const float2 x = s_b[n];
                acc.x = fmaf(x.x, w.x, fmaf(-x.y, w.y, acc.x));
                acc.y = fmaf(x.x, w.y, fmaf(x.y, w.x, acc.y));
                idx += k;
                if (idx >= nfft) idx -= nfft;
            }
            s_a[k] = acc;
        }
        __syncthreads();
    }
    for (int k = tid; k < nfft; k += blockDim.x) {
        ore[base + k] = s_a[k].x;
        oim[base + k] = s_a[k].y;
    }
}

__global__ void __launch_bounds__(256) copy_strided_kernel(float *dst, int ds, const float *__restrict__ src, int ss, long long n) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i * ds] = src[i * ss];
}

}  // namespace

extern "C" int nntk_cuda_dft(int nfft, int inverse, const float *d_twiddle, const float *d_in_re, const float *d_in_im,
                             float *d_out_re, float *d_out_im, long long batch) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (nfft < 1 || !d_twiddle || !d_in_re || !d_in_im || !d_out_re || !d_out_im || batch < 0) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "dft: bad arguments (nfft %d)", nfft);
        return NNTK_ERR_ARG_;
    }
    if (batch == 0) return 0;
    cudaStream_t st = nntk_stream();
    const float2 *tw = reinterpret_cast<const float2 *>(d_twiddle);
    if (nfft == NNTK_SPEC_N) {
        long long grid = (batch + DFT_WARPS - 1) / DFT_WARPS;
        const long long cap = 8LL * nntk_cuda_sm_count();
        if (grid > cap) grid = cap;
        if (inverse) dft512_kernel<true><<<(unsigned)grid, DFT_WARPS * 32, 0, st>>>(d_in_re, d_in_im, d_out_re, d_out_im, tw, batch);
        else dft512_kernel<false><<<(unsigned)grid, DFT_WARPS * 32, 0, st>>>(d_in_re, d_in_im, d_out_re, d_out_im, tw, batch);
        NNTK_LAUNCH_CHECK("dft512_kernel");
        return 0;
    }
    int log2n = -1;
    if ((nfft & (nfft - 1)) == 0) {
        log2n = 0;
        while ((1 << log2n) < nfft) ++log2n;
    }
    const size_t bytes = (size_t)2 * nfft * sizeof(float2);
    if (bytes > 200 * 1024 || batch > 0x7fffffffLL) {
        nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "dft: nfft %d / batch %lld too large", nfft, batch);
        return NNTK_ERR_UNSUPPORTED_;
    }
    static size_t attr_bytes = 48 * 1024;
    if (bytes > attr_bytes) {
        NNTK_CUDA_TRY(cudaFuncSetAttribute(dft_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
        attr_bytes = bytes;
    }
    dft_generic_kernel<<<(unsigned)batch, 128, bytes, st>>>(nfft, log2n, inverse ? 1 : 0, tw, d_in_re, d_in_im, d_out_re, d_out_im);
    NNTK_LAUNCH_CHECK("dft_generic_kernel");
    return 0;
}

extern "C" int nntk_cuda_copy_strided(float *d_dst, int dst_stride, const float *d_src, int src_stride, long long n) {
    int rc = nntk_cuda_ensure_init();
    if (rc) return rc;
    if (!d_dst || !d_src || n < 0 || dst_stride < 1 || src_stride < 1) {
        nntk_cuda_set_error(NNTK_ERR_ARG_, "copy_strided: bad arguments");
        return NNTK_ERR_ARG_;
    }
    if (n == 0) return 0;
    long long grid = (n + 255) / 256;
    if (grid > 148 * 16) grid = 148 * 16;
    copy_strided_kernel<<<(unsigned)grid, 256, 0, nntk_stream()>>>(d_dst, dst_stride, d_src, src_stride, n);
    NNTK_LAUNCH_CHECK("copy_strided_kernel");
    return 0;
}
