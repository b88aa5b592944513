// spectrogram.cu -- K1: batched STFT front-end (replaces signal/spectrogram.c:113-135 + signal/dft.c:34-47 +
// third_party/kissfft/kiss_fft.c:236-300 of the reference for the inference path).
//
// Fast path (nfft == 512): one CTA owns a tile of F consecutive frames of one clip.  The clip span the tile needs,
// (F-1)*step + window_size samples, is read from HBM ONCE with 16-byte loads into shared memory; the frame overlap
// (hop 160 vs window 512 => every sample belongs to 3.2 frames) is served from shared memory.  One warp transforms
// TWO real frames as one 512-point complex FFT held in registers (csrc/spectrogram_core.h), the window multiply is
// fused into the register load, the two frames are separated IN REGISTERS (no third exchange), magnitude / PSD scaling
// is applied there, and the [F, 257] result tile -- a flat
// contiguous span of the [n_clips, ntime, nfreq] output, whose 257-float rows are never 16-byte aligned -- is staged
// in shared memory with the alignment of its global address and written back with 16-byte stores.
//
// Generic path (any other nfft): one CTA per frame, shared-memory radix-2 FFT when nfft is a power of two, direct
// O(nfft^2) DFT from the twiddle table otherwise.  Correct for every nfft the reference accepts, not tuned.
//
// HBM roofline: algorithmic bytes per clip = 4*input_size + 4*ntime*nfreq.
#include "common.cuh"
#include "spectrogram_core.h"

namespace {

struct SpecParams {
    int window_size, step, input_size, ntime, nfreq;
    int frames_per_tile, tiles_per_clip;
    int mode;           // 0 magnitude, 1 PSD
    float mag_mul;      // magnitude: |norm| * 0.5 / scale            (the split pass yields 2|X|)
    float psd_mid;      // PSD bins 1..nfreq-2: norm^2 * 0.25 * 2/scale (the split pass yields 4|X|^2)
    float psd_edge;     // PSD bins 0, nfreq-1: norm^2 * 0.25 / scale
    int in_vec_ok;      // input base pointer 16-byte aligned
    int out_vec_ok;     // output base pointer 16-byte aligned
};

__device__ __forceinline__ float sqrt_approx(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// Shared memory layout (floats):
//   [0, in_cap)                    clip span, stored at offset (global index & 3) so 16-byte global loads line up
//   [in_cap, in_cap + out_cap)     output stage, stored at offset (global index & 3)
//   then 8 warps x 576 float2      FFT exchange buffers
__global__ void __launch_bounds__(256, 2)
spectrogram512_kernel(SpecParams p, int in_cap, int out_cap, const float *__restrict__ window,
                      const float2 *__restrict__ tw512, const float *__restrict__ in, float *__restrict__ out) {
    extern __shared__ __align__(16) float smem[];
    float *s_in = smem;
    float *s_out = smem + in_cap;
    float2 *s_x = reinterpret_cast<float2 *>(smem + in_cap + out_cap);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long long clip = blockIdx.x / p.tiles_per_clip;
    const int tile = blockIdx.x - (int)(clip * p.tiles_per_clip);
    const int f0 = tile * p.frames_per_tile;
    const int nf = min(p.frames_per_tile, p.ntime - f0);

    // ---- 1. clip span -> shared memory (each sample read from HBM once per tile) -------------------------------
    const long long gi0 = clip * (long long)p.input_size + (long long)f0 * p.step;
    const int span = (nf - 1) * p.step + p.window_size;
    const int ia = p.in_vec_ok ? (int)(gi0 & 3) : 0;
    if (p.in_vec_ok) {
        // smem index s <-> global gi0 - ia + s ; vector v covers s = 4v..4v+3
        const float *gbase = in + (gi0 - ia);
        const int total = ia + span;
        const int nvec = total >> 2;
        for (int v = tid; v < nvec; v += 256) {
            float4 x = __ldg(reinterpret_cast<const float4 *>(gbase) + v);
            *reinterpret_cast<float4 *>(s_in + 4 * v) = x;
        }
        for (int s = 4 * nvec + tid; s < total; s += 256) s_in[s] = __ldg(gbase + s);
    } else {
        const float *gbase = in + gi0;
        for (int s = tid; s < span; s += 256) s_in[s] = __ldg(gbase + s);
    }

    // per-lane FFT constants (registers) while the loads are in flight
    SpecLaneConst lc;
    spec_lane_const(lane, tw512, lc);
    __syncthreads();

    // ---- 2. FFTs: warp w handles frame pairs w, w+8, ... --------------------------------------------------------
    const long long go0 = (clip * (long long)p.ntime + f0) * (long long)p.nfreq;
    const int oa = p.out_vec_ok ? (int)(go0 & 3) : 0;
    float2 *sx = s_x + warp * NNTK_SPEC_SMEM_F2;
    const int npairs = (nf + 1) >> 1;
    for (int pr = warp; pr < npairs; pr += 8) {
        const int fa = 2 * pr, fb = 2 * pr + 1;
        const bool has_b = fb < nf;
        const float *xa = s_in + ia + fa * p.step;
        const float *xb = s_in + ia + (has_b ? fb : fa) * p.step;
        float2 v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int n = lane + 32 * j;
            float w = 0.0f, a = 0.0f, b = 0.0f;
            if (n < p.window_size) {
                w = __ldg(window + n);
                a = xa[n];
                b = xb[n];
            }
            v[j] = make_float2(w * a, has_b ? w * b : 0.0f);
        }
        spec_pass1(v, lc);
        spec_xchg1_store(lane, v, sx);
        __syncwarp();
        spec_xchg1_load(lane, v, sx);
        __syncwarp();
        spec_pass2(v, lc);
        spec_xchg2_store(lane, v, sx);
        __syncwarp();
        spec_xchg2_load(lane, v, sx);
        __syncwarp();
        spec_pass3(v);
        // split the packed pair in registers, scale, stage
        float pa[9], pb[9];
        spec_split_regs(lane, v, pa, pb);
        float *oa_row = s_out + oa + fa * p.nfreq;
        float *ob_row = oa_row + p.nfreq;
        const int c1 = lane, c2 = lane == 0 ? 32 : 64 - lane;
#pragma unroll
        for (int slot = 0; slot < 9; ++slot) {
            if (slot == 8 && lane != 0) break;
            const int k = slot == 8 ? 256 : 64 * (slot & 3) + ((slot >> 2) ? c2 : c1);
            float ra, rb;
            if (p.mode == 0) {
                ra = sqrt_approx(pa[slot]) * p.mag_mul;
                rb = sqrt_approx(pb[slot]) * p.mag_mul;
            } else {
                const float m = (k == 0 || k == p.nfreq - 1) ? p.psd_edge : p.psd_mid;
                ra = pa[slot] * m;
                rb = pb[slot] * m;
            }
            oa_row[k] = ra;
            if (has_b) ob_row[k] = rb;
        }
        __syncwarp();
    }
    __syncthreads();

    // ---- 3. output stage -> HBM, one flat span -------------------------------------------------------------------
    const int n_out = nf * p.nfreq;
    if (p.out_vec_ok) {
        float *gbase = out + (go0 - oa);   // 16-byte aligned
        const int head = oa ? (4 - oa) : 0;                 // scalars before the first aligned vector
        const int first_vec = oa ? 1 : 0;
        const int end = oa + n_out;
        const int last_vec = end >> 2;                      // vectors [first_vec, last_vec) are complete
        for (int s = oa + tid; s < oa + head && s < end; s += 256) gbase[s] = s_out[s];
        for (int v = first_vec + tid; v < last_vec; v += 256)
            reinterpret_cast<float4 *>(gbase)[v] = *reinterpret_cast<const float4 *>(s_out + 4 * v);
        for (int s = max(4 * last_vec, oa + head) + tid; s < end; s += 256) gbase[s] = s_out[s];
    } else {
        float *gbase = out + go0;
        for (int s = tid; s < n_out; s += 256) gbase[s] = s_out[s];
    }
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

extern "C" int nntk_cuda_spectrogram(const nntk_spectrogram_args *a, const float *d_in, float *d_out, int n_clips) {
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
    p.out_vec_ok = (((uintptr_t)d_out) & 15) == 0;
    cudaStream_t st = nntk_stream();

    if (a->nfft == NNTK_SPEC_N) {
        // frames per tile: largest of 32/16/8/4/2 whose shared memory lets two CTAs share an SM
        const size_t xchg = (size_t)8 * NNTK_SPEC_SMEM_F2 * sizeof(float2);
        int F = 32;
        size_t bytes = 0;
        int in_cap = 0, out_cap = 0;
        for (;; F >>= 1) {
            in_cap = (((F - 1) * a->step + a->window_size + 3 + 3) / 4) * 4;
            out_cap = ((F * a->nfreq + 3 + 3) / 4) * 4;
            bytes = (size_t)(in_cap + out_cap) * sizeof(float) + xchg;
            if (bytes <= 112 * 1024 || F == 2) break;
        }
        if (bytes > 227 * 1024) {
            nntk_cuda_set_error(NNTK_ERR_UNSUPPORTED_, "spectrogram: step %d too large for the tiled kernel", a->step);
            return NNTK_ERR_UNSUPPORTED_;
        }
        p.frames_per_tile = F;
        p.tiles_per_clip = (a->ntime + F - 1) / F;
        static size_t attr_bytes = 0;
        if (bytes > attr_bytes) {
            NNTK_CUDA_TRY(cudaFuncSetAttribute(spectrogram512_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               (int)bytes));
            attr_bytes = bytes;
        }
        long long grid = (long long)n_clips * p.tiles_per_clip;
        if (grid > 0x7fffffffLL) {
            nntk_cuda_set_error(NNTK_ERR_ARG_, "spectrogram: %d clips exceed one launch", n_clips);
            return NNTK_ERR_ARG_;
        }
        spectrogram512_kernel<<<(unsigned)grid, 256, bytes, st>>>(p, in_cap, out_cap, a->d_window,
                                                                 reinterpret_cast<const float2 *>(a->d_twiddle), d_in,
                                                                 d_out);
        NNTK_LAUNCH_CHECK("spectrogram512_kernel");
        return 0;
    }
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
