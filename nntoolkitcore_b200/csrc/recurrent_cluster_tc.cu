// recurrent_cluster_tc.cu -- K4, persistent tensor-core recurrence (NNTK_MATH_TF32, H <= 256): ONE launch runs all T steps
// of GRU / LSTM for the whole batch.  Replaces the per-timestep GEMV pairs of layers/gru.c:129-204 and lstm.c:185-268.
//
// A thread-block CLUSTER of C = H/32 CTAs owns a slice of 128 independent sequences for the whole sequence length:
//   * CTA `rank` owns hidden units [32 rank, 32 rank + 32): its columns of U (all G gates: N = 32 G accumulator columns)
//     are loaded ONCE by TMA, K-major / SWIZZLE_128B / rounded to nearest TF32, and stay resident in shared memory;
//   * h_{t-1} of the 128 sequences (rounded to nearest TF32) is the A operand: a [128 x H] tile in EVERY CTA's shared
//     memory, k-block j (32 floats = one 128-byte swizzle row per sequence) being exactly the slice CTA j produces;
//   * per step: one elected thread issues H/8 tcgen05.mma (M = 128, N = 32 G, K = 8) into TMEM; tcgen05.commit signals
//     mma_done; 8 epilogue warps read the accumulators (thread = one sequence x 16 hidden units, ALL gates of a unit in
//     one thread), add b_h and the hoisted projection xw_t (prefetched during the previous exchange), apply the gate
//     nonlinearities and the state update (gru.c:156-186, lstm.c:199-238 op order), keep h and c in REGISTERS for the
//     whole sequence, store h_t to HBM (return_sequences), write the rounded h_t slice into their own A k-block and push
//     that 16 KB k-block to every peer with one DSMEM bulk copy each (cp.async.bulk shared::cta -> shared::cluster) that
//     completes on the peer's a_full mbarrier;
//   * flow control is two mbarriers per CTA, no cluster-wide hardware barrier inside the loop:
//       a_full  (2 arrivals + (C-1) x 16 KB tx per phase)  "all k-blocks of h_t have landed here"   -> MMA of step t+1
//       a_free  (C remote arrivals per phase)             "every CTA's MMA of step t has read its A" -> push h_t
// Slices are independent, so a grid of ceil(B/128) clusters needs no grid-level synchronisation; c never touches HBM, h
// only as the layer output.  Shared memory: 128 H x 4 (A) + 128 G H (U slice) bytes, e.g. GRU-256: 128 + 96 KB.
// Eligibility: H % 32 == 0, H <= 256, C <= 16, the above within 227 KB (GRU/RNN-256, LSTM-128; LSTM-256 does not fit).

#include "tc_common.cuh"

namespace {

constexpr int CT_THREADS = 320;              // warp 0: a_free signalling, warp 1: MMA issuer, warps 2..9: epilogue
constexpr int CT_BM = 128;                   // sequences per cluster
constexpr int CT_KB_BYTES = CT_BM * 128;     // one k-block of A: 128 rows x 32 floats

struct CtParams {
    int B, H, T, C;
    int return_sequences;
    const float *bh;                // [G*H] or null (not added)
    const float *xw;                // [B][T][G*H], b_i already added
    float *h_state, *c_state;       // [B][H] in/out
    float *out;                     // [B][T][H] or [B][H]
    int act[5];
    float act_a[5];
};

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t saddr, uint32_t rank) {
    uint32_t d;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(saddr), "r"(rank));
    return d;
}
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void dsmem_bulk_copy(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t mbar_cluster) {
    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst_cluster), "r"(src_cta), "r"(bytes), "r"(mbar_cluster) : "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// DEF: the default gate activations (GRU sigmoid, sigmoid, tanh; LSTM sigmoid, sigmoid, tanh, sigmoid, tanh) compiled in --
// the run-time kind switch per call cost ~3x the instructions of the gate math itself (ncu, r1).
template <bool DEF, int KIND>
__device__ __forceinline__ float ct_act(const CtParams &p, int slot, float x) {
    return DEF ? act_fast(KIND, 1.0f, x) : act_fast(p.act[slot], p.act_a[slot], x);
}

template <int G, bool DEF>
__global__ void __launch_bounds__(CT_THREADS, 1)
recurrent_cluster_tc_kernel(const __grid_constant__ CUtensorMap map_u, CtParams p) {
    constexpr int N = 32 * G;
    constexpr uint32_t TMEM_COLS = N <= 32 ? 32 : (N <= 64 ? 64 : 128);
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    const int C = p.C, H = p.H, T = p.T, GH = G * H;
    uint8_t *sA = smem;                                   // [C k-blocks][128 rows][32 floats], swizzled
    uint8_t *sB = smem + (size_t)C * CT_KB_BYTES;         // [C k-blocks][N rows][32 floats], swizzled
    uint64_t *bars = reinterpret_cast<uint64_t *>(sB + (size_t)C * N * 128);
    uint64_t *b_full = bars, *a_full = bars + 1, *mma_done = bars + 2, *a_free = bars + 3;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 4);
    float *s_bh = reinterpret_cast<float *>(bars + 6);    // [G][32]

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_rank();
    const int row0 = (blockIdx.x / C) * CT_BM;

    if (tid == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_u) : "memory");
        mbar_init(b_full, 1); mbar_init(a_full, 2); mbar_init(mma_done, 1); mbar_init(a_free, (uint32_t)C);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        mbar_expect_tx(b_full, (uint32_t)(C * N * 128));
        for (int kb = 0; kb < C; ++kb) tma_load_2d(sB + (size_t)kb * N * 128, &map_u, b_full, kb * 32, (int)rank * N);
        mbar_expect_tx(a_full, (uint32_t)((C - 1) * CT_KB_BYTES));      // phase 0: the initial exchange of h_0
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
    cluster_sync_all();          // every CTA's barriers exist before anybody arrives on / copies into a peer
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== a_free signalling: "my MMA of step t is complete, my A tile may be overwritten" -> every CTA =====
        if (lane < C) {
            const uint32_t remote = map_to_rank(smem_u32(a_free), (uint32_t)lane);
            for (int t = 0; t + 1 < T; ++t) {
                mbar_wait(mma_done, (uint32_t)(t & 1));
                mbar_arrive_remote(remote);
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            constexpr uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(CT_BM >> 4) << 24);
            const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
            mbar_wait(b_full, 0);
            for (int t = 0; t < T; ++t) {
                mbar_wait(a_full, (uint32_t)(t & 1));
                if (t + 1 < T) mbar_expect_tx(a_full, (uint32_t)((C - 1) * CT_KB_BYTES));   // arm the next phase
                tc_fence_after();
                for (int kb = 0; kb < C; ++kb) {
                    const uint64_t da = make_desc(a0 + kb * CT_KB_BYTES), db = make_desc(b0 + kb * N * 128);
#pragma unroll
                    for (int k = 0; k < 4; ++k) tc_mma_tf32(tmem_base, da + 2 * k, db + 2 * k, idesc, (kb | k) ? 1u : 0u);
                }
                tc_commit(mma_done);
            }
        }
    } else {
        // ===== epilogue: thread = sequence (row) x 16 hidden units =====
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int r = q * 32 + lane;
        const int row = row0 + r;
        const bool row_ok = row < p.B;
        const long long rr = row_ok ? row : 0;
        const int ul = half * 16;                          // first unit inside the CTA's 32
        const int unit0 = (int)rank * 32 + ul;             // first hidden unit (global index)
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + ul;
        // own A k-block, this thread's four 16-byte chunks (SWIZZLE_128B: chunk index ^ (row & 7))
        uint8_t *a_row = sA + (size_t)rank * CT_KB_BYTES + (r >> 3) * 1024 + (r & 7) * 128;
        const uint32_t my_kb = smem_u32(sA) + rank * CT_KB_BYTES;
        const uint32_t a_full_s = smem_u32(a_full);

        float h[16], c[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 hv = row_ok ? *reinterpret_cast<const float4 *>(p.h_state + rr * H + unit0 + 4 * j) : make_float4(0.f, 0.f, 0.f, 0.f);
            h[4 * j] = hv.x; h[4 * j + 1] = hv.y; h[4 * j + 2] = hv.z; h[4 * j + 3] = hv.w;
            float4 cv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (G == 4 && row_ok) cv = *reinterpret_cast<const float4 *>(p.c_state + rr * H + unit0 + 4 * j);
            c[4 * j] = cv.x; c[4 * j + 1] = cv.y; c[4 * j + 2] = cv.z; c[4 * j + 3] = cv.w;
        }
        float4 xr[G][4];
        const float *xw_row = p.xw + rr * (long long)T * GH + unit0;
        auto load_xw = [&](int t) {
#pragma unroll
            for (int g = 0; g < G; ++g)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    xr[g][j] = row_ok ? __ldg(reinterpret_cast<const float4 *>(xw_row + (long long)t * GH + g * H) + j)
                                      : make_float4(0.f, 0.f, 0.f, 0.f);
        };
        // publishes the rounded h registers as k-block `rank` of every CTA's A tile; e = exchange index (0 = h_0)
        auto exchange = [&](int e) {
            mbar_wait(a_free, (uint32_t)((e & 1) ^ 1));      // e = 0 passes at once; e = t+1 waits for every MMA of step t
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int chunk = (half * 4 + j) ^ (r & 7);
                *reinterpret_cast<float4 *>(a_row + chunk * 16) =
                    make_float4(round_tf32(h[4 * j]), round_tf32(h[4 * j + 1]), round_tf32(h[4 * j + 2]), round_tf32(h[4 * j + 3]));
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            tc_fence_before();
            epi_bar_sync();
            if (warp == 2 && lane < C) {
                if (lane == (int)rank) mbar_arrive(a_full);
                else dsmem_bulk_copy(map_to_rank(my_kb, (uint32_t)lane), my_kb, CT_KB_BYTES, map_to_rank(a_full_s, (uint32_t)lane));
            }
        };
        exchange(0);
        load_xw(0);
        for (int t = 0; t < T; ++t) {
            mbar_wait(mma_done, (uint32_t)(t & 1));
            tc_fence_after();
            float acc[G][16];
#pragma unroll
            for (int g = 0; g < G; ++g) tc_ld_32x16(taddr + g * 32, acc[g]);
#pragma unroll
            for (int u = 0; u < 16; ++u) {
                float hu[G], x[G];
#pragma unroll
                for (int g = 0; g < G; ++g) {
                    hu[g] = acc[g][u] + s_bh[g * 32 + ul + u];
                    x[g] = (&xr[g][u >> 2].x)[u & 3];
                }
                if (G == 3) {
                    const float z = ct_act<DEF, NNTK_ACT_SIGMOID>(p, 0, x[0] + hu[0]);
                    const float rg = ct_act<DEF, NNTK_ACT_SIGMOID>(p, 1, x[1] + hu[1]);
                    const float hh = ct_act<DEF, NNTK_ACT_TANH>(p, 2, fmaf(rg, hu[2], x[2]));
                    h[u] = fmaf(z, h[u] - hh, hh);                    // (1 - z) hh + z h_{t-1}
                } else if (G == 4) {
                    const float ig = ct_act<DEF, NNTK_ACT_SIGMOID>(p, 0, x[0] + hu[0]);
                    const float fg = ct_act<DEF, NNTK_ACT_SIGMOID>(p, 1, x[1] + hu[1]);
                    const float gg = ct_act<DEF, NNTK_ACT_TANH>(p, 2, x[2] + hu[2]);
                    const float og = ct_act<DEF, NNTK_ACT_SIGMOID>(p, 3, x[3] + hu[3]);
                    c[u] = fmaf(fg, c[u], ig * gg);
                    h[u] = og * ct_act<DEF, NNTK_ACT_TANH>(p, 4, c[u]);
                }
            }
            if (p.return_sequences && row_ok) {
                float4 *o = reinterpret_cast<float4 *>(p.out + (rr * T + t) * H + unit0);
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = make_float4(h[4 * j], h[4 * j + 1], h[4 * j + 2], h[4 * j + 3]);
            }
            if (t + 1 < T) {
                exchange(t + 1);
                load_xw(t + 1);
            }
        }
        if (row_ok) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 hv = make_float4(h[4 * j], h[4 * j + 1], h[4 * j + 2], h[4 * j + 3]);
                *reinterpret_cast<float4 *>(p.h_state + rr * H + unit0 + 4 * j) = hv;
                if (!p.return_sequences) *reinterpret_cast<float4 *>(p.out + rr * H + unit0 + 4 * j) = hv;
                if (G == 4)
                    *reinterpret_cast<float4 *>(p.c_state + rr * H + unit0 + 4 * j) = make_float4(c[4 * j], c[4 * j + 1], c[4 * j + 2], c[4 * j + 3]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();          // nobody leaves while a peer could still address this CTA's shared memory
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

// Bt[(rank*G + g)*32 + u][k] = tf32(U[k][g*H + 32 rank + u])
__global__ void pack_cluster_weights_kernel(const float *__restrict__ U, float *__restrict__ Bt, int H, int G) {
    const long long total = (long long)G * H * H;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int n = (int)(i / H), k = (int)(i - (long long)n * H);
        const int rank = n / (G * 32), rem = n - rank * G * 32, g = rem >> 5, u = rem & 31;
        Bt[i] = round_tf32(__ldg(U + (long long)k * G * H + g * H + rank * 32 + u));
    }
}

float *g_ct_scratch = nullptr;
size_t g_ct_floats = 0;

template <int G, bool DEF>
int launch_ct(const CUtensorMap &mu, const CtParams &p, int slices, size_t smem, cudaStream_t st) {
    auto kern = recurrent_cluster_tc_kernel<G, DEF>;
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
    cfg.gridDim = dim3((unsigned)(slices * p.C));
    cfg.blockDim = dim3(CT_THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)p.C;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    NNTK_CUDA_TRY(cudaLaunchKernelEx(&cfg, kern, mu, p));
    nntk_cuda_count_launch(1);
    return 0;
}

}  // namespace

// Runs all T steps of one batch chunk in ONE launch.  *used = 0 when the shape is not eligible (the caller falls back to
// the per-step tensor-core kernels).  xw: hoisted projection [nb, T, G*H]; h / c: state [nb, H] in/out.
extern "C" int nntk_cuda_recurrent_cluster_tc(const nntk_recurrent_args *r, int nb, const float *xw, float *h, float *c,
                                              float *out, int *used) {
    *used = 0;
    const int G = r->cell, H = r->H, T = r->T;
    if (G == 1 || (H & 31) || H > 256 || T < 1 || nb < 1) return 0;
    const int C = H / 32;
    const size_t smem = (size_t)C * CT_KB_BYTES + (size_t)C * G * 32 * 128 + 64 + (size_t)G * 32 * 4 + 1024;
    if (C > 16 || smem > 227 * 1024) return 0;
    if ((((uintptr_t)h) | ((uintptr_t)out) | ((uintptr_t)xw) | ((uintptr_t)c)) & 15) return 0;
    EncodeTiledFn encode = get_encode();
    if (!encode) {
        nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_cluster_tc: cuTensorMapEncodeTiled is not available from this driver");
        return NNTK_ERR_CUDA_;
    }
    cudaStream_t st = nntk_stream();
    const size_t need = (size_t)G * H * H;
    if (need > g_ct_floats) {
        NNTK_CUDA_TRY(cudaStreamSynchronize(st));
        if (g_ct_scratch) cudaFree(g_ct_scratch);
        g_ct_scratch = nullptr; g_ct_floats = 0;
        NNTK_CUDA_TRY(cudaMalloc(&g_ct_scratch, need * sizeof(float)));
        g_ct_floats = need;
    }
    {
        const long long total = (long long)need;
        const int blocks = (int)((total + 255) / 256 < 1024 ? (total + 255) / 256 : 1024);
        pack_cluster_weights_kernel<<<blocks, 256, 0, st>>>(r->d_U, g_ct_scratch, H, G);
        NNTK_LAUNCH_CHECK("pack_cluster_weights_kernel");
    }
    CUtensorMap mu;
    {
        cuuint64_t dims[2] = {(cuuint64_t)H, (cuuint64_t)G * H};
        cuuint64_t strides[1] = {(cuuint64_t)H * 4};
        cuuint32_t box[2] = {32, (cuuint32_t)(G * 32)};
        cuuint32_t estr[2] = {1, 1};
        CUresult cr = encode(&mu, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void *)g_ct_scratch, dims, strides, box, estr,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) {
            nntk_cuda_set_error(NNTK_ERR_CUDA_, "recurrent_cluster_tc: tensor map for U rejected (CUresult %d)", (int)cr);
            return NNTK_ERR_CUDA_;
        }
    }
    CtParams p;
    p.B = nb; p.H = H; p.T = T; p.C = C; p.return_sequences = r->return_sequences;
    p.bh = ((G == 3) || r->v2) ? r->d_bh : nullptr;
    p.xw = xw; p.h_state = h; p.c_state = c; p.out = out;
    for (int i = 0; i < 5; ++i) { p.act[i] = r->act[i]; p.act_a[i] = r->act_a[i]; }
    const int slices = (nb + CT_BM - 1) / CT_BM;
    const int S = NNTK_ACT_SIGMOID, TH = NNTK_ACT_TANH;
    const bool def = (G == 3) ? (p.act[0] == S && p.act[1] == S && p.act[2] == TH)
                              : (p.act[0] == S && p.act[1] == S && p.act[2] == TH && p.act[3] == S && p.act[4] == TH);
    int rc;
    if (G == 3) rc = def ? launch_ct<3, true>(mu, p, slices, smem, st) : launch_ct<3, false>(mu, p, slices, smem, st);
    else rc = def ? launch_ct<4, true>(mu, p, slices, smem, st) : launch_ct<4, false>(mu, p, slices, smem, st);
    if (rc) return rc;
    *used = 1;
    return 0;
}
