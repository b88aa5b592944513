// probe_cg2.cu -- development probe (not product code): one CTA pair issues tcgen05.mma.cta_group::2 with M = 128 (64 rows per
// CTA) and checks (1) where D[m][n] lands in each CTA's TMEM, (2) that B's N rows are split across the two CTAs' shared memory,
// (3) TMA loads with .cta_group::2 signalling the leader's mbarrier, (4) the multicast tcgen05.commit, and (5) the MMA rate
// (cycles per instruction for several N) -- the facts csrc/recurrent_fused.cu is built on.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I nntoolkitcore_b200/csrc -I include scripts/probe_cg2.cu -o scripts/build/probe_cg2
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cluster_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync() { asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t rank) { uint32_t d; asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(rank)); return d; }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c)); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tWL:\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t@p bra WD;\n\tbra WL;\n\tWD:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// TMA 2-D load, executed by BOTH CTAs of the pair; the transaction bytes are counted on the barrier at cluster address `bar_cl`
__device__ __forceinline__ void tma_load_2d_cg2(void *dst, const CUtensorMap *map, uint32_t bar_cl, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(bar_cl), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void mma_cg2(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void commit_mc2(uint64_t *bar, uint16_t mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {     // K-major SWIZZLE_128B, 8-row groups 1024 B apart
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fff);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
constexpr int NMAX = 256;
struct Params { int N; float *out; long long *cycles; int reps; int nacc; };

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1) probe_kernel(const __grid_constant__ CUtensorMap map_b, Params p) {
    extern __shared__ uint8_t raw[];
    uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t *sA = smem;                       // [64 rows][128 B]
    uint8_t *sB = smem + 8192;                // [N/2 rows][128 B]
    uint64_t *bars = (uint64_t *)(smem + 8192 + NMAX / 2 * 128);
    uint64_t *full = bars, *done = bars + 1;
    uint32_t *slot = (uint32_t *)(bars + 2);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t rank = cluster_rank();
    const int N = p.N;
    if (tid == 0) {
        mbar_init(full, 1); mbar_init(done, 1); mbar_init(bars + 4, 1); mbar_init(bars + 5, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    // A[m][k] = (m % 7) - 3 + (k % 5) as fp16, m = global row (64 rank + i); swizzled 16-byte chunks
    for (int idx = tid; idx < 64 * 8; idx += 128) {
        const int r = idx >> 3, c = idx & 7, m = 64 * (int)rank + r;
        __half v[8];
        for (int j = 0; j < 8; ++j) { const int k = c * 8 + j; v[j] = __float2half((float)((m % 7) - 3 + (k % 5))); }
        *reinterpret_cast<uint4 *>(sA + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<uint4 *>(v);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *slot;
    if (tid == 0) {
        if (rank == 0) mbar_expect_tx(full, (uint32_t)(N * 128));       // both halves
        tma_load_2d_cg2(sB, &map_b, mapa(smem_u32(full), 0), 0, (int)rank * (N / 2));
    }
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    if (rank == 0 && tid == 0) {
        mbar_wait(full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint64_t da = make_desc(smem_u32(sA)), db = make_desc(smem_u32(sB));
        for (int k = 0; k < 4; ++k) mma_cg2(tmem, da + 2 * k, db + 2 * k, idesc, k ? 1u : 0u);
        commit_mc2(done, 3);
    }
    mbar_wait(done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // dump: every warp reads its lane quarter, columns 0 .. 127
    for (int c0 = 0; c0 < 128; c0 += 8) {
        uint32_t u[8];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                     : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7])
                     : "r"(tmem + ((uint32_t)(warp * 32) << 16) + c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int j = 0; j < 8; ++j) p.out[((size_t)rank * 128 + warp * 32 + lane) * 128 + c0 + j] = __uint_as_float(u[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync();
    // rate: `reps` MMAs back to back on the same operands, issued the CUTLASS way: the WHOLE warp runs the loop converged with
    // warp-uniform values (so descriptors live in uniform registers) and only the instruction itself is predicated on elect.sync
    const int warp_u = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const uint32_t rank_u = __shfl_sync(0xffffffffu, rank, 0);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
    if (rank_u == 0 && warp_u == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t sa_u = __shfl_sync(0xffffffffu, smem_u32(sA), 0), sb_u = __shfl_sync(0xffffffffu, smem_u32(sB), 0);
        const uint64_t da = make_desc(sa_u), db = make_desc(sb_u);
        const long long t0 = clock64();
        int a = 0;
        for (int i = 0; i < p.reps; i += 4) {      // round-robin over `nacc` independent accumulators (64 columns apart)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (elect_one()) mma_cg2(tmem_u + 128 + a * 64, da + 2 * k, db + 2 * k, idesc, 1u);
            }
            if (++a == p.nacc) a = 0;
        }
        const long long t1 = clock64();
        if (elect_one()) commit_mc2(done, 3);
        mbar_wait(done, 1);
        const long long t2 = clock64();
        if (lane == 0) { p.cycles[0] = t1 - t0; p.cycles[1] = t2 - t0; }
        // commit cost: the same MMAs with a multicast commit (to a barrier nobody waits on) after every 4 / with a non-multicast one
        __syncwarp();
        const long long t3 = clock64();
        for (int i = 0; i < p.reps; i += 4) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (elect_one()) mma_cg2(tmem_u + 128, da + 2 * k, db + 2 * k, idesc, 1u);
            if (elect_one()) commit_mc2(bars + 4, 3);
        }
        const long long t4 = clock64();
        for (int i = 0; i < p.reps; i += 4) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (elect_one()) mma_cg2(tmem_u + 128, da + 2 * k, db + 2 * k, idesc, 1u);
            if (elect_one()) asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bars + 5)) : "memory");
        }
        const long long t5 = clock64();
        if (lane == 0) { p.cycles[2] = t4 - t3; p.cycles[3] = t5 - t4; }
    } else {
        mbar_wait(done, 1);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    cluster_sync();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}

// ---- single-CTA rate probe: cta_group::1, M = 64 / 128, operands = whatever is in shared memory (timing only)
__device__ __forceinline__ void mma_cg1(uint32_t d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
__global__ void __launch_bounds__(128, 1) probe1_kernel(int M, int N, int reps, int nacc, long long *cycles) {
    extern __shared__ uint8_t raw[];
    uint8_t *smem = (uint8_t *)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    uint8_t *sA = smem, *sB = smem + 16384;
    uint64_t *done = (uint64_t *)(smem + 16384 + 32768);
    uint32_t *slot = (uint32_t *)(done + 1);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (16384 + 32768) / 4; i += 128) ((uint32_t *)smem)[i] = 0;
    if (tid == 0) { mbar_init(done, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *slot;
    const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    const int warp_u = __shfl_sync(0xffffffffu, warp, 0);
    const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
    if (warp_u == 0) {
        const uint32_t sa_u = __shfl_sync(0xffffffffu, smem_u32(sA), 0), sb_u = __shfl_sync(0xffffffffu, smem_u32(sB), 0);
        const uint64_t da = make_desc(sa_u), db = make_desc(sb_u);
        const long long t0 = clock64();
        int a = 0;
        for (int i = 0; i < reps; i += 4) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (elect_one()) mma_cg1(tmem_u + a * 256, da + 2 * k, db + 2 * k, idesc, 1u);
            if (++a == nacc) a = 0;
        }
        if (elect_one()) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(done)) : "memory");
        mbar_wait(done, 0);
        const long long t2 = clock64();
        if ((tid & 31) == 0) cycles[0] = t2 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void *fp = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
    EncodeTiledFn encode = (EncodeTiledFn)fp;
    float *d_out; long long *d_cyc;
    CK(cudaMalloc(&d_out, 2 * 128 * 128 * 4)); CK(cudaMalloc(&d_cyc, 64));
    const int smem = 8192 + NMAX / 2 * 128 + 256 + 1024;
    CK(cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    {
        const int sm1 = 16384 + 32768 + 64 + 1024;
        CK(cudaFuncSetAttribute(probe1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, sm1));
        for (int M : {64, 128}) for (int N : {32, 64, 96, 128, 192, 256}) for (int na : {1, 2}) {
            probe1_kernel<<<1, 128, sm1>>>(M, N, 512, na, d_cyc);
            CK(cudaDeviceSynchronize());
            long long c; CK(cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost));
            printf("cta_group::1 M=%3d N=%3d, %d accumulator(s): %.1f cyc/MMA (model max(M,128)*N/256 = %.0f)\n", M, N, na, (double)c / 512, (M > 128 ? M : 128) * N / 256.0);
        }
    }
    for (int N : {96, 32, 64, 128, 192, 256}) {
        std::vector<__half> B((size_t)N * 64);
        for (int n = 0; n < N; ++n) for (int k = 0; k < 64; ++k) B[(size_t)n * 64 + k] = __float2half((float)((n % 11) - 5 + (k % 3)));
        __half *d_B; CK(cudaMalloc(&d_B, B.size() * 2)); CK(cudaMemcpy(d_B, B.data(), B.size() * 2, cudaMemcpyHostToDevice));
        CUtensorMap mb;
        cuuint64_t dims[2] = {64, (cuuint64_t)N}; cuuint64_t strides[1] = {128}; cuuint32_t box[2] = {64, (cuuint32_t)(N / 2)}; cuuint32_t es[2] = {1, 1};
        CUresult cr = encode(&mb, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, d_B, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) { printf("encode failed %d\n", (int)cr); return 1; }
        CK(cudaMemset(d_out, 0xff, 2 * 128 * 128 * 4));
        Params p{N, d_out, d_cyc, 512, 1};
        probe_kernel<<<2, 128, smem>>>(mb, p);
        CK(cudaDeviceSynchronize());
        if (N <= 128) for (int na = 2; na <= 6; ++na) {
            Params q{N, d_out, d_cyc, 480, na};
            long long c2[2];
            probe_kernel<<<2, 128, smem>>>(mb, q);
            CK(cudaDeviceSynchronize());
            CK(cudaMemcpy(c2, d_cyc, 16, cudaMemcpyDeviceToHost));
            printf("   N=%3d round-robin over %d accumulators: %.1f cyc/MMA\n", N, na, (double)c2[1] / q.reps);
            if (na == 2) { long long c4[4]; CK(cudaMemcpy(c4, d_cyc, 32, cudaMemcpyDeviceToHost));
                printf("   N=%3d 4 MMAs + multicast commit: %.1f cyc per group; + plain commit: %.1f cyc per group (4 MMAs alone: %.1f)\n", N, (double)c4[2] / (q.reps / 4), (double)c4[3] / (q.reps / 4), 4.0 * c2[1] / q.reps); }
            p.nacc = 1;
            probe_kernel<<<2, 128, smem>>>(mb, p);      // leave the single-accumulator result in d_out / d_cyc for the report below
            CK(cudaDeviceSynchronize());
        }
        std::vector<float> out(2 * 128 * 128); long long cyc[2];
        CK(cudaMemcpy(out.data(), d_out, out.size() * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(cyc, d_cyc, 16, cudaMemcpyDeviceToHost));
        // expectation under the assumed layout: CTA r, lane L, column c: row m = 64 r + (L % 64), n = (L / 64) * (N / 2) + c
        int bad = 0, shown = 0;
        for (int r = 0; r < 2; ++r) for (int L = 0; L < 128; ++L) for (int c = 0; c < N / 2; ++c) {
            const int m = 64 * r + (L % 64), n = (L / 64) * (N / 2) + c;
            float want = 0;
            for (int k = 0; k < 64; ++k) want += (float)((m % 7) - 3 + (k % 5)) * (float)((n % 11) - 5 + (k % 3));
            const float got = out[((size_t)r * 128 + L) * 128 + c];
            if (got != want) { ++bad; if (shown++ < 6) printf("  N=%d cta %d lane %d col %d: got %g want %g (m %d n %d)\n", N, r, L, c, got, want, m, n); }
        }
        printf("N=%3d layout %s (%d mismatches); %d MMAs: issue %lld cyc (%.1f/MMA), complete %lld cyc (%.1f/MMA; model N/4 = %.0f)\n", N,
               bad ? "MISMATCH" : "ok", bad, p.reps, cyc[0], (double)cyc[0] / p.reps, cyc[1], (double)cyc[1] / p.reps, N / 4.0);
        cudaFree(d_B);
    }
    return 0;
}
