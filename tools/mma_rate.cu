// Issue-rate probe for the tcgen05 k-block used by K2's tensor-core layers: one thread per CTA
// issues, per k-block of 32 floats, four k-steps of [M=128,N=64 (A_hi x B_hi|B_lo)] + [M=128,N=32
// (A_lo x B_hi)] kind::tf32 MMAs on operands already in shared memory (contents irrelevant), with
// no producer work at all.  Reports cycles per k-block for several shared-memory layouts, i.e. the
// tensor-core floor the staging pipeline has to keep up with.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
        : "memory");
}

// mode 0: no swizzle, LBO = lbo bytes, SBO = 8*lbo;  mode 1: SWIZZLE_128B (rows of 128 B, SBO = 1024)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, int mode, int lbo) {
    if (mode == 0)
        return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)((8 * lbo) >> 4) << 32) |
               ((uint64_t)1 << 46);
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}

__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred;
}
__device__ __forceinline__ void mma_ts_e(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t elected) {
    asm volatile(
        "{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 e, %5, 0;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc), "r"(elected) : "memory");
}
// K2's k-block, issued by a whole warp with uniform operands (one elected lane): four N = n_main MMAs into rotating
// accumulators, then (with_cross) four N = 32 MMAs into their upper halves; A from tensor memory
__global__ void __launch_bounds__(256) mma_rate_uniform_kernel(int nkb, int n_main, int with_cross, int commit_every, int fence_every, long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar, bar2;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 24576; i += 256) reinterpret_cast<float*>(smem)[i] = 1.0f;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar2)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    const uint64_t bh = make_desc(smem_u32(smem), 0, 144);
    const uint32_t idm = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n_main >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t id32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    if (warp == 7) {
        const uint32_t el = elect_one();
        long long t0 = clock64();
        for (int kb = 0; kb < nkb; ++kb) {
            const uint32_t ta = tmem + 256 + 64 * (kb & 3);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) mma_ts_e(tmem + 64 * ks, ta + 8 * ks, bh + ks * 18, idm, kb ? 1u : 0u, el);
            if (with_cross) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) mma_ts_e(tmem + 64 * ks + 32, ta + 32 + 8 * ks, bh + ks * 18, id32, 1u, el);
            }
            if (commit_every) {                              // predicated commit: no divergent branch in the issuing warp
                const uint32_t ec = ((kb % commit_every) == commit_every - 1) ? el : 0u;
                asm volatile("{\n\t.reg .pred e;\n\tsetp.ne.b32 e, %1, 0;\n\t"
                             "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(&bar2)), "r"(ec) : "memory");
            }
            if (fence_every && (kb % fence_every) == fence_every - 1) {
                __syncwarp();
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
        }
        long long t1 = clock64();
        if (el) {
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        }
        unsigned ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
        long long t2 = clock64();
        if (blockIdx.x == 0 && el) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// rotate: main accumulator = k-step (not k-block), cross accumulator alternates; ts: A operand from tensor memory
__global__ void __launch_bounds__(256) mma_rate_kernel(int mode, int lbo, int nkb, int n_main, int with_cross, int rotate, int ts,
                                                       long long* out) {
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < 24576; i += 256) reinterpret_cast<float*>(smem)[i] = 1.0f;      // 96 KB of finite data
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_s;
    const int tile = (mode == 0) ? 16 * 8 * lbo : 16384;               // bytes of one 128-row operand tile
    const uint32_t base = smem_u32(smem);
    const uint64_t ah = make_desc(base, mode, lbo), al = make_desc(base + tile, mode, lbo);
    const uint64_t bh = make_desc(base + 2 * tile, mode, lbo);
    const uint64_t kstep = (mode == 0) ? (uint64_t)((2 * lbo) >> 4) : (uint64_t)(32 >> 4);
    const uint32_t idesc_main = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n_main >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t idesc32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    long long t0 = 0, t1 = 0, t2 = 0;
    if (tid == 0) {
        t0 = clock64();
        for (int kb = 0; kb < nkb; ++kb) {
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                const uint32_t dm = tmem + 64 * (rotate ? ks : (kb & 3)), dc = tmem + 256 + (rotate ? 32 * (ks & 1) : 0);
                if (ts) {
                    mma_ts(dm, tmem + 320 + 8 * ks, bh + ks * kstep, idesc_main, (kb | ks) ? 1u : 0u);
                    if (with_cross) mma_ts(dc, tmem + 352 + 8 * ks, bh + ks * kstep, idesc32, (kb | ks) ? 1u : 0u);
                } else {
                    mma_tf32(dm, ah + ks * kstep, bh + ks * kstep, idesc_main, (kb | ks) ? 1u : 0u);
                    if (with_cross) mma_tf32(dc, al + ks * kstep, bh + ks * kstep, idesc32, (kb | ks) ? 1u : 0u);
                }
            }
        }
        t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        unsigned ok = 0;
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
        t2 = clock64();
        if (blockIdx.x == 0) { out[0] = t1 - t0; out[1] = t2 - t0; }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

int main() {
    long long* out;
    cudaMalloc(&out, 16);
    cudaFuncSetAttribute(mma_rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    struct Cfg { int mode, lbo, n_main, cross; const char* name; int rotate, ts; } cfgs[] = {
        {0, 144, 64, 1, "no-swizzle LBO=144 (K2 today), N=64 main + N=32 cross"},
        {0, 128, 64, 1, "no-swizzle LBO=128,            N=64 main + N=32 cross"},
        {1, 0, 64, 1, "SWIZZLE_128B,                  N=64 main + N=32 cross"},
        {0, 144, 64, 0, "no-swizzle LBO=144, N=64 main only"},
        {0, 128, 64, 0, "no-swizzle LBO=128, N=64 main only"},
        {1, 0, 64, 0, "SWIZZLE_128B,       N=64 main only"},
        {0, 128, 32, 0, "no-swizzle LBO=128, N=32 only"},
        {1, 0, 32, 0, "SWIZZLE_128B,       N=32 only"},
        {1, 0, 128, 0, "SWIZZLE_128B,       N=128 only"},
        {1, 0, 256, 0, "SWIZZLE_128B,       N=256 only (reads past B tile: rate only)"},
        {0, 144, 64, 1, "SS LBO=144 main+cross, accumulators rotate per k-step", 1, 0},
        {0, 128, 64, 1, "SS LBO=128 main+cross, accumulators rotate per k-step", 1, 0},
        {1, 0, 64, 1, "SS SWIZZLE_128B main+cross, accumulators rotate", 1, 0},
        {0, 144, 64, 1, "TS (A in TMEM) main+cross, accumulators rotate", 1, 1},
        {0, 144, 64, 1, "TS (A in TMEM) main+cross, one accumulator per k-block", 0, 1},
        {0, 144, 64, 0, "TS (A in TMEM) main only, accumulators rotate", 1, 1},
    };
    const int nkb = 256;
    for (auto& c : cfgs) {
        for (int rep = 0; rep < 2; ++rep) {
            mma_rate_kernel<<<148, 256, 100 * 1024>>>(c.mode, c.lbo, nkb, c.n_main, c.cross, c.rotate, c.ts, out);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("%s: %s\n", c.name, cudaGetErrorString(e)); return 1; }
        }
        long long h[2];
        cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
        printf("%-62s issue %7.1f  complete %7.1f cycles per k-block (32 k)\n", c.name, (double)h[0] / nkb, (double)h[1] / nkb);
    }
    cudaFuncSetAttribute(mma_rate_uniform_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    const int ucfg[][4] = {{64, 1, 0, 0}, {64, 1, 2, 0}, {64, 1, 1, 0}, {64, 1, 2, 2}, {64, 0, 0, 0}, {32, 0, 0, 0}, {128, 0, 0, 0}};
    for (auto& u : ucfg) {
        for (int rep = 0; rep < 2; ++rep) {
            mma_rate_uniform_kernel<<<148, 256, 100 * 1024>>>(nkb, u[0], u[1], u[2], u[3], out);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("uniform: %s\n", cudaGetErrorString(e)); return 1; }
        }
        long long h[2];
        cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
        printf("uniform issue, TS, 4 x N=%-3d %-22s commit every %d, fence every %d blocks: issue %7.1f  complete %7.1f cycles per k-block\n", u[0],
               u[1] ? "+ 4 x N=32 (K2 block)" : "", u[2], u[3], (double)h[0] / nkb, (double)h[1] / nkb);
    }
    return 0;
}
