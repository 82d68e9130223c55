// Standalone check of the tcgen05 building blocks used by K2's tensor-core layers:
//   D[128 x 32] = A[128 x K] * B[32 x K]^T   with A, B fp32 in global memory (K-major rows),
// computed as 3xTF32 (hi*hi + hi*lo + lo*hi) with tcgen05.mma kind::tf32, operands staged into the
// canonical K-major no-swizzle shared-memory layout, accumulator in TMEM, read back with tcgen05.ld.
// Compares against an fp64 CPU reference.   nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include <vector>
#include <cuda_runtime.h>

constexpr int M = 128, N = 32, KB = 32;                 // k-block of 32 floats = 8 16-byte chunks

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, SWIZZLE_NONE ("interleave") canonical layout of a [rows][KB] tile:
//   16-byte chunk (r, kc) at  (r/8)*SBO + kc*LBO + (r%8)*16,  LBO = 128, SBO = (KB/4)*128
constexpr int LBO = 144, SBO = (KB / 4) * LBO;          // padded: chunk stores with kc fastest are bank-conflict free
__device__ __forceinline__ uint32_t tile_off(int r, int kc) { return (r >> 3) * SBO + kc * LBO + (r & 7) * 16; }

__device__ __forceinline__ uint64_t make_desc(uint32_t saddr) {
    const uint64_t lbo = LBO >> 4, sbo = SBO >> 4;
    return (uint64_t)((saddr >> 4) & 0x3fff) | (lbo << 16) | (sbo << 32) | ((uint64_t)1 << 46);   // version 1, no swizzle
}

__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    uint32_t h, l;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));         // round to nearest tf32 (low 13 bits zero)
    hi = __uint_as_float(h);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(x - hi));    // the remainder, again rounded: |x-hi-lo| <= 2^-24 |x|
    lo = __uint_as_float(l);
}

__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %6, %7, %8}, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(0), "r"(0), "r"(0), "r"(0)
        : "memory");
}

__global__ void __launch_bounds__(256) tc_probe_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ D,
                                                       int K) {
    extern __shared__ __align__(1024) unsigned char smem[];
    float* sAh = reinterpret_cast<float*>(smem);                 // [128][KB] canonical, 16 KB
    float* sAl = sAh + (M / 8) * SBO / 4;
    float* sBh = sAl + (M / 8) * SBO / 4;                        // [32][KB]
    float* sBl = sBh + (N / 8) * SBO / 4;
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_s;
    // instruction descriptor: D=F32, A=B=TF32, both K-major, N=32, M=128
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    unsigned phase = 0;
    const int nkb = (K + KB - 1) / KB;
    for (int kb = 0; kb < nkb; ++kb) {
        // stage A (128 x 8 chunks = 1024 chunks, 4 per thread) and B (32 x 8 = 256 chunks, 1 per thread)
        for (int e = tid; e < M * (KB / 4); e += 256) {
            const int r = e >> 3, kc = e & 7, k = kb * KB + 4 * kc;
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k < K) x = *reinterpret_cast<const float4*>(A + (size_t)r * K + k);      // K % 4 == 0
            float4 h, l;
            split_tf32(x.x, h.x, l.x); split_tf32(x.y, h.y, l.y); split_tf32(x.z, h.z, l.z); split_tf32(x.w, h.w, l.w);
            *reinterpret_cast<float4*>(reinterpret_cast<char*>(sAh) + tile_off(r, kc)) = h;
            *reinterpret_cast<float4*>(reinterpret_cast<char*>(sAl) + tile_off(r, kc)) = l;
        }
        {
            const int r = tid >> 3, kc = tid & 7, k = kb * KB + 4 * kc;
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (k < K) x = *reinterpret_cast<const float4*>(B + (size_t)r * K + k);
            float4 h, l;
            split_tf32(x.x, h.x, l.x); split_tf32(x.y, h.y, l.y); split_tf32(x.z, h.z, l.z); split_tf32(x.w, h.w, l.w);
            *reinterpret_cast<float4*>(reinterpret_cast<char*>(sBh) + tile_off(r, kc)) = h;
            *reinterpret_cast<float4*>(reinterpret_cast<char*>(sBl) + tile_off(r, kc)) = l;
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic writes -> visible to the tensor core
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            for (int ks = 0; ks < KB / 8; ++ks) {                       // 8 tf32 (32 bytes = 2 chunks) per MMA
                const uint32_t adv = ks * 2 * LBO;
                const uint64_t ah = make_desc(smem_u32(sAh) + adv), al = make_desc(smem_u32(sAl) + adv);
                const uint64_t bh = make_desc(smem_u32(sBh) + adv), bl = make_desc(smem_u32(sBl) + adv);
                // hi*hi into one of 8 accumulators (round robin over k-blocks), the two small cross terms
                // into a ninth: the tensor core truncates when it accumulates, so fewer and smaller
                // additions per accumulator keep the error at the fp32-FMA level
                mma_tf32(tmem_d + 32 * (kb & 7), ah, bh, idesc, (kb >= 8 || ks) ? 1u : 0u);
                mma_tf32(tmem_d + 32 * 8, al, bh, idesc, (kb | ks) ? 1u : 0u);
                mma_tf32(tmem_d + 32 * 8, ah, bl, idesc, 1u);
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        }
        // wait for the MMAs before the stage is overwritten (single-buffered here)
        {
            unsigned ok = 0;
            while (!ok)
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(ok) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
            phase ^= 1;
        }
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // epilogue: warp w reads TMEM lanes 32*(w%4)..+31; warps 0-3 columns 0-15, warps 4-7 columns 16-31
    {
        const int row = 32 * (warp & 3) + lane, c0 = (warp >> 2) * 16;
        float sum[16];
        for (int j = 0; j < 16; ++j) sum[j] = 0.f;
        const int nacc = nkb < 8 ? nkb : 8;
        for (int a = 8; a >= 0; --a) {                          // cross terms first, then the main accumulators
            if (a < 8 && a >= nacc) continue;
            uint32_t v[16];
            const uint32_t taddr = tmem_d + ((uint32_t)(32 * (warp & 3)) << 16) + 32 * a + c0;
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                         : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                           "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                         : "r"(taddr));
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int j = 0; j < 16; ++j) sum[j] += __uint_as_float(v[j]);
        }
        for (int j = 0; j < 16; ++j) D[(size_t)row * N + c0 + j] = sum[j];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512));
}

int main() {
    for (int K : {32, 100, 500}) {
        std::vector<float> hA((size_t)M * K), hB((size_t)N * K), hD((size_t)M * N);
        srand(K);
        for (auto& x : hA) x = (rand() / (float)RAND_MAX - 0.5f) * 2.f;
        for (auto& x : hB) x = (rand() / (float)RAND_MAX - 0.5f) * 2.f;
        float *dA, *dB, *dD;
        cudaMalloc(&dA, hA.size() * 4); cudaMalloc(&dB, hB.size() * 4); cudaMalloc(&dD, hD.size() * 4);
        cudaMemcpy(dA, hA.data(), hA.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dB, hB.data(), hB.size() * 4, cudaMemcpyHostToDevice);
        cudaMemset(dD, 0xff, hD.size() * 4);
        const size_t smem = 2 * (M / 8) * SBO + 2 * (N / 8) * SBO + 1024;
        cudaFuncSetAttribute(tc_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        tc_probe_kernel<<<1, 256, smem>>>(dA, dB, dD, K);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("K=%d CUDA error: %s\n", K, cudaGetErrorString(e)); return 1; }
        cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
        double maxerr = 0, maxref = 0, maxerr32 = 0;
        for (int m = 0; m < M; ++m)
            for (int n = 0; n < N; ++n) {
                double ref = 0; float r32 = 0.f;
                for (int k = 0; k < K; ++k) { ref += (double)hA[(size_t)m * K + k] * hB[(size_t)n * K + k]; r32 = fmaf(hA[(size_t)m * K + k], hB[(size_t)n * K + k], r32); }
                maxerr = fmax(maxerr, fabs(hD[(size_t)m * N + n] - ref));
                maxerr32 = fmax(maxerr32, fabs((double)r32 - ref));
                maxref = fmax(maxref, fabs(ref));
            }
        printf("K=%3d  max|D-ref| = %.3e  (fp32 fma loop: %.3e)  max|ref| = %.3f  rel = %.2e\n", K, maxerr, maxerr32, maxref, maxerr / maxref);
        cudaFree(dA); cudaFree(dB); cudaFree(dD);
    }
    return 0;
}
