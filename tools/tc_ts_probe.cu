// Probe of the "A operand through tensor memory" pipeline for K2's wide layers (forward shape):
//   D[n][m] = sum_k W[n][k] X[m][k],   n < N <= 128 weight rows (MMA M = 128), m < 32 samples, 3xTF32.
// Weights:  global --TMA 2D tile (SWIZZLE_128B, OOB rows/columns zero-filled)--> raw fp32 tile in shared memory
//           --LDS (thread = row), hi/lo split in registers--> tcgen05.st --> A_hi | A_lo in TENSOR MEMORY.
// Samples:  the small operand, split and staged in shared memory (K-major, no swizzle): B = [X_hi | X_lo].
// MMAs:     per k-step (8 k)  main[ks] += A_hi x [X_hi | X_lo]  (N = 64)   and   cross[ks&1] += A_lo x X_hi  (N = 32)
// so consecutive MMAs never hit the same accumulator (a dependent MMA costs ~90 cycles, tools/mma_rate.cu).
// Checks the result against fp64 and reports cycles per k-block with all 148 SMs running the same loop.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tc_ts_probe tc_ts_probe.cu
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

constexpr int NS = 4;                                    // raw weight stages (TMA lead)
constexpr int NSA = 4;                                   // A (tensor memory) / B (shared memory) operand stages
constexpr int KB = 32;                                   // k-block (floats)
constexpr int RAW_BYTES = 128 * KB * 4;                  // 16 KB raw weight tile
constexpr int LBO = 144, SBO = 8 * LBO;                  // B tiles: K-major no-swizzle, padded chunk stride
constexpr int BT_BYTES = 4 * SBO;                        // one 32-row B tile (hi or lo)
constexpr int TM_MAIN = 0, TM_A = 256;                   // TMEM columns: 4 x 64 accumulators [hi*hi | hi*lo + lo*hi], NSA x 64 A stages

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float rn_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
__device__ __forceinline__ void split(float x, uint32_t& hi, uint32_t& lo) {
    const float h = rn_tf32(x);
    hi = __float_as_uint(h);
    lo = __float_as_uint(x - h);                          // exact (<= 13 bits); the tensor core truncates it to tf32
}
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n));
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, unsigned parity) {
    unsigned ok = 0;
    while (!ok)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ unsigned mbar_test(uint64_t* b, unsigned parity) {
    unsigned ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(b)), "r"(parity) : "memory");
    return ok;
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ uint64_t b_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)(LBO >> 4) << 16) | ((uint64_t)(SBO >> 4) << 32) | ((uint64_t)1 << 46);
}
// D[tmem] (+)= A[tmem] * B[smem desc].  Executed by a WHOLE warp with warp-uniform operands; `elected` (from
// elect.sync) predicates the instruction, so ptxas keeps the operands in uniform registers and emits the
// MMAs back to back (a lane-0 branch makes it wrap every MMA in a per-thread loop: ~60 cycles each)
__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred;
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t elected) {
    asm volatile(
        "{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 e, %5, 0;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc), "r"(elected) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
                 "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                   "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
                   "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}

// Roles: warps 0-3 transform the weights (warp = TMEM lane quadrant, thread = weight row), warps 4-6 stage the
// samples, warp 7 issues the MMAs and the TMA refills.  Nobody waits for anybody inside a k-block except through
// the three barrier rings.
__global__ void __launch_bounds__(256, 1) ts_fwd_kernel(const __grid_constant__ CUtensorMap wmap, const float* __restrict__ X,
                                                        float* __restrict__ D, int N, int K, long long* clocks) {
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* raw = smem;                                           // NS x 16 KB, 1024-byte aligned (swizzle atom)
    unsigned char* bst = smem + NS * RAW_BYTES;                          // NSA x [X_hi tile | X_lo tile]
    __shared__ uint64_t raw_full[NS], raw_free[NS], a_full[NSA], mma_done[NSA];
    __shared__ uint32_t tmem_slot;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (tid == 0) {
        for (int s = 0; s < NS; ++s) { mbar_init(&raw_full[s], 1); mbar_init(&raw_free[s], 4); }
        for (int s = 0; s < NSA; ++s) { mbar_init(&a_full[s], 5); mbar_init(&mma_done[s], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_slot;
    const uint32_t elected = elect_one();
    const int nkb = (K + KB - 1) / KB;
    // the samples live in shared memory, as K2's activations do
    float* xs = reinterpret_cast<float*>(smem + NS * RAW_BYTES + NSA * 2 * BT_BYTES);
    const int ldx = ((K + 3) & ~3) + ((((K + 3) >> 2) & 1) ? 0 : 4);   // odd number of 16-byte units, like K2's activations
    for (int e = tid; e < 32 * K; e += 256) xs[(e / K) * ldx + (e % K)] = X[e];
    __syncthreads();
    const uint32_t idesc64 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t idesc32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    long long t0 = clock64();
    long long ta_[6] = {0, 0, 0, 0, 0, 0}, tc_ = t0;
#ifdef PROBE_TRACE
#define TR(role, kb, ev) do { if (blockIdx.x == 0 && lane == 0) clocks[64 + ((role) * 32 + (kb)) * 4 + (ev)] = clock64() - t0; } while (0)
#else
#define TR(role, kb, ev) do { } while (0)
#endif
#ifdef PROBE_TIMING
#define TACC(i) do { } while (0)
#define TSET() do { } while (0)
#else
#define TACC(i) do { } while (0)
#define TSET() do { } while (0)
#endif
    if (warp < 4) {
        // ---------------- weight transform, two k-blocks per loop iteration ----------------
        const int r = 32 * warp + lane;
        const int sw = r & 7;
        float4 x0[8], x1[8];
        auto load_row = [&](float4 (&x)[8], int kb) {
            const unsigned char* rowp = raw + (kb % NS) * RAW_BYTES + r * 128;
#pragma unroll
            for (int c = 0; c < 8; ++c) x[c] = *reinterpret_cast<const float4*>(rowp + ((c ^ sw) << 4));
            // the values are consumed by the split before this warp's next arrival on the same slot; signalling
            // right after the loads were ISSUED would be too early, so the arrival follows the first use (below)
        };
        auto free_slot = [&](int kb) { __syncwarp(); if (lane == 0) mbar_arrive(&raw_free[kb % NS]); };
        auto split_store = [&](const float4 (&x)[8], int kb) {
#ifdef PROBE_EMPTY_STAGES
            { const int sa = kb % NSA; if (kb >= NSA) mbar_wait(&mma_done[sa], ((kb / NSA) & 1u) ^ 1u); return; }
#endif
            uint32_t hi[32], lo[32];
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                split(x[c].x, hi[4 * c + 0], lo[4 * c + 0]); split(x[c].y, hi[4 * c + 1], lo[4 * c + 1]);
                split(x[c].z, hi[4 * c + 2], lo[4 * c + 2]); split(x[c].w, hi[4 * c + 3], lo[4 * c + 3]);
            }
            const int sa = kb % NSA;
            if (kb >= NSA) mbar_wait(&mma_done[sa], ((kb / NSA) & 1u) ^ 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16) + TM_A + 64 * sa;
#ifndef PROBE_NO_ST
            tmem_st32(ta, hi);
            tmem_st32(ta + 32, lo);
#else
            if (hi[0] == 0x12345u && lo[5] == 0x777u) tmem_st32(ta, hi);
#endif
        };
        mbar_wait(&raw_full[0], 0u);
        load_row(x0, 0);
        if (nkb > 1) { mbar_wait(&raw_full[1 % NS], 0u); load_row(x1, 1); }
        for (int kb = 0; kb < nkb; kb += 2) {
            if (warp == 0) TR(0, kb, 0);
            const bool two = kb + 1 < nkb;
            split_store(x0, kb);
            free_slot(kb);
            if (kb + 2 < nkb) { mbar_wait(&raw_full[(kb + 2) % NS], ((kb + 2) / NS) & 1u); load_row(x0, kb + 2); }
            if (two) { split_store(x1, kb + 1); free_slot(kb + 1); }
            if (kb + 3 < nkb) { mbar_wait(&raw_full[(kb + 3) % NS], ((kb + 3) / NS) & 1u); load_row(x1, kb + 3); }
            if (warp == 0) TR(0, kb, 1);
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&a_full[kb % NSA]);
                if (two) mbar_arrive(&a_full[(kb + 1) % NSA]);
            }
            if (warp == 0) TR(0, kb, 3);
        }
    } else if (warp == 6) {
        // ---------------- TMA producer: refill a raw slot as soon as the four transform warps have consumed it ----------------
        if (lane == 0) {
            for (int kb = 0; kb < NS && kb < nkb; ++kb) {
                mbar_expect_tx(&raw_full[kb], RAW_BYTES);
                tma_2d(raw + kb * RAW_BYTES, &wmap, kb * KB, 0, &raw_full[kb]);
            }
            for (int kb = 0; kb + NS < nkb; ++kb) {
                const int s = kb % NS;
                mbar_wait(&raw_free[s], (kb / NS) & 1u);
                mbar_expect_tx(&raw_full[s], RAW_BYTES);
                tma_2d(raw + s * RAW_BYTES, &wmap, (kb + NS) * KB, 0, &raw_full[s]);
            }
        }
        __syncwarp();
    } else if (warp < 6) {
        // ---------------- sample staging: the two warps alternate blocks ----------------
        const int kc = lane & 7, m0 = lane >> 3;
        for (int kb = warp - 4; kb < nkb; kb += 2) {
            const int sa = kb % NSA;
            TR(1, kb, 0);
            if (kb >= NSA) mbar_wait(&mma_done[sa], ((kb / NSA) & 1u) ^ 1u);
            TR(1, kb, 1);
            unsigned char* bh = bst + sa * 2 * BT_BYTES;
            const int k = kb * KB + 4 * kc;
#ifndef PROBE_EMPTY_STAGES
            float4 xv[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                xv[i] = (k < K) ? *reinterpret_cast<const float4*>(xs + (m0 + 4 * i) * ldx + k) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int m = m0 + 4 * i;
                uint32_t hh[4], ll[4];
                split(xv[i].x, hh[0], ll[0]); split(xv[i].y, hh[1], ll[1]); split(xv[i].z, hh[2], ll[2]); split(xv[i].w, hh[3], ll[3]);
                const uint32_t o = (m >> 3) * SBO + kc * LBO + (m & 7) * 16;
                *reinterpret_cast<uint4*>(bh + o) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
                *reinterpret_cast<uint4*>(bh + BT_BYTES + o) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
            }
#endif
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&a_full[sa]);
            TR(1, kb, 3);
        }
    } else {
        // ---------------- issuer: two k-blocks per loop iteration ----------------
        for (int kb = 0; kb < nkb; kb += 2) {
            const bool two = kb + 1 < nkb;
            const int sa0 = kb % NSA, sa1 = (kb + 1) % NSA;
            TR(2, kb, 0);
            mbar_wait(&a_full[sa0], (kb / NSA) & 1u);
            if (two) mbar_wait(&a_full[sa1], ((kb + 1) / NSA) & 1u);
            TR(2, kb, 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t bd0 = b_desc(smem_u32(bst + sa0 * 2 * BT_BYTES)), bd1 = b_desc(smem_u32(bst + sa1 * 2 * BT_BYTES));
            const uint32_t ta0 = tmem + TM_A + 64 * sa0, ta1 = tmem + TM_A + 64 * sa1;
            const uint32_t e1 = two ? elected : 0u;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                mma_ts(tmem + TM_MAIN + 64 * ks, ta0 + 8 * ks, bd0 + ks * ((2 * LBO) >> 4), idesc64, kb ? 1u : 0u, elected);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                mma_ts(tmem + TM_MAIN + 64 * ks + 32, ta0 + 32 + 8 * ks, bd0 + ks * ((2 * LBO) >> 4), idesc32, 1u, elected);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                mma_ts(tmem + TM_MAIN + 64 * ks, ta1 + 8 * ks, bd1 + ks * ((2 * LBO) >> 4), idesc64, 1u, e1);
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
                mma_ts(tmem + TM_MAIN + 64 * ks + 32, ta1 + 32 + 8 * ks, bd1 + ks * ((2 * LBO) >> 4), idesc32, 1u, e1);
            if (elected) {
                // one commit covers both blocks (it arrives when every MMA issued so far is done)
                asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mma_done[sa0])) : "memory");
                if (two) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&mma_done[sa1])) : "memory");
            }
            __syncwarp();
            TR(2, kb, 3);
        }
    }
    // the last commit covers every MMA; only the issuer has seen every earlier phase of that barrier (a warp that
    // skipped phases could pass a parity wait early), so it waits and the others meet it at the block barrier
    if (warp == 7) mbar_wait(&mma_done[(nkb - 1) % NSA], ((nkb - 1) / NSA) & 1u);
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    long long t1 = clock64();
    // epilogue: lane = weight row, this warp's 16 samples; small terms first
    {
        const int q = warp & 3, h = warp >> 2, r = 32 * q + lane;
        const int c0 = 16 * h;
        float sum[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) sum[j] = 0.f;
        const uint32_t tb = tmem + ((uint32_t)(32 * q) << 16);
        for (int t = 0; t < 8; ++t) {
            const uint32_t col = t < 4 ? TM_MAIN + 64 * t + 32 : TM_MAIN + 64 * (t - 4);
            uint32_t v[16];
            tmem_ld16(tb + col + c0, v);
#pragma unroll
            for (int j = 0; j < 16; ++j) sum[j] += __uint_as_float(v[j]);
        }
        if (r < N && blockIdx.x == 0)
            for (int j = 0; j < 16; ++j) D[(size_t)r * 32 + c0 + j] = sum[j];
    }
    if (blockIdx.x == 0 && lane == 0 && (warp == 0 || warp == 4 || warp == 7)) {
        long long* o = clocks + (warp == 0 ? 0 : (warp == 4 ? 8 : 16));
        o[0] = t1 - t0; o[1] = nkb;
        for (int i = 0; i < 5; ++i) o[2 + i] = ta_[i];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    EncodeTiledFn encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &qres) != cudaSuccess || !encode) {
        printf("cuTensorMapEncodeTiled not available\n");
        return 1;
    }
    long long* dclk;
    cudaMalloc(&dclk, 8 * (64 + 3 * 32 * 4)); cudaMemset(dclk, 0, 8 * (64 + 3 * 32 * 4));
    const size_t smem = NS * RAW_BYTES + NSA * 2 * BT_BYTES + 32 * 516 * 4 + 1024;
    cudaFuncSetAttribute(ts_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int cfg[][2] = {{100, 500}, {100, 100}, {128, 32}, {64, 256}, {100, 288}, {100, 320}, {100, 384}, {100, 416}, {100, 512}};
    for (auto& c : cfg) {
        const int N = c[0], K = c[1];
        std::vector<float> hW((size_t)N * K), hX((size_t)32 * K), hD((size_t)128 * 32);
        srand(N * 1000 + K);
        for (auto& x : hW) x = (rand() / (float)RAND_MAX - 0.5f) * 2.f;
        for (auto& x : hX) x = (rand() / (float)RAND_MAX - 0.5f) * 2.f;
        float *dW, *dX, *dD;
        cudaMalloc(&dW, hW.size() * 4); cudaMalloc(&dX, hX.size() * 4); cudaMalloc(&dD, hD.size() * 4);
        cudaMemcpy(dW, hW.data(), hW.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dX, hX.data(), hX.size() * 4, cudaMemcpyHostToDevice);
        cudaMemset(dD, 0xff, hD.size() * 4);
        CUtensorMap map;
        const cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)N};
        const cuuint64_t gstr[1] = {(cuuint64_t)K * 4};
        const cuuint32_t box[2] = {KB, 128};
        const cuuint32_t estr[2] = {1, 1};
        CUresult cr = encode(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dW, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                             CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (cr != CUDA_SUCCESS) { printf("encode failed: %d\n", (int)cr); return 1; }
        if (K == 500)
            for (int grid : {1, 8, 32, 74}) {
                for (int rep = 0; rep < 3; ++rep) ts_fwd_kernel<<<grid, 256, smem>>>(map, dX, dD, N, K, dclk);
                cudaDeviceSynchronize();
                long long c2[2];
                cudaMemcpy(c2, dclk, 16, cudaMemcpyDeviceToHost);
                printf("    grid %3d: k-loop %lld cycles = %.1f per k-block\n", grid, c2[0], (double)c2[0] / c2[1]);
            }
        for (int rep = 0; rep < 3; ++rep) {
            ts_fwd_kernel<<<148, 256, smem>>>(map, dX, dD, N, K, dclk);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("N=%d K=%d CUDA error: %s\n", N, K, cudaGetErrorString(e)); return 1; }
        }
        long long clk[32];
        cudaMemcpy(clk, dclk, 192, cudaMemcpyDeviceToHost);
        cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
        double maxerr = 0, maxref = 0, maxerr32 = 0;
        for (int n = 0; n < N; ++n)
            for (int m = 0; m < 32; ++m) {
                double ref = 0; float r32 = 0.f;
                for (int k = 0; k < K; ++k) { ref += (double)hW[(size_t)n * K + k] * hX[(size_t)m * K + k]; r32 = fmaf(hW[(size_t)n * K + k], hX[(size_t)m * K + k], r32); }
                maxerr = fmax(maxerr, fabs(hD[(size_t)n * 32 + m] - ref));
                maxerr32 = fmax(maxerr32, fabs((double)r32 - ref));
                maxref = fmax(maxref, fabs(ref));
            }
        printf("N=%3d K=%3d  max|D-ref| = %.3e (fp32 fma loop %.3e, max|ref| %.2f)   k-loop %lld cycles = %.1f per k-block (%lld blocks)\n",
               N, K, maxerr, maxerr32, maxref, clk[0], (double)clk[0] / clk[1], clk[1]);
#ifdef PROBE_TRACE
        if (K == 500) {
            std::vector<long long> tr(64 + 3 * 32 * 4);
            cudaMemcpy(tr.data(), dclk, tr.size() * 8, cudaMemcpyDeviceToHost);
            printf("  kb |  A: start split+lds  md-seen arrived |  B: start md-seen arrived |  I: start af-seen committed\n");
            for (int kb = 0; kb < 16; ++kb) {
                const long long* a = &tr[64 + (0 * 32 + kb) * 4], *b = &tr[64 + (1 * 32 + kb) * 4], *i = &tr[64 + (2 * 32 + kb) * 4];
                printf("  %2d | %8lld %8lld %8lld %8lld | %8lld %8lld %8lld | %8lld %8lld %8lld\n", kb, a[0], a[1], a[2], a[3], b[0], b[1], b[3], i[0], i[1], i[3]);
            }
        }
#endif
        for (int t = 0; t < 3; ++t) printf("    role %d (0 = A transform, 1 = B staging, 2 = issuer) per k-block: [A: wait raw + lds] %.0f, wait mma_done %.0f, [B: stage+arrive] %.0f, [A: polls+split | I: wait a_full] %.0f, [A: st+wait::st+arrive | I: tma+mma+commit] %.0f\n", t, (double)clk[8 * t + 2] / clk[1], (double)clk[8 * t + 3] / clk[1], (double)clk[8 * t + 4] / clk[1], (double)clk[8 * t + 5] / clk[1], (double)clk[8 * t + 6] / clk[1]);
        cudaFree(dW); cudaFree(dX); cudaFree(dD);
    }
    return 0;
}
