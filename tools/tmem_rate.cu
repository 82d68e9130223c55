// Tensor-memory read-out rate: 8 warps (2 per lane quadrant) read COLS columns of their quadrant with
// tcgen05.ld.32x32b.x8 / x16 / x32 (one wait per load, or one wait per batch of four loads).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_rate tmem_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
template <int NC> __device__ __forceinline__ void ld(uint32_t a, uint32_t* v);
template <> __device__ __forceinline__ void ld<8>(uint32_t a, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "r"(a));
}
template <> __device__ __forceinline__ void ld<16>(uint32_t a, uint32_t* v) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(a));
}
template <> __device__ __forceinline__ void ld<32>(uint32_t a, uint32_t* v) { ld<16>(a, v); ld<16>(a + 16, v + 16); }
__device__ __forceinline__ void ldwait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

template <int NC, int BATCH>
__global__ void __launch_bounds__(256) k(int cols, int reps, int nwarps, long long* out, float* sink) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot + ((uint32_t)(32 * (warp & 3)) << 16);
    float acc = 0.f;
    __syncthreads();
    long long t0 = clock64();
    if (warp < nwarps)
        for (int r = 0; r < reps; ++r)
            for (int c = 0; c < cols; c += NC * BATCH) {
                uint32_t v[BATCH][NC];
#pragma unroll
                for (int b = 0; b < BATCH; ++b) ld<NC>(tmem + c + NC * b, v[b]);
                ldwait();
#pragma unroll
                for (int b = 0; b < BATCH; ++b)
#pragma unroll
                    for (int j = 0; j < NC; ++j) acc += __uint_as_float(v[b][j]);
            }
    __syncthreads();
    long long t1 = clock64();
    if (acc == 12345.f) sink[threadIdx.x] = acc;
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512));
}
template <int NC, int BATCH> void run(const char* name, int nwarps, long long* out, float* sink) {
    const int cols = 256, reps = 16;
    for (int i = 0; i < 2; ++i) { k<NC, BATCH><<<148, 256>>>(cols, reps, nwarps, out, sink); cudaDeviceSynchronize(); }
    long long h; cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost);
    const double bytes = (double)reps * cols * 4 * 32 * nwarps;
    printf("%-34s %d warps: %8lld cycles for %.0f KB -> %.1f B/clk/SM\n", name, nwarps, h, bytes / 1024, bytes / h);
}
int main() {
    long long* out; float* sink; cudaMalloc(&out, 8); cudaMalloc(&sink, 1024);
    run<8, 1>("x8, wait each", 8, out, sink);
    run<16, 1>("x16, wait each", 8, out, sink);
    run<32, 1>("2 x x16, one wait", 8, out, sink);
    run<16, 4>("4 x x16, one wait", 8, out, sink);
    run<8, 4>("4 x x8, one wait", 8, out, sink);
    run<16, 4>("4 x x16, one wait", 4, out, sink);
    run<16, 4>("4 x x16, one wait", 1, out, sink);
    return 0;
}
