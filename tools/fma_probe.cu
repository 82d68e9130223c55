// FP32 FMA issue-rate probe: which operand forms reach 128 FMA/clk/SM on sm_100a?
#include <cstdio>
#include <cuda_runtime.h>
__global__ void __launch_bounds__(256) chain_kernel(float* out, const float* in, int iters) {
    float a = in[0], b = in[1];           // registers, not constants
    float x[8];
    for (int u = 0; u < 8; ++u) x[u] = threadIdx.x + u;
    for (int i = 0; i < iters; ++i)
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int u = 0; u < 8; ++u) x[u] = fmaf(x[u], a, b);
    float s = 0; for (int u = 0; u < 8; ++u) s += x[u];
    if (s == 123.456f) out[0] = s;
}
template <int RM, int RN>
__global__ void __launch_bounds__(256) outer_kernel(float* out, const float* in, int iters) {
    float a[RM], b[RN], acc[RM][RN];
    for (int i = 0; i < RM; ++i) a[i] = in[i] + threadIdx.x;
    for (int j = 0; j < RN; ++j) b[j] = in[8 + j] * threadIdx.x;
    for (int i = 0; i < RM; ++i) for (int j = 0; j < RN; ++j) acc[i][j] = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
#pragma unroll
            for (int i = 0; i < RM; ++i)
#pragma unroll
                for (int j = 0; j < RN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            // perturb operands so nothing is hoisted (cheap: RM+RN adds per RM*RN fmas)
#pragma unroll
            for (int i = 0; i < RM; ++i) a[i] += 1e-3f;
#pragma unroll
            for (int j = 0; j < RN; ++j) b[j] -= 1e-3f;
        }
    }
    float s = 0; for (int i = 0; i < RM; ++i) for (int j = 0; j < RN; ++j) s += acc[i][j];
    if (s == 123.456f) out[0] = s;
}
template <typename F> float timeit(F f) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    float best = 1e9;
    for (int r = 0; r < 5; ++r) { cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); best = ms < best ? ms : best; }
    return best;
}
int main() {
    float *out, *in; cudaMalloc(&out, 64); cudaMalloc(&in, 256);
    float h[64]; for (int i = 0; i < 64; ++i) h[i] = 0.5f + 0.01f * i; cudaMemcpy(in, h, 256, cudaMemcpyHostToDevice);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int iters = 2048;
    for (int bps : {1, 2, 4, 8}) {
        int grid = sms * bps;
        float ms = timeit([&] { chain_kernel<<<grid, 256>>>(out, in, iters); });
        printf("chain   blocks/SM %d: %.1f TFLOP/s\n", bps, (double)grid * 256 * iters * 64 * 2 / (ms * 1e-3) / 1e12);
        ms = timeit([&] { outer_kernel<4, 7><<<grid, 256>>>(out, in, iters); });
        printf("outer4x7 blocks/SM %d: %.1f TFLOP/s\n", bps, (double)grid * 256 * iters * 4 * 28 * 2 / (ms * 1e-3) / 1e12);
        ms = timeit([&] { outer_kernel<4, 8><<<grid, 256>>>(out, in, iters); });
        printf("outer4x8 blocks/SM %d: %.1f TFLOP/s\n", bps, (double)grid * 256 * iters * 4 * 32 * 2 / (ms * 1e-3) / 1e12);
        ms = timeit([&] { outer_kernel<8, 8><<<grid, 256>>>(out, in, iters); });
        printf("outer8x8 blocks/SM %d: %.1f TFLOP/s\n", bps, (double)grid * 256 * iters * 4 * 64 * 2 / (ms * 1e-3) / 1e12);
    }
    return 0;
}
