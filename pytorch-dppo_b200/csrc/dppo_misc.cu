// version / error strings of the C ABI
#include "common.cuh"

extern "C" int dppo_version(void) { return 100; }   // 0.1.0

extern "C" const char* dppo_error_string(int code) {
    switch (code) {
        case 0: return "ok";
        case DPPO_E_BADARG: return "dppo: bad argument (null pointer, negative size or unknown flavour)";
        case DPPO_E_TOOLARGE: return "dppo: dimension larger than the kernels support";
        case DPPO_E_SCRATCH: return "dppo: scratch buffer too small (see *_scratch_bytes)";
        case DPPO_E_ALIGN: return "dppo: pointer not 16-byte aligned";
        case DPPO_E_COMM: return "dppo: communicator misuse";
        case DPPO_E_NOALLOC_LIMIT:
            return "dppo: size exceeds the built-in scratch of this *_noalloc entry point (dppo_gae_scan_noalloc: "
                   "T <= 4194304 steps); call the variant that takes caller-provided scratch";
        default: break;
    }
    if (code > 0) return cudaGetErrorString((cudaError_t)code);
    return "dppo: unknown error";
}

// ---------------------------------------------------------------------------
// FP32 FMA throughput probe (bench only): the fused update kernel is bound by
// the FP32 pipe, and MEASURED_PEAKS.json has no fp32 number, so bench.py measures
// one: 8 independent FMA chains per thread, grid = 8 CTAs per SM.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fp32_probe_kernel(float* out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            x0 = fmaf(x0, a, b); x1 = fmaf(x1, a, b); x2 = fmaf(x2, a, b); x3 = fmaf(x3, a, b);
            x4 = fmaf(x4, a, b); x5 = fmaf(x5, a, b); x6 = fmaf(x6, a, b); x7 = fmaf(x7, a, b);
        }
    }
    const float s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 123.456f) out[0] = s;                  // keeps the chains alive
}

extern "C" int dppo_fp32_probe(int iters, float* out, double* flops_host, void* stream) {
    if (iters < 1 || !out || !flops_host) return DPPO_E_BADARG;
    const int grid = dppo_sm_count() * 8;
    fp32_probe_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(out, iters, 0.999f, 0.001f);
    DPPO_LAUNCH_CHECK();
    *flops_host = (double)grid * 256.0 * (double)iters * 64.0 * 2.0;
    return 0;
}
