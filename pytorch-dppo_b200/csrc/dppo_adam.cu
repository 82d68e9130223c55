// K3 (single-GPU part) - deterministic sum of the per-CTA gradient partials,
// optional global-norm clip, and the Adam step, fused into one pass over the
// parameter bucket (28 B/param + 4 B/param per partial).
//
// Reference: Shared_grad_buffers.add_gradient/reset (model.py:53-59), the
// chief's optimizer.step() on the summed gradient (chief.py:13-20),
// clip_grad_norm + optimizer.step (ppo.py:173-174); Adam arithmetic is torch
// 2.11 torch/optim/adam.py _single_tensor_adam:
//     m.lerp_(g, 1-b1)                 -> m = fma(w, g-m, m)
//     v.mul_(b2).addcmul_(g, g, 1-b2)  -> v = fma((1-b2)*g, g, v*b2)
//     denom = v.sqrt()/sqrt(bc2) + eps
//     p.addcdiv_(m, denom, -lr/bc1)    -> p = p + ((-lr/bc1)*m)/denom
// (association pinned bitwise against torch CPU apart from torch's own sqrt,
// which is not correctly rounded).
#include <math.h>
#include "common.cuh"

namespace {

constexpr int ADAM_THREADS = 256;

struct AdamScalars {
    float w1, b2, a2, bc2_sqrt, eps, alpha;
};

__device__ __forceinline__ float adam_one(float g, float& p, float& m, float& v, const AdamScalars& s) {
    m = fmaf(s.w1, __fsub_rn(g, m), m);
    v = fmaf(__fmul_rn(s.a2, g), g, __fmul_rn(v, s.b2));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), s.bc2_sqrt), s.eps);
    p = __fadd_rn(p, __fdiv_rn(__fmul_rn(s.alpha, m), denom));
    return p;
}

__device__ __forceinline__ float4 sum_partials(const float4* __restrict__ partials, int nparts, int n4, int i,
                                               const float4* __restrict__ gshared) {
    float4 G = gshared ? gshared[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    // fixed order p = 0..nparts-1 (bit-reproducible; model.py:53-55 adds in arrival order)
#pragma unroll 4
    for (int p = 0; p < nparts; ++p) {
        const float4 x = __ldcs(&partials[(size_t)p * n4 + i]);
        G.x = __fadd_rn(G.x, x.x); G.y = __fadd_rn(G.y, x.y); G.z = __fadd_rn(G.z, x.z); G.w = __fadd_rn(G.w, x.w);
    }
    return G;
}

__device__ __forceinline__ void reduce_losses(const float* __restrict__ loss_partials, int nparts, float* loss_out) {
    if (loss_partials && loss_out && blockIdx.x == 0 && threadIdx.x < 4) {
        float s = 0.f;
        for (int p = 0; p < nparts; ++p) s = __fadd_rn(s, loss_partials[p * 4 + threadIdx.x]);
        loss_out[threadIdx.x] = s;
    }
}

// ---- many partials (one per K2 CTA): 8 warps of a block each sum every 8th partial for the
// same 32 float4 columns, then the eight group sums are added in fixed order.  8x more loads in
// flight and 8x more blocks than one thread per column walking all partials.
constexpr int RG = 8;                      // partial groups (= warps per block)
constexpr int RQ = 32;                     // float4 columns per block

__device__ __forceinline__ float4 sum_partials_grouped(const float4* __restrict__ partials, int nparts, int n4, int i,
                                                       const float4* __restrict__ gshared, float4 (*red)[RQ]) {
    const int q = threadIdx.x & 31, g = threadIdx.x >> 5;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n4) {
#pragma unroll 6
        for (int p = g; p < nparts; p += RG) {
            const float4 x = __ldcs(&partials[(size_t)p * n4 + i]);
            a.x = __fadd_rn(a.x, x.x); a.y = __fadd_rn(a.y, x.y); a.z = __fadd_rn(a.z, x.z); a.w = __fadd_rn(a.w, x.w);
        }
    }
    red[g][q] = a;
    __syncthreads();
    float4 G = make_float4(0.f, 0.f, 0.f, 0.f);
    if (g == 0 && i < n4) {
        // this GPU's gradient first, the shared bucket's content on top of it: the same association as
        // the multi-GPU kernel with one rank (dppo_comm.cu), so W = 1 is bit-identical to this path
#pragma unroll
        for (int r = 0; r < RG; ++r) {
            const float4 x = red[r][q];
            G.x = __fadd_rn(G.x, x.x); G.y = __fadd_rn(G.y, x.y); G.z = __fadd_rn(G.z, x.z); G.w = __fadd_rn(G.w, x.w);
        }
        if (gshared) {
            const float4 s0 = gshared[i];
            G.x = __fadd_rn(s0.x, G.x); G.y = __fadd_rn(s0.y, G.y); G.z = __fadd_rn(s0.z, G.z); G.w = __fadd_rn(s0.w, G.w);
        }
    }
    return G;
}

__global__ void __launch_bounds__(RG * 32)
reduce_adam_grouped_kernel(const float4* __restrict__ partials, int nparts, float4* __restrict__ gshared,
                           float4* __restrict__ theta, float4* __restrict__ theta_prev, float4* __restrict__ m,
                           float4* __restrict__ v, int n4, AdamScalars s, const AdamScalars* __restrict__ sdev,
                           float4* __restrict__ grad_out, const float* __restrict__ loss_partials,
                           float* __restrict__ loss_out) {
    __shared__ float4 red[RG][RQ];
    if (sdev) s = *sdev;
    reduce_losses(loss_partials, nparts, loss_out);
    const int i = blockIdx.x * RQ + (threadIdx.x & 31);
    const float4 G = sum_partials_grouped(partials, nparts, n4, i, gshared, red);
    if ((threadIdx.x >> 5) != 0 || i >= n4) return;
    float4 p = theta[i], mm = m[i], vv = v[i];
    if (theta_prev) theta_prev[i] = p;
    adam_one(G.x, p.x, mm.x, vv.x, s);
    adam_one(G.y, p.y, mm.y, vv.y, s);
    adam_one(G.z, p.z, mm.z, vv.z, s);
    adam_one(G.w, p.w, mm.w, vv.w, s);
    theta[i] = p; m[i] = mm; v[i] = vv;
    if (grad_out) grad_out[i] = G;
    if (gshared) gshared[i] = make_float4(0.f, 0.f, 0.f, 0.f);          // chief.py:19
}

// clipping, pass 1 (grouped): G -> gtmp, per-block sum of squares (fp64, fixed order)
__global__ void __launch_bounds__(RG * 32)
reduce_sumsq_grouped_kernel(const float4* __restrict__ partials, int nparts, float4* __restrict__ gshared, int n4,
                            float4* __restrict__ gtmp, double* __restrict__ blocksums,
                            const float* __restrict__ loss_partials, float* __restrict__ loss_out) {
    __shared__ float4 red[RG][RQ];
    __shared__ double sq[RQ];
    reduce_losses(loss_partials, nparts, loss_out);
    const int i = blockIdx.x * RQ + (threadIdx.x & 31);
    const float4 G = sum_partials_grouped(partials, nparts, n4, i, gshared, red);
    if ((threadIdx.x >> 5) == 0) {
        double ss = 0.0;
        if (i < n4) {
            gtmp[i] = G;
            if (gshared) gshared[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            ss = (double)G.x * G.x + (double)G.y * G.y + (double)G.z * G.z + (double)G.w * G.w;
        }
        sq[threadIdx.x] = ss;
        __syncwarp();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int k = 0; k < RQ; ++k) t += sq[k];
            blocksums[blockIdx.x] = t;
        }
    }
}

// no clipping: everything in one kernel
__global__ void __launch_bounds__(ADAM_THREADS)
reduce_adam_kernel(const float4* __restrict__ partials, int nparts, float4* __restrict__ gshared,
                   float4* __restrict__ theta, float4* __restrict__ theta_prev, float4* __restrict__ m,
                   float4* __restrict__ v, int n4, AdamScalars s, const AdamScalars* __restrict__ sdev,
                   float4* __restrict__ grad_out, const float* __restrict__ loss_partials,
                   float* __restrict__ loss_out) {
    if (sdev) s = *sdev;
    reduce_losses(loss_partials, nparts, loss_out);
    const int i = blockIdx.x * ADAM_THREADS + threadIdx.x;
    if (i >= n4) return;
    const float4 G = sum_partials(partials, nparts, n4, i, gshared);
    float4 p = theta[i], mm = m[i], vv = v[i];
    if (theta_prev) theta_prev[i] = p;
    adam_one(G.x, p.x, mm.x, vv.x, s);
    adam_one(G.y, p.y, mm.y, vv.y, s);
    adam_one(G.z, p.z, mm.z, vv.z, s);
    adam_one(G.w, p.w, mm.w, vv.w, s);
    theta[i] = p; m[i] = mm; v[i] = vv;
    if (grad_out) grad_out[i] = G;
    if (gshared) gshared[i] = make_float4(0.f, 0.f, 0.f, 0.f);          // chief.py:19
}

// clipping, pass 1: G -> gtmp, per-block sum of squares (fp64, fixed tree)
__global__ void __launch_bounds__(ADAM_THREADS)
reduce_sumsq_kernel(const float4* __restrict__ partials, int nparts, float4* __restrict__ gshared, int n4,
                    float4* __restrict__ gtmp, double* __restrict__ blocksums,
                    const float* __restrict__ loss_partials, float* __restrict__ loss_out) {
    __shared__ double red[ADAM_THREADS];
    reduce_losses(loss_partials, nparts, loss_out);
    const int i = blockIdx.x * ADAM_THREADS + threadIdx.x;
    double ss = 0.0;
    if (i < n4) {
        const float4 G = sum_partials(partials, nparts, n4, i, gshared);
        gtmp[i] = G;
        if (gshared) gshared[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        ss = (double)G.x * G.x + (double)G.y * G.y + (double)G.z * G.z + (double)G.w * G.w;
    }
    red[threadIdx.x] = ss;
    __syncthreads();
    for (int s = ADAM_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) blocksums[blockIdx.x] = red[0];
}

// clipping, pass 2: coef = min(1, max_norm/(||G||+1e-6)) (ppo.py:173), then Adam
__global__ void __launch_bounds__(ADAM_THREADS)
clip_adam_kernel(const float4* __restrict__ gtmp, const double* __restrict__ blocksums, int nblocks, float max_norm,
                 float4* __restrict__ theta, float4* __restrict__ theta_prev, float4* __restrict__ m,
                 float4* __restrict__ v, int n4, AdamScalars s, const AdamScalars* __restrict__ sdev,
                 float4* __restrict__ grad_out, float* __restrict__ norm_out) {
    __shared__ double red[ADAM_THREADS];
    if (sdev) s = *sdev;
    double ss = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += ADAM_THREADS) ss += blocksums[b];
    red[threadIdx.x] = ss;
    __syncthreads();
    for (int st = ADAM_THREADS / 2; st > 0; st >>= 1) {
        if (threadIdx.x < st) red[threadIdx.x] += red[threadIdx.x + st];
        __syncthreads();
    }
    const float total = (float)sqrt(red[0]);
    float coef = __fdiv_rn(max_norm, __fadd_rn(total, 1e-6f));
    coef = fminf(coef, 1.0f);
    if (norm_out && blockIdx.x == 0 && threadIdx.x == 0) *norm_out = total;
    const int i = blockIdx.x * ADAM_THREADS + threadIdx.x;
    if (i >= n4) return;
    float4 G = gtmp[i];
    G.x = __fmul_rn(G.x, coef); G.y = __fmul_rn(G.y, coef); G.z = __fmul_rn(G.z, coef); G.w = __fmul_rn(G.w, coef);
    float4 p = theta[i], mm = m[i], vv = v[i];
    if (theta_prev) theta_prev[i] = p;
    adam_one(G.x, p.x, mm.x, vv.x, s);
    adam_one(G.y, p.y, mm.y, vv.y, s);
    adam_one(G.z, p.z, mm.z, vv.z, s);
    adam_one(G.w, p.w, mm.w, vv.w, s);
    theta[i] = p; m[i] = mm; v[i] = vv;
    if (grad_out) grad_out[i] = G;
}

}  // namespace

// scratch: float gtmp[P_pad]; double blocksums[nblocks]; float norm
extern "C" size_t dppo_reduce_adam_scratch_bytes(int P_pad) {
    if (P_pad <= 0) return 16;
    const int nblocks = ceil_div(P_pad / 4, RQ);                 // the grouped grid is the larger one
    return align_up(sizeof(float) * (size_t)P_pad, 16) + sizeof(double) * (size_t)nblocks + 16;
}

static AdamScalars dppo_adam_scalars(int step, double lr, double beta1, double beta2, double eps) {
    AdamScalars s;
    const double bc1 = 1.0 - pow(beta1, (double)step);
    const double bc2 = 1.0 - pow(beta2, (double)step);
    s.w1 = (float)(1.0 - beta1);
    s.b2 = (float)beta2;
    s.a2 = (float)(1.0 - beta2);
    s.bc2_sqrt = (float)sqrt(bc2);
    s.eps = (float)eps;
    s.alpha = (float)(-(lr / bc1));
    return s;
}

extern "C" int dppo_adam_scalars_host(int step, double lr, double beta1, double beta2, double eps, float* out6_host) {
    if (step < 1 || !out6_host) return DPPO_E_BADARG;
    const AdamScalars s = dppo_adam_scalars(step, lr, beta1, beta2, eps);
    out6_host[0] = s.w1; out6_host[1] = s.b2; out6_host[2] = s.a2;
    out6_host[3] = s.bc2_sqrt; out6_host[4] = s.eps; out6_host[5] = s.alpha;
    return 0;
}

extern "C" int dppo_reduce_adam(const float* partials, int nparts, float* gshared, float* theta, float* theta_prev,
                                float* m, float* v, int P_pad, int step, double lr, double beta1, double beta2,
                                double eps, const float* scalars_dev, double max_grad_norm, float* grad_out,
                                const float* loss_partials, float* loss_out, void* scratch, size_t scratch_bytes,
                                void* stream_) {
    if (P_pad <= 0 || (P_pad & 3) || nparts < 0 || (step < 1 && !scalars_dev)) return DPPO_E_BADARG;
    if (!theta || !m || !v || (nparts > 0 && !partials)) return DPPO_E_BADARG;
    if (!aligned16(partials) || !aligned16(gshared) || !aligned16(theta) || !aligned16(theta_prev) || !aligned16(m) ||
        !aligned16(v) || !aligned16(grad_out))
        return DPPO_E_ALIGN;
    cudaStream_t stream = (cudaStream_t)stream_;
    const int n4 = P_pad / 4;
    const int nblocks = ceil_div(n4, ADAM_THREADS);
    const AdamScalars s = dppo_adam_scalars(step < 1 ? 1 : step, lr, beta1, beta2, eps);
    const bool grouped = nparts > 2;                          // many K2 partials: spread them over 8 warps
    const int gblocks = ceil_div(n4, RQ);
    if (max_grad_norm > 0.0) {
        if (!scratch || scratch_bytes < dppo_reduce_adam_scratch_bytes(P_pad)) return DPPO_E_SCRATCH;
        if (!aligned16(scratch)) return DPPO_E_ALIGN;
        float4* gtmp = (float4*)scratch;
        double* blocksums = (double*)((char*)scratch + align_up(sizeof(float) * (size_t)P_pad, 16));
        const int nsums = grouped ? gblocks : nblocks;
        float* norm = (float*)(blocksums + gblocks);
        if (grouped)
            reduce_sumsq_grouped_kernel<<<gblocks, RG * 32, 0, stream>>>((const float4*)partials, nparts,
                                                                         (float4*)gshared, n4, gtmp, blocksums,
                                                                         loss_partials, loss_out);
        else
            reduce_sumsq_kernel<<<nblocks, ADAM_THREADS, 0, stream>>>((const float4*)partials, nparts, (float4*)gshared,
                                                                       n4, gtmp, blocksums, loss_partials, loss_out);
        DPPO_LAUNCH_CHECK();
        clip_adam_kernel<<<nblocks, ADAM_THREADS, 0, stream>>>(gtmp, blocksums, nsums, (float)max_grad_norm,
                                                                (float4*)theta, (float4*)theta_prev, (float4*)m,
                                                                (float4*)v, n4, s, (const AdamScalars*)scalars_dev,
                                                                (float4*)grad_out, norm);
        DPPO_LAUNCH_CHECK();
    } else if (grouped) {
        reduce_adam_grouped_kernel<<<gblocks, RG * 32, 0, stream>>>((const float4*)partials, nparts, (float4*)gshared,
                                                                    (float4*)theta, (float4*)theta_prev, (float4*)m,
                                                                    (float4*)v, n4, s, (const AdamScalars*)scalars_dev,
                                                                    (float4*)grad_out, loss_partials, loss_out);
        DPPO_LAUNCH_CHECK();
    } else {
        reduce_adam_kernel<<<nblocks, ADAM_THREADS, 0, stream>>>((const float4*)partials, nparts, (float4*)gshared,
                                                                  (float4*)theta, (float4*)theta_prev, (float4*)m,
                                                                  (float4*)v, n4, s, (const AdamScalars*)scalars_dev,
                                                                  (float4*)grad_out, loss_partials, loss_out);
        DPPO_LAUNCH_CHECK();
    }
    return 0;
}
