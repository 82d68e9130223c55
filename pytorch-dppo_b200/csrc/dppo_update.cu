// K2 - fused minibatch gather + actor-critic forward/backward + PPO losses.
//
// Reference: ReplayMemory.sample (train.py:31-33 / ppo.py:58-60), Model.forward
// (model.py:35-45), the DPPO loss train.py:140-161 (+ normal() 41-44) or the
// ppo.py loss ppo.py:147-170, and backward (train.py:166 / ppo.py:172).
//
// CUDA-core path (all widths).  One CTA owns a tile of TM <= 32 gathered
// samples and carries them through old-net forward, new-net forward, the loss
// and the whole backward with every activation resident in shared memory;
// weights stream L2 -> shared memory in chunks (they are shared by all CTAs and
// stay L2-resident).  Each CTA writes ONE partial of every gradient tensor
// (its samples' contribution); K3 sums the partials in CTA order, so results
// are bit-reproducible (no float atomics).
//
// Register tiling: forward  C[m][n] += in[m][k..k+3] . W[n][k..k+3]   4x4 per thread
//                  dX       C[m][k] += dOut[m][n..n+3] . W[n..n+3][k] 4x(4*KT)
//                  dW       C[n][k] += dOut[m][n] * in[m][k]          4x4
// all operands read as 128-bit shared-memory vectors along their contiguous
// dimension; rows are interleaved over threads and leading dimensions are odd
// multiples of 16 B so the quarter-warp phases are bank-conflict free.
#include <math.h>
#include "common.cuh"

namespace {

constexpr int UPD_THREADS = 256;
constexpr int UPD_TM_MAX = 32;
constexpr int UPD_STAGE = 5120;             // floats per weight staging buffer (20 KB)
constexpr int UPD_MAX_A = 64;
constexpr int UPD_MAX_WIDTH = 1024;

// smallest multiple of 4 >= k whose quarter (in 16-byte units) is odd
__host__ __device__ inline int ld_of(int k) {
    int l = (k + 3) & ~3;
    if (((l >> 2) & 1) == 0) l += 4;
    return l;
}

struct Net {                                 // pointers into one flat parameter bucket
    const float *ls, *W1, *b1, *W2, *b2, *V1, *c1, *V2, *c2, *Wmu, *bmu, *Wv, *bv;
};
struct GradOut {                             // pointers into this CTA's partial bucket
    float *ls, *W1, *b1, *W2, *b2, *V1, *c1, *V2, *c2, *Wmu, *bmu, *Wv, *bv;
};

struct Layout { int off[DPPO_N_PARAMS]; int size[DPPO_N_PARAMS]; int P_pad; };

struct UpdArgs {
    int flavour;
    dppo_dims d;
    Layout lay;
    const float *obs, *act, *adv, *ret, *logp_old;
    const int32_t* idx;
    int B, TM, ntiles;
    const float *theta, *theta_old;
    float clip, ent_coeff;
    float* partials;
    float* loss_partials;
    // shared-memory carve-up (float offsets)
    int ldO, ldH1, ldH2, ldG1, ldG2, ldA;
    int o_x, o_h1, o_h2, o_g1, o_g2, o_act, o_mu, o_muo, o_dmu, o_v, o_vo, o_dv, o_adv, o_ret, o_lpo, o_red, o_stage;
    int smem_floats;
};

__device__ __forceinline__ Net net_of(const float* t, const Layout& L) {
    Net n;
    n.ls = t + L.off[0]; n.W1 = t + L.off[1]; n.b1 = t + L.off[2]; n.W2 = t + L.off[3]; n.b2 = t + L.off[4];
    n.V1 = t + L.off[5]; n.c1 = t + L.off[6]; n.V2 = t + L.off[7]; n.c2 = t + L.off[8];
    n.Wmu = t + L.off[9]; n.bmu = t + L.off[10]; n.Wv = t + L.off[11]; n.bv = t + L.off[12];
    return n;
}
__device__ __forceinline__ GradOut grad_of(float* t, const Layout& L) {
    GradOut n;
    n.ls = t + L.off[0]; n.W1 = t + L.off[1]; n.b1 = t + L.off[2]; n.W2 = t + L.off[3]; n.b2 = t + L.off[4];
    n.V1 = t + L.off[5]; n.c1 = t + L.off[6]; n.V2 = t + L.off[7]; n.c2 = t + L.off[8];
    n.Wmu = t + L.off[9]; n.bmu = t + L.off[10]; n.Wv = t + L.off[11]; n.bv = t + L.off[12];
    return n;
}

// ---------------------------------------------------------------------------
// forward layer: out[m][n] = act(sum_k in[m][k] W[n][k] + b[n]),  m < TM, n < N
// in: smem [TM][ldin] (zero beyond K); W: global row-major [N][K]; out: smem [TM][ldout]
// ---------------------------------------------------------------------------
template <bool TANH>
__device__ void fwd_layer(const float* __restrict__ W, const float* __restrict__ bias, int N, int K,
                          const float* in, int ldin, float* out, int ldout, int TM, float* stage) {
    const int MG = TM >> 2;                                  // row groups (TM is a multiple of 4)
    const int Kp = (K + 3) & ~3;
    const bool vecW = ((K & 3) == 0) && aligned16_dev(W);
    // column chunk: as many 4-column groups as threads allow, bounded by the stage
    int NGmax = UPD_THREADS / MG;
    for (int n0 = 0; n0 < N; n0 += 4 * NGmax) {
        const int NCH = min(N - n0, 4 * NGmax);
        const int NG = (NCH + 3) >> 2;
        const int NCHp = NG * 4;
        int KC = Kp;                                         // k-chunk so that NCHp*ld(KC) fits the stage
        while (NCHp * ld_of(KC) > UPD_STAGE) KC = ((KC >> 1) + 3) & ~3;
        const int ni = threadIdx.x % NG, mi = threadIdx.x / NG;
        const bool active = mi < MG;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        for (int k0 = 0; k0 < Kp; k0 += KC) {
            const int kc = min(KC, Kp - k0);
            const int lds = ld_of(kc);
            __syncthreads();                                 // previous chunk fully consumed
            // stage W[n0 .. n0+NCHp)[k0 .. k0+kc) -> stage[nn][lds], zero-filled outside (N,K)
            const int q = kc >> 2;
            for (int e = threadIdx.x; e < NCHp * q; e += UPD_THREADS) {
                const int nn = e / q, kq = e - nn * q;
                const int n = n0 + nn, k = k0 + 4 * kq;
                float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (n < N) {
                    const float* src = W + (size_t)n * K + k;
                    if (vecW) w = __ldg(reinterpret_cast<const float4*>(src));
                    else {
                        if (k + 0 < K) w.x = __ldg(src + 0);
                        if (k + 1 < K) w.y = __ldg(src + 1);
                        if (k + 2 < K) w.z = __ldg(src + 2);
                        if (k + 3 < K) w.w = __ldg(src + 3);
                    }
                }
                *reinterpret_cast<float4*>(stage + nn * lds + 4 * kq) = w;
            }
            __syncthreads();
            if (active) {
                const float* a0 = in + mi * ldin + k0;
                const float* b0 = stage + ni * lds;
                for (int kk = 0; kk < kc; kk += 4) {
                    float4 a[4], b[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(a0 + (i * MG) * ldin + kk);
#pragma unroll
                    for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(b0 + (j * NG) * lds + kk);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
                            acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
                            acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
                            acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
                        }
                }
            }
        }
        if (active) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int n = n0 + ni + j * NG;
                if (n < N) {
                    const float bb = __ldg(bias + n);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float y = acc[i][j] + bb;
                        if (TANH) y = tanhf(y);
                        out[(mi + i * MG) * ldout + n] = y;
                    }
                }
            }
        }
    }
    __syncthreads();
}

// tiny head (N <= A or 1): one warp per output, lanes over k
__device__ void head_fwd(const float* __restrict__ W, const float* __restrict__ bias, int N, int K, const float* in,
                         int ldin, float* out, int ldout, int TM) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int o = warp; o < TM * N; o += UPD_THREADS / 32) {
        const int m = o / N, n = o - m * N;
        float s = 0.f;
        for (int k = lane; k < K; k += 32) s = fmaf(in[m * ldin + k], __ldg(W + (size_t)n * K + k), s);
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
        if (lane == 0) out[m * ldout + n] = s + __ldg(bias + n);
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// dX through a tanh layer, in place:
//   act[m][k] <- (sum_n dOut[m][n] W[n][k]) * (1 - act[m][k]^2),  k < K
// dOut: smem [TM][ldd] (zero beyond N); W global [N][K]
// ---------------------------------------------------------------------------
constexpr int DX_KT = 4;                                     // k-quads per thread
__device__ void dx_layer(const float* __restrict__ W, int N, int K, const float* dOut, int ldd, float* act, int lda,
                         int TM, float* stage) {
    const int MG = TM >> 2;
    const int KQ = (K + 3) >> 2;                             // k-quads
    const int Kp = KQ * 4;
    const bool vecW = ((K & 3) == 0) && aligned16_dev(W);
    const int KGL = UPD_THREADS / MG;                        // quad lanes per row group
    const int Np = (N + 3) & ~3;
    int NC = (UPD_STAGE / Kp) & ~3;                          // rows of W per chunk
    if (NC > Np) NC = Np;
    const int ki = threadIdx.x % KGL, mi = threadIdx.x / KGL;
    const bool active = mi < MG;
    for (int kq0 = 0; kq0 < KQ; kq0 += KGL * DX_KT) {        // usually a single pass
        float acc[4][DX_KT][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < DX_KT; ++j)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[i][j][c] = 0.f;
        for (int n0 = 0; n0 < Np; n0 += NC) {
            const int nc = min(NC, Np - n0);
            __syncthreads();
            for (int e = threadIdx.x; e < nc * KQ; e += UPD_THREADS) {
                const int nn = e / KQ, kq = e - nn * KQ;
                const int n = n0 + nn, k = 4 * kq;
                float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
                if (n < N) {
                    const float* src = W + (size_t)n * K + k;
                    if (vecW) w = __ldg(reinterpret_cast<const float4*>(src));
                    else {
                        if (k + 0 < K) w.x = __ldg(src + 0);
                        if (k + 1 < K) w.y = __ldg(src + 1);
                        if (k + 2 < K) w.z = __ldg(src + 2);
                        if (k + 3 < K) w.w = __ldg(src + 3);
                    }
                }
                *reinterpret_cast<float4*>(stage + nn * Kp + 4 * kq) = w;
            }
            __syncthreads();
            if (active) {
                for (int nn = 0; nn < nc; nn += 4) {
                    float4 a[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        a[i] = *reinterpret_cast<const float4*>(dOut + (mi + i * MG) * ldd + n0 + nn);
#pragma unroll
                    for (int j = 0; j < DX_KT; ++j) {
                        const int kq = kq0 + ki + j * KGL;
                        if (kq < KQ) {
                            const float* wp = stage + nn * Kp + 4 * kq;
                            const float4 w0 = *reinterpret_cast<const float4*>(wp);
                            const float4 w1 = *reinterpret_cast<const float4*>(wp + Kp);
                            const float4 w2 = *reinterpret_cast<const float4*>(wp + 2 * Kp);
                            const float4 w3 = *reinterpret_cast<const float4*>(wp + 3 * Kp);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                acc[i][j][0] = fmaf(a[i].x, w0.x, acc[i][j][0]);
                                acc[i][j][1] = fmaf(a[i].x, w0.y, acc[i][j][1]);
                                acc[i][j][2] = fmaf(a[i].x, w0.z, acc[i][j][2]);
                                acc[i][j][3] = fmaf(a[i].x, w0.w, acc[i][j][3]);
                                acc[i][j][0] = fmaf(a[i].y, w1.x, acc[i][j][0]);
                                acc[i][j][1] = fmaf(a[i].y, w1.y, acc[i][j][1]);
                                acc[i][j][2] = fmaf(a[i].y, w1.z, acc[i][j][2]);
                                acc[i][j][3] = fmaf(a[i].y, w1.w, acc[i][j][3]);
                                acc[i][j][0] = fmaf(a[i].z, w2.x, acc[i][j][0]);
                                acc[i][j][1] = fmaf(a[i].z, w2.y, acc[i][j][1]);
                                acc[i][j][2] = fmaf(a[i].z, w2.z, acc[i][j][2]);
                                acc[i][j][3] = fmaf(a[i].z, w2.w, acc[i][j][3]);
                                acc[i][j][0] = fmaf(a[i].w, w3.x, acc[i][j][0]);
                                acc[i][j][1] = fmaf(a[i].w, w3.y, acc[i][j][1]);
                                acc[i][j][2] = fmaf(a[i].w, w3.z, acc[i][j][2]);
                                acc[i][j][3] = fmaf(a[i].w, w3.w, acc[i][j][3]);
                            }
                        }
                    }
                }
            }
        }
        if (active) {
#pragma unroll
            for (int j = 0; j < DX_KT; ++j) {
                const int kq = kq0 + ki + j * KGL;
                if (kq < KQ) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float* ap = act + (mi + i * MG) * lda + 4 * kq;
                        float4 h = *reinterpret_cast<float4*>(ap);
                        h.x = acc[i][j][0] * (1.f - h.x * h.x);
                        h.y = acc[i][j][1] * (1.f - h.y * h.y);
                        h.z = acc[i][j][2] * (1.f - h.z * h.z);
                        h.w = acc[i][j][3] * (1.f - h.w * h.w);
                        // columns k >= K hold act = 0 and W = 0, so they stay 0
                        *reinterpret_cast<float4*>(ap) = h;
                    }
                }
            }
        }
    }
    __syncthreads();
}

// dX through a tiny head (N small): act[m][k] <- (sum_n dOut[m][n] W[n][k]) (1 - act^2)
__device__ void head_dx(const float* __restrict__ W, int N, int K, const float* dOut, int ldd, float* act, int lda,
                        int TM) {
    for (int e = threadIdx.x; e < TM * K; e += UPD_THREADS) {
        const int m = e / K, k = e - m * K;
        float s = 0.f;
        for (int n = 0; n < N; ++n) s = fmaf(dOut[m * ldd + n], __ldg(W + (size_t)n * K + k), s);
        const float h = act[m * lda + k];
        act[m * lda + k] = s * (1.f - h * h);
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// dW[n][k] (+)= sum_m dOut[m][n] in[m][k];  db[n] (+)= sum_m dOut[m][n]
// written to this CTA's partial bucket (global).  m ascending: fixed order.
// ---------------------------------------------------------------------------
__device__ void dw_layer(float* __restrict__ gW, float* __restrict__ gb, int N, int K, const float* dOut, int ldd,
                         const float* in, int ldin, int TM, bool accumulate) {
    const int NQ = (N + 3) >> 2, KQ = (K + 3) >> 2;
    const bool vecG = ((K & 3) == 0) && aligned16_dev(gW);
    for (int t = threadIdx.x; t < NQ * KQ; t += UPD_THREADS) {
        const int nq = t / KQ, kq = t - nq * KQ;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
        const float* dp = dOut + 4 * nq;
        const float* ip = in + 4 * kq;
#pragma unroll 4
        for (int m = 0; m < TM; ++m) {
            const float4 a = *reinterpret_cast<const float4*>(dp + m * ldd);
            const float4 b = *reinterpret_cast<const float4*>(ip + m * ldin);
            acc[0][0] = fmaf(a.x, b.x, acc[0][0]); acc[0][1] = fmaf(a.x, b.y, acc[0][1]);
            acc[0][2] = fmaf(a.x, b.z, acc[0][2]); acc[0][3] = fmaf(a.x, b.w, acc[0][3]);
            acc[1][0] = fmaf(a.y, b.x, acc[1][0]); acc[1][1] = fmaf(a.y, b.y, acc[1][1]);
            acc[1][2] = fmaf(a.y, b.z, acc[1][2]); acc[1][3] = fmaf(a.y, b.w, acc[1][3]);
            acc[2][0] = fmaf(a.z, b.x, acc[2][0]); acc[2][1] = fmaf(a.z, b.y, acc[2][1]);
            acc[2][2] = fmaf(a.z, b.z, acc[2][2]); acc[2][3] = fmaf(a.z, b.w, acc[2][3]);
            acc[3][0] = fmaf(a.w, b.x, acc[3][0]); acc[3][1] = fmaf(a.w, b.y, acc[3][1]);
            acc[3][2] = fmaf(a.w, b.z, acc[3][2]); acc[3][3] = fmaf(a.w, b.w, acc[3][3]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int n = 4 * nq + i;
            if (n < N) {
                float* dst = gW + (size_t)n * K + 4 * kq;
                if (vecG) {
                    float4 o = make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
                    if (accumulate) {
                        const float4 p = *reinterpret_cast<const float4*>(dst);
                        o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
                    }
                    *reinterpret_cast<float4*>(dst) = o;
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (4 * kq + j < K) dst[j] = accumulate ? dst[j] + acc[i][j] : acc[i][j];
                }
            }
        }
    }
    for (int n = threadIdx.x; n < N; n += UPD_THREADS) {
        float s = 0.f;
        for (int m = 0; m < TM; ++m) s += dOut[m * ldd + n];
        gb[n] = accumulate ? gb[n] + s : s;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// whole forward of one network on the tile (model.py:35-45)
// ---------------------------------------------------------------------------
__device__ void net_forward(const Net& nw, const UpdArgs& P, float* S, float* mu_out, float* v_out) {
    const dppo_dims& d = P.d;
    float* stage = S + P.o_stage;
    fwd_layer<true>(nw.W1, nw.b1, d.H1, d.O, S + P.o_x, P.ldO, S + P.o_h1, P.ldH1, P.TM, stage);
    fwd_layer<true>(nw.W2, nw.b2, d.H2, d.H1, S + P.o_h1, P.ldH1, S + P.o_h2, P.ldH2, P.TM, stage);
    head_fwd(nw.Wmu, nw.bmu, d.A, d.H2, S + P.o_h2, P.ldH2, mu_out, P.ldA, P.TM);
    fwd_layer<true>(nw.V1, nw.c1, d.G1, d.O, S + P.o_x, P.ldO, S + P.o_g1, P.ldG1, P.TM, stage);
    fwd_layer<true>(nw.V2, nw.c2, d.G2, d.G1, S + P.o_g1, P.ldG1, S + P.o_g2, P.ldG2, P.TM, stage);
    head_fwd(nw.Wv, nw.bv, 1, d.G2, S + P.o_g2, P.ldG2, v_out, 1, P.TM);
}

// deterministic block sum (fixed tree), result broadcast
__device__ float block_sum(float x, float* red) {
    __syncthreads();
    red[threadIdx.x] = x;
    __syncthreads();
    for (int s = UPD_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    const float r = red[0];
    __syncthreads();
    return r;
}

__global__ void __launch_bounds__(UPD_THREADS, 1) update_fused_kernel(const UpdArgs P) {
    extern __shared__ __align__(16) float S[];
    const dppo_dims& d = P.d;
    const int TM = P.TM;
    const Net nw = net_of(P.theta, P.lay);
    float* part = P.partials + (size_t)blockIdx.x * P.lay.P_pad;
    const GradOut go = grad_of(part, P.lay);
    float* red = S + P.o_red;
    if (threadIdx.x < DPPO_N_PARAMS)                                   // padding floats of the bucket stay 0
        for (int j = P.lay.size[threadIdx.x]; j < ((P.lay.size[threadIdx.x] + 3) & ~3); ++j)
            part[P.lay.off[threadIdx.x] + j] = 0.f;
    float lsum_clip = 0.f, lsum_val = 0.f, lsum_ent = 0.f;           // this thread's share of the loss sums
    bool first = true;

    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, first = false) {
        const int b0 = tile * TM;
        const int valid = min(TM, P.B - b0);
        // ---- zero the activation area (pads must read as 0), then gather --------------------
        for (int i = threadIdx.x; i < P.o_stage; i += UPD_THREADS) S[i] = 0.f;
        __syncthreads();
        for (int e = threadIdx.x; e < valid * d.O; e += UPD_THREADS) {
            const int m = e / d.O, k = e - m * d.O;
            S[P.o_x + m * P.ldO + k] = __ldg(P.obs + (size_t)__ldg(P.idx + b0 + m) * d.O + k);
        }
        for (int e = threadIdx.x; e < valid * d.A; e += UPD_THREADS) {
            const int m = e / d.A, a = e - m * d.A;
            S[P.o_act + m * d.A + a] = __ldg(P.act + (size_t)__ldg(P.idx + b0 + m) * d.A + a);
        }
        for (int m = threadIdx.x; m < valid; m += UPD_THREADS) {
            const size_t r = (size_t)__ldg(P.idx + b0 + m);
            S[P.o_adv + m] = __ldg(P.adv + r);
            S[P.o_ret + m] = __ldg(P.ret + r);
            if (P.flavour == DPPO_FLAVOUR_PPO) S[P.o_lpo + m] = __ldg(P.logp_old + r);
        }
        __syncthreads();

        // ---- forward(s) -----------------------------------------------------------------------
        if (P.flavour == DPPO_FLAVOUR_DPPO) {
            const Net no = net_of(P.theta_old, P.lay);
            net_forward(no, P, S, S + P.o_muo, S + P.o_vo);              // train.py:142
        }
        net_forward(nw, P, S, S + P.o_mu, S + P.o_v);                    // train.py:145 / ppo.py:148

        // ---- losses and their derivatives w.r.t. mu, v, log_std ---------------------------------
        const float lo = 1.f - P.clip, hi = 1.f + P.clip;               // clamp bounds (train.py:151)
        if (P.flavour == DPPO_FLAVOUR_DPPO) {
            const float g0 = -1.f / ((float)P.B * (float)d.A);          // d loss_clip / d min(.,.)
            const float ge = -P.ent_coeff / ((float)P.B * (float)d.A);
            const Net no = net_of(P.theta_old, P.lay);
            for (int e = threadIdx.x; e < TM * d.A; e += UPD_THREADS) {
                const int m = e / d.A, a = e - m * d.A;
                float dmu = 0.f, dls = 0.f;
                if (m < valid) {
                    const float x = S[P.o_act + m * d.A + a];
                    const float A_ = S[P.o_adv + m];
                    // normal(x, mu, sigma_sq), train.py:41-44, new and old parameters
                    const float s = expf(__ldg(nw.ls + a)), so = expf(__ldg(no.ls + a));
                    const float u = x - S[P.o_mu + m * P.ldA + a];
                    const float w = -(u * u);
                    const float two_s = 2.f * s;
                    const float qq = w / two_s;
                    const float ea = expf(qq);
                    const float z = two_s * 3.14159274101257324f;       // fp32(np.pi)
                    const float rt = sqrtf(z);
                    const float bb = 1.f / rt;
                    const float p = ea * bb;
                    const float uo = x - S[P.o_muo + m * P.ldA + a];
                    const float po = expf(-(uo * uo) / (2.f * so)) * (1.f / sqrtf(2.f * so * 3.14159274101257324f));
                    const float den = 1e-10f + po;                       // train.py:148
                    const float ratio = p / den;
                    const float surr1 = ratio * A_;
                    const float rc = fminf(fmaxf(ratio, lo), hi);
                    const float surr2 = rc * A_;
                    lsum_clip += fminf(surr1, surr2);
                    lsum_ent += p * logf(p + 1e-5f);                     // train.py:159
                    // torch.min backward: ties split the gradient in halves
                    float d1, d2;
                    if (surr1 < surr2) { d1 = g0; d2 = 0.f; }
                    else if (surr1 > surr2) { d1 = 0.f; d2 = g0; }
                    else { d1 = 0.5f * g0; d2 = 0.5f * g0; }
                    float dratio = d1 * A_;
                    if (ratio >= lo && ratio <= hi) dratio += d2 * A_;   // clamp backward (inclusive)
                    float dp = dratio / den;
                    dp += ge * (logf(p + 1e-5f) + p / (p + 1e-5f));
                    // back through normal()
                    const float dea = dp * bb, dbb = dp * ea;
                    const float dq = dea * ea;
                    const float dw = dq / two_s;
                    float dtwo_s = -dq * w / (two_s * two_s);
                    dmu = 2.f * u * dw;                                  // u = x - mu, w = -u^2
                    const float drt = -dbb / (rt * rt);
                    const float dz = drt / (2.f * rt);
                    dtwo_s += dz * 3.14159274101257324f;
                    dls = 2.f * dtwo_s * s;                              // sigma_sq = exp(log_std)
                }
                S[P.o_dmu + m * P.ldA + a] = dmu;
                S[P.o_muo + m * P.ldA + a] = dls;                        // mu_old no longer needed: reuse for dls
            }
            const float gv = 0.5f / (float)P.B;                          // train.py:157
            for (int m = threadIdx.x; m < TM; m += UPD_THREADS) {
                float dv = 0.f;
                if (m < valid) {
                    const float v = S[P.o_v + m], vo = S[P.o_vo + m], R = S[P.o_ret + m];
                    const float e1 = v - R;
                    const float f1 = e1 * e1;                            // train.py:154
                    const float dl = v - vo;
                    const float vc = vo + fminf(fmaxf(dl, -P.clip), P.clip);   // train.py:155
                    const float e2 = vc - R;
                    const float f2 = e2 * e2;                            // train.py:156
                    lsum_val += fmaxf(f1, f2);
                    const bool inr = (dl >= -P.clip) && (dl <= P.clip);
                    float c1, c2;
                    if (f1 > f2) { c1 = gv; c2 = 0.f; }
                    else if (f1 < f2) { c1 = 0.f; c2 = gv; }
                    else { c1 = 0.5f * gv; c2 = 0.5f * gv; }
                    dv = c1 * 2.f * e1 + (inr ? c2 * 2.f * e2 : 0.f);
                }
                S[P.o_dv + 4 * m] = dv;
            }
        } else {
            const float g0 = -1.f / (float)P.B;                          // ppo.py:161
            const float gv = 1.f / (float)P.B;                           // ppo.py:164
            const float half_log_2pi = 0.918938533204672742f;            // 0.5*math.log(2*math.pi)
            for (int m = threadIdx.x; m < TM; m += UPD_THREADS) {
                float dv = 0.f, dlogp = 0.f;
                if (m < valid) {
                    float logp = 0.f;
                    for (int a = 0; a < d.A; ++a) {                      // ppo.py:150-151
                        const float ls = __ldg(nw.ls + a);
                        const float zz = (S[P.o_act + m * d.A + a] - S[P.o_mu + m * P.ldA + a]) / expf(ls);
                        const float t = (-0.5f * (zz * zz) - half_log_2pi) - ls;
                        logp = (a == 0) ? t : logp + t;
                    }
                    const float ratio = expf(logp - S[P.o_lpo + m]);    // ppo.py:156
                    const float A_ = S[P.o_adv + m];
                    const float surr1 = ratio * A_;
                    const float surr2 = fminf(fmaxf(ratio, lo), hi) * A_;
                    lsum_clip += fminf(surr1, surr2);
                    float d1, d2;
                    if (surr1 < surr2) { d1 = g0; d2 = 0.f; }
                    else if (surr1 > surr2) { d1 = 0.f; d2 = g0; }
                    else { d1 = 0.5f * g0; d2 = 0.5f * g0; }
                    float dratio = d1 * A_;
                    if (ratio >= lo && ratio <= hi) dratio += d2 * A_;
                    dlogp = dratio * ratio;
                    const float e1 = S[P.o_v + m] - S[P.o_ret + m];
                    lsum_val += e1 * e1;
                    dv = gv * 2.f * e1;
                }
                S[P.o_dv + 4 * m] = dv;
                for (int a = 0; a < d.A; ++a) {
                    float dmu = 0.f, dls = 0.f;
                    if (m < valid) {
                        const float sg = expf(__ldg(nw.ls + a));
                        const float zz = (S[P.o_act + m * d.A + a] - S[P.o_mu + m * P.ldA + a]) / sg;
                        dmu = dlogp * zz / sg;
                        dls = dlogp * (zz * zz - 1.f);
                    }
                    S[P.o_dmu + m * P.ldA + a] = dmu;
                    S[P.o_muo + m * P.ldA + a] = dls;
                }
            }
        }
        __syncthreads();
        // d log_std[a] = sum over the tile's samples (ascending m)
        for (int a = threadIdx.x; a < d.A; a += UPD_THREADS) {
            float s = 0.f;
            for (int m = 0; m < TM; ++m) s += S[P.o_muo + m * P.ldA + a];
            if (P.flavour == DPPO_FLAVOUR_PPO && blockIdx.x == 0 && first) s -= P.ent_coeff;   // ppo.py:152-153,167
            go.ls[a] = first ? s : go.ls[a] + s;
        }
        __syncthreads();

        // ---- backward, policy trunk -------------------------------------------------------------
        const bool acc = !first;
        float* stage = S + P.o_stage;
        dw_layer(go.Wmu, go.bmu, d.A, d.H2, S + P.o_dmu, P.ldA, S + P.o_h2, P.ldH2, TM, acc);
        head_dx(nw.Wmu, d.A, d.H2, S + P.o_dmu, P.ldA, S + P.o_h2, P.ldH2, TM);          // h2 <- dz2
        dw_layer(go.W2, go.b2, d.H2, d.H1, S + P.o_h2, P.ldH2, S + P.o_h1, P.ldH1, TM, acc);
        dx_layer(nw.W2, d.H2, d.H1, S + P.o_h2, P.ldH2, S + P.o_h1, P.ldH1, TM, stage);  // h1 <- dz1
        dw_layer(go.W1, go.b1, d.H1, d.O, S + P.o_h1, P.ldH1, S + P.o_x, P.ldO, TM, acc);
        // ---- backward, value trunk --------------------------------------------------------------
        dw_layer(go.Wv, go.bv, 1, d.G2, S + P.o_dv, 4, S + P.o_g2, P.ldG2, TM, acc);
        head_dx(nw.Wv, 1, d.G2, S + P.o_dv, 4, S + P.o_g2, P.ldG2, TM);
        dw_layer(go.V2, go.c2, d.G2, d.G1, S + P.o_g2, P.ldG2, S + P.o_g1, P.ldG1, TM, acc);
        dx_layer(nw.V2, d.G2, d.G1, S + P.o_g2, P.ldG2, S + P.o_g1, P.ldG1, TM, stage);
        dw_layer(go.V1, go.c1, d.G1, d.O, S + P.o_g1, P.ldG1, S + P.o_x, P.ldO, TM, acc);
    }

    // ---- per-CTA loss sums (scaled so that K3 only has to add them) -----------------------------
    const float sc = block_sum(lsum_clip, red);
    const float sv = block_sum(lsum_val, red);
    const float se = block_sum(lsum_ent, red);
    if (threadIdx.x == 0) {
        float* lp = P.loss_partials + 4 * blockIdx.x;
        if (P.flavour == DPPO_FLAVOUR_DPPO) {
            const float BA = (float)P.B * (float)d.A;
            lp[0] = -sc / BA;
            lp[1] = 0.5f * sv / (float)P.B;
            lp[2] = -P.ent_coeff * se / BA;
        } else {
            lp[0] = -sc / (float)P.B;
            lp[1] = sv / (float)P.B;
            float ent = 0.f;                                             // ppo.py:152-153
            if (blockIdx.x == 0)
                for (int a = 0; a < d.A; ++a) ent += (0.5f + 0.918938533204672742f) + __ldg(nw.ls + a);
            lp[2] = -P.ent_coeff * ent;
        }
        lp[3] = 0.f;
    }
}

int fill_layout(const dppo_dims& d, Layout& L) {
    const int sizes[DPPO_N_PARAMS] = {d.A, d.H1 * d.O, d.H1, d.H2 * d.H1, d.H2, d.G1 * d.O, d.G1,
                                      d.G2 * d.G1, d.G2, d.A * d.H2, d.A, d.G2, 1};
    int o = 0;
    for (int i = 0; i < DPPO_N_PARAMS; ++i) {
        L.off[i] = o;
        L.size[i] = sizes[i];
        o += (sizes[i] + 3) & ~3;
    }
    L.P_pad = o;
    return o;
}

int check_dims(const dppo_dims* d) {
    if (!d) return DPPO_E_BADARG;
    if (d->O < 1 || d->A < 1 || d->H1 < 1 || d->H2 < 1 || d->G1 < 1 || d->G2 < 1) return DPPO_E_BADARG;
    if (d->A > UPD_MAX_A || d->O > UPD_MAX_WIDTH || d->H1 > UPD_MAX_WIDTH || d->H2 > UPD_MAX_WIDTH ||
        d->G1 > UPD_MAX_WIDTH || d->G2 > UPD_MAX_WIDTH)
        return DPPO_E_TOOLARGE;
    return 0;
}

// samples per CTA and number of CTAs (= partials) for a minibatch of B
void pick_tiling(int B, int* TM, int* nparts) {
    const int sms = dppo_sm_count();
    int tm = ((B + sms - 1) / sms + 3) & ~3;
    if (tm < 4) tm = 4;
    if (tm > UPD_TM_MAX) tm = UPD_TM_MAX;
    const int ntiles = (B + tm - 1) / tm;
    *TM = tm;
    *nparts = ntiles < sms ? ntiles : sms;
}

int plan(UpdArgs& P) {
    const dppo_dims& d = P.d;
    P.ldO = ld_of(d.O); P.ldH1 = ld_of(d.H1); P.ldH2 = ld_of(d.H2); P.ldG1 = ld_of(d.G1); P.ldG2 = ld_of(d.G2);
    P.ldA = ld_of(d.A);
    const int TM = P.TM;
    int o = 0;
    auto take = [&](int n) { int r = o; o += (n + 3) & ~3; return r; };
    P.o_x = take(TM * P.ldO);
    P.o_h1 = take(TM * P.ldH1);
    P.o_h2 = take(TM * P.ldH2);
    P.o_g1 = take(TM * P.ldG1);
    P.o_g2 = take(TM * P.ldG2);
    P.o_act = take(TM * d.A);
    P.o_mu = take(TM * P.ldA);
    P.o_muo = take(TM * P.ldA);
    P.o_dmu = take(TM * P.ldA);
    P.o_v = take(TM);
    P.o_vo = take(TM);
    P.o_dv = take(TM * 4);                      // [TM][4]: column 0 = dv, so rows are 16-byte vectors
    P.o_adv = take(TM);
    P.o_ret = take(TM);
    P.o_lpo = take(TM);
    P.o_stage = o;                              // everything before the stage is zeroed per tile
    o += UPD_STAGE;
    P.o_red = take(UPD_THREADS);
    P.smem_floats = o;
    return o;
}

}  // namespace

int dppo_sm_count() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess) return 148;
        if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
    }
    return sms;
}

extern "C" int dppo_param_layout(const dppo_dims* d, int* offsets_host, int* sizes_host) {
    const int rc = check_dims(d);
    if (rc) return rc;
    Layout L;
    fill_layout(*d, L);
    for (int i = 0; i < DPPO_N_PARAMS; ++i) {
        if (offsets_host) offsets_host[i] = L.off[i];
        if (sizes_host) sizes_host[i] = L.size[i];
    }
    return L.P_pad;
}

extern "C" int dppo_update_nparts(const dppo_dims* d, int B) {
    const int rc = check_dims(d);
    if (rc) return rc;
    if (B < 1) return DPPO_E_BADARG;
    int TM, np;
    pick_tiling(B, &TM, &np);
    return np;
}

extern "C" size_t dppo_update_partials_bytes(const dppo_dims* d, int B) {
    if (check_dims(d) || B < 1) return 0;
    Layout L;
    fill_layout(*d, L);
    int TM, np;
    pick_tiling(B, &TM, &np);
    return sizeof(float) * (size_t)np * L.P_pad;
}

extern "C" int dppo_update_fused(int flavour, const dppo_dims* d, const float* obs, const float* act, const float* adv,
                                 const float* ret, const float* logp_old, const int32_t* idx, int B, const float* theta,
                                 const float* theta_old, float clip, float ent_coeff, float* partials,
                                 float* loss_partials, void* stream_) {
    int rc = check_dims(d);
    if (rc) return rc;
    if (flavour != DPPO_FLAVOUR_DPPO && flavour != DPPO_FLAVOUR_PPO) return DPPO_E_BADARG;
    if (B < 1 || !obs || !act || !adv || !ret || !idx || !theta || !partials || !loss_partials) return DPPO_E_BADARG;
    if (flavour == DPPO_FLAVOUR_DPPO && !theta_old) return DPPO_E_BADARG;
    if (flavour == DPPO_FLAVOUR_PPO && !logp_old) return DPPO_E_BADARG;
    if (!aligned16(theta) || !aligned16(theta_old) || !aligned16(partials)) return DPPO_E_ALIGN;
    UpdArgs P;
    P.flavour = flavour;
    P.d = *d;
    fill_layout(*d, P.lay);
    P.obs = obs; P.act = act; P.adv = adv; P.ret = ret; P.logp_old = logp_old; P.idx = idx;
    P.B = B;
    int nparts;
    pick_tiling(B, &P.TM, &nparts);
    P.ntiles = (B + P.TM - 1) / P.TM;
    P.theta = theta; P.theta_old = theta_old;
    P.clip = clip; P.ent_coeff = ent_coeff;
    P.partials = partials; P.loss_partials = loss_partials;
    plan(P);
    const size_t smem = sizeof(float) * (size_t)P.smem_floats;
    if (smem > 227 * 1024) return DPPO_E_TOOLARGE;
    static size_t smem_set = 0;
    if (smem > smem_set) {
        DPPO_CUDA_TRY(cudaFuncSetAttribute(update_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set = smem;
    }
    update_fused_kernel<<<nparts, UPD_THREADS, smem, (cudaStream_t)stream_>>>(P);
    DPPO_LAUNCH_CHECK();
    return 0;
}
