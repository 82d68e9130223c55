// K2 - fused minibatch gather + actor-critic forward/backward + PPO losses.
//
// Reference: ReplayMemory.sample (train.py:31-33 / ppo.py:58-60), Model.forward
// (model.py:35-45), the DPPO loss train.py:140-161 (+ normal() 41-44) or the
// ppo.py loss ppo.py:147-170, and backward (train.py:166 / ppo.py:172).
//
// One CTA (256 threads, one per SM) owns a tile of TM <= 32 gathered samples and carries them through the old-net
// forward, the new-net forward, the loss and the whole backward with every activation resident in shared memory.
// Each CTA writes ONE partial of every gradient tensor (its samples' contribution); K3 sums the partials in CTA
// order, so results are bit-reproducible (no float atomics).
//
// Tensor-core path (tcgen05, 3xTF32 with fp32 accumulation in tensor memory) for the reference-sized layers:
//   wide layers (N <= 128 outputs, K >= 64, K % 4 == 0): forward and dX through ts_layer (TMA tiles -> registers ->
//     tensor memory -> MMA, role-specialised warps), dW through dw_layer_ts;
//   first layers (K = observation width <= 32, N <= 512): fwd_first_ts / dw_first_ts from weight copies that a bulk
//     copy fetched one phase earlier.
// FP32-pipe path for every other shape (wide nets such as 256x256, odd widths): register-tiled FFMA layers
//   forward  C[m][n] += in[m][k..k+3] . W[n][k..k+3]   4x4 per thread
//   dX       C[m][k] += dOut[m][n..n+3] . W[n..n+3][k] 4x(4*KT)
//   dW       C[n][k] += dOut[m][n] * in[m][k]          4x4
// with weights streamed L2 -> shared memory by TMA bulk copies on mbarriers, all operands read as 128-bit
// shared-memory vectors; leading dimensions are odd multiples of 16 B so quarter-warp phases are conflict free.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <cuda.h>
#include "common.cuh"

// per-phase clock64() stamps of CTA 0 (tools/k2clock.py); compiled in with -DDPPO_K2_TIMING only
#ifdef DPPO_K2_TIMING
__device__ long long g_k2_clock[64];
#define K2_TICK(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) g_k2_clock[i] = clock64(); } while (0)
#define K2_ACC_DECL long long acc_t0 = 0; long long acc_r[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define K2_ACC_T0 acc_t0 = clock64()
#define K2_ACC(i) do { long long t = clock64(); acc_r[(i) - 40] += t - acc_t0; acc_t0 = t; } while (0)
#define K2_ACC_FLUSH do { if (blockIdx.x == 0 && (threadIdx.x == 0 || threadIdx.x == 224)) for (int q = 0; q < 8; ++q) g_k2_clock[(threadIdx.x ? 48 : 40) + q] += acc_r[q]; } while (0)
#else
#define K2_TICK(i) do { } while (0)
#define K2_ACC_DECL
#define K2_ACC_T0
#define K2_ACC(i) do { } while (0)
#define K2_ACC_FLUSH do { } while (0)
#endif
#ifndef DPPO_K2_EARLY_RELEASE
#define DPPO_K2_EARLY_RELEASE 0
#endif
#ifdef DPPO_K2_TRACE
__device__ long long g_k2_trace[4 * 64 * 8];
__device__ const float* g_k2_trace_bias;            // trace the forward layer with this bias vector (null: the wide ones)
#define K2_TR(role, gp, ev) do { if (trace_on && lane == 0) g_k2_trace[(((role) * 64 + (gp)) << 3) + (ev)] = clock64() - trace_t0; } while (0)
#else
#define K2_TR(role, gp, ev) do { } while (0)
#endif

namespace {

constexpr int UPD_THREADS = 256;                  // the tensor-core layers assign roles to exactly eight warps
constexpr int UPD_TM_MAX = 32;
constexpr int UPD_STAGE_MAX = 12800;        // floats per staging buffer, two buffers (together the 100 KB tensor-core staging region)
constexpr int UPD_STAGE_MIN = 4096;
constexpr int UPD_ROWS = 32;                // sample rows per CTA tile (TM <= 32 of them are real)
constexpr int UPD_MAX_A = 64;
constexpr int UPD_MAX_WIDTH = 1024;


// tanh for the hidden layers: branch-free, 1 - 2/(1 + 2^(2x log2 e)) with the SFU exp2/rcp approximations
// (2^-22 relative each), and the odd Taylor polynomial below |x| < 0.04 where the subtraction would cancel.
// Absolute error <= ~2.5e-7 over the whole range (tanhf: ~1.2e-7) at a third of the instructions; the parity
// tests (1e-5 per tensor) hold with margin.
__device__ __forceinline__ float tanh_act(float x) {
    const float ax = fabsf(x);
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(ax * 2.885390081777927f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
    const float big = fmaf(-2.0f, r, 1.0f);
    const float x2 = ax * ax;
    const float small = ax * fmaf(x2, fmaf(x2, 0.13333334f, -0.33333334f), 1.0f);
    return copysignf(ax < 0.04f ? small : big, x);
}

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
    int SF;                                     // floats per weight staging buffer
    int o_zero_end, o_heads, head_floats, o_bars, heads_in_smem;
    // tensor maps of the wide layers' weights (TMA tiles for the tensor-core path); use_ts: maps valid and staging region big enough
    int use_ts;
    int first_ts;                                   // ... and on the tensor cores (small_first, room for the sample tiles)
    int small_first;                                // first layers (K = O <= 32) run from prefetched flat weight copies
    int trunk_serial;                               // policy trunk start to end, then the value trunk, in the same activation buffers
    // forward-only mode (dppo_policy_forward: the rollout's batched Model.forward + action sampling, train.py:84-90)
    int fwd_only, std_is_variance;
    const float* eps;                               // [B][A] standard-normal draws (nullable: action = mu)
    float *f_mu, *f_v, *f_act, *f_logp, *f_obs;     // output columns (nullable), row r of the batch lands in row
    long long f_row0, f_stride;                     //   f_row0 + r * f_stride  (straight into the SoA rollout buffer)
    // batch-level dW (dwg): 0 off.  1: the update kernel forms no dW; it leaves every layer's input and every dz
    // feature-major in Z ([rows][Bp], Bp = 32 * tiles) and its loss sums in zloss.  2: the dW pass over Z - CTA =
    // (range of samples, parameter, 32 input features, 128 outputs), one wide-layer call whose "weights" are dz^T.
    int dwg, Bp, dwg_units, dwg_chunk, zloss_rows;   // dwg_chunk: samples per range (the grid is units x ranges)
    float* Z;
    float* zloss;                                   // [zloss_rows][4]
    int zr_x, zr_h1, zr_g1, zr_h2, zr_g2;           // first rows of the input operands (Kf rows, the ones row, pad to 32)
    int zr_d1p, zr_d1v, zr_d2p, zr_d2v, zr_dmu, zr_dv;   // first rows of dz: p_fc1, v_fc1, p_fc2, v_fc2, [dmu | dls], dv
    // tile operands (box 32 samples x 128 rows, SWIZZLE_128B): dz^T of v_fc2, p_fc2, v_fc1, p_fc1; for the heads h2^T, g2^T (+ ones)
    alignas(64) CUtensorMap map_dz[6];
    // 32-row operands (box 32 x 32, no swizzle): x^T, h1^T, g1^T (+ ones rows); for the heads [dmu | dls]^T and dv^T
    alignas(64) CUtensorMap map_in[5];
    alignas(64) CUtensorMap map_fwd[2][4];          // [theta, theta_old][W2, V2, W1, V1]: box 32 (k) x 128 (n), SWIZZLE_128B
    alignas(64) CUtensorMap map_dx[2];              // theta's W2, V2 for dX: box 128 (k) x 32 (n), no swizzle
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
// TMA bulk copies (cp.async.bulk, global -> shared, completion on an mbarrier): one
// instruction moves a whole weight row (or a whole block of rows), so staging costs no
// address arithmetic per element.  bar_phase[b] tracks the parity of stage buffer b.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, unsigned parity) {
    unsigned ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
#ifdef DPPO_K2_SPINTRAP
    long long spin = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spin > 4000000) { printf("mbar_wait stuck: block %d thread %d parity %u\n", blockIdx.x, threadIdx.x, parity); __trap(); }
    }
#else
    while (!mbar_try_wait(bar, parity)) { }
#endif
}
__device__ __forceinline__ void bulk_g2s(float* smem, const float* gmem, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (unsigned)__cvta_generic_to_shared(smem)),
                 "l"(gmem), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar))
                 : "memory");
}

// rows [n0, n0+nrows) x columns [k0, k0+kc) of W[N][K] -> dst[nn][lds]; K % 4 == 0, k0 % 4 == 0,
// W 16-byte aligned.  Rows >= N and columns >= K are NOT written (callers never let them reach a
// kept accumulator).  One bulk copy per row, issued by as many threads as there are rows.
__device__ __forceinline__ void stage_rows_bulk(float* dst, int lds, const float* __restrict__ W, int N, int K, int n0,
                                                int nrows, int k0, int kc, uint64_t* bar) {
    const int vr = min(nrows, N - n0);
    const int cols = min(kc, K - k0);
    if (vr <= 0 || cols <= 0) { if (threadIdx.x == 0) mbar_expect_tx(bar, 0); return; }
    if (threadIdx.x == 0) mbar_expect_tx(bar, (unsigned)(vr * cols * 4));
    for (int r = threadIdx.x; r < vr; r += UPD_THREADS)
        bulk_g2s(dst + r * lds, W + (size_t)(n0 + r) * K + k0, (unsigned)(cols * 4), bar);
}
// nfloats contiguous floats (multiple of 4, 16-byte aligned source) -> dst, one bulk copy
__device__ __forceinline__ void stage_flat_bulk(float* dst, const float* __restrict__ src, int nfloats, uint64_t* bar) {
    if (threadIdx.x == 0) {
        mbar_expect_tx(bar, (unsigned)(nfloats * 4));
        if (nfloats > 0) bulk_g2s(dst, src, (unsigned)(nfloats * 4), bar);
    }
}

// ---------------------------------------------------------------------------
// forward layer: out[m][n] = act(sum_k in[m][k] W[n][k] + b[n]),  m < 32, n < N
// in: smem [32][ldin] (finite beyond K); W: global row-major [N][K]; out: smem [32][ldout]
//
// Warp tile: 32 rows x 4*CN columns; lanes are 8 (rows) x 4 (columns), a thread owns rows
// mi+8i and columns c0+ni+4j, so per 4-deep k step a warp issues 4 + CN 128-bit shared loads
// (one wavefront each: 8 resp. 4 distinct 16-byte rows) for 16*CN FFMAs.  The 8 warps form
// NW column blocks x KW k-splits; k-split partners are summed through shared memory.
// ---------------------------------------------------------------------------
template <int CN, bool TANH>
__device__ __noinline__ unsigned fwd_layer_t(const float* __restrict__ W, const float* __restrict__ bias, int N, int K, const float* in,
                            int ldin, float* out, int ldout, float* stage, int SF, uint64_t* bars, unsigned ph) {
    constexpr int NWARPS = UPD_THREADS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int mi = lane >> 2, ni = lane & 3;
    const int Kp = (K + 3) & ~3;
    const bool vecW = (K & 3) == 0;                          // buckets are 16-byte aligned by construction
    int NW = (N + 4 * CN - 1) / (4 * CN);                     // column blocks needed
    if (NW > NWARPS) NW = NWARPS;
    int KW = 1;                                               // k-splits: largest power of two with NW*KW <= 8
    while (NW * KW * 2 <= NWARPS && (Kp >> 2) >= KW * 2) KW *= 2;
    const int cw = warp % NW, kw = warp / NW;
    const bool wactive = kw < KW;
    const int colsPerPass = NW * 4 * CN;
    for (int n0 = 0; n0 < N; n0 += colsPerPass) {
        const int nrows = min(colsPerPass, ((N - n0 + 4 * CN - 1) / (4 * CN)) * 4 * CN);
        // k-chunk: multiple of 4*KW, stage rows padded to an odd number of 16-byte units.
        // flat mode (K % 4 != 0, first layers): the whole row block is one contiguous piece of W,
        // copied as is (row stride K) and read with scalar loads.
        int KC = (SF / nrows - 4) & ~(4 * KW - 1);
        if (KC > ((Kp + 4 * KW - 1) / (4 * KW)) * 4 * KW) KC = ((Kp + 4 * KW - 1) / (4 * KW)) * 4 * KW;
        const bool flat = !vecW;
        const int vr = min(nrows, N - n0);
        if (flat) KC = Kp;
        const int lds = flat ? K : ld_of(KC);
        const int nchunks = flat ? 1 : (Kp + KC - 1) / KC;
        float acc[4][CN];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < CN; ++j) acc[i][j] = 0.f;
        float bb[CN];
#pragma unroll
        for (int j = 0; j < CN; ++j) {
            const int n = n0 + cw * 4 * CN + ni + 4 * j;
            bb[j] = (wactive && kw == 0 && n < N) ? __ldg(bias + n) : 0.f;
        }
        __syncthreads();                                      // stage buffers free, `in` complete, barriers ready
        if (flat) {
            const int nfl = vr * K;
            stage_flat_bulk(stage, W + (size_t)n0 * K, nfl & ~3, &bars[0]);
            if (threadIdx.x == 32)
                for (int e = nfl & ~3; e < nfl; ++e) stage[e] = __ldg(W + (size_t)n0 * K + e);
        } else {
            stage_rows_bulk(stage, lds, W, N, K, n0, nrows, 0, KC, &bars[0]);
        }
        for (int c = 0; c < nchunks; ++c) {
            mbar_wait(&bars[c & 1], (ph >> (c & 1)) & 1u);    // chunk c has landed
            ph ^= 1u << (c & 1);
            __syncthreads();                                  // everyone is done with chunk c-1
            if (c + 1 < nchunks)
                stage_rows_bulk(stage + ((c + 1) & 1) * SF, lds, W, N, K, n0, nrows, (c + 1) * KC, KC, &bars[(c + 1) & 1]);
            if (wactive && cw * 4 * CN < nrows) {
                if (flat) {
                    // rows beyond N were not staged: clamp them (their accumulators are dropped)
                    const float* sbj[CN];
#pragma unroll
                    for (int j = 0; j < CN; ++j) sbj[j] = stage + min(cw * 4 * CN + ni + 4 * j, vr - 1) * K;
                    const int ksub = (K + KW - 1) / KW;
                    const int kend = min(kw * ksub + ksub, K);
                    const float* a0 = in + mi * ldin;
                    for (int k = kw * ksub; k < kend; ++k) {
                        float a[4], b[CN];
#pragma unroll
                        for (int i = 0; i < 4; ++i) a[i] = a0[(8 * i) * ldin + k];
#pragma unroll
                        for (int j = 0; j < CN; ++j) b[j] = sbj[j][k];
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int j = 0; j < CN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
                    }
                } else {
                    const float* sb = stage + (c & 1) * SF + (cw * 4 * CN + ni) * lds;
                    const int ksub = KC / KW;                 // this warp's slice of the chunk
                    const int kbeg = kw * ksub;
                    const int kend = min(kbeg + ksub, Kp - c * KC);
                    const float* a0 = in + mi * ldin + c * KC;
                    for (int kk = kbeg; kk < kend; kk += 4) {
                        float4 a[4], b[CN];
#pragma unroll
                        for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(a0 + (8 * i) * ldin + kk);
#pragma unroll
                        for (int j = 0; j < CN; ++j) b[j] = *reinterpret_cast<const float4*>(sb + (4 * j) * lds + kk);
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int j = 0; j < CN; ++j) {
                                acc[i][j] = fmaf(a[i].x, b[j].x, acc[i][j]);
                                acc[i][j] = fmaf(a[i].y, b[j].y, acc[i][j]);
                                acc[i][j] = fmaf(a[i].z, b[j].z, acc[i][j]);
                                acc[i][j] = fmaf(a[i].w, b[j].w, acc[i][j]);
                            }
                    }
                }
            }
        }
        // k-split partners: kw = KW-1 .. 1 hand their partial tiles over through shared memory
        for (int r = KW - 1; r >= 1; --r) {
            __syncthreads();                                  // stage buffers no longer read / previous round consumed
            if (wactive && kw == r) {
                float* sc = stage + (cw * 32) * (4 * CN);
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < CN; ++j) sc[(mi + 8 * i) * (4 * CN) + ni + 4 * j] = acc[i][j];
            }
            __syncthreads();
            if (wactive && kw == 0) {
                const float* sc = stage + (cw * 32) * (4 * CN);
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < CN; ++j) acc[i][j] += sc[(mi + 8 * i) * (4 * CN) + ni + 4 * j];
            }
        }
        if (wactive && kw == 0) {
#pragma unroll
            for (int j = 0; j < CN; ++j) {
                const int n = n0 + cw * 4 * CN + ni + 4 * j;
                if (n < N) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float y = acc[i][j] + bb[j];
                        if (TANH) y = tanh_act(y);
                        out[(mi + 8 * i) * ldout + n] = y;
                    }
                }
            }
        }
    }
    __syncthreads();
    return ph;
}


// ---------------------------------------------------------------------------
// tcgen05 building blocks shared by the tensor-core layers below.
// Shared-memory operand tiles use the canonical K-major no-swizzle UMMA layout: 16-byte chunk (row r, chunk kc) at
// (r/8)*SBO + kc*LBO + (r%8)*16 with LBO = 144 B (padded: chunk stores with kc fastest are bank-conflict free).
// ---------------------------------------------------------------------------
constexpr int TC_KB = 32;                                   // k-block: 32 floats = 8 chunks
constexpr int TC_LBO = 144, TC_SBO = (TC_KB / 4) * TC_LBO;  // bytes

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t tc_off(int r, int kc) { return (r >> 3) * TC_SBO + kc * TC_LBO + (r & 7) * 16; }
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr) {
    return (uint64_t)((saddr >> 4) & 0x3fff) | ((uint64_t)(TC_LBO >> 4) << 16) | ((uint64_t)(TC_SBO >> 4) << 32) |
           ((uint64_t)1 << 46);                             // version 1, SWIZZLE_NONE, K-major
}
// round-to-nearest (ties away) of an fp32 number to tf32 with integer ALU ops: cvt.rna.tf32.f32 gives the same
// result but runs on the conversion pipe at 1/8 of the ALU rate (checked against cvt.rna in tools/tc_probe.cu)
__device__ __forceinline__ float tc_rn_tf32(float x) {
    return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}

// tmem_ld_issue only starts the load: several can be in flight before one tmem_ld_wait
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
template <int NC>
__device__ __forceinline__ void tmem_ld_issue(uint32_t taddr, uint32_t* v) {
    static_assert(NC == 8 || NC == 16, "tmem_ld: 8 or 16 columns");
    if constexpr (NC == 16) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                     : "r"(taddr));
    } else {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(taddr));
    }
}
template <int NC>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t (&v)[NC]) {
    static_assert(NC == 8 || NC == 16, "tmem_ld: 8 or 16 columns");
    if constexpr (NC == 16) {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                       "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                     : "r"(taddr));
    } else {
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                     : "r"(taddr));
    }
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// ---------------------------------------------------------------------------
// First layers (narrow input, K <= 32): out[m][n] = tanh(sum_k in[m][k] W[n][k] + b[n]) with the WHOLE weight
// matrix already in shared memory, flat [N][K], fetched by a bulk copy that was issued a phase earlier
// (first_layers_prefetch), so nothing is staged or waited for here.  Lanes are 8 (rows) x 4 (columns); a thread owns
// rows mi+8i and columns n0+ni+4j; the input row is read as 128-bit vectors, the weights as scalars (rows of K floats
// are not 16-byte aligned; consecutive columns fall into different banks because K is odd or the rows are short).
// ---------------------------------------------------------------------------
template <int CN>
__device__ __noinline__ void fwd_small_k(const float* Ws, const float* __restrict__ bias, int N, int K, const float* in,
                                         int ldin, float* out, int ldout) {
    constexpr int NWARPS = UPD_THREADS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int mi = lane >> 2, ni = lane & 3;
    const int K4 = K & ~3;
    const float* a0 = in + mi * ldin;
    for (int n0 = warp * 4 * CN; n0 < N; n0 += NWARPS * 4 * CN) {
        const float* wp[CN];
#pragma unroll
        for (int j = 0; j < CN; ++j) wp[j] = Ws + min(n0 + ni + 4 * j, N - 1) * K;   // clamped columns are dropped below
        float acc[4][CN];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < CN; ++j) acc[i][j] = 0.f;
        for (int k = 0; k < K4; k += 4) {
            float4 a[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(a0 + (8 * i) * ldin + k);
#pragma unroll
            for (int j = 0; j < CN; ++j) {
                const float b0 = wp[j][k], b1 = wp[j][k + 1], b2 = wp[j][k + 2], b3 = wp[j][k + 3];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    acc[i][j] = fmaf(a[i].x, b0, acc[i][j]);
                    acc[i][j] = fmaf(a[i].y, b1, acc[i][j]);
                    acc[i][j] = fmaf(a[i].z, b2, acc[i][j]);
                    acc[i][j] = fmaf(a[i].w, b3, acc[i][j]);
                }
            }
        }
        for (int k = K4; k < K; ++k) {
            float a[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = a0[(8 * i) * ldin + k];
#pragma unroll
            for (int j = 0; j < CN; ++j) {
                const float b = wp[j][k];
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[i][j] = fmaf(a[i], b, acc[i][j]);
            }
        }
#pragma unroll
        for (int j = 0; j < CN; ++j) {
            const int n = n0 + ni + 4 * j;
            if (n < N) {
                const float bb = __ldg(bias + n);
#pragma unroll
                for (int i = 0; i < 4; ++i) out[(mi + 8 * i) * ldout + n] = tanh_act(acc[i][j] + bb);
            }
        }
    }
    __syncthreads();
}
// both first-layer weight matrices of one network -> flat copies in the staging region (W1 at 0, V1 behind it),
// completion on `bar`; the caller orders this after the last use of the region (block barrier)
__device__ __forceinline__ int first_v1_offset(const dppo_dims& d) { return (d.H1 * d.O + 3) & ~3; }
__device__ __forceinline__ void first_layers_prefetch(float* stage, const float* __restrict__ W1, const float* __restrict__ V1,
                                                      const dppo_dims& d, uint64_t* bar) {
    const int n1 = d.H1 * d.O, n2 = d.G1 * d.O;
    float* v1s = stage + first_v1_offset(d);
    if (threadIdx.x == 0) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the region was last written through the generic proxy
        mbar_expect_tx(bar, (unsigned)(((n1 & ~3) + (n2 & ~3)) * 4));
        if (n1 & ~3) bulk_g2s(stage, W1, (unsigned)((n1 & ~3) * 4), bar);
        if (n2 & ~3) bulk_g2s(v1s, V1, (unsigned)((n2 & ~3) * 4), bar);
    }
    if (threadIdx.x == 32) {                                            // tails (visible after the next block barrier)
        for (int e = n1 & ~3; e < n1; ++e) stage[e] = __ldg(W1 + e);
        for (int e = n2 & ~3; e < n2; ++e) v1s[e] = __ldg(V1 + e);
    }
}

// ---------------------------------------------------------------------------
// Tensor-core layers: the big operand goes through TENSOR MEMORY (tcgen05.mma with A in TMEM, B in shared memory).
//   big operand:    global --TMA 2D tile, zero-filled out of bounds--> raw fp32 tile in shared memory (or, for the
//                   activation operands of the dW layers, straight from the activations in shared memory)
//                   --LDS (thread = TMEM lane), hi/lo split in registers, tcgen05.st--> A_hi | A_lo in tensor memory
//   small operand:  32 samples (or <= 128 output features for dW), split into K-major no-swizzle tiles [hi | lo]
//   MMAs:           per k-step (8 k)  acc += A_hi x [B_hi | B_lo]  (hi*hi | hi*lo in adjacent columns), then
//                   acc[hi*lo half] += A_lo x B_hi: 3xTF32, fp32 accumulation in tensor memory.  The tensor core
//                   truncates when it accumulates, so long reductions rotate over four accumulators; a dependent
//                   MMA also waits ~90 cycles (tools/mma_rate.cu), so consecutive MMAs go to different ones.
// Shared-memory traffic per k-block is 16 KB (TMA) + 16 KB (LDS) + the 32-sample tiles, against 64 KB when both split
// operands are staged in shared memory, and no weight passes through more than one thread's registers.
// Issue rules found the hard way (tools/mma_rate.cu, tools/tc_ts_probe.cu): tcgen05.mma / commit / the TMA copy are
// executed by a whole converged warp with elect.sync right next to the instruction - a lane-0 branch or a predicate
// handed in from elsewhere makes ptxas wrap each one in a per-thread loop (60-80 cycles per MMA instead of N/2);
// every branch in an eight-warp CTA costs tens of cycles, so epilogues and issue sequences are straight-line.
// ---------------------------------------------------------------------------
constexpr int TS_RAW_BYTES = 128 * TC_KB * 4;               // 16 KB: one raw weight tile
constexpr int TS_BT_BYTES = 4 * TC_SBO;                     // one 32-row sample tile (hi or lo)
constexpr int TS_BYTES = 4 * TS_RAW_BYTES + 4 * 2 * TS_BT_BYTES;   // 102 400: two ring slots of two k-blocks each
constexpr int TS_TM_A = 256;                                // TMEM: 4 x 64 accumulator columns, then 256 A columns
constexpr int TS_BAR0 = 2;                                  // first pipeline barrier in the barrier array (0, 1: FP32-path stages)
constexpr int TS_NBARS = 10;                                // raw_full[2] +0, raw_free[2] +2, a_full[2] +4, mma_done[2] +8
constexpr int TS_PH_RAW = 8, TS_PH_AF = 12, TS_PH_MD = 16;  // parity bits in `ph` (raw_free: TS_PH_RF = 10)

// D[tmem] (+)= A[tmem] * B[smem descriptor]
// The issuing warp takes elect.sync ONCE and hands the plain result to every MMA: ptxas recognises that pattern,
// moves the operands to uniform registers up front and emits the MMAs back to back.  Any predicate derived from it
// (and-ed with a condition) or produced elsewhere makes it wrap each MMA in a per-thread loop, so conditions are
// uniform branches around groups of MMAs instead.
__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
    return pred;
}
__device__ __forceinline__ void ts_mma_e(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc, uint32_t elected) {
    asm volatile(
        "{\n\t.reg .pred p, e;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 e, %5, 0;\n\t"
        "@e tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc), "r"(elected) : "memory");
}
// whole converged warp; one elected lane commits (no divergent branch in the issuing warp)
__device__ __forceinline__ void ts_commit(uint64_t* bar) {
    asm volatile("{\n\t.reg .pred e;\n\telect.sync _|e, 0xffffffff;\n\t"
                 "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
// whole converged warp: one elected lane arms `bar` with the byte count and issues one or two 2D tile loads
// (the second one when `two` is non-zero): a raw tile pair of the wide-layer pipeline
__device__ __forceinline__ void tma_pair_elect(uint32_t dst0, uint32_t dst1, const CUtensorMap* map, int a0, int b0, int a1, int b1,
                                               uint32_t bar, uint32_t bytes, uint32_t two) {
    asm volatile(
        "{\n\t.reg .pred e, t;\n\tsetp.ne.b32 t, %9, 0;\n\telect.sync _|e, 0xffffffff;\n\tand.pred t, t, e;\n\t"
        "@e mbarrier.arrive.expect_tx.shared::cta.b64 _, [%3], %8;\n\t"
        "@e cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%2, {%4, %5}], [%3];\n\t"
        "@t cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%1], [%2, {%6, %7}], [%3];\n\t}"
        ::"r"(dst0), "r"(dst1), "l"(map), "r"(bar), "r"(a0), "r"(b0), "r"(a1), "r"(b1), "r"(bytes), "r"(two) : "memory");
}
// two more tiles onto a barrier whose arrival comes later (tma_pair_elect): only the byte count is announced here
__device__ __forceinline__ void tma_pair_more_elect(uint32_t dst0, uint32_t dst1, const CUtensorMap* map, int a0, int b0, int a1, int b1,
                                                    uint32_t bar, uint32_t bytes, uint32_t two) {
    asm volatile(
        "{\n\t.reg .pred e, t;\n\tsetp.ne.b32 t, %9, 0;\n\telect.sync _|e, 0xffffffff;\n\tand.pred t, t, e;\n\t"
        "@e mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%3], %8;\n\t"
        "@e cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%2, {%4, %5}], [%3];\n\t"
        "@t cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%1], [%2, {%6, %7}], [%3];\n\t}"
        ::"r"(dst0), "r"(dst1), "l"(map), "r"(bar), "r"(a0), "r"(b0), "r"(a1), "r"(b1), "r"(bytes), "r"(two) : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&v)[32]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
                 "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                   "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
                   "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31]) : "memory");
}
// compiler barrier that needs 16 registers to hold their final values (used by the DPPO_K2_EARLY_RELEASE experiment)
__device__ __forceinline__ void ts_pin16(const uint32_t* v) {
    asm volatile("" ::"r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]),
                 "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]) : "memory");
}
// hi = rn_tf32(x); lo = x - hi exactly (<= 13 significant bits): the tensor core truncates it to tf32 (error <= 2^-22 |x|)
__device__ __forceinline__ void ts_split(float x, uint32_t& hi, uint32_t& lo) {
    const float h = tc_rn_tf32(x);
    hi = __float_as_uint(h);
    lo = __float_as_uint(x - h);
}
// the same block by ONE warp, straight-line: lane = (row group of 4, chunk), eight rows per lane.  Load and store are
// separate so that the loads of the next block can be in flight while the warp waits for its slot (the 32-row operand
// of the batch-level dW pass comes from global memory)
__device__ __forceinline__ void ts_block_load(float4 (&xv)[8], const float* src, int ld, int k0, int K, int lane) {
    const int kc = lane & 7, m0 = lane >> 3, k = k0 + 4 * kc;
#pragma unroll
    for (int i = 0; i < 8; ++i)
        xv[i] = (k < K) ? *reinterpret_cast<const float4*>(src + (size_t)(m0 + 4 * i) * ld + k) : make_float4(0.f, 0.f, 0.f, 0.f);
}
__device__ __forceinline__ void ts_block_store(unsigned char* bh, const float4 (&xv)[8], int lane) {
    const int kc = lane & 7, m0 = lane >> 3;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        uint32_t hh[4], ll[4];
        ts_split(xv[i].x, hh[0], ll[0]); ts_split(xv[i].y, hh[1], ll[1]); ts_split(xv[i].z, hh[2], ll[2]); ts_split(xv[i].w, hh[3], ll[3]);
        const uint32_t o = tc_off(m0 + 4 * i, kc);
        *reinterpret_cast<uint4*>(bh + o) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        *reinterpret_cast<uint4*>(bh + TS_BT_BYTES + o) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
    }
}
__device__ __forceinline__ void ts_stage_block_warp(unsigned char* bh, const float* src, int ld, int k0, int K, int lane) {
    float4 xv[8];
    ts_block_load(xv, src, ld, k0, K, lane);
    ts_block_store(bh, xv, lane);
}
// the 32 samples' columns [k0, k0+32) of src[32][ld] -> [hi | lo] K-major tiles, by `nthr` threads (t = 0..nthr-1)
__device__ __forceinline__ void ts_stage_samples(unsigned char* bh, const float* src, int ld, int k0, int K, int t, int nthr) {
    for (int e = t; e < 8 * UPD_ROWS; e += nthr) {
        const int m = e >> 3, kc = e & 7, k = k0 + 4 * kc;
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < K) x = *reinterpret_cast<const float4*>(src + m * ld + k);
        uint32_t hh[4], ll[4];
        ts_split(x.x, hh[0], ll[0]); ts_split(x.y, hh[1], ll[1]); ts_split(x.z, hh[2], ll[2]); ts_split(x.w, hh[3], ll[3]);
        const uint32_t o = tc_off(m, kc);
        *reinterpret_cast<uint4*>(bh + o) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        *reinterpret_cast<uint4*>(bh + TS_BT_BYTES + o) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
    }
}
// ---------------------------------------------------------------------------
// The wide-layer pipeline (forward and dX share it).  Ring granule = a PAIR of k-blocks (64 reduction elements):
// every barrier round trip, tcgen05.commit (which drains the MMA queue) and loop iteration is paid once per pair.
// Two ring slots; slot s holds two raw tiles (2 x 16 KB), two sample tile pairs, 128 A columns of tensor memory.
//   warps 0-3  transform: raw tile -> registers -> hi/lo -> tensor memory (warp = TMEM lane quadrant)
//   warps 4,5  samples:   one block of the pair each (activations -> [hi | lo] K-major tiles)
//   warp 6     TMA producer: refills a slot as soon as the transform warps have consumed it (raw_free)
//   warp 7     MMA issuer (whole warp, uniform operands, elected lane), one commit per pair
// Barriers (index from TS_BAR0; parity bit in `ph`):  raw_full[2] +0 (bits 8,9)   raw_free[2] +2 (bits 10,11)
//                                                      a_full[2]  +4 (bits 12,13)  mma_done[2] +8 (bits 16,17)
// DX = false: out[m][n] = tanh(sum_k in[m][k] W[n][k] + b[n]); map box 32 (k) x 128 (n), SWIZZLE_128B; tiles of 128 n.
// DX = true : act[m][k] <- (sum_n dOut[m][n] W[n][k]) (1 - act[m][k]^2); map box 128 (k) x 32 (n), no swizzle;
//             tiles of 128 k, the rings run across tiles, the accumulators are drained between tiles.
// ---------------------------------------------------------------------------
constexpr int TS_PH_RF = 10;
// Extras of a wide-layer call (all zero / null for a plain layer):
//   kofs/row0  offsets of the TMA coordinates (reduction index / first tile row)
//   dwg        batch-level dW pass: the "weights" are dz^T [n][samples] (feature-major, TMA), the 32-row operand is a
//              block of 32 input features in^T [kf][samples] read from global memory, the reduction runs over a range
//              of samples, and out[kf][n] = dW[n][kf] goes to this split's partial bucket instead of the activations:
//              gW[n * ldw + kf0 + kf] for kf0 + kf < Kf, the column kf0 + kf == Kf (the ones row) is the bias gradient
//              gb[n]; rows n >= n1 keep only that column (gb2[n - n1] + add2: d log_std rides with the mu head).
//              tr: the roles are swapped (the heads: few dz rows, many inputs): the tile's rows are INPUT features (row Kf =
//              the ones row) and the 32-row operand is a block of dz rows, out[m][n] = dW[m][n] goes to gW[m * ldw + n].
//              The 32-row operand's blocks arrive by TMA too (bmap, rows brow0.., raw tiles in braw) on the same barrier
//              as the dz tiles; the sample warps split them from shared memory as in every other layer.
struct TsX {
    int kofs, row0;
    int dwg, tr;
    const CUtensorMap* bmap; int brow0; float* braw;   // dwg: the 32-row operand's blocks by TMA (box 32 x 32) into braw[slot][block][32][32]
    float *gW, *gb, *gb2;
    int ldw, kf0, Kf, n1;
    float add2;
};
template <bool DX>
__device__ __noinline__ unsigned ts_layer(const CUtensorMap* map, const float* __restrict__ bias, int N, int K,
                                          const float* bsrc, int ldb, float* dst, int ldo, unsigned char* ts, uint64_t* bars,
                                          unsigned ph, uint32_t tmem, const TsX& Xref) {
    const TsX X = Xref;                                     // in registers: nothing below may reload it after a store
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int R = DX ? N : K;                               // reduction length
    const int nrb = (R + TC_KB - 1) / TC_KB;                // k-blocks per tile
    const int npt = (nrb + 1) >> 1;                         // pairs per tile
    const int ntiles = DX ? (K + 127) >> 7 : (N + 127) >> 7;   // tiles of 128 TMEM lanes: k (dX) / output features (forward)
    const int NP = npt * ntiles;                            // pairs in all
    uint64_t* raw_full = bars + TS_BAR0;
    uint64_t* raw_free = raw_full + 2;
    uint64_t* a_full = raw_full + 4;
    uint64_t* mma_done = raw_full + 8;
    unsigned char* raw = ts;
    unsigned char* bst = ts + 4 * TS_RAW_BYTES;
    // parity of the u-th use (u = 0, 1, ...) of a barrier whose tracked bit is `bit`
    auto par = [&](int bit, int u) -> unsigned { return ((ph >> bit) ^ (unsigned)u) & 1u; };
    auto tma_pair = [&](int gp) {                           // (whole warp 6) both raw tiles of pair gp into slot gp & 1
        const int s = gp & 1, mt = gp / npt, p = gp - mt * npt;
        const uint32_t two = (nrb - 2 * p >= 2) ? 1u : 0u;
        const uint32_t d0 = smem_u32(raw + (2 * s) * TS_RAW_BYTES), bar = smem_u32(&raw_full[s]);
        const int blk = 2 * p;
        if (!DX && X.bmap != nullptr) {                      // dwg: the 32-row operand's two blocks (4 KB each) first
            const uint32_t b0 = smem_u32(X.braw + (2 * s) * 1024);
            tma_pair_more_elect(b0, b0 + 4096, X.bmap, TC_KB * blk + X.kofs, X.brow0, TC_KB * (blk + 1) + X.kofs, X.brow0, bar,
                                (1u + two) * 4096u, two);
        }
        if (DX) tma_pair_elect(d0, d0 + TS_RAW_BYTES, map, 128 * mt, TC_KB * blk, 128 * mt, TC_KB * (blk + 1), bar, (1u + two) * TS_RAW_BYTES, two);
        else tma_pair_elect(d0, d0 + TS_RAW_BYTES, map, TC_KB * blk + X.kofs, 128 * mt + X.row0, TC_KB * (blk + 1) + X.kofs, 128 * mt + X.row0, bar, (1u + two) * TS_RAW_BYTES, two);
    };
    __syncthreads();                                        // previous users of the staging region are done
#ifdef DPPO_K2_TRACE
    const bool trace_on = blockIdx.x == 0 && !DX && (g_k2_trace_bias ? bias == g_k2_trace_bias : K > 256);
    const long long trace_t0 = clock64();
#endif
    if (warp == 6) {
        tma_pair(0);
        if (NP > 1) tma_pair(1);
    }
    for (int mt = 0; mt < ntiles; ++mt) {
        const int gp0 = mt * npt;
        if (warp < 4) {
            // ---------------- transform ----------------
            const int r = 32 * warp + lane;                 // TMEM lane: weight row (forward) / k within the tile (dX)
            for (int gp = gp0; gp < gp0 + npt; ++gp) {
                const int s = gp & 1, u = gp >> 1;
                const int nb2 = min(2, nrb - 2 * (gp - gp0));
                if (warp == 0) K2_TR(0, gp, 0);
                mbar_wait(&raw_full[s], par(TS_PH_RAW + s, u));
                if (warp == 0) K2_TR(0, gp, 1);
                const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16) + TS_TM_A + 128 * s;
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    if (i < nb2) {
                        uint32_t hi[32], lo[32];
                        const unsigned char* tile = raw + (2 * s + i) * TS_RAW_BYTES;
                        if (DX) {
                            const float* colp = reinterpret_cast<const float*>(tile) + r;
#pragma unroll
                            for (int n = 0; n < 32; ++n) ts_split(colp[n * 128], hi[n], lo[n]);
                        } else {
                            const unsigned char* rowp = tile + r * 128;
#pragma unroll
                            for (int c = 0; c < 8; ++c) {   // 128-byte swizzle: chunk c of row r sits at c ^ (r % 8)
                                const float4 x = *reinterpret_cast<const float4*>(rowp + ((c ^ (r & 7)) << 4));
                                ts_split(x.x, hi[4 * c + 0], lo[4 * c + 0]); ts_split(x.y, hi[4 * c + 1], lo[4 * c + 1]);
                                ts_split(x.z, hi[4 * c + 2], lo[4 * c + 2]); ts_split(x.w, hi[4 * c + 3], lo[4 * c + 3]);
                            }
                        }
#if DPPO_K2_EARLY_RELEASE
                        // EXPERIMENT (never the shipped build; see "raw tile release" below): release the slot as soon as
                        // its last tile has been split.  =1: as ptxas schedules it (the arrive overtakes the split);
                        // =2: with the split results pinned in registers first.
                        if (i == nb2 - 1) {
#if DPPO_K2_EARLY_RELEASE == 2
                            ts_pin16(hi); ts_pin16(hi + 16); ts_pin16(lo); ts_pin16(lo + 16);
#endif
                            __syncwarp();
                            if (lane == 0) mbar_arrive(&raw_free[s]);
                        }
#endif
                        if (i == 0) {                       // the slot's previous MMAs have read these A columns
                            if (warp == 0) K2_TR(0, gp, 2);
                            if (gp >= 2) mbar_wait(&mma_done[s], par(TS_PH_MD + s, u - 1));
                            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                            if (warp == 0) K2_TR(0, gp, 3);
                        }
                        tmem_st32(ta + 64 * i, hi);
                        tmem_st32(ta + 64 * i + 32, lo);
                    }
                }
                if (warp == 0) K2_TR(0, gp, 4);
                // Raw tile release.  The producer warp refills the slot (TMA, async proxy) as soon as all four transform
                // warps have arrived on raw_free, and nothing orders a shared-memory load that has been ISSUED but has not
                // yet read the array against that refill: the arrive does not depend on the loaded registers, and ptxas
                // sinks the split arithmetic (the only consumer of the loads) below it.  Round 1's per-tile release right
                // after the split did exactly that and lost the race in the dX pass of the 500-wide layer (21-25 of 300
                // launches; the same kernel with the split results pinned in registers before the arrive is bit-clean:
                // profiles/r02_k2_release_race.md).  Rule: release a TMA destination only after an instruction that
                // consumes every value loaded from it.  Here the arrive follows both tiles' tcgen05.st in program order;
                // the st read the split results, the split results need every LDS destination register, so all loads of
                // the slot have completed when the arrive issues - a data dependence, not a time window.
#if !DPPO_K2_EARLY_RELEASE
                __syncwarp();
                if (lane == 0) mbar_arrive(&raw_free[s]);
#endif
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&a_full[s]);
                if (warp == 0) K2_TR(0, gp, 5);
            }
        } else if (warp < 6) {
            // ---------------- samples: warp 4 stages the first block of every pair, warp 5 the second ----------------
            const int i = warp - 4;
            for (int gp = gp0; gp < gp0 + npt; ++gp) {
                const int s = gp & 1, u = gp >> 1, blk = 2 * (gp - gp0) + i;
                K2_TR(1, gp, 0);
                if (gp >= 2) mbar_wait(&mma_done[s], par(TS_PH_MD + s, u - 1));
                K2_TR(1, gp, 1);
                if (!DX && X.bmap != nullptr) {             // dwg: this pair's raw blocks have landed with the dz tiles
                    mbar_wait(&raw_full[s], par(TS_PH_RAW + s, u));
                    if (blk < nrb) ts_stage_block_warp(bst + (2 * s + i) * 2 * TS_BT_BYTES, X.braw + (2 * s + i) * 1024, 32, 0, 32, lane);
                } else if (blk < nrb) {
                    ts_stage_block_warp(bst + (2 * s + i) * 2 * TS_BT_BYTES, bsrc, ldb, blk * TC_KB, R, lane);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) mbar_arrive(&a_full[s]);
                K2_TR(1, gp, 2);
            }
        } else if (warp == 6) {
            // ---------------- TMA producer ----------------
            for (int gp = gp0; gp < gp0 + npt; ++gp)
                if (gp + 2 < NP) {
                    K2_TR(2, gp, 0);
                    mbar_wait(&raw_free[gp & 1], par(TS_PH_RF + (gp & 1), gp >> 1));
                    if (!DX && X.bmap != nullptr)           // dwg: the sample warps have read their raw blocks too
                        mbar_wait(&a_full[gp & 1], par(TS_PH_AF + (gp & 1), gp >> 1));
                    K2_TR(2, gp, 1);
                    tma_pair(gp + 2);
                    K2_TR(2, gp, 2);
                }
        } else {
            // ---------------- MMA issuer ----------------
            const uint32_t idesc64 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            const uint32_t idesc32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            constexpr uint64_t KSTEP = (2 * TC_LBO) >> 4;
            const uint32_t el = elect_one();
            for (int gp = gp0; gp < gp0 + npt; ++gp) {
                const int s = gp & 1, u = gp >> 1;
                const bool two = nrb - 2 * (gp - gp0) >= 2;                         // second block of the pair present
                K2_TR(3, gp, 0);
                mbar_wait(&a_full[s], par(TS_PH_AF + s, u));
                K2_TR(3, gp, 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint64_t bd0 = tc_desc(smem_u32(bst + (2 * s) * 2 * TS_BT_BYTES));
                const uint64_t bd1 = tc_desc(smem_u32(bst + (2 * s + 1) * 2 * TS_BT_BYTES));
                const uint32_t ta = tmem + TS_TM_A + 128 * s;
                // main terms into the four rotating accumulators, the cross terms (into their hi*lo halves) four MMAs
                // later: consecutive MMAs never touch the same accumulator
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) ts_mma_e(tmem + 64 * ks, ta + 8 * ks, bd0 + ks * KSTEP, idesc64, gp > gp0 ? 1u : 0u, el);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) ts_mma_e(tmem + 64 * ks + 32, ta + 32 + 8 * ks, bd0 + ks * KSTEP, idesc32, 1u, el);
                if (two) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) ts_mma_e(tmem + 64 * ks, ta + 64 + 8 * ks, bd1 + ks * KSTEP, idesc64, 1u, el);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) ts_mma_e(tmem + 64 * ks + 32, ta + 96 + 8 * ks, bd1 + ks * KSTEP, idesc32, 1u, el);
                }
                ts_commit(&mma_done[s]);
                K2_TR(3, gp, 2);
            }
            // the tile's last commit covers all its MMAs; the issuer has seen every phase of these barriers, the other
            // warps have not (a parity wait could let them through early), so they meet it at the block barrier
            const int gl = gp0 + npt - 1;
            mbar_wait(&mma_done[gl & 1], par(TS_PH_MD + (gl & 1), gl >> 1));
            K2_TR(3, gl, 3);
        }
        __syncthreads();
        if (warp == 0) K2_TR(0, 63, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // ---------------- epilogue: lane = feature, this warp's 16 samples; the four small halves first ----------------
        {
            const int q = warp & 3, f = 128 * mt + 32 * q + lane, c0 = 16 * (warp >> 2);
            const float bb = (!DX && bias != nullptr && f < N) ? __ldg(bias + f) : 0.f;   // forward: bias of my feature (in flight during the loads)
            float sum[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) sum[j] = 0.f;
            const uint32_t tb = tmem + ((uint32_t)(32 * q) << 16) + c0;
#pragma unroll
            for (int hf = 0; hf < 2; ++hf) {
                uint32_t v[4][16];
#pragma unroll
                for (int t = 0; t < 4; ++t) tmem_ld_issue<16>(tb + 64 * t + (hf ? 0 : 32), v[t]);
                tmem_ld_wait();
#pragma unroll
                for (int t = 0; t < 4; ++t)
#pragma unroll
                    for (int j = 0; j < 16; ++j) sum[j] += __uint_as_float(v[t][j]);
            }
            if (DX) {
                if (f < K) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        float* ap = dst + (c0 + j) * ldo + f;
                        const float h = *ap;
                        sum[j] *= 1.f - h * h;
                        *ap = sum[j];
                    }
                }
            } else if (X.dwg && X.tr) {
                const int n = X.row0 + f;                   // input feature (n == Kf: the ones row -> bias gradients)
                if (n <= X.Kf) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const int m = X.kf0 + c0 + j;       // dz row
                        if (m < X.n1) {
                            if (n < X.Kf) X.gW[(size_t)m * X.ldw + n] = sum[j];
                            else X.gb[m] = sum[j];
                        } else if (X.gb2 != nullptr && m < 2 * X.n1 && n == X.Kf) {
                            X.gb2[m - X.n1] = sum[j] + X.add2;
                        }
                    }
                }
            } else if (X.dwg) {
                const int n = X.row0 + f;                   // f counts from the first tile row of this call
                if (n < X.n1) {
                    float* row = X.gW + (size_t)n * X.ldw;
                    const int kfa = X.kf0 + c0;
                    if ((X.ldw & 3) == 0 && kfa + 16 <= X.Kf) {       // whole 64 bytes inside the row, 16-byte aligned
#pragma unroll
                        for (int j = 0; j < 4; ++j)
                            reinterpret_cast<float4*>(row + kfa)[j] = make_float4(sum[4 * j], sum[4 * j + 1], sum[4 * j + 2], sum[4 * j + 3]);
                    } else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            if (kfa + j < X.Kf) row[kfa + j] = sum[j];
                            else if (kfa + j == X.Kf) X.gb[n] = sum[j];
                        }
                    }
                } else if (n < N + X.row0 && X.gb2 != nullptr) {
#pragma unroll
                    for (int j = 0; j < 16; ++j)
                        if (X.kf0 + c0 + j == X.Kf) X.gb2[n - X.n1] = sum[j] + X.add2;
                }
            } else {
                if (f < N) {
#pragma unroll
                    for (int j = 0; j < 16; ++j) { sum[j] = tanh_act(sum[j] + bb); dst[(c0 + j) * ldo + f] = sum[j]; }
                }
            }
        }
        if (warp == 0) K2_TR(0, 63, 1);
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
    }
    // every phase of every ring barrier has completed: advance the tracked parities by the number of uses
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        const unsigned uses = (unsigned)(((NP + 1 - s) >> 1) & 1);
        ph ^= (uses << (TS_PH_RAW + s)) | (uses << (TS_PH_RF + s)) | (uses << (TS_PH_AF + s)) | (uses << (TS_PH_MD + s));
    }
    return ph;
}

// MMAs of the first-layer routines for NT tiles (compile time) and nks k-steps: k-step outer / tile inner, so that
// consecutive MMAs go to different accumulators (a dependent MMA waits ~90 cycles); plain elected predicate
template <int NT>
__device__ __forceinline__ void first_issue(uint32_t tmem, uint64_t bd, int nks, uint32_t el) {
    const uint32_t idesc64 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    const uint32_t idesc32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(32 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
    constexpr uint64_t KSTEP = (2 * TC_LBO) >> 4;
    for (int ks = 0; ks < nks; ++ks)
#pragma unroll
        for (int t = 0; t < NT; ++t)
            ts_mma_e(tmem + 64 * t, tmem + TS_TM_A + 64 * t + 8 * ks, bd + ks * KSTEP, idesc64, ks ? 1u : 0u, el);
    for (int ks = 0; ks < nks; ++ks)
#pragma unroll
        for (int t = 0; t < NT; ++t)
            ts_mma_e(tmem + 64 * t + 32, tmem + TS_TM_A + 64 * t + 32 + 8 * ks, bd + ks * KSTEP, idesc32, 1u, el);
}
__device__ __forceinline__ void first_issue_all(int ntiles, uint32_t tmem, uint64_t bd, int nks) {
    const uint32_t el = elect_one();
    if (ntiles == 1) first_issue<1>(tmem, bd, nks, el);
    else if (ntiles == 2) first_issue<2>(tmem, bd, nks, el);
    else if (ntiles == 3) first_issue<3>(tmem, bd, nks, el);
    else first_issue<4>(tmem, bd, nks, el);
}

// ---------------------------------------------------------------------------
// First layers (K = O <= 32 inputs) on the tensor cores.  Up to four tiles of 128 output features are transformed
// into tensor memory at once (4 x 64 A columns next to 4 x 64 accumulator columns), so there is no pipeline: stage,
// one block barrier, all MMAs, one commit, epilogue.  The weights come from the flat copy in shared memory that
// first_layers_prefetch fetched a phase earlier; the samples' tiles live behind those copies in the staging region.
//   forward  D[n][m] = W[n][:] . x[m][:]        A = W rows (lane = n, columns = k),   B = x       [hi | lo], K-major over k
//   dW       D[n][k] = dOut[:][n] . x[:][k]     A = dOut^T (lane = n, columns = s),   B = x^T     [hi | lo], K-major over s
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned char* first_b_tiles(float* stage, const dppo_dims& d) {
    const int floats = (first_v1_offset(d) + d.G1 * d.O + 31) & ~31;
    return reinterpret_cast<unsigned char*>(stage + floats);
}
__device__ __noinline__ unsigned fwd_first_ts(const float* Ws, const float* __restrict__ bias, int N, int K, const float* in,
                                              int ldin, float* out, int ldout, unsigned char* bt, uint64_t* bars, unsigned ph,
                                              uint32_t tmem) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntiles = (N + 127) >> 7, nks = (K + 7) >> 3;
    uint64_t* mma_done = bars + TS_BAR0 + 8;
    float bb[4];                                            // biases of my output feature in each tile: fetched now, used in the epilogue
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int n = 128 * t + 32 * (warp & 3) + lane;
        bb[t] = n < N ? __ldg(bias + n) : 0.f;
    }
    // ---- B: the samples, one 16-byte chunk per thread ----
    ts_stage_samples(bt, in, ldin, 0, K, tid, UPD_THREADS);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    // ---- A: weight rows -> tensor memory, all tiles ----
    if (warp < 4) {
        for (int t = 0; t < ntiles; ++t) {
            const int n = 128 * t + 32 * warp + lane;
            const float* wr = Ws + min(n, N - 1) * K;
            uint32_t hi[32], lo[32];
#pragma unroll
            for (int k = 0; k < 32; ++k) ts_split((n < N && k < K) ? wr[k] : 0.f, hi[k], lo[k]);
            const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16) + TS_TM_A + 64 * t;
            tmem_st32(ta, hi);
            tmem_st32(ta + 32, lo);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 7) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint64_t bd = tc_desc(smem_u32(bt));
        first_issue_all(ntiles, tmem, bd, nks);
        ts_commit(&mma_done[0]);
    }
    mbar_wait(&mma_done[0], (ph >> TS_PH_MD) & 1u);
    ph ^= 1u << TS_PH_MD;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    {
        const int q = warp & 3, c0 = 16 * (warp >> 2);
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            if (t >= ntiles) break;
            const int n = 128 * t + 32 * q + lane;
            uint32_t v[2][16];
            const uint32_t tb = tmem + ((uint32_t)(32 * q) << 16) + 64 * t + c0;
            tmem_ld_issue<16>(tb + 32, v[0]);
            tmem_ld_issue<16>(tb, v[1]);
            tmem_ld_wait();
            if (n < N) {
                float o[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    o[j] = tanh_act((__uint_as_float(v[0][j]) + __uint_as_float(v[1][j])) + bb[t]);
                    out[(c0 + j) * ldout + n] = o[j];
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    return ph;
}

// dW / db of a first layer (see above).  img: N*K floats of shared memory (the flat image is flushed coalesced).
__device__ __noinline__ unsigned dw_first_ts(float* __restrict__ gW, float* __restrict__ gb, int N, int K, const float* dOut,
                                             int ldd, const float* in, int ldin, bool accumulate, float* img, unsigned char* bt,
                                             uint64_t* bars, unsigned ph, uint32_t tmem) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntiles = (N + 127) >> 7;
    uint64_t* mma_done = bars + TS_BAR0 + 8;
    __syncthreads();                                        // staging region free, dOut complete
    // ---- B = x^T: row k (< 32), chunk kc = samples 4kc..4kc+3 ----
    {
        const int k = tid & 31, kc = tid >> 5;              // 256 chunks, one per thread
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < K) {
            const float* p = in + (4 * kc) * ldin + k;
            x = make_float4(p[0], p[ldin], p[2 * ldin], p[3 * ldin]);
        }
        uint32_t hh[4], ll[4];
        ts_split(x.x, hh[0], ll[0]); ts_split(x.y, hh[1], ll[1]); ts_split(x.z, hh[2], ll[2]); ts_split(x.w, hh[3], ll[3]);
        const uint32_t o = tc_off(k, kc);
        *reinterpret_cast<uint4*>(bt + o) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
        *reinterpret_cast<uint4*>(bt + TS_BT_BYTES + o) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp < 4) {
        for (int t = 0; t < ntiles; ++t) {
            const int n = 128 * t + 32 * warp + lane;
            const float* cp = dOut + min(n, N - 1);
            uint32_t hi[32], lo[32];
#pragma unroll
            for (int sidx = 0; sidx < 32; ++sidx) ts_split(n < N ? cp[sidx * ldd] : 0.f, hi[sidx], lo[sidx]);
            const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16) + TS_TM_A + 64 * t;
            tmem_st32(ta, hi);
            tmem_st32(ta + 32, lo);
        }
        asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 7) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint64_t bd = tc_desc(smem_u32(bt));
        first_issue_all(ntiles, tmem, bd, 4);
        ts_commit(&mma_done[0]);
    }
    mbar_wait(&mma_done[0], (ph >> TS_PH_MD) & 1u);
    ph ^= 1u << TS_PH_MD;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    {
        const int q = warp & 3, c0 = 16 * (warp >> 2);      // this warp's 16 of the 32 k columns
        for (int t = 0; t < ntiles; ++t) {
            const int n = 128 * t + 32 * q + lane;
            uint32_t v[2][16];
            const uint32_t tb = tmem + ((uint32_t)(32 * q) << 16) + 64 * t + c0;
            tmem_ld_issue<16>(tb + 32, v[0]);
            tmem_ld_issue<16>(tb, v[1]);
            tmem_ld_wait();
            if (n < N) {
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (c0 + j < K) img[n * K + c0 + j] = __uint_as_float(v[0][j]) + __uint_as_float(v[1][j]);
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    const int tot = N * K;
    if ((tot & 3) == 0) {
        float4* g4 = reinterpret_cast<float4*>(gW);
        const float4* s4 = reinterpret_cast<const float4*>(img);
        for (int e = tid; e < (tot >> 2); e += UPD_THREADS) {
            float4 o = s4[e];
            if (accumulate) { const float4 p4 = g4[e]; o.x += p4.x; o.y += p4.y; o.z += p4.z; o.w += p4.w; }
            g4[e] = o;
        }
    } else {
        for (int e = tid; e < tot; e += UPD_THREADS) gW[e] = accumulate ? gW[e] + img[e] : img[e];
    }
    for (int n = tid; n < N; n += UPD_THREADS) {
        float sb = 0.f;
        for (int m = 0; m < 32; ++m) sb += dOut[m * ldd + n];
        gb[n] = accumulate ? gb[n] + sb : sb;
    }
    __syncthreads();
    return ph;
}

template <bool TANH>
__device__ __forceinline__ unsigned fwd_layer(const float* __restrict__ W, const float* __restrict__ bias, int N, int K,
                                          const float* in, int ldin, float* out, int ldout, float* stage, int SF,
                                          uint64_t* bars, unsigned ph, uint32_t tmem, const CUtensorMap* wmap) {
    // wide GEMM-shaped layers go to the tensor cores (tcgen05), the rest stays on the FP32 pipe
    if (wmap != nullptr && N >= 32 && K >= 64 && (K & 3) == 0) {        // = ts_eligible() on the host
        TsX x = {};
        return ts_layer<false>(wmap, bias, N, K, in, ldin, out, ldout, reinterpret_cast<unsigned char*>(stage), bars, ph, tmem, x);
    }
    // 4*CN columns per warp: 28 fits N = 100 (reference widths) in four blocks, 32 otherwise
    const int blocks32 = (N + 31) / 32;
    if (N > 28 * (blocks32 - 1) && N <= 28 * blocks32 && blocks32 <= 8)
        return fwd_layer_t<7, TANH>(W, bias, N, K, in, ldin, out, ldout, stage, SF, bars, ph);
    return fwd_layer_t<8, TANH>(W, bias, N, K, in, ldin, out, ldout, stage, SF, bars, ph);
}

// head weights (mu: [A][ldH2] + bias[A], v: [ldG2] + bias) of one network, resident in shared
// memory for the whole kernel (they are tiny and every tile needs them four times)
struct HeadS { const float *Wmu, *bmu, *Wv, *bv; int ldw; };   // ldw: row stride of Wmu

__device__ __forceinline__ void load_heads(float* dst, const Net& nw, const UpdArgs& P) {
    const dppo_dims& d = P.d;
    for (int e = threadIdx.x; e < d.A * d.H2; e += UPD_THREADS) {
        const int n = e / d.H2, k = e - n * d.H2;
        dst[n * P.ldH2 + k] = __ldg(nw.Wmu + e);
    }
    float* b = dst + d.A * P.ldH2;
    for (int e = threadIdx.x; e < d.A; e += UPD_THREADS) b[e] = __ldg(nw.bmu + e);
    float* wv = b + ((d.A + 3) & ~3);
    for (int e = threadIdx.x; e < d.G2; e += UPD_THREADS) wv[e] = __ldg(nw.Wv + e);
    if (threadIdx.x == 0) wv[P.ldG2] = __ldg(nw.bv);
}
__device__ __forceinline__ HeadS heads_at(const float* base, const UpdArgs& P) {
    HeadS h;
    h.ldw = P.ldH2;
    h.Wmu = base;
    h.bmu = base + P.d.A * P.ldH2;
    h.Wv = h.bmu + ((P.d.A + 3) & ~3);
    h.bv = h.Wv + P.ldG2;
    return h;
}

// both heads of one network in one pass: thread (m, n) does one dot product from shared memory
//   n < A : mu[m][n] = h2[m] . Wmu[n] + bmu[n]         n == A : v[m] = g2[m] . Wv + bv
__device__ __noinline__ void heads_fwd(HeadS hs, int A, int H2, int G2, const float* h2, int ldh2, const float* g2,
                                       int ldg2, float* mu_out, int ldA, float* v_out, int mode = 0) {
    for (int o = threadIdx.x; o < UPD_ROWS * (A + 1); o += UPD_THREADS) {
        const int m = o / (A + 1), n = o - m * (A + 1);
        const bool isv = n == A;
        if ((mode == 1 && isv) || (mode == 2 && !isv)) continue;   // one trunk at a time (trunk-serial tiles)
        const float* x = isv ? g2 + m * ldg2 : h2 + m * ldh2;
        const float* w = isv ? hs.Wv : hs.Wmu + n * hs.ldw;
        const int K4 = ((isv ? G2 : H2) + 3) >> 2;          // pad columns: activations finite, weights 0
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
        for (int k = 0; k < K4; ++k) {
            const float4 a = *reinterpret_cast<const float4*>(x + 4 * k);
            const float4 b = *reinterpret_cast<const float4*>(w + 4 * k);
            s0 = fmaf(a.x, b.x, s0); s1 = fmaf(a.y, b.y, s1); s2 = fmaf(a.z, b.z, s2); s3 = fmaf(a.w, b.w, s3);
        }
        const float s = (s0 + s1) + (s2 + s3);
        if (isv) v_out[m] = s + hs.bv[0];
        else mu_out[m * ldA + n] = s + hs.bmu[n];
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// dX through a tanh layer, in place:
//   act[m][k] <- (sum_n dOut[m][n] W[n][k]) * (1 - act[m][k]^2),  k < K
// dOut: smem [32][ldd] (zero beyond N); W global [N][K]; W rows stream through the stage.
// Lanes are 8 (rows) x 4 (k-quads); a thread owns rows mi+8i and quads q0+ki+4j (j < KT).
// ---------------------------------------------------------------------------
// rows [n0, n0+nc) of W[N][K] -> dst[nn][Kp] for the dX pass.  K % 4 == 0: the rows are one
// contiguous block (one bulk copy); otherwise element-wise.  Rows >= N are zero-filled (they meet
// dOut's zero padding; garbage there could be NaN).
__device__ __forceinline__ void dx_stage(float* dst, const float* __restrict__ W, int N, int K, int Kp, int n0, int nc,
                                         bool vec, uint64_t* bar) {
    const int vr = max(0, min(nc, N - n0));
    for (int e = vr * Kp + threadIdx.x; e < nc * Kp; e += UPD_THREADS) dst[e] = 0.f;
    if (vec) {
        stage_flat_bulk(dst, W + (size_t)n0 * K, vr * K, bar);
    } else {
        for (int e = threadIdx.x; e < vr * Kp; e += UPD_THREADS) {
            const int r = e / Kp, k = e - r * Kp;
            dst[e] = k < K ? __ldg(W + (size_t)(n0 + r) * K + k) : 0.f;
        }
        if (threadIdx.x == 0) mbar_expect_tx(bar, 0);
    }
}

template <int KT>
__device__ __noinline__ unsigned dx_layer_t(const float* __restrict__ W, int N, int K, const float* dOut, int ldd, float* act, int lda,
                           float* stage, int SF, uint64_t* bars, unsigned ph) {
    constexpr int NWARPS = UPD_THREADS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int mi = lane >> 2, ki = lane & 3;
    const int KQ = (K + 3) >> 2, Kp = KQ * 4;
    const bool vecW = (K & 3) == 0;
    const int Np = (N + 3) & ~3;
    int NC = (SF / Kp) & ~3;                                 // rows of W per chunk
    if (NC > Np) NC = Np;
    const int nchunks = (Np + NC - 1) / NC;
    for (int q0 = 0; q0 < KQ; q0 += NWARPS * 4 * KT) {       // usually a single pass
        const int qb = q0 + warp * 4 * KT + ki;
        float acc[4][KT][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < KT; ++j)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[i][j][c] = 0.f;
        __syncthreads();
        dx_stage(stage, W, N, K, Kp, 0, NC, vecW, &bars[0]);
        for (int c = 0; c < nchunks; ++c) {
            mbar_wait(&bars[c & 1], (ph >> (c & 1)) & 1u);
            ph ^= 1u << (c & 1);
            __syncthreads();
            if (c + 1 < nchunks) dx_stage(stage + ((c + 1) & 1) * SF, W, N, K, Kp, (c + 1) * NC, NC, vecW, &bars[(c + 1) & 1]);
            const int n0 = c * NC;
            const int nc = min(NC, Np - n0);
            const float* sw = stage + (c & 1) * SF;
            if (qb < KQ) {
                for (int nn = 0; nn < nc; nn += 4) {
                    float4 a[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        a[i] = *reinterpret_cast<const float4*>(dOut + (mi + 8 * i) * ldd + n0 + nn);
#pragma unroll
                    for (int j = 0; j < KT; ++j) {
                        const int kq = qb + 4 * j;
                        if (kq < KQ) {
                            const float* wp = sw + nn * Kp + 4 * kq;
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
#pragma unroll
        for (int j = 0; j < KT; ++j) {
            const int kq = qb + 4 * j;
            if (kq < KQ) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    float* ap = act + (mi + 8 * i) * lda + 4 * kq;
                    float4 h = *reinterpret_cast<float4*>(ap);
                    h.x = acc[i][j][0] * (1.f - h.x * h.x);
                    h.y = acc[i][j][1] * (1.f - h.y * h.y);
                    h.z = acc[i][j][2] * (1.f - h.z * h.z);
                    h.w = acc[i][j][3] * (1.f - h.w * h.w);
                    // columns k >= K see W = 0 (zero-filled stage), so they stay finite
                    *reinterpret_cast<float4*>(ap) = h;
                }
            }
        }
    }
    __syncthreads();
    return ph;
}

__device__ __forceinline__ unsigned dx_layer(const float* __restrict__ W, int N, int K, const float* dOut, int ldd,
                                         float* act, int lda, float* stage, int SF, uint64_t* bars, unsigned ph,
                                         const CUtensorMap* xmap, uint32_t tmem) {
    if (xmap != nullptr && N >= 32 && K >= 64 && (K & 3) == 0) {             // = ts_eligible() on the host
        TsX x = {};
        return ts_layer<true>(xmap, nullptr, N, K, dOut, ldd, act, lda, reinterpret_cast<unsigned char*>(stage), bars, ph, tmem, x);
    }
    const int KQ = (K + 3) >> 2;
    if (KQ > UPD_THREADS / 4) return dx_layer_t<4>(W, N, K, dOut, ldd, act, lda, stage, SF, bars, ph);
    if (KQ > UPD_THREADS / 8) return dx_layer_t<2>(W, N, K, dOut, ldd, act, lda, stage, SF, bars, ph);
    return dx_layer_t<1>(W, N, K, dOut, ldd, act, lda, stage, SF, bars, ph);
}

// dX through a tiny head (N small), weights in shared memory [N][ldw] (zero beyond K):
//   act[m][k] <- (sum_n dOut[m][n] W[n][k]) (1 - act[m][k]^2), four k per thread
__device__ __noinline__ void head_dx(const float* Ws, int ldw, int N, int K, const float* dOut, int ldd, float* act,
                                     int lda) {
    const int KQ = (K + 3) >> 2;
    for (int e = threadIdx.x; e < UPD_ROWS * KQ; e += UPD_THREADS) {
        const int m = e / KQ, kq = e - m * KQ;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int n = 0; n < N; ++n) {
            const float dd = dOut[m * ldd + n];
            const float4 w = *reinterpret_cast<const float4*>(Ws + n * ldw + 4 * kq);
            s.x = fmaf(dd, w.x, s.x); s.y = fmaf(dd, w.y, s.y); s.z = fmaf(dd, w.z, s.z); s.w = fmaf(dd, w.w, s.w);
        }
        float4 h = *reinterpret_cast<float4*>(act + m * lda + 4 * kq);
        h.x = s.x * (1.f - h.x * h.x); h.y = s.y * (1.f - h.y * h.y);
        h.z = s.z * (1.f - h.z * h.z); h.w = s.w * (1.f - h.w * h.w);
        *reinterpret_cast<float4*>(act + m * lda + 4 * kq) = h;
    }
    __syncthreads();
}

// both heads in ONE pass (every separately compiled phase costs ~2 k cycles of cold instruction fetch whatever its work):
// items [0, 32*KQa) are the mu head's (m, k-quad) pairs, the rest the value head's; same arithmetic as head_dx
__device__ __noinline__ void head_dx2(const float* Wa, int ldwa, int Na, int Ka, const float* dOa, int ldda, float* acta, int ldaa,
                                      const float* Wb, int Kb, const float* dOb, int lddb, float* actb, int ldab) {
    const int KQa = (Ka + 3) >> 2, KQb = (Kb + 3) >> 2;
    for (int e = threadIdx.x; e < UPD_ROWS * (KQa + KQb); e += UPD_THREADS) {
        const bool second = e >= UPD_ROWS * KQa;
        const int e2 = second ? e - UPD_ROWS * KQa : e, KQ = second ? KQb : KQa;
        const int m = e2 / KQ, kq = e2 - m * KQ;
        const int N = second ? 1 : Na, ldw = second ? 0 : ldwa, ldd = second ? lddb : ldda;
        const float* Ws = second ? Wb : Wa;
        const float* dOut = second ? dOb : dOa;
        float* act = (second ? actb + m * ldab : acta + m * ldaa) + 4 * kq;
        float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int n = 0; n < N; ++n) {
            const float dd = dOut[m * ldd + n];
            const float4 w = *reinterpret_cast<const float4*>(Ws + n * ldw + 4 * kq);
            s.x = fmaf(dd, w.x, s.x); s.y = fmaf(dd, w.y, s.y); s.z = fmaf(dd, w.z, s.z); s.w = fmaf(dd, w.w, s.w);
        }
        float4 h = *reinterpret_cast<float4*>(act);
        h.x = s.x * (1.f - h.x * h.x); h.y = s.y * (1.f - h.y * h.y);
        h.z = s.z * (1.f - h.z * h.z); h.w = s.w * (1.f - h.w * h.w);
        *reinterpret_cast<float4*>(act) = h;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// dW[n][k] (+)= sum_m dOut[m][n] in[m][k];  db[n] (+)= sum_m dOut[m][n]   (m < 32, ascending:
// fixed order) written to this CTA's partial bucket (global).
// Warp tile 16 n x 64 k; lanes are 4 (n-quads) x 8 (k-quads), a thread owns one n-quad and two
// k-quads: 3 shared loads (one wavefront each) per 32 FFMAs.
// ---------------------------------------------------------------------------
__device__ __noinline__ void dw_layer(float* __restrict__ gW, float* __restrict__ gb, int N, int K, const float* dOut, int ldd,
                         const float* in, int ldin, bool accumulate) {
    constexpr int NWARPS = UPD_THREADS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nql = lane >> 3, kql = lane & 7;
    const int NQ = (N + 3) >> 2, KQ = (K + 3) >> 2;
    const int NB = (NQ + 3) >> 2, KB = (KQ + 15) >> 4;
    const bool vecG = ((K & 3) == 0) && aligned16_dev(gW);
    for (int t = warp; t < NB * KB; t += NWARPS) {
        const int nb = t / KB, kb = t - nb * KB;
        const int nq = nb * 4 + nql;
        const int kq0 = kb * 16 + kql, kq1 = kq0 + 8;
        // clamp the read positions of out-of-range tiles (their results are dropped)
        const float* dp = dOut + 4 * min(nq, NQ - 1);
        const float* ip0 = in + 4 * min(kq0, KQ - 1);
        const float* ip1 = in + 4 * min(kq1, KQ - 1);
        float acc[4][8];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll 4
        for (int m = 0; m < 32; ++m) {
            const float4 a = *reinterpret_cast<const float4*>(dp + m * ldd);
            const float4 b0 = *reinterpret_cast<const float4*>(ip0 + m * ldin);
            const float4 b1 = *reinterpret_cast<const float4*>(ip1 + m * ldin);
            const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                acc[i][0] = fmaf(av[i], b0.x, acc[i][0]); acc[i][1] = fmaf(av[i], b0.y, acc[i][1]);
                acc[i][2] = fmaf(av[i], b0.z, acc[i][2]); acc[i][3] = fmaf(av[i], b0.w, acc[i][3]);
                acc[i][4] = fmaf(av[i], b1.x, acc[i][4]); acc[i][5] = fmaf(av[i], b1.y, acc[i][5]);
                acc[i][6] = fmaf(av[i], b1.z, acc[i][6]); acc[i][7] = fmaf(av[i], b1.w, acc[i][7]);
            }
        }
        if (nq < NQ) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int kq = h ? kq1 : kq0;
                if (kq >= KQ) continue;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int n = 4 * nq + i;
                    if (n < N) {
                        float* dst = gW + (size_t)n * K + 4 * kq;
                        if (vecG) {
                            float4 o = make_float4(acc[i][4 * h], acc[i][4 * h + 1], acc[i][4 * h + 2], acc[i][4 * h + 3]);
                            if (accumulate) {
                                const float4 p = *reinterpret_cast<const float4*>(dst);
                                o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w;
                            }
                            *reinterpret_cast<float4*>(dst) = o;
                        } else {
#pragma unroll
                            for (int j = 0; j < 4; ++j)
                                if (4 * kq + j < K) dst[j] = accumulate ? dst[j] + acc[i][4 * h + j] : acc[i][4 * h + j];
                        }
                    }
                }
            }
        }
    }
    for (int n = threadIdx.x; n < N; n += UPD_THREADS) {
        float s = 0.f;
        for (int m = 0; m < 32; ++m) s += dOut[m * ldd + n];
        gb[n] = accumulate ? gb[n] + s : s;
    }
    __syncthreads();
}

// dW for a narrow input (K <= 24, first layers): lanes are 16 (n-quads) x 2 (k-groups), a thread owns
// one n-quad and the k-quads g*KH .. g*KH+KH-1 (KH = ceil(KQ/2) <= 3), so nothing is padded to 64 columns.
__device__ __noinline__ void dw_layer_smallk(float* __restrict__ gW, float* __restrict__ gb, int N, int K,
                                             const float* dOut, int ldd, const float* in, int ldin, bool accumulate,
                                             float* img) {
    constexpr int NWARPS = UPD_THREADS / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nql = lane >> 1, g = lane & 1;
    const int NQ = (N + 3) >> 2, KQ = (K + 3) >> 2;
    const int KH = (KQ + 1) >> 1;
    for (int nq0 = warp * 16; nq0 < NQ; nq0 += NWARPS * 16) {
        const int nq = nq0 + nql;
        const float* dp = dOut + 4 * min(nq, NQ - 1);
        float acc[4][12];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 12; ++j) acc[i][j] = 0.f;
#pragma unroll 2
        for (int m = 0; m < 32; ++m) {
            const float4 a = *reinterpret_cast<const float4*>(dp + m * ldd);
            const float av[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
            for (int h = 0; h < 3; ++h) {
                if (h < KH) {
                    const int kq = min(g * KH + h, KQ - 1);
                    const float4 b = *reinterpret_cast<const float4*>(in + m * ldin + 4 * kq);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        acc[i][4 * h + 0] = fmaf(av[i], b.x, acc[i][4 * h + 0]);
                        acc[i][4 * h + 1] = fmaf(av[i], b.y, acc[i][4 * h + 1]);
                        acc[i][4 * h + 2] = fmaf(av[i], b.z, acc[i][4 * h + 2]);
                        acc[i][4 * h + 3] = fmaf(av[i], b.w, acc[i][4 * h + 3]);
                    }
                }
            }
        }
        // the tile goes to a flat [N][K] image in shared memory first: the tensor is contiguous in
        // the bucket, so the flush below is fully coalesced (rows of 17 floats are not, one by one)
        if (nq < NQ) {
#pragma unroll
            for (int h = 0; h < 3; ++h) {
                const int kq = g * KH + h;
                if (h < KH && kq < KQ) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int n = 4 * nq + i;
                        if (n < N) {
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int k = 4 * kq + j;
                                if (k < K) img[n * K + k] = acc[i][4 * h + j];
                            }
                        }
                    }
                }
            }
        }
    }
    __syncthreads();
    const int tot = N * K;
    if ((tot & 3) == 0) {
        float4* g4 = reinterpret_cast<float4*>(gW);
        const float4* s4 = reinterpret_cast<const float4*>(img);
        for (int e = threadIdx.x; e < (tot >> 2); e += UPD_THREADS) {
            float4 o = s4[e];
            if (accumulate) { const float4 p = g4[e]; o.x += p.x; o.y += p.y; o.z += p.z; o.w += p.w; }
            g4[e] = o;
        }
    } else {
        for (int e = threadIdx.x; e < tot; e += UPD_THREADS) gW[e] = accumulate ? gW[e] + img[e] : img[e];
    }
    for (int n = threadIdx.x; n < N; n += UPD_THREADS) {
        float s = 0.f;
        for (int m = 0; m < 32; ++m) s += dOut[m * ldd + n];
        gb[n] = accumulate ? gb[n] + s : s;
    }
    __syncthreads();
}


// dW of a wide layer on the tensor cores, operands through tensor memory:
//   dW[n][k] (+)= sum_s dOut[s][n] in[s][k]      D[k][n] = A[k][s] . B[n][s]^T  per tile of 128 k and chunk of <= 128 n
//   A = in^T: TMEM lane = k, columns = the 32 samples (read down a column of the activations, conflict-free)
//   B = dOut^T as [hi rows 0..Np-1 | lo rows Np..2Np-1] K-major tiles (Np = chunk width rounded up to 16), staged per chunk
//   MMAs per k-step: D[:, 0..2Np) += A_hi x [B_hi | B_lo]  (N = 2 Np),  D[:, Np..2Np) += A_lo x B_hi  (N = Np)
// Only 4 k-steps per tile, so one accumulator is accurate.  The epilogue has lane = k: for a fixed n a warp writes
// 128 contiguous bytes of the partial.  Tiles are processed one after the other (transform, MMA, epilogue); the
// epilogue - 2 Np columns of tensor memory and Np x 128 floats to global memory - dominates.  Layers with more than
// 128 outputs (256-wide nets) run chunk by chunk; the activation tile is transformed again for every chunk.
__device__ __noinline__ unsigned dw_layer_ts(float* __restrict__ gW, float* __restrict__ gb, int N, int K, const float* dOut,
                                             int ldd, const float* in, int ldin, bool accumulate, unsigned char* ts,
                                             uint64_t* bars, unsigned ph, uint32_t tmem) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ntiles = (K + 127) >> 7;
    uint64_t* mma_done = bars + TS_BAR0 + 8;
    K2_ACC_DECL;
    K2_ACC_T0;
    constexpr uint64_t KSTEP = (2 * TC_LBO) >> 4;
    for (int nbase = 0; nbase < N; nbase += 128) {
        const int Nc = min(128, N - nbase);                 // this chunk's outputs
        const int Np = (Nc + 15) & ~15;
        __syncthreads();                                    // staging region free, dOut / in complete
        // ---- B: dOut^T of the chunk, split ----
        for (int e = tid; e < Np * 8; e += UPD_THREADS) {
            const int n = e % Np, kc = e / Np;              // consecutive threads -> consecutive n
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (n < Nc) {
                const float* p = dOut + (4 * kc) * ldd + nbase + n;
                x = make_float4(p[0], p[ldd], p[2 * ldd], p[3 * ldd]);
            }
            uint32_t hh[4], ll[4];
            ts_split(x.x, hh[0], ll[0]); ts_split(x.y, hh[1], ll[1]); ts_split(x.z, hh[2], ll[2]); ts_split(x.w, hh[3], ll[3]);
            *reinterpret_cast<uint4*>(ts + tc_off(n, kc)) = make_uint4(hh[0], hh[1], hh[2], hh[3]);
            *reinterpret_cast<uint4*>(ts + tc_off(Np + n, kc)) = make_uint4(ll[0], ll[1], ll[2], ll[3]);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        K2_ACC(40);
        const uint32_t idm = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)((2 * Np) >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t idc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(Np >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint64_t bd = tc_desc(smem_u32(ts));
        for (int t = 0; t < ntiles; ++t) {
            if (warp < 4) {                                 // ---- A: 32 samples of my activation column ----
                const int k = 128 * t + 32 * warp + lane;
                uint32_t hi[32], lo[32];
                const float* cp = in + min(k, K - 1);
#pragma unroll
                for (int sidx = 0; sidx < 32; ++sidx) ts_split(k < K ? cp[sidx * ldin] : 0.f, hi[sidx], lo[sidx]);
                const uint32_t ta = tmem + ((uint32_t)(32 * warp) << 16) + TS_TM_A;
                tmem_st32(ta, hi);
                tmem_st32(ta + 32, lo);
                asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
            }
            K2_ACC(41);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();
            K2_ACC(42);
            if (warp == 7) {
                const uint32_t el = elect_one();
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) ts_mma_e(tmem, tmem + TS_TM_A + 8 * ks, bd + ks * KSTEP, idm, ks ? 1u : 0u, el);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) ts_mma_e(tmem + Np, tmem + TS_TM_A + 32 + 8 * ks, bd + ks * KSTEP, idc, 1u, el);
                ts_commit(&mma_done[0]);
            }
            mbar_wait(&mma_done[0], (ph >> TS_PH_MD) & 1u);     // every thread sees every phase of this barrier
            ph ^= 1u << TS_PH_MD;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            K2_ACC(43);
            {
                const int q = warp & 3, k = 128 * t + 32 * q + lane;
                const int nh = Np >> 1, n0 = (warp >> 2) * nh;  // this warp's half of the chunk's columns
                const uint32_t tb = tmem + ((uint32_t)(32 * q) << 16);
                for (int c = 0; c < nh; c += 32) {              // up to 32 columns of both halves in flight, one wait
                    uint32_t v1[32], v2[32];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (c + 8 * i < nh) {
                            tmem_ld_issue<8>(tb + n0 + c + 8 * i, v1 + 8 * i);
                            tmem_ld_issue<8>(tb + Np + n0 + c + 8 * i, v2 + 8 * i);
                        }
                    tmem_ld_wait();
                    // straight-line stores: one running pointer, the column limit as a predicate (a branch per element
                    // costs more than the store here)
                    const int lim = min(nh - c, Nc - n0 - c);
                    if (k < K && lim > 0) {
                        float* dst = gW + (size_t)(nbase + n0 + c) * K + k;
                        if (!accumulate) {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const float o = __uint_as_float(v2[j]) + __uint_as_float(v1[j]);
                                if (j < lim) *dst = o;
                                dst += K;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < 32; ++j) {
                                const float o = __uint_as_float(v2[j]) + __uint_as_float(v1[j]);
                                if (j < lim) *dst += o;
                                dst += K;
                            }
                        }
                    }
                }
            }
            K2_ACC(44);
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncthreads();                                    // accumulator and A columns free for the next tile
            K2_ACC(45);
        }
    }
    for (int n = tid; n < N; n += UPD_THREADS) {
        float sb = 0.f;
        for (int m = 0; m < 32; ++m) sb += dOut[m * ldd + n];
        gb[n] = accumulate ? gb[n] + sb : sb;
    }
    K2_ACC(46);
    K2_ACC_FLUSH;
    __syncthreads();
    return ph;
}

// dW / db of both heads in one pass, one thread per output element (A*H2 + G2 weights, A + 1 biases):
//   dWmu[n][k] = sum_s dmu[s][n] h2[s][k],   dWv[k] = sum_s dv[s] g2[s][k]    (s ascending: fixed order)
__device__ __noinline__ void heads_dw(float* __restrict__ gWmu, float* __restrict__ gbmu, float* __restrict__ gWv,
                                      float* __restrict__ gbv, int A, int H2, int G2, const float* dmu, int ldA,
                                      const float* dv4, const float* h2, int ldh2, const float* g2, int ldg2,
                                      bool accumulate, int mode = 0) {
    const int nmu = A * H2;
    for (int o = threadIdx.x; o < nmu + G2 + A + 1; o += UPD_THREADS) {
        const bool policy = o < nmu || (o >= nmu + G2 && o < nmu + G2 + A);
        if ((mode == 1 && !policy) || (mode == 2 && policy)) continue;   // one trunk at a time (trunk-serial tiles)
        float s = 0.f;
        float* dst;
        if (o < nmu) {
            const int n = o / H2, k = o - n * H2;
#pragma unroll 8
            for (int m = 0; m < UPD_ROWS; ++m) s = fmaf(dmu[m * ldA + n], h2[m * ldh2 + k], s);
            dst = gWmu + o;
        } else if (o < nmu + G2) {
            const int k = o - nmu;
#pragma unroll 8
            for (int m = 0; m < UPD_ROWS; ++m) s = fmaf(dv4[4 * m], g2[m * ldg2 + k], s);
            dst = gWv + k;
        } else if (o < nmu + G2 + A) {
            const int n = o - nmu - G2;
            for (int m = 0; m < UPD_ROWS; ++m) s += dmu[m * ldA + n];
            dst = gbmu + n;
        } else {
            for (int m = 0; m < UPD_ROWS; ++m) s += dv4[4 * m];
            dst = gbv;
        }
        *dst = accumulate ? *dst + s : s;
    }
    __syncthreads();
}

// ---------------------------------------------------------------------------
// whole forward of one network on the tile (model.py:35-45)
// ---------------------------------------------------------------------------
// `next`: the network whose first layers are prefetched into the staging region while this one's heads run
// features [0, nf) of a sample-major shared-memory array (32 rows of ld floats, ld a multiple of 4) -> feature-major
// rows of Z, this tile's 32 columns.  A warp takes four features at a time: lane = sample, one 128-bit load along the
// features, four stores of a whole 128-byte row each.
__device__ __forceinline__ void z_dump(float* zrow0, int zld, const float* src, int ld, int nf) {
    const int warp = threadIdx.x >> 5, m = threadIdx.x & 31;
    for (int f = 4 * warp; f < nf; f += 4 * (UPD_THREADS / 32)) {
        const float4 v = *reinterpret_cast<const float4*>(src + m * ld + f);
        float* z = zrow0 + (size_t)f * zld + m;
        z[0] = v.x;
        if (f + 1 < nf) z[zld] = v.y;
        if (f + 2 < nf) z[2 * (size_t)zld] = v.z;
        if (f + 3 < nf) z[3 * (size_t)zld] = v.w;
    }
}

__device__ __forceinline__ unsigned net_forward(const Net& nw, const HeadS& hs, const UpdArgs& P, float* S, float* mu_out,
                                            float* v_out, int tick, unsigned ph, uint32_t tmem, const CUtensorMap* maps,
                                            const Net* next, int zcol = -1) {
    const dppo_dims& d = P.d;
    float* stage = S + P.o_stage;
    uint64_t* bars = reinterpret_cast<uint64_t*>(S + P.o_bars);
    // batch-level dW (zcol >= 0: this tile's columns of Z): every hidden activation is copied out feature-major right
    // after its layer, so that the stores drain while the next layer computes instead of in one burst
    auto zd = [&](int zr, const float* src, int ld, int nf) {
        if (zcol >= 0) z_dump(P.Z + (size_t)zr * P.Bp + zcol, P.Bp, src, ld, nf);
    };
    if (P.small_first) {
        mbar_wait(&bars[0], ph & 1u);                       // this network's W1 and V1 have landed (prefetched)
        ph ^= 1u;
        if (P.first_ts && d.H1 <= 512)
            ph = fwd_first_ts(stage, nw.b1, d.H1, d.O, S + P.o_x, P.ldO, S + P.o_h1, P.ldH1, first_b_tiles(stage, d), bars, ph, tmem);
        else if (d.H1 <= 128) fwd_small_k<4>(stage, nw.b1, d.H1, d.O, S + P.o_x, P.ldO, S + P.o_h1, P.ldH1);
        else fwd_small_k<8>(stage, nw.b1, d.H1, d.O, S + P.o_x, P.ldO, S + P.o_h1, P.ldH1);
        K2_TICK(tick + 0);
        if (P.first_ts && d.G1 <= 512)
            ph = fwd_first_ts(stage + first_v1_offset(d), nw.c1, d.G1, d.O, S + P.o_x, P.ldO, S + P.o_g1, P.ldG1, first_b_tiles(stage, d), bars, ph, tmem);
        else if (d.G1 <= 128) fwd_small_k<4>(stage + first_v1_offset(d), nw.c1, d.G1, d.O, S + P.o_x, P.ldO, S + P.o_g1, P.ldG1);
        else fwd_small_k<8>(stage + first_v1_offset(d), nw.c1, d.G1, d.O, S + P.o_x, P.ldO, S + P.o_g1, P.ldG1);
        K2_TICK(tick + 1);
    } else {
        ph = fwd_layer<true>(nw.W1, nw.b1, d.H1, d.O, S + P.o_x, P.ldO, S + P.o_h1, P.ldH1, stage, P.SF, bars, ph, tmem,
                             maps ? maps + 2 : nullptr);
        K2_TICK(tick + 0);
        ph = fwd_layer<true>(nw.V1, nw.c1, d.G1, d.O, S + P.o_x, P.ldO, S + P.o_g1, P.ldG1, stage, P.SF, bars, ph, tmem,
                             maps ? maps + 3 : nullptr);
        K2_TICK(tick + 1);
    }
    zd(P.zr_h1, S + P.o_h1, P.ldH1, d.H1);
    zd(P.zr_g1, S + P.o_g1, P.ldG1, d.G1);
    ph = fwd_layer<true>(nw.W2, nw.b2, d.H2, d.H1, S + P.o_h1, P.ldH1, S + P.o_h2, P.ldH2, stage, P.SF, bars, ph, tmem,
                         maps ? maps + 0 : nullptr);
    K2_TICK(tick + 2);
    zd(P.zr_h2, S + P.o_h2, P.ldH2, d.H2);
    ph = fwd_layer<true>(nw.V2, nw.c2, d.G2, d.G1, S + P.o_g1, P.ldG1, S + P.o_g2, P.ldG2, stage, P.SF, bars, ph, tmem,
                         maps ? maps + 1 : nullptr);
    K2_TICK(tick + 3);
    zd(P.zr_g2, S + P.o_g2, P.ldG2, d.G2);
    if (P.small_first && next != nullptr) first_layers_prefetch(stage, next->W1, next->V1, d, &bars[0]);
    heads_fwd(hs, d.A, d.H2, d.G2, S + P.o_h2, P.ldH2, S + P.o_g2, P.ldG2, mu_out, P.ldA, v_out);
    K2_TICK(tick + 4);
    return ph;
}

// One trunk of one network on the tile, in the shared activation buffers a1 / a2 (trunk-serial tiles: wide nets whose
// two trunks do not fit shared memory side by side).  trunk 0: policy (p_fc1, p_fc2, mu), trunk 1: value (v_fc1, v_fc2, v).
__device__ __forceinline__ unsigned trunk_forward(const Net& nw, const HeadS& hs, const UpdArgs& P, float* S, int trunk,
                                              float* mu_out, float* v_out, unsigned ph, uint32_t tmem, const CUtensorMap* maps) {
    const dppo_dims& d = P.d;
    float* stage = S + P.o_stage;
    uint64_t* bars = reinterpret_cast<uint64_t*>(S + P.o_bars);
    const int N1 = trunk ? d.G1 : d.H1, N2 = trunk ? d.G2 : d.H2, ld1 = trunk ? P.ldG1 : P.ldH1, ld2 = trunk ? P.ldG2 : P.ldH2;
    float* a1 = S + P.o_h1;
    float* a2 = S + P.o_h2;
    ph = fwd_layer<true>(trunk ? nw.V1 : nw.W1, trunk ? nw.c1 : nw.b1, N1, d.O, S + P.o_x, P.ldO, a1, ld1, stage, P.SF, bars, ph,
                         tmem, maps ? maps + 2 + trunk : nullptr);
    ph = fwd_layer<true>(trunk ? nw.V2 : nw.W2, trunk ? nw.c2 : nw.b2, N2, N1, a1, ld1, a2, ld2, stage, P.SF, bars, ph, tmem,
                         maps ? maps + trunk : nullptr);
    heads_fwd(hs, d.A, d.H2, d.G2, a2, P.ldH2, a2, P.ldG2, mu_out, P.ldA, v_out, 1 + trunk);
    return ph;
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

// three deterministic block sums (the same fixed tree as block_sum) sharing their barriers; r1 / r2: two more scratch
// arrays of UPD_THREADS floats (the mu / mu_old area of the tile, dead by then)
__device__ void block_sum3(float a, float b, float c, float* r0, float* r12, float& oa, float& ob, float& oc) {
    float* r1 = r12;
    float* r2 = r12 + UPD_THREADS;
    __syncthreads();
    r0[threadIdx.x] = a; r1[threadIdx.x] = b; r2[threadIdx.x] = c;
    __syncthreads();
    for (int s = UPD_THREADS / 2; s > 0; s >>= 1) {
        if (threadIdx.x < s) {
            r0[threadIdx.x] += r0[threadIdx.x + s];
            r1[threadIdx.x] += r1[threadIdx.x + s];
            r2[threadIdx.x] += r2[threadIdx.x + s];
        }
        __syncthreads();
    }
    oa = r0[0]; ob = r1[0]; oc = r2[0];
    __syncthreads();
}

__global__ void __launch_bounds__(UPD_THREADS, 1) update_fused_kernel(const __grid_constant__ UpdArgs P) {
    extern __shared__ __align__(1024) float S[];
    const dppo_dims& d = P.d;
    const int TM = P.TM;
    const Net nw = net_of(P.theta, P.lay);
    float* part = P.partials + (size_t)blockIdx.x * P.lay.P_pad;
    const GradOut go = grad_of(part, P.lay);
    float* red = S + P.o_red;
    // the two stage-buffer mbarriers are initialised ONCE; `ph` carries their phase parities
    if (threadIdx.x == 0) {
        uint64_t* b = reinterpret_cast<uint64_t*>(S + P.o_bars);
        mbar_init(&b[0], 1);                               // weight stage buffers (TMA bulk copies)
        mbar_init(&b[1], 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&b[TS_BAR0 + i], 1);                  // raw_full (TMA tx)
            mbar_init(&b[TS_BAR0 + 2 + i], 4);              // raw_free: the four transform warps
            mbar_init(&b[TS_BAR0 + 4 + i], 6);              // a_full: four transform warps + the two sample warps
            mbar_init(&b[TS_BAR0 + 8 + i], 1);              // mma_done (tcgen05.commit)
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // tensor memory: all 512 columns (one CTA per SM), 9 accumulators of 32 columns are used
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(S + P.o_bars + 2 * (TS_BAR0 + TS_NBARS));
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = *tmem_slot;
    unsigned ph = 0;
    if (P.dwg == 2) {
        // ---- batch-level dW pass: dW[n][kf] = sum over this CTA's range of samples of dz[s][n] * in[s][kf] ----------
        // One wide-layer call per CTA: "weights" = dz^T (rows n, TMA from Z), 32-row operand = the 32 input features
        // kf0.. of in^T (rows of Z, read by the sample warps), reduction over the samples; the ones row behind every
        // input operand makes column Kf the bias gradient.  The split's partial bucket gets every element once.
        const int split = blockIdx.x / P.dwg_units;
        int u = blockIdx.x - split * P.dwg_units;
        const int s0 = split * P.dwg_chunk, ns = min(P.dwg_chunk, P.Bp - s0);
        float* bucket = P.partials + (size_t)split * P.lay.P_pad;
        const GradOut gs = grad_of(bucket, P.lay);
        if (u == 0) {
            if (threadIdx.x < DPPO_N_PARAMS)                           // padding floats of the bucket stay 0
                for (int j = P.lay.size[threadIdx.x]; j < ((P.lay.size[threadIdx.x] + 3) & ~3); ++j)
                    bucket[P.lay.off[threadIdx.x] + j] = 0.f;
            if (threadIdx.x < 4) {                                    // the update kernel's loss sums, rows in order
                const int nsp = gridDim.x / P.dwg_units;
                const int r0 = (int)((long long)P.zloss_rows * split / nsp), r1 = (int)((long long)P.zloss_rows * (split + 1) / nsp);
                float a = 0.f;
                for (int r = r0; r < r1; ++r) a += P.zloss[4 * r + threadIdx.x];
                P.loss_partials[4 * split + threadIdx.x] = a;
            }
        }
        for (int p = 0; p < 6 && ns > 0; ++p) {
            // hidden layers (p < 4): tile rows = dz rows (outputs n), 32-row operand = input features.
            // heads (p >= 4): swapped - tile rows = input features (+ the ones row), 32-row operand = the few dz rows.
            int rows_t, rows_b, Kf, n1, im;                           // tile-operand rows, 32-row-operand rows
            float *gW, *gb, *gb2 = nullptr;
            if (p == 0) { rows_t = d.G2; Kf = d.G1; rows_b = Kf + 1; im = 2; gW = gs.V2; gb = gs.c2; n1 = rows_t; }
            else if (p == 1) { rows_t = d.H2; Kf = d.H1; rows_b = Kf + 1; im = 1; gW = gs.W2; gb = gs.b2; n1 = rows_t; }
            else if (p == 2) { rows_t = d.G1; Kf = d.O; rows_b = Kf + 1; im = 0; gW = gs.V1; gb = gs.c1; n1 = rows_t; }
            else if (p == 3) { rows_t = d.H1; Kf = d.O; rows_b = Kf + 1; im = 0; gW = gs.W1; gb = gs.b1; n1 = rows_t; }
            else if (p == 4) { Kf = d.H2; rows_t = Kf + 1; rows_b = 2 * d.A; im = 3; gW = gs.Wmu; gb = gs.bmu; gb2 = gs.ls; n1 = d.A; }
            else { Kf = d.G2; rows_t = Kf + 1; rows_b = 1; im = 4; gW = gs.Wv; gb = gs.bv; n1 = 1; }
            const int kfb = (rows_b + 31) >> 5, nt = (rows_t + 127) >> 7;
            if (u >= kfb * nt) { u -= kfb * nt; continue; }
            const int t = u / kfb, b = u - t * kfb;
            TsX x = {};
            x.dwg = 1; x.tr = p >= 4 ? 1 : 0; x.kofs = s0; x.row0 = 128 * t;
            x.gW = gW; x.gb = gb; x.gb2 = gb2; x.ldw = Kf; x.kf0 = 32 * b; x.Kf = Kf; x.n1 = n1;
            x.add2 = (P.flavour == DPPO_FLAVOUR_PPO && split == 0) ? -P.ent_coeff : 0.f;     // ppo.py:152-153,167
            x.bmap = &P.map_in[im]; x.brow0 = 32 * b; x.braw = S;                           // (the activation area is free in this pass)
            ph = ts_layer<false>(&P.map_dz[p], nullptr, min(128, rows_t - 128 * t), ns, nullptr, 0, nullptr, 0,
                                 reinterpret_cast<unsigned char*>(S + P.o_stage), reinterpret_cast<uint64_t*>(S + P.o_bars), ph, tmem, x);
            break;
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        if (threadIdx.x < 32)
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
        return;
    }
    // head weights of both networks stay in shared memory for the whole kernel
    // (when shared memory is tight - Humanoid with 256-wide nets - they are read from global memory instead)
    if (P.heads_in_smem) {
        for (int i = threadIdx.x; i < 2 * P.head_floats; i += UPD_THREADS) S[P.o_heads + i] = 0.f;
        __syncthreads();
        load_heads(S + P.o_heads, nw, P);
        if (P.flavour == DPPO_FLAVOUR_DPPO) load_heads(S + P.o_heads + P.head_floats, net_of(P.theta_old, P.lay), P);
    }
    K2_TICK(0);
    HeadS hs_new = heads_at(S + P.o_heads, P), hs_old = heads_at(S + P.o_heads + P.head_floats, P);
    if (!P.heads_in_smem) {
        hs_new.Wmu = nw.Wmu; hs_new.bmu = nw.bmu; hs_new.Wv = nw.Wv; hs_new.bv = nw.bv; hs_new.ldw = d.H2;
        if (P.flavour == DPPO_FLAVOUR_DPPO) {
            const Net no = net_of(P.theta_old, P.lay);
            hs_old.Wmu = no.Wmu; hs_old.bmu = no.bmu; hs_old.Wv = no.Wv; hs_old.bv = no.bv; hs_old.ldw = d.H2;
        }
    }
    if (threadIdx.x < DPPO_N_PARAMS && !P.fwd_only && !P.dwg)          // padding floats of the bucket stay 0
        for (int j = P.lay.size[threadIdx.x]; j < ((P.lay.size[threadIdx.x] + 3) & ~3); ++j)
            part[P.lay.off[threadIdx.x] + j] = 0.f;
    float lsum_clip = 0.f, lsum_val = 0.f, lsum_ent = 0.f;           // this thread's share of the loss sums
    bool first = true;

    for (int tile = blockIdx.x; tile < P.ntiles; tile += gridDim.x, first = false) {
        K2_TICK(1);
        const int b0 = tile * TM;
        const int valid = min(TM, P.B - b0);
        const bool same_old = P.flavour == DPPO_FLAVOUR_DPPO && P.theta_old == P.theta;
        const bool run_old = P.flavour == DPPO_FLAVOUR_DPPO && !same_old;
        if (P.small_first) {                                 // first layers of the first network to run: in flight during the gather
            const Net nf = net_of(run_old ? P.theta_old : P.theta, P.lay);
            first_layers_prefetch(S + P.o_stage, nf.W1, nf.V1, d, reinterpret_cast<uint64_t*>(S + P.o_bars));
        }
        // ---- gather + zero the activation area (pads must read as 0) ---------------------------
        // The gathered words are fetched into registers first (two dependent global loads each: position, then value), the
        // activation area is zeroed with 128-bit stores while they are in flight, then the words are dropped in.
        constexpr int GQ = 4;                                 // gathered obs words per thread and round
        const int nobs = valid * d.O;
        float gx[GQ];
        int ge[GQ];
#pragma unroll
        for (int q = 0; q < GQ; ++q) {
            const int e = threadIdx.x + q * UPD_THREADS;
            ge[q] = e;
            gx[q] = 0.f;
            if (e < nobs) {
                const int m = e / d.O, k = e - m * d.O;
                const size_t r = P.idx ? (size_t)__ldg(P.idx + b0 + m) : (size_t)(b0 + m);
                gx[q] = __ldg(P.obs + r * d.O + k);
            }
        }
        {
            float4* z4 = reinterpret_cast<float4*>(S);
            const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int i = threadIdx.x; i < (P.o_zero_end >> 2); i += UPD_THREADS) z4[i] = zero;   // o_zero_end is a multiple of 4
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < GQ; ++q)
            if (ge[q] < nobs) { const int m = ge[q] / d.O, k = ge[q] - m * d.O; S[P.o_x + m * P.ldO + k] = gx[q]; }
        for (int e = threadIdx.x + GQ * UPD_THREADS; e < nobs; e += UPD_THREADS) {      // wide observations: the rest
            const int m = e / d.O, k = e - m * d.O;
            const size_t r = P.idx ? (size_t)__ldg(P.idx + b0 + m) : (size_t)(b0 + m);
            S[P.o_x + m * P.ldO + k] = __ldg(P.obs + r * d.O + k);
        }
        for (int e = threadIdx.x; e < valid * d.A && !P.fwd_only; e += UPD_THREADS) {
            const int m = e / d.A, a = e - m * d.A;
            S[P.o_act + m * d.A + a] = __ldg(P.act + (size_t)__ldg(P.idx + b0 + m) * d.A + a);
        }
        for (int m = threadIdx.x; m < valid && !P.fwd_only; m += UPD_THREADS) {
            const size_t r = (size_t)__ldg(P.idx + b0 + m);
            S[P.o_adv + m] = __ldg(P.adv + r);
            S[P.o_ret + m] = __ldg(P.ret + r);
            if (P.flavour == DPPO_FLAVOUR_PPO) S[P.o_lpo + m] = __ldg(P.logp_old + r);
        }
        __syncthreads();
        if (P.dwg) {                                                      // batch-level dW: x^T and the ones rows of the five inputs
            z_dump(P.Z + (size_t)P.zr_x * P.Bp + tile * UPD_ROWS, P.Bp, S + P.o_x, P.ldO, d.O);
            for (int e = threadIdx.x; e < 5 * UPD_ROWS; e += UPD_THREADS) {
                const int o = e >> 5, m = e & 31;
                const int r = o == 0 ? P.zr_x + d.O : o == 1 ? P.zr_h1 + d.H1 : o == 2 ? P.zr_g1 + d.G1 : o == 3 ? P.zr_h2 + d.H2 : P.zr_g2 + d.G2;
                P.Z[(size_t)r * P.Bp + tile * UPD_ROWS + m] = 1.f;
            }
        }

        K2_TICK(2);
        const float lo = 1.f - P.clip, hi = 1.f + P.clip;               // clamp bounds (train.py:151)
        const Net no = net_of(P.flavour == DPPO_FLAVOUR_DPPO ? P.theta_old : P.theta, P.lay);
        float* stage = S + P.o_stage;
        uint64_t* bars = reinterpret_cast<uint64_t*>(S + P.o_bars);
        const bool acc = !first;

        // ---- losses and their derivatives w.r.t. mu, log_std (policy part) and v (value part) --------------------
        // The two parts share nothing: the policy loss needs mu, mu_old, the actions and the advantages, the value loss
        // v, v_old and the returns.  They run back to back for ordinary tiles and one per trunk for trunk-serial tiles.
        auto loss_policy = [&]() {
            if (P.flavour == DPPO_FLAVOUR_DPPO) {
                const float g0 = -1.f / ((float)P.B * (float)d.A);          // d loss_clip / d min(.,.)
                const float ge = -P.ent_coeff / ((float)P.B * (float)d.A);
                for (int e = threadIdx.x; e < UPD_ROWS * d.A; e += UPD_THREADS) {
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
            } else {
                const float g0 = -1.f / (float)P.B;                          // ppo.py:161
                const float half_log_2pi = 0.918938533204672742f;            // 0.5*math.log(2*math.pi)
                for (int m = threadIdx.x; m < UPD_ROWS; m += UPD_THREADS) {
                    float dlogp = 0.f;
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
                    }
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
        };
        auto loss_value = [&]() {
            if (P.flavour == DPPO_FLAVOUR_DPPO) {
                const float gv = 0.5f / (float)P.B;                          // train.py:157
                for (int m = threadIdx.x; m < UPD_ROWS; m += UPD_THREADS) {
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
                const float gv = 1.f / (float)P.B;                           // ppo.py:164
                for (int m = threadIdx.x; m < UPD_ROWS; m += UPD_THREADS) {
                    float dv = 0.f;
                    if (m < valid) {
                        const float e1 = S[P.o_v + m] - S[P.o_ret + m];
                        lsum_val += e1 * e1;
                        dv = gv * 2.f * e1;
                    }
                    S[P.o_dv + 4 * m] = dv;
                }
            }
        };
        // d log_std[a] = sum over the tile's samples (ascending m)
        auto dls_reduce = [&]() {
            for (int a = threadIdx.x; a < d.A; a += UPD_THREADS) {
                float s = 0.f;
                for (int m = 0; m < UPD_ROWS; ++m) s += S[P.o_muo + m * P.ldA + a];
                if (P.flavour == DPPO_FLAVOUR_PPO && blockIdx.x == 0 && first) s -= P.ent_coeff;   // ppo.py:152-153,167
                go.ls[a] = first ? s : go.ls[a] + s;
            }
        };
        // dW / db of a layer: tensor cores when the shape allows (= ts_eligible() on the host), else the FP32 pipe
        auto dw_any = [&](float* gW, float* gb, int N, int K, const float* dOut, int ldd, const float* in, int ldin) {
            if (P.use_ts && N >= 32 && K >= 64 && (K & 3) == 0)
                ph = dw_layer_ts(gW, gb, N, K, dOut, ldd, in, ldin, acc, reinterpret_cast<unsigned char*>(stage), bars, ph, tmem);
            else
                dw_layer(gW, gb, N, K, dOut, ldd, in, ldin, acc);
        };

        // forward-only mode: mu, v, the sampled action and its log-probability of this tile's rows, then the next tile
        auto write_policy_outputs = [&]() {
            const float half_log_2pi = 0.918938533204672742f;
            for (int e = threadIdx.x; e < valid * d.A; e += UPD_THREADS) {
                const int m = e / d.A, a = e - m * d.A;
                const size_t r = (size_t)(P.f_row0 + (long long)(b0 + m) * P.f_stride);
                const float mu = S[P.o_mu + m * P.ldA + a];
                const float s = expf(__ldg(nw.ls + a));                              // sigma_sq = exp(log_std), model.py:40
                const float sd = P.std_is_variance ? sqrtf(s) : s;                   // train.py:89 / ppo.py:93
                const float act = P.eps ? mu + sd * __ldg(P.eps + (size_t)(b0 + m) * d.A + a) : mu;
                S[P.o_dmu + m * P.ldA + a] = act;
                if (P.f_mu) P.f_mu[r * d.A + a] = mu;
                if (P.f_act) P.f_act[r * d.A + a] = act;
            }
            for (int e = threadIdx.x; e < valid * d.O && P.f_obs != nullptr; e += UPD_THREADS) {
                const int m = e / d.O, k = e - m * d.O;
                P.f_obs[(size_t)(P.f_row0 + (long long)(b0 + m) * P.f_stride) * d.O + k] = S[P.o_x + m * P.ldO + k];
            }
            __syncthreads();
            for (int m = threadIdx.x; m < valid; m += UPD_THREADS) {
                const size_t r = (size_t)(P.f_row0 + (long long)(b0 + m) * P.f_stride);
                if (P.f_v) P.f_v[r] = S[P.o_v + m];
                if (P.f_logp) {                                                       // ppo.py:95-98 (sigma_sq used as the std)
                    float lp = 0.f;
                    for (int a = 0; a < d.A; ++a) {
                        const float ls = __ldg(nw.ls + a);
                        const float zz = (S[P.o_dmu + m * P.ldA + a] - S[P.o_mu + m * P.ldA + a]) / expf(ls);
                        const float t = (-0.5f * (zz * zz) - half_log_2pi) - ls;
                        lp = (a == 0) ? t : lp + t;
                    }
                    P.f_logp[r] = lp;
                }
            }
            __syncthreads();
        };

        if (P.trunk_serial) {
            // ---- wide nets: one trunk from its forward passes to its last gradient, then the other, in the same
            // activation buffers a1 (first hidden layer) and a2 (second hidden layer) --------------------------------
            float* a1 = S + P.o_h1;
            float* a2 = S + P.o_h2;
            for (int trunk = 0; trunk < 2; ++trunk) {
                const int N1 = trunk ? d.G1 : d.H1, N2 = trunk ? d.G2 : d.H2;
                const int ld1 = trunk ? P.ldG1 : P.ldH1, ld2 = trunk ? P.ldG2 : P.ldH2;
                if (trunk) {                                  // pad columns of the buffers must read as 0 again
                    for (int i = P.o_h1 + threadIdx.x; i < P.o_act; i += UPD_THREADS) S[i] = 0.f;
                    __syncthreads();
                }
                if (run_old)
                    ph = trunk_forward(no, hs_old, P, S, trunk, S + P.o_muo, S + P.o_vo, ph, tmem, P.use_ts ? P.map_fwd[1] : nullptr);
                ph = trunk_forward(nw, hs_new, P, S, trunk, S + P.o_mu, S + P.o_v, ph, tmem, P.use_ts ? P.map_fwd[0] : nullptr);
                if (P.fwd_only) continue;
                if (same_old) {
                    if (trunk == 0) for (int e = threadIdx.x; e < UPD_ROWS * P.ldA; e += UPD_THREADS) S[P.o_muo + e] = S[P.o_mu + e];
                    else for (int e = threadIdx.x; e < UPD_ROWS; e += UPD_THREADS) S[P.o_vo + e] = S[P.o_v + e];
                    __syncthreads();
                }
                if (P.dwg) {
                    // batch-level dW: this trunk's layer inputs (a1, a2) and dz leave for Z as they appear; no dW here
                    float* zc = P.Z + tile * UPD_ROWS;
                    z_dump(zc + (size_t)(trunk ? P.zr_g1 : P.zr_h1) * P.Bp, P.Bp, a1, ld1, N1);
                    z_dump(zc + (size_t)(trunk ? P.zr_g2 : P.zr_h2) * P.Bp, P.Bp, a2, ld2, N2);
                    if (trunk == 0) loss_policy(); else loss_value();
                    __syncthreads();
                    if (trunk == 0) {
                        z_dump(zc + (size_t)P.zr_dmu * P.Bp, P.Bp, S + P.o_dmu, P.ldA, d.A);
                        z_dump(zc + (size_t)(P.zr_dmu + d.A) * P.Bp, P.Bp, S + P.o_muo, P.ldA, d.A);
                        head_dx(hs_new.Wmu, hs_new.ldw, d.A, d.H2, S + P.o_dmu, P.ldA, a2, ld2);              // a2 <- dz2
                    } else {
                        z_dump(zc + (size_t)P.zr_dv * P.Bp, P.Bp, S + P.o_dv, 4, 1);
                        head_dx(hs_new.Wv, 0, 1, d.G2, S + P.o_dv, 4, a2, ld2);
                    }
                    __syncthreads();
                    z_dump(zc + (size_t)(trunk ? P.zr_d2v : P.zr_d2p) * P.Bp, P.Bp, a2, ld2, N2);
                    ph = dx_layer(trunk ? nw.V2 : nw.W2, N2, N1, a2, ld2, a1, ld1, stage, P.SF, bars, ph, &P.map_dx[trunk], tmem);   // a1 <- dz1
                    z_dump(zc + (size_t)(trunk ? P.zr_d1v : P.zr_d1p) * P.Bp, P.Bp, a1, ld1, N1);
                    __syncthreads();                          // the buffers are zeroed for the other trunk / the next tile
                    continue;
                }
                if (trunk == 0) {
                    loss_policy();
                    __syncthreads();
                    dls_reduce();
                } else {
                    loss_value();
                }
                __syncthreads();
                heads_dw(go.Wmu, go.bmu, go.Wv, go.bv, d.A, d.H2, d.G2, S + P.o_dmu, P.ldA, S + P.o_dv, a2, P.ldH2, a2, P.ldG2,
                         acc, 1 + trunk);
                if (trunk == 0) head_dx(hs_new.Wmu, hs_new.ldw, d.A, d.H2, S + P.o_dmu, P.ldA, a2, ld2);      // a2 <- dz2
                else head_dx(hs_new.Wv, 0, 1, d.G2, S + P.o_dv, 4, a2, ld2);
                dw_any(trunk ? go.V2 : go.W2, trunk ? go.c2 : go.b2, N2, N1, a2, ld2, a1, ld1);
                ph = dx_layer(trunk ? nw.V2 : nw.W2, N2, N1, a2, ld2, a1, ld1, stage, P.SF, bars, ph,
                              P.use_ts ? &P.map_dx[trunk] : nullptr, tmem);                                   // a1 <- dz1
                dw_any(trunk ? go.V1 : go.W1, trunk ? go.c1 : go.b1, N1, d.O, a1, ld1, S + P.o_x, P.ldO);
            }
            if (P.fwd_only) write_policy_outputs();
            continue;
        }

        // ---- forward(s) -----------------------------------------------------------------------
        // epoch 0 of a cycle: model_old is a copy of model (train.py:128-129), so the caller passes
        // theta_old == theta and the old network's outputs are the new network's, bit for bit
        if (run_old)
            ph = net_forward(no, hs_old, P, S, S + P.o_muo, S + P.o_vo, 3, ph, tmem, P.use_ts ? P.map_fwd[1] : nullptr, &nw);              // train.py:142
        ph = net_forward(nw, hs_new, P, S, S + P.o_mu, S + P.o_v, 8, ph, tmem, P.use_ts ? P.map_fwd[0] : nullptr, nullptr,
                         P.dwg ? tile * UPD_ROWS : -1);                    // train.py:145 / ppo.py:148
        if (P.fwd_only) { write_policy_outputs(); continue; }
        if (same_old) {
            for (int e = threadIdx.x; e < UPD_ROWS * P.ldA; e += UPD_THREADS) S[P.o_muo + e] = S[P.o_mu + e];
            for (int e = threadIdx.x; e < UPD_ROWS; e += UPD_THREADS) S[P.o_vo + e] = S[P.o_v + e];
            __syncthreads();
        }
        loss_policy();
        loss_value();
        __syncthreads();
        K2_TICK(13);
        if (P.dwg) {
            // ---- backward, batch-level dW: only the dz chain runs here; every dz goes to Z next to the layer inputs
            // the forward pass left there, and the dW pass (second launch) forms all weight gradients from them
            float* zc = P.Z + tile * UPD_ROWS;                             // this tile's 32 columns of Z
            z_dump(zc + (size_t)P.zr_dmu * P.Bp, P.Bp, S + P.o_dmu, P.ldA, d.A);
            z_dump(zc + (size_t)(P.zr_dmu + d.A) * P.Bp, P.Bp, S + P.o_muo, P.ldA, d.A);      // dls (kept in the mu_old buffer)
            z_dump(zc + (size_t)P.zr_dv * P.Bp, P.Bp, S + P.o_dv, 4, 1);
            head_dx2(hs_new.Wmu, hs_new.ldw, d.A, d.H2, S + P.o_dmu, P.ldA, S + P.o_h2, P.ldH2,       // h2 <- dz2, g2 <- dz2 (value)
                     hs_new.Wv, d.G2, S + P.o_dv, 4, S + P.o_g2, P.ldG2);
            z_dump(zc + (size_t)P.zr_d2p * P.Bp, P.Bp, S + P.o_h2, P.ldH2, d.H2);
            z_dump(zc + (size_t)P.zr_d2v * P.Bp, P.Bp, S + P.o_g2, P.ldG2, d.G2);
            K2_TICK(17);
            // the wider trunk first: its rows of dz1 drain while the other trunk's dX runs
            const bool vfirst = d.G1 >= d.H1;
            for (int pass = 0; pass < 2; ++pass) {
                if ((pass == 0) == vfirst) {
                    ph = dx_layer(nw.V2, d.G2, d.G1, S + P.o_g2, P.ldG2, S + P.o_g1, P.ldG1, stage, P.SF, bars, ph, &P.map_dx[1], tmem);   // g1 <- dz1 (value)
                    z_dump(zc + (size_t)P.zr_d1v * P.Bp, P.Bp, S + P.o_g1, P.ldG1, d.G1);
                } else {
                    ph = dx_layer(nw.W2, d.H2, d.H1, S + P.o_h2, P.ldH2, S + P.o_h1, P.ldH1, stage, P.SF, bars, ph, &P.map_dx[0], tmem);   // h1 <- dz1
                    z_dump(zc + (size_t)P.zr_d1p * P.Bp, P.Bp, S + P.o_h1, P.ldH1, d.H1);
                }
                K2_TICK(20 + pass);
            }
            __syncthreads();                                              // the next tile zeroes the activation area
            K2_TICK(21);
            continue;
        }
        dls_reduce();
        __syncthreads();

        K2_TICK(14);
        // ---- backward ---------------------------------------------------------------------------
        heads_dw(go.Wmu, go.bmu, go.Wv, go.bv, d.A, d.H2, d.G2, S + P.o_dmu, P.ldA, S + P.o_dv, S + P.o_h2, P.ldH2,
                 S + P.o_g2, P.ldG2, acc);
        K2_TICK(15);
        // The two trunks are independent, so every layer routine runs for both of them back to back: its code is
        // fetched once (the routines together do not fit the 32 KB instruction cache).
        head_dx(hs_new.Wmu, hs_new.ldw, d.A, d.H2, S + P.o_dmu, P.ldA, S + P.o_h2, P.ldH2);          // h2 <- dz2
        K2_TICK(16);
        head_dx(hs_new.Wv, 0, 1, d.G2, S + P.o_dv, 4, S + P.o_g2, P.ldG2);                          // g2 <- dz2 (value)
        K2_TICK(17);
        dw_any(go.W2, go.b2, d.H2, d.H1, S + P.o_h2, P.ldH2, S + P.o_h1, P.ldH1);
        K2_TICK(18);
        dw_any(go.V2, go.c2, d.G2, d.G1, S + P.o_g2, P.ldG2, S + P.o_g1, P.ldG1);
        K2_TICK(19);
        ph = dx_layer(nw.W2, d.H2, d.H1, S + P.o_h2, P.ldH2, S + P.o_h1, P.ldH1, stage, P.SF, bars, ph,
                      P.use_ts ? &P.map_dx[0] : nullptr, tmem);                                     // h1 <- dz1
        K2_TICK(20);
        ph = dx_layer(nw.V2, d.G2, d.G1, S + P.o_g2, P.ldG2, S + P.o_g1, P.ldG1, stage, P.SF, bars, ph,
                      P.use_ts ? &P.map_dx[1] : nullptr, tmem);                                     // g1 <- dz1 (value)
        K2_TICK(21);
        if (P.first_ts && d.H1 <= 512)
            ph = dw_first_ts(go.W1, go.b1, d.H1, d.O, S + P.o_h1, P.ldH1, S + P.o_x, P.ldO, acc, stage, first_b_tiles(stage, d), bars, ph, tmem);
        else if (d.O <= 24 && d.H1 >= 256 && d.H1 * d.O <= P.SF) dw_layer_smallk(go.W1, go.b1, d.H1, d.O, S + P.o_h1, P.ldH1, S + P.o_x, P.ldO, acc, stage);
        else dw_any(go.W1, go.b1, d.H1, d.O, S + P.o_h1, P.ldH1, S + P.o_x, P.ldO);
        K2_TICK(22);
        if (P.first_ts && d.G1 <= 512)
            ph = dw_first_ts(go.V1, go.c1, d.G1, d.O, S + P.o_g1, P.ldG1, S + P.o_x, P.ldO, acc, stage, first_b_tiles(stage, d), bars, ph, tmem);
        else if (d.O <= 24 && d.G1 >= 256 && d.G1 * d.O <= P.SF) dw_layer_smallk(go.V1, go.c1, d.G1, d.O, S + P.o_g1, P.ldG1, S + P.o_x, P.ldO, acc, stage);
        else dw_any(go.V1, go.c1, d.G1, d.O, S + P.o_g1, P.ldG1, S + P.o_x, P.ldO);
        K2_TICK(23);
        K2_TICK(24);
    }

    K2_TICK(25);
    // ---- per-CTA loss sums (scaled so that K3 only has to add them) -----------------------------
    float sc, sv, se;                                                 // the three loss sums in one pass over the tree
    block_sum3(lsum_clip, lsum_val, lsum_ent, red, S + P.o_mu, sc, sv, se);
    if (threadIdx.x == 0 && !P.fwd_only) {
        float* lp = (P.dwg ? P.zloss : P.loss_partials) + 4 * blockIdx.x;
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
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

// layer-routine selection (dppo_update_set_path); the environment is consulted once, for the initial value
int g_k2_path = -1;
int k2_path() {
    if (g_k2_path < 0) g_k2_path = getenv("DPPO_K2_NO_TS") != nullptr ? DPPO_K2_PATH_FP32 : DPPO_K2_PATH_AUTO;
    return g_k2_path;
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
    // a tile always occupies 32 MMA columns: once full tiles fit in one wave, fill them (B = 4096: 128 CTAs of 32
    // samples instead of 147 of 28 - the same time per CTA, 13 % fewer gradient partials for K3 to add)
    if (tm > 16 && (B + UPD_TM_MAX - 1) / UPD_TM_MAX <= sms) tm = UPD_TM_MAX;
    const int ntiles = (B + tm - 1) / tm;
    *TM = tm;
    *nparts = ntiles < sms ? ntiles : sms;
}

// layers that can run on the tensor cores (mirrored by the conditions in fwd_layer / dx_layer / dw_any)
bool ts_eligible(int N, int K) { return N >= 32 && K >= 64 && (K & 3) == 0; }

// Shared-memory carve-up.  Preference order: everything resident with the 100 KB staging region that the tensor-core
// layers need; if that does not fit (wide nets), first the head weights move to global memory, then the two trunks
// share their activation buffers (trunk-serial tiles: P.o_g1 = P.o_h1, P.o_g2 = P.o_h2); if the staging region still
// does not fit, the FP32-pipe layers run with whatever staging is left (>= UPD_STAGE_MIN).
int plan(UpdArgs& P) {
    const dppo_dims& d = P.d;
    P.ldO = ld_of(d.O); P.ldH1 = ld_of(d.H1); P.ldH2 = ld_of(d.H2); P.ldG1 = ld_of(d.G1); P.ldG2 = ld_of(d.G2);
    P.ldA = ld_of(d.A);
    const int TM = UPD_ROWS;                    // tiles always hold 32 rows; P.TM of them are samples
    P.head_floats = (d.A * P.ldH2 + ((d.A + 3) & ~3) + P.ldG2 + 4 + 3) & ~3;
    const bool heads_global_ok = (d.H2 & 3) == 0 && (d.G2 & 3) == 0;     // global head rows must be 16-byte aligned
    const bool want_ts = k2_path() == DPPO_K2_PATH_AUTO &&
                         (ts_eligible(d.H2, d.H1) || ts_eligible(d.G2, d.G1) || ts_eligible(d.H1, d.O) || ts_eligible(d.G1, d.O));
    auto layout = [&](bool heads_smem, bool serial) {
        int o = 0;
        auto take = [&](int n) { int r = o; o += (n + 3) & ~3; return r; };
        P.o_x = take(TM * P.ldO);
        if (serial) {
            P.o_h1 = take(TM * (P.ldH1 > P.ldG1 ? P.ldH1 : P.ldG1));
            P.o_h2 = take(TM * (P.ldH2 > P.ldG2 ? P.ldH2 : P.ldG2));
            P.o_g1 = P.o_h1;
            P.o_g2 = P.o_h2;
        } else {
            P.o_h1 = take(TM * P.ldH1);
            P.o_h2 = take(TM * P.ldH2);
            P.o_g1 = take(TM * P.ldG1);
            P.o_g2 = take(TM * P.ldG2);
        }
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
        P.o_zero_end = o;                           // everything before this is zeroed per tile
        P.o_red = take(UPD_THREADS);
        P.o_bars = take(2 * (TS_BAR0 + TS_NBARS) + 4);   // the mbarriers (8 bytes each) + the TMEM base address
        P.heads_in_smem = heads_smem ? 1 : 0;
        P.o_heads = take(heads_smem ? 2 * P.head_floats : 4);
        o = (o + 255) & ~255;                       // 1024-byte aligned stage buffers (UMMA descriptors, TMA swizzle atoms)
        P.o_stage = o;
        int sf = ((227 * 1024 / 4 - o) / 2) & ~127;  // two weight staging buffers from what is left of the 227 KB
        if (sf > UPD_STAGE_MAX) sf = UPD_STAGE_MAX;
        P.SF = sf;
        P.trunk_serial = serial ? 1 : 0;
        P.smem_floats = o + 2 * (sf > 0 ? sf : 0);
        return sf;
    };
    const int ts_floats = TS_BYTES / 8;             // per staging buffer
    int sf = layout(true, false);
    if (want_ts && sf < ts_floats) {
        bool done = false;
        if (heads_global_ok && layout(false, false) >= ts_floats) done = true;
        if (!done && layout(true, true) >= ts_floats) done = true;
        if (!done && heads_global_ok && layout(false, true) >= ts_floats) done = true;
        if (!done) sf = layout(true, false);
    }
    if (P.SF < UPD_STAGE_MIN && !P.trunk_serial && heads_global_ok) layout(false, false);   // FP32 pipe, tight shared memory
    const int o = P.smem_floats;
    // first layers from flat weight copies prefetched a phase ahead (narrow observations); not for trunk-serial tiles
    P.small_first = (!P.trunk_serial && d.O <= 32 && ((d.H1 * d.O + 3) & ~3) + d.G1 * d.O <= 2 * P.SF) ? 1 : 0;
    P.first_ts = (P.small_first && ((((d.H1 * d.O + 3) & ~3) + d.G1 * d.O + 31) & ~31) + 2 * TS_BT_BYTES / 4 <= 2 * P.SF &&
                  k2_path() == DPPO_K2_PATH_AUTO) ? 1 : 0;
    return o;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
// tensor map of a row-major fp32 matrix W[rows][cols] with a box of box_c x box_r elements
bool make_wmap(CUtensorMap* map, const float* W, int rows, int cols, int box_c, int box_r, CUtensorMapSwizzle swz) {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) return false;
    const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t gstr[1] = {(cuuint64_t)cols * 4};
    const cuuint32_t box[2] = {(cuuint32_t)box_c, (cuuint32_t)box_r};
    const cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(W), gdim, gstr, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, swz, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// The six tensor maps depend only on (device, theta, theta_old, dims): encoded once per combination and kept.
// An engine alternates between at most two (theta_old = theta in epoch 0, the lagging copy afterwards).
struct MapEntry {
    int dev;
    const float *theta, *theta_old;
    dppo_dims d;
    bool ok;
    CUtensorMap fwd[2][4], dx[2];
};
constexpr int MAP_CACHE = 16;
MapEntry g_maps[MAP_CACHE];
int g_maps_used = 0, g_maps_next = 0;

const MapEntry* wide_layer_maps(int dev, const float* theta, const float* theta_old, const dppo_dims& d, const Layout& lay) {
    for (int i = 0; i < g_maps_used; ++i) {
        const MapEntry& e = g_maps[i];
        if (e.dev == dev && e.theta == theta && e.theta_old == theta_old && memcmp(&e.d, &d, sizeof(d)) == 0) return &e;
    }
    MapEntry& e = g_maps[g_maps_next];
    g_maps_next = (g_maps_next + 1) % MAP_CACHE;
    if (g_maps_used < MAP_CACHE) ++g_maps_used;
    e.dev = dev; e.theta = theta; e.theta_old = theta_old; e.d = d;
    bool ok = true;
    for (int net = 0; net < 2 && ok; ++net) {
        const float* t = net == 0 ? theta : (theta_old ? theta_old : theta);
        if (ts_eligible(d.H2, d.H1))
            ok = ok && make_wmap(&e.fwd[net][0], t + lay.off[3], d.H2, d.H1, TC_KB, 128, CU_TENSOR_MAP_SWIZZLE_128B);
        if (ts_eligible(d.G2, d.G1))
            ok = ok && make_wmap(&e.fwd[net][1], t + lay.off[7], d.G2, d.G1, TC_KB, 128, CU_TENSOR_MAP_SWIZZLE_128B);
        if (ts_eligible(d.H1, d.O))                              // wide observations: the first layers too
            ok = ok && make_wmap(&e.fwd[net][2], t + lay.off[1], d.H1, d.O, TC_KB, 128, CU_TENSOR_MAP_SWIZZLE_128B);
        if (ts_eligible(d.G1, d.O))
            ok = ok && make_wmap(&e.fwd[net][3], t + lay.off[5], d.G1, d.O, TC_KB, 128, CU_TENSOR_MAP_SWIZZLE_128B);
    }
    if (ts_eligible(d.H2, d.H1))
        ok = ok && make_wmap(&e.dx[0], theta + lay.off[3], d.H2, d.H1, 128, TC_KB, CU_TENSOR_MAP_SWIZZLE_NONE);
    if (ts_eligible(d.G2, d.G1))
        ok = ok && make_wmap(&e.dx[1], theta + lay.off[7], d.G2, d.G1, 128, TC_KB, CU_TENSOR_MAP_SWIZZLE_NONE);
    e.ok = ok;
    return &e;
}

constexpr int MAX_DEVICES = 64;

// ---- batch-level dW (see UpdArgs::dwg) ------------------------------------------------------------------------
// Used when every hidden layer runs on the tensor-core routines (their epilogues hold a feature's 16 samples in
// registers, which is the layout Z wants) and the minibatch is large enough for the second launch to pay.  The decision depends on (dims, B, the K2 path, DPPO_K2_DWG) only, so
// dppo_update_nparts and dppo_update_fused agree.
constexpr int DWG_MAX_SPLIT = 8;                    // ranges of samples = gradient partials handed to K3 (chosen per shape)
constexpr int DWG_MIN_B = 1024;
struct DwgGeom { int on, Bp, rows, units, chunk, nk2, ksplit; int zr[11]; };
int g_dwg_enabled = -1;
bool dwg_enabled() {
    if (g_dwg_enabled < 0) g_dwg_enabled = getenv("DPPO_K2_DWG") ? atoi(getenv("DPPO_K2_DWG")) : 1;   // (2: timing experiment, first launch only)
    return g_dwg_enabled != 0;
}
DwgGeom dwg_geom(const dppo_dims& d, int B) {
    DwgGeom G;
    memset(&G, 0, sizeof(G));
    if (!dwg_enabled() || k2_path() != DPPO_K2_PATH_AUTO || B < DWG_MIN_B) return G;
    UpdArgs Q;                                      // the shared-memory plan decides which routines run
    memset(&Q, 0, sizeof(Q));
    Q.d = d; Q.B = B;
    int np;
    pick_tiling(B, &Q.TM, &np);
    plan(Q);
    if ((size_t)2 * Q.SF * sizeof(float) < (size_t)TS_BYTES) return G;
    if (!ts_eligible(d.H2, d.H1) || !ts_eligible(d.G2, d.G1)) return G;
    const bool first_small = Q.small_first && Q.first_ts && d.H1 <= 512 && d.G1 <= 512;
    if (!first_small && !(ts_eligible(d.H1, d.O) && ts_eligible(d.G1, d.O))) return G;
    const int ntiles = (B + Q.TM - 1) / Q.TM;
    G.on = 1;
    G.nk2 = np;
    G.Bp = ntiles * UPD_ROWS;
    int r = 0;
    auto in_rows = [&](int kf) { const int at = r; r += (kf + 1 + 31) & ~31; return at; };   // Kf rows + the ones row, pad to 32
    G.zr[0] = in_rows(d.O); G.zr[1] = in_rows(d.H1); G.zr[2] = in_rows(d.G1); G.zr[3] = in_rows(d.H2); G.zr[4] = in_rows(d.G2);
    auto dz_rows = [&](int n) { const int at = r; r += n; return at; };
    G.zr[5] = dz_rows(d.H1); G.zr[6] = dz_rows(d.G1); G.zr[7] = dz_rows(d.H2); G.zr[8] = dz_rows(d.G2);
    G.zr[9] = dz_rows(2 * d.A); G.zr[10] = dz_rows(1);
    G.rows = r;
    auto units = [](int rows_t, int rows_b) { return ((rows_b + 31) >> 5) * ((rows_t + 127) >> 7); };
    G.units = units(d.G2, d.G1 + 1) + units(d.H2, d.H1 + 1) + units(d.G1, d.O + 1) + units(d.H1, d.O + 1) +
              units(d.H2 + 1, 2 * d.A) + units(d.G2 + 1, 1);
    // ranges of samples: every CTA of the pass costs ~5 k cycles (fill, epilogue) + ~1.3 k per pair of k-blocks of its
    // range, the grid runs in waves of one CTA per SM, K3 pays ~0.3 k per partial bucket
    const int sms = dppo_sm_count();
    long best = -1;
    for (int k = 1; k <= DWG_MAX_SPLIT; ++k) {
        const int chunk = ((G.Bp + k - 1) / k + 63) & ~63;
        if ((long)chunk * (k - 1) >= G.Bp) continue;               // the last range would be empty
        const long waves = ((long)G.units * k + sms - 1) / sms;
        const long cost = waves * (5000 + 1300L * (chunk / 64)) + 300L * k;
        if (best < 0 || cost < best) { best = cost; G.ksplit = k; G.chunk = chunk; }
    }
    return G;
}
size_t dwg_scratch_bytes(const DwgGeom& G) { return sizeof(float) * ((size_t)G.rows * G.Bp + 4 * (size_t)G.nk2) + 256; }

// The per-device Z buffer: grown on demand, never while the stream is being captured into a graph, and a buffer that
// has been handed out is never freed (a captured graph may still point at it - the tensor maps and the pointer are
// baked into its kernel nodes), so growing leaves the smaller buffer allocated.
void* g_zbuf[64] = {nullptr};
size_t g_zcap[64] = {0};
int dwg_scratch(int dev, size_t bytes, cudaStream_t stream, float** out) {
    if (g_zcap[dev] < bytes) {
        cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
        if (stream != nullptr && cudaStreamIsCapturing(stream, &st) == cudaSuccess && st != cudaStreamCaptureStatusNone)
            return DPPO_E_SCRATCH;                  // dppo_update_prepare (or one plain call) must come before the capture
        void* p = nullptr;
        DPPO_CUDA_TRY(cudaMalloc(&p, bytes));
        g_zbuf[dev] = p;
        g_zcap[dev] = bytes;
    }
    *out = (float*)g_zbuf[dev];
    return 0;
}
struct DzMaps { int dev; const float* Z; dppo_dims d; int Bp; bool ok; CUtensorMap m[6]; CUtensorMap in[5]; };
DzMaps g_dzmaps[64];
const DzMaps* dz_maps(int dev, float* Z, const dppo_dims& d, const DwgGeom& G) {
    DzMaps& e = g_dzmaps[dev];
    if (e.Z == Z && e.Bp == G.Bp && e.dev == dev && memcmp(&e.d, &d, sizeof(d)) == 0 && e.ok) return &e;
    e.dev = dev; e.Z = Z; e.d = d; e.Bp = G.Bp;
    // tile operands: dz^T of v_fc2, p_fc2, v_fc1, p_fc1; the heads take their INPUTS (h2^T, g2^T with the ones row) here
    const int rows[6] = {G.zr[8], G.zr[7], G.zr[6], G.zr[5], G.zr[3], G.zr[4]};
    const int n[6] = {d.G2, d.H2, d.G1, d.H1, d.H2 + 1, d.G2 + 1};
    bool ok = true;
    for (int i = 0; i < 6 && ok; ++i)
        ok = make_wmap(&e.m[i], Z + (size_t)rows[i] * G.Bp, n[i], G.Bp, TC_KB, 128, CU_TENSOR_MAP_SWIZZLE_128B);
    // 32-row operands: the inputs x^T, h1^T, g1^T (Kf rows + ones, padded to 32); for the heads [dmu | dls]^T and dv^T
    const int brow[5] = {G.zr[0], G.zr[1], G.zr[2], G.zr[9], G.zr[10]};
    const int bn[5] = {(d.O + 1 + 31) & ~31, (d.H1 + 1 + 31) & ~31, (d.G1 + 1 + 31) & ~31, 2 * d.A, 1};
    for (int i = 0; i < 5 && ok; ++i)
        ok = make_wmap(&e.in[i], Z + (size_t)brow[i] * G.Bp, bn[i], G.Bp, TC_KB, 32, CU_TENSOR_MAP_SWIZZLE_NONE);
    e.ok = ok;
    return &e;
}

}  // namespace

extern "C" int dppo_update_prepare(const dppo_dims* d, int B) {
    const int rc = check_dims(d);
    if (rc) return rc;
    if (B < 1) return DPPO_E_BADARG;
    const DwgGeom G = dwg_geom(*d, B);
    if (!G.on) return 0;
    int dev = 0;
    DPPO_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return DPPO_E_BADARG;
    float* z;
    return dwg_scratch(dev, dwg_scratch_bytes(G), nullptr, &z);
}

namespace {

}  // namespace

// per-device caches below are indexed by the CUDA device ordinal: one host thread per device, any number of
// devices per process
int dppo_sm_count() {
    static int sms[MAX_DEVICES] = {0};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) return 148;
    if (sms[dev] == 0) {
        if (cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms[dev] <= 0)
            sms[dev] = 148;
    }
    return sms[dev];
}

extern "C" int dppo_update_set_path(int path) {
    if (path != DPPO_K2_PATH_AUTO && path != DPPO_K2_PATH_FP32) return DPPO_E_BADARG;
    const int prev = k2_path();
    g_k2_path = path;
    return prev;
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
    const DwgGeom G = dwg_geom(*d, B);
    return G.on ? G.ksplit : np;
}

extern "C" int dppo_update_set_batch_dw(int on) {
    const int prev = dwg_enabled() ? 1 : 0;
    g_dwg_enabled = on ? 1 : 0;
    return prev;
}

extern "C" int dppo_update_launches(const dppo_dims* d, int B) {
    const int rc = check_dims(d);
    if (rc) return rc;
    if (B < 1) return DPPO_E_BADARG;
    return dwg_geom(*d, B).on ? 2 : 1;
}

extern "C" size_t dppo_update_partials_bytes(const dppo_dims* d, int B) {
    if (check_dims(d) || B < 1) return 0;
    Layout L;
    fill_layout(*d, L);
    int TM, np;
    pick_tiling(B, &TM, &np);
    const DwgGeom G = dwg_geom(*d, B);
    if (G.on) np = G.ksplit;
    return sizeof(float) * (size_t)np * L.P_pad;
}

// shared tail of the two entry points: tiling, shared-memory plan, tensor maps, launch
static int plan_and_launch(UpdArgs& P, const dppo_dims* d, int B, const float* theta, const float* theta_old, void* stream_) {
    fill_layout(*d, P.lay);
    P.d = *d;
    P.B = B;
    int nparts;
    pick_tiling(B, &P.TM, &nparts);
    P.ntiles = (B + P.TM - 1) / P.TM;
    P.theta = theta; P.theta_old = theta_old;
    plan(P);
    const size_t smem = sizeof(float) * (size_t)P.smem_floats;
    if (smem > 227 * 1024 || P.SF < UPD_STAGE_MIN) return DPPO_E_TOOLARGE;
    int dev = 0;
    DPPO_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return DPPO_E_BADARG;
    // tensor-core path of the wide layers: needs the 100 KB staging region and tensor maps of the weights
    P.use_ts = 0;
    if ((size_t)2 * P.SF * sizeof(float) >= (size_t)TS_BYTES && k2_path() == DPPO_K2_PATH_AUTO) {
        const MapEntry* e = wide_layer_maps(dev, theta, theta_old, *d, P.lay);
        if (e->ok) {
            memcpy(P.map_fwd, e->fwd, sizeof(P.map_fwd));
            memcpy(P.map_dx, e->dx, sizeof(P.map_dx));
            P.use_ts = 1;
        }
    }
    static size_t smem_set[MAX_DEVICES] = {0};                  // the attribute is per device
    if (smem > smem_set[dev]) {
        DPPO_CUDA_TRY(cudaFuncSetAttribute(update_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set[dev] = smem;
    }
    const DwgGeom G = P.fwd_only ? DwgGeom{} : dwg_geom(*d, B);
    if (G.on) {
        if (!P.use_ts) return DPPO_E_TOOLARGE;                  // (the tensor maps could not be made)
        float* z;
        const int rc = dwg_scratch(dev, dwg_scratch_bytes(G), (cudaStream_t)stream_, &z);
        if (rc) return rc;
        const DzMaps* zm = dz_maps(dev, z, *d, G);
        if (!zm->ok) return DPPO_E_TOOLARGE;
        memcpy(P.map_dz, zm->m, sizeof(P.map_dz));
        memcpy(P.map_in, zm->in, sizeof(P.map_in));
        P.dwg = 1; P.Bp = G.Bp; P.dwg_units = G.units; P.dwg_chunk = G.chunk; P.zloss_rows = nparts;
        P.Z = z; P.zloss = z + (size_t)G.rows * G.Bp;
        P.zr_x = G.zr[0]; P.zr_h1 = G.zr[1]; P.zr_g1 = G.zr[2]; P.zr_h2 = G.zr[3]; P.zr_g2 = G.zr[4];
        P.zr_d1p = G.zr[5]; P.zr_d1v = G.zr[6]; P.zr_d2p = G.zr[7]; P.zr_d2v = G.zr[8]; P.zr_dmu = G.zr[9]; P.zr_dv = G.zr[10];
    }
    update_fused_kernel<<<nparts, UPD_THREADS, smem, (cudaStream_t)stream_>>>(P);
    DPPO_LAUNCH_CHECK();
    if (G.on && g_dwg_enabled != 2) {                           // second launch: all weight gradients from Z
        P.dwg = 2;
        update_fused_kernel<<<G.units * G.ksplit, UPD_THREADS, smem, (cudaStream_t)stream_>>>(P);
        DPPO_LAUNCH_CHECK();
    }
    return 0;
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
    memset(&P, 0, sizeof(P));
    P.flavour = flavour;
    P.obs = obs; P.act = act; P.adv = adv; P.ret = ret; P.logp_old = logp_old; P.idx = idx;
    P.clip = clip; P.ent_coeff = ent_coeff;
    P.partials = partials; P.loss_partials = loss_partials;
    return plan_and_launch(P, d, B, theta, theta_old, stream_);
}

// Forward-only use of the same kernel: the tile's gather and net_forward, then mu / v / action / log-prob instead of
// the loss and the backward pass.
extern "C" int dppo_policy_forward(const dppo_dims* d, const float* states, int N, const float* theta, const float* eps,
                                   int std_is_variance, float* mu_out, float* v_out, float* act_out, float* logp_out,
                                   float* states_out, int64_t out_row0, int64_t out_row_stride, void* stream_) {
    int rc = check_dims(d);
    if (rc) return rc;
    if (N < 1 || !states || !theta || out_row0 < 0 || out_row_stride < 1) return DPPO_E_BADARG;
    if (!mu_out && !v_out && !act_out && !logp_out) return DPPO_E_BADARG;
    if (!aligned16(theta)) return DPPO_E_ALIGN;
    UpdArgs P;
    memset(&P, 0, sizeof(P));
    P.flavour = DPPO_FLAVOUR_PPO;                               // one network, no lagging copy
    P.obs = states;
    P.fwd_only = 1;
    P.std_is_variance = std_is_variance ? 1 : 0;
    P.eps = eps;
    P.f_mu = mu_out; P.f_v = v_out; P.f_act = act_out; P.f_logp = logp_out; P.f_obs = states_out;
    P.f_row0 = out_row0; P.f_stride = out_row_stride;
    return plan_and_launch(P, d, N, theta, nullptr, stream_);
}

#ifdef DPPO_K2_TRACE
extern "C" int dppo_debug_k2_trace_bias(const float* bias_dev) {
    return (int)cudaMemcpyToSymbol(g_k2_trace_bias, &bias_dev, sizeof(bias_dev));
}
extern "C" int dppo_debug_k2_trace(long long* out_host) {
    return (int)cudaMemcpyFromSymbol(out_host, g_k2_trace, sizeof(long long) * 4 * 64 * 8);
}
#endif
#ifdef DPPO_K2_TIMING
extern "C" int dppo_debug_k2_clocks(long long* out64_host) {
    return (int)cudaMemcpyFromSymbol(out64_host, g_k2_clock, sizeof(long long) * 64);
}
#endif
