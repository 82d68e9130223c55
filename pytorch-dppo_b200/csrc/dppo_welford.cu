// K4 - running observation statistics + normalisation over a whole rollout.
//
// Reference: Shared_obs_stats.observes / normalize (model.py:68-80), called
// once per environment step (train.py:84-85): n += 1; mean += (x-mean)/n;
// M2 += (x-mean_old)(x-mean); var = max(M2/n, 1e-2);
// xhat = clamp((x-mean)/sqrt(var), -5, 5) - row t is normalised with the
// statistics that already include row t.
//
// Batched form: with the pivot p = the mean before the rollout, the state after
// row t follows from two prefix sums S1_t = sum_{s<=t}(x_s-p), S2_t =
// sum_{s<=t}(x_s-p)^2:   n_t = n0+t+1,  mean_t = p + S1_t/n_t,
// M2_t = M2_0 + S2_t - S1_t^2/n_t.   (x-p) is formed in fp32 (exact pivot), the
// sums are carried in fp64, mean_t / M2_t are rounded to fp32 and the
// normalisation is the reference's fp32 arithmetic.
//
// Three launches: (a) per-tile column sums, (b) exclusive scan of the tile sums per column + new
// state, (c) per-row prefix + normalise.  obs is read in (a) and (c): the second read is an L2 hit
// while the rollout is smaller than L2 (every named config), else HBM sees 12*O B/row.
#include "common.cuh"

namespace {

constexpr int WF_THREADS = 256;
constexpr int WF_TILE_ELEMS = 8192;                 // 32 KB of fp32 per CTA (16 KB and 64 KB were slower)
constexpr int64_t WF_CHUNK_BYTES = (int64_t)1 << 40;  // rows per launch group (chunking for L2 reuse did not pay: see DESIGN.md)

struct WfGeom {
    int O, R, G, RPG;                               // rows per tile, row groups per tile, rows per group
    int64_t T;
    int ntiles;
};

__host__ __device__ inline int wf_rows_per_tile(int O) {
    int R = (WF_TILE_ELEMS / O) & ~3;               // multiple of 4 rows keeps tiles 16-byte aligned
    if (R > 2048) R = 2048;                         // bounds the per-row reciprocal tables
    return R < 4 ? 4 : R;
}

// KEEP: the aggregate pass leaves the rows in L2 for the apply pass (which streams them out)
template <bool KEEP>
__device__ __forceinline__ void wf_load_tile(float* s, const float* __restrict__ g, int64_t off, int n, bool vec) {
    if (vec && (n & 3) == 0) {
        const float4* g4 = reinterpret_cast<const float4*>(g + off);
        float4* s4 = reinterpret_cast<float4*>(s);
        for (int i = threadIdx.x; i < n / 4; i += WF_THREADS) s4[i] = KEEP ? __ldg(&g4[i]) : __ldcs(&g4[i]);
    } else {
        for (int i = threadIdx.x; i < n; i += WF_THREADS) s[i] = g[off + i];
    }
}

// group-local sums: item (g,c) walks rows [g*RPG, min((g+1)*RPG, rows)) of column c
__device__ __forceinline__ void wf_group_sums(const float* s, const float* s_p, const WfGeom& q, int rows,
                                              double* s_g1, double* s_g2) {
    for (int it = threadIdx.x; it < q.G * q.O; it += WF_THREADS) {
        const int g = it / q.O, c = it - g * q.O;
        const float p = s_p[c];
        double a1 = 0.0, a2 = 0.0;
        const int r0 = g * q.RPG, r1 = min(r0 + q.RPG, rows);
        for (int r = r0; r < r1; ++r) {
            const double d = (double)__fsub_rn(s[r * q.O + c], p);
            a1 += d;
            a2 = fma(d, d, a2);
        }
        s_g1[it] = a1;
        s_g2[it] = a2;
    }
}

// dynamic smem: float tile[R*O] | float p[O] | double g1[G*O] | double g2[G*O]
__global__ void __launch_bounds__(WF_THREADS)
wf_aggregate_kernel(const float* __restrict__ obs, WfGeom q, const float* __restrict__ n, const float* __restrict__ mean,
                    const float* __restrict__ mean_diff, float* __restrict__ state0, double* __restrict__ agg, bool vec) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* s = reinterpret_cast<float*>(smem_raw);
    float* s_p = s + q.R * q.O;
    double* s_g1 = reinterpret_cast<double*>(s_p + ((q.O + 3) & ~3));
    double* s_g2 = s_g1 + q.G * q.O;
    const int tile = blockIdx.x;
    const int64_t row0 = (int64_t)tile * q.R;
    const int rows = (int)min((int64_t)q.R, q.T - row0);
    for (int c = threadIdx.x; c < q.O; c += WF_THREADS) {
        s_p[c] = mean[c];
        if (tile == 0) {                              // keep the pre-rollout state for pass (c)
            state0[c] = mean[c];
            state0[q.O + c] = mean_diff[c];
            if (c == 0) state0[2 * q.O] = n[0];
        }
    }
    wf_load_tile<true>(s, obs, row0 * q.O, rows * q.O, vec);
    __syncthreads();
    wf_group_sums(s, s_p, q, rows, s_g1, s_g2);
    __syncthreads();
    for (int c = threadIdx.x; c < q.O; c += WF_THREADS) {
        double a1 = 0.0, a2 = 0.0;
        for (int g = 0; g < q.G; ++g) { a1 += s_g1[g * q.O + c]; a2 += s_g2[g * q.O + c]; }
        agg[(size_t)c * q.ntiles + tile] = a1;                                   // [2][O][ntiles]
        agg[(size_t)(q.O + c) * q.ntiles + tile] = a2;
    }
}

// one CTA per column: exclusive scan of the tile sums (in place) and the new state
__global__ void __launch_bounds__(WF_THREADS)
wf_scan_kernel(WfGeom q, double* __restrict__ agg, const float* __restrict__ state0, float* __restrict__ n,
               float* __restrict__ mean, float* __restrict__ mean_diff, float* __restrict__ var) {
    __shared__ double part[2][WF_THREADS];
    const int c = blockIdx.x;
    double* a1 = agg + (size_t)c * q.ntiles;
    double* a2 = agg + (size_t)(q.O + c) * q.ntiles;
    const int L = (q.ntiles + WF_THREADS - 1) / WF_THREADS;
    const int t0 = min(threadIdx.x * L, q.ntiles), t1 = min(t0 + L, q.ntiles);
    double s1 = 0.0, s2 = 0.0;
    for (int t = t0; t < t1; ++t) { s1 += a1[t]; s2 += a2[t]; }
    part[0][threadIdx.x] = s1;
    part[1][threadIdx.x] = s2;
    __syncthreads();
    if (threadIdx.x < 2) {                             // serial exclusive scan of 256 partials, fixed order
        double run = 0.0;
        for (int i = 0; i < WF_THREADS; ++i) { const double x = part[threadIdx.x][i]; part[threadIdx.x][i] = run; run += x; }
        if (threadIdx.x == 0) part[0][0] = 0.0;
    }
    __syncthreads();
    double r1 = part[0][threadIdx.x], r2 = part[1][threadIdx.x];
    for (int t = t0; t < t1; ++t) {
        const double x1 = a1[t], x2 = a2[t];
        a1[t] = r1; a2[t] = r2;
        r1 += x1; r2 += x2;
    }
    __syncthreads();
    if (threadIdx.x == WF_THREADS - 1) {               // r1, r2 of the last thread = column totals
        // (threads whose chunk is empty carry the prefix through unchanged)
        const float p = state0[c];
        const double M0 = (double)state0[q.O + c];
        const double n1 = (double)state0[2 * q.O] + (double)q.T;
        const double inv = 1.0 / n1;
        const float nn = (float)n1;
        const float mm = (float)((double)p + r1 * inv);
        const float md = (float)(M0 + r2 - r1 * r1 * inv);
        n[c] = nn;
        mean[c] = mm;
        mean_diff[c] = md;
        var[c] = fmaxf(__fdiv_rn(md, nn), 1e-2f);       // model.py:75
    }
}

__global__ void __launch_bounds__(WF_THREADS)
wf_apply_kernel(const float* __restrict__ obs, float* __restrict__ out, WfGeom q, const float* __restrict__ state0,
                const double* __restrict__ agg, bool vec) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float* s = reinterpret_cast<float*>(smem_raw);
    float* s_p = s + q.R * q.O;
    double* s_g1 = reinterpret_cast<double*>(s_p + ((q.O + 3) & ~3));
    double* s_g2 = s_g1 + q.G * q.O;
    const int tile = blockIdx.x;
    const int64_t row0 = (int64_t)tile * q.R;
    const int rows = (int)min((int64_t)q.R, q.T - row0);
    for (int c = threadIdx.x; c < q.O; c += WF_THREADS) s_p[c] = state0[c];
    wf_load_tile<false>(s, obs, row0 * q.O, rows * q.O, vec);
    __syncthreads();
    wf_group_sums(s, s_p, q, rows, s_g1, s_g2);
    __syncthreads();
    // exclusive prefix over the row groups of each column, seeded with the tile prefix
    for (int c = threadIdx.x; c < q.O; c += WF_THREADS) {
        double r1 = agg[(size_t)c * q.ntiles + tile], r2 = agg[(size_t)(q.O + c) * q.ntiles + tile];
        for (int g = 0; g < q.G; ++g) {
            const double x1 = s_g1[g * q.O + c], x2 = s_g2[g * q.O + c];
            s_g1[g * q.O + c] = r1; s_g2[g * q.O + c] = r2;
            r1 += x1; r2 += x2;
        }
    }
    __syncthreads();
    const double n0 = (double)state0[2 * q.O];
    // per-row 1/n_t (fp64 and fp32), once per row instead of once per element: the int->fp64 and
    // fp64->fp32 conversions run at 1/8 of the FP32 rate and were the bottleneck of this kernel
    double* s_inv = s_g2 + q.G * q.O;
    float* s_nf = reinterpret_cast<float*>(s_inv + q.R);
    for (int r = threadIdx.x; r < rows; r += WF_THREADS) {
        const double nt = n0 + (double)(row0 + r + 1);
        const double r0 = (double)__frcp_rn((float)nt);       // fp32 reciprocal + one Newton step in fp64
        s_inv[r] = r0 * (2.0 - nt * r0);
        s_nf[r] = (float)s_inv[r];                             // 1/n_t in fp32
    }
    __syncthreads();
    for (int it = threadIdx.x; it < q.G * q.O; it += WF_THREADS) {
        const int g = it / q.O, c = it - g * q.O;
        const float p = s_p[c];
        const double M0 = (double)state0[q.O + c];
        double a1 = s_g1[it], a2 = s_g2[it];
        const int r0 = g * q.RPG, r1 = min(r0 + q.RPG, rows);
        for (int r = r0; r < r1; ++r) {
            const float x = s[r * q.O + c];
            const double d = (double)__fsub_rn(x, p);
            a1 += d;
            a2 = fma(d, d, a2);
            const double t = a1 * s_inv[r];
            const float mean_t = (float)((double)p + t);
            const float md_t = (float)(M0 + a2 - a1 * t);
            // var = max(M2/n, 1e-2), xhat = (x-mean)/sqrt(var) (model.py:75-80) as multiply-by-reciprocal and
            // rsqrt: within 2 ulp of the divide/sqrt form, an order of magnitude fewer instructions
            const float var_t = fmaxf(md_t * s_nf[r], 1e-2f);
            const float xh = __fsub_rn(x, mean_t) * rsqrtf(var_t);
            s[r * q.O + c] = fminf(fmaxf(xh, -5.f), 5.f);
        }
    }
    __syncthreads();
    const int nel = rows * q.O;
    const int64_t off = row0 * q.O;
    if (vec && (nel & 3) == 0) {
        float4* o4 = reinterpret_cast<float4*>(out + off);
        const float4* s4 = reinterpret_cast<const float4*>(s);
        for (int i = threadIdx.x; i < nel / 4; i += WF_THREADS) __stcs(&o4[i], s4[i]);
    } else {
        for (int i = threadIdx.x; i < nel; i += WF_THREADS) out[off + i] = s[i];
    }
}

WfGeom wf_geom(int64_t T, int O) {
    WfGeom q;
    q.O = O; q.T = T;
    q.R = wf_rows_per_tile(O);
    q.G = WF_THREADS / O;
    if (q.G < 1) q.G = 1;
    if (q.G > q.R) q.G = q.R;
    q.RPG = (q.R + q.G - 1) / q.G;
    q.ntiles = (int)ceil_div64(T, q.R);
    return q;
}

size_t wf_smem_bytes(const WfGeom& q) {
    return sizeof(float) * ((size_t)q.R * q.O + ((q.O + 3) & ~3)) + 2 * sizeof(double) * (size_t)q.G * q.O +
           (sizeof(double) + sizeof(float)) * (size_t)q.R + 16;            // + per-row 1/n (fp64) and n (fp32)
}

}  // namespace

// scratch: float state0[2*O+1 (padded)] | double agg[2][O][ntiles]
extern "C" size_t dppo_welford_scratch_bytes(int64_t T, int O) {
    if (T <= 0 || O <= 0) return 16;
    const WfGeom q = wf_geom(T, O);
    return align_up(sizeof(float) * (size_t)(2 * O + 1), 16) + 2 * sizeof(double) * (size_t)O * q.ntiles;
}

extern "C" int dppo_welford(const float* obs, float* out, int64_t T, int O, float* n, float* mean, float* mean_diff,
                            float* var, void* scratch, size_t scratch_bytes, void* stream_) {
    if (T < 0 || O <= 0) return DPPO_E_BADARG;
    if (T == 0) return 0;
    if (!obs || !out || !n || !mean || !mean_diff || !var || !scratch) return DPPO_E_BADARG;
    if (O > 4096) return DPPO_E_TOOLARGE;
    if (ceil_div64(T, wf_rows_per_tile(O)) > 0x7fffffff) return DPPO_E_TOOLARGE;
    if (scratch_bytes < dppo_welford_scratch_bytes(T, O)) return DPPO_E_SCRATCH;
    if (!aligned16(scratch)) return DPPO_E_ALIGN;
    cudaStream_t stream = (cudaStream_t)stream_;
    float* state0 = (float*)scratch;
    double* agg = (double*)((char*)scratch + align_up(sizeof(float) * (size_t)(2 * O + 1), 16));
    const bool vec = aligned16(obs) && aligned16(out);
    // Row chunks (WF_CHUNK_BYTES each; one chunk in practice).  The running state is carried from chunk
    // to chunk exactly as between two calls.
    const int R = wf_rows_per_tile(O);
    int64_t chunk_rows = ((int64_t)WF_CHUNK_BYTES / ((int64_t)O * 4) / R) * R;
    if (chunk_rows < R) chunk_rows = R;
    for (int64_t t0 = 0; t0 < T; t0 += chunk_rows) {
        const int64_t tc = (T - t0 < chunk_rows) ? T - t0 : chunk_rows;
        const WfGeom q = wf_geom(tc, O);
        const size_t smem = wf_smem_bytes(q);
        if (smem > 48 * 1024 && t0 == 0) {
            DPPO_CUDA_TRY(cudaFuncSetAttribute(wf_aggregate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            DPPO_CUDA_TRY(cudaFuncSetAttribute(wf_apply_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        }
        const float* in = obs + t0 * O;
        float* o = out + t0 * O;
        wf_aggregate_kernel<<<q.ntiles, WF_THREADS, smem, stream>>>(in, q, n, mean, mean_diff, state0, agg, vec);
        DPPO_LAUNCH_CHECK();
        wf_scan_kernel<<<O, WF_THREADS, 0, stream>>>(q, agg, state0, n, mean, mean_diff, var);
        DPPO_LAUNCH_CHECK();
        wf_apply_kernel<<<q.ntiles, WF_THREADS, smem, stream>>>(in, o, q, state0, agg, vec);
        DPPO_LAUNCH_CHECK();
    }
    return 0;
}
