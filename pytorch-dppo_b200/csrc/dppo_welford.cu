// K4 - running observation statistics + normalisation over a whole rollout, ONE pass over HBM.
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
// One launch.  A CTA owns a tile of R rows x CB columns (~33 KB of fp32; CB = O for O <= 32, else column
// blocks of 32), reads it from HBM once into shared memory, forms its column sums, obtains the sums of all
// rows above it from the tiles before it, normalises the tile in shared memory and streams it out: 8*O
// bytes per row of HBM traffic.  Tiles are handed out by an atomic ticket, so every tile above a running
// tile has started and publishes its sums without depending on anything below it (no deadlock whatever
// the residency).
//
// The prefix over the tiles of a column block is DETERMINISTIC and has NO serial chain.  fp64 addition is not
// associative, so the classic decoupled look-back (which adds whatever mix of aggregates and prefixes it happens to
// find) would make results depend on timing; and a chain of inclusive prefixes handed from tile to tile costs an L2
// round trip per link.  Instead the tiles form a fixed radix-16 hierarchy: level 0 = tiles, level l = groups of 16^l
// tiles.  The last tile of a group publishes the group's sums (tot[l]) from the 16 sums one level down - that needs
// nothing from outside the group - and every tile adds, in a fixed order (highest level first),
//     for each level l:  tot[l][the d_l groups before mine inside my level-(l+1) group]      (d_l = digit l of the
// tile number in base 16: at most 15 payloads per level, 5 levels cover 10^6 tiles).  Every wait is on a flag with
// acquire/release semantics and only ever on tiles with smaller ticket numbers, which have started.
#include "common.cuh"

namespace {

constexpr int WF_THREADS = 256;
constexpr int WF_CB = 32;                           // widest column block
constexpr int WF_L1 = 16;                           // radix of the hierarchy
constexpr int WF_MAX_LEV = 6;                       // 16^6 tiles
constexpr int WF_LOOK = 16;                         // payloads staged per pass of the look-back (smem per CTA decides the occupancy)
constexpr int WF_MAX_ROWS = 8448;                   // bounds the per-row reciprocal table (256 groups x 33 rows, O = 1)

struct WfGeom {
    int O, CB, ncb;                                 // columns, columns per block (last block may be narrower), blocks
    int R, G, RPG;                                  // rows per tile, row groups per tile, rows per group
    int64_t T;
    int ntr;                                        // row tiles; tiles in all = ntr * ncb
    int nlev;                                       // levels of the radix-16 hierarchy (level 0 = tiles)
    int lev_n[WF_MAX_LEV];                          // groups per column block at each level (lev_n[0] = ntr)
    int lev_off[WF_MAX_LEV];                        // offset of level l in the per-column-block arrays
    int per_cb;                                     // entries per column block over all levels
};

struct WfScratch {                                  // carved from the caller's scratch
    int* ticket;
    uint4* pay;                                     // [ncb][per_cb][2*CB] LL words: tot[l][g] of column block cb at cb*per_cb + lev_off[l] + g
};

// A published double travels as ONE 16-byte word {lo, 1, hi, 1} (the "LL" idea of NCCL's low-latency protocol, as in
// dppo_comm.cu): 8-byte halves are single-copy atomic, so a reader that sees the tag in both halves has the value -
// no flag array, no fence on either side, one L2 round trip per look.  The payload area is cleared per call.
__device__ __forceinline__ void wf_publish(uint4* p, double x) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(x);
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"((unsigned)b), "r"(1u), "r"((unsigned)(b >> 32)), "r"(1u)
                 : "memory");
}
__device__ __forceinline__ double wf_fetch(const uint4* p) {
    uint4 v;
    do {
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    } while (v.y != 1u || v.w != 1u);
    return __longlong_as_double((long long)((unsigned long long)v.x | ((unsigned long long)v.z << 32)));
}

// dynamic smem: float tile[R*CB] | float p[CB], M0[CB] | double g1[G*CB], g2[G*CB] | double base[2*CB], agg[2*CB] |
//               union { double look[WF_LOOK][2*CB];  float nf[R] }
__global__ void __launch_bounds__(WF_THREADS, 5)
wf_fused_kernel(const float* __restrict__ obs, float* __restrict__ out, const WfGeom q, float* __restrict__ n,
                float* __restrict__ mean, float* __restrict__ mean_diff, float* __restrict__ var, const WfScratch w,
                const bool vec, const double* __restrict__ totals, const int rank, const int world) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_ticket;
    float* s = reinterpret_cast<float*>(smem_raw);
    float* s_p = s + q.R * q.CB;
    float* s_m0 = s_p + WF_CB;
    double* s_g1 = reinterpret_cast<double*>(s_m0 + WF_CB);
    double* s_g2 = s_g1 + q.G * q.CB;
    double* s_base = s_g2 + q.G * q.CB;
    double* s_agg = s_base + 2 * WF_CB;
    double* s_look = s_agg + 2 * WF_CB;
    float* s_nf = reinterpret_cast<float*>(s_look);   // the look-back staging is dead when the table is built

    if (threadIdx.x == 0) s_ticket = atomicAdd(w.ticket, 1);
    __syncthreads();
    const int rt = s_ticket / q.ncb, cb = s_ticket - rt * q.ncb;     // row tile, column block
    const int col0 = cb * q.CB, cw = min(q.CB, q.O - col0);           // this block's columns
    const int V = 2 * q.CB;                                           // payload doubles per tile
    const int64_t row0 = (int64_t)rt * q.R;
    const int rows = (int)min((int64_t)q.R, q.T - row0);
    const float n0f = n[col0];                                        // read before this tile publishes anything:
    for (int c = threadIdx.x; c < cw; c += WF_THREADS) {              // the state is rewritten only by the LAST row
        s_p[c] = mean[col0 + c];                                      // tile, which waits (transitively) for every
        s_m0[c] = mean_diff[col0 + c];                                // other tile's sums
    }
    // ---- tile -> shared memory (the only read of obs) ----
    if (cw == q.O) {
        const int nel = rows * q.O;
        const int64_t off = row0 * q.O;
        if (vec && (nel & 3) == 0) {
            const float4* g4 = reinterpret_cast<const float4*>(obs + off);
            float4* s4 = reinterpret_cast<float4*>(s);
            for (int i = threadIdx.x; i < nel / 4; i += WF_THREADS) s4[i] = __ldcs(&g4[i]);
        } else {
            for (int i = threadIdx.x; i < nel; i += WF_THREADS) s[i] = __ldcs(obs + off + i);
        }
    } else {
        for (int i = threadIdx.x; i < rows * cw; i += WF_THREADS) {
            const int r = i / cw, c = i - r * cw;
            s[r * q.CB + c] = __ldcs(obs + (row0 + r) * q.O + col0 + c);
        }
    }
    const int ld = (cw == q.O) ? q.O : q.CB;                          // row stride of the tile in shared memory
    __syncthreads();
    // ---- group-local sums: item (g,c) walks rows [g*RPG, min((g+1)*RPG, rows)) of column c ----
    for (int it = threadIdx.x; it < q.G * cw; it += WF_THREADS) {
        const int g = it / cw, c = it - g * cw;
        const float p = s_p[c];
        double a1 = 0.0, a2 = 0.0;
        const int r0 = g * q.RPG, r1 = min(r0 + q.RPG, rows);
#pragma unroll 4
        for (int r = r0; r < r1; ++r) {
            const double d = (double)__fsub_rn(s[r * ld + c], p);
            a1 += d;
            a2 = fma(d, d, a2);
        }
        s_g1[g * q.CB + c] = a1;
        s_g2[g * q.CB + c] = a2;
    }
    __syncthreads();
    // ---- tile sums (fixed order over the groups), exclusive prefix of the groups in place, publish ----
    const size_t cb0 = (size_t)cb * q.per_cb;                         // this column block's slice of pay[]
    if (threadIdx.x < q.CB) {                                         // every word of the payload is published (readers
        const int c = threadIdx.x;                                    // poll all 2*CB of them): zeros beyond this block's columns
        double r1 = 0.0, r2 = 0.0;
        for (int g = 0; g < q.G && c < cw; ++g) {
            const double x1 = s_g1[g * q.CB + c], x2 = s_g2[g * q.CB + c];
            s_g1[g * q.CB + c] = r1; s_g2[g * q.CB + c] = r2;
            r1 += x1; r2 += x2;
        }
        s_agg[c] = r1; s_agg[q.CB + c] = r2;
        wf_publish(&w.pay[(cb0 + rt) * V + c], r1);
        wf_publish(&w.pay[(cb0 + rt) * V + q.CB + c], r2);
    }
    __syncthreads();
    // ---- sums of all rows above this tile.  While this tile is the LAST of its group at a level it handles that
    // level on its own and publishes the group's sums at once (a group sum needs nothing from outside the group); all
    // remaining levels are then fetched together: one L2 round trip for a tile that completes nothing (15 of 16). ----
    double own = threadIdx.x < V ? s_agg[threadIdx.x] : 0.0;
    double lev_sum[WF_MAX_LEV];
#pragma unroll
    for (int l = 0; l < WF_MAX_LEV; ++l) lev_sum[l] = 0.0;
    int lfirst = 0;                                                   // first level not yet summed
    {
        int idx = rt;
#pragma unroll
        for (int l = 0; l < WF_MAX_LEV; ++l) {
            if (l + 1 >= q.nlev || idx % WF_L1 != WF_L1 - 1) break;   // I complete a group of level l+1
            const size_t base = cb0 + q.lev_off[l] + (idx - (WF_L1 - 1));
            for (int e = threadIdx.x; e < (WF_L1 - 1) * V; e += WF_THREADS) s_look[e] = wf_fetch(&w.pay[base * V + e]);
            __syncthreads();
            if (threadIdx.x < V) {
                double sum = 0.0;
                for (int j = 0; j < WF_L1 - 1; ++j) sum += s_look[j * V + threadIdx.x];
                lev_sum[l] = sum;
                own = sum + own;                                      // the 15 before me + mine
                wf_publish(&w.pay[(cb0 + q.lev_off[l + 1] + idx / WF_L1) * V + threadIdx.x], own);
            }
            __syncthreads();
            idx /= WF_L1;
            lfirst = l + 1;
        }
    }
    {
        // the remaining levels: payload j of the combined list belongs to level L(j); staged WF_LOOK at a time
        int cnt[WF_MAX_LEV], start[WF_MAX_LEV], total = 0;
        size_t base[WF_MAX_LEV];
        int idx = rt;
#pragma unroll
        for (int l = 0; l < WF_MAX_LEV; ++l) {
            const int dgt = idx % WF_L1;
            cnt[l] = (l >= lfirst && l < q.nlev) ? dgt : 0;
            base[l] = cb0 + (l < q.nlev ? q.lev_off[l] : 0) + (idx - dgt);
            start[l] = total;
            total += cnt[l];
            idx /= WF_L1;
        }
        for (int p0 = 0; p0 < total; p0 += WF_LOOK) {
            const int np = min(WF_LOOK, total - p0);
            for (int e = threadIdx.x; e < np * V; e += WF_THREADS) {
                const int j = p0 + e / V, v = e % V;
                int l = 0;
#pragma unroll
                for (int k = 1; k < WF_MAX_LEV; ++k) if (j >= start[k] && cnt[k] > 0) l = k;
                s_look[e] = wf_fetch(&w.pay[(base[l] + (j - start[l])) * V + v]);
            }
            __syncthreads();
            if (threadIdx.x < V) {
#pragma unroll
                for (int l = 0; l < WF_MAX_LEV; ++l) {                 // within a level in order; the levels are combined below
                    const int j0 = max(start[l], p0), j1 = min(start[l] + cnt[l], p0 + np);
                    for (int j = j0; j < j1; ++j) lev_sum[l] += s_look[(j - p0) * V + threadIdx.x];
                }
            }
            __syncthreads();
        }
    }
    if (threadIdx.x < V) {
        const int v = threadIdx.x;
        double above = 0.0;                                           // highest level first: a fixed order
#pragma unroll
        for (int l = WF_MAX_LEV - 1; l >= 0; --l) above += lev_sum[l];
        // statistics shared by several workers: the rollouts of the workers before this one (rank order) come first
        double prior = 0.0;
        if (totals != nullptr) {
            const int c = v < q.CB ? v : v - q.CB;
            if (c < cw)
                for (int r = 0; r < rank; ++r)
                    prior += totals[(size_t)r * (1 + 2 * q.O) + 1 + (v < q.CB ? 0 : q.O) + col0 + c];
        }
        s_base[v] = prior + above;
    }
    __syncthreads();
    // ---- per-row 1/n_t in fp32, once per row instead of once per element ----
    double n0 = (double)n0f, rows_all = (double)q.T;               // rows before this rollout / rows of all workers
    if (totals != nullptr) {
        rows_all = 0.0;
        for (int r = 0; r < world; ++r) {
            const double tr = totals[(size_t)r * (1 + 2 * q.O)];
            if (r < rank) n0 += tr;
            rows_all += tr;
        }
    }
    for (int r = threadIdx.x; r < rows; r += WF_THREADS)
        s_nf[r] = __frcp_rn((float)(n0 + (double)(row0 + r + 1)));   // n_t < 2^24 is exact in fp32
    __syncthreads();
    // ---- the new state, by the tile that holds the last row of this column block ----
    if (rt == q.ntr - 1 && threadIdx.x < cw) {
        const int c = threadIdx.x;
        double t1 = s_base[c] + s_agg[c], t2 = s_base[q.CB + c] + s_agg[q.CB + c];
        if (totals != nullptr) {                                     // every worker's rows, in rank order: the same
            t1 = 0.0; t2 = 0.0;                                      // state on every replica
            for (int r = 0; r < world; ++r) {
                t1 += totals[(size_t)r * (1 + 2 * q.O) + 1 + col0 + c];
                t2 += totals[(size_t)r * (1 + 2 * q.O) + 1 + q.O + col0 + c];
            }
        }
        const double n1 = (double)n0f + rows_all, inv = 1.0 / n1;
        const float nn = (float)n1;
        const float md = (float)((double)s_m0[c] + t2 - t1 * t1 * inv);
        n[col0 + c] = nn;
        mean[col0 + c] = (float)((double)s_p[c] + t1 * inv);
        mean_diff[col0 + c] = md;
        var[col0 + c] = fmaxf(__fdiv_rn(md, nn), 1e-2f);             // model.py:75
    }
    // ---- per-row statistics + normalise, in place in shared memory ----
    // The state at the first row of a group comes from the fp64 prefix sums; from there the group's <= 33 rows follow
    // the reference's own fp32 recurrence (model.py:71-75: mean += (x-mean)/n; M2 += (x-mean_old)(x-mean)), so no
    // fp64 arithmetic or conversion is spent per element and no error accumulates beyond a group.
    for (int it = threadIdx.x; it < q.G * cw; it += WF_THREADS) {
        const int g = it / cw, c = it - g * cw;
        const int r0 = g * q.RPG, r1 = min(r0 + q.RPG, rows);
        if (r0 >= r1) continue;
        const double ng = n0 + (double)(row0 + r0);                  // rows before the group
        float mean_f = s_p[c], m2_f = s_m0[c];
        if (ng > 0.0) {
            const double a1 = s_base[c] + s_g1[g * q.CB + c], a2 = s_base[q.CB + c] + s_g2[g * q.CB + c];
            const double t = a1 / ng;
            mean_f = (float)((double)s_p[c] + t);
            m2_f = (float)((double)s_m0[c] + a2 - a1 * t);
        }
#pragma unroll 4
        for (int r = r0; r < r1; ++r) {
            const float x = s[r * ld + c], inv = s_nf[r];
            const float d = __fsub_rn(x, mean_f);                    // x - mean_old
            mean_f = fmaf(d, inv, mean_f);                           // model.py:73
            const float e = __fsub_rn(x, mean_f);                    // x - mean
            m2_f = fmaf(d, e, m2_f);                                 // model.py:74
            // var = max(M2/n, 1e-2), xhat = (x-mean)/sqrt(var) (model.py:75-80) as multiply-by-reciprocal and
            // rsqrt.approx (2 ulp): var >= 1e-2, so no denormal handling is needed
            const float var_t = fmaxf(m2_f * inv, 1e-2f);
            float rs;
            asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(var_t));
            s[r * ld + c] = fminf(fmaxf(e * rs, -5.f), 5.f);
        }
    }
    __syncthreads();
    // ---- stream the tile out (the only write) ----
    if (cw == q.O) {
        const int nel = rows * q.O;
        const int64_t off = row0 * q.O;
        if (vec && (nel & 3) == 0) {
            float4* o4 = reinterpret_cast<float4*>(out + off);
            const float4* s4 = reinterpret_cast<const float4*>(s);
            for (int i = threadIdx.x; i < nel / 4; i += WF_THREADS) __stcs(&o4[i], s4[i]);
        } else {
            for (int i = threadIdx.x; i < nel; i += WF_THREADS) __stcs(out + off + i, s[i]);
        }
    } else {
        for (int i = threadIdx.x; i < rows * cw; i += WF_THREADS) {
            const int r = i / cw, c = i - r * cw;
            __stcs(out + (row0 + r) * q.O + col0 + c, s[r * q.CB + c]);
        }
    }
}

// ---- column sums of a whole rollout relative to the current mean (the worker's contribution to statistics that
// several workers share): per-tile sums, then one CTA per column adds them in a fixed order ----
__global__ void __launch_bounds__(WF_THREADS)
wf_tile_sums_kernel(const float* __restrict__ obs, int64_t T, int O, int R, const float* __restrict__ mean,
                    double* __restrict__ part, int ntiles) {
    __shared__ double sh1[WF_THREADS], sh2[WF_THREADS];
    const int tile = blockIdx.x, c = blockIdx.y * 32 + (threadIdx.x & 31), g = threadIdx.x >> 5;   // 8 row groups x 32 columns
    const int64_t row0 = (int64_t)tile * R;
    const int rows = (int)min((int64_t)R, T - row0);
    double a1 = 0.0, a2 = 0.0;
    if (c < O) {
        const float p = mean[c];
        for (int r = g; r < rows; r += WF_THREADS / 32) {
            const double d = (double)__fsub_rn(obs[(row0 + r) * O + c], p);
            a1 += d;
            a2 = fma(d, d, a2);
        }
    }
    sh1[threadIdx.x] = a1; sh2[threadIdx.x] = a2;
    __syncthreads();
    if (g == 0 && c < O) {
        for (int k = 1; k < WF_THREADS / 32; ++k) { a1 += sh1[k * 32 + (threadIdx.x & 31)]; a2 += sh2[k * 32 + (threadIdx.x & 31)]; }
        part[(size_t)c * ntiles + tile] = a1;
        part[(size_t)(O + c) * ntiles + tile] = a2;
    }
}
__global__ void __launch_bounds__(WF_THREADS)
wf_totals_kernel(const double* __restrict__ part, int64_t T, int O, int ntiles, double* __restrict__ totals) {
    __shared__ double sh[WF_THREADS];
    const int v = blockIdx.x;                                         // 0..2*O-1
    const double* a = part + (size_t)v * ntiles;
    const int L = (ntiles + WF_THREADS - 1) / WF_THREADS;
    const int t0 = min(threadIdx.x * L, ntiles), t1 = min(t0 + L, ntiles);
    double sum = 0.0;
    for (int t = t0; t < t1; ++t) sum += a[t];
    sh[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        double run = 0.0;
        for (int i = 0; i < WF_THREADS; ++i) run += sh[i];
        totals[1 + v] = run;
        if (v == 0) totals[0] = (double)T;
    }
}

WfGeom wf_geom(int64_t T, int O) {
    WfGeom q;
    q.O = O; q.T = T;
    q.CB = O <= WF_CB ? O : WF_CB;
    q.ncb = (O + q.CB - 1) / q.CB;
    // G row groups of RPG rows; RPG = 33 makes consecutive groups start 33*CB = CB (mod 32) banks apart, so a warp
    // whose 32 (group, column) items span two or more groups reads 32 different banks (RPG = 32 put every group on
    // the same banks: half of all shared-memory wavefronts were conflicts)
    q.G = WF_THREADS / q.CB;
    if (q.G < 1) q.G = 1;
    q.RPG = 33;
    while (q.RPG > 1 && (int64_t)q.G * q.RPG > WF_MAX_ROWS) q.RPG -= 2;
    q.R = (q.G * q.RPG) & ~3;                       // multiple of 4 rows keeps whole-row tiles 16-byte aligned
    if (q.R < 4) q.R = 4;
    if ((int64_t)q.R > ((T + 3) & ~(int64_t)3)) {   // short rollouts: one tile, fewer rows per group
        q.R = (int)((T + 3) & ~(int64_t)3);
        q.RPG = (q.R + q.G - 1) / q.G;
    }
    q.ntr = (int)ceil_div64(T, q.R);
    q.nlev = 0;
    int n = q.ntr, off = 0;
    for (int l = 0; l < WF_MAX_LEV; ++l) { q.lev_n[l] = 0; q.lev_off[l] = 0; }
    while (q.nlev < WF_MAX_LEV) {
        q.lev_n[q.nlev] = n;
        q.lev_off[q.nlev] = off;
        off += n;
        ++q.nlev;
        if (n <= WF_L1) break;                      // every tile's group number at this level is < 16: the top level
        n = (n + WF_L1 - 1) / WF_L1;                // groups one level up (the last may be partial: never read)
    }
    q.per_cb = off;
    return q;
}

size_t wf_smem_bytes(const WfGeom& q) {
    const size_t look = sizeof(double) * WF_LOOK * 2 * q.CB;
    const size_t tables = sizeof(float) * (size_t)q.R;
    return sizeof(float) * ((size_t)q.R * q.CB + 2 * WF_CB) + sizeof(double) * (2 * (size_t)q.G * q.CB + 4 * WF_CB) +
           (look > tables ? look : tables) + 16;
}

size_t wf_pay_bytes(const WfGeom& q) { return sizeof(uint4) * 2 * q.CB * (size_t)q.ncb * (size_t)q.per_cb; }

}  // namespace

// scratch: int ticket | LL payload words (see WfScratch), cleared per call
extern "C" size_t dppo_welford_scratch_bytes(int64_t T, int O) {
    if (T <= 0 || O <= 0) return 16;
    const WfGeom q = wf_geom(T, O);
    return 16 + wf_pay_bytes(q);
}

extern "C" size_t dppo_welford_totals_scratch_bytes(int64_t T, int O) {
    if (T <= 0 || O <= 0) return 16;
    return sizeof(double) * 2 * (size_t)O * (size_t)ceil_div64(T, 512);
}

extern "C" int dppo_welford_totals(const float* obs, int64_t T, int O, const float* mean, double* totals,
                                   void* scratch, size_t scratch_bytes, void* stream_) {
    if (T <= 0 || O <= 0 || !obs || !mean || !totals || !scratch) return DPPO_E_BADARG;
    if (O > 4096 || ceil_div64(T, 512) > 0x7fffffff) return DPPO_E_TOOLARGE;
    if (scratch_bytes < dppo_welford_totals_scratch_bytes(T, O)) return DPPO_E_SCRATCH;
    if (!aligned16(scratch)) return DPPO_E_ALIGN;
    const int ntiles = (int)ceil_div64(T, 512);
    cudaStream_t stream = (cudaStream_t)stream_;
    wf_tile_sums_kernel<<<dim3(ntiles, ceil_div(O, 32)), WF_THREADS, 0, stream>>>(obs, T, O, 512, mean, (double*)scratch, ntiles);
    DPPO_LAUNCH_CHECK();
    wf_totals_kernel<<<2 * O, WF_THREADS, 0, stream>>>((const double*)scratch, T, O, ntiles, totals);
    DPPO_LAUNCH_CHECK();
    return 0;
}

static int welford_launch(const float* obs, float* out, int64_t T, int O, float* n, float* mean, float* mean_diff,
                          float* var, const double* totals, int rank, int world, void* scratch, size_t scratch_bytes,
                          void* stream_);

extern "C" int dppo_welford(const float* obs, float* out, int64_t T, int O, float* n, float* mean, float* mean_diff,
                            float* var, void* scratch, size_t scratch_bytes, void* stream_) {
    return welford_launch(obs, out, T, O, n, mean, mean_diff, var, nullptr, 0, 1, scratch, scratch_bytes, stream_);
}

extern "C" int dppo_welford_shared(const float* obs, float* out, int64_t T, int O, float* n, float* mean,
                                   float* mean_diff, float* var, const double* totals_all, int rank, int world,
                                   void* scratch, size_t scratch_bytes, void* stream_) {
    if (!totals_all || world < 1 || rank < 0 || rank >= world) return DPPO_E_BADARG;
    return welford_launch(obs, out, T, O, n, mean, mean_diff, var, totals_all, rank, world, scratch, scratch_bytes, stream_);
}

static int welford_launch(const float* obs, float* out, int64_t T, int O, float* n, float* mean, float* mean_diff,
                          float* var, const double* totals, int rank, int world, void* scratch, size_t scratch_bytes,
                          void* stream_) {
    if (T < 0 || O <= 0) return DPPO_E_BADARG;
    if (T == 0) return 0;
    if (!obs || !out || !n || !mean || !mean_diff || !var || !scratch) return DPPO_E_BADARG;
    if (O > 4096) return DPPO_E_TOOLARGE;
    const WfGeom q = wf_geom(T, O);
    if ((int64_t)q.ntr * q.ncb > 0x3fffffff) return DPPO_E_TOOLARGE;
    if (scratch_bytes < dppo_welford_scratch_bytes(T, O)) return DPPO_E_SCRATCH;
    if (!aligned16(scratch)) return DPPO_E_ALIGN;
    cudaStream_t stream = (cudaStream_t)stream_;
    WfScratch w;
    w.ticket = (int*)scratch;
    w.pay = (uint4*)((char*)scratch + 16);
    const bool vec = aligned16(obs) && aligned16(out);
    const size_t smem = wf_smem_bytes(q);
    static size_t smem_set[64] = {0};                             // per device
    int dev = 0;
    DPPO_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return DPPO_E_BADARG;
    if (smem > smem_set[dev]) {
        DPPO_CUDA_TRY(cudaFuncSetAttribute(wf_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set[dev] = smem;
    }
    DPPO_CUDA_TRY(cudaMemsetAsync(scratch, 0, 16 + wf_pay_bytes(q), stream));      // ticket + the LL tags
    wf_fused_kernel<<<q.ntr * q.ncb, WF_THREADS, smem, stream>>>(obs, out, q, n, mean, mean_diff, var, w, vec, totals,
                                                                  rank, world);
    DPPO_LAUNCH_CHECK();
    return 0;
}
