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
#ifndef WF_RPG_V
#define WF_RPG_V 33
#define WF_UNROLL_V 11
#define WF_MINB_V 5
#endif
constexpr int WF_UNROLL = WF_UNROLL_V;
constexpr int WF_RPG = WF_RPG_V;                          // rows per group of a full tile (the unrolled case of both passes)
constexpr int WF_MAX_ROWS = 256 * WF_RPG;                   // bounds the per-row reciprocal table (256 groups x 33 rows, O = 1)
constexpr int WF_INFL = 4;                          // look-back loads a thread keeps in flight (4 parts x 4 >= 15)

struct WfGeom {
    int O, CB, ncb;                                 // columns, columns per block (last block may be narrower), blocks
    int R, G, RPG;                                  // rows per tile, row groups per tile, rows per group
    int64_t T;
    int ntr;                                        // row tiles; tiles in all = ntr * ncb
    int nlev;                                       // levels of the radix-16 hierarchy (level 0 = tiles)
    int lev_n[WF_MAX_LEV];                          // groups per column block at each level (lev_n[0] = ntr)
    int lev_off[WF_MAX_LEV];                        // offset of level l in the per-column-block arrays
    int per_cb;                                     // entries per column block over all levels
    unsigned magic;                                 // floor(2^32 / CB) + 1: it / CB == umulhi(it, magic) for it < 2^32 / CB
    int nblk;                                       // ceil(2*CB / 8): blocks of 8 payload words (look-back slots)
    // shared memory carve-up (bytes from the start of the dynamic region); the tile is at 0
    int o_p, o_m0, o_g1, o_g2, o_agg, o_base, o_lsum, o_part, o_nf, o_ctl, smem;
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
__device__ __forceinline__ uint4 wf_look(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ bool wf_there(const uint4& v) { return v.y == 1u && v.w == 1u; }
__device__ __forceinline__ double wf_value(const uint4& v) {
    return __longlong_as_double((long long)((unsigned long long)v.x | ((unsigned long long)v.z << 32)));
}

// Sums of lists of consecutive payloads.  List u = payloads [lb[u], lb[u] + lc[u]) of V words each, lc <= 15.  A warp
// covers 8 words x 4 parts of one list: lane = part*8 + word, part h adds entries 4h..4h+3 in index order with all its
// loads in flight at once (one L2 round trip for everything that is already published, which is the normal case:
// tickets are handed out in order; 8 lanes read 128 contiguous bytes), and two butterfly steps add the parts as
// (p0+p1)+(p2+p3) - a fixed tree, so the result does not depend on timing.   part[u*V + word].
__device__ __forceinline__ void wf_gather(const uint4* __restrict__ pay, const long long* lb, const int* lc, const int nlists,
                                          const int V, const int nblk, double* part) {
    const int slots = nlists * nblk * 32;                             // lists x blocks of 8 words x 4 parts: whole warps
    const unsigned inv = 65536u / (unsigned)nblk + 1u;                // blk / nblk for blk < 6 * 8
    for (int e = threadIdx.x; e < slots; e += WF_THREADS) {
        const int h = (e >> 3) & 3, blk = e >> 5;
        const int u = (int)(((unsigned)blk * inv) >> 16);
        const int v = ((blk - u * nblk) << 3) + (e & 7);
        const int n = v < V ? min(lc[u] - h * WF_INFL, WF_INFL) : 0;
        double sum = 0.0;
        const uint4* p = pay + (lb[u] + h * WF_INFL) * V + v;
        uint4 wv[WF_INFL];
#pragma unroll
        for (int k = 0; k < WF_INFL; ++k)
            if (k < n) wv[k] = wf_look(p + k * V);
#pragma unroll
        for (int k = 0; k < WF_INFL; ++k) {                           // polling keeps the warp converged (the butterfly
            bool need = k < n && !wf_there(wv[k]);                    // below wants all 32 lanes)
#ifdef WF_X_NO_WAIT
            need = false;
#endif
            while (__any_sync(0xffffffffu, need))
                if (need) {
                    wv[k] = wf_look(p + k * V);
                    need = !wf_there(wv[k]);
                }
            if (k < n) sum += wf_value(wv[k]);
        }
        sum += __shfl_xor_sync(0xffffffffu, sum, 8);
        sum += __shfl_xor_sync(0xffffffffu, sum, 16);
        if (h == 0 && v < V) part[u * V + v] = sum;
    }
}

// Experiment builds (tools/build_variant.py --file dppo_welford.cu <name> -D...; never the shipped library; the
// WF_X_* ones give wrong results and exist for timing only - profiles/r02_k4_summary.md has the numbers):
//   DPPO_K4_TIMING      phase clocks      WF_X_COPY_ONLY=1   tile in, tile out, nothing else
//   WF_X_NO_LOOKBACK    no look-back      WF_X_NO_WAIT       look-back loads without waiting for the tags
//   WF_X_LEVEL0_ONLY    look-back of level 0 only
#ifdef DPPO_K4_TIMING
// phase clocks of an experiment build (tools/build_variant.py --file dppo_welford.cu k4timing -DDPPO_K4_TIMING):
// thread 0 of every CTA adds the nanoseconds between marks
__device__ unsigned long long wf_phase_ns[8];
__device__ __forceinline__ unsigned long long wf_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define WF_MARK(i) do { if (threadIdx.x == 0) { const unsigned long long t_ = wf_now(); atomicAdd(&wf_phase_ns[i], t_ - t_mark); t_mark = t_; } } while (0)
#else
#define WF_MARK(i) do { } while (0)
#endif

// it / d for it < 2^32 / d with magic = floor(2^32 / d) + 1 (d > 1)
__device__ __forceinline__ int wf_div(int it, int d, unsigned magic) { return d == 1 ? it : (int)__umulhi((unsigned)it, magic); }

// dynamic smem: float tile[R*CB] | float p[CB], M0[CB] | double g1[G*CB], g2[G*CB], agg[2*CB], base[2*CB],
//               lsum[nlev][2*CB], part[nlev][2*CB] | float nf[R] | long long lb[WF_MAX_LEV]; int lc[WF_MAX_LEV]; mbarrier
__global__ void __launch_bounds__(WF_THREADS, WF_MINB_V)
wf_fused_kernel(const float* __restrict__ obs, float* __restrict__ out, const WfGeom q, float* __restrict__ n,
                float* __restrict__ mean, float* __restrict__ mean_diff, float* __restrict__ var, const WfScratch w,
                const bool vec, const double* __restrict__ totals, const int rank, const int world) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ int s_ticket;
    float* s = reinterpret_cast<float*>(smem_raw);
    float* s_p = reinterpret_cast<float*>(smem_raw + q.o_p);
    float* s_m0 = reinterpret_cast<float*>(smem_raw + q.o_m0);
    double* s_g1 = reinterpret_cast<double*>(smem_raw + q.o_g1);
    double* s_g2 = reinterpret_cast<double*>(smem_raw + q.o_g2);
    double* s_agg = reinterpret_cast<double*>(smem_raw + q.o_agg);
    double* s_base = reinterpret_cast<double*>(smem_raw + q.o_base);
    double* s_lsum = reinterpret_cast<double*>(smem_raw + q.o_lsum);
    double* s_part = reinterpret_cast<double*>(smem_raw + q.o_part);
    float* s_nf = reinterpret_cast<float*>(smem_raw + q.o_nf);
    long long* s_lb = reinterpret_cast<long long*>(smem_raw + q.o_ctl);
    int* s_lc = reinterpret_cast<int*>(s_lb + WF_MAX_LEV);
    unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_lb + WF_MAX_LEV + WF_MAX_LEV / 2);
    const int V = 2 * q.CB;
#ifdef DPPO_K4_TIMING
    unsigned long long t_mark = wf_now();
#endif

    if (threadIdx.x == 0) {
        s_ticket = atomicAdd(w.ticket, 1);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(s_bar)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    WF_MARK(0);                                                       // ticket
    // Tiles are handed out in ticket order, so every tile above this one has started
    const int ticket = s_ticket;
    int rt = ticket, cb = 0;                                          // row tile, column block
    if (q.ncb > 1) { rt = ticket / q.ncb; cb = ticket - rt * q.ncb; }
    const int col0 = cb * q.CB, cw = min(q.CB, q.O - col0);
    const int64_t row0 = (int64_t)rt * q.R;
    const int rows = (int)min((int64_t)q.R, q.T - row0);
    const int ld = q.CB;                                              // (CB == O for whole-row tiles)
    const long long cb0 = (long long)cb * q.per_cb;
    const unsigned magic = cw == q.CB ? q.magic : (cw > 1 ? 0xffffffffu / (unsigned)cw + 1u : 0u);
    // whole-row tiles of aligned arrays travel as ONE bulk async copy each way (no registers, no LSU instructions)
    const int nel = rows * cw;
    const bool bulk = vec && cw == q.O && (nel & 3) == 0;
    if (bulk) {
        if (threadIdx.x == 0) {
            const unsigned bar = (unsigned)__cvta_generic_to_shared(s_bar), bytes = (unsigned)nel * 4u;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                             (unsigned)__cvta_generic_to_shared(s)), "l"(obs + row0 * q.O), "r"(bytes), "r"(bar) : "memory");
        }
    } else if (cw == q.O) {
        for (int i = threadIdx.x; i < nel; i += WF_THREADS) s[i] = __ldcs(obs + row0 * q.O + i);
    } else {
        for (int i = threadIdx.x; i < nel; i += WF_THREADS) {
            const int r = wf_div(i, cw, magic), c = i - r * cw;
            s[r * ld + c] = __ldcs(obs + (row0 + r) * q.O + col0 + c);
        }
    }
    const float n0f = n[col0];                                        // read before this tile publishes anything:
    for (int c = threadIdx.x; c < cw; c += WF_THREADS) {              // the state is rewritten only by the LAST row
        s_p[c] = mean[col0 + c];                                      // tile, which waits (transitively) for every
        s_m0[c] = mean_diff[col0 + c];                                // other tile's sums
    }
    // rows of all workers before this rollout / in all (statistics shared by several workers); 1/n_t per row in fp32,
    // once per row instead of once per element
    double n0 = (double)n0f, rows_all = (double)q.T;
    if (totals != nullptr) {
        rows_all = 0.0;
        for (int r = 0; r < world; ++r) {
            const double tr = totals[(size_t)r * (1 + 2 * q.O)];
            if (r < rank) n0 += tr;
            rows_all += tr;
        }
    }
    {
        const double nb = n0 + (double)row0 + 1.0;
        for (int r = threadIdx.x; r < rows; r += WF_THREADS) s_nf[r] = __frcp_rn((float)(nb + (double)r));
    }
    // the lists of the look-back (digit l of the tile number counts the level-l groups before mine inside my
    // level-(l+1) group); a tile whose low digits are 15 completes those groups and sums them itself below
    int lfirst = 0;
#pragma unroll
    for (int l = 0; l + 1 < WF_MAX_LEV; ++l)
        if (lfirst == l && l + 1 < q.nlev && ((rt >> (4 * l)) & (WF_L1 - 1)) == WF_L1 - 1) lfirst = l + 1;
    if (threadIdx.x < WF_MAX_LEV) {
        const int l = threadIdx.x, idx = rt >> (4 * l), dgt = idx & (WF_L1 - 1);
        s_lb[l] = cb0 + (l < q.nlev ? q.lev_off[l] : 0) + (idx - dgt);
        s_lc[l] = l < q.nlev ? dgt : 0;                               // (levels below lfirst: 15, used by the completion)
    }
    if (bulk) {
        unsigned ok = 0;
        const unsigned bar = (unsigned)__cvta_generic_to_shared(s_bar);
        while (!ok)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(bar), "r"(0u) : "memory");
    }
    __syncthreads();
    WF_MARK(1);                                                       // tile in shared memory

#ifdef WF_X_COPY_ONLY                                                 // timing experiment: load + store only
    if (bulk) {
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + row0 * q.O),
                         "r"((unsigned)__cvta_generic_to_shared(s)), "r"((unsigned)nel * 4u) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        return;
    }
#endif
    // ---- pass A: group-local sums.  Item (g,c) walks rows [g*RPG, min((g+1)*RPG, rows)) of column c ----
    for (int it = threadIdx.x; it < q.G * cw; it += WF_THREADS) {
        const int g = wf_div(it, cw, magic), c = it - g * cw;
        const float p = s_p[c];
        double a1 = 0.0, a2 = 0.0;
        const int r0 = g * q.RPG, r1 = min(r0 + q.RPG, rows);
        const float* ps = s + r0 * ld + c;
        if (r1 - r0 == WF_RPG) {
#pragma unroll WF_UNROLL
            for (int k = 0; k < WF_RPG; ++k) {
                const double d = (double)__fsub_rn(ps[k * ld], p);
                a1 += d;
                a2 = fma(d, d, a2);
            }
        } else {
            for (int k = 0; k < r1 - r0; ++k) {
                const double d = (double)__fsub_rn(ps[k * ld], p);
                a1 += d;
                a2 = fma(d, d, a2);
            }
        }
        s_g1[g * q.CB + c] = a1;
        s_g2[g * q.CB + c] = a2;
    }
    __syncthreads();
    WF_MARK(2);                                                       // pass A
    // tile sums (fixed order over the groups), exclusive prefix of the groups in place, publish
    double own = 0.0;
    if (threadIdx.x < V) {                                            // every word of the payload is published (readers
        const int v = threadIdx.x, c = v < q.CB ? v : v - q.CB;       // poll all 2*CB of them): zeros beyond this block's columns
        double* a = v < q.CB ? s_g1 : s_g2;
        double run = 0.0;
        if (c < cw)
            for (int g = 0; g < q.G; ++g) {
                const double x1 = a[g * q.CB + c];
                a[g * q.CB + c] = run;
                run += x1;
            }
        s_agg[v] = run;
        own = run;
        wf_publish(&w.pay[(cb0 + rt) * V + v], run);
    }
    // While this tile is the LAST of its group at a level it adds up that level now and publishes the group's sums
    // at once (a group sum needs nothing from outside the group)
    for (int l = 0; l < lfirst; ++l) {
        __syncthreads();
        wf_gather(w.pay, s_lb + l, s_lc + l, 1, V, q.nblk, s_part);
        __syncthreads();
        if (threadIdx.x < V) {
            const double sum = s_part[threadIdx.x];
            s_lsum[l * V + threadIdx.x] = sum;
            own = sum + own;                                          // the 15 before me + mine
            wf_publish(&w.pay[(cb0 + q.lev_off[l + 1] + (rt >> (4 * (l + 1)))) * V + threadIdx.x], own);
        }
    }
    __syncthreads();
    WF_MARK(3);                                                       // published (+ completed groups)
    // ---- the sums of all rows above this tile: the remaining levels, all in flight together ----
#if defined(WF_X_LEVEL0_ONLY)
    wf_gather(w.pay + 0, s_lb + lfirst, s_lc + lfirst, lfirst == 0 ? 1 : 0, V, q.nblk, s_part);
#elif !defined(WF_X_NO_LOOKBACK)
    wf_gather(w.pay + 0, s_lb + lfirst, s_lc + lfirst, q.nlev - lfirst, V, q.nblk, s_part);
#endif
    __syncthreads();
    WF_MARK(4);                                                       // look-back
    if (threadIdx.x < V) {
        const int v = threadIdx.x;
        double above = 0.0;                                           // highest level first: a fixed order
        for (int l = q.nlev - 1; l >= lfirst; --l) above += s_part[(l - lfirst) * V + v];
        for (int l = lfirst - 1; l >= 0; --l) above += s_lsum[l * V + v];
        // statistics shared by several workers: the rollouts of the workers before this one (rank order) come first
        double prior = 0.0;
        if (totals != nullptr) {
            const int c = v < q.CB ? v : v - q.CB;
            if (c < cw)
                for (int r = 0; r < rank; ++r) prior += totals[(size_t)r * (1 + 2 * q.O) + 1 + (v < q.CB ? 0 : q.O) + col0 + c];
        }
        s_base[v] = prior + above;
    }
    __syncthreads();
    // the new state, by the tile that holds the last row of this column block
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
        var[col0 + c] = fmaxf(__fdiv_rn(md, nn), 1e-2f);               // model.py:75
    }
    // ---- pass B: per-row statistics + normalise, in place in shared memory.  The state at the first row of a group
    // comes from the fp64 prefix sums; from there the group's <= 33 rows follow the reference's own fp32 recurrence
    // (model.py:71-75: mean += (x-mean)/n; M2 += (x-mean_old)(x-mean)), so no fp64 arithmetic or conversion is spent per
    // element and no error accumulates beyond a group.
    for (int it = threadIdx.x; it < q.G * cw; it += WF_THREADS) {
        const int g = wf_div(it, cw, magic), c = it - g * cw;
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
        float* ps = s + r0 * ld + c;
        const float* pn = s_nf + r0;
        auto row = [&](int k) {
            const float xv = ps[k * ld], inv = pn[k];
            const float d = __fsub_rn(xv, mean_f);                   // x - mean_old
            mean_f = fmaf(d, inv, mean_f);                           // model.py:73
            const float e = __fsub_rn(xv, mean_f);                   // x - mean
            m2_f = fmaf(d, e, m2_f);                                 // model.py:74
            // var = max(M2/n, 1e-2), xhat = (x-mean)/sqrt(var) (model.py:75-80) as multiply-by-reciprocal and
            // rsqrt.approx (2 ulp): var >= 1e-2, so no denormal handling is needed
            const float var_t = fmaxf(m2_f * inv, 1e-2f);
            float rs;
            asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(var_t));
            ps[k * ld] = fminf(fmaxf(e * rs, -5.f), 5.f);
        };
        if (r1 - r0 == WF_RPG) {
#pragma unroll WF_UNROLL
            for (int k = 0; k < WF_RPG; ++k) row(k);
        } else {
            for (int k = 0; k < r1 - r0; ++k) row(k);
        }
    }
    // stream the tile out (the only write)
    if (bulk) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the rows were written through the generic proxy
        __syncthreads();
        WF_MARK(5);                                                   // pass B
        if (threadIdx.x == 0) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + row0 * q.O),
                         "r"((unsigned)__cvta_generic_to_shared(s)), "r"((unsigned)nel * 4u) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");   // the tile must outlive the copy's reads
        }
        WF_MARK(6);                                                   // store read out
        return;
    }
    __syncthreads();
    if (cw == q.O) {
        for (int i = threadIdx.x; i < nel; i += WF_THREADS) __stcs(out + row0 * q.O + i, s[i]);
    } else {
        for (int i = threadIdx.x; i < nel; i += WF_THREADS) {
            const int r = wf_div(i, cw, magic), c = i - r * cw;
            __stcs(out + (row0 + r) * q.O + col0 + c, s[r * ld + c]);
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
    q.RPG = WF_RPG;
    while (q.RPG > 1 && (int64_t)q.G * q.RPG > WF_MAX_ROWS) q.RPG -= 2;
    // rollouts that would not even give every SM two tiles (the benchmark's 65 536 x 17: 134 tiles) are latency bound
    // - one tile's load, sums, look-back, normalise, store in a row: shorter groups, more tiles in flight
    {
        const int64_t tiles33 = ceil_div64(T, (q.G * q.RPG) & ~3) * q.ncb;
        if (q.RPG == WF_RPG && tiles33 < 2 * (int64_t)dppo_sm_count() && T >= 4 * q.G * 17) q.RPG = 17;
    }
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
    q.magic = q.CB > 1 ? 0xffffffffu / (unsigned)q.CB + 1u : 0u;
    q.nblk = (2 * q.CB + 7) / 8;
    const int V = 2 * q.CB;
    size_t o = sizeof(float) * (((size_t)q.R * q.CB + 3) & ~(size_t)3);
    q.o_p = (int)o; o += sizeof(float) * WF_CB;
    q.o_m0 = (int)o; o += sizeof(float) * WF_CB;
    q.o_g1 = (int)o; o += sizeof(double) * q.G * q.CB;
    q.o_g2 = (int)o; o += sizeof(double) * q.G * q.CB;
    q.o_agg = (int)o; o += sizeof(double) * V;
    q.o_base = (int)o; o += sizeof(double) * V;
    q.o_lsum = (int)o; o += sizeof(double) * q.nlev * V;
    q.o_part = (int)o; o += sizeof(double) * q.nlev * V;
    q.o_nf = (int)o; o += sizeof(float) * (((size_t)q.R + 3) & ~(size_t)3);
    o = (o + 15) & ~(size_t)15;
    q.o_ctl = (int)o; o += sizeof(long long) * (WF_MAX_LEV + WF_MAX_LEV / 2) + 16;
    q.smem = (int)o;
    return q;
}

size_t wf_pay_bytes(const WfGeom& q) { return sizeof(uint4) * 2 * q.CB * (size_t)q.ncb * (size_t)q.per_cb; }

}  // namespace

#ifdef DPPO_K4_TIMING
extern "C" int dppo_welford_phase_ns(unsigned long long* out8, int reset) {
    DPPO_CUDA_TRY(cudaMemcpyFromSymbol(out8, wf_phase_ns, sizeof(unsigned long long) * 8));
    if (reset) { unsigned long long z[8] = {0}; DPPO_CUDA_TRY(cudaMemcpyToSymbol(wf_phase_ns, z, sizeof(z))); }
    return 0;
}
#endif

// scratch: int ticket | LL payload words (see WfScratch), cleared per call
extern "C" size_t dppo_welford_scratch_bytes(int64_t T, int O) {
    if (T <= 0 || O <= 0) return 16;
    return 16 + wf_pay_bytes(wf_geom(T, O));
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
    const bool vec = aligned16(obs) && aligned16(out);
    const WfGeom q = wf_geom(T, O);
    if ((int64_t)q.ntr * q.ncb > 0x3fffffff) return DPPO_E_TOOLARGE;
    if (scratch_bytes < dppo_welford_scratch_bytes(T, O)) return DPPO_E_SCRATCH;
    if (!aligned16(scratch)) return DPPO_E_ALIGN;
    cudaStream_t stream = (cudaStream_t)stream_;
    WfScratch w;
    w.ticket = (int*)scratch;
    w.pay = (uint4*)((char*)scratch + 16);
    const size_t smem = (size_t)q.smem;
    static size_t smem_set[64] = {0};                             // per device
    int dev = 0;
    DPPO_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return DPPO_E_BADARG;
    if (smem > smem_set[dev]) {
        DPPO_CUDA_TRY(cudaFuncSetAttribute(wf_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set[dev] = smem;
    }
    DPPO_CUDA_TRY(cudaMemsetAsync(scratch, 0, 16 + wf_pay_bytes(q), stream));      // ticket + the LL tags
    wf_fused_kernel<<<q.ntr * q.ncb, WF_THREADS, smem, stream>>>(obs, out, q, n, mean, mean_diff, var, w, vec, totals, rank, world);
    DPPO_LAUNCH_CHECK();
    return 0;
}
