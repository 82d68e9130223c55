// K1 - GAE / returns as a reverse segmented linear-recurrence scan.
//
// Reference: train.py:113-122 / ppo.py:124-133
//     td = r[i] + gamma*V[i+1] - V[i];  A = td + (gamma*lam)*A;  R = A + V[i]
// per rollout segment, A reset to 0 and V[i+1] := bootstrap at a segment end.
//
// Formulation: step i is the affine map  f_i(x) = td_i + a_i * x  with
// a_i = 0 at a segment end (this is how the done-mask enters: it zeroes the
// coefficient, and the zero propagates through the warp shuffles) and
// a_i = fp32(gamma*lam) elsewhere.  A_i = (f_i o f_{i+1} o ...)(0).
//
// One pass over HBM (20 B/step + 1 B flag): every thread owns 8 consecutive
// steps, composes them, a warp-shuffle suffix scan and a block scan give each
// thread its carry-in.  Tiles are independent when gamma*lam decays fast enough
// (see HALO below); otherwise they are chained right-to-left: tile order is handed
// out by an atomic ticket (a tile's right-hand neighbours always started earlier)
// and the value entering a tile is obtained from a FIXED two-level hierarchy of
// composed maps - blocks of 16 tiles, super-blocks of 16 blocks - applied in a fixed
// order, so the result does not depend on timing (fp32 composition is not
// associative; the usual decoupled look-back, which composes whatever mix of tile
// maps and resolved values it finds, is not bit-repeatable when the decay is slow).
// With its carry known, each thread REPLAYS its
// 8 steps with the reference's exact fp32 operation order (no FMA), so the
// only rounding difference to the sequential loop sits in the carries.
#include <math.h>
#include "common.cuh"

namespace {

constexpr int GAE_THREADS = 512;
constexpr int GAE_K = 8;                          // steps per thread
constexpr int GAE_TILE = GAE_THREADS * GAE_K;     // 4096 steps per CTA

struct Aff { float a, b; };                       // x -> b + a*x
struct GaeChain {                                 // chained mode: carved from the caller's scratch
    int *ticket, *flagA, *flagB, *flagG;
    float2 *mapA, *mapB;
    float* valG;
};
// (f o g)(x) = f.b + f.a*(g.b + g.a*x)
__device__ __forceinline__ Aff compose(Aff f, Aff g) {
    Aff r;
    r.a = f.a * g.a;
    r.b = fmaf(f.a, g.b, f.b);
    return r;
}

// Chained mode, scratch layout (GaeChain): tiles are numbered u = 0, 1, ... from the RIGHT end of the buffer.
//   mapA[u]  the tile's own map f_u               flagA[u]   published by every tile, depends on nothing
//   mapB[q]  f_{16q+15} o ... o f_{16q}           flagB[q]   published by the block's last tile, needs only mapA
//   valG[a]  the advantage entering tile 256a     flagG[a]   published by the last tile of super-block a-1: the
//                                                            only serial chain, one link per 256 tiles
// The value entering tile u = 256a + 16b + c is  (f_{u-1} o ... o f_{u-c}) ( mapB[16a+b-1] ( ... mapB[16a] ( valG[a] ) ) ).
// HALO > 0: the tile's window is GAE_TILE steps of which the last HALO are a read-only look-ahead:
// |gamma*lam|^HALO is below 2^-40, so the advantage entering the window from the right cannot change
// any fp32 result in the first GAE_TILE-HALO steps and the tiles become independent (no ticket, no
// look-back, no inter-CTA traffic).  HALO == 0: chained tiles with decoupled look-back (exact for
// any gamma*lam, used when the decay is too slow for a halo).
template <bool VEC, bool HALO>
__global__ void __launch_bounds__(GAE_THREADS)
gae_scan_kernel(const float* __restrict__ rew, const float* __restrict__ val, const float* __restrict__ boot,
                const uint8_t* __restrict__ seg_end, float* __restrict__ adv, float* __restrict__ ret,
                int64_t T, float g, float gl, int ntiles, int halo, const GaeChain ch) {
    __shared__ int s_tile;
    __shared__ float2 s_look[32];
    __shared__ Aff s_warp[GAE_THREADS / 32];
    __shared__ float s_carry;

    int tile;
    int64_t tile_base, out_end;
    if (HALO) {
        tile = blockIdx.x;
        tile_base = (int64_t)tile * (GAE_TILE - halo);
        out_end = tile_base + (GAE_TILE - halo);          // steps >= out_end belong to the next tile
    } else {
        if (threadIdx.x == 0) s_tile = ntiles - 1 - atomicAdd(ch.ticket, 1);   // right-most tile first
        __syncthreads();
        tile = s_tile;
        tile_base = (int64_t)tile * GAE_TILE;
        out_end = tile_base + GAE_TILE;
    }
    const int64_t base = tile_base + (int64_t)threadIdx.x * GAE_K;

    float r[GAE_K], v[GAE_K + 1], bt[GAE_K];
    uint8_t se[GAE_K];
    if (VEC && base + GAE_K <= T) {
        const float4* r4 = reinterpret_cast<const float4*>(rew + base);
        const float4* v4 = reinterpret_cast<const float4*>(val + base);
        const float4* b4 = reinterpret_cast<const float4*>(boot + base);
        float4 x0 = r4[0], x1 = r4[1], y0 = v4[0], y1 = v4[1], z0 = b4[0], z1 = b4[1];
        uint2 e = *reinterpret_cast<const uint2*>(seg_end + base);
        r[0] = x0.x; r[1] = x0.y; r[2] = x0.z; r[3] = x0.w; r[4] = x1.x; r[5] = x1.y; r[6] = x1.z; r[7] = x1.w;
        v[0] = y0.x; v[1] = y0.y; v[2] = y0.z; v[3] = y0.w; v[4] = y1.x; v[5] = y1.y; v[6] = y1.z; v[7] = y1.w;
        bt[0] = z0.x; bt[1] = z0.y; bt[2] = z0.z; bt[3] = z0.w; bt[4] = z1.x; bt[5] = z1.y; bt[6] = z1.z; bt[7] = z1.w;
#pragma unroll
        for (int k = 0; k < 4; ++k) { se[k] = (e.x >> (8 * k)) & 0xff; se[4 + k] = (e.y >> (8 * k)) & 0xff; }
    } else {
#pragma unroll
        for (int k = 0; k < GAE_K; ++k) {
            const int64_t i = base + k;
            const bool in = i < T;
            r[k] = in ? rew[i] : 0.f;
            v[k] = in ? val[i] : 0.f;
            bt[k] = in ? boot[i] : 0.f;
            se[k] = in ? seg_end[i] : (uint8_t)1;
        }
    }
    v[GAE_K] = (base + GAE_K < T) ? val[base + GAE_K] : 0.f;
#pragma unroll
    for (int k = 0; k < GAE_K; ++k)
        if (base + k >= T - 1) se[k] = 1;          // the buffer's last step always closes a segment

    // td_i with the reference's roundings: fl(fl(r + fl(g*Vnext)) - V)
    float td[GAE_K];
#pragma unroll
    for (int k = 0; k < GAE_K; ++k) {
        const float vn = se[k] ? bt[k] : v[k + 1];
        td[k] = __fsub_rn(__fadd_rn(r[k], __fmul_rn(g, vn)), v[k]);
    }
    // thread-local composition f_0 o f_1 o ... o f_7
    Aff loc;
    loc.a = 1.f; loc.b = 0.f;
#pragma unroll
    for (int k = GAE_K - 1; k >= 0; --k) {
        Aff f; f.a = se[k] ? 0.f : gl; f.b = td[k];
        loc = compose(f, loc);
    }
    // inclusive suffix scan over the lanes of a warp: inc_l = f_l o f_{l+1} o ... o f_31
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    Aff inc = loc;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        Aff o;
        o.a = __shfl_down_sync(0xffffffffu, inc.a, d);
        o.b = __shfl_down_sync(0xffffffffu, inc.b, d);
        if (lane + d < 32) inc = compose(inc, o);
    }
    if (lane == 0) s_warp[warp] = inc;
    __syncthreads();
    // composition of everything to the right of this thread inside the tile
    Aff right;                                       // lanes right of me in my warp
    right.a = __shfl_down_sync(0xffffffffu, inc.a, 1);
    right.b = __shfl_down_sync(0xffffffffu, inc.b, 1);
    if (lane == 31) { right.a = 1.f; right.b = 0.f; }
#pragma unroll
    for (int w = 0; w < GAE_THREADS / 32; ++w)
        if (w > warp) right = compose(right, s_warp[w]);

    // tile map and the value entering the tile from the right (warp 0; see GaeChain)
    if (HALO) {
        if (threadIdx.x == 0) s_carry = 0.f;              // what enters from beyond the halo has decayed away
    } else if (warp == 0) {
        const int u = ntiles - 1 - tile;
        const int sa = u >> 8, sb = (u >> 4) & 15, sc = u & 15;
        const Aff agg = compose(loc, right);              // (lane 0 holds the whole tile's map)
        if (lane == 0) {
            __stcg(&ch.mapA[u], make_float2(agg.a, agg.b));
            __threadfence();
            st_release_gpu(&ch.flagA[u], 1);
        }
        // round 0: the maps of the tiles before this one in its block, fetched side by side, composed in order
        if (lane < sc) {
            while (ld_acquire_gpu(&ch.flagA[u - sc + lane]) == 0) { }
            s_look[lane] = __ldcg(&ch.mapA[u - sc + lane]);
        }
        __syncwarp();
        Aff M; M.a = 1.f; M.b = 0.f;                      // f_{u-1} o ... o f_{u-sc}
        if (lane == 0) {
            for (int j = 0; j < sc; ++j) { Aff f; f.a = s_look[j].x; f.b = s_look[j].y; M = compose(f, M); }
            if (sc == 15) {                               // last tile of its block: publish the block's map
                const Aff Bm = compose(agg, M);
                __stcg(&ch.mapB[u >> 4], make_float2(Bm.a, Bm.b));
                __threadfence();
                st_release_gpu(&ch.flagB[u >> 4], 1);
            }
        }
        __syncwarp();
        // round 1: the maps of the blocks before this one in its super-block and the value entering the super-block
        if (lane < sb) {
            while (ld_acquire_gpu(&ch.flagB[(sa << 4) + lane]) == 0) { }
            s_look[lane] = __ldcg(&ch.mapB[(sa << 4) + lane]);
        } else if (lane == 16 && sa > 0) {
            while (ld_acquire_gpu(&ch.flagG[sa]) == 0) { }
            s_look[16] = make_float2(__ldcg(&ch.valG[sa]), 0.f);
        }
        __syncwarp();
        if (lane == 0) {
            float x = sa > 0 ? s_look[16].x : 0.f;
            for (int j = 0; j < sb; ++j) x = (s_look[j].x == 0.f) ? s_look[j].y : fmaf(s_look[j].x, x, s_look[j].y);
            s_carry = (M.a == 0.f) ? M.b : fmaf(M.a, x, M.b);
            if (sc == 15 && sb == 15) {                   // last tile of its super-block: the value entering the next one
                const Aff Bm = compose(agg, M);
                __stcg(&ch.valG[sa + 1], (Bm.a == 0.f) ? Bm.b : fmaf(Bm.a, x, Bm.b));
                __threadfence();
                st_release_gpu(&ch.flagG[sa + 1], 1);
            }
        }
    }
    __syncthreads();
    // A at the first step right of this thread
    float A = (right.a == 0.f) ? right.b : fmaf(right.a, s_carry, right.b);
    // replay with the reference's op order: A = fl(td + fl(gl*A)); R = fl(A + V)
    float oa[GAE_K], orr[GAE_K];
#pragma unroll
    for (int k = GAE_K - 1; k >= 0; --k) {
        A = se[k] ? td[k] : __fadd_rn(td[k], __fmul_rn(gl, A));
        oa[k] = A;
        orr[k] = __fadd_rn(A, v[k]);
    }
    if (HALO && base >= out_end) return;                  // look-ahead steps are not written (out_end % 8 == 0)
    if (VEC && base + GAE_K <= T) {
        float4* a4 = reinterpret_cast<float4*>(adv + base);
        float4* q4 = reinterpret_cast<float4*>(ret + base);
        a4[0] = make_float4(oa[0], oa[1], oa[2], oa[3]);
        a4[1] = make_float4(oa[4], oa[5], oa[6], oa[7]);
        q4[0] = make_float4(orr[0], orr[1], orr[2], orr[3]);
        q4[1] = make_float4(orr[4], orr[5], orr[6], orr[7]);
    } else {
#pragma unroll
        for (int k = 0; k < GAE_K; ++k)
            if (base + k < T) { adv[base + k] = oa[k]; ret[base + k] = orr[k]; }
    }
}

}  // namespace

// look-ahead length for the independent-tile mode, 0 if the decay is too slow
static int gae_halo(float gl) {
    const double a = fabs((double)gl);
    if (a == 0.0) return 8;
    if (a >= 1.0) return 0;
    const double h = ceil(40.0 * log(2.0) / -log(a));
    if (h > GAE_TILE / 2) return 0;
    return ((int)h + 7) & ~7;
}

static size_t gae_flag_bytes(int64_t ntiles) {
    const int64_t nb1 = ceil_div64(ntiles, 16), nb2 = ceil_div64(nb1, 16);
    return align_up(sizeof(int) * (size_t)(1 + ntiles + nb1 + nb2 + 1), 16);
}
extern "C" size_t dppo_gae_scratch_bytes(int64_t T) {
    if (T <= 0) return 16;
    const int64_t ntiles = ceil_div64(T, GAE_TILE), nb1 = ceil_div64(ntiles, 16), nb2 = ceil_div64(nb1, 16);
    return gae_flag_bytes(ntiles) + sizeof(float2) * (size_t)(ntiles + nb1) + align_up(sizeof(float) * (size_t)(nb2 + 1), 16);
}

extern "C" int dppo_gae_scan(const float* rewards, const float* values, const float* boot, const uint8_t* seg_end,
                             float* adv, float* ret, int64_t T, double gamma, double lam,
                             void* scratch, size_t scratch_bytes, void* stream_) {
    if (T < 0) return DPPO_E_BADARG;
    if (T == 0) return 0;
    if (!rewards || !values || !boot || !seg_end || !adv || !ret || !scratch) return DPPO_E_BADARG;
    if (ceil_div64(T, GAE_TILE / 2) > 0x7fffffff) return DPPO_E_TOOLARGE;
    if (scratch_bytes < dppo_gae_scratch_bytes(T)) return DPPO_E_SCRATCH;
    if (!aligned16(scratch)) return DPPO_E_ALIGN;
    cudaStream_t stream = (cudaStream_t)stream_;
    // gamma multiplies a 0-dim fp32 tensor, gamma*lam is a Python double product
    // that multiplies an fp32 tensor (train.py:118-119): both end up as fp32 scalars
    const float g = (float)gamma, gl = (float)(gamma * lam);
    const bool vec = aligned16(rewards) && aligned16(values) && aligned16(boot) && aligned16(adv) && aligned16(ret) &&
                     ((reinterpret_cast<uintptr_t>(seg_end) & 7u) == 0);
    const int halo = gae_halo(gl);
    GaeChain ch = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    if (halo > 0) {
        const int step = GAE_TILE - halo;
        const int ntiles = (int)ceil_div64(T, step);
        if (vec)
            gae_scan_kernel<true, true><<<ntiles, GAE_THREADS, 0, stream>>>(rewards, values, boot, seg_end, adv, ret, T,
                                                                             g, gl, ntiles, halo, ch);
        else
            gae_scan_kernel<false, true><<<ntiles, GAE_THREADS, 0, stream>>>(rewards, values, boot, seg_end, adv, ret, T,
                                                                              g, gl, ntiles, halo, ch);
        DPPO_LAUNCH_CHECK();
        return 0;
    }
    const int ntiles = (int)ceil_div64(T, GAE_TILE);
    const int nb1 = ceil_div(ntiles, 16), nb2 = ceil_div(nb1, 16);
    const size_t flag_bytes = gae_flag_bytes(ntiles);
    ch.ticket = (int*)scratch;
    ch.flagA = ch.ticket + 1;
    ch.flagB = ch.flagA + ntiles;
    ch.flagG = ch.flagB + nb1;
    ch.mapA = (float2*)((char*)scratch + flag_bytes);
    ch.mapB = ch.mapA + ntiles;
    ch.valG = (float*)(ch.mapB + nb1);
    (void)nb2;
    DPPO_CUDA_TRY(cudaMemsetAsync(scratch, 0, flag_bytes, stream));
    if (vec)
        gae_scan_kernel<true, false><<<ntiles, GAE_THREADS, 0, stream>>>(rewards, values, boot, seg_end, adv, ret, T, g,
                                                                          gl, ntiles, 0, ch);
    else
        gae_scan_kernel<false, false><<<ntiles, GAE_THREADS, 0, stream>>>(rewards, values, boot, seg_end, adv, ret, T, g,
                                                                           gl, ntiles, 0, ch);
    DPPO_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------
// dppo_gae_scan_noalloc - the same scan for callers that cannot hand us scratch.
//
// dppo_gae_scan needs a few bytes per 4096-step tile for the look-back chain, and
// it expects the caller to allocate them.  Some callers cannot: a host program
// that was given a fixed set of device buffers by someone else's allocator and
// has no allocator of its own; a control loop that must not call cudaMalloc
// because cudaMalloc synchronises the whole device and may fail late; code that
// runs after the application has frozen its memory pool.  For them the chain
// lives in the array below, which is part of the library image: every GPU that
// loads libdppo_b200.so keeps these ~13 KB resident until the process exits.
//
// Limits and rules
//   * T <= DPPO_GAE_NOALLOC_MAX_T (4 194 304 steps); longer buffers are refused
//     with DPPO_E_NOALLOC_LIMIT - use dppo_gae_scan with your own scratch.
//   * The first call on each device looks up the device address of the array
//     (cudaGetSymbolAddress, with that device current) and caches it.  Make that
//     first call outside of any stream capture.
//   * Nothing is allocated or freed, here or in anything this calls.
//   * The array is used only when the scan chains its tiles, i.e. when gamma*lam
//     is too close to 1 for independent tiles (gae_halo() == 0, |gamma*lam| >= ~0.987).
//     With a faster decay (the reference's 0.99*0.95) the tiles need no scratch at
//     all and the array is left alone.
//   * All calls on a device share the one array.  In the chained mode the ticket
//     and the tile flags, which must start at zero, are cleared on `stream` at the
//     start of every call; the carries are overwritten before they are read.  Two
//     calls that can overlap on the GPU (different streams, no dependency between
//     them) race on it - serialise them.
// ---------------------------------------------------------------------------
#define DPPO_GAE_NOALLOC_MAX_TILES ((DPPO_GAE_NOALLOC_MAX_T + GAE_TILE - 1) / GAE_TILE)
#define DPPO_GAE_NOALLOC_NB1 ((DPPO_GAE_NOALLOC_MAX_TILES + 15) / 16)
#define DPPO_GAE_NOALLOC_NB2 ((DPPO_GAE_NOALLOC_NB1 + 15) / 16)
#define DPPO_GAE_NOALLOC_FLAG_BYTES \
    ((sizeof(int) * (1 + DPPO_GAE_NOALLOC_MAX_TILES + DPPO_GAE_NOALLOC_NB1 + DPPO_GAE_NOALLOC_NB2 + 1) + 15) / 16 * 16)
#define DPPO_GAE_NOALLOC_BYTES                                                                                  \
    (DPPO_GAE_NOALLOC_FLAG_BYTES + sizeof(float2) * (DPPO_GAE_NOALLOC_MAX_TILES + DPPO_GAE_NOALLOC_NB1) + \
     (sizeof(float) * (DPPO_GAE_NOALLOC_NB2 + 1) + 15) / 16 * 16)

static __device__ __align__(256) unsigned char g_gae_noalloc_scratch[DPPO_GAE_NOALLOC_BYTES];

namespace {
constexpr int NOALLOC_MAX_DEVICES = 64;
void* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};      // device address of the array, per device
}  // namespace

extern "C" int dppo_gae_scan_noalloc(const float* rewards, const float* values, const float* boot,
                                     const uint8_t* seg_end, float* adv, float* ret, int64_t T, double gamma,
                                     double lam, void* stream) {
    if (T < 0) return DPPO_E_BADARG;
    if (T > (int64_t)DPPO_GAE_NOALLOC_MAX_T) return DPPO_E_NOALLOC_LIMIT;
    if (T == 0) return 0;
    int dev = -1;
    DPPO_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return DPPO_E_BADARG;
    if (g_noalloc_ptr[dev] == nullptr) {
        // In host code `g_gae_noalloc_scratch` is only a handle for the symbol, not an address the
        // GPU can use: ask the runtime where the array lives on THIS device.
        void* p = nullptr;
        DPPO_CUDA_TRY(cudaGetSymbolAddress(&p, g_gae_noalloc_scratch));
        if (p == nullptr) return DPPO_E_BADARG;
        g_noalloc_ptr[dev] = p;
    }
    static_assert(DPPO_GAE_NOALLOC_MAX_T % GAE_TILE == 0, "maximum must be a whole number of tiles");
    return dppo_gae_scan(rewards, values, boot, seg_end, adv, ret, T, gamma, lam, g_noalloc_ptr[dev],
                         (size_t)DPPO_GAE_NOALLOC_BYTES, stream);
}
