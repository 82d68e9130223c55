// K3, multi-GPU part - gradient SUM over workers fused with Adam in ONE kernel over
// NVLink peer memory (no NCCL call on the data path).
//
// Reference: every worker does `shared_grad_buffers.add_gradient(model)` into one
// shared-memory bucket (model.py:53-55, a SUM, not a mean), the chief runs Adam on it
// once all workers have counted in and zeroes it (chief.py:13-20); Counter/TrafficLight
// (utils.py) are the barrier.  Here worker r lives on GPU r and every GPU applies the
// identical update to its replica:
//
//   each CTA owns 32 float4 columns of the bucket:
//     1. sum this GPU's K2 partials for those columns (8 warps, as in dppo_adam.cu)
//     2. PUSH the slice into slot [parity][rank] of EVERY GPU's receive buffer with peer stores
//        over NVLink.  Every 8-byte word carries 4 bytes of payload and the 4-byte epoch number
//        (the "LL" idea of NCCL's low-latency protocol): 8-byte stores are single-copy atomic, so
//        a reader that sees the epoch in a word also sees its payload - no fence, no separate flag.
//     3. poll the W slots of this column block until every word carries this epoch, add the W
//        slices in RANK ORDER on top of the shared bucket's content (ones before the first update,
//        model.py:51) - identical values in identical order on every GPU, so the replicas stay
//        bit-identical - and run Adam on the columns.
//   Waiting for the words is the barrier that Counter + TrafficLight + sleep(1) implement in the
//   reference.  Receive slots are double-buffered by epoch parity; a slot is rewritten two epochs
//   later, after its reader has pushed the epoch in between (which it does after reading).
//
// Two exchange algorithms in the same kernel (dppo_comm_set_algorithm; default by world size):
//   DPPO_COMM_ALGO_PUSH_ALL  as above: one NVLink hop, every rank sends its whole bucket to every peer
//                            (W-1)*P floats out per rank - the lowest latency, right for W <= 4;
//   DPPO_COMM_ALGO_OWNER     column block jb belongs to rank jb mod W: every rank pushes its slice of the block to the
//                            owner only, the owner adds the W slices in rank order (plus the shared bucket's content)
//                            and pushes the SUM to every rank, all ranks run Adam on it.  Two hops, but 2*(W-1)/W * P
//                            floats out per rank instead of (W-1)*P: at W = 8 a quarter of the bytes, which is what
//                            the exchange was bound by there.  The sum is formed by one rank and copied, so replicas
//                            stay bit-identical by construction.
//
// CTAs walk the column blocks in index order on every GPU and push ALL their blocks before they
// poll any, so a poll only ever waits for pushes that depend on nothing.  That argument needs every
// rank's CTAs to actually run: the grid is capped at what is co-resident on this GPU (occupancy x SMs),
// and a rank that never shows up (a dead process, an unequal number of calls) cannot be waited out -
// every poll therefore carries a clock64() budget (CM_SPIN_BUDGET cycles, a few seconds); when it runs
// out the CTA records DPPO_E_COMM in the communicator's error word, stops polling and the results of that
// call are undefined.  dppo_comm_check() reads the word (it synchronises the device).
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"

namespace {

constexpr int CM_MAX_WORLD = 8;
constexpr int CM_THREADS = 256;            // 8 warps: partial groups
constexpr int CM_RQ = 32;                  // float4 columns per block
constexpr int CM_AG_MAX = 1 + 2 * 4096;    // doubles per rank in dppo_comm_allgather (rows + 2 sums per observation column)
constexpr long long CM_SPIN_BUDGET = 8000000000LL;   // clock64() cycles a poll may wait (~4 s at 1.9 GHz)

struct AdamScalars { float w1, b2, a2, bc2_sqrt, eps, alpha; };

struct CommDev {                            // everything a kernel needs, by value
    int rank, world, n4, gblocks;
    uint4* recv[CM_MAX_WORLD];              // recv[r]: rank r's receive buffer [2][world][n4][2] LL words (peer-mapped)
    unsigned* epoch;                        // this rank: completed epochs
    unsigned* done;                         // this rank: CTAs finished in the current launch
    unsigned* err;                          // this rank: non-zero once a poll has timed out
    uint4* ag[CM_MAX_WORLD];                // ag[r]: rank r's all-gather buffer [2][world][CM_AG_MAX] LL words (peer-mapped)
    unsigned* ag_epoch;                     // this rank: completed all-gathers
    uint4* gs[CM_MAX_WORLD];                // gs[r]: rank r's buffer for the owners' sums [2][n4][2] LL words (OWNER algorithm)
    int algo;
};

}  // namespace

struct dppo_comm {
    int rank, world, P_pad, n4, gblocks, loopback, grid;
    void* base;                             // this rank's allocation
    size_t bytes, recv_bytes, flag_bytes, gs_bytes;
    void* peer_base[CM_MAX_WORLD];          // mapped bases (peer_base[rank] == base)
    CommDev dev;
};

namespace {

__device__ __forceinline__ void st_ll(uint4* p, float d0, float d1, unsigned flag) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(__float_as_uint(d0)), "r"(flag),
                 "r"(__float_as_uint(d1)), "r"(flag)
                 : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p)
                 : "memory");
    return v;
}

__device__ __forceinline__ float adam_one(float g, float& p, float& m, float& v, const AdamScalars& s) {
    m = fmaf(s.w1, __fsub_rn(g, m), m);
    v = fmaf(__fmul_rn(s.a2, g), g, __fmul_rn(v, s.b2));
    const float denom = __fadd_rn(__fdiv_rn(__fsqrt_rn(v), s.bc2_sqrt), s.eps);
    p = __fadd_rn(p, __fdiv_rn(__fmul_rn(s.alpha, m), denom));
    return p;
}

// One launch serves `nvirt` ranks: nvirt == 1 is the real thing (one process per GPU);
// nvirt == world is the single-GPU loopback used by the tests, where blockIdx.y is the rank and
// the launch is cooperative so that all ranks' CTAs are co-resident.
struct RankArgs {
    CommDev c;
    const float4* partials;
    int nparts;
    float4 *gshared, *theta, *theta_prev, *m, *v, *grad_out;
    const AdamScalars* sdev;
    const float* loss_partials;
    float* loss_out;
};
struct LaunchArgs {
    RankArgs r[CM_MAX_WORLD];
    AdamScalars s;
};

__global__ void __launch_bounds__(CM_THREADS) allreduce_adam_kernel(const LaunchArgs L) {
    __shared__ float4 red[CM_THREADS / 32][CM_RQ];
    const RankArgs& A = L.r[blockIdx.y];
    const CommDev& C = A.c;
    AdamScalars s = L.s;
    if (A.sdev) s = *A.sdev;
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const unsigned epoch = *reinterpret_cast<volatile unsigned*>(C.epoch) + 1u;   // same for every CTA of this launch
    const int par = epoch & 1u;
    const int n4 = C.n4;
    if (A.loss_partials && A.loss_out && blockIdx.x == 0 && threadIdx.x < 4) {
        float t = 0.f;
        for (int p = 0; p < A.nparts; ++p) t = __fadd_rn(t, A.loss_partials[p * 4 + threadIdx.x]);
        A.loss_out[threadIdx.x] = t;
    }
    // ---- phase 1: sum my partials per column block and push the slices to every rank ------------
    for (int jb = blockIdx.x; jb < C.gblocks; jb += gridDim.x) {
        const int i = jb * CM_RQ + lane;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < n4) {
#pragma unroll 6
            for (int p = g; p < A.nparts; p += CM_THREADS / 32) {
                const float4 x = __ldcs(&A.partials[(size_t)p * n4 + i]);
                a.x = __fadd_rn(a.x, x.x); a.y = __fadd_rn(a.y, x.y); a.z = __fadd_rn(a.z, x.z); a.w = __fadd_rn(a.w, x.w);
            }
        }
        __syncthreads();                                  // red[] of the previous column block consumed
        red[g][lane] = a;
        __syncthreads();
        // PUSH_ALL: warps 0..world-1, warp w pushes to rank w (all warps hold the same sum); OWNER: warp 0 pushes to the owner
        const int owner = jb % C.world;
        if ((C.algo == DPPO_COMM_ALGO_OWNER ? g == 0 : g < C.world) && i < n4) {
            float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int r = 0; r < CM_THREADS / 32; ++r) {
                const float4 x = red[r][lane];
                mine.x = __fadd_rn(mine.x, x.x); mine.y = __fadd_rn(mine.y, x.y);
                mine.z = __fadd_rn(mine.z, x.z); mine.w = __fadd_rn(mine.w, x.w);
            }
            uint4* dst = C.recv[C.algo == DPPO_COMM_ALGO_OWNER ? owner : g] + (((size_t)par * C.world + C.rank) * n4 + i) * 2;
            st_ll(dst, mine.x, mine.y, epoch);
            st_ll(dst + 1, mine.z, mine.w, epoch);
        }
    }
    // ---- phase 2: poll every rank's slice, sum in rank order, Adam -----------------------------------
    // warp r (< world) polls rank r's slice of the column block - the W round trips to the words that the
    // peers wrote run side by side instead of one after the other - and parks it in shared memory; warp 0
    // then adds the W slices in rank order and runs Adam.
    for (int jb = blockIdx.x; jb < C.gblocks; jb += gridDim.x) {
        const int i = jb * CM_RQ + lane;
        const bool owner_algo = C.algo == DPPO_COMM_ALGO_OWNER;
        const bool gather_here = !owner_algo || (jb % C.world) == C.rank;     // this rank adds the W slices of the block
        __syncthreads();                                  // red[] of phase 1 / of the previous column block consumed
        if (gather_here && g < C.world && i < n4) {
            const uint4* q = C.recv[C.rank] + (((size_t)par * C.world + g) * n4 + i) * 2;
            uint4 lo = ld_ll(q), hi = ld_ll(q + 1);
            if (lo.y != epoch || lo.w != epoch || hi.y != epoch || hi.w != epoch) {
                const long long t0 = clock64();
                while (true) {
                    lo = ld_ll(q); hi = ld_ll(q + 1);
                    if (lo.y == epoch && lo.w == epoch && hi.y == epoch && hi.w == epoch) break;
                    if (clock64() - t0 > CM_SPIN_BUDGET) { atomicExch(C.err, 1u); break; }
                }
            }
            red[g][lane] = make_float4(__uint_as_float(lo.x), __uint_as_float(lo.z), __uint_as_float(hi.x), __uint_as_float(hi.z));
        }
        __syncthreads();
        float4 G = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gather_here && i < n4) {                      // (every warp forms the same sum: warp w pushes it to rank w below)
            for (int r = 0; r < C.world; ++r) {
                const float4 x = red[r][lane];
                G.x = __fadd_rn(G.x, x.x); G.y = __fadd_rn(G.y, x.y); G.z = __fadd_rn(G.z, x.z); G.w = __fadd_rn(G.w, x.w);
            }
            if (A.gshared) {                              // shared bucket content on top (as dppo_adam.cu)
                const float4 s0 = A.gshared[i];
                G.x = __fadd_rn(s0.x, G.x); G.y = __fadd_rn(s0.y, G.y); G.z = __fadd_rn(s0.z, G.z); G.w = __fadd_rn(s0.w, G.w);
            }
        }
        if (owner_algo) {
            if (gather_here && g < C.world && i < n4) {   // second hop: the owner's sum to every rank (its own copy included)
                uint4* dst = C.gs[g] + ((size_t)par * n4 + i) * 2;
                st_ll(dst, G.x, G.y, epoch);
                st_ll(dst + 1, G.z, G.w, epoch);
            }
            if (g == 0 && i < n4) {
                const uint4* q = C.gs[C.rank] + ((size_t)par * n4 + i) * 2;
                uint4 lo = ld_ll(q), hi = ld_ll(q + 1);
                if (lo.y != epoch || lo.w != epoch || hi.y != epoch || hi.w != epoch) {
                    const long long t0 = clock64();
                    while (true) {
                        lo = ld_ll(q); hi = ld_ll(q + 1);
                        if (lo.y == epoch && lo.w == epoch && hi.y == epoch && hi.w == epoch) break;
                        if (clock64() - t0 > CM_SPIN_BUDGET) { atomicExch(C.err, 1u); break; }
                    }
                }
                G = make_float4(__uint_as_float(lo.x), __uint_as_float(lo.z), __uint_as_float(hi.x), __uint_as_float(hi.z));
            }
        }
        if (g != 0 || i >= n4) continue;
        float4 p = A.theta[i], mm = A.m[i], vv = A.v[i];
        if (A.theta_prev) A.theta_prev[i] = p;
        adam_one(G.x, p.x, mm.x, vv.x, s);
        adam_one(G.y, p.y, mm.y, vv.y, s);
        adam_one(G.z, p.z, mm.z, vv.z, s);
        adam_one(G.w, p.w, mm.w, vv.w, s);
        A.theta[i] = p; A.m[i] = mm; A.v[i] = vv;
        if (A.grad_out) A.grad_out[i] = G;
        if (A.gshared) A.gshared[i] = make_float4(0.f, 0.f, 0.f, 0.f);      // chief.py:19
    }
    // the last CTA of this rank to finish closes the epoch (every CTA read it at its start)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned fin = atomicAdd(C.done, 1u) + 1u;
        if (fin == gridDim.x) {
            *reinterpret_cast<volatile unsigned*>(C.done) = 0u;
            *reinterpret_cast<volatile unsigned*>(C.epoch) = epoch;
            __threadfence();
        }
    }
}

AdamScalars make_scalars(int step, double lr, double beta1, double beta2, double eps) {
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

size_t layout(dppo_comm* c) {
    c->recv_bytes = align_up(sizeof(uint4) * 2 * 2 * (size_t)c->world * c->n4, 256);    // LL: 2x the payload
    c->flag_bytes = align_up(sizeof(uint4) * 2 * (size_t)c->world * CM_AG_MAX, 256);   // the small all-gather's slots
    c->gs_bytes = align_up(sizeof(uint4) * 2 * 2 * (size_t)c->n4, 256);                // the owners' sums (OWNER algorithm)
    return c->recv_bytes + c->flag_bytes + c->gs_bytes + 256;
}

void fill_dev(dppo_comm* c) {
    c->dev.rank = c->rank; c->dev.world = c->world; c->dev.n4 = c->n4; c->dev.gblocks = c->gblocks;
    for (int r = 0; r < c->world; ++r) {
        c->dev.recv[r] = (uint4*)c->peer_base[r];
        c->dev.ag[r] = (uint4*)((char*)c->peer_base[r] + c->recv_bytes);
        c->dev.gs[r] = (uint4*)((char*)c->peer_base[r] + c->recv_bytes + c->flag_bytes);
    }
    if (c->dev.algo != DPPO_COMM_ALGO_PUSH_ALL && c->dev.algo != DPPO_COMM_ALGO_OWNER)
        c->dev.algo = c->world > 4 ? DPPO_COMM_ALGO_OWNER : DPPO_COMM_ALGO_PUSH_ALL;
    c->dev.epoch = (unsigned*)((char*)c->base + c->recv_bytes + c->flag_bytes + c->gs_bytes);
    c->dev.done = c->dev.epoch + 1;
    c->dev.err = c->dev.epoch + 2;
    c->dev.ag_epoch = c->dev.epoch + 3;
}

// every CTA of the launch must be resident at once (a CTA polls words that the same-numbered CTA of each
// peer pushes; CTAs are dispatched in index order, so equal grids of co-resident CTAs cannot wait on a CTA
// that has not started)
int grid_for(const dppo_comm* c, int* out) {
    int per_sm = 0;
    DPPO_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, allreduce_adam_kernel, CM_THREADS, 0));
    const int cap = per_sm * dppo_sm_count();
    if (cap < 1) return DPPO_E_TOOLARGE;
    *out = c->gblocks < cap ? c->gblocks : cap;
    return 0;
}

}  // namespace

// ---- life cycle (the one place where the library allocates: peer-visible memory must come
// from cudaMalloc so that it can be exported with cudaIpcGetMemHandle) ---------------------------
extern "C" int dppo_comm_create(int rank, int world, int P_pad, dppo_comm** out) {
    if (!out || world < 1 || world > CM_MAX_WORLD || rank < 0 || rank >= world || P_pad <= 0 || (P_pad & 3))
        return DPPO_E_BADARG;
    dppo_comm* c = (dppo_comm*)calloc(1, sizeof(dppo_comm));
    if (!c) return DPPO_E_BADARG;
    c->rank = rank; c->world = world; c->P_pad = P_pad; c->n4 = P_pad / 4;
    c->gblocks = ceil_div(c->n4, CM_RQ);
    c->bytes = layout(c);
    cudaError_t e = cudaMalloc(&c->base, c->bytes);
    if (e != cudaSuccess) { free(c); return (int)e; }
    e = cudaMemset(c->base, 0, c->bytes);
    if (e != cudaSuccess) { cudaFree(c->base); free(c); return (int)e; }
    c->peer_base[rank] = c->base;
    if (world == 1) fill_dev(c);
    *out = c;
    return 0;
}

// PUSH_ALL / OWNER (see the top of this file); every rank of a communicator must choose the same, between epochs.
extern "C" int dppo_comm_set_algorithm(dppo_comm* c, int algo) {
    if (!c || (algo != DPPO_COMM_ALGO_PUSH_ALL && algo != DPPO_COMM_ALGO_OWNER)) return DPPO_E_BADARG;
    c->dev.algo = algo;
    return 0;
}
extern "C" int dppo_comm_get_algorithm(dppo_comm* c) { return c ? c->dev.algo : DPPO_E_BADARG; }

extern "C" int dppo_comm_handle_bytes(void) { return (int)sizeof(cudaIpcMemHandle_t); }

extern "C" int dppo_comm_local_handle(dppo_comm* c, void* handle_host) {
    if (!c || !handle_host) return DPPO_E_BADARG;
    cudaIpcMemHandle_t h;
    DPPO_CUDA_TRY(cudaIpcGetMemHandle(&h, c->base));
    memcpy(handle_host, &h, sizeof(h));
    return 0;
}

// handles_host: world consecutive handles in rank order (this rank's own entry is ignored)
extern "C" int dppo_comm_open_peers(dppo_comm* c, const void* handles_host) {
    if (!c || !handles_host) return DPPO_E_BADARG;
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)handles_host + (size_t)r * sizeof(h), sizeof(h));
        void* p = nullptr;
        DPPO_CUDA_TRY(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        c->peer_base[r] = p;
    }
    fill_dev(c);
    return 0;
}

// Test helper: `world` communicators in ONE process on ONE GPU whose "peers" are each other's
// buffers; only usable with dppo_allreduce_adam_loopback.
extern "C" int dppo_comm_create_loopback(int world, int P_pad, dppo_comm** out_array) {
    if (!out_array || world < 1 || world > CM_MAX_WORLD) return DPPO_E_BADARG;
    for (int r = 0; r < world; ++r) {
        const int rc = dppo_comm_create(r, world, P_pad, &out_array[r]);
        if (rc) return rc;
        out_array[r]->loopback = 1;
    }
    for (int r = 0; r < world; ++r) {
        for (int q = 0; q < world; ++q) out_array[r]->peer_base[q] = out_array[q]->base;
        fill_dev(out_array[r]);
    }
    return 0;
}

extern "C" int dppo_comm_destroy(dppo_comm* c) {
    if (!c) return 0;
    if (!c->loopback)
        for (int r = 0; r < c->world; ++r)
            if (r != c->rank && c->peer_base[r]) cudaIpcCloseMemHandle(c->peer_base[r]);
    if (c->base) cudaFree(c->base);
    free(c);
    return 0;
}

static int fill_rank_args(RankArgs& R, dppo_comm* c, const float* partials, int nparts, float* gshared, float* theta,
                          float* theta_prev, float* m, float* v, int P_pad, const float* scalars_dev, float* grad_out,
                          const float* loss_partials, float* loss_out) {
    if (!c || P_pad != c->P_pad || nparts < 0 || !theta || !m || !v || (nparts > 0 && !partials)) return DPPO_E_BADARG;
    if (!c->dev.recv[c->world - 1] || !c->dev.recv[0]) return DPPO_E_COMM;       // peers not opened
    if (!aligned16(partials) || !aligned16(gshared) || !aligned16(theta) || !aligned16(theta_prev) || !aligned16(m) ||
        !aligned16(v) || !aligned16(grad_out))
        return DPPO_E_ALIGN;
    R.c = c->dev;
    R.partials = (const float4*)partials; R.nparts = nparts;
    R.gshared = (float4*)gshared; R.theta = (float4*)theta; R.theta_prev = (float4*)theta_prev;
    R.m = (float4*)m; R.v = (float4*)v; R.grad_out = (float4*)grad_out;
    R.sdev = (const AdamScalars*)scalars_dev;
    R.loss_partials = loss_partials; R.loss_out = loss_out;
    return 0;
}

extern "C" int dppo_allreduce_adam(dppo_comm* c, const float* partials, int nparts, float* gshared, float* theta,
                                   float* theta_prev, float* m, float* v, int P_pad, int step, double lr, double beta1,
                                   double beta2, double eps, const float* scalars_dev, float* grad_out,
                                   const float* loss_partials, float* loss_out, void* stream) {
    if (!c || c->loopback) return DPPO_E_COMM;
    if (step < 1 && !scalars_dev) return DPPO_E_BADARG;
    LaunchArgs L;                                             // by value into the launch: nothing shared between calls
    int rc = fill_rank_args(L.r[0], c, partials, nparts, gshared, theta, theta_prev, m, v, P_pad, scalars_dev,
                            grad_out, loss_partials, loss_out);
    if (rc) return rc;
    L.s = make_scalars(step < 1 ? 1 : step, lr, beta1, beta2, eps);
    if (c->grid == 0 && (rc = grid_for(c, &c->grid)) != 0) return rc;
    allreduce_adam_kernel<<<dim3(c->grid, 1), CM_THREADS, 0, (cudaStream_t)stream>>>(L);
    DPPO_LAUNCH_CHECK();
    return 0;
}

// 0 while no poll of this communicator has timed out, DPPO_E_COMM afterwards (sticky).  Synchronises the device.
extern "C" int dppo_comm_check(dppo_comm* c) {
    if (!c || !c->dev.err) return DPPO_E_BADARG;
    unsigned e = 0;
    DPPO_CUDA_TRY(cudaMemcpy(&e, c->dev.err, sizeof(e), cudaMemcpyDeviceToHost));
    return e ? DPPO_E_COMM : 0;
}

// All `world` virtual ranks in one cooperative launch on the current GPU (tests).  Arrays are
// indexed by rank.
extern "C" int dppo_allreduce_adam_loopback(dppo_comm** comms, int world, const float* const* partials,
                                            const int* nparts, float* const* gshared, float* const* theta,
                                            float* const* theta_prev, float* const* m, float* const* v, int P_pad,
                                            int step, double lr, double beta1, double beta2, double eps,
                                            float* const* grad_out, void* stream) {
    if (!comms || world < 1 || world > CM_MAX_WORLD || step < 1) return DPPO_E_BADARG;
    LaunchArgs L;
    for (int r = 0; r < world; ++r) {
        if (!comms[r] || !comms[r]->loopback || comms[r]->world != world) return DPPO_E_COMM;
        const int rc = fill_rank_args(L.r[r], comms[r], partials[r], nparts[r], gshared ? gshared[r] : nullptr, theta[r],
                                      theta_prev ? theta_prev[r] : nullptr, m[r], v[r], P_pad, nullptr,
                                      grad_out ? grad_out[r] : nullptr, nullptr, nullptr);
        if (rc) return rc;
    }
    L.s = make_scalars(step, lr, beta1, beta2, eps);
    int per_sm = 0;
    DPPO_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, allreduce_adam_kernel, CM_THREADS, 0));
    int gx = per_sm * dppo_sm_count() / world;                 // all ranks' CTAs must be co-resident
    if (gx < 1) return DPPO_E_TOOLARGE;
    if (gx > comms[0]->gblocks) gx = comms[0]->gblocks;
    void* args[] = {(void*)&L};
    DPPO_CUDA_TRY(cudaLaunchCooperativeKernel((const void*)allreduce_adam_kernel, dim3(gx, world), dim3(CM_THREADS), args,
                                              0, (cudaStream_t)stream));
    return 0;
}

// ---- small all-gather over the same peer channel: every rank contributes `count` doubles and receives all
// ranks' vectors in rank order.  Used once per rollout for the observation statistics that the reference's
// workers share (one Shared_obs_stats object in shared memory, main.py:54): each worker's (rows, S1, S2).
// One CTA per rank; a double travels as ONE 16-byte LL word {lo, epoch, hi, epoch}. ----
struct AgArgs {
    CommDev c[CM_MAX_WORLD];
    const double* src[CM_MAX_WORLD];
    double* dst[CM_MAX_WORLD];
    int count;
};
__global__ void __launch_bounds__(CM_THREADS) allgather_kernel(const AgArgs A) {
    const CommDev& C = A.c[blockIdx.x];
    const unsigned epoch = *reinterpret_cast<volatile unsigned*>(C.ag_epoch) + 1u;
    const int par = epoch & 1u;
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(A.src[blockIdx.x]);
    for (int i = threadIdx.x; i < A.count; i += CM_THREADS) {
        const unsigned long long x = src[i];
        for (int r = 0; r < C.world; ++r) {
            uint4* d = C.ag[r] + ((size_t)par * C.world + C.rank) * CM_AG_MAX + i;
            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(d), "r"((unsigned)x), "r"(epoch),
                         "r"((unsigned)(x >> 32)), "r"(epoch) : "memory");
        }
    }
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(A.dst[blockIdx.x]);
    for (int e = threadIdx.x; e < A.count * C.world; e += CM_THREADS) {
        const int r = e / A.count, i = e - r * A.count;
        const uint4* q = C.ag[C.rank] + ((size_t)par * C.world + r) * CM_AG_MAX + i;
        uint4 v = ld_ll(q);
        if (v.y != epoch || v.w != epoch) {
            const long long t0 = clock64();
            while (true) {
                v = ld_ll(q);
                if (v.y == epoch && v.w == epoch) break;
                if (clock64() - t0 > CM_SPIN_BUDGET) { atomicExch(C.err, 1u); break; }
            }
        }
        dst[e] = (unsigned long long)v.x | ((unsigned long long)v.z << 32);
    }
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); *reinterpret_cast<volatile unsigned*>(C.ag_epoch) = epoch; }
}

extern "C" int dppo_comm_allgather(dppo_comm* c, const double* src, double* dst, int count, void* stream) {
    if (!c || c->loopback || !src || !dst || count < 1) return DPPO_E_BADARG;
    if (count > CM_AG_MAX) return DPPO_E_TOOLARGE;
    if (!c->dev.ag[c->world - 1] || !c->dev.ag[0]) return DPPO_E_COMM;
    AgArgs A;
    A.c[0] = c->dev; A.src[0] = src; A.dst[0] = dst; A.count = count;
    allgather_kernel<<<1, CM_THREADS, 0, (cudaStream_t)stream>>>(A);
    DPPO_LAUNCH_CHECK();
    return 0;
}
// all virtual ranks of a loopback communicator set in one launch (tests): src[r] / dst[r] per rank
extern "C" int dppo_comm_allgather_loopback(dppo_comm** comms, int world, const double* const* src, double* const* dst,
                                            int count, void* stream) {
    if (!comms || world < 1 || world > CM_MAX_WORLD || !src || !dst || count < 1) return DPPO_E_BADARG;
    if (count > CM_AG_MAX) return DPPO_E_TOOLARGE;
    AgArgs A;
    for (int r = 0; r < world; ++r) {
        if (!comms[r] || !comms[r]->loopback || comms[r]->world != world) return DPPO_E_COMM;
        A.c[r] = comms[r]->dev; A.src[r] = src[r]; A.dst[r] = dst[r];
    }
    A.count = count;
    allgather_kernel<<<world, CM_THREADS, 0, (cudaStream_t)stream>>>(A);      // <= 8 CTAs: always co-resident
    DPPO_LAUNCH_CHECK();
    return 0;
}

// sum of the K2 partials only (building block of the NCCL baseline path, see b200_comm.py)
__global__ void __launch_bounds__(256) reduce_partials_kernel(const float4* __restrict__ partials, int nparts, int n4,
                                                              float4* __restrict__ out) {
    __shared__ float4 red[8][32];
    const int lane = threadIdx.x & 31, g = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + lane;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < n4)
        for (int p = g; p < nparts; p += 8) {
            const float4 x = __ldcs(&partials[(size_t)p * n4 + i]);
            a.x = __fadd_rn(a.x, x.x); a.y = __fadd_rn(a.y, x.y); a.z = __fadd_rn(a.z, x.z); a.w = __fadd_rn(a.w, x.w);
        }
    red[g][lane] = a;
    __syncthreads();
    if (g == 0 && i < n4) {
        float4 G = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int r = 0; r < 8; ++r) {
            const float4 x = red[r][lane];
            G.x = __fadd_rn(G.x, x.x); G.y = __fadd_rn(G.y, x.y); G.z = __fadd_rn(G.z, x.z); G.w = __fadd_rn(G.w, x.w);
        }
        out[i] = G;
    }
}

extern "C" int dppo_reduce_partials(const float* partials, int nparts, float* out, int P_pad, void* stream) {
    if (!partials || !out || nparts < 1 || P_pad <= 0 || (P_pad & 3)) return DPPO_E_BADARG;
    if (!aligned16(partials) || !aligned16(out)) return DPPO_E_ALIGN;
    const int n4 = P_pad / 4;
    reduce_partials_kernel<<<ceil_div(n4, 32), 256, 0, (cudaStream_t)stream>>>((const float4*)partials, nparts, n4,
                                                                               (float4*)out);
    DPPO_LAUNCH_CHECK();
    return 0;
}
