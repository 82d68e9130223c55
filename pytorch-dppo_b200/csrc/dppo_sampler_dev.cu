// On-device minibatch index generation, bit-compatible with CPython's random.sample - the call the reference makes
// at train.py:32 / ppo.py:59 (``random.sample(self.memory, batch_size)``) - so that the index block of a step never
// has to cross PCIe: the Mersenne Twister state lives in device memory, the positions are written where K2 reads them.
//
// The algorithm is not in /root/reference: it lives in CPython (3.12.3 in this image), Lib/random.py ``Random.sample`` +
// ``_randbelow_with_getrandbits`` and Modules/_randommodule.c ``genrand_uint32`` (MT19937); csrc/dppo_sampler.cu is the
// host restatement, this file the device one.  tests/test_gpu_sampler.py pins both against random.sample itself,
// including the generator state they leave behind.
//
// One CTA.  The generator's 624-word regeneration is done by all threads (three dependent sections of the twist, each
// parallel), tempering included, so a batch of 624 outputs costs a few hundred cycles.
//   * set branch (population larger than CPython's ``setsize``; C4: 4096 of 65536): random.sample draws
//     ``j = randbelow(n)`` until j is new.  Accepting "the first occurrence of a value among the valid draws" is order
//     independent, so a whole batch is resolved in parallel: every valid candidate atomicMin's its draw number into
//     first[value]; a candidate is accepted iff it holds the minimum; a block scan ranks the accepted ones; the round
//     ends at the k-th acceptance and the generator position is left exactly after that draw.
//   * pool branch (n <= setsize; C2/C3: a permutation of 2048): ``result[i] = pool[j]; pool[j] = pool[n-i-1]`` with
//     j = randbelow(n-i) is sequential by construction (every step reads what earlier steps wrote); one thread walks
//     it with the pool in shared memory (global memory beyond 48 K entries).  ~30 cycles per draw; the engine runs the
//     kernel on a side stream one step ahead.
#include <limits.h>
#include "common.cuh"

namespace {

constexpr int SD_THREADS = 640;                    // >= 624
constexpr int MT_N = 624, MT_M = 397;
constexpr int SD_POOL_SMEM_MAX = 49152;            // pool entries held in shared memory (192 KB)

struct SdShared {
    uint32_t mt[MT_N];                             // generator state
    uint32_t out[MT_N];                            // tempered outputs of the current state
    int pos;                                       // outputs consumed (CPython's ``index``)
    int status, round, have;                       // control words written by thread 0 / the scan
    int scan[SD_THREADS / 32];
    int cut;
};

__device__ __forceinline__ uint32_t mt_temper(uint32_t y) {
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9d2c5680u;
    y ^= (y << 15) & 0xefc60000u;
    y ^= (y >> 18);
    return y;
}
__device__ __forceinline__ uint32_t mt_mix(uint32_t a, uint32_t b, uint32_t far) {
    const uint32_t y = (a & 0x80000000u) | (b & 0x7fffffffu);
    return far ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}
// genrand_uint32's state regeneration, all threads.  Section by section every thread first reads its three inputs
// (the OLD s[i], s[i+1] and the far word, which is old for i < 227 and NEW afterwards), then all write.
__device__ void mt_regenerate(SdShared& sh) {
    const int t = threadIdx.x;
    const int lo[3] = {0, MT_N - MT_M, 2 * (MT_N - MT_M)};
    const int hi[3] = {MT_N - MT_M, 2 * (MT_N - MT_M), MT_N};
    for (int sec = 0; sec < 3; ++sec) {
        const int i = lo[sec] + t;
        uint32_t v = 0;
        const bool mine = i < hi[sec];
        if (mine) {
            const uint32_t far = sec == 0 ? sh.mt[i + MT_M] : sh.mt[i + MT_M - MT_N];
            v = mt_mix(sh.mt[i], i == MT_N - 1 ? sh.mt[0] : sh.mt[i + 1], far);
        }
        __syncthreads();
        if (mine) sh.mt[i] = v;
        __syncthreads();
    }
    if (t < MT_N) sh.out[t] = mt_temper(sh.mt[t]);
    if (t == 0) sh.pos = 0;
    __syncthreads();
}

__device__ __forceinline__ int bit_length_dev(uint32_t n) { return 32 - __clz(n); }

// exclusive block scan of a 0/1 flag; returns this thread's rank, total in *total (all threads)
__device__ int block_rank(int flag, SdShared& sh, int* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned m = __ballot_sync(0xffffffffu, flag);
    const int in_warp = __popc(m & ((1u << lane) - 1u));
    if (lane == 0) sh.scan[warp] = __popc(m);
    __syncthreads();
    int before = 0, tot = 0;
    for (int w = 0; w < SD_THREADS / 32; ++w) {
        const int c = sh.scan[w];
        if (w < warp) before += c;
        tot += c;
    }
    __syncthreads();
    *total = tot;
    return before + in_warp;
}

__global__ void __launch_bounds__(SD_THREADS)
sample_kernel(uint32_t* __restrict__ g_state, int* __restrict__ g_index, const uint32_t n, const int k, const int rounds,
              int32_t* __restrict__ out, int* __restrict__ first, int32_t* __restrict__ pool_g, const int pool_branch) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SdShared& sh = *reinterpret_cast<SdShared*>(smem_raw);
    int32_t* pool = pool_g != nullptr ? pool_g : reinterpret_cast<int32_t*>(smem_raw + ((sizeof(SdShared) + 15) & ~(size_t)15));
    const int t = threadIdx.x;
    if (t < MT_N) { sh.mt[t] = g_state[t]; sh.out[t] = mt_temper(sh.mt[t]); }
    if (t == 0) { sh.pos = *g_index; sh.round = 0; sh.have = 0; }
    __syncthreads();

    if (!pool_branch) {
        // ---------------- set branch ----------------
        const int bits = bit_length_dev(n);
        int round = 0, have = 0, drawn = 0;                   // uniform: accepted / valid-or-not draws so far in this round
        while (round < rounds) {
            if (sh.pos >= MT_N) mt_regenerate(sh);
            const int pos = sh.pos;
            const int i = pos + t;
            uint32_t cand = 0;
            bool valid = false;
            if (i < MT_N) {
                cand = sh.out[i] >> (32 - bits);              // getrandbits(bits)
                valid = cand < n;                             // _randbelow: redraw while r >= n
            }
            if (valid) atomicMin(&first[cand], drawn + t);
            __syncthreads();
            const bool acc = valid && __ldcg(&first[cand]) == drawn + t;   // first occurrence among the valid draws
            int total;
            const int rank = block_rank(acc ? 1 : 0, sh, &total);
            int32_t* dst = out + (size_t)round * k;
            if (have + total < k) {                           // the round goes on after this batch
                if (acc) dst[have + rank] = (int32_t)cand;
                have += total;
                drawn += MT_N - pos;
                __syncthreads();
                if (t == 0) sh.pos = MT_N;
                __syncthreads();
                continue;
            }
            // the k-th acceptance is in this batch: everything after it belongs to the next round
            if (acc && have + rank == k - 1) sh.cut = t;
            __syncthreads();
            const int cut = sh.cut;
            if (acc && t <= cut) dst[have + rank] = (int32_t)cand;
            // undo what the draws after the cut (not consumed in this round) left in the table
            if (valid && t > cut && __ldcg(&first[cand]) == drawn + t) first[cand] = INT_MAX;
            __syncthreads();
            for (int e = t; e < k; e += SD_THREADS) first[dst[e]] = INT_MAX;   // and this round's own entries
            __syncthreads();
            if (t == 0) sh.pos = pos + cut + 1;
            ++round; have = 0; drawn = 0;
            __syncthreads();
        }
    } else {
        // ---------------- pool branch ----------------
        int i = 0, round = 0;                                 // thread 0's progress (kept in registers across the loop)
        for (int e = t; e < (int)n; e += SD_THREADS) pool[e] = e;
        __syncthreads();
        while (true) {
            if (t == 0) {
                int status = 0;                               // 0 done, 1 out of random words, 2 next round
                int pos = sh.pos;
                int32_t* dst = out + (size_t)round * k;
                while (true) {
                    if (i == k) { ++round; i = 0; status = round < rounds ? 2 : 0; break; }
                    const uint32_t m = n - (uint32_t)i;
                    const int bits = bit_length_dev(m);
                    bool got = false;
                    uint32_t j = 0;
                    while (pos < MT_N) {                      // _randbelow(m)
                        j = sh.out[pos++] >> (32 - bits);
                        if (j < m) { got = true; break; }
                    }
                    if (!got) { status = 1; break; }
                    dst[i] = pool[j];
                    pool[j] = pool[m - 1];
                    ++i;
                }
                sh.pos = pos;
                sh.status = status;
            }
            __syncthreads();
            const int status = sh.status;
            if (status == 0) break;
            if (status == 1) mt_regenerate(sh);
            else {
                for (int e = t; e < (int)n; e += SD_THREADS) pool[e] = e;
                __syncthreads();
            }
        }
    }
    __syncthreads();
    if (t < MT_N) g_state[t] = sh.mt[t];
    if (t == 0) *g_index = sh.pos;
}

bool pool_branch_for(int64_t n, int k) {
    // setsize = 21 (+ 4 ** ceil(log(3k, 4)) if k > 5)        Lib/random.py, Random.sample
    double setsize = 21;
    if (k > 5) setsize += pow(4.0, ceil(log((double)k * 3.0) / log(4.0)));
    return (double)n <= setsize;
}

}  // namespace

extern "C" size_t dppo_sample_scratch_bytes(int64_t n, int k) {
    if (n < 1 || k < 0) return 16;
    return align_up(sizeof(int32_t) * (size_t)n, 16);         // first[] (set branch) or the pool (pool branch, large n)
}

extern "C" int dppo_sample_indices(uint32_t* mt_state624, int* mt_index, int64_t n, int k, int rounds, int32_t* out,
                                   void* scratch, size_t scratch_bytes, void* stream_) {
    if (!mt_state624 || !mt_index || k < 0 || rounds < 0) return DPPO_E_BADARG;
    if (n < 1 || n > 0x7fffffff || k > n) return DPPO_E_BADARG;       // random.sample raises ValueError for k > n
    if (k == 0 || rounds == 0) return 0;                               // random.sample(pop, 0) draws nothing
    if (!out || !scratch) return DPPO_E_BADARG;
    if (scratch_bytes < dppo_sample_scratch_bytes(n, k)) return DPPO_E_SCRATCH;
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool pool = pool_branch_for(n, k);
    size_t smem = (sizeof(SdShared) + 15) & ~(size_t)15;
    int32_t* pool_g = nullptr;
    if (pool) {
        if (n <= SD_POOL_SMEM_MAX) smem += sizeof(int32_t) * (size_t)n;
        else pool_g = (int32_t*)scratch;
    } else {
        DPPO_CUDA_TRY(cudaMemsetAsync(scratch, 0x7f, sizeof(int32_t) * (size_t)n, stream));   // first[] = 0x7f7f7f7f > any draw number
    }
    static size_t smem_set[64] = {0};
    int dev = 0;
    DPPO_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return DPPO_E_BADARG;
    if (smem > smem_set[dev]) {
        DPPO_CUDA_TRY(cudaFuncSetAttribute(sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        smem_set[dev] = smem;
    }
    sample_kernel<<<1, SD_THREADS, smem, stream>>>(mt_state624, mt_index, (uint32_t)n, k, rounds, out, (int*)scratch, pool_g,
                                                   pool ? 1 : 0);
    DPPO_LAUNCH_CHECK();
    return 0;
}
