// Host-side minibatch index generation, bit-compatible with CPython's
// random.sample - the call the reference makes at train.py:32 / ppo.py:59
// (``random.sample(self.memory, batch_size)``).
//
// The algorithm is not in /root/reference: it lives in CPython (3.12.3 in this
// image), Lib/random.py ``Random.sample`` + ``_randbelow_with_getrandbits`` and
// Modules/_randommodule.c ``genrand_uint32`` / ``getrandbits`` (MT19937).
// Restated here so that E x B positions cost microseconds instead of tens of
// milliseconds of interpreter time; the caller passes the 624-word Mersenne
// Twister state taken from ``random.getstate()`` and stores the advanced state
// back with ``random.setstate()``, so Python's own generator stays in step.
// tests/test_sampler_cpu.py pins it against random.sample itself.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "common.cuh"

namespace {

constexpr int MT_N = 624, MT_M = 397;

struct MT {
    uint32_t* s;
    int idx;
    uint32_t next() {                                   // genrand_uint32
        static const uint32_t mag01[2] = {0x0u, 0x9908b0dfu};
        if (idx >= MT_N) {
            int kk;
            uint32_t y;
            for (kk = 0; kk < MT_N - MT_M; kk++) {
                y = (s[kk] & 0x80000000u) | (s[kk + 1] & 0x7fffffffu);
                s[kk] = s[kk + MT_M] ^ (y >> 1) ^ mag01[y & 1u];
            }
            for (; kk < MT_N - 1; kk++) {
                y = (s[kk] & 0x80000000u) | (s[kk + 1] & 0x7fffffffu);
                s[kk] = s[kk + (MT_M - MT_N)] ^ (y >> 1) ^ mag01[y & 1u];
            }
            y = (s[MT_N - 1] & 0x80000000u) | (s[0] & 0x7fffffffu);
            s[MT_N - 1] = s[MT_M - 1] ^ (y >> 1) ^ mag01[y & 1u];
            idx = 0;
        }
        uint32_t y = s[idx++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    // _randbelow_with_getrandbits(n), n >= 1, n < 2^31: k = n.bit_length() bits per draw
    uint32_t below(uint32_t n, int bits) {
        uint32_t r = next() >> (32 - bits);
        while (r >= n) r = next() >> (32 - bits);
        return r;
    }
};

int bit_length(uint32_t n) {
    int b = 0;
    while (n) { ++b; n >>= 1; }
    return b;
}

}  // namespace

extern "C" int dppo_sample_indices_host(uint32_t* mt_state624_host, int* mt_index_host, int64_t n, int k, int rounds,
                                        int32_t* out_host) {
    if (!mt_state624_host || !mt_index_host || !out_host || k < 0 || rounds < 0) return DPPO_E_BADARG;
    if (n < 1 || n > 0x7fffffff || k > n) return DPPO_E_BADARG;        // random.sample raises ValueError for k > n
    MT mt{mt_state624_host, *mt_index_host};
    // setsize = 21 (+ 4 ** ceil(log(3k, 4)) if k > 5)        Lib/random.py, Random.sample
    double setsize = 21;
    if (k > 5) setsize += pow(4.0, ceil(log((double)k * 3.0) / log(4.0)));
    const uint32_t un = (uint32_t)n;
    if ((double)n <= setsize) {
        // pool branch: partial Fisher-Yates on a copy of the population
        int32_t* pool = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
        if (!pool) return DPPO_E_BADARG;
        for (int r = 0; r < rounds; ++r) {
            for (int64_t i = 0; i < n; ++i) pool[i] = (int32_t)i;
            int32_t* out = out_host + (size_t)r * k;
            for (int i = 0; i < k; ++i) {
                const uint32_t m = un - (uint32_t)i;
                const uint32_t j = mt.below(m, bit_length(m));
                out[i] = pool[j];
                pool[j] = pool[m - 1];
            }
        }
        free(pool);
    } else {
        // set branch: rejection of already selected positions.  Membership is a bitmap of n bits
        // while that stays cache-sized (8 KB for a 65536-step rollout), else an open-addressing set.
        const int bits = bit_length(un);
        if (n <= (int64_t)1 << 24) {
            const size_t words = ((size_t)n + 63) / 64;
            uint64_t* bm = (uint64_t*)malloc(sizeof(uint64_t) * words);
            if (!bm) return DPPO_E_BADARG;
            for (int r = 0; r < rounds; ++r) {
                memset(bm, 0, sizeof(uint64_t) * words);
                int32_t* out = out_host + (size_t)r * k;
                for (int i = 0; i < k; ++i) {
                    uint32_t j;
                    for (;;) {
                        j = mt.below(un, bits);
                        uint64_t& w = bm[j >> 6];
                        const uint64_t bit = (uint64_t)1 << (j & 63);
                        if (w & bit) continue;
                        w |= bit;
                        break;
                    }
                    out[i] = (int32_t)j;
                }
            }
            free(bm);
        } else {
            size_t cap = 16;
            while (cap < (size_t)k * 2 + 2) cap <<= 1;
            int32_t* tab = (int32_t*)malloc(sizeof(int32_t) * cap);
            if (!tab) return DPPO_E_BADARG;
            for (int r = 0; r < rounds; ++r) {
                memset(tab, 0xff, sizeof(int32_t) * cap);
                int32_t* out = out_host + (size_t)r * k;
                for (int i = 0; i < k; ++i) {
                    for (;;) {
                        const uint32_t j = mt.below(un, bits);
                        size_t h = ((size_t)j * 2654435761u) & (cap - 1);
                        bool seen = false;
                        while (tab[h] != -1) {
                            if ((uint32_t)tab[h] == j) { seen = true; break; }
                            h = (h + 1) & (cap - 1);
                        }
                        if (seen) continue;
                        tab[h] = (int32_t)j;
                        out[i] = (int32_t)j;
                        break;
                    }
                }
            }
            free(tab);
        }
    }
    *mt_index_host = mt.idx;
    return 0;
}
