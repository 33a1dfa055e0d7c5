// hostgen.cpp -- host-side synthetic input: fresh random batches written straight into pinned
// staging buffers by a few threads.  Host code only.
//
// The reference builds every synthetic batch with torch.rand_like / torch.randint_like on the
// host (src/data.py:129-132): one serial Mersenne-Twister stream, 69 ms for a 128x3x224x224 fp32
// batch on the B200 box, more than three times the training step it feeds.  Here the generator is
// counter based -- element i of batch b is a hash of (seed, b, i) -- so any number of threads can
// fill disjoint ranges of the same buffer without sharing state, the result does not depend on
// the thread count, and the loop vectorises.  The destination is the pinned slot the H2D copy will
// read, so there is no intermediate tensor and no extra pass over the data.
//
// The values are uniform in [0, 1) with 24 random bits (the same resolution as torch.rand for
// fp32); labels are uniform in [0, high).  This is input for throughput runs, not a statistical
// generator: two rounds of a 32-bit integer hash over a 64-bit counter folded to 32 bits.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <thread>
#include <vector>

#include "../include/dlb200.h"

void dlb_set_error(const char* fmt, ...);

namespace {

inline uint64_t splitmix64(uint64_t x) {
    x += 0x9e3779b97f4a7c15ull;
    x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
    x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
    return x ^ (x >> 31);
}

// 32-bit finaliser with good avalanche (two multiply-xorshift rounds)
inline uint32_t mix32(uint32_t x) {
    x ^= x >> 16;
    x *= 0x7feb352du;
    x ^= x >> 15;
    x *= 0x846ca68bu;
    x ^= x >> 16;
    return x;
}

// One contiguous range [lo, hi) of the batch.  `key` already folds seed and batch index; ranges
// beyond 2^32 elements get a fresh key per 2^32 block so the 32-bit counter never repeats.
__attribute__((target_clones("avx2", "default"))) void fill_uniform_f32(float* dst, int64_t lo, int64_t hi, uint64_t key) {
    while (lo < hi) {
        const int64_t block = lo >> 32;
        const int64_t end = std::min<int64_t>(hi, (block + 1) << 32);
        const uint64_t k = splitmix64(key + (uint64_t)block);
        const uint32_t k0 = (uint32_t)k, k1 = (uint32_t)(k >> 32);
        for (int64_t i = lo; i < end; ++i) {
            const uint32_t r = mix32(((uint32_t)i ^ k0) * 0x9e3779b1u + k1);
            dst[i] = (float)(r >> 8) * (1.0f / 16777216.0f);
        }
        lo = end;
    }
}

// bf16 output: the upper 16 bits of the fp32 value (truncation keeps the value in [0, 1))
__attribute__((target_clones("avx2", "default"))) void fill_uniform_bf16(uint16_t* dst, int64_t lo, int64_t hi, uint64_t key) {
    while (lo < hi) {
        const int64_t block = lo >> 32;
        const int64_t end = std::min<int64_t>(hi, (block + 1) << 32);
        const uint64_t k = splitmix64(key + (uint64_t)block);
        const uint32_t k0 = (uint32_t)k, k1 = (uint32_t)(k >> 32);
        for (int64_t i = lo; i < end; ++i) {
            const uint32_t r = mix32(((uint32_t)i ^ k0) * 0x9e3779b1u + k1);
            const float f = (float)(r >> 8) * (1.0f / 16777216.0f);
            uint32_t bits;
            memcpy(&bits, &f, 4);
            dst[i] = (uint16_t)(bits >> 16);
        }
        lo = end;
    }
}

void fill_randint_i64(int64_t* dst, int64_t lo, int64_t hi, uint64_t key, int64_t high) {
    for (int64_t i = lo; i < hi; ++i) {
        const uint64_t r = splitmix64(key ^ (uint64_t)i * 0xd1342543de82ef95ull);
        // multiply-shift range reduction: unbiased enough for labels (high << 2^32)
        dst[i] = (int64_t)(((r >> 32) * (uint64_t)high) >> 32);
    }
}

template <typename Fn>
void run_threads(int64_t n, int nthreads, int64_t grain, Fn fn) {
    const int64_t max_useful = std::max<int64_t>(1, (n + grain - 1) / grain);
    const int t = (int)std::max<int64_t>(1, std::min<int64_t>(nthreads, max_useful));
    if (t == 1) {
        fn(0, n);
        return;
    }
    // ranges are multiples of 64 elements so no two threads share a cache line of the destination
    const int64_t per = ((n + t - 1) / t + 63) & ~int64_t(63);
    std::vector<std::thread> pool;
    pool.reserve(t - 1);
    for (int i = 1; i < t; ++i) {
        const int64_t lo = std::min(n, per * i), hi = std::min(n, per * (i + 1));
        if (lo < hi) pool.emplace_back([=] { fn(lo, hi); });
    }
    fn(0, std::min(n, per));
    for (auto& th : pool) th.join();
}

}  // namespace

extern "C" int dlb_host_fill_uniform(void* dst, int dtype, int64_t numel, uint64_t seed, uint64_t batch_index,
                                     int nthreads) {
    if (dst == nullptr || numel < 0 || (dtype != DLB_F32 && dtype != DLB_BF16)) {
        dlb_set_error("dlb_host_fill_uniform: null destination, negative count, or dtype other than fp32/bf16");
        return DLB_E_ARG;
    }
    const uint64_t key = splitmix64(splitmix64(seed) ^ (batch_index * 0xa0761d6478bd642full + 0x1234));
    if (dtype == DLB_F32) {
        float* p = static_cast<float*>(dst);
        run_threads(numel, nthreads, 1 << 18, [=](int64_t lo, int64_t hi) { fill_uniform_f32(p, lo, hi, key); });
    } else {
        uint16_t* p = static_cast<uint16_t*>(dst);
        run_threads(numel, nthreads, 1 << 18, [=](int64_t lo, int64_t hi) { fill_uniform_bf16(p, lo, hi, key); });
    }
    return 0;
}

extern "C" int dlb_host_fill_randint(int64_t* dst, int64_t numel, int64_t high, uint64_t seed, uint64_t batch_index) {
    if (dst == nullptr || numel < 0 || high <= 0 || high > (int64_t)1 << 31) {
        dlb_set_error("dlb_host_fill_randint: null destination, negative count, or high outside 1..2^31");
        return DLB_E_ARG;
    }
    const uint64_t key = splitmix64(splitmix64(seed ^ 0x6c62272e07bb0142ull) + batch_index);
    fill_randint_i64(dst, 0, numel, key, high);
    return 0;
}
