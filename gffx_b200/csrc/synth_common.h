// synth_common.h — counter-based generators shared by host (synth.cpp) and device (engine kernels):
// every value is a pure function of (seed, index), so any shard can produce its slice anywhere.
#pragma once
#include <cstdint>
#if defined(__CUDACC__)
#define GFFX_HD __host__ __device__ __forceinline__
#else
#define GFFX_HD inline
#endif

namespace gffx {

GFFX_HD uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
GFFX_HD uint64_t rnd2(uint64_t seed, uint64_t a, uint64_t b) {
    return splitmix64(splitmix64(seed ^ (a * 0xD1B54A32D192ED03ull)) + b * 0x9E3779B97F4A7C15ull);
}
GFFX_HD uint64_t mulhi64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}

// Region i of a synthetic BED: sequence drawn proportionally to its extent, start uniform in
// [0, extent-length], end = start+length (mirrors `bedtools random -l`, benchmark/run_benchmark.sh:24).
// cum[k] = sum of extents of sequences < k (n_seq+1 entries).
GFFX_HD void synth_region(uint64_t seed, uint64_t i, uint32_t length, const uint64_t* cum, const uint32_t* extent, uint32_t n_seq,
                          uint32_t* chr, uint32_t* start, uint32_t* end) {
    const uint64_t u = rnd2(seed, i, 1), v = rnd2(seed, i, 2);
    const uint64_t t = mulhi64(u, cum[n_seq]);
    uint32_t lo = 0, hi = n_seq;  // last k with cum[k] <= t
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (cum[mid] <= t) lo = mid; else hi = mid;
    }
    const uint32_t ext = extent[lo];
    const uint32_t span = ext > length ? ext - length + 1 : 1;
    const uint32_t s = (uint32_t)mulhi64(v, (uint64_t)span);
    *chr = lo;
    *start = s;
    *end = s + length;
}

}  // namespace gffx
