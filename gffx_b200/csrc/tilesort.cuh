// tilesort.cuh — CTA-local counting sort of a tile of window-bucketed regions by a finer position key.
//
// partition.cuh groups the regions by coarse windows (a few Mb) because finer grouping would turn its
// stores into scattered sectors.  The kernels that look things up per region (root join, rank-space scatter,
// depth / coverage scatter) want a warp's 32 lanes inside a few hundred kb so that a look-up touches one or
// two cache lines; they get that by sorting their own tile once more, entirely in shared memory:
// 256 sub-buckets of the window (key = bits [shift-8, shift) of the start), one shared atomic per region
// for its rank, one CTA scan of the 256 counters, one shared-memory round trip of the (chr,start,end) triple.
// Afterwards thread t holds the sorted positions k*THREADS + t, so at every step of the unrolled loops a warp
// works on 32 ADJACENT regions.
#pragma once
#include "scan.cuh"

namespace gffx {

template <int THREADS, int QPT> struct TileSortSmem {
    uint32_t c[THREADS * QPT], s[THREADS * QPT], e[THREADS * QPT];
    uint32_t hist[512];
    uint32_t scan[THREADS / 32 + 1];
    uint32_t first_c, first_s;
};

// in : thread t holds tile elements t*QPT + k for k < n_in (the vector-load arrangement)
// out: thread t holds the sorted elements k*THREADS + t (BLOCKED = false: at every step of an unrolled loop a warp
//      works on 32 adjacent regions) or t*QPT + k (BLOCKED = true); valid[k] says whether that position exists (< m)
// key: the 256 sub-buckets of the tile's FIRST window (key_shift = window shift - 8), then the same for whatever
//      follows it in the tile (a tile that straddles a window boundary must not interleave the two windows)
template <int THREADS, int QPT, bool BLOCKED = false>
__device__ __forceinline__ void tile_sort(TileSortSmem<THREADS, QPT>& S, uint32_t key_shift, uint32_t m, int n_in, uint32_t (&c)[QPT],
                                          uint32_t (&s)[QPT], uint32_t (&e)[QPT], bool (&valid)[QPT]) {
    static_assert(THREADS == 256, "two histogram counters per thread");
    S.hist[threadIdx.x] = 0;
    S.hist[threadIdx.x + 256] = 0;
    if (threadIdx.x == 0) { S.first_c = c[0]; S.first_s = s[0]; }  // the tile is never empty
    __syncthreads();
    const uint32_t fc = S.first_c, fw = S.first_s >> (key_shift + 8);
    uint32_t r[QPT], key[QPT];
#pragma unroll
    for (int k = 0; k < QPT; k++)
        if (k < n_in) {
            key[k] = ((s[k] >> key_shift) & 255u) + ((c[k] != fc || (s[k] >> (key_shift + 8)) != fw) ? 256u : 0u);
            r[k] = atomicAdd(&S.hist[key[k]], 1u);
        }
    __syncthreads();
    const uint32_t m0 = S.hist[2 * threadIdx.x], m1 = S.hist[2 * threadIdx.x + 1];
    uint32_t tot;
    const uint32_t ex = block_scan_excl(m0 + m1, OpSum(), 0u, S.scan, &tot);
    __syncthreads();  // every thread has read its two counters
    S.hist[2 * threadIdx.x] = ex;
    S.hist[2 * threadIdx.x + 1] = ex + m0;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < QPT; k++)
        if (k < n_in) {
            const uint32_t p = S.hist[key[k]] + r[k];
            S.c[p] = c[k]; S.s[p] = s[k]; S.e[p] = e[k];
        }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < QPT; k++) {
        const uint32_t p = BLOCKED ? threadIdx.x * QPT + (uint32_t)k : (uint32_t)k * THREADS + threadIdx.x;
        valid[k] = p < m;
        if (valid[k]) { c[k] = S.c[p]; s[k] = S.s[p]; e[k] = S.e[p]; }
    }
}

}  // namespace gffx
