// partition.cuh — K0: bucket the query regions by (sequence, start >> shift) "windows" before any join.
//
// Why: the reference walks a pointer tree per region (src/utils/tree.rs:98-121); a GPU thread per region
// doing the same look-ups in region order touches a different part of the index in every lane: the r01a
// profile (profiles/r01a_ncu_full_summary.txt) shows ~8.5 L1 sectors and ~21 dependent loads per region
// and the SMs idle on L1 wavefronts / L2 latency at <3 % of the HBM roofline.  After this pass a warp's
// 32 regions lie in the same ~1 Mb window, so every index look-up of the later kernels (root join,
// rank-space scatter, depth / coverage scatter) falls into the same one or two 128-byte lines, which sit
// in L1.  The reference's own pair order is "per chromosome (hash-map order), then BED order"
// (src/commands/intersect.rs:124-165); nothing downstream depends on it.
//
// Cost model, measured (profiles/r01c, r01d): shared-memory atomics over a few hundred counters are cheap
// (~4 wavefronts per warp), what is expensive is HBM seeing scattered 8-byte stores: with one thread storing
// one region straight to its window, the 120 MB output thrashes the L2 and DRAM reads/writes partial sectors
// 2-3 times over (r01d: 338 MB written for 120 MB of output, 310 us).  So every tile is sorted by window in
// shared memory first and stored run by run with consecutive lanes on consecutive addresses, and the window
// count is kept small enough (<= 511) for the runs to be whole 32-byte sectors; the kernels that need finer
// locality than a window sort their own tile again in shared memory (rootjoin.cuh) or count into shared memory (counts.cuh).
//   part_count_kernel    tile of 4096 regions: shared-memory histogram over the windows, added to the global window totals
//   part_offsets_kernel  one CTA: exclusive scan of the <= 512 totals = first bucketed position of every window (`poff`)
//                        and the windows' write cursors
//   part_scatter_kernel  the tile again: ranks from the shared histogram, one atomicAdd per (tile, window) on the
//                        window's cursor claims the run's place, tile sorted in shared memory, runs stored coalesced
//                        as (start,end) pairs (+ the sequence numbers where a consumer wants them)
// The order of the runs inside a window is the arrival order of the tiles, the order inside a run that of the shared
// atomics: arbitrary, and nothing downstream depends on it (the join counts, the filters take max/min/or).
// (Until r01l the runs were placed by a window x tile count matrix and its device-wide scan: three more launches, 10 MB of
// strided matrix traffic per 10 M regions.)
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gffx {

constexpr int PART_THREADS = 512;           // the scatter kernel runs 2 CTAs of 61 KB shared memory per SM (51 registers per thread)
constexpr int PART_TILE = 4096;             // regions per CTA
constexpr int PART_UNROLL = 4;              // independent loads in flight per thread
constexpr uint32_t PART_MAX_WINDOWS = 511;  // + 1 bucket for regions whose sequence number is out of range
constexpr uint32_t PART_NB_MAX = PART_MAX_WINDOWS + 1;

// seqmap[c] of real sequence c (VSeqMap, host_index.h): x = base on its virtual sequence, y = clamp, z = first window of the
// virtual sequence | its window count << 16, w = the virtual sequence.  The partition pass is where regions leave the
// caller's coordinates: (c, s, e) -> virtual (base + min(s, clamp), base + min(e, clamp)) in window z.lo + min(vs >> shift, nw - 1).
struct WinGrid {
    const uint32_t* win0;  // n_vseqs + 1: windows of virtual sequence v are [win0[v], win0[v+1])
    const uint4* seqmap;   // n_seqs
    uint32_t n_seqs, n_vseqs, shift, n_win;
};

constexpr uint32_t PART_MAP_SMEM = 512;   // sequences whose map entries a CTA keeps in shared memory (8 KB); more: read through L1

// thread t of a tile holds elements 4 * (k * THREADS + t) .. + 3, k < PART_TILE / (4 * THREADS): 16-byte loads when the arrays allow
constexpr int PART_VEC = PART_TILE / (4 * PART_THREADS);
__device__ __forceinline__ void part_load4(const uint32_t* __restrict__ a, uint32_t i, uint32_t end, bool al16, uint32_t (&v)[4]) {
    if (al16 && i + 4 <= end) {
        const uint4 x = __ldcs(reinterpret_cast<const uint4*>(a + i));
        v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) v[j] = i + j < end ? __ldcs(a + i + j) : 0u;
    }
}

__device__ __forceinline__ void part_stage_map(const WinGrid& G, uint4* s_map) {
    if (G.n_seqs <= PART_MAP_SMEM)
        for (uint32_t i = threadIdx.x; i < G.n_seqs; i += PART_THREADS) s_map[i] = __ldg(G.seqmap + i);
}
// window of region (c, s, .) and its virtual start; c >= n_seqs: the error bucket n_win, coordinates untouched
__device__ __forceinline__ uint32_t part_window(const WinGrid& G, const uint4* s_map, uint32_t c, uint32_t s, uint4* m_out) {
    if (c >= G.n_seqs) { *m_out = make_uint4(0u, 0xFFFFFFFFu, 0u, 0xFFFFFFFFu); return G.n_win; }
    const uint4 m = G.n_seqs <= PART_MAP_SMEM ? s_map[c] : __ldg(G.seqmap + c);
    *m_out = m;
    const uint32_t b = (m.x + min(s, m.y)) >> G.shift, nw = m.z >> 16;
    return (m.z & 0xFFFFu) + (b < nw ? b : nw - 1);
}

#ifndef GFFX_PART_COUNT_TILES
#define GFFX_PART_COUNT_TILES 2
#endif
constexpr int PART_COUNT_TILES = GFFX_PART_COUNT_TILES;   // tiles per CTA of the count pass: fewer global atomics on the window totals

static __global__ void __launch_bounds__(PART_THREADS)
    part_count_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, uint32_t n, uint32_t* __restrict__ tot) {
    __shared__ uint32_t sh[PART_NB_MAX];
    __shared__ uint4 s_map[PART_MAP_SMEM];
    const uint32_t nb = G.n_win + 1;
    const bool al16 = ((reinterpret_cast<uintptr_t>(chr) | reinterpret_cast<uintptr_t>(qs)) & 15u) == 0;
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS) sh[i] = 0;
    part_stage_map(G, s_map);
    __syncthreads();
    for (int t = 0; t < PART_COUNT_TILES; t++) {
        const uint64_t base64 = ((uint64_t)blockIdx.x * PART_COUNT_TILES + t) * PART_TILE;
        if (base64 >= n) break;
        const uint32_t base = (uint32_t)base64;
        const uint32_t end = min(n, base + (uint32_t)PART_TILE);
        uint32_t c[PART_VEC][4], s[PART_VEC][4];
#pragma unroll
        for (int k = 0; k < PART_VEC; k++) {
            const uint32_t i = base + 4u * ((uint32_t)k * PART_THREADS + threadIdx.x);
            part_load4(chr, i, end, al16, c[k]);
            part_load4(qs, i, end, al16, s[k]);
        }
#pragma unroll
        for (int k = 0; k < PART_VEC; k++)
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (base + 4u * ((uint32_t)k * PART_THREADS + threadIdx.x) + j < end) {
                    uint4 m;
                    atomicAdd(&sh[part_window(G, s_map, c[k][j], s[k][j], &m)], 1u);
                }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS)
        if (sh[i]) atomicAdd(tot + i, sh[i]);
}

// one CTA of PART_NB_MAX threads: poff[w] = cursor[w] = first bucketed position of window w, poff[nb] = n; the totals are
// zeroed for the next partition pass on the stream
static __global__ void __launch_bounds__(PART_NB_MAX) part_offsets_kernel(uint32_t* __restrict__ tot, uint32_t nb, uint32_t n,
                                                                         uint32_t* __restrict__ poff, uint32_t* __restrict__ cursor) {
    __shared__ uint32_t warp_tot[PART_NB_MAX / 32];
    const uint32_t w = threadIdx.x, lane = w & 31u;
    const uint32_t v = w < nb ? tot[w] : 0u;
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, x, d);
        if (lane >= (uint32_t)d) x += o;
    }
    if (lane == 31) warp_tot[w >> 5] = x;
    __syncthreads();
    uint32_t pre = 0;
    for (uint32_t k = 0; k < (w >> 5); k++) pre += warp_tot[k];
    const uint32_t ex = pre + x - v;
    if (w < nb) { poff[w] = ex; cursor[w] = ex; tot[w] = 0; }
    if (w == 0) poff[nb] = n;
}

constexpr size_t PART_SCATTER_SMEM = (size_t)PART_TILE * 14 + (size_t)PART_NB_MAX * 8;

// ose takes the regions in VIRTUAL coordinates; oc (unless nullptr: a window lies in one virtual sequence) the virtual sequence
// numbers, 0xFFFFFFFF for a region whose sequence number is out of range
static __global__ void __launch_bounds__(PART_THREADS, 2)
    part_scatter_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, const uint32_t* __restrict__ qe,
                        uint32_t n, uint32_t* __restrict__ cursor, uint32_t* __restrict__ oc, uint2* __restrict__ ose) {
    extern __shared__ __align__(16) uint32_t part_dyn[];  // PART_SCATTER_SMEM bytes
    uint2* s_se = reinterpret_cast<uint2*>(part_dyn);                  // the tile sorted by window
    uint32_t* s_c = part_dyn + 2 * PART_TILE;
    uint16_t* s_w = reinterpret_cast<uint16_t*>(s_c + PART_TILE);
    uint32_t* s_hist = s_c + PART_TILE + PART_TILE / 2;                // regions per window -> first sorted position
    uint32_t* s_delta = s_hist + PART_NB_MAX;                          // global position minus sorted position of the window's run
    __shared__ uint32_t s_warp[PART_THREADS / 32];
    __shared__ uint4 s_map[PART_MAP_SMEM];
    const uint32_t nb = G.n_win + 1;
    const uint32_t base = blockIdx.x * (uint32_t)PART_TILE;
    const uint32_t end = min(n, base + (uint32_t)PART_TILE);
    const bool al16 = ((reinterpret_cast<uintptr_t>(chr) | reinterpret_cast<uintptr_t>(qs) | reinterpret_cast<uintptr_t>(qe)) & 15u) == 0;
    static_assert(PART_NB_MAX == PART_THREADS, "one window per thread");
    s_hist[threadIdx.x] = 0;
    part_stage_map(G, s_map);
    uint32_t c[PART_VEC][4], s[PART_VEC][4], e[PART_VEC][4];
#pragma unroll
    for (int k = 0; k < PART_VEC; k++) {
        const uint32_t i = base + 4u * ((uint32_t)k * PART_THREADS + threadIdx.x);
        part_load4(chr, i, end, al16, c[k]);
        part_load4(qs, i, end, al16, s[k]);
        part_load4(qe, i, end, al16, e[k]);
    }
    __syncthreads();
    uint32_t wr[PART_VEC][4];   // window << 12 | rank inside the tile's run of that window
#pragma unroll
    for (int k = 0; k < PART_VEC; k++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            wr[k][j] = 0xFFFFFFFFu;
            if (base + 4u * ((uint32_t)k * PART_THREADS + threadIdx.x) + j < end) {
                uint4 m;
                const uint32_t w = part_window(G, s_map, c[k][j], s[k][j], &m);
                wr[k][j] = (w << 12) | atomicAdd(&s_hist[w], 1u);
                s[k][j] = m.x + min(s[k][j], m.y);   // from here on: virtual coordinates
                e[k][j] = m.x + min(e[k][j], m.y);
                c[k][j] = m.w;
            }
        }
    static_assert(PART_TILE <= 4096, "12 bits of rank");
    __syncthreads();
    {   // exclusive scan of the histogram; every window with regions claims its run on the window's cursor
        const uint32_t w = threadIdx.x, lane = w & 31u;
        const uint32_t v = s_hist[w];
        uint32_t x = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= (uint32_t)d) x += o;
        }
        if (lane == 31) s_warp[w >> 5] = x;
        const uint32_t g = (v && w < nb) ? atomicAdd(cursor + w, v) : 0u;
        __syncthreads();
        uint32_t pre = 0;
        for (uint32_t k = 0; k < (w >> 5); k++) pre += s_warp[k];
        const uint32_t ex = pre + x - v;
        s_hist[w] = ex;
        s_delta[w] = g - ex;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < PART_VEC; k++)
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (wr[k][j] != 0xFFFFFFFFu) {
                const uint32_t w = wr[k][j] >> 12, lp = s_hist[w] + (wr[k][j] & 0xFFFu);
                s_se[lp] = make_uint2(s[k][j], e[k][j]);
                s_w[lp] = (uint16_t)w;
                if (oc) s_c[lp] = c[k][j];
            }
    __syncthreads();
    const uint32_t cnt = end - base;
    for (uint32_t i = threadIdx.x; i < cnt; i += PART_THREADS) {
        const uint32_t g = i + s_delta[s_w[i]];
        ose[g] = s_se[i];
        if (oc) oc[g] = s_c[i];
    }
}

}  // namespace gffx
