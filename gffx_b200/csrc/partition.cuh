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
#include <algorithm>
#include <cstdint>

#include "bulkcopy.cuh"

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

// sequences whose map entries a CTA keeps in shared memory (4 KB); more: read through L1.  (A 4-byte "fast" entry per sequence —
// first window and a limit below which a coordinate is its own virtual coordinate — halves the kernels' shared-memory
// wavefronts and makes both slower, r02y: 25 -> 28.5 and 61 -> 65 us; they are bound by instruction issue and dependent
// latency, not by the shared-memory pipe, and the one LDS.128 is the fewest instructions.)
constexpr uint32_t PART_MAP_SMEM = 256;

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
// window of region (c, s, .) and its map entry; c >= n_seqs: the error bucket n_win, coordinates untouched
__device__ __forceinline__ uint32_t part_window(const WinGrid& G, const uint4* s_map, uint32_t c, uint32_t s, uint4* m_out) {
    if (c >= G.n_seqs) { *m_out = make_uint4(0u, 0xFFFFFFFFu, 0u, 0xFFFFFFFFu); return G.n_win; }
    const uint4 m = G.n_seqs <= PART_MAP_SMEM ? s_map[c] : __ldg(G.seqmap + c);
    *m_out = m;
    const uint32_t b = (m.x + min(s, m.y)) >> G.shift, nw = m.z >> 16;
    return (m.z & 0xFFFFu) + (b < nw ? b : nw - 1);
}

// The count pass is persistent too: a CTA walks tiles blockIdx.x, + gridDim.x, ... with TWO tiles' columns (sequence number,
// start: 2 x 16 KB each) in flight to shared memory by cp.async.bulk, counts into one shared histogram for all its tiles and adds
// the non-zero cells to the global window totals once (a CTA per two tiles did that 1221 times per 10 M regions and waited for
// its loads twice: 26 us where the traffic needs 12).
constexpr int PART_COUNT_CTAS = 3;   // per SM
constexpr size_t PART_COUNT_SMEM = (size_t)2 * 2 * PART_TILE * 4;

static __global__ void __launch_bounds__(PART_THREADS, PART_COUNT_CTAS)
    part_count_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, uint32_t n, uint32_t* __restrict__ tot) {
    extern __shared__ __align__(16) uint32_t part_dyn[];   // [2 stages][chr, start][PART_TILE]
    __shared__ uint32_t sh[PART_NB_MAX];
    __shared__ uint4 s_map[PART_MAP_SMEM];
    __shared__ __align__(8) uint64_t mbar[2];
    const uint32_t nb = G.n_win + 1;
    const uint32_t tiles = (n + PART_TILE - 1) / PART_TILE;
    const bool al16 = ((reinterpret_cast<uintptr_t>(chr) | reinterpret_cast<uintptr_t>(qs)) & 15u) == 0;
    auto bulk_ok = [&](uint32_t t) { return al16 && t < tiles && (uint64_t)(t + 1) * PART_TILE <= n; };
    auto issue = [&](uint32_t t, uint32_t st) {   // thread 0
        const size_t o = (size_t)t * PART_TILE;
        mbar_expect_tx(&mbar[st], 2u * PART_TILE * 4u);
        bulk_g2s(part_dyn + st * 2 * PART_TILE, chr + o, PART_TILE * 4u, &mbar[st]);
        bulk_g2s(part_dyn + st * 2 * PART_TILE + PART_TILE, qs + o, PART_TILE * 4u, &mbar[st]);
    };
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS) sh[i] = 0;
    part_stage_map(G, s_map);
    if (threadIdx.x == 0) {
        mbar_init(&mbar[0], 1);
        mbar_init(&mbar[1], 1);
        if (bulk_ok(blockIdx.x)) issue(blockIdx.x, 0);
        if (bulk_ok(blockIdx.x + gridDim.x)) issue(blockIdx.x + gridDim.x, 1);
    }
    uint32_t phases = 0;   // bit st: the parity stage st's barrier completes next
    __syncthreads();
    uint32_t st = 0;
    for (uint32_t tile = blockIdx.x; tile < tiles; tile += gridDim.x, st ^= 1u) {
        const uint32_t base = tile * (uint32_t)PART_TILE;
        const uint32_t end = min(n, base + (uint32_t)PART_TILE);
        uint32_t c[PART_VEC][4], s[PART_VEC][4];
        const bool staged = bulk_ok(tile);
        if (staged) {
            mbar_wait(&mbar[st], (phases >> st) & 1u);
            phases ^= 1u << st;
            const uint32_t* raw = part_dyn + st * 2 * PART_TILE;
#pragma unroll
            for (int k = 0; k < PART_VEC; k++) {
                const uint32_t i = 4u * ((uint32_t)k * PART_THREADS + threadIdx.x);
                const uint4 a = *reinterpret_cast<const uint4*>(raw + i), b = *reinterpret_cast<const uint4*>(raw + PART_TILE + i);
                c[k][0] = a.x; c[k][1] = a.y; c[k][2] = a.z; c[k][3] = a.w;
                s[k][0] = b.x; s[k][1] = b.y; s[k][2] = b.z; s[k][3] = b.w;
            }
            __syncthreads();   // the stage is in registers: hand it back to the copy engine
            if (threadIdx.x == 0 && bulk_ok(tile + 2 * gridDim.x)) {
                fence_proxy_async_smem();
                issue(tile + 2 * gridDim.x, st);
            }
        } else {
#pragma unroll
            for (int k = 0; k < PART_VEC; k++) {
                const uint32_t i = base + 4u * ((uint32_t)k * PART_THREADS + threadIdx.x);
                part_load4(chr, i, end, al16, c[k]);
                part_load4(qs, i, end, al16, s[k]);
            }
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
// zeroed for the next partition pass on the stream.  (Folding this into the count pass — its last CTA to finish, found by a
// ticket, does the scan — removes the launch and measures the same, r03o: the count pass's tail grows by what the launch cost.)
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

#ifndef GFFX_PART_STILE
#define GFFX_PART_STILE 4096
#endif
constexpr int PART_STILE = GFFX_PART_STILE;   // regions per tile of the scatter pass (2048 with three or four CTAs per SM measured slower:
                                              // 0.397 / 0.428 against 0.372 ms per four passes — shorter runs, twice the per-tile scans and cursor claims)
constexpr int PART_SVEC = PART_STILE / (4 * PART_THREADS);
// shared memory of the scatter kernel: the sorted tile (10 bytes per region), the window histogram / deltas, and the NEXT tile's
// three input columns as the copy engine delivers them (3 x 16 KB): 92 KB, two CTAs per SM
constexpr size_t PART_SORT_SMEM = (size_t)PART_STILE * 10 + (size_t)PART_NB_MAX * 8;
constexpr size_t PART_SCATTER_SMEM = PART_SORT_SMEM + (size_t)PART_STILE * 12;
#ifndef GFFX_PART_SCTAS
#define GFFX_PART_SCTAS (GFFX_PART_STILE <= 2048 ? 3 : 2)
#endif
constexpr int PART_SCATTER_CTAS = GFFX_PART_SCTAS;   // per SM; the kernel is persistent: grid = min(tiles, SMs * PART_SCATTER_CTAS)

// ose takes the regions in VIRTUAL coordinates (a window lies in one virtual sequence: the consumers take the sequence from the
// window; regions whose sequence number is out of range land in the bucket behind the last window).
// Persistent (r02): a CTA walks tiles blockIdx.x, + gridDim.x, ...; while it ranks, sorts and stores one tile, the next tile's
// columns arrive in shared memory by three cp.async.bulk copies (16 KB each, completion on an mbarrier) — the load latency that
// a CTA per tile paid in front of every tile (r02k: 64 us per 10 M regions, DRAM 28 % busy) is off the critical path.  A tile
// that the copy engine cannot take (arrays not 16-byte aligned, the partial last tile) is loaded by the threads as before.
static __global__ void __launch_bounds__(PART_THREADS, PART_SCATTER_CTAS)
    part_scatter_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, const uint32_t* __restrict__ qe,
                        uint32_t n, uint32_t* __restrict__ cursor, uint2* __restrict__ ose) {
    extern __shared__ __align__(16) uint32_t part_dyn[];  // PART_SCATTER_SMEM bytes
    uint2* s_se = reinterpret_cast<uint2*>(part_dyn);                  // the tile sorted by window
    uint16_t* s_w = reinterpret_cast<uint16_t*>(part_dyn + 2 * PART_STILE);
    uint32_t* s_hist = part_dyn + 2 * PART_STILE + PART_STILE / 2;       // regions per window -> first sorted position
    uint32_t* s_delta = s_hist + PART_NB_MAX;                          // global position minus sorted position of the window's run
    uint32_t* s_raw = s_delta + PART_NB_MAX;                           // [3][PART_STILE]: chr, start, end of the tile in flight
    __shared__ uint32_t s_warp[PART_THREADS / 32];
    __shared__ uint4 s_map[PART_MAP_SMEM];
    __shared__ __align__(8) uint64_t mbar;
    const uint32_t nb = G.n_win + 1;
    const uint32_t tiles = (n + PART_STILE - 1) / PART_STILE;
    const bool al16 = ((reinterpret_cast<uintptr_t>(chr) | reinterpret_cast<uintptr_t>(qs) | reinterpret_cast<uintptr_t>(qe)) & 15u) == 0;
    static_assert(PART_NB_MAX == PART_THREADS, "one window per thread");
    auto bulk_ok = [&](uint32_t t) { return al16 && t < tiles && (uint64_t)(t + 1) * PART_STILE <= n; };
    auto issue = [&](uint32_t t) {   // thread 0
        const size_t o = (size_t)t * PART_STILE;
        mbar_expect_tx(&mbar, 3u * PART_STILE * 4u);
        bulk_g2s(s_raw, chr + o, PART_STILE * 4u, &mbar);
        bulk_g2s(s_raw + PART_STILE, qs + o, PART_STILE * 4u, &mbar);
        bulk_g2s(s_raw + 2 * PART_STILE, qe + o, PART_STILE * 4u, &mbar);
    };
    part_stage_map(G, s_map);
    if (threadIdx.x == 0) {
        mbar_init(&mbar, 1);
        if (bulk_ok(blockIdx.x)) issue(blockIdx.x);
    }
    uint32_t phase = 0;
    __syncthreads();
    for (uint32_t tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const uint32_t base = tile * (uint32_t)PART_STILE;
        const uint32_t end = min(n, base + (uint32_t)PART_STILE);
        s_hist[threadIdx.x] = 0;
        uint32_t c[PART_SVEC][4], s[PART_SVEC][4], e[PART_SVEC][4];
        if (bulk_ok(tile)) {
            mbar_wait(&mbar, phase);
            phase ^= 1u;
#pragma unroll
            for (int k = 0; k < PART_SVEC; k++) {
                const uint32_t i = 4u * ((uint32_t)k * PART_THREADS + threadIdx.x);
                const uint4 a = *reinterpret_cast<const uint4*>(s_raw + i), b = *reinterpret_cast<const uint4*>(s_raw + PART_STILE + i),
                            d = *reinterpret_cast<const uint4*>(s_raw + 2 * PART_STILE + i);
                c[k][0] = a.x; c[k][1] = a.y; c[k][2] = a.z; c[k][3] = a.w;
                s[k][0] = b.x; s[k][1] = b.y; s[k][2] = b.z; s[k][3] = b.w;
                e[k][0] = d.x; e[k][1] = d.y; e[k][2] = d.z; e[k][3] = d.w;
            }
        } else {
#pragma unroll
            for (int k = 0; k < PART_SVEC; k++) {
                const uint32_t i = base + 4u * ((uint32_t)k * PART_THREADS + threadIdx.x);
                part_load4(chr, i, end, al16, c[k]);
                part_load4(qs, i, end, al16, s[k]);
                part_load4(qe, i, end, al16, e[k]);
            }
        }
        __syncthreads();   // the histogram is clear, the staged columns are in registers
        if (threadIdx.x == 0 && bulk_ok(tile + gridDim.x)) {
            fence_proxy_async_smem();   // the threads' reads of s_raw come before the copy engine's writes
            issue(tile + gridDim.x);
        }
        uint32_t wr[PART_SVEC][4];   // window << 12 | rank inside the tile's run of that window
#pragma unroll
        for (int k = 0; k < PART_SVEC; k++)
#pragma unroll
            for (int j = 0; j < 4; j++) {
                wr[k][j] = 0xFFFFFFFFu;
                if (base + 4u * ((uint32_t)k * PART_THREADS + threadIdx.x) + j < end) {
                    uint4 m;
                    const uint32_t w = part_window(G, s_map, c[k][j], s[k][j], &m);
                    wr[k][j] = (w << 12) | atomicAdd(&s_hist[w], 1u);
                    s[k][j] = m.x + min(s[k][j], m.y);   // from here on: virtual coordinates
                    e[k][j] = m.x + min(e[k][j], m.y);
                }
            }
        static_assert(PART_STILE <= 4096, "12 bits of rank");
        __syncthreads();
        uint32_t g, ex;
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
            g = (v && w < nb) ? atomicAdd(cursor + w, v) : 0u;   // its answer is first needed after the sort below
            __syncthreads();
            uint32_t pre = 0;
            for (uint32_t k = 0; k < (w >> 5); k++) pre += s_warp[k];
            ex = pre + x - v;
            s_hist[w] = ex;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < PART_SVEC; k++)
#pragma unroll
            for (int j = 0; j < 4; j++)
                if (wr[k][j] != 0xFFFFFFFFu) {
                    const uint32_t w = wr[k][j] >> 12, lp = s_hist[w] + (wr[k][j] & 0xFFFu);
                    s_se[lp] = make_uint2(s[k][j], e[k][j]);
                    s_w[lp] = (uint16_t)w;
                }
        s_delta[threadIdx.x] = g - ex;
        __syncthreads();
        const uint32_t cnt = end - base;
        for (uint32_t i = threadIdx.x; i < cnt; i += PART_THREADS) {
            const uint32_t gp = i + s_delta[s_w[i]];
            ose[gp] = s_se[i];
        }
        __syncthreads();   // the sorted tile and the deltas are read: the next tile may overwrite them
    }
}

}  // namespace gffx
