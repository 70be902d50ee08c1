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
// locality than a window sort their own tile again in shared memory (tilesort.cuh).
//   part_count_kernel    tile of 4096 regions: shared-memory histogram over the windows -> column `tile`
//                        of a window-major count matrix
//   device_scan          exclusive scan of the matrix = first output position of every (window, tile)
//   part_scatter_kernel  ranks from a second shared histogram pass, tile sorted in shared memory, runs
//                        stored coalesced: (start,end) pairs and the sequence numbers
// The order inside a (tile, window) run is the arrival order of the atomics, i.e. arbitrary.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gffx {

constexpr int PART_THREADS = 512;           // 3 CTAs of 70 KB per SM: 512 threads each keep 1536 threads in flight
constexpr int PART_TILE = 4096;             // regions per CTA
constexpr int PART_UNROLL = 4;              // independent loads in flight per thread
constexpr uint32_t PART_MAX_WINDOWS = 511;  // + 1 bucket for regions whose sequence number is out of range
constexpr uint32_t PART_NB_MAX = PART_MAX_WINDOWS + 1;
constexpr size_t PART_SCATTER_SMEM = (size_t)PART_TILE * 14 + (size_t)PART_NB_MAX * 12;

struct WinGrid {
    const uint32_t* win0;  // n_seqs + 1: windows of sequence s are [win0[s], win0[s+1])
    uint32_t n_seqs, shift, n_win;
};

__device__ __forceinline__ uint32_t window_of(const WinGrid& G, uint32_t c, uint32_t s) {
    if (c >= G.n_seqs) return G.n_win;  // the join kernels report the error
    const uint32_t w0 = __ldg(G.win0 + c), nw = __ldg(G.win0 + c + 1) - w0;
    const uint32_t b = s >> G.shift;
    return w0 + (b < nw ? b : nw - 1);
}

static __global__ void __launch_bounds__(PART_THREADS)
    part_count_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, uint32_t n, uint32_t n_tiles,
                      uint32_t* __restrict__ counts) {
    __shared__ uint32_t sh[PART_NB_MAX];
    const uint32_t nb = G.n_win + 1;
    const uint32_t base = blockIdx.x * (uint32_t)PART_TILE;
    const uint32_t end = min(n, base + (uint32_t)PART_TILE);
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS) sh[i] = 0;
    __syncthreads();
    for (uint32_t i0 = base + threadIdx.x; i0 < end; i0 += PART_THREADS * PART_UNROLL) {
        uint32_t c[PART_UNROLL], s[PART_UNROLL];
#pragma unroll
        for (int k = 0; k < PART_UNROLL; k++) {
            const uint32_t i = i0 + k * PART_THREADS;
            if (i < end) { c[k] = __ldg(chr + i); s[k] = __ldg(qs + i); }
        }
#pragma unroll
        for (int k = 0; k < PART_UNROLL; k++)
            if (i0 + k * PART_THREADS < end) atomicAdd(&sh[window_of(G, c[k], s[k])], 1u);
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS) counts[(uint64_t)i * n_tiles + blockIdx.x] = sh[i];
}

static __global__ void __launch_bounds__(PART_THREADS)
    part_scatter_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, const uint32_t* __restrict__ qe,
                        uint32_t n, uint32_t n_tiles, const uint32_t* __restrict__ counts, const uint32_t* __restrict__ scanned,
                        uint32_t* __restrict__ oc, uint2* __restrict__ ose) {
    extern __shared__ uint32_t part_dyn[];  // PART_SCATTER_SMEM bytes
    uint32_t* s_c = part_dyn;
    uint32_t* s_s = s_c + PART_TILE;
    uint32_t* s_e = s_s + PART_TILE;
    uint16_t* s_w = reinterpret_cast<uint16_t*>(s_e + PART_TILE);
    uint32_t* s_loff = s_e + PART_TILE + PART_TILE / 2;
    uint32_t* s_goff = s_loff + PART_NB_MAX;
    uint32_t* s_cur = s_goff + PART_NB_MAX;
    __shared__ uint32_t s_warp[PART_THREADS / 32];
    const uint32_t nb = G.n_win + 1;
    const uint32_t base = blockIdx.x * (uint32_t)PART_TILE;
    const uint32_t end = min(n, base + (uint32_t)PART_TILE);
    // local offsets of the windows inside the sorted tile: exclusive scan of this tile's column (2 windows per thread)
    {
        const uint32_t w0 = 2 * threadIdx.x, w1 = w0 + 1;
        const uint32_t c0 = w0 < nb ? __ldg(counts + (uint64_t)w0 * n_tiles + blockIdx.x) : 0u;
        const uint32_t c1 = w1 < nb ? __ldg(counts + (uint64_t)w1 * n_tiles + blockIdx.x) : 0u;
        if (w0 < nb) s_goff[w0] = __ldg(scanned + (uint64_t)w0 * n_tiles + blockIdx.x);
        if (w1 < nb) s_goff[w1] = __ldg(scanned + (uint64_t)w1 * n_tiles + blockIdx.x);
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        uint32_t x = c0 + c1;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += o;
        }
        if (lane == 31) s_warp[warp] = x;
        __syncthreads();
        uint32_t pre = 0;
        for (int k = 0; k < warp; k++) pre += s_warp[k];
        const uint32_t ex = pre + x - (c0 + c1);
        if (w0 < nb) { s_loff[w0] = ex; s_cur[w0] = ex; }
        if (w1 < nb) { s_loff[w1] = ex + c0; s_cur[w1] = ex + c0; }
    }
    __syncthreads();
    for (uint32_t i0 = base + threadIdx.x; i0 < end; i0 += PART_THREADS * PART_UNROLL) {
        uint32_t c[PART_UNROLL], s[PART_UNROLL], e[PART_UNROLL];
#pragma unroll
        for (int k = 0; k < PART_UNROLL; k++) {
            const uint32_t i = i0 + k * PART_THREADS;
            if (i < end) { c[k] = __ldcs(chr + i); s[k] = __ldcs(qs + i); e[k] = __ldcs(qe + i); }
        }
#pragma unroll
        for (int k = 0; k < PART_UNROLL; k++)
            if (i0 + k * PART_THREADS < end) {
                const uint32_t w = window_of(G, c[k], s[k]);
                const uint32_t lp = atomicAdd(&s_cur[w], 1u);
                s_c[lp] = c[k]; s_s[lp] = s[k]; s_e[lp] = e[k]; s_w[lp] = (uint16_t)w;
            }
    }
    __syncthreads();
    const uint32_t cnt = end - base;
    for (uint32_t i = threadIdx.x; i < cnt; i += PART_THREADS) {
        const uint32_t w = s_w[i];
        const uint32_t g = s_goff[w] + (i - s_loff[w]);
        ose[g] = make_uint2(s_s[i], s_e[i]);
        oc[g] = s_c[i];
    }
}

// first bucketed position of every window: off[w] = scanned[w][0], off[nb] = n
static __global__ void part_offsets_kernel(const uint32_t* __restrict__ scanned, uint32_t nb, uint32_t n_tiles, uint32_t n,
                                           uint32_t* __restrict__ off) {
    const uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w < nb) off[w] = scanned[(uint64_t)w * n_tiles];
    else if (w == nb) off[w] = n;
}

}  // namespace gffx
