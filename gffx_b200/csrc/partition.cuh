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
// Cost model (B300_MICROARCH.md): a shared-memory atomic costs ~2 cycles per lane and a divergent global
// store ~1 cycle per lane, against ~0.03 cycles per lane for coalesced traffic, so the pass is built to need
// exactly ONE shared atomic and ONE scattered 8-byte store per region:
//   part_rank_kernel     tile of 16384 regions: shared-memory histogram over the windows; the value each
//                        atomicAdd returns IS the region's rank inside (tile, window) and is kept (u16,
//                        coalesced); the tile's counts go to column `tile` of a window-major matrix
//   device_scan          exclusive scan of the matrix = first output position of every (window, tile)
//   part_scatter_kernel  position = scanned[window][tile] + rank: no atomics, one (start,end) store; the
//                        sequence number goes to a second array for the kernels that want it
// The order inside a (tile, window) run is the arrival order of the atomics, i.e. arbitrary.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gffx {

constexpr int PART_THREADS = 512;
constexpr int PART_TILE = 16384;             // regions per CTA: ranks fit 16 bits
constexpr uint32_t PART_MAX_WINDOWS = 4095;  // + 1 bucket for regions whose sequence number is out of range

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
    part_rank_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, uint32_t n, uint32_t n_tiles,
                     uint16_t* __restrict__ rank, uint32_t* __restrict__ counts) {
    extern __shared__ uint32_t sh[];
    const uint32_t nb = G.n_win + 1;
    const uint32_t base = blockIdx.x * (uint32_t)PART_TILE;
    const uint32_t end = min(n, base + (uint32_t)PART_TILE);
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS) sh[i] = 0;
    __syncthreads();
    for (uint32_t i = base + threadIdx.x; i < end; i += PART_THREADS)
        rank[i] = (uint16_t)atomicAdd(&sh[window_of(G, __ldcs(chr + i), __ldcs(qs + i))], 1u);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS) counts[(uint64_t)i * n_tiles + blockIdx.x] = sh[i];
}

static __global__ void __launch_bounds__(PART_THREADS)
    part_scatter_kernel(WinGrid G, const uint32_t* __restrict__ chr, const uint32_t* __restrict__ qs, const uint32_t* __restrict__ qe,
                        uint32_t n, uint32_t n_tiles, const uint16_t* __restrict__ rank, const uint32_t* __restrict__ scanned,
                        uint32_t* __restrict__ oc, uint2* __restrict__ ose) {
    extern __shared__ uint32_t sh[];
    const uint32_t nb = G.n_win + 1;
    const uint32_t base = blockIdx.x * (uint32_t)PART_TILE;
    const uint32_t end = min(n, base + (uint32_t)PART_TILE);
    for (uint32_t i = threadIdx.x; i < nb; i += PART_THREADS) sh[i] = __ldg(scanned + (uint64_t)i * n_tiles + blockIdx.x);
    __syncthreads();
    for (uint32_t i = base + threadIdx.x; i < end; i += PART_THREADS) {
        const uint32_t c = __ldcs(chr + i), s = __ldcs(qs + i), e = __ldcs(qe + i);
        const uint32_t p = sh[window_of(G, c, s)] + __ldcs(rank + i);
        ose[p] = make_uint2(s, e);
        oc[p] = c;
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
