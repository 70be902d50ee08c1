// winjoin.cuh — K1: the root-level join over window-bucketed regions, replacing
// IntervalTree::query_interval (src/utils/tree.rs:98-121) and the per-region loop of query_features
// (src/commands/intersect.rs:139-165).
//
// Index layout (winjoin_build.h; all of it L2-resident, the slice one CTA touches L1-resident):
//   ent_se[j]   (start,end) of the SHORT root intervals, per sequence sorted by start   8 B
//   sdir[w,b]   first short entry whose running max-end exceeds the first base of directory bin b of
//               window w: every earlier entry ends at or before the bin                 4 B
//   alw_*[a]    the LONG intervals (those reaching past their long_k-th successor, e.g. a
//               chromosome-spanning `region` root) listed once per window they overlap
// A region (seq, rs, re) of window w = win(seq, rs) reads sdir[w, bin(rs)], scans ent_se forward while
// start < re testing end > rs, then tests the window's long list; if it reaches into later windows it
// also takes the long intervals that START there.  Together that is exactly the set
// { iv : iv.start < re && iv.end > rs } of tree.rs:110 for any rs/re (rs >= re included).
// The chain is two dependent loads deep (sdir -> ent_se) instead of the tree's ~12 node hops, and a warp's
// 32 regions share the same few cache lines because partition.cuh put them in the same window.
//
// One fused kernel per call (wj_join_kernel below).
#pragma once
#include "scan.cuh"
#include "tilesort.cuh"

namespace gffx {

constexpr int WJ_THREADS = 256;
constexpr int WJ_QPT = 4;  // consecutive bucketed regions per thread: one 16 B load of chr, two of (start,end)
constexpr int WJ_TILE = WJ_THREADS * WJ_QPT;
constexpr int WJ_STAGE_PAIRS = 2048;  // 24 KB of shared memory

struct WinJoin {
    const uint32_t* win0;      // n_seqs+1
    const uint32_t* seq_send;  // n_seqs
    const uint2* ent_se;       // n_ent (shorts then longs)
    const uint32_t* ent_fid;   // n_ent
    const uint32_t* sdir;      // n_win << sub_bits
    const uint32_t* alw0;      // n_win+1
    const uint2* alw_se;
    const uint32_t* alw_id;
    uint32_t n_seqs, shift, sub_bits, n_ent;
};

// mode test of src/commands/intersect.rs:145-158 and the XOR of :161
template <int MODE, bool INVERT> __device__ __forceinline__ bool wj_keep(uint32_t s, uint32_t e, uint32_t rs, uint32_t re) {
    const bool keep = MODE == 0 ? (s >= rs && e <= re) : MODE == 1 ? (s <= rs && e >= re) : true;
    return INVERT != keep;
}

template <int MODE, bool INVERT, class F>
__device__ __forceinline__ void wj_query(const WinJoin& J, uint32_t c, uint32_t rs, uint32_t re, F&& f) {
    const uint32_t w0 = __ldg(J.win0 + c), nw = __ldg(J.win0 + c + 1) - w0;
    const uint32_t jend = __ldg(J.seq_send + c);
    uint32_t wl = rs >> J.shift;
    if (wl >= nw) wl = nw - 1;  // beyond the sequence's extent: its last window
    const uint32_t w = w0 + wl;
    uint32_t b = (rs - (wl << J.shift)) >> (J.shift - J.sub_bits);
    const uint32_t bmax = (1u << J.sub_bits) - 1u;
    if (b > bmax) b = bmax;
    uint32_t j = __ldg(J.sdir + (((size_t)w) << J.sub_bits) + b);
    const uint32_t a0 = __ldg(J.alw0 + w), a1 = __ldg(J.alw0 + w + 1);
    for (; j < jend; j++) {
        const uint2 se = __ldg(J.ent_se + j);
        if (se.x >= re) break;
        if (se.y > rs && wj_keep<MODE, INVERT>(se.x, se.y, rs, re)) f(j, se);
    }
    for (uint32_t a = a0; a < a1; a++) {
        const uint2 se = __ldg(J.alw_se + a);
        if (se.x < re && se.y > rs && wj_keep<MODE, INVERT>(se.x, se.y, rs, re)) f(__ldg(J.alw_id + a), se);
    }
    if (re != 0 && wl + 1 < nw) {  // long intervals that START in the later windows the region reaches into
        uint32_t wr = (re - 1u) >> J.shift;
        if (wr >= nw) wr = nw - 1;
        for (uint32_t ww = wl + 1; ww <= wr; ww++) {
            const uint32_t lo2 = ww << J.shift;
            const uint32_t b0 = __ldg(J.alw0 + w0 + ww), b1 = __ldg(J.alw0 + w0 + ww + 1);
            for (uint32_t a = b0; a < b1; a++) {
                const uint2 se = __ldg(J.alw_se + a);
                if (se.x >= lo2 && se.x < re && se.y > rs && wj_keep<MODE, INVERT>(se.x, se.y, rs, re)) f(__ldg(J.alw_id + a), se);
            }
        }
    }
}

// First pass of the fused kernel: the same search, but instead of calling back it RECORDS which candidates hit, so
// that the emit pass walks hits only (no second directory look-up, no failed candidates).  Bit i (< 16) of the mask:
// short entry j0+i hit; bit 16+i (i < 15): entry a0+i of the window's long list hit; bit 31: something did not fit
// (more candidates than bits, or a hit in a later window's long list) and the emit pass repeats the full search.
template <int MODE, bool INVERT, class F>
__device__ __forceinline__ uint32_t wj_probe(const WinJoin& J, uint32_t c, uint32_t rs, uint32_t re, uint32_t& j0, uint32_t& a0, F&& on_hit) {
    const uint32_t w0 = __ldg(J.win0 + c), nw = __ldg(J.win0 + c + 1) - w0;
    uint32_t wl = rs >> J.shift;
    if (wl >= nw) wl = nw - 1;  // beyond the sequence's extent: its last window
    const uint32_t w = w0 + wl;
    uint32_t b = (rs - (wl << J.shift)) >> (J.shift - J.sub_bits);
    const uint32_t bmax = (1u << J.sub_bits) - 1u;
    if (b > bmax) b = bmax;
    j0 = __ldg(J.sdir + (((size_t)w) << J.sub_bits) + b);
    a0 = __ldg(J.alw0 + w);
    const uint32_t a1 = __ldg(J.alw0 + w + 1);
    uint32_t mask = 0;
    uint32_t bit = 1u;
    for (uint32_t j = j0;; j++, bit <<= 1) {  // ends at the sequence's sentinel entry (start 0xFFFFFFFF) at the latest
        const uint2 se = __ldg(J.ent_se + j);
        if (se.x >= re) break;
        if (se.y > rs && wj_keep<MODE, INVERT>(se.x, se.y, rs, re)) {
            mask |= (bit & 0xFFFFu) ? bit : 0x80000000u;
            on_hit(j);
        }
    }
    bit = 1u << 16;
    for (uint32_t a = a0; a < a1; a++, bit <<= 1) {
        const uint2 se = __ldg(J.alw_se + a);
        if (se.x < re && se.y > rs && wj_keep<MODE, INVERT>(se.x, se.y, rs, re)) {
            mask |= (bit & 0x7FFF0000u) ? bit : 0x80000000u;
            on_hit(__ldg(J.alw_id + a));
        }
    }
    if (re != 0 && wl + 1 < nw) {  // long intervals that START in the later windows the region reaches into
        uint32_t wr = (re - 1u) >> J.shift;
        if (wr >= nw) wr = nw - 1;
        for (uint32_t ww = wl + 1; ww <= wr; ww++) {
            const uint32_t lo2 = ww << J.shift;
            const uint32_t b0 = __ldg(J.alw0 + w0 + ww), b1 = __ldg(J.alw0 + w0 + ww + 1);
            for (uint32_t a = b0; a < b1; a++) {
                const uint2 se = __ldg(J.alw_se + a);
                if (se.x >= lo2 && se.x < re && se.y > rs && wj_keep<MODE, INVERT>(se.x, se.y, rs, re)) {
                    mask |= 0x80000000u;
                    on_hit(__ldg(J.alw_id + a));
                }
            }
        }
    }
    return mask;
}

// thread t loads tile elements t*4 .. t*4+3 (one 16 B load of chr, two of (start,end)); returns how many exist
__device__ __forceinline__ int wj_load(const uint32_t* __restrict__ chr, const uint2* __restrict__ se, uint64_t n, uint64_t q0,
                                       uint32_t (&c)[WJ_QPT], uint32_t (&s)[WJ_QPT], uint32_t (&e)[WJ_QPT]) {
    if (q0 >= n) return 0;
    if (q0 + WJ_QPT <= n) {
        const uint4 cc = __ldcs(reinterpret_cast<const uint4*>(chr + q0));
        const uint4 a = __ldcs(reinterpret_cast<const uint4*>(se + q0));
        const uint4 b = __ldcs(reinterpret_cast<const uint4*>(se + q0) + 1);
        c[0] = cc.x; c[1] = cc.y; c[2] = cc.z; c[3] = cc.w;
        s[0] = a.x; e[0] = a.y; s[1] = a.z; e[1] = a.w;
        s[2] = b.x; e[2] = b.y; s[3] = b.z; e[3] = b.w;
        return WJ_QPT;
    }
    int m = 0;
#pragma unroll
    for (int k = 0; k < WJ_QPT; k++)
        if (q0 + k < n) {
            const uint2 v = se[q0 + k];
            c[k] = chr[q0 + k]; s[k] = v.x; e[k] = v.y;
            m = k + 1;
        }
    return m;
}

// EMIT = false: pair count + per-entry hit flags only (the CLI path needs no pairs).
// EMIT = true : also writes the pairs; a tile whose claimed range does not fit `cap` writes nothing (the host
//               grows the buffer to the reported total and runs the kernel again).
// Per tile: load -> shared-memory sort -> probe (count, hit masks, hit-root flags) -> CTA exclusive scan -> ONE global
// atomic claims the tile's output range -> the recorded hits are staged as (root_fid,start,end) in shared memory ->
// coalesced stores.
template <int MODE, bool INVERT, bool EMIT>
__global__ void __launch_bounds__(WJ_THREADS)
    wj_join_kernel(WinJoin J, const uint32_t* __restrict__ chr, const uint2* __restrict__ qse, uint64_t n, uint32_t key_shift,
                   uint8_t* ent_hit, unsigned long long* __restrict__ n_pairs, uint32_t* __restrict__ pairs, unsigned long long cap,
                   uint32_t* __restrict__ err) {
    union Smem {
        TileSortSmem<WJ_THREADS, WJ_QPT> sort;
        uint32_t stage[EMIT ? WJ_STAGE_PAIRS * 3 : 1];
    };
    __shared__ Smem U;
    __shared__ uint32_t sm[WJ_THREADS / 32 + 1];
    __shared__ unsigned long long sbase;
    const uint64_t tile0 = (uint64_t)blockIdx.x * WJ_TILE;
    const uint32_t m = (uint32_t)min((uint64_t)WJ_TILE, n - tile0);
    uint32_t c[WJ_QPT], s[WJ_QPT], e[WJ_QPT];
    bool valid[WJ_QPT];
    const int n_in = wj_load(chr, qse, n, tile0 + (uint64_t)threadIdx.x * WJ_QPT, c, s, e);
    tile_sort<WJ_THREADS, WJ_QPT>(U.sort, key_shift, m, n_in, c, s, e, valid);
    uint32_t cnt = 0;
    uint32_t mask[WJ_QPT], j0[WJ_QPT], a0[WJ_QPT];
#pragma unroll
    for (int k = 0; k < WJ_QPT; k++) {
        mask[k] = 0;
        if (!valid[k]) continue;
        if (c[k] >= J.n_seqs) { *err = 1; continue; }
        mask[k] = wj_probe<MODE, INVERT>(J, c[k], s[k], e[k], j0[k], a0[k], [&](uint32_t id) {
            cnt++;
            // ~230 regions hit every root: an L1-cached test keeps the store count O(1) per root and SM
            // (a stale 0 only repeats an idempotent store)
            if (!ent_hit[id]) ent_hit[id] = 1;
        });
    }
    uint32_t tot;
    uint32_t pos = block_scan_excl(cnt, OpSum(), 0u, sm, &tot);  // its barriers also retire U.sort before U.stage is written
    if (tot == 0) return;
    if (threadIdx.x == 0) sbase = atomicAdd(n_pairs, (unsigned long long)tot);
    if (!EMIT) return;
    __syncthreads();
    const unsigned long long gbase = sbase;
    if (gbase + tot > cap) return;
    if (tot <= WJ_STAGE_PAIRS) {
        uint32_t* dst = U.stage + 3 * pos;
#pragma unroll
        for (int k = 0; k < WJ_QPT; k++) {
            if (mask[k] & 0x80000000u) {
                wj_query<MODE, INVERT>(J, c[k], s[k], e[k], [&](uint32_t id, uint2 se) {
                    dst[0] = __ldg(J.ent_fid + id); dst[1] = se.x; dst[2] = se.y;
                    dst += 3;
                });
                continue;
            }
            for (uint32_t mm = mask[k] & 0xFFFFu; mm; mm &= mm - 1) {
                const uint32_t id = j0[k] + (uint32_t)__ffs(mm) - 1u;
                const uint2 se = __ldg(J.ent_se + id);
                dst[0] = __ldg(J.ent_fid + id); dst[1] = se.x; dst[2] = se.y;
                dst += 3;
            }
            for (uint32_t mm = mask[k] >> 16; mm; mm &= mm - 1) {
                const uint32_t a = a0[k] + (uint32_t)__ffs(mm) - 1u;
                const uint2 se = __ldg(J.alw_se + a);
                dst[0] = __ldg(J.ent_fid + __ldg(J.alw_id + a)); dst[1] = se.x; dst[2] = se.y;
                dst += 3;
            }
        }
        __syncthreads();
        uint32_t* g = pairs + 3 * gbase;
        for (uint32_t i = threadIdx.x; i < 3 * tot; i += WJ_THREADS) __stcs(g + i, U.stage[i]);
    } else {  // a tile denser than the staging buffer: the same walk straight to global memory
        uint32_t* dst = pairs + 3 * (gbase + pos);
#pragma unroll
        for (int k = 0; k < WJ_QPT; k++)
            if (mask[k])
                wj_query<MODE, INVERT>(J, c[k], s[k], e[k], [&](uint32_t id, uint2 se) {
                    dst[0] = __ldg(J.ent_fid + id); dst[1] = se.x; dst[2] = se.y;
                    dst += 3;
                });
    }
}

// per-entry hit flags -> per-block selection flags; entries whose fid has no .gof record are
// appended to warn_fids (the reference warns about them, intersect.rs:269-272).
static __global__ void wj_select_blocks_kernel(const uint8_t* __restrict__ ent_hit, const uint32_t* __restrict__ ent_blk,
                                               const uint32_t* __restrict__ ent_fid, uint32_t n_ent, uint8_t* __restrict__ blk_sel,
                                               uint32_t* __restrict__ warn_fids, uint32_t* __restrict__ n_warn, uint32_t warn_cap) {
    uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_ent || !ent_hit[j]) return;
    uint32_t b = ent_blk[j];
    if (b != 0xFFFFFFFFu) blk_sel[b] = 1;
    else {
        uint32_t k = atomicAdd(n_warn, 1u);
        if (k < warn_cap) warn_fids[k] = ent_fid[j];
    }
}

}  // namespace gffx
