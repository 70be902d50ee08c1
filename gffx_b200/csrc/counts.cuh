// counts.cuh — K6 / K7: depth (src/commands/depth.rs:121-293) and coverage breadth
// (src/commands/coverage.rs:208-450) as a difference-array / histogram scatter over the regions,
// prefix scans, and a per-ID reduction over the units of counts_index.h.
//
//   depth     per region and space: Hs[ub(start)]++, He[lb(end)]++ (rank slots of the space's line
//             coordinates, L2-resident).  With S, E the prefix sums:  #regions overlapping
//             [X_i, X_j) = S[j] - E[i] (every region has start < end, so "ends at or before X_i" and
//             "starts at or after X_j" exclude each other).  An ID of several segments is hit by
//             sum over its segments (S[end] - E[start]) - sum over its gaps Span[gap], Span = the
//             regions that reach across the whole gap (start < a && end > b: counted by both
//             neighbours).  Until r02 the kernel counted the regions lying INSIDE each gap instead
//             (G = E[b] - S[a] + Span is the same number): every read inside an intron walked the gap
//             index and bumped a counter, half of the kernel's time (profiles/r02k_counts); a gap
//             inside a 150-base read is rare, and the walk starts at the slot the start look-up
//             already found.
//   coverage  per region and space: the partial first/last 64-base words are OR-ed into a bitmap, the
//             fully covered words in between go through a word-granular difference array; a scan turns
//             positive counts into all-ones words, a second scan gives the covered-base prefix P, and
//             covered bases in [s,e) = C(e) - C(s) with C(x) = P[word] + popc(bits below x).
#pragma once
#include "filter.cuh"
#include "rootjoin.cuh"
#include "counts_index.h"

namespace gffx {

struct CovSpace {  // 16 B, same layout as CovSpaceHost
    uint32_t lo, nbits;
    unsigned long long word0;
};

struct CountsDev {
    RankSpace R;            // spaces as "sequences": X, dir, descriptors
    const CovSpace* cov;
    const uint32_t* gsp0;       // per space: its gaps are [gsp0[s], gsp0[s+1]) of the arrays below, the last one a sentinel
    const uint32_t* gap_bits;   // per rank slot: its coordinate is the `a` of a gap (counts_index.h)
    const uint32_t* gs_first;   // per rank slot: first gap of the space with a >= the slot's coordinate
    const uint32_t* gap_a;      // gaps by (space, a), one 0xFFFFFFFF sentinel per space
    const uint32_t* gap_b;
    const uint32_t* gap_seg;    // left segment of the gap
    // tree intervals of irregular blocks (counts_index.h): by (sequence, start), with the running maximum of the ends
    const uint32_t* irr_seq0;
    const uint32_t* irr_s;
    const uint32_t* irr_e;
    const uint32_t* irr_pm;
    const uint32_t* irr_space;
    const uint32_t* irr_grp;
    const uint32_t* irr_pos;
    const uint32_t* irr_base;   // per irregular interval: base of its sequence on the virtual sequence (a private space keeps the lines' own coordinates)
    const uint32_t* grp_off;
    const uint32_t* grp_seq;
    const uint32_t* grp_s;
    const uint32_t* grp_e;
    const uint32_t* irr_blk;    // per irregular interval: its block

    uint32_t n_seqs;
    // accumulators
    uint32_t* Hs;
    uint32_t* He;
    uint32_t* G;
    unsigned long long* bits;
    int* diff;
    uint32_t* diff_used;   // set once a region went to the difference array: its scan is skipped while this is 0 (reads never get there)
    uint8_t* blk_routed;
};

// the irregular tree intervals of sequence `seq` that a region hits: start < re && end > rs (src/utils/tree.rs:110, any rs / re).
// A handful of intervals per index: the last start below re by binary search, then backwards while the running maximum of the
// ends still exceeds rs.
template <class F> __device__ __forceinline__ void irr_query(const CountsDev& D, uint32_t seq, uint32_t rs, uint32_t re, F&& f) {
    const uint32_t i0 = __ldg(D.irr_seq0 + seq), i1 = __ldg(D.irr_seq0 + seq + 1);
    if (i0 == i1) return;
    uint32_t lo = i0, hi = i1;   // first interval with start >= re
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(D.irr_s + mid) < re) lo = mid + 1; else hi = mid;
    }
    for (uint32_t j = lo; j-- > i0;) {
        if (__ldg(D.irr_pm + j) <= rs) break;
        if (__ldg(D.irr_e + j) > rs) f(j);
    }
}

__device__ __forceinline__ void depth_space_update(const CountsDev& D, uint32_t sp, uint32_t rs, uint32_t re) {
    const RankSeq S = load_rank_seq(D.R, sp);
    if (S.cnt == 0) return;
    bool eq;
    uint32_t lb = rank_lb(D.R, S, rs, &eq);
    const uint32_t ubl = lb + (eq ? 1u : 0u), ub = S.slot0 + ubl;
    atomicAdd(D.Hs + ub, 1u);                             // slot ub(start): counted by S[k] for every X_k > start
    // lb(end) = ub(start) + the coordinates inside (start, end): a short region has none or a few, next to the ones the start
    // look-up read — walk them instead of a second look-up (which a long region still does)
    uint32_t lbe = ubl;
    bool gaps = re - rs >= GAP_SHORT;   // some coordinate strictly inside the region starts a gap the region is long enough to reach across
    {
        const uint32_t* X = D.R.X + S.x0;
        int steps = 0;
        while (lbe < S.cnt && steps < 8 && __ldg(X + lbe) < re) {
            const uint32_t slot = S.slot0 + lbe;
            gaps = gaps || ((__ldg(D.gap_bits + (slot >> 5)) >> (slot & 31)) & 1u);
            lbe++; steps++;
        }
        if (steps == 8) { lbe = rank_lb(D.R, S, re, &eq); gaps = true; }
    }
    atomicAdd(D.He + S.slot0 + lbe, 1u);                  // slot lb(end):   counted by E[k] for every X_k >= end
    // gaps [a,b] the region reaches across: start < a and b < end; a is a coordinate of the space strictly inside the region
    if (gaps) {
        const uint32_t lo = __ldg(D.gs_first + ub);   // first gap with a > start: a >= X[ub(start)] (the sentinel at the latest)
        for (uint32_t j = lo;; j++) {
            if (__ldg(D.gap_a + j) >= re) break;              // sorted by a; the space's sentinel ends it
            if (__ldg(D.gap_b + j) < re) atomicAdd(D.G + __ldg(D.gap_seg + j), 1u);
        }
    }
}

constexpr unsigned long long COV_DIRECT_WORDS = 4;   // whole 64-base words a region may cover and still set them directly

__device__ __forceinline__ void cov_space_update(const CountsDev& D, uint32_t sp, uint32_t rs, uint32_t re) {
    const uint4 raw = __ldg(reinterpret_cast<const uint4*>(D.cov + sp));
    const uint32_t lo = raw.x, nbits = raw.y;
    if (nbits == 0) return;
    const unsigned long long word0 = ((unsigned long long)raw.w << 32) | raw.z;
    const uint32_t a = rs > lo ? rs : lo, hi = lo + nbits, b = re < hi ? re : hi;
    if (a >= b) return;
    const uint32_t p0 = a - lo, p1 = b - lo - 1u;  // first and last covered bit
    const unsigned long long wa = word0 + (p0 >> 6), wb = word0 + (p1 >> 6);
    const unsigned long long ma = ~0ull << (p0 & 63), mb = ~0ull >> (63 - (p1 & 63));
    // bits are only ever set during a job, so a word that already shows the mask needs no atomic
    if (wb - wa - 1 <= COV_DIRECT_WORDS || wa == wb) {
        // a short region (a read): its few whole words are OR-ed like the partial ones — once a word is full, which at
        // sequencing depth it is after the first reads, this is a load from the sector the partial words came from, where
        // the difference array costs two atomics per region whatever the bitmap already holds (profiles/r02v: 1.1 G L2
        // sectors per 100 M reads, the kernel L2-bound).  The words are read as aligned groups of four — one 32-byte sector,
        // one request (LDG.256) — so a 150-base read costs one or two L2 requests, not three or four.
        for (unsigned long long g = wa & ~3ull; g <= wb; g += 4) {
            unsigned long long q[4];
            asm volatile("ld.global.cg.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(q[0]), "=l"(q[1]), "=l"(q[2]), "=l"(q[3]) : "l"(D.bits + g));
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const unsigned long long w = g + i;
                if (w < wa || w > wb) continue;
                const unsigned long long m = (w == wa ? ma : ~0ull) & (w == wb ? mb : ~0ull);
                if ((q[i] & m) != m) atomicOr(D.bits + w, m);
            }
        }
    } else {
        if ((__ldcg(D.bits + wa) & ma) != ma) atomicOr(D.bits + wa, ma);
        if ((__ldcg(D.bits + wb) & mb) != mb) atomicOr(D.bits + wb, mb);
        atomicAdd(D.diff + wa + 1, 1);
        atomicAdd(D.diff + wb, -1);
        if (__ldcg(D.diff_used) == 0u) *D.diff_used = 1u;
    }
}

// Coverage scatter: a thread per region of the window-bucketed batch; the window gives the sequence (= space).  Regions with
// start >= end are ignored (both BED dialects of the counting commands drop them: depth.rs:488, coverage.rs:253).
// A regular block (one tree interval, its lines inside it, on the sequence's own space) needs no routing: "some region
// overlaps its tree interval" is "some base of the interval is covered", which counts_rows_kernel reads off the bitmap.  Only
// the tree intervals of irregular blocks are looked up (a handful, usually none): their private spaces take the regions
// routed to them (OR / +-1 updates tolerate repeats: no dedup needed).
constexpr uint32_t CV_IRR_CAP = 16;
constexpr uint32_t CV_THREADS = 256;
constexpr uint32_t CV_TILE = 4096;

static __global__ void __launch_bounds__(CV_THREADS)
    coverage_scatter_kernel(CountsDev D, const uint32_t* __restrict__ win_seq, const uint32_t* __restrict__ win0, uint32_t shift,
                            const uint32_t* __restrict__ poff, uint32_t n_win, const uint2* __restrict__ se, uint64_t n, uint32_t* __restrict__ err) {
    __shared__ uint32_t s_wf, s_wl;
    __shared__ uint32_t s_irr_j[CV_IRR_CAP], s_irr_s[CV_IRR_CAP], s_irr_e[CV_IRR_CAP], s_nirr;   // the irregular tree intervals that touch the window at hand
    const uint32_t tile0 = blockIdx.x * CV_TILE, m = (uint32_t)min((uint64_t)CV_TILE, n - tile0);
    {   // windows of the tile's first and last region: (number of window offsets <= position) - 1; poff has n_win + 2 cells
        int nf = 0, nl = 0;
        for (uint32_t i0 = 0; i0 < n_win + 2; i0 += CV_THREADS) {
            const uint32_t i = i0 + threadIdx.x;
            const uint32_t o = i < n_win + 2 ? __ldg(poff + i) : 0xFFFFFFFFu;
            nf += __syncthreads_count(o <= tile0);
            nl += __syncthreads_count(o <= tile0 + m - 1);
        }
        if (threadIdx.x == 0) { s_wf = (uint32_t)nf - 1u; s_wl = (uint32_t)nl - 1u; }
    }
    __syncthreads();
    for (uint32_t w = s_wf; w <= s_wl; w++) {
        const uint32_t p_lo = max(__ldg(poff + w), tile0), p_hi = min(__ldg(poff + w + 1), tile0 + m);
        if (p_lo >= p_hi) continue;
        if (w >= n_win) { if (threadIdx.x == 0) *err = 1; continue; }   // regions whose sequence number is out of range
        const uint32_t sp = __ldg(win_seq + w);
        const bool irr = __ldg(D.irr_seq0 + sp) != __ldg(D.irr_seq0 + sp + 1);
        // a region that ends inside the window can only hit irregular tree intervals that touch the window: found once per
        // (tile, window) instead of a search over the sequence's list per region (depth_window_kernel does the same)
        const uint32_t lo_w = (w - __ldg(win0 + sp)) << shift;
        const uint32_t hi_w = (w + 1 == __ldg(win0 + sp + 1)) ? 0xFFFFFFFFu : lo_w + (1u << shift);
        if (irr) {
            __syncthreads();   // the previous window's candidates are no longer read
            if (threadIdx.x == 0) {
                uint32_t k = 0;
                irr_query(D, sp, lo_w, hi_w, [&](uint32_t j) {
                    if (k < CV_IRR_CAP) { s_irr_j[k] = j; s_irr_s[k] = __ldg(D.irr_s + j); s_irr_e[k] = __ldg(D.irr_e + j); }
                    k++;
                });
                s_nirr = k;
            }
            __syncthreads();
        }
        for (uint32_t p = p_lo + threadIdx.x; p < p_hi; p += CV_THREADS) {
            const uint2 r = __ldcs(se + p);
            if (r.x >= r.y) continue;
            cov_space_update(D, sp, r.x, r.y);
            if (irr) {
                auto hit = [&](uint32_t j) {
                    const uint32_t blk = __ldg(D.irr_blk + j);
                    if (!D.blk_routed[blk]) D.blk_routed[blk] = 1;
                    const uint32_t vb = __ldg(D.irr_base + j);
                    cov_space_update(D, __ldg(D.irr_space + j), r.x - vb, r.y - vb);
                };
                const uint32_t nirr = s_nirr;
                if (nirr <= CV_IRR_CAP && r.y <= hi_w) {
                    for (uint32_t k = 0; k < nirr; k++)
                        if (s_irr_s[k] < r.y && s_irr_e[k] > r.x) hit(s_irr_j[k]);
                } else {
                    irr_query(D, sp, r.x, r.y, hit);
                }
            }
        }
    }
}

// ---- depth, one CTA per RUN of a window's reads (r02) ----------------------------------------------------------------------
// A thread per read against global memory (r01's counts_scatter_kernel<0>) pays ~3.7 scattered L2 operations per read (two histogram atomics, the look-up's sectors): 2.4 ms per
// 100 M reads, the L2's rate for that many operations (profiles/exp_atomics.cu).  After K0 a window's reads are
// contiguous, and a window has only a few thousand rank slots — so a CTA takes a RUN of up to 61 440 reads of ONE window, brings
// the window's part of the coordinate set into shared memory (the bulk copies of rootjoin.cuh), counts into two shared-memory
// histograms over the window's slots (16-bit cells: a run cannot overflow them) and adds the non-zero cells to the global
// histograms once: ~0.25 L2 atomics per read instead of two, no global look-ups.  Reads whose end lies beyond the window's part,
// gap updates (rare, counts.cuh above) and the irregular blocks' join keep their global paths; a window whose part does not fit
// (WR_SLICE_CAP values) is done read by read with depth_space_update.
// the counts coordinate set is twice as dense as the line filter's (every counted line's start and end): its windows' offset tables
// take one more bit of fine bin, so that a look-up's binary search has as few steps as the join's
#ifndef GFFX_DW_FB
#define GFFX_DW_FB (GFFX_WR_FB + 1)
#endif
constexpr uint32_t DW_FB = GFFX_DW_FB, DW_NBF = 1u << DW_FB;
constexpr uint32_t DW_XOFF_STRIDE = DW_NBF + 8;   // u16 cells per window: DW_NBF + 1 used, rows 16-byte multiples
constexpr uint32_t DW_THREADS = 512;
constexpr uint32_t DW_RUN = 15u * 4096u;   // reads per CTA: below 2^16 (16-bit histogram cells), a multiple of the partition tile
constexpr uint32_t DW_GAPW = (WR_SLICE_CAP + 8) / 32 + 2;   // words of the window's gap-start bits
constexpr size_t DW_SMEM = (size_t)(WR_SLICE_CAP + 8) * 4 + (size_t)DW_XOFF_STRIDE * 2 + 2 * (size_t)(WR_SLICE_CAP + 8) * 2 + (size_t)DW_GAPW * 4;

constexpr uint32_t DW_IRR_CAP = 16;   // irregular tree intervals of one window kept in shared memory (more: every read searches the sequence's list)

struct DepthWin {
    const uint32_t* win_seq;   // window -> (virtual) sequence = space
    const uint32_t* win0;      // first window of every sequence
    const uint32_t* poff;      // n_win + 2 window offsets of the bucketed reads
    uint32_t shift, n_win;
    RankSlices XS;             // the windows' parts of the COUNTS coordinate set (D.R.X)
    bool any_irregular;
};

static __global__ void __launch_bounds__(DW_THREADS, 2)
    depth_window_kernel(CountsDev D, DepthWin W, const uint2* __restrict__ se, uint32_t* __restrict__ err) {
    extern __shared__ __align__(16) uint32_t dw_dyn[];
    uint32_t* const sx_raw = dw_dyn;
    uint16_t* const soff = reinterpret_cast<uint16_t*>(sx_raw + WR_SLICE_CAP + 8);
    uint32_t* const hs = reinterpret_cast<uint32_t*>(soff + DW_XOFF_STRIDE);            // pairs of 16-bit cells
    uint32_t* const he = hs + (WR_SLICE_CAP + 8) / 2;
    uint32_t* const sgap = he + (WR_SLICE_CAP + 8) / 2;   // bit j: the part's value j starts a gap (counts_index.h: gap_bits)
    __shared__ uint32_t scan[DW_THREADS / 32 + 1];
    __shared__ uint32_t s_w, s_run;
    __shared__ uint32_t s_irr_j[DW_IRR_CAP], s_irr_s[DW_IRR_CAP], s_irr_e[DW_IRR_CAP], s_nirr;   // the irregular tree intervals that touch this window
    __shared__ __align__(8) uint64_t mbar;
    const uint32_t tid = threadIdx.x;
    // run table: a block scan of the windows' run counts (n_win <= 511: a window per thread)
    const uint32_t cnt_w = tid < W.n_win ? __ldg(W.poff + tid + 1) - __ldg(W.poff + tid) : 0u, nr = (cnt_w + DW_RUN - 1) / DW_RUN;
    uint32_t total;
    const uint32_t r0 = block_scan_excl(nr, OpSum(), 0u, scan, &total);
    if (tid == 0 && __ldg(W.poff + W.n_win + 1) != __ldg(W.poff + W.n_win) && blockIdx.x == 0) *err = 1;   // reads in the bucket of unknown sequences
    if (blockIdx.x >= total) return;
    if (nr != 0 && r0 <= blockIdx.x && blockIdx.x < r0 + nr) { s_w = tid; s_run = blockIdx.x - r0; }
    if (tid == 0) mbar_init(&mbar, 1);
    __syncthreads();
    const uint32_t w = s_w;
    const uint32_t p_lo = __ldg(W.poff + w) + s_run * DW_RUN, p_hi = min(__ldg(W.poff + w + 1), p_lo + DW_RUN);
    const uint32_t sp = __ldg(W.win_seq + w);
    const uint32_t lo = (w - __ldg(W.win0 + sp)) << W.shift, fsh = W.shift - DW_FB;
    const RankSeq RS = load_rank_seq(D.R, sp);
    // a read that ends inside the window can only hit irregular tree intervals that touch the window: usually none, a few at most —
    // found once per CTA instead of a binary search over the sequence's list per read (10 % of the kernel's instructions, r03j)
    const uint32_t hi_w = (w + 1 == __ldg(W.win0 + sp + 1)) ? 0xFFFFFFFFu : lo + (1u << W.shift);   // the last window takes every later start
    if (W.any_irregular && tid == 0) {
        uint32_t k = 0;
        irr_query(D, sp, lo, hi_w, [&](uint32_t j) {
            if (k < DW_IRR_CAP) { s_irr_j[k] = j; s_irr_s[k] = __ldg(D.irr_s + j); s_irr_e[k] = __ldg(D.irr_e + j); }
            k++;
        });
        s_nirr = k;
    }
    if (W.any_irregular) __syncthreads();
    const uint32_t xlen = __ldg(W.XS.xlen + w), xw0 = __ldg(W.XS.xw0 + w);
    const bool sliced = RS.cnt != 0 && xlen != 0xFFFFFFFFu;
    if (sliced) {
        if (tid == 0) {
            const uint32_t bytes = (((xw0 & 3u) + xlen + 1u) * 4u + 15u) & ~15u;   // X is padded: reading a few values on is safe
            mbar_expect_tx(&mbar, bytes + DW_XOFF_STRIDE * 2u);
            bulk_g2s(sx_raw, D.R.X + (xw0 & ~3u), bytes, &mbar);
            bulk_g2s(soff, W.XS.xoff + (size_t)w * DW_XOFF_STRIDE, DW_XOFF_STRIDE * 2u, &mbar);
        }
        for (uint32_t i = tid; i < (xlen + 2u + 1u) / 2u; i += DW_THREADS) { hs[i] = 0; he[i] = 0; }
        {   // the gap-start bits of the part's slots, realigned to bit 0 = the part's first value
            const uint32_t s0 = RS.slot0 + (xw0 - RS.x0), sh = s0 & 31u;
            for (uint32_t i = tid; i < (xlen + 31u) / 32u; i += DW_THREADS) {
                const uint32_t a = __ldg(D.gap_bits + (s0 >> 5) + i), b = __ldg(D.gap_bits + (s0 >> 5) + i + 1);   // the array has a spare word
                sgap[i] = __funnelshift_r(a, b, sh);
            }
        }
        __syncthreads();
        mbar_wait(&mbar, 0);
    }
    const uint32_t* sx = sx_raw + (xw0 & 3u);
    const uint32_t rank0 = xw0 - RS.x0;   // rank of the part's first value within the space
    uint2 r_next = p_lo + tid < p_hi ? __ldcs(se + p_lo + tid) : make_uint2(0, 0);
    for (uint32_t p = p_lo + tid; p < p_hi; p += DW_THREADS) {
        const uint2 r = r_next;   // the next read is already on its way while this one is counted
        if (p + DW_THREADS < p_hi) r_next = __ldcs(se + p + DW_THREADS);
        const uint32_t rs = r.x, re = r.y;
        if (rs >= re) continue;
        if (!sliced) {
            depth_space_update(D, sp, rs, re);
        } else {
            bool eq;
            const uint32_t lb = rank_lb_slice(sx, soff, xlen, min((rs - lo) >> fsh, DW_NBF - 1u), rs, &eq);
            const uint32_t ub = lb + (eq ? 1u : 0u);   // <= xlen: the slot after the part's last value is the next window's first
            atomicAdd(hs + (ub >> 1), (ub & 1u) ? 0x10000u : 1u);
            uint32_t j = ub;
            bool gaps = re - rs >= GAP_SHORT;
            while (j < xlen && sx[j] < re) {
                gaps = gaps || ((sgap[j >> 5] >> (j & 31u)) & 1u);
                j++;
            }
            if (j < xlen || rank0 + xlen == RS.cnt) atomicAdd(he + (j >> 1), (j & 1u) ? 0x10000u : 1u);
            else {   // the read runs past the window's part: its end's rank lies in a later window
                bool eq2;
                atomicAdd(D.He + RS.slot0 + rank_lb(D.R, RS, re, &eq2), 1u);
                gaps = true;
            }
            if (gaps)
                for (uint32_t g = __ldg(D.gs_first + RS.slot0 + rank0 + ub);; g++) {
                    if (__ldg(D.gap_a + g) >= re) break;              // sorted by a; the space's sentinel ends it
                    if (__ldg(D.gap_b + g) < re) atomicAdd(D.G + __ldg(D.gap_seg + g), 1u);
                }
        }
        if (W.any_irregular) {
            auto hit = [&](uint32_t j) {
                const uint32_t g = __ldg(D.irr_grp + j);   // a root fid with several tree intervals is routed once per region (depth.rs:239-241)
                if (g != 0xFFFFFFFFu) {
                    const uint32_t m0 = __ldg(D.grp_off + g), m1 = m0 + __ldg(D.irr_pos + j);
                    for (uint32_t m = m0; m < m1; m++)
                        if (__ldg(D.grp_seq + m) == sp && __ldg(D.grp_s + m) < re && __ldg(D.grp_e + m) > rs) return;
                }
                const uint32_t vb = __ldg(D.irr_base + j);
                depth_space_update(D, __ldg(D.irr_space + j), rs - vb, re - vb);
            };
            const uint32_t nirr = s_nirr;
            if (nirr <= DW_IRR_CAP && re <= hi_w) {   // start >= the window's first base: only the window's own candidates can be hit
                for (uint32_t k = 0; k < nirr; k++)
                    if (s_irr_s[k] < re && s_irr_e[k] > rs) hit(s_irr_j[k]);
            } else {
                irr_query(D, sp, rs, re, hit);
            }
        }
    }
    if (!sliced) return;
    __syncthreads();
    for (uint32_t i = tid; i <= xlen; i += DW_THREADS) {   // the run's counts into the global histograms: one atomic per non-zero cell
        const uint32_t a = (hs[i >> 1] >> (16u * (i & 1u))) & 0xFFFFu, b = (he[i >> 1] >> (16u * (i & 1u))) & 0xFFFFu;
        if (a) atomicAdd(D.Hs + RS.slot0 + rank0 + i, a);
        if (b) atomicAdd(D.He + RS.slot0 + rank0 + i, b);
    }
}

// ---- finish -----------------------------------------------------------------------------------
struct DiffLoad {
    const int* d;
    __device__ int operator()(uint64_t i) const { return d[i]; }
};
struct CoverStore {  // inclusive prefix of the word difference array > 0  =>  the whole word is covered
    unsigned long long* bits;
    __device__ void operator()(uint64_t i, int v) const {
        if (v > 0) bits[i] = ~0ull;
    }
};
// covered bases per GROUP of eight bitmap words (one 64-byte line): the prefix over groups is an eighth of a per-word prefix
// to write and to read back, and a query adds the few words of its own group, which share the line of the word it reads anyway
struct PopcLoad8 {
    const unsigned long long* bits;   // padded to whole groups
    __device__ uint32_t operator()(uint64_t g) const {
        const ulonglong2* p = reinterpret_cast<const ulonglong2*>(bits + 8 * g);
        uint32_t c = 0;
#pragma unroll
        for (int k = 0; k < 4; k++) { const ulonglong2 v = p[k]; c += (uint32_t)__popcll(v.x) + (uint32_t)__popcll(v.y); }
        return c;
    }
};

struct CountsFin {
    const uint32_t* id_unit0;
    const uint32_t* id_units;
    const uint32_t* u_space;
    const uint32_t* u_chrom;
    const uint32_t* u_blk;
    const uint32_t* u_base;
    const uint32_t* u_rs;       // the one tree interval of a regular block's units (0xFFFFFFFF: irregular block)
    const uint32_t* u_re;
    const uint32_t* u_line0;
    const uint32_t* u_seg0;
    const uint32_t* l_fs;
    const uint32_t* l_fe;
    const uint32_t* l_sfs;
    const uint32_t* l_sfe;
    const uint32_t* g_s;
    const uint32_t* g_e;
    const uint32_t* g_ss;
    const uint32_t* g_se;
    const CovSpace* cov;
    // scanned accumulators
    const uint32_t* S;
    const uint32_t* E;
    const uint32_t* G;
    const unsigned long long* bits;
    const uint32_t* P;          // exclusive prefix of the covered bases per group of eight words
    const uint8_t* blk_routed;
    uint32_t n_ids;
};

struct RowArrays {
    uint32_t* chrom;
    uint32_t* start;
    uint32_t* end;
    unsigned long long* value;
    uint8_t* flag;
};

__device__ __forceinline__ uint32_t covered_before(const CountsFin& F, const CovSpace& cv, uint32_t x) {
    const uint32_t p = x - cv.lo;
    const unsigned long long w = cv.word0 + (p >> 6), g = w >> 3;
    const ulonglong2* q = reinterpret_cast<const ulonglong2*>(F.bits + 8 * g);
    uint32_t c = F.P[g];
    const uint32_t k = (uint32_t)(w & 7u);   // whole words of the group below w, then the bits of w below x
#pragma unroll
    for (uint32_t i = 0; i < 4; i++) {
        const ulonglong2 v = q[i];
        const uint32_t a = 2 * i, b = 2 * i + 1;
        c += (uint32_t)__popcll(a < k ? v.x : a == k ? v.x & ((1ull << (p & 63)) - 1ull) : 0ull);
        c += (uint32_t)__popcll(b < k ? v.y : b == k ? v.y & ((1ull << (p & 63)) - 1ull) : 0ull);
    }
    return c;
}

// one thread per ID: reduce its units (ascending block order) into one row
template <int KIND>
__global__ void __launch_bounds__(256) counts_rows_kernel(CountsFin F, RowArrays R) {
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= F.n_ids) return;
    unsigned long long value = 0;
    uint32_t mn = 0xFFFFFFFFu, mx = 0, chrom = 0xFFFFFFFFu;
    bool any = false;
    for (uint32_t k = F.id_unit0[id]; k < F.id_unit0[id + 1]; k++) {
        const uint32_t u = F.id_units[k];
        const uint32_t s0 = F.u_seg0[u], s1 = F.u_seg0[u + 1];
        const uint32_t vb = F.u_base[u];   // rows carry the sequence's own coordinates
        if (KIND == 0) {
            uint32_t d = 0;   // mod 2^32; the true value is below 2^32
            for (uint32_t g = s0; g < s1; g++) d += F.S[F.g_se[g]] - F.E[F.g_ss[g]];   // regions overlapping segment g
            for (uint32_t g = s0; g + 1 < s1; g++) d -= F.G[g];                         // ... once, not per segment they reach
            if (d == 0) continue;
            value += d;
            any = true;
            if (chrom == 0xFFFFFFFFu) chrom = F.u_chrom[u];
            // hull over the lines that were hit (depth.rs:198-199)
            for (uint32_t l = F.u_line0[u]; l < F.u_line0[u + 1]; l++)
                if (F.S[F.l_sfe[l]] - F.E[F.l_sfs[l]] != 0) {
                    mn = min(mn, F.l_fs[l] - vb);
                    mx = max(mx, F.l_fe[l] - vb);
                }
        } else {
            const CovSpace cv = F.cov[F.u_space[u]];
            const uint32_t rs = F.u_rs[u];
            // routed: some region overlaps the block's tree interval — covered bases inside it (regular), the scatter's flag (irregular)
            if (rs != 0xFFFFFFFFu ? covered_before(F, cv, F.u_re[u]) == covered_before(F, cv, rs) : !F.blk_routed[F.u_blk[u]]) continue;
            any = true;
            if (chrom == 0xFFFFFFFFu) chrom = F.u_chrom[u];
            for (uint32_t g = s0; g < s1; g++) value += (uint32_t)(covered_before(F, cv, F.g_e[g]) - covered_before(F, cv, F.g_s[g]));
            mn = min(mn, F.g_s[s0] - vb);       // hull over ALL lines (coverage.rs:345-346): segments are sorted and disjoint
            mx = max(mx, F.g_e[s1 - 1] - vb);
        }
    }
    R.flag[id] = any ? 1 : 0;
    if (any) {
        R.chrom[id] = chrom;
        R.start[id] = mn == 0xFFFFFFFFu ? 0u : mn;
        R.end[id] = mx;
        R.value[id] = value;
    }
}

struct FlagLoad {
    const uint8_t* f;
    __device__ uint32_t operator()(uint64_t i) const { return f[i]; }
};

// order-preserving compaction of the flagged rows (ID dictionary order)
static __global__ void __launch_bounds__(256)
    counts_compact_kernel(RowArrays R, const uint32_t* __restrict__ pos, uint32_t n_ids, uint32_t* __restrict__ o_id,
                          uint32_t* __restrict__ o_chrom, uint32_t* __restrict__ o_start, uint32_t* __restrict__ o_end,
                          unsigned long long* __restrict__ o_value) {
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n_ids || !R.flag[id]) return;
    const uint32_t p = pos[id];
    o_id[p] = id;
    o_chrom[p] = R.chrom[id];
    o_start[p] = R.start[id];
    o_end[p] = R.end[id];
    o_value[p] = R.value[id];
}

}  // namespace gffx
