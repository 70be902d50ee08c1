// rootjoin.cuh — K1 in its root-parallel form: the same per-root hit counts as wj_count_kernel (winjoin.cuh), i.e. the result
// of IntervalTree::query_interval (src/utils/tree.rs:98-121, predicate :110) filtered by the mode test of query_features
// (src/commands/intersect.rs:145-161), with the roles swapped: the ROOTS walk the regions.
//
// Why: what a region contributes is "+1 on every root it hits", and with 10 M regions against ~120 roots per window the
// region-parallel rounds of wj_count_kernel spend ~7 warp instructions per region on candidate loops whose result is a
// handful of counters (profiles/r02f: 52 % of the kernel's instructions).  Here a (tile, window) segment of regions is
// counting-sorted by FINE bin of its start in shared memory (2^10 bins per window, ~4 regions per bin and tile), and every root
// of the window gets its count from that sorted tile:
//   * a region whose start lies in a bin strictly inside the root, (bin(s), bin(e-1)), overlaps it (its end is beyond its
//     start): those come from the histogram's prefix sums, no region is looked at;
//   * the bins around the root's two ends — [bin(s - maxlen), bin(s)] and bin(e-1), maxlen the longest region of the segment —
//     hold the only regions whose answer depends on their exact coordinates: one warp tests them 32 at a time with the
//     literal predicate.  --contained needs only the left range (the region must start at or before s), --contains-region
//     counts the inside bins up to bin(e - maxlen) and tests the rest, -I is overlap minus the mode count.
// A root costs ~3 warp iterations however many regions hit it.  What the scheme assumes of a region — start < end and a
// length of at most WR_TAIL_BINS fine bins — is decided per region: the others (inverted, empty, very long regions) are
// sorted into a tail part of the same tile and resolved region-parallel against the bin lists (wj_round_global, the rule
// of wj_small_segment), so any input gives the reference's counts; roots with end <= start test every region of the tile.
#pragma once
#include "bulkcopy.cuh"
#include "winjoin.cuh"

namespace gffx {

#ifndef GFFX_WR_FB
#define GFFX_WR_FB 10
#endif
constexpr int WR_FB = GFFX_WR_FB;           // fine bins per window: 2^WR_FB (the partition grid's shift is >= 12)
constexpr uint32_t WR_NBF = 1u << WR_FB;
constexpr uint32_t WR_TAIL_BINS = 8;        // a region longer than this many fine bins belongs to the tail part
constexpr uint32_t WR_SERIAL_MAX = 48;      // a root with more positions to test than this gets the whole warp
constexpr int WR_CPT = (int)(2 * WR_NBF / WJ_THREADS);   // counters per thread in the histogram scan
static_assert(WR_CPT * WJ_THREADS == 2 * WR_NBF && (WR_CPT == 2 || WR_CPT == 4), "the histogram scan gives every thread two or four counters");

struct WrSmem {
    // regions per key (fine bin; + WR_NBF for the tail part) -> first sorted position; [2 * WR_NBF] = all.  16-bit cells (a tile has
    // 4096 regions), counted by 32-bit shared atomics on the word that holds two of them: 4 KB instead of 8, which is what lets
    // three CTAs share an SM next to the rank slice
    uint16_t hist[2 * WR_NBF + 8];
    uint32_t scan[WJ_WARPS + 1];
    unsigned long long tot;
    uint32_t w_first, w_last;
    uint32_t maxlen, re_max;         // over the normal part: longest region, largest end
    uint32_t next_tail, next_scat;   // tail rounds and scatter chunks are handed to the warps in arrival order
};

// mode test of src/commands/intersect.rs:145-158 alone
template <int MODE> __device__ __forceinline__ bool wj_mode(uint32_t s, uint32_t e, uint32_t rs, uint32_t re) {
    return MODE == 0 ? (s >= rs && e <= re) : MODE == 1 ? (s <= rs && e >= re) : true;
}

// one round of <= 32 regions of window w (lane t holds region t; a lane without one holds (0xFFFFFFFF, 0)) against the bin
// lists, straight from global memory: the list of the smallest bin among them, then the run while start < the largest end
template <int MODE, bool INVERT>
__device__ __noinline__ unsigned long long wj_round_global(const WinBins& C, uint32_t w, uint32_t lo, uint2 r, bool have, uint32_t lane, uint8_t* ent_hit,
                                                           unsigned long long* ent_cnt) {
    const uint32_t nb = 1u << C.sub, bsh = C.shift - C.sub;
    const size_t row = (size_t)w * (nb + 1);
    const uint32_t b = __reduce_min_sync(0xffffffffu, have ? min((r.x - lo) >> bsh, nb - 1) : 0xFFFFFFFFu);
    const uint32_t o0 = __ldg(C.boff + row + b), o1 = __ldg(C.boff + row + b + 1);
    unsigned long long tot = 0;
    for (uint32_t i = o0; i < o1; i++) {
        const uint4 v = __ldg(C.list4 + i);
        const bool hit = v.x < r.y && v.y > r.x && wj_keep<MODE, INVERT>(v.x, v.y, r.x, r.y);
        const uint32_t nh = __popc(__ballot_sync(0xffffffffu, hit));
        if (nh && lane == 0) { atomicAdd(ent_cnt + v.w, (unsigned long long)nh); ent_hit[v.w] = 1; }
        tot += nh;
    }
    const uint32_t re_max = __reduce_max_sync(0xffffffffu, r.y);
    for (uint32_t j = __ldg(C.first + row + b + 1);; j++) {   // the sequence's sentinel (start 0xFFFFFFFF) ends it at the latest
        const uint4 v = __ldg(C.ent4 + j);
        if (v.x >= re_max) break;
        const bool hit = v.x < r.y && v.y > r.x && wj_keep<MODE, INVERT>(v.x, v.y, r.x, r.y);
        const uint32_t nh = __popc(__ballot_sync(0xffffffffu, hit));
        if (nh && lane == 0) { atomicAdd(ent_cnt + j, (unsigned long long)nh); ent_hit[j] = 1; }
        tot += nh;
    }
    return tot;
}

// ---- the window's slice of the rank space in shared memory ---------------------------------------------------------------
// The line filter's region scatter (filter.cuh) needs lb(qs) = #(X < qs) over the sequence's coordinate set X.  From global
// memory that is a directory cell and then the values it points at: two dependent L2 (often DRAM) round trips per region,
// 44 % of this kernel's stall samples (profiles/r02i).  A (tile, window) segment only ever asks about ONE window's part of X
// (a few thousand values), so the CTA brings that part and a per-fine-bin offset table into shared memory with two bulk
// asynchronous copies (cp.async.bulk, completion on an mbarrier: the copy engine moves the bytes while the warps sort the
// tile) and the look-up becomes a short binary search in shared memory.  A window whose part exceeds WR_SLICE_CAP values
// keeps the global look-up.
#ifndef GFFX_WR_SLICE_CAP
#define GFFX_WR_SLICE_CAP 9216
#endif
constexpr uint32_t WR_SLICE_CAP = GFFX_WR_SLICE_CAP;   // values of X a CTA can hold (36 KB)
constexpr uint32_t WR_XOFF_STRIDE = WR_NBF + 8;   // u16 cells per window in the offset table: WR_NBF + 1 used, rows 16-byte multiples
constexpr size_t WR_SLICE_SMEM = (size_t)(WR_SLICE_CAP + 8) * 4 + (size_t)WR_XOFF_STRIDE * 2;
struct RankSlices {
    const uint32_t* xw0;    // n_win: index into RankSpace::X of the window's first value
    const uint32_t* xlen;   // n_win: values in the window's part; 0xFFFFFFFF: too many for shared memory
    const uint16_t* xoff;   // n_win * WR_XOFF_STRIDE: [b] = values of the part below fine bin b's first base, [WR_NBF] = xlen
};
// lb = #(part < v) for a v of fine bin b; *eq = (part[lb] == v).  sx[k] = value k of the window's part
__device__ __forceinline__ uint32_t rank_lb_slice(const uint32_t* sx, const uint16_t* soff, uint32_t xlen, uint32_t b, uint32_t v, bool* eq) {
    uint32_t lo = soff[b], hi = soff[b + 1];
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (sx[mid] < v) lo = mid + 1; else hi = mid;
    }
    *eq = lo < xlen && sx[lo] == v;   // beyond the part: the next value lies in a later window, above v
    return lo;
}

#ifndef WR_MIN_CTAS
#define WR_MIN_CTAS 3
#endif
template <int MODE, bool INVERT, bool FILT, bool SLICE>
#ifndef WR_SLICE_CTAS
#define WR_SLICE_CTAS 3
#endif
__global__ void __launch_bounds__(WJ_THREADS, SLICE ? WR_SLICE_CTAS : WR_MIN_CTAS)
    wj_root_kernel(WinBins C, const uint32_t* __restrict__ poff, const uint2* __restrict__ qse, uint64_t n, uint8_t* ent_hit,
                   unsigned long long* ent_cnt, unsigned long long* __restrict__ n_pairs, uint32_t* __restrict__ err, RankSpace R, FilterArrays F,
                   uint32_t tail_bins, RankSlices XS) {
    extern __shared__ __align__(16) uint2 wj_sorted[];   // WJ_DYN_SMEM bytes: the (tile, window) regions ordered by key; SLICE: + WR_SLICE_SMEM
    __shared__ __align__(16) WrSmem S;
    __shared__ __align__(8) uint64_t mbar;
    uint32_t* const sx_raw = reinterpret_cast<uint32_t*>(wj_sorted + WJ_TILE);                  // the window's part of X, from a 16-byte boundary
    uint16_t* const soff = reinterpret_cast<uint16_t*>(sx_raw + WR_SLICE_CAP + 8);
    uint32_t mbar_phase = 0;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t tile0 = blockIdx.x * (uint32_t)WJ_TILE;
    const uint32_t m = (uint32_t)min((uint64_t)WJ_TILE, n - tile0);
    const uint32_t n_win = C.n_win, fsh = C.shift - WR_FB, nb_old = 1u << C.sub;
    const uint32_t taillen = (uint32_t)min((unsigned long long)tail_bins << fsh, 0xFFFFFFFFull);   // tail_bins: WR_TAIL_BINS unless a test says otherwise
    const bool al16 = (reinterpret_cast<uintptr_t>(qse) & 15u) == 0;
    {   // windows of the tile's first and last region: (number of window offsets <= position) - 1; poff has n_win + 2 cells
        int nf = 0, nl = 0;
        for (uint32_t i0 = 0; i0 < n_win + 2; i0 += WJ_THREADS) {
            const uint32_t i = i0 + tid;
            const uint32_t o = i < n_win + 2 ? __ldg(poff + i) : 0xFFFFFFFFu;
            nf += __syncthreads_count(o <= tile0);
            nl += __syncthreads_count(o <= tile0 + m - 1);
        }
        if (tid == 0) {
            S.w_first = (uint32_t)nf - 1u; S.w_last = (uint32_t)nl - 1u;
            if (SLICE) mbar_init(&mbar, 1);
        }
    }
    __syncthreads();
    const uint32_t w_first = S.w_first, w_last = S.w_last;
    for (uint32_t w = w_first; w <= w_last; w++) {
        const uint32_t p_lo = max(__ldg(poff + w), tile0), p_hi = min(__ldg(poff + w + 1), tile0 + m);
        if (p_lo >= p_hi) continue;   // empty window (uniform)
        if (w >= n_win) {             // bucket of the regions whose sequence number is out of range
            if (tid == 0) *err = 1;
            continue;
        }
        const uint32_t seq = __ldg(C.win_seq + w);
        const uint32_t lo = (w - __ldg(C.win0 + seq)) << C.shift;
        const bool is_last = w + 1 == __ldg(C.win0 + seq + 1);   // its last bin also holds the regions that start beyond the grid
        RankSeq RS{};
        bool filt = false, use_slice = false;   // no line of a sequence without rank space can ever match
        uint32_t xlen = 0, xw0 = 0;
        if (FILT) {
            RS = load_rank_seq(R, seq);
            filt = RS.canon && RS.cnt != 0;
            if (SLICE && filt) {
                xlen = __ldg(XS.xlen + w);
                xw0 = __ldg(XS.xw0 + w);
                use_slice = xlen != 0xFFFFFFFFu;
                if (use_slice && tid == 0) {   // the previous window's readers of the buffers are behind the barrier that ended it
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    const uint32_t bytes = (((xw0 & 3u) + xlen + 1u) * 4u + 15u) & ~15u;   // X is padded: reading a few values on is safe
                    mbar_expect_tx(&mbar, bytes + WR_XOFF_STRIDE * 2u);
                    bulk_g2s(sx_raw, R.X + (xw0 & ~3u), bytes, &mbar);
                    bulk_g2s(soff, XS.xoff + (size_t)w * WR_XOFF_STRIDE, WR_XOFF_STRIDE * 2u, &mbar);
                }
            }
        }
        // thread t holds the tile elements 2 * (k * THREADS + t) and the next one, k < QPT / 2 (as wj_count_kernel does); they are
        // loaded per window — a tile lies in one window but for a few, and the registers are free again after the sort
        uint32_t s[WJ_QPT], e[WJ_QPT];
#pragma unroll
        for (int k = 0; k < WJ_QPT / 2; k++) {
            const uint32_t i = 2u * ((uint32_t)k * WJ_THREADS + tid);
            uint4 v = make_uint4(0, 0, 0, 0);
            if (i + 1 < m && al16) v = __ldg(reinterpret_cast<const uint4*>(qse + tile0 + i));
            else if (i < m) {
                const uint2 u = __ldg(qse + tile0 + i);
                v.x = u.x; v.y = u.y;
                if (i + 1 < m) { const uint2 u1 = __ldg(qse + tile0 + i + 1); v.z = u1.x; v.w = u1.y; }
            }
            s[2 * k] = v.x; e[2 * k] = v.y; s[2 * k + 1] = v.z; e[2 * k + 1] = v.w;
        }
        // ---- 1. counting sort of the segment by key = fine bin of the start (+ WR_NBF: tail part) -------------------------
#pragma unroll
        for (int k = 0; k < WR_CPT; k++) S.hist[WR_CPT * tid + k] = 0;   // (the compiler joins these into word stores)
        if (tid == 0) { S.hist[2 * WR_NBF] = 0; S.maxlen = 0; S.re_max = 0; S.tot = 0; S.next_tail = 0; S.next_scat = 0; }
        __syncthreads();
        uint32_t kr[WJ_QPT];   // key << 12 | rank among the segment's regions of that key
        uint32_t my_len = 0, my_re = 0;
#pragma unroll
        for (int k = 0; k < WJ_QPT; k++) {
            const uint32_t q = tile0 + 2u * ((uint32_t)(k >> 1) * WJ_THREADS + tid) + (uint32_t)(k & 1);
            kr[k] = 0xFFFFFFFFu;
            if (q >= p_lo && q < p_hi) {
                const bool tail = s[k] >= e[k] || e[k] - s[k] > taillen;
                const uint32_t key = min((s[k] - lo) >> fsh, WR_NBF - 1u) + (tail ? WR_NBF : 0u);
                const uint32_t old = atomicAdd(reinterpret_cast<uint32_t*>(S.hist) + (key >> 1), (key & 1u) ? 0x10000u : 1u);
                kr[k] = (key << 12) | ((key & 1u) ? old >> 16 : old & 0xFFFFu);
                if (!tail) { my_len = max(my_len, e[k] - s[k]); my_re = max(my_re, e[k]); }
            }
        }
        my_len = __reduce_max_sync(0xffffffffu, my_len);
        my_re = __reduce_max_sync(0xffffffffu, my_re);
        if (lane == 0 && my_re) { atomicMax(&S.maxlen, my_len); atomicMax(&S.re_max, my_re); }
        __syncthreads();
        {
            uint32_t c[WR_CPT], sum = 0;
#pragma unroll
            for (int k = 0; k < WR_CPT; k++) { c[k] = S.hist[WR_CPT * tid + k]; sum += c[k]; }
            uint32_t tot;
            uint32_t ex = block_scan_excl(sum, OpSum(), 0u, S.scan, &tot);
#pragma unroll
            for (int k = 0; k < WR_CPT; k++) { S.hist[WR_CPT * tid + k] = (uint16_t)ex; ex += c[k]; }
            if (tid == 0) S.hist[2 * WR_NBF] = (uint16_t)tot;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < WJ_QPT; k++)
            if (kr[k] != 0xFFFFFFFFu) wj_sorted[S.hist[kr[k] >> 12] + (kr[k] & 0xFFFu)] = make_uint2(s[k], e[k]);
        __syncthreads();
        const uint32_t m_seg = p_hi - p_lo, n_norm = S.hist[WR_NBF];
        unsigned long long mytot = 0;   // this lane's part of the segment's hit total
        // ---- 2. the roots of the window over the normal part ------------------------------------------------------------
        // One root per lane, 32 consecutive roots per warp: the lane works out its root's bins, the two position ranges to test
        // and the count of the inside bins, tests its positions (a dozen on average) and updates the root's counter.  The warps
        // without roots (a window has a few hundred at most) go on to the scatter below.
        if (n_norm) {
            const size_t row = (size_t)w * (nb_old + 1);
            const uint32_t l0 = __ldg(C.boff + row), nl0 = __ldg(C.boff + row + 1) - l0, j0 = __ldg(C.first + row);
            // the run ends where no normal region of the window can reach (run_end: start >= window end + tail length, or the
            // sequence's sentinel): the indices below never leave the sequence's part of the array
            const uint32_t n_roots = nl0 + (__ldg(C.run_end + w) - j0);
            const uint32_t maxlen = S.maxlen, re_max = S.re_max;
            const uint16_t* H = S.hist;
            for (uint32_t base = warp * 32u; base < n_roots; base += WJ_THREADS) {
                const uint32_t i = base + lane;
                bool valid = i < n_roots;
                uint4 v = make_uint4(0, 0, 0, 0);
                if (valid) {
                    if (i < nl0) {   // the roots that reach into the window from the left: members of its first bin's list
                        v = __ldg(C.list4 + l0 + i);
                        valid = v.x < lo;   // one that starts inside the window is in the run
                    } else {
                        v = __ldg(C.ent4 + j0 + (i - nl0));
                        valid = v.x < re_max;   // no region of the segment ends beyond its start otherwise
                    }
                }
                const uint32_t rs = v.x, re = v.y;   // the root's interval
                uint32_t ra = 0, rb = 0, mid = 0;    // positions to test: [ra & 0xFFFF, ra >> 16) and [rb & 0xFFFF, rb >> 16); regions of the inside bins
                if (valid) {
                    if (re <= rs) {
                        ra = n_norm << 16;   // an empty or inverted root interval: nothing can be said per bin, every region is tested
                    } else {
                        const int bs = rs < lo ? -1 : (int)min((rs - lo) >> fsh, WR_NBF - 1u);
                        uint32_t be = (re - 1u - lo) >> fsh;   // re > lo: the root overlaps the window or starts in it
                        if (be >= WR_NBF) be = is_last ? WR_NBF - 1u : WR_NBF;
                        uint32_t bq = be;   // first bin after the inside bins whose regions pass the mode test without a look
                        if (MODE == 1) {
                            const uint32_t el = re > maxlen ? re - maxlen : 0u;   // a region that starts before el ends at or before re
                            bq = min(be, el <= lo ? 0u : min((el - lo) >> fsh, WR_NBF - 1u));
                        }
                        const uint32_t mstart = (uint32_t)(bs + 1), mend = max(bq, mstart);
                        if (bs >= 0) {
                            const uint32_t sl = rs > maxlen ? rs - maxlen : 0u;   // a region that reaches rs from the left starts after sl
                            const uint32_t bl = sl <= lo ? 0u : min((sl - lo) >> fsh, WR_NBF - 1u);
                            ra = (uint32_t)H[bl] | ((uint32_t)H[bs + 1] << 16);
                        }
                        if (MODE != 0 || INVERT) {   // --contained alone: a region that starts after rs cannot contain the root
                            mid = (uint32_t)H[mend] - (uint32_t)H[mstart];
                            const uint32_t b_hi = min(be, WR_NBF - 1u);
                            if (mend <= b_hi) rb = (uint32_t)H[mend] | ((uint32_t)H[b_hi + 1] << 16);
                        }
                    }
                }
                uint32_t mine = 0;   // overlap hits | mode hits << 16 among the tested positions (a segment has <= 4096 regions)
                const uint32_t a0 = ra & 0xFFFFu, n1 = (ra >> 16) - a0, b0 = rb & 0xFFFFu, nn = n1 + (rb >> 16) - b0;
                const bool big = valid && nn > WR_SERIAL_MAX;
                // a root with few positions to test (the rule: ~a dozen) walks them itself, 32 roots side by side ...
                if (valid && !big)
#pragma unroll 1
                    for (uint32_t l = 0; l < nn; l++) {
                        const uint2 q = wj_sorted[l < n1 ? a0 + l : b0 + (l - n1)];
                        const bool ov = rs < q.y && re > q.x;
                        mine += (ov ? 1u : 0u) + ((ov && wj_mode<MODE>(rs, re, q.x, q.y)) ? 0x10000u : 0u);
                    }
                // ... one with many (an inverted root interval, a tile crowded into a few bins) has the warp's 32 lanes test them
                for (uint32_t todo = __ballot_sync(0xffffffffu, big); todo; todo &= todo - 1u) {
                    const int t = __ffs(todo) - 1;
                    const uint32_t ts = __shfl_sync(0xffffffffu, rs, t), te = __shfl_sync(0xffffffffu, re, t);
                    const uint32_t ta0 = __shfl_sync(0xffffffffu, a0, t), tn1 = __shfl_sync(0xffffffffu, n1, t);
                    const uint32_t tb0 = __shfl_sync(0xffffffffu, b0, t), tnn = __shfl_sync(0xffffffffu, nn, t);
                    uint32_t acc = 0;
#pragma unroll 1
                    for (uint32_t l = lane; l < tnn; l += 32) {
                        const uint2 q = wj_sorted[l < tn1 ? ta0 + l : tb0 + (l - tn1)];
                        const bool ov = ts < q.y && te > q.x;
                        acc += (ov ? 1u : 0u) + ((ov && wj_mode<MODE>(ts, te, q.x, q.y)) ? 0x10000u : 0u);
                    }
                    acc = __reduce_add_sync(0xffffffffu, acc);
                    if ((int)lane == t) mine = acc;
                }
                const uint32_t n_ov = (mine & 0xFFFFu) + mid, n_md = MODE == 2 ? n_ov : (mine >> 16) + (MODE == 1 ? mid : 0u);
                const uint32_t res = INVERT ? n_ov - n_md : n_md;
                if (valid && res) {
                    atomicAdd(ent_cnt + v.w, (unsigned long long)res);
                    ent_hit[v.w] = 1;
                    mytot += res;
                }
            }
        }
        // ---- 3. the tail part, region-parallel against the bin lists ------------------------------------------------------
        if (n_norm < m_seg) {
            for (;;) {
                uint32_t t = 0;
                if (lane == 0) t = atomicAdd(&S.next_tail, 1u);
                t = __shfl_sync(0xffffffffu, t, 0);
                const uint32_t p = n_norm + 32u * t + lane;
                if (n_norm + 32u * t >= m_seg) break;
                const bool have = p < m_seg;
                const uint2 r = have ? wj_sorted[p] : make_uint2(0xFFFFFFFFu, 0u);
                const unsigned long long nh = wj_round_global<MODE, INVERT>(C, w, lo, r, have, lane, ent_hit, ent_cnt);
                if (lane == 0) mytot += nh;
            }
        }
        // ---- 4. the rank-space scatter of the line filter (K3a), in sorted order: neighbours share the look-up's cache lines ----
        // Chunks of 64 positions go to the warps in arrival order (the warps that had roots or tail rounds come later and take
        // less); a lane looks up two regions side by side, so that two chains of dependent loads are in flight per thread.
        if (FILT && filt) {
            const uint32_t* sx = sx_raw + (xw0 & 3u);
            const uint32_t rank0 = xw0 - RS.x0;   // rank of the part's first value within the sequence
            if (SLICE && use_slice) {
                mbar_wait(&mbar, mbar_phase);
                mbar_phase ^= 1u;
            }
            for (;;) {
                uint32_t c = 0;
                if (lane == 0) c = atomicAdd(&S.next_scat, 1u);
                c = __shfl_sync(0xffffffffu, c, 0);
                const uint32_t p = 64u * c + lane;
                if (64u * c >= m_seg) break;
                const bool h0 = p < m_seg, h1 = p + 32u < m_seg;
                const uint2 r0 = wj_sorted[h0 ? p : 64u * c], r1 = wj_sorted[h1 ? p + 32u : 64u * c];
                bool eq0, eq1;
                uint32_t lb0, lb1;
                if (SLICE && use_slice) {
                    lb0 = rank0 + rank_lb_slice(sx, soff, xlen, min((r0.x - lo) >> fsh, WR_NBF - 1u), r0.x, &eq0);
                    lb1 = rank0 + rank_lb_slice(sx, soff, xlen, min((r1.x - lo) >> fsh, WR_NBF - 1u), r1.x, &eq1);
                } else {
                    lb0 = rank_lb(R, RS, r0.x, &eq0);
                    lb1 = rank_lb(R, RS, r1.x, &eq1);
                }
                if (h0) scatter_apply<MODE>(R, RS, F, seq, r0.x, r0.y, lb0, eq0);
                if (h1) scatter_apply<MODE>(R, RS, F, seq, r1.x, r1.y, lb1, eq1);
            }
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) mytot += __shfl_xor_sync(0xffffffffu, mytot, d);
        if (lane == 0 && mytot) atomicAdd(&S.tot, mytot);
        __syncthreads();
        if (tid == 0 && S.tot) atomicAdd(n_pairs, S.tot);
        __syncthreads();   // the histogram and the sorted tile are reused by the next window of this tile
    }
}

}  // namespace gffx
