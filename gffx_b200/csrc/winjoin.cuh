// winjoin.cuh — K1: the root-level join over window-bucketed regions, replacing
// IntervalTree::query_interval (src/utils/tree.rs:98-121) and the per-region loop of query_features
// (src/commands/intersect.rs:139-165).
//
// The index, the search and the kernel are described at wj_join_kernel below; the result is exactly the set
// { iv : iv.start < re && iv.end > rs } of tree.rs:110 for any rs/re (rs >= re included), filtered by the mode
// test of intersect.rs:145-163.
#pragma once
#include "scan.cuh"
#include "filter.cuh"

namespace gffx {

constexpr int WJ_THREADS = 512;
constexpr int WJ_QPT = 8;            // regions per thread
constexpr int WJ_TILE = WJ_THREADS * WJ_QPT;
constexpr int WJ_WARPS = WJ_THREADS / 32;

// mode test of src/commands/intersect.rs:145-158 and the XOR of :161
template <int MODE, bool INVERT> __device__ __forceinline__ bool wj_keep(uint32_t s, uint32_t e, uint32_t rs, uint32_t re) {
    const bool keep = MODE == 0 ? (s >= rs && e <= re) : MODE == 1 ? (s <= rs && e >= re) : true;
    return INVERT != keep;
}

// ------------------------------------------------------------------------------------------------
// K1 = wj_count_kernel (regions -> hits per root) + wj_expand_kernel (hits per root -> the pair array).
//
// What query_features returns per hit is (iv.root_fid, iv.start, iv.end) (src/commands/intersect.rs:162): the pair
// names the ROOT only, so the result is the multiset { root r, once per region that hits r }.  The join therefore
// counts, per root, the regions that pass the tests of tree.rs:110 and intersect.rs:145-161, and the pair array is
// the run-length expansion of those counts (cnt[r] copies of r's 12-byte record, roots in start order per sequence —
// the reference's own order is hash-map order per chromosome, nothing downstream depends on it).
//
// Index (winjoin_build.h, WinBinsHost): every window of the partition grid is cut into 2^sub bins; a bin owns the list
// of the intervals overlapping it, and `first` points into the sequence's start-sorted interval array.  The intervals
// that overlap a region starting in bin b are the members of b's list that pass the test, then the run of the sorted
// array from first[b+1] while start < re.  Long genes, nested genes and chromosome-spanning roots cost nothing
// extra: no running max-end to back-track over (r01h/r01i profiles).
//
// After K0 the bucketed regions of one window are CONTIGUOUS, so a tile of WJ_TILE regions lies in one window
// (rarely a few: the sum over all tiles of extra windows is <= n_win).  Per (tile, window) the CTA
//   1. copies the window's slice — bin offsets, `first`, the list members and the run, as 16-byte
//      (start,end,fid,id) records — into shared memory; whatever exceeds the slice capacity is read from global
//      memory, so the result is the same for any input;
//   2. counting-sorts its regions by bin in shared memory (one shared atomic per region, one scan of <= 512 counters);
//   3. hands rounds of 32 consecutive sorted regions to its warps: the regions of a round start in the same bin or the
//      next few, so ONE candidate set serves the round (the list of the round's first bin, then the run while start <
//      the round's largest region end) and the candidate loops are uniform across the warp — one broadcast 16-byte
//      shared load per candidate, the exact per-lane test, one ballot, and lane t keeps the popcount of candidate t in a
//      register until the round ends.  No lane divergence, no per-hit stores (r01k / s17 profiles of the
//      per-region-thread versions: 480-490 instructions per region in divergent candidate loops and per-pair
//      bookkeeping);
//   4. adds the non-zero slot counters to the global per-root counters (one 64-bit atomic per hit root and CTA).
constexpr int WJ_BINS_MAX = 512;    // bins per window (sub <= 9)
constexpr int WJ_LIST_CAP = 768;    // list members of the window held in shared memory (12 KB)
constexpr int WJ_RUN_CAP = 256;     // start-sorted intervals from the window's first held in shared memory (4 KB)

struct WinBins {
    const uint32_t* win_seq;   // n_win
    const uint32_t* win0;      // n_seqs + 1
    const uint4* ent4;         // start-sorted intervals per sequence + sentinel: (start, end, root_fid, id)
    const uint32_t* ent_fid;   // the same root_fid by id (block selection / warn list)
    const uint4* list4;        // bin lists as records
    const uint32_t* first;     // n_win * (nb + 1)
    const uint32_t* boff;      // n_win * (nb + 1)
    const uint32_t* run_end;   // n_win: end of the window's root run for the root-parallel kernel (rootjoin.cuh)
    uint32_t shift, sub, n_win, n_ent;
};

struct WjSmem {   // static shared memory (26 KB); the sorted tile (WJ_DYN_SMEM bytes) is the dynamic part
    uint4 list[WJ_LIST_CAP];
    uint4 run[WJ_RUN_CAP];
    uint2 bf[WJ_BINS_MAX + 1];                     // (boff - l0, first - j0): slice-local indices
    uint32_t hist[WJ_BINS_MAX + 1];                // regions per bin -> first sorted position of the bin
    uint32_t lcnt[WJ_LIST_CAP], rcnt[WJ_RUN_CAP];  // hits per slice member
    uint32_t scan[WJ_WARPS + 1];
    unsigned long long tot;                        // hits of the (tile, window): 64-bit, a region may hit every interval of a huge index
    uint32_t w_first, w_last;
    uint32_t next_round;                           // rounds are handed to the warps in arrival order
    uint32_t sent;                                 // slice-local index of the sequence's sentinel if the run holds it, else 0xFFFFFFFF
};
constexpr size_t WJ_DYN_SMEM = (size_t)WJ_TILE * sizeof(uint2);   // the (tile, window) regions ordered by bin

// A (tile, window) segment of at most 32 regions — a small job, or the edge of a tile — is not worth a slice in shared memory:
// one warp resolves it as a single round straight from the global index (the same candidate rule as the rounds below: the
// list of the smallest bin among its regions, then the run while start < the largest region end).  Warps take such
// segments side by side, so a 10 k-region job no longer walks the windows one after the other with CTA-wide barriers
// (4.3 ms for 1000 regions before, profiles/host_overhead.py).
template <int MODE, bool INVERT, bool FILT>
__device__ __noinline__ void wj_small_segment(const WinBins& C, uint32_t w, const uint2* __restrict__ qse, uint32_t p_lo, uint32_t m_seg, uint32_t lane,
                                              uint8_t* ent_hit, unsigned long long* ent_cnt, unsigned long long* n_pairs, const RankSpace& R,
                                              const FilterArrays& F) {
    const uint32_t nb = 1u << C.sub, bsh = C.shift - C.sub;
    const size_t row = (size_t)w * (nb + 1);
    const uint32_t seq = __ldg(C.win_seq + w);
    const uint32_t lo = (w - __ldg(C.win0 + seq)) << C.shift;
    const bool have = lane < m_seg;
    const uint2 r = have ? qse[p_lo + lane] : make_uint2(0xFFFFFFFFu, 0u);
    if (FILT) {
        const RankSeq RS = load_rank_seq(R, seq);
        if (RS.canon && RS.cnt != 0 && have) scatter_region<MODE>(R, RS, F, seq, r.x, r.y);
    }
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
    if (lane == 0 && tot) atomicAdd(n_pairs, tot);
}
constexpr uint32_t WJ_SMALL_SEG = 32;

#ifndef WJ_MIN_CTAS
#define WJ_MIN_CTAS 3     // 40 registers: the join alone fits without spills
#endif
#ifndef WJ_MIN_CTAS_FILT
#define WJ_MIN_CTAS_FILT 3   // 96 bytes of spills, still faster than 2 CTAs (measured)
#endif
// FILT: the same pass also scatters every region onto the line filter's rank space (K3a, filter.cuh: scatter_region) —
// the regions are in shared memory already, and a round's 32 regions start inside one bin, so their rank look-ups
// walk the same few cache lines.  (As its own kernel, with its own load and tile sort, K3a took 120-135 us per 10 M
// regions, L1-bound on those look-ups: profiles/r01k.)
// SMALL: few regions per window (a small job): segments of <= 32 regions go to wj_small_segment, one warp each.
template <int MODE, bool INVERT, bool FILT, bool SMALL>
__global__ void __launch_bounds__(WJ_THREADS, FILT ? WJ_MIN_CTAS_FILT : WJ_MIN_CTAS)
    wj_count_kernel(WinBins C, const uint32_t* __restrict__ poff, const uint2* __restrict__ qse, uint64_t n, uint8_t* ent_hit,
                    unsigned long long* ent_cnt, unsigned long long* __restrict__ n_pairs, uint32_t* __restrict__ err, RankSpace R,
                    FilterArrays F) {
    extern __shared__ __align__(16) uint2 wj_sorted[];
    __shared__ WjSmem S;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t tile0 = blockIdx.x * (uint32_t)WJ_TILE;
    const uint32_t m = (uint32_t)min((uint64_t)WJ_TILE, n - tile0);
    const uint32_t n_win = C.n_win, nb = 1u << C.sub, bsh = C.shift - C.sub;
    // thread t holds the tile elements 2 * (k * THREADS + t) and the next one, k < QPT / 2: 16-byte loads, a warp on 512 contiguous bytes
    uint32_t s[WJ_QPT], e[WJ_QPT];
    const bool al16 = (reinterpret_cast<uintptr_t>(qse) & 15u) == 0;   // a batch appended at an odd offset is only 8-byte aligned
#pragma unroll
    for (int k = 0; k < WJ_QPT / 2; k++) {
        const uint32_t i = 2u * ((uint32_t)k * WJ_THREADS + tid);
        uint4 v = make_uint4(0, 0, 0, 0);
        if (i + 1 < m && al16) v = __ldcs(reinterpret_cast<const uint4*>(qse + tile0 + i));
        else if (i < m) {
            const uint2 u = __ldcs(qse + tile0 + i);
            v.x = u.x; v.y = u.y;
            if (i + 1 < m) { const uint2 u1 = __ldcs(qse + tile0 + i + 1); v.z = u1.x; v.w = u1.y; }
        }
        s[2 * k] = v.x; e[2 * k] = v.y; s[2 * k + 1] = v.z; e[2 * k + 1] = v.w;
    }
    // windows of the tile's first and last region: (number of window offsets <= position) - 1; poff has n_win + 2 cells
    {
        int nf = 0, nl = 0;
        for (uint32_t i0 = 0; i0 < n_win + 2; i0 += WJ_THREADS) {
            const uint32_t i = i0 + tid;
            const uint32_t o = i < n_win + 2 ? __ldg(poff + i) : 0xFFFFFFFFu;
            nf += __syncthreads_count(o <= tile0);
            nl += __syncthreads_count(o <= tile0 + m - 1);
        }
        if (tid == 0) { S.w_first = (uint32_t)nf - 1u; S.w_last = (uint32_t)nl - 1u; S.sent = 0xFFFFFFFFu; }
    }
    __syncthreads();
    const uint32_t w_first = S.w_first, w_last = S.w_last;
    // small segments first, one warp each; a SMALL launch spreads them over gridDim.y CTAs per tile, of which only y = 0 goes on
    for (uint32_t w = w_first + warp + WJ_WARPS * blockIdx.y; SMALL && w <= w_last && w < n_win; w += WJ_WARPS * gridDim.y) {
        const uint32_t p_lo = max(__ldg(poff + w), tile0), p_hi = min(__ldg(poff + w + 1), tile0 + m);
        if (p_lo < p_hi && p_hi - p_lo <= WJ_SMALL_SEG) wj_small_segment<MODE, INVERT, FILT>(C, w, qse, p_lo, p_hi - p_lo, lane, ent_hit, ent_cnt, n_pairs, R, F);
    }
    if (SMALL && blockIdx.y != 0) return;
    for (uint32_t w = w_first; w <= w_last; w++) {
        const uint32_t p_lo = max(__ldg(poff + w), tile0), p_hi = min(__ldg(poff + w + 1), tile0 + m);
        if (p_lo >= p_hi) continue;   // empty window (uniform)
        if (w >= n_win) {             // bucket of the regions whose sequence number is out of range
            if (tid == 0) *err = 1;
            continue;
        }
        if (SMALL && p_hi - p_lo <= WJ_SMALL_SEG) continue;   // done above
        // ---- 1. the window's slice -----------------------------------------------------------------
        const size_t row = (size_t)w * (nb + 1);
        const uint32_t seq = __ldg(C.win_seq + w);
        const uint32_t lo = (w - __ldg(C.win0 + seq)) << C.shift;
        RankSeq RS{};
        bool filt = false;   // no line of a sequence without rank space can ever match
        if (FILT) {
            RS = load_rank_seq(R, seq);
            filt = RS.canon && RS.cnt != 0;
        }
        const uint32_t l0 = __ldg(C.boff + row), j0 = __ldg(C.first + row);
        const uint32_t nlist = min(__ldg(C.boff + row + nb) - l0, (uint32_t)WJ_LIST_CAP);
        const uint32_t nrun = min(min(__ldg(C.first + row + nb) - j0 + 8u, (uint32_t)WJ_RUN_CAP), C.n_ent - j0);
        for (uint32_t i = tid; i <= nb; i += WJ_THREADS) {
            S.bf[i] = make_uint2(__ldg(C.boff + row + i) - l0, __ldg(C.first + row + i) - j0);
            S.hist[i] = 0;
        }
        for (uint32_t i = tid; i < nlist; i += WJ_THREADS) { S.list[i] = __ldg(C.list4 + l0 + i); S.lcnt[i] = 0; }
        for (uint32_t i = tid; i < nrun; i += WJ_THREADS) {
            const uint4 v = __ldg(C.ent4 + j0 + i);
            S.run[i] = v; S.rcnt[i] = 0;
            if (v.x == 0xFFFFFFFFu) atomicMin(&S.sent, i);   // what follows the sentinel belongs to the next sequence
        }
        if (tid == 0) { S.tot = 0; S.next_round = 0; }
        __syncthreads();
        // ---- 2. counting sort of the window's regions by bin ----------------------------------------
        uint32_t rank[WJ_QPT];
#pragma unroll
        for (int k = 0; k < WJ_QPT; k++) {
            const uint32_t q = tile0 + 2u * ((uint32_t)(k >> 1) * WJ_THREADS + tid) + (uint32_t)(k & 1);
            rank[k] = 0xFFFFFFFFu;
            if (q >= p_lo && q < p_hi) rank[k] = atomicAdd(&S.hist[min((s[k] - lo) >> bsh, nb - 1)], 1u);
        }
        __syncthreads();
        {
            static_assert(WJ_BINS_MAX <= WJ_THREADS, "one bin counter per thread");
            const uint32_t v = tid < nb ? S.hist[tid] : 0u;
            uint32_t tot;
            const uint32_t ex = block_scan_excl(v, OpSum(), 0u, S.scan, &tot);
            if (tid < nb) S.hist[tid] = ex;   // every thread read its counter before the scan's first barrier
            if (tid == 0) S.hist[nb] = tot;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < WJ_QPT; k++)
            if (rank[k] != 0xFFFFFFFFu) wj_sorted[S.hist[min((s[k] - lo) >> bsh, nb - 1)] + rank[k]] = make_uint2(s[k], e[k]);
        __syncthreads();
        // ---- 3. probe: one bin per warp at a time, 32 of its regions per round, candidates in lock step ----------
        const bool list_full = __ldg(C.boff + row + nb) - l0 <= (uint32_t)WJ_LIST_CAP;   // every list member of the window is in shared memory
        const uint32_t run_end = S.sent == 0xFFFFFFFFu ? nrun : min(nrun, S.sent + 1u);   // the shared-memory run is ascending in start up to here
        unsigned long long mytot = 0;   // this lane's part of the (tile, window) hit total
        // the general walks: any list length, members beyond the shared-memory slice read from global memory
        auto walk_list = [&](const uint2 r, uint32_t i0, uint32_t i1) {
            for (uint32_t i = i0; i < i1; i++) {
                const bool in = i < nlist;
                const uint4 v = in ? S.list[i] : __ldg(C.list4 + l0 + i);
                const bool hit = v.x < r.y && v.y > r.x && wj_keep<MODE, INVERT>(v.x, v.y, r.x, r.y);
                const uint32_t nh = __popc(__ballot_sync(0xffffffffu, hit));
                if (nh && lane == 0) {
                    mytot += nh;
                    if (in) atomicAdd(&S.lcnt[i], nh);
                    else { atomicAdd(ent_cnt + v.w, (unsigned long long)nh); ent_hit[v.w] = 1; }
                }
            }
        };
        auto walk_run = [&](const uint2 r, uint32_t i0, uint32_t re_max) {
            for (uint32_t i = i0;; i++) {   // the sequence's sentinel (start 0xFFFFFFFF) ends it at the latest
                const bool in = i < nrun;
                const uint4 v = in ? S.run[i] : __ldg(C.ent4 + j0 + i);
                if (v.x >= re_max) break;
                const bool hit = v.x < r.y && v.y > r.x && wj_keep<MODE, INVERT>(v.x, v.y, r.x, r.y);
                const uint32_t nh = __popc(__ballot_sync(0xffffffffu, hit));
                if (nh && lane == 0) {
                    mytot += nh;
                    if (in) atomicAdd(&S.rcnt[i], nh);
                    else { atomicAdd(ent_cnt + j0 + i, (unsigned long long)nh); ent_hit[j0 + i] = 1; }
                }
            }
        };
        // A round = 32 consecutive positions of the sorted tile.  Its regions start in bins b_first <= b: an interval that hits one
        // of them either overlaps bin b_first (it starts before that bin's end and ends after the region's start, which is
        // not before the bin's start) or starts later, at or after the bin's end — so the candidates of the whole round are
        // the list of b_first and the run from first[b_first + 1] while start < the round's largest region end.  The test
        // per candidate and lane is the exact one, so a candidate too many costs a compare, never a wrong count.
        const uint32_t m_seg = p_hi - p_lo, n_rounds = (m_seg + 31u) >> 5;
        for (;;) {
            uint32_t rd = 0;
            if (lane == 0) rd = atomicAdd(&S.next_round, 1u);   // rounds go to the warps in arrival order: no warp waits long at the barrier
            rd = __shfl_sync(0xffffffffu, rd, 0);
            if (rd >= n_rounds) break;
            const bool have = rd * 32u + lane < m_seg;
            // a lane without a region holds (rs, re) = (max, 0): no interval has start < 0
            const uint2 r = have ? wj_sorted[rd * 32u + lane] : make_uint2(0xFFFFFFFFu, 0u);
            if (FILT && filt && have) scatter_region<MODE>(R, RS, F, seq, r.x, r.y);
            const uint32_t b = __shfl_sync(0xffffffffu, min((r.x - lo) >> bsh, nb - 1), 0);   // lane 0 always has a region, the smallest bin
            const uint2 b0 = S.bf[b], b1 = S.bf[b + 1];
            const uint32_t nL = b1.x - b0.x;
            // lean form: lane t keeps the hits of the bin's list member t and run member t in registers
            uint32_t lacc = 0, racc = 0;
            if (list_full && nL <= 32u) {
                for (uint32_t t = 0; t < nL; t++) {
                    const uint4 v = S.list[b0.x + t];
                    const bool hit = v.x < r.y && v.y > r.x && wj_keep<MODE, INVERT>(v.x, v.y, r.x, r.y);
                    const uint32_t nh = __popc(__ballot_sync(0xffffffffu, hit));
                    if (lane == t) lacc += nh;
                }
            } else {
                walk_list(r, b0.x, b1.x);
            }
            const uint32_t re_max = __reduce_max_sync(0xffffffffu, r.y);
            // within 32 run members of the first one the start reaches the round's largest region end, all in shared memory
            const uint32_t jl = min(b1.y + 31u, run_end - 1u);
            const uint32_t run_reach = (run_end && b1.y < run_end) ? S.run[jl].x : 0u;
            if (run_reach >= re_max) {
                for (uint32_t t = 0;; t++) {
                    const uint4 v = S.run[b1.y + t];
                    if (v.x >= re_max) break;
                    const bool hit = v.x < r.y && v.y > r.x && wj_keep<MODE, INVERT>(v.x, v.y, r.x, r.y);
                    const uint32_t nh = __popc(__ballot_sync(0xffffffffu, hit));
                    if (lane == t) racc += nh;
                }
            } else {
                walk_run(r, b1.y, re_max);
            }
            if (lacc) atomicAdd(&S.lcnt[b0.x + lane], lacc);
            if (racc) atomicAdd(&S.rcnt[b1.y + lane], racc);
            mytot += lacc + racc;
        }
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) mytot += __shfl_xor_sync(0xffffffffu, mytot, d);
        if (lane == 0 && mytot) atomicAdd(&S.tot, mytot);
        __syncthreads();
        // ---- 4. slot counters -> per-root counters and hit flags ---------------------------------------------
        for (uint32_t i = tid; i < nlist; i += WJ_THREADS) {
            const uint32_t c = S.lcnt[i];
            if (c) {
                const uint32_t id = S.list[i].w;
                atomicAdd(ent_cnt + id, (unsigned long long)c);
                ent_hit[id] = 1;   // a plain byte store: testing the flag first would stall on the load
            }
        }
        for (uint32_t i = tid; i < nrun; i += WJ_THREADS) {
            const uint32_t c = S.rcnt[i];
            if (c) {
                atomicAdd(ent_cnt + j0 + i, (unsigned long long)c);
                ent_hit[j0 + i] = 1;
            }
        }
        if (tid == 0) {
            if (S.tot) atomicAdd(n_pairs, S.tot);
            S.sent = 0xFFFFFFFFu;
        }
        __syncthreads();  // the slice and the sorted tile are reused by the next window of this tile
    }
}

// element i of the scan that turns the per-root counters into first-pair offsets: n_ent counters, then one 0
struct WjCntLoad {
    const unsigned long long* cnt;
    uint32_t n_ent;
    __device__ unsigned long long operator()(uint64_t i) const { return i < n_ent ? cnt[i] : 0ull; }
};

// K1b: pairs[off[j] .. off[j+1]) = cnt[j] copies of (root_fid, start, end) of sorted interval j.  Output-centric: a
// warp owns WJ_EXP_CHUNK consecutive pairs, finds the interval of its first pair by binary search over `off`
// (n_ent + 1 ascending offsets, off[n_ent] = total) and walks forward, skipping the intervals nobody hit 32 at a
// time; the stores are 16-byte st.cs (the 12-byte record repeats with period three vectors), so a root with
// millions of hits (a chromosome-spanning `region` root) is spread over as many warps as its bytes need.
constexpr int WJ_EXP_THREADS = 256;
constexpr uint32_t WJ_EXP_CHUNK = 2048;   // pairs per warp and step: 24 KB

static __global__ void __launch_bounds__(WJ_EXP_THREADS)
    wj_expand_kernel(const unsigned long long* __restrict__ off, const uint4* __restrict__ ent4, const uint32_t* __restrict__ ent_base,
                     uint32_t n_ent, uint32_t* __restrict__ pairs, unsigned long long cap) {
    const unsigned long long total = off[n_ent];
    if (total > cap) return;   // the host grows the buffer and launches again
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t n_warps = (uint64_t)gridDim.x * (WJ_EXP_THREADS / 32);
    // pairs per warp and step: the whole grid gets work even for a small result (the walk over the intervals of a chunk
    // is a chain of dependent loads), at most WJ_EXP_CHUNK
    const unsigned long long per = min((unsigned long long)WJ_EXP_CHUNK, max(32ull, (total + n_warps - 1) / n_warps));
    for (uint64_t chunk = (uint64_t)blockIdx.x * (WJ_EXP_THREADS / 32) + (threadIdx.x >> 5); chunk * per < total; chunk += n_warps) {
        const unsigned long long p0 = chunk * per, p1 = min(p0 + per, total);
        // the interval of the chunk's first pair: off[lo] <= p0 < off[hi], 32 pivots per step (a lane each), so a table of 60 k
        // offsets is settled in four dependent loads instead of sixteen
        uint32_t lo = 0, hi = n_ent;
        while (hi - lo > 1) {
            const uint32_t span = hi - lo, step = (span + 32u) / 33u;   // pivots lo + step * (lane + 1), all < hi for the lanes that count
            const uint32_t piv = lo + step * (lane + 1u);
            const bool le = piv < hi && __ldg(off + piv) <= p0;
            const uint32_t m = __ballot_sync(0xffffffffu, le);          // monotone: the lanes with off[piv] <= p0 form a prefix
            const uint32_t k = (uint32_t)__popc(m);                     // lo moves to the last such pivot, hi to the first other one
            const uint32_t nlo = k ? lo + step * k : lo, nhi = min(hi, lo + step * (k + 1u));
            lo = nlo; hi = nhi;
        }
        uint32_t j = lo;   // off[j] <= p0 < off[j + 1]
        unsigned long long p = p0;
        // the records of the next 32 intervals with pairs are fetched lane-parallel (offset, interval, base: three independent
        // loads per lane) and handed round by shuffle: the chain of dependent loads per interval was most of a warp's time
        while (p < p1) {
            // lane t looks at interval j + t: does it own pairs beyond p?
            const uint32_t jj = j + lane;
            const unsigned long long oe = jj < n_ent ? (unsigned long long)__ldg(off + jj + 1) : 0ull;
            uint4 v = make_uint4(0, 0, 0, 0);
            uint32_t vb = 0;
            const bool has = jj < n_ent && oe > p;
            if (has) { v = __ldg(ent4 + jj); vb = __ldg(ent_base + jj); }
            uint32_t todo = __ballot_sync(0xffffffffu, has);
            if (!todo) { j += 32; continue; }   // 32 intervals nobody hit
            for (; todo && p < p1; todo &= todo - 1u) {
                const int t = __ffs(todo) - 1;
                const unsigned long long te = __shfl_sync(0xffffffffu, oe, t);
                const uint32_t c0 = __shfl_sync(0xffffffffu, v.z, t);
                const uint32_t c1 = __shfl_sync(0xffffffffu, v.x - vb, t), c2 = __shfl_sync(0xffffffffu, v.y - vb, t);   // the sequence's own coordinates
                const unsigned long long pe = min(te, p1);   // this interval's pairs in the chunk: [p, pe)
                const unsigned long long w0 = 3ull * p, w1 = 3ull * pe;   // word w of the pair array holds component w % 3
                unsigned long long wa = min((w0 + 3ull) & ~3ull, w1);   // first 16-byte boundary at or after w0
                if (w0 + lane < wa) {
                    const uint32_t k = (uint32_t)((w0 + lane) % 3ull);
                    __stcs(pairs + w0 + lane, k == 0 ? c0 : k == 1 ? c1 : c2);
                }
                const unsigned long long wb = wa + ((w1 - wa) & ~3ull);   // end of the whole vectors
                if (wa < wb) {
                    uint32_t k = (uint32_t)((wa + 4ull * lane) % 3ull);
                    uint4* g = reinterpret_cast<uint4*>(pairs + wa);
                    const uint64_t nv = (wb - wa) >> 2;
                    for (uint64_t u = lane; u < nv; u += 32) {
                        // components k, k+1, k+2, k (mod 3); the next vector of this lane is 128 words on: k advances by 128 % 3 = 2
                        const uint4 x = k == 0 ? make_uint4(c0, c1, c2, c0) : k == 1 ? make_uint4(c1, c2, c0, c1) : make_uint4(c2, c0, c1, c2);
                        __stcs(g + u, x);
                        k = k >= 1 ? k - 1 : 2;
                    }
                }
                if (wb + lane < w1) {
                    const uint32_t k = (uint32_t)((wb + lane) % 3ull);
                    __stcs(pairs + wb + lane, k == 0 ? c0 : k == 1 ? c1 : c2);
                }
                p = pe;
            }
            j += 32;
        }
    }
}

// per-root hit flags -> per-block selection flags, by GATHER: block b reads the flags of the tree intervals that name it
// (blk_ent[blk_ent0[b] .. blk_ent0[b+1]): one interval, or several with a duplicate root fid), so no flag array has to be
// cleared first and no two threads write one byte.  n_sel (may be NULL: the -e merge counts them itself) += selected blocks
// with a usable start.  Intervals whose fid has no .gof record (`orphans`) are appended to warn_fids when hit (the
// reference warns about them, intersect.rs:269-272).
static __global__ void __launch_bounds__(256)
    wj_select_blocks_kernel(const uint8_t* __restrict__ ent_hit, const uint32_t* __restrict__ blk_ent0, const uint32_t* __restrict__ blk_ent,
                            uint32_t n_blocks, const uint64_t* __restrict__ bstart, const uint32_t* __restrict__ orphans, uint32_t n_orphans,
                            const uint32_t* __restrict__ ent_fid, uint8_t* __restrict__ blk_sel, unsigned long long* __restrict__ n_sel,
                            uint32_t* __restrict__ warn_fids, uint32_t* __restrict__ n_warn, uint32_t warn_cap) {
    __shared__ uint32_t sm[256 / 32 + 1];
    const uint32_t b = blockIdx.x * 256 + threadIdx.x;
    uint32_t k = 0;
    if (b < n_blocks) {
        bool sel = false;
        for (uint32_t i = blk_ent0[b]; i < blk_ent0[b + 1]; i++) sel |= ent_hit[blk_ent[i]] != 0;
        blk_sel[b] = sel ? 1 : 0;
        k = (sel && bstart[b] != 0xFFFFFFFFFFFFFFFFull) ? 1u : 0u;
    }
    if (b < n_orphans) {
        const uint32_t j = orphans[b];
        if (ent_hit[j]) {
            const uint32_t p = atomicAdd(n_warn, 1u);
            if (p < warn_cap) warn_fids[p] = ent_fid[j];
        }
    }
    uint32_t tot;
    block_scan_excl(k, OpSum(), 0u, sm, &tot);
    if (n_sel && threadIdx.x == 0 && tot) atomicAdd(n_sel, (unsigned long long)tot);
}

}  // namespace gffx
