// winjoin.cuh — K1: the root-level join over window-bucketed regions, replacing
// IntervalTree::query_interval (src/utils/tree.rs:98-121) and the per-region loop of query_features
// (src/commands/intersect.rs:139-165).
//
// The index, the search and the kernel are described at wj_join_kernel below; the result is exactly the set
// { iv : iv.start < re && iv.end > rs } of tree.rs:110 for any rs/re (rs >= re included), filtered by the mode
// test of intersect.rs:145-163.
#pragma once
#include "scan.cuh"

namespace gffx {

constexpr int WJ_THREADS = 256;
constexpr int WJ_QPT = 4;  // consecutive bucketed regions per thread: one 16 B load of chr, two of (start,end)
constexpr int WJ_TILE = WJ_THREADS * WJ_QPT;
constexpr int WJ_STAGE_PAIRS = 1792;  // 21 KB of shared memory (5 CTAs per SM with the 20 KB window slice)

// mode test of src/commands/intersect.rs:145-158 and the XOR of :161
template <int MODE, bool INVERT> __device__ __forceinline__ bool wj_keep(uint32_t s, uint32_t e, uint32_t rs, uint32_t re) {
    const bool keep = MODE == 0 ? (s >= rs && e <= re) : MODE == 1 ? (s <= rs && e >= re) : true;
    return INVERT != keep;
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

// ------------------------------------------------------------------------------------------------
// K1, bin lists in shared memory.
//
// Index (winjoin_build.h, WinBinsHost): every window of the partition grid is cut into 2^sub bins; a bin owns the list
// of the intervals overlapping it, and `first` points into the sequence's start-sorted interval array.  The intervals
// that overlap a region starting in bin b are the members of b's list that pass the test, then the run of the sorted
// array from first[b+1] while start < re.  Long genes, nested genes and chromosome-spanning roots cost nothing
// extra: no running max-end to back-track over (r01h/r01i profiles: 400-900 instructions per region went into
// candidates that an earlier long gene dragged in, or into probing one chain per nesting level).
//
// After K0 the bucketed regions of one window are CONTIGUOUS, so a tile of WJ_TILE regions lies in one window
// (rarely a few: the sum over all tiles of extra windows is <= n_win).  The CTA copies the window's slice — bin
// offsets, `first`, the list members and the run, as 16-byte (start,end,fid,id) records — into shared memory once,
// and every region resolves with shared-memory reads only; whatever exceeds the slice capacity is read from global
// memory, so the result is the same set for any input.
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
    uint32_t shift, sub, n_win, n_ent;
};

struct WjSlice {
    uint4 list[WJ_LIST_CAP];
    uint4 run[WJ_RUN_CAP];
    uint32_t boff[WJ_BINS_MAX + 1];
    uint32_t first[WJ_BINS_MAX + 1];
    uint8_t lhit[WJ_LIST_CAP], rhit[WJ_RUN_CAP];   // hit flags of the slice members: plain byte stores, no atomics
    uint32_t fits;   // every list member of the window is in the slice
    uint32_t w_first, w_last, lo, l0, nlist, j0, nrun;
};

// visits, in order, the intervals with start < re && end > rs that pass the mode test: f(record, bit) where bit is the
// position code of the hit (list member i < 16 -> bit i, run member i < 15 -> bit 16+i, else bit 31 = "does not fit")
template <int MODE, bool INVERT, class F>
__device__ __forceinline__ void wj_visit(const WinBins& C, const WjSlice& S, uint32_t b, uint32_t rs, uint32_t re, F&& f) {
    const uint32_t o0 = S.boff[b], o1 = S.boff[b + 1];
    uint32_t bit = 1u;
    for (uint32_t a = o0; a < o1; a++, bit <<= 1) {
        const uint32_t i = a - S.l0;
        const uint4 v = i < S.nlist ? S.list[i] : __ldg(C.list4 + a);
        if (v.x < re && v.y > rs && wj_keep<MODE, INVERT>(v.x, v.y, rs, re)) f(v, (bit & 0xFFFFu) ? bit : 0x80000000u, i, true);
    }
    bit = 1u << 16;
    for (uint32_t j = S.first[b + 1];; j++, bit <<= 1) {   // the sequence's sentinel (start 0xFFFFFFFF) ends it at the latest
        const uint32_t i = j - S.j0;
        const uint4 v = i < S.nrun ? S.run[i] : __ldg(C.ent4 + j);
        if (v.x >= re) break;
        if (v.y > rs && wj_keep<MODE, INVERT>(v.x, v.y, rs, re)) f(v, (bit & 0x7FFF0000u) ? bit : 0x80000000u, i, false);
    }
}

// EMIT = false: pair count + per-entry hit flags only (the CLI path needs no pairs).
// EMIT = true : also writes the pairs; a (tile, window) whose claimed range does not fit `cap` writes nothing (the
//               host grows the buffer to the reported total and runs the kernel again).
template <int MODE, bool INVERT, bool EMIT>
__global__ void __launch_bounds__(WJ_THREADS)
    wj_join_kernel(WinBins C, const uint32_t* __restrict__ poff, const uint32_t* __restrict__ chr, const uint2* __restrict__ qse,
                   uint64_t n, uint8_t* ent_hit, unsigned long long* __restrict__ n_pairs, uint32_t* __restrict__ pairs,
                   unsigned long long cap, uint32_t* __restrict__ err) {
    __shared__ WjSlice S;
    __shared__ __align__(16) uint32_t stage[EMIT ? WJ_STAGE_PAIRS * 3 : 4];
    __shared__ uint32_t sm[WJ_THREADS / 32 + 1];
    __shared__ unsigned long long sbase;
    const uint32_t tile0 = blockIdx.x * (uint32_t)WJ_TILE;
    const uint32_t m = (uint32_t)min((uint64_t)WJ_TILE, n - tile0);
    const uint32_t n_win = C.n_win, nb = 1u << C.sub, bsh = C.shift - C.sub;
    uint32_t c[WJ_QPT], s[WJ_QPT], e[WJ_QPT];
    const uint32_t q0 = tile0 + threadIdx.x * WJ_QPT;
    wj_load(chr, qse, n, q0, c, s, e);
    // windows of the tile's first and last region: (number of window offsets <= position) - 1; poff has n_win + 2 cells
    {
        int nf = 0, nl = 0;
        for (uint32_t i0 = 0; i0 < n_win + 2; i0 += WJ_THREADS) {
            const uint32_t i = i0 + threadIdx.x;
            const uint32_t o = i < n_win + 2 ? __ldg(poff + i) : 0xFFFFFFFFu;
            nf += __syncthreads_count(o <= tile0);
            nl += __syncthreads_count(o <= tile0 + m - 1);
        }
        if (threadIdx.x == 0) { S.w_first = (uint32_t)nf - 1u; S.w_last = (uint32_t)nl - 1u; }
    }
    __syncthreads();
    const uint32_t w_first = S.w_first, w_last = S.w_last;
    for (uint32_t w = w_first; w <= w_last; w++) {
        const uint32_t p_lo = max(__ldg(poff + w), tile0), p_hi = min(__ldg(poff + w + 1), tile0 + m);
        if (p_lo >= p_hi) continue;   // empty window (uniform)
        if (w >= n_win) {             // bucket of the regions whose sequence number is out of range
            if (threadIdx.x == 0) *err = 1;
            continue;
        }
        const size_t row = (size_t)w * (nb + 1);
        if (threadIdx.x == 0) {
            const uint32_t cw = __ldg(C.win_seq + w);
            S.lo = (w - __ldg(C.win0 + cw)) << C.shift;
            const uint32_t l0 = __ldg(C.boff + row), l1 = __ldg(C.boff + row + nb);
            const uint32_t j0 = __ldg(C.first + row), j1 = __ldg(C.first + row + nb);
            S.l0 = l0; S.nlist = min(l1 - l0, (uint32_t)WJ_LIST_CAP);
            S.fits = (l1 - l0 <= (uint32_t)WJ_LIST_CAP) ? 1u : 0u;
            S.j0 = j0; S.nrun = min(min(j1 - j0 + 8u, (uint32_t)WJ_RUN_CAP), C.n_ent - j0);
        }
        for (uint32_t i = threadIdx.x; i <= nb; i += WJ_THREADS) {
            S.boff[i] = __ldg(C.boff + row + i);
            S.first[i] = __ldg(C.first + row + i);
        }
        static_assert(WJ_LIST_CAP / 4 <= WJ_THREADS && WJ_RUN_CAP / 4 <= WJ_THREADS, "one word of hit flags per thread");
        if (threadIdx.x < WJ_LIST_CAP / 4) reinterpret_cast<uint32_t*>(S.lhit)[threadIdx.x] = 0;
        if (threadIdx.x < WJ_RUN_CAP / 4) reinterpret_cast<uint32_t*>(S.rhit)[threadIdx.x] = 0;
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < S.nlist; i += WJ_THREADS) S.list[i] = __ldg(C.list4 + S.l0 + i);
        for (uint32_t i = threadIdx.x; i < S.nrun; i += WJ_THREADS) S.run[i] = __ldg(C.ent4 + S.j0 + i);
        __syncthreads();
        const uint32_t lo = S.lo;
        uint32_t cnt = 0;
        uint32_t mask[WJ_QPT], bin[WJ_QPT];
#pragma unroll
        for (int k = 0; k < WJ_QPT; k++) {
            mask[k] = 0;
            const uint32_t q = q0 + k;
            if (q < p_lo || q >= p_hi) continue;
            uint32_t b = (s[k] - lo) >> bsh;
            if (b > nb - 1) b = nb - 1;
            bin[k] = b;
            uint32_t mk = 0;
            const uint32_t rs = s[k], re = e[k];
            const uint32_t o0 = S.boff[b], nL = S.boff[b + 1] - o0;
            if (S.fits && nL <= 16u) {
                // the common case, branch-free: the first four list members predicated, then the (rare) rest
                const uint32_t il = o0 - S.l0;
#pragma unroll
                for (uint32_t t = 0; t < 4; t++) {
                    const bool ok = t < nL;
                    const uint4 v = S.list[ok ? il + t : 0u];
                    if (ok && v.x < re && v.y > rs && wj_keep<MODE, INVERT>(v.x, v.y, rs, re)) { mk |= 1u << t; S.lhit[il + t] = 1; }
                }
                for (uint32_t t = 4; t < nL; t++) {
                    const uint4 v = S.list[il + t];
                    if (v.x < re && v.y > rs && wj_keep<MODE, INVERT>(v.x, v.y, rs, re)) { mk |= 1u << t; S.lhit[il + t] = 1; }
                }
            } else {
                uint32_t bit = 1u;
                for (uint32_t a = o0; a < o0 + nL; a++, bit <<= 1) {
                    const uint32_t i = a - S.l0;
                    const uint4 v = i < S.nlist ? S.list[i] : __ldg(C.list4 + a);
                    if (v.x < re && v.y > rs && wj_keep<MODE, INVERT>(v.x, v.y, rs, re)) {
                        mk |= (bit & 0xFFFFu) ? bit : 0x80000000u;
                        if (bit & 0xFFFF0000u || bit == 0) cnt++;   // beyond the position bits: counted here, not by the popcount below
                        if (i < S.nlist) S.lhit[i] = 1;
                        else if (!ent_hit[v.w]) ent_hit[v.w] = 1;
                    }
                }
            }
            {
                uint32_t bit = 1u << 16;
                for (uint32_t j = S.first[b + 1];; j++, bit <<= 1) {   // the sequence's sentinel (start 0xFFFFFFFF) ends it at the latest
                    const uint32_t i = j - S.j0;
                    const uint4 v = i < S.nrun ? S.run[i] : __ldg(C.ent4 + j);
                    if (v.x >= re) break;
                    if (v.y > rs && wj_keep<MODE, INVERT>(v.x, v.y, rs, re)) {
                        mk |= (bit & 0x7FFF0000u) ? bit : 0x80000000u;
                        if (!(bit & 0x7FFF0000u)) cnt++;
                        if (i < S.nrun) S.rhit[i] = 1;
                        else if (!ent_hit[j]) ent_hit[j] = 1;
                    }
                }
            }
            cnt += __popc(mk & 0x7FFFFFFFu);
            mask[k] = mk;
        }
        uint32_t tot;
        const uint32_t pos = block_scan_excl(cnt, OpSum(), 0u, sm, &tot);
        if (tot != 0) {   // uniform
            // hit flags of the slice: one idempotent byte store per hit root and CTA
            for (uint32_t i = threadIdx.x; i < S.nlist; i += WJ_THREADS)
                if (S.lhit[i]) {
                    uint8_t* h = ent_hit + S.list[i].w;
                    if (!*h) *h = 1;
                }
            for (uint32_t i = threadIdx.x; i < S.nrun; i += WJ_THREADS)
                if (S.rhit[i]) {
                    uint8_t* h = ent_hit + S.j0 + i;
                    if (!*h) *h = 1;
                }
            if (threadIdx.x == 0) sbase = atomicAdd(n_pairs, (unsigned long long)tot);
            if (EMIT) {
                __syncthreads();
                const unsigned long long gbase = sbase;
                if (gbase + tot <= cap) {
                    const bool staged = tot <= WJ_STAGE_PAIRS;
                    uint32_t* dst = staged ? stage + 3 * pos : pairs + 3 * (gbase + pos);
#pragma unroll
                    for (int k = 0; k < WJ_QPT; k++) {
                        const uint32_t mk = mask[k];
                        if (!mk) continue;
                        if (mk & 0x80000000u) {   // more hits than position bits: the same walk again
                            wj_visit<MODE, INVERT>(C, S, bin[k], s[k], e[k], [&](const uint4& v, uint32_t, uint32_t, bool) {
                                dst[0] = v.z; dst[1] = v.x; dst[2] = v.y;
                                dst += 3;
                            });
                            continue;
                        }
                        const uint32_t o0 = S.boff[bin[k]], f1 = S.first[bin[k] + 1];
                        for (uint32_t mm = mk & 0xFFFFu; mm; mm &= mm - 1) {
                            const uint32_t a = o0 + (uint32_t)__ffs(mm) - 1u, i = a - S.l0;
                            const uint4 v = i < S.nlist ? S.list[i] : __ldg(C.list4 + a);
                            dst[0] = v.z; dst[1] = v.x; dst[2] = v.y;
                            dst += 3;
                        }
                        for (uint32_t mm = mk >> 16; mm; mm &= mm - 1) {
                            const uint32_t j = f1 + (uint32_t)__ffs(mm) - 1u, i = j - S.j0;
                            const uint4 v = i < S.nrun ? S.run[i] : __ldg(C.ent4 + j);
                            dst[0] = v.z; dst[1] = v.x; dst[2] = v.y;
                            dst += 3;
                        }
                    }
                    if (staged) {
                        __syncthreads();
                        // 3*tot words to pairs + 3*gbase: scalar head up to 16-byte alignment, 16-byte body, scalar tail
                        uint32_t* g = pairs + 3 * gbase;
                        const uint32_t nwords = 3 * tot;
                        const uint32_t head = min(nwords, (uint32_t)((4u - ((uint32_t)(3ull * gbase) & 3u)) & 3u));
                        if (threadIdx.x < head) __stcs(g + threadIdx.x, stage[threadIdx.x]);
                        const uint32_t nvec = (nwords - head) >> 2;
                        uint4* gv = reinterpret_cast<uint4*>(g + head);
                        for (uint32_t i = threadIdx.x; i < nvec; i += WJ_THREADS) {
                            const uint32_t* sp = stage + head + 4 * i;   // shared reads are word-aligned only
                            __stcs(gv + i, make_uint4(sp[0], sp[1], sp[2], sp[3]));
                        }
                        const uint32_t done = head + 4 * nvec;
                        if (threadIdx.x < nwords - done) __stcs(g + done + threadIdx.x, stage[done + threadIdx.x]);
                    }
                }
            }
        }
        __syncthreads();  // the slice, the stage and sbase are reused by the next window of this tile
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
