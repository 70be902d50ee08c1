// filter.cuh — K3/K4/K5: the line-level filter of write_gff_match_only_by_coords
// (src/commands/intersect.rs:232-438; predicate gff_line_overlaps_queries :441-523, type filter
// gff_type_allowed :80-102) and the block merge of write_gff_output (src/utils/common.rs:199-229).
//
// The reference tests every line of every selected block against EVERY region of the line's
// sequence (O(L_sel x Q_seq)).  Here the fixed side (the lines) is turned into a rank space per
// sequence at index-build time and the regions are scattered onto it in one streaming pass:
//
//   X_seq   sorted distinct raw column-4/5 values of the sequence's lines (L2-resident)
//   slot k  of a sequence, k in [0,|X|], stands for "lb = #(X < v) == k"
//   A[slot] (u64, atomicMax)  tag(seq)|(qe+1) at slot lb(qs)   -> prefix-max  PMX
//   B[slot] (u64, atomicMin)  tag(seq)|qe     at slot ub(qs)   -> suffix-min  SMN
//   Ms/Me[cell] (u8 marks)    cell = 2*slot + (v == X[k])      -> next-set    NS/NE
//
// and each line (rank iS, iE of its raw S, E precomputed) is decided with O(1) look-ups:
//   Contained       exists q: qs<=S && E<=qe   <=>  PMX[iS] carries seq and value >= E+1
//   ContainsRegion  exists q: S<=qs && qe<=E   <=>  SMN[iS+1] carries seq and value <= E
//   Overlap  c1 qs<=S<=qe <=> PMX[iS]>=S+1 ; c2 same with E ; c3 some qs in [S,E] <=> NS[2iS+1] <= 2iE+1 ;
//            c4 some qe in [S,E] <=> NE[..] (only regions with qs>qe are marked in Me: for qs<=qe
//            c4 implies c1 or c3).
// Every clause is the reference's clause translated literally, so inverted regions (qs>qe) and
// inverted lines (S>E) behave as in the reference.  The sequence tag in the high bits lets one
// plain (unsegmented) scan serve all sequences.
//
// Nothing is cleared between jobs: every key carries the job's EPOCH above the sequence tag (A: epoch << 41, so a newer
// job's key beats any stale one under max; B: ~epoch << 40, likewise under min) and a mark byte IS the job's 8-bit
// epoch.  A reader compares the whole tag, so a cell that only holds an older job's value reads as empty.  When a
// counter wraps (2^23 - 1 jobs; 255 for the marks) the arrays are cleared once and the count restarts (engine.cu).
#pragma once
#include "scan.cuh"

namespace gffx {

struct RankSeq {  // per sequence, 32 B
    uint32_t x0, cnt, dir_base, nb, shift, slot0, canon, pad;
};

struct RankSpace {
    const uint32_t* X;
    const uint32_t* dir;
    const RankSeq* seqs;
    uint32_t n_seqs;
};

// lb = #(X_seq < v) as a local index in [0,cnt]; *eq = (X_seq[lb] == v)
__device__ __forceinline__ uint32_t rank_lb(const RankSpace& R, const RankSeq& S, uint32_t v, bool* eq) {
    const uint32_t b = v >> S.shift;
    *eq = false;
    if (b >= S.nb) return S.cnt;
    const uint2 d = make_uint2(__ldg(R.dir + S.dir_base + b), __ldg(R.dir + S.dir_base + b + 1));
    uint32_t lo = d.x, hi = d.y;  // X[lo..hi) are the values inside bin b
    const uint32_t* X = R.X + S.x0;
    while (hi - lo > 4) {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(X + mid) < v) lo = mid + 1; else hi = mid;
    }
    // at most 4 values left: five independent loads (the array is padded by 8 entries) instead of a chain of dependent ones;
    // the values are sorted and distinct, those past `hi` belong to later bins (> v), those past cnt to the next sequence
    uint32_t x[5];
#pragma unroll
    for (int k = 0; k < 5; k++) x[k] = __ldg(X + lo + k);
    uint32_t c = 0;
    bool e = false;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        if (k < 4) c += (lo + k < hi && x[k] < v) ? 1u : 0u;
        e |= lo + k < S.cnt && x[k] == v;
    }
    *eq = e;
    return lo + c;
}

__device__ __forceinline__ RankSeq load_rank_seq(const RankSpace& R, uint32_t seq) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(R.seqs + seq));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(R.seqs + seq) + 1);
    RankSeq S;
    S.x0 = a.x; S.cnt = a.y; S.dir_base = a.z; S.nb = a.w;
    S.shift = b.x; S.slot0 = b.y; S.canon = b.z; S.pad = 0;
    return S;
}

constexpr int FILT_THREADS = 256;
constexpr int FILT_QPT = 4;
constexpr uint64_t TAG_MAX_SHIFT = 33;  // A keys: (epoch<<8 | seq)<<33 | (qe+1)
constexpr uint64_t TAG_MIN_SHIFT = 32;  // B keys: ((~epoch & EPOCH_MAX)<<8 | seq)<<32 | qe
constexpr uint32_t EPOCH_MAX = (1u << 23) - 1u;   // 23 bits of epoch above 8 bits of (virtual) sequence

struct FilterArrays {
    unsigned long long* A;  // n_slots
    unsigned long long* B;  // n_slots
    uint8_t* Ms;            // 2*n_slots
    uint8_t* Me;            // 2*n_slots
    uint32_t* any_inv;      // set when some region has qs > qe: only then Me holds marks and NE is scanned / consulted
    uint32_t epoch;         // 1 .. EPOCH_MAX: the job's number since the arrays were last cleared
    uint32_t epoch8;        // 1 .. 255: the same for the mark bytes
};
__device__ __forceinline__ unsigned long long tag_max(uint32_t epoch, uint32_t seq) { return ((unsigned long long)epoch << 8) | seq; }
__device__ __forceinline__ unsigned long long tag_min(uint32_t epoch, uint32_t seq) { return ((unsigned long long)(~epoch & EPOCH_MAX) << 8) | seq; }

// one region (qs, qe) of sequence `seq` onto the rank space: 1-3 scattered L2 updates; (lb, eq) = rank_lb of qs
template <int MODE>
__device__ __forceinline__ void scatter_apply(const RankSpace& R, const RankSeq& S, const FilterArrays& F, uint32_t seq, uint32_t qs, uint32_t qe,
                                              uint32_t lb, bool eq) {
    const uint32_t slot = S.slot0 + lb;
    if (MODE == 0) {
        atomicMax(F.A + slot, (tag_max(F.epoch, seq) << TAG_MAX_SHIFT) | ((unsigned long long)qe + 1ull));
    } else if (MODE == 1) {
        atomicMin(F.B + slot + (eq ? 1u : 0u), (tag_min(F.epoch, seq) << TAG_MIN_SHIFT) | (unsigned long long)qe);
    } else {
        atomicMax(F.A + slot, (tag_max(F.epoch, seq) << TAG_MAX_SHIFT) | ((unsigned long long)qe + 1ull));
        F.Ms[2ull * slot + (eq ? 1u : 0u)] = (uint8_t)F.epoch8;
        if (qs > qe) {
            bool eq2;
            const uint32_t lb2 = rank_lb(R, S, qe, &eq2);
            F.Me[2ull * (S.slot0 + lb2) + (eq2 ? 1u : 0u)] = (uint8_t)F.epoch8;
            *F.any_inv = 1u;
        }
    }
}
template <int MODE>
__device__ __forceinline__ void scatter_region(const RankSpace& R, const RankSeq& S, const FilterArrays& F, uint32_t seq, uint32_t qs, uint32_t qe) {
    bool eq;
    const uint32_t lb = rank_lb(R, S, qs, &eq);
    scatter_apply<MODE>(R, S, F, seq, qs, qe, lb, eq);
}

struct LineArrays {
    const uint8_t* flags;    // LN_I_OK bit
    const uint32_t* blk;
    const uint32_t* type_id;
    const uint32_t* iseq;
    const uint32_t* S;
    const uint32_t* E;
    const uint32_t* iS;      // global slot of rank(S)  (slot0 + local rank)
    const uint32_t* iE;
    const uint64_t* off;
    const uint32_t* len;
    uint64_t n;
};

struct FilterScans {
    const unsigned long long* PMX;  // inclusive prefix max of A
    const unsigned long long* SMN;  // inclusive suffix min of B
    const uint32_t* NS;             // next marked cell >= c in Ms
    const uint32_t* NE;             // same for Me
    const uint32_t* ne_on;          // FilterArrays::any_inv: 0 = no mark in Me, NE was not scanned
    uint32_t epoch;                 // the job's epoch: part of every key's tag
};

// The per-line decision in two rounds of independent loads (a thread is latency bound here: one line, a handful of loads):
// round 1 = the line's own columns, round 2 = the scan cells they point at; nothing is loaded behind a branch.
struct LineCols {
    uint32_t seq, S, E, iS, iE;
};

template <int MODE>
__device__ __forceinline__ bool line_matches(const LineCols& c, const FilterScans& P) {
    const uint32_t seq = c.seq, S = c.S, E = c.E, iS = c.iS, iE = c.iE;
    if (MODE == 0) {
        const unsigned long long k = P.PMX[iS];
        return (k >> TAG_MAX_SHIFT) == tag_max(P.epoch, seq) && (k & ((1ull << TAG_MAX_SHIFT) - 1)) >= (unsigned long long)E + 1ull;
    } else if (MODE == 1) {
        const unsigned long long k = P.SMN[iS + 1];
        return (k >> TAG_MIN_SHIFT) == tag_min(P.epoch, seq) && (uint32_t)k <= E;
    } else {
        const unsigned long long mask = (1ull << TAG_MAX_SHIFT) - 1;
        const uint32_t c0 = 2u * iS + 1u, c1 = 2u * iE + 1u;
        const unsigned long long k1 = P.PMX[iS], k2 = P.PMX[iE];
        const uint32_t ns = P.NS[c0];
        const bool ne_on = *P.ne_on != 0u;
        const unsigned long long tag = tag_max(P.epoch, seq);
        if ((k1 >> TAG_MAX_SHIFT) == tag && (k1 & mask) >= (unsigned long long)S + 1ull) return true;  // qs<=S<=qe
        if ((k2 >> TAG_MAX_SHIFT) == tag && (k2 & mask) >= (unsigned long long)E + 1ull) return true;  // qs<=E<=qe
        if (S <= E) {
            if (ns <= c1) return true;  // S<=qs<=E
            if (ne_on && P.NE[c0] <= c1) return true;  // S<=qe<=E
        }
        return false;
    }
}

constexpr int LINE_THREADS = 256;
constexpr int LINE_ITEMS = 4;                            // consecutive lines per thread: 16-byte loads of the 4-byte columns
constexpr int LINE_TILE = LINE_THREADS * LINE_ITEMS;     // lines per CTA of the decision / compaction kernels
constexpr uint64_t LINE_FOLD_TILES = 16384;              // up to 16.8 M lines the compaction CTAs add up the tile counts themselves

// four consecutive elements of a column, i a multiple of 4 (the arrays are cudaMalloc'ed: 16-byte aligned)
__device__ __forceinline__ void load4(const uint32_t* __restrict__ a, uint64_t i, uint64_t n, uint32_t (&v)[4]) {
    if (i + 4 <= n) {
        const uint4 x = *reinterpret_cast<const uint4*>(a + i);
        v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
    } else {
#pragma unroll
        for (int j = 0; j < 4; j++) v[j] = i + j < n ? a[i + j] : 0u;
    }
}
__device__ __forceinline__ uint32_t load4b(const uint8_t* __restrict__ a, uint64_t i, uint64_t n) {   // four bytes, byte j in bits 8j..
    if (i + 4 <= n) return *reinterpret_cast<const uint32_t*>(a + i);
    uint32_t w = 0;
    for (int j = 0; j < 4; j++)
        if (i + j < n) w |= (uint32_t)a[i + j] << (8 * j);
    return w;
}
__device__ __forceinline__ void store4b(uint8_t* __restrict__ a, uint64_t i, uint64_t n, uint32_t w) {
    if (i + 4 <= n) { *reinterpret_cast<uint32_t*>(a + i) = w; return; }
    for (int j = 0; j < 4; j++)
        if (i + j < n) a[i + j] = (uint8_t)(w >> (8 * j));
}

// pass 1: per-line decision -> keep flag (byte) + per-CTA count.  A thread decides 4 consecutive lines: the columns come in
// as 16-byte loads, the scan cells they point at as four independent gathers.
template <int MODE>
__global__ void __launch_bounds__(LINE_THREADS)
    eval_lines_kernel(LineArrays L, FilterScans P, const uint8_t* __restrict__ blk_sel, const uint8_t* __restrict__ type_allow,
                      bool have_types, uint8_t* __restrict__ keep, uint32_t* __restrict__ tile_count) {
    __shared__ uint32_t sm[LINE_THREADS / 32 + 1];
    const uint64_t i = ((uint64_t)blockIdx.x * LINE_THREADS + threadIdx.x) * LINE_ITEMS;
    uint32_t k = 0;
    if (i < L.n) {
        const uint32_t fl = load4b(L.flags, i, L.n);
        uint32_t blk[4], ty[4] = {0, 0, 0, 0}, seq[4], S[4], E[4], iS[4], iE[4];
        load4(L.blk, i, L.n, blk);
        if (have_types) load4(L.type_id, i, L.n, ty);
        load4(L.iseq, i, L.n, seq); load4(L.S, i, L.n, S); load4(L.E, i, L.n, E); load4(L.iS, i, L.n, iS); load4(L.iE, i, L.n, iE);
        uint32_t w = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (i + j >= L.n) break;   // padded entries are 0: valid indices, but not lines
            bool ok = ((fl >> (8 * j)) & 1u) && blk_sel[blk[j]];
            if (have_types) ok = ok && ty[j] != 0xFFFFFFFFu && type_allow[ty[j]];
            const bool mt = line_matches<MODE>(LineCols{seq[j], S[j], E[j], iS[j], iE[j]}, P);   // iS / iE are 0 (a valid slot) without usable coordinates
            if (ok && mt) { w |= 1u << (8 * j); k++; }
        }
        store4b(keep, i, L.n, w);
    }
    uint32_t tot;
    block_scan_excl(k, OpSum(), 0u, sm, &tot);
    if (threadIdx.x == 0) tile_count[blockIdx.x] = tot;
}

// The same decision in two steps, for jobs spread over several GPUs: "does any region match this line" is an OR over the
// regions, hence over the shards, and so is "was this root hit".  Each shard evaluates its own regions (match_lines), the
// byte flags are MAX-reduced across the shards, and keep_lines applies the block selection and the type filter.
template <int MODE>
__global__ void __launch_bounds__(LINE_THREADS) match_lines_kernel(LineArrays L, FilterScans P, uint8_t* __restrict__ match) {
    const uint64_t i = ((uint64_t)blockIdx.x * LINE_THREADS + threadIdx.x) * LINE_ITEMS;
    if (i >= L.n) return;
    const uint32_t fl = load4b(L.flags, i, L.n);
    uint32_t seq[4], S[4], E[4], iS[4], iE[4];
    load4(L.iseq, i, L.n, seq); load4(L.S, i, L.n, S); load4(L.E, i, L.n, E); load4(L.iS, i, L.n, iS); load4(L.iE, i, L.n, iE);
    uint32_t w = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if (i + j >= L.n) break;
        const bool mt = line_matches<MODE>(LineCols{seq[j], S[j], E[j], iS[j], iE[j]}, P);
        if (((fl >> (8 * j)) & 1u) && mt) w |= 1u << (8 * j);
    }
    store4b(match, i, L.n, w);
}

static __global__ void __launch_bounds__(LINE_THREADS)
    keep_lines_kernel(LineArrays L, const uint8_t* __restrict__ match, const uint8_t* __restrict__ blk_sel,
                      const uint8_t* __restrict__ type_allow, bool have_types, uint8_t* __restrict__ keep, uint32_t* __restrict__ tile_count) {
    __shared__ uint32_t sm[LINE_THREADS / 32 + 1];
    const uint64_t i = ((uint64_t)blockIdx.x * LINE_THREADS + threadIdx.x) * LINE_ITEMS;
    uint32_t k = 0;
    if (i < L.n) {
        const uint32_t mt = load4b(match, i, L.n);
        uint32_t blk[4], ty[4] = {0, 0, 0, 0};
        load4(L.blk, i, L.n, blk);
        if (have_types) load4(L.type_id, i, L.n, ty);
        uint32_t w = 0;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            if (i + j >= L.n) break;
            bool ok = ((mt >> (8 * j)) & 0xFFu) && blk_sel[blk[j]];
            if (have_types) ok = ok && ty[j] != 0xFFFFFFFFu && type_allow[ty[j]];
            if (ok) { w |= 1u << (8 * j); k++; }
        }
        store4b(keep, i, L.n, w);
    }
    uint32_t tot;
    block_scan_excl(k, OpSum(), 0u, sm, &tot);
    if (threadIdx.x == 0) tile_count[blockIdx.x] = tot;
}

// pass 2: order-preserving compaction of (offset, offset+len).  The CTA adds up its predecessors' counts itself (<= a few
// thousand tiles of 1024 lines: a dozen loads per thread), so no scan pass runs between the two kernels; the last CTA
// reports the number of ranges.
static __global__ void __launch_bounds__(LINE_THREADS)
    emit_ranges_kernel(LineArrays L, const uint8_t* __restrict__ keep, const uint32_t* __restrict__ tile_count,
                       const unsigned long long* __restrict__ tile_off, ulonglong2* __restrict__ ranges, unsigned long long* __restrict__ bytes_out,
                       unsigned long long* __restrict__ n_out) {
    __shared__ unsigned long long sb[LINE_THREADS / 32 + 1];
    const uint64_t i = ((uint64_t)blockIdx.x * LINE_THREADS + threadIdx.x) * LINE_ITEMS;
    // offset, length and flag of the thread's 4 lines: loaded side by side, before anything is decided
    uint32_t kw = 0, len[4] = {0, 0, 0, 0};
    unsigned long long off[4] = {0, 0, 0, 0};
    if (i < L.n) {
        kw = load4b(keep, i, L.n);
        load4(L.len, i, L.n, len);
        if (i + 4 <= L.n) {
            const ulonglong2 a = *reinterpret_cast<const ulonglong2*>(L.off + i), b = *reinterpret_cast<const ulonglong2*>(L.off + i + 2);
            off[0] = a.x; off[1] = a.y; off[2] = b.x; off[3] = b.y;
        } else {
            for (int j = 0; j < 4; j++)
                if (i + j < L.n) off[j] = L.off[i + j];
        }
    }
    // ranges of the tiles in front of this one: added up here, or (tile_off != nullptr: a line table of > LINE_FOLD_TILES tiles,
    // where adding up would be quadratic) taken from a scan the host ran in between
    unsigned long long before = 0;
    if (tile_off) before = threadIdx.x == 0 ? tile_off[blockIdx.x] : 0ull;
    else for (uint32_t t = threadIdx.x; t < blockIdx.x; t += LINE_THREADS) before += tile_count[t];
    // one block scan over (count, bytes) packed per thread: count < 2^11 per CTA, bytes < 2^53
    uint32_t k = 0;
    unsigned long long bytes = 0;
#pragma unroll
    for (int j = 0; j < 4; j++)
        if ((kw >> (8 * j)) & 0xFFu) { k++; bytes += len[j]; }
    unsigned long long tb;
    block_scan_excl(before, OpSum(), 0ull, sb, &tb);
    const unsigned long long t0 = tb;
    __syncthreads();
    unsigned long long tot;
    unsigned long long pos = t0 + block_scan_excl((unsigned long long)k, OpSum(), 0ull, sb, &tot);
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 4; j++)
        if ((kw >> (8 * j)) & 0xFFu) ranges[pos++] = make_ulonglong2(off[j], off[j] + len[j]);
    unsigned long long btot;
    block_scan_excl(bytes, OpSum(), 0ull, sb, &btot);
    if (threadIdx.x == 0) {
        if (btot) atomicAdd(bytes_out, btot);
        if (blockIdx.x == gridDim.x - 1) *n_out = t0 + tot;
    }
}

// K5, one CTA: selected blocks (sorted by start) -> merged byte ranges, rules of common.rs:199-245:
// drop sentinel starts, merge while s <= current_end, keep cs < ce and ce <= file_len.
// Also counts the selected blocks.  out_counts[0] = n_ranges, [1] = n_selected, [2..3] = bytes (u64).
static __global__ void __launch_bounds__(1024)
    merge_blocks_kernel(const uint64_t* __restrict__ bstart, const uint64_t* __restrict__ bend, const uint8_t* __restrict__ blk_sel,
                        uint32_t n_blocks, uint64_t file_len, bool emit, ulonglong2* __restrict__ ranges,
                        unsigned long long* __restrict__ out_counts) {
    __shared__ unsigned long long smax[1024 / 32 + 1];
    __shared__ uint32_t ssum[1024 / 32 + 1];
    __shared__ unsigned long long sbytes[1024 / 32 + 1];
    // carried across chunks (same value in every thread)
    unsigned long long run_max = 0;      // max end over selected blocks so far
    uint32_t n_sel = 0, n_out = 0;
    unsigned long long cur_start = 0;    // start of the open segment
    bool open = false;
    unsigned long long bytes = 0;
    for (uint32_t base = 0; base < n_blocks; base += 1024) {
        const uint32_t b = base + threadIdx.x;
        const bool sel = b < n_blocks && blk_sel[b] && bstart[b] != 0xFFFFFFFFFFFFFFFFull;
        const unsigned long long s = sel ? bstart[b] : 0, e = sel ? bend[b] : 0;
        unsigned long long tot_max;
        unsigned long long ex_max = block_scan_excl((unsigned long long)e, OpMax(), 0ull, smax, &tot_max);
        __syncthreads();
        uint32_t tot_sel;
        const uint32_t ex_sel = block_scan_excl(sel ? 1u : 0u, OpSum(), 0u, ssum, &tot_sel);
        __syncthreads();
        const unsigned long long prev_max = ex_max > run_max ? ex_max : run_max;   // max end of selected blocks before b
        const bool first = sel && (n_sel + ex_sel == 0);
        const bool head = sel && (first || s > prev_max);
        // a head closes the previous segment [seg_start, prev_max); find seg_start: it is the start of the
        // previous head — carried via a second scan over (head ? index : 0) would need another pass, so
        // instead every head publishes itself and the NEXT head (or the tail) closes it.
        // Step 1: exclusive count of heads -> slot of each segment in `ranges` (unfiltered).
        uint32_t tot_head;
        const uint32_t ex_head = block_scan_excl(head ? 1u : 0u, OpSum(), 0u, ssum, &tot_head);
        __syncthreads();
        (void)ex_head;
        // Close segments: a segment ends right before each head (with end prev_max) and at the very end.
        // Thread with a head at b closes the segment that was open before it.
        // Its start is the most recent head start before b: exclusive max-scan of (head ? s+1 : 0).
        unsigned long long tot_hs;
        unsigned long long ex_hs = block_scan_excl(head ? s + 1ull : 0ull, OpMax(), 0ull, smax, &tot_hs);
        __syncthreads();
        unsigned long long seg_start_p1 = ex_hs ? ex_hs : (open ? cur_start + 1ull : 0ull);
        const bool closes = head && seg_start_p1 != 0;
        const unsigned long long cs = seg_start_p1 - 1ull, ce = prev_max;
        const bool good = closes && cs < ce && ce <= file_len;
        uint32_t tot_good;
        const uint32_t ex_good = block_scan_excl(good ? 1u : 0u, OpSum(), 0u, ssum, &tot_good);
        __syncthreads();
        if (good && emit) ranges[n_out + ex_good] = make_ulonglong2(cs, ce);
        unsigned long long tot_b;
        block_scan_excl(good ? ce - cs : 0ull, OpSum(), 0ull, sbytes, &tot_b);
        __syncthreads();
        bytes += tot_b;
        n_out += tot_good;
        // carry
        if (tot_hs) { cur_start = tot_hs - 1ull; open = true; }
        run_max = tot_max > run_max ? tot_max : run_max;
        n_sel += tot_sel;
    }
    if (threadIdx.x == 0) {
        if (open) {
            const unsigned long long cs = cur_start, ce = run_max;
            if (cs < ce && ce <= file_len) {
                if (emit) ranges[n_out] = make_ulonglong2(cs, ce);
                n_out++;
                bytes += ce - cs;
            }
        }
        out_counts[0] = n_out;
        out_counts[1] = n_sel;
        out_counts[2] = bytes;
    }
}

}  // namespace gffx
