// join.cuh — K1: the root-level join, replacing IntervalTree::query_interval (src/utils/tree.rs:98-121)
// and the per-region loop of query_features (src/commands/intersect.rs:139-165).
//
// Layout in HBM (built once per index, L2-resident: <= ~20 B per root interval):
//   per sequence the root intervals are sorted by start and split AIList-style into a few lists so
//   that no interval in a list spans more than JOIN_K following starts (a chromosome-long root would
//   otherwise drag every forward scan back to index 0).  Per list:
//     ent_se[j]  = (start,end)  packed 8 B, sorted by start       -> one 8 B load per candidate
//     ent_fid[j] = root_fid                                        -> read only for emitted pairs
//     dir[b]     = first j whose running max-end exceeds b<<shift  -> direct-address replacement of the
//                  lower_bound on the running max-end: every earlier interval ends at or before the bin
//   A query (seq, rs, re) probes dir[rs>>shift] and scans forward while start < re, testing end > rs:
//   exactly `iv.start < end && iv.end > start` (tree.rs:110) for any rs/re, including rs >= re.
//
// Two passes: join_count_kernel (per-tile pair counts + per-entry hit flags), a device scan of the
// tile counts, join_emit_kernel (re-search, CTA-level exclusive scan, shared-memory staging,
// coalesced 4-byte stores of the 12-byte pairs).
#pragma once
#include "scan.cuh"

namespace gffx {

constexpr int JOIN_THREADS = 256;
constexpr int JOIN_QPT = 4;                       // consecutive queries per thread (one 16 B load per column)
constexpr int JOIN_TILE = JOIN_THREADS * JOIN_QPT;
constexpr int JOIN_STAGE_PAIRS = 3072;            // 36 KB of shared memory for pair staging
constexpr int JOIN_LONG_K = 16;     // an interval that reaches past the start of its 16th successor is "long"
constexpr int JOIN_MAX_LISTS = 8;   // per sequence

struct JoinList {
    uint32_t base, count, dir_base, nbins, shift, pad0, pad1, pad2;
};

struct JoinIndex {
    const uint32_t* seq_list0;  // n_seqs+1: lists of sequence s are [seq_list0[s], seq_list0[s+1])
    const JoinList* lists;
    const uint2* ent_se;
    const uint32_t* ent_fid;
    const uint32_t* dir;
    uint32_t n_seqs;
};

// mode test of src/commands/intersect.rs:145-158 and the XOR of :161
template <int MODE, bool INVERT> __device__ __forceinline__ bool keep_pair(uint32_t s, uint32_t e, uint32_t rs, uint32_t re) {
    bool keep = MODE == 0 ? (s >= rs && e <= re) : MODE == 1 ? (s <= rs && e >= re) : true;
    return INVERT != keep;
}

template <int MODE, bool INVERT, class F>
__device__ __forceinline__ void join_query(const JoinIndex& J, uint32_t seq, uint32_t rs, uint32_t re, F&& f) {
    const uint32_t l0 = __ldg(J.seq_list0 + seq), l1 = __ldg(J.seq_list0 + seq + 1);
    for (uint32_t l = l0; l < l1; l++) {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(J.lists + l));      // base,count,dir_base,nbins
        const uint32_t shift = __ldg(&J.lists[l].shift);
        const uint32_t b = rs >> shift;
        if (b >= a.w) continue;  // rs lies beyond every end of this list
        uint32_t j = __ldg(J.dir + a.z + b);
        const uint32_t jend = a.x + a.y;
        for (; j < jend; j++) {
            const uint2 se = __ldg(J.ent_se + j);
            if (se.x >= re) break;
            if (se.y > rs && keep_pair<MODE, INVERT>(se.x, se.y, rs, re)) f(j, se);
        }
    }
}

struct Query4 {
    uint32_t c[JOIN_QPT], s[JOIN_QPT], e[JOIN_QPT];
    int n;
};

// thread t of tile `tile` owns queries [tile*JOIN_TILE + t*4, +4) of the window-bucketed arrays
// (partition.cuh: chr[] and (start,end)[]; cudaMalloc'ed, so the vector loads are aligned)
__device__ __forceinline__ void load_queries(const uint32_t* __restrict__ chr, const uint2* __restrict__ se, uint64_t n, uint64_t q0,
                                             Query4& Q) {
    if (q0 + JOIN_QPT <= n) {
        const uint4 c = __ldcs(reinterpret_cast<const uint4*>(chr + q0));
        const uint4 a = __ldcs(reinterpret_cast<const uint4*>(se + q0));
        const uint4 b = __ldcs(reinterpret_cast<const uint4*>(se + q0) + 1);
        Q.c[0] = c.x; Q.c[1] = c.y; Q.c[2] = c.z; Q.c[3] = c.w;
        Q.s[0] = a.x; Q.e[0] = a.y; Q.s[1] = a.z; Q.e[1] = a.w;
        Q.s[2] = b.x; Q.e[2] = b.y; Q.s[3] = b.z; Q.e[3] = b.w;
        Q.n = JOIN_QPT;
    } else {
        Q.n = 0;
#pragma unroll
        for (int k = 0; k < JOIN_QPT; k++)
            if (q0 + k < n) {
                const uint2 v = se[q0 + k];
                Q.c[k] = chr[q0 + k]; Q.s[k] = v.x; Q.e[k] = v.y;
                Q.n = k + 1;
            }
    }
}

template <int MODE, bool INVERT>
__global__ void __launch_bounds__(JOIN_THREADS)
    join_count_kernel(JoinIndex J, const uint32_t* __restrict__ chr, const uint2* __restrict__ se,
                      uint64_t n, uint8_t* __restrict__ ent_hit, uint32_t* __restrict__ tile_count,
                      uint32_t* __restrict__ err) {
    __shared__ uint32_t sm[JOIN_THREADS / 32 + 1];
    const uint64_t q0 = (uint64_t)blockIdx.x * JOIN_TILE + (uint64_t)threadIdx.x * JOIN_QPT;
    Query4 Q;
    Q.n = 0;
    if (q0 < n) load_queries(chr, se, n, q0, Q);
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < JOIN_QPT; k++) {
        if (k < Q.n) {
            if (Q.c[k] >= J.n_seqs) { *err = 1; continue; }
            join_query<MODE, INVERT>(J, Q.c[k], Q.s[k], Q.e[k], [&](uint32_t j, uint2) {
                cnt++;
                // ~230 regions hit every root: test in L2 first so that the byte is stored O(1) times, not O(hits)
                if (!__ldcg(ent_hit + j)) ent_hit[j] = 1;
            });
        }
    }
    uint32_t tot;
    block_scan_excl(cnt, OpSum(), 0u, sm, &tot);
    if (threadIdx.x == 0) tile_count[blockIdx.x] = tot;
}

template <int MODE, bool INVERT>
__global__ void __launch_bounds__(JOIN_THREADS)
    join_emit_kernel(JoinIndex J, const uint32_t* __restrict__ chr, const uint2* __restrict__ se,
                     uint64_t n, const uint64_t* __restrict__ tile_off, uint32_t* __restrict__ pairs) {
    __shared__ uint32_t sm[JOIN_THREADS / 32 + 1];
    __shared__ uint32_t stage[JOIN_STAGE_PAIRS * 3];
    const uint64_t q0 = (uint64_t)blockIdx.x * JOIN_TILE + (uint64_t)threadIdx.x * JOIN_QPT;
    Query4 Q;
    Q.n = 0;
    if (q0 < n) load_queries(chr, se, n, q0, Q);
    uint32_t cnt = 0;
#pragma unroll
    for (int k = 0; k < JOIN_QPT; k++)
        if (k < Q.n && Q.c[k] < J.n_seqs) join_query<MODE, INVERT>(J, Q.c[k], Q.s[k], Q.e[k], [&](uint32_t, uint2) { cnt++; });
    uint32_t tot;
    uint32_t pos = block_scan_excl(cnt, OpSum(), 0u, sm, &tot);
    const uint64_t gbase = tile_off[blockIdx.x];
    if (tot == 0) return;
    const bool staged = tot <= JOIN_STAGE_PAIRS;
    uint32_t* dst = staged ? stage : pairs + 3 * gbase;
#pragma unroll
    for (int k = 0; k < JOIN_QPT; k++)
        if (k < Q.n && Q.c[k] < J.n_seqs)
            join_query<MODE, INVERT>(J, Q.c[k], Q.s[k], Q.e[k], [&](uint32_t j, uint2 se) {
                dst[3 * pos + 0] = __ldg(J.ent_fid + j);
                dst[3 * pos + 1] = se.x;
                dst[3 * pos + 2] = se.y;
                pos++;
            });
    if (staged) {
        __syncthreads();
        uint32_t* g = pairs + 3 * gbase;
        for (uint32_t i = threadIdx.x; i < 3 * tot; i += JOIN_THREADS) __stcs(g + i, stage[i]);
    }
}

// per-entry hit flags -> per-block selection flags; entries whose fid has no .gof record are
// appended to warn_fids (the reference warns about them, intersect.rs:269-272).
static __global__ void select_blocks_kernel(const uint8_t* __restrict__ ent_hit, const uint32_t* __restrict__ ent_blk,
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
