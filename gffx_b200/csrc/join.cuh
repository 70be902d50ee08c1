// join.cuh — the small interval-list join left from r01's root join: depth / coverage use it for the tree intervals of
// IRREGULAR blocks only (blocks with several tree intervals or lines outside their one interval; a handful per index, none in
// a clean annotation — counts.cuh), plus the tile loader of those two kernels.  The root join itself is rootjoin.cuh /
// winjoin.cuh.
//
// Layout (join_build.h): per sequence the intervals sorted by start, split AIList-style into a few lists so that no interval
// of a list spans more than JOIN_LONG_K following starts; per list ent_se[j] = (start, end), ent_fid[j] = payload, and
// dir[b] = first j whose running max-end exceeds b << shift.  A query (seq, rs, re) probes dir[rs >> shift] and scans forward
// while start < re, testing end > rs: exactly `iv.start < end && iv.end > start` (src/utils/tree.rs:110) for any rs / re.
#pragma once
#include "scan.cuh"

namespace gffx {

constexpr int JOIN_THREADS = 256;
constexpr int JOIN_QPT = 4;                       // consecutive queries per thread (one 16 B load per column)
constexpr int JOIN_TILE = JOIN_THREADS * JOIN_QPT;
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

}  // namespace gffx
