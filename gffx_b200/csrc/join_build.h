// join_build.h — host-side construction of the flattened interval lists searched by join.cuh.
// One builder serves every interval set the engine joins regions against: the root intervals of
// the .rit trees (intersect), the intron-like gaps of multi-segment IDs and the tree intervals of
// irregular blocks (depth / coverage, counts.cu).
#pragma once
#include <cstdint>
#include <vector>

namespace gffx {

struct JoinItem {
    uint32_t seq, start, end, payload;
};

struct JoinListHost {  // mirrors JoinList of join.cuh (32 B)
    uint32_t base, count, dir_base, nbins, shift, pad0, pad1, pad2;
};

struct JoinHost {
    std::vector<uint32_t> seq_list0;      // n_seqs + 1
    std::vector<JoinListHost> lists;
    std::vector<uint32_t> ent_start, ent_end;  // interleaved on upload as uint2
    std::vector<uint32_t> ent_payload;    // payload of the item behind each entry
    std::vector<uint32_t> ent_item;       // index of the item behind each entry
    std::vector<uint32_t> dir;
    size_t n_entries() const { return ent_start.size(); }
};

// Per sequence: stable sort by (start, end); peel intervals that reach past the start of their
// `long_k`-th successor into further lists (at most `max_lists` per sequence); per list a
// direct-address directory over the running max-end with about 4 bins per entry.
void build_join_host(uint32_t n_seqs, const std::vector<JoinItem>& items, int long_k, int max_lists, JoinHost* out);

}  // namespace gffx
