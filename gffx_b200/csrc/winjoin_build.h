// winjoin_build.h — host-side construction of the window-addressed root join index searched by
// winjoin.cuh (K1).  Replaces the boxed centered tree of src/utils/tree.rs:30-64 with flat arrays
// that a window-bucketed region (partition.cuh) reaches in two dependent loads.
#pragma once
#include <cstdint>
#include <vector>
#include "join_build.h"

namespace gffx {

struct WinJoinHost {
    uint32_t shift = 0;       // window = 2^shift bases (same grid as the partition pass)
    uint32_t sub_bits = 0;    // each window has 2^sub_bits directory bins of 2^(shift-sub_bits) bases
    uint32_t n_short = 0;     // unified entry ids: [0,n_short) short entries, [n_short,n_ent) long entries
    std::vector<uint32_t> win0;        // n_seqs+1 (copy of the grid)
    std::vector<uint32_t> seq_send;    // n_seqs: one past the last short entry of the sequence
    // unified entry arrays (short entries per sequence sorted by (start,end), then the long ones)
    std::vector<uint32_t> ent_start, ent_end, ent_payload, ent_item;
    std::vector<uint32_t> sdir;        // n_win << sub_bits: first short entry whose running max-end exceeds the bin's first base
    std::vector<uint32_t> alw0;        // n_win+1: always-list of window w is [alw0[w], alw0[w+1])
    std::vector<uint32_t> alw_start, alw_end, alw_id;  // long entries overlapping the window (unified id)
    size_t n_entries() const { return ent_start.size(); }
};

// items: (seq,start,end,payload).  win0/shift: the window grid.  An entry is "long" when it reaches past
// the start of its `long_k`-th successor (peeled `levels` times); long entries are listed in every window
// they overlap, the rest stay in one start-sorted list per sequence with a running-max-end directory.
void build_win_join(uint32_t n_seqs, const std::vector<JoinItem>& items, const std::vector<uint32_t>& win0, uint32_t shift, int long_k,
                    int levels, WinJoinHost* out);

}  // namespace gffx
