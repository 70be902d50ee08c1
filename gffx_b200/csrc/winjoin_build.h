// winjoin_build.h — host-side construction of the window-addressed root join index searched by
// winjoin.cuh (K1).  Replaces the boxed centered tree of src/utils/tree.rs:30-64 with flat arrays
// whose per-window slices fit in shared memory.
#pragma once
#include <cstdint>
#include <vector>

namespace gffx {

struct JoinItem {   // an interval of (virtual) sequence `seq` with a payload
    uint32_t seq, start, end, payload;
};

// Bin-list form (K1, winjoin.cuh).  Every window of the partition grid is cut into 2^sub bins; bin b of window w owns
//   list[boff[w,b] .. boff[w,b+1])  the intervals that overlap the bin (start < bin end && end > bin start), by start
//   first[w,b]                      the first interval of the sequence's start-sorted array with start >= bin start
// so the intervals overlapping a region that starts in bin b are: members of bin b's list (tested), followed by the
// run of the start-sorted array from first[w,b+1] while start < region end (their start lies beyond the bin, hence
// beyond the region start).  No tree, no running max-end, no back-tracking over nested or chromosome-spanning
// intervals: they simply sit in the list of every bin they cover.
struct WinBinsHost {
    uint32_t shift = 0, sub = 0, n_win = 0;
    std::vector<uint32_t> win_seq;   // n_win: sequence of each window
    // per sequence the intervals sorted by (start,end), one (0xFFFFFFFF,0xFFFFFFFF) sentinel after each sequence
    std::vector<uint32_t> ent_start, ent_end, ent_payload, ent_item;
    std::vector<uint32_t> first;     // n_win * (nb + 1), nb = 2^sub: cell nb of a window = first interval beyond the window
    std::vector<uint32_t> boff;      // n_win * (nb + 1): offsets into `list`; cell nb of a window = end of its last bin
    std::vector<uint32_t> list;      // entry ids
    size_t n_entries() const { return ent_start.size(); }
};

// max_list: the bins are made coarser until the lists hold at most this many ids (long intervals are listed once per
// bin they cover)
void build_win_bins(uint32_t n_seqs, const std::vector<JoinItem>& items, const std::vector<uint32_t>& win0, uint32_t shift,
                    uint32_t max_sub, uint64_t max_list, WinBinsHost* out);

}  // namespace gffx
