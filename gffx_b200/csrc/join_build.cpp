#include "join_build.h"
#include <algorithm>

namespace gffx {

void build_join_host(uint32_t n_seqs, const std::vector<JoinItem>& items, int long_k, int max_lists, JoinHost* out) {
    *out = JoinHost();
    std::vector<std::vector<uint32_t>> by_seq(n_seqs);
    for (size_t i = 0; i < items.size(); i++) by_seq[items[i].seq].push_back((uint32_t)i);
    out->seq_list0.assign((size_t)n_seqs + 1, 0);
    std::vector<uint32_t> cur, keep, longs, pmax;
    for (uint32_t s = 0; s < n_seqs; s++) {
        out->seq_list0[s] = (uint32_t)out->lists.size();
        std::vector<uint32_t>& ids = by_seq[s];
        std::stable_sort(ids.begin(), ids.end(), [&](uint32_t a, uint32_t b) {
            if (items[a].start != items[b].start) return items[a].start < items[b].start;
            return items[a].end < items[b].end;
        });
        cur = ids;
        for (int level = 0; !cur.empty(); level++) {
            keep.clear();
            longs.clear();
            if (level == max_lists - 1 || cur.size() <= (size_t)4 * long_k) keep = cur;
            else
                for (size_t i = 0; i < cur.size(); i++) {
                    const bool is_long = i + long_k < cur.size() && items[cur[i]].end > items[cur[i + long_k]].start;
                    (is_long ? longs : keep).push_back(cur[i]);
                }
            if (!keep.empty()) {
                JoinListHost L{};
                L.base = (uint32_t)out->ent_start.size();
                L.count = (uint32_t)keep.size();
                uint32_t maxend = 0;
                for (uint32_t i : keep) maxend = std::max(maxend, items[i].end);
                uint32_t shift = 6;
                while (((uint64_t)maxend >> shift) + 1 > std::max<uint64_t>(64, 4ull * keep.size())) shift++;
                L.shift = shift;
                L.nbins = (maxend >> shift) + 1;
                L.dir_base = (uint32_t)out->dir.size();
                pmax.resize(keep.size());
                uint32_t pm = 0;
                for (size_t k = 0; k < keep.size(); k++) {
                    const JoinItem& it = items[keep[k]];
                    out->ent_start.push_back(it.start);
                    out->ent_end.push_back(it.end);
                    out->ent_payload.push_back(it.payload);
                    out->ent_item.push_back(keep[k]);
                    pm = std::max(pm, it.end);
                    pmax[k] = pm;
                }
                // dir[b] = first entry (absolute) whose inclusive running max-end exceeds b<<shift:
                // every earlier entry of the list ends at or before the bin's first coordinate
                uint32_t j = 0;
                for (uint32_t b = 0; b < L.nbins; b++) {
                    const uint64_t lo = (uint64_t)b << shift;
                    while (j < keep.size() && (uint64_t)pmax[j] <= lo) j++;
                    out->dir.push_back(L.base + j);
                }
                out->lists.push_back(L);
            }
            cur.swap(longs);
        }
    }
    out->seq_list0[n_seqs] = (uint32_t)out->lists.size();
}

}  // namespace gffx
