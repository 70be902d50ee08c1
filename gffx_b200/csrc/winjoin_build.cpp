#include "winjoin_build.h"
#include <algorithm>

namespace gffx {

void build_win_join(uint32_t n_seqs, const std::vector<JoinItem>& items, const std::vector<uint32_t>& win0, uint32_t shift, int long_k,
                    int levels, WinJoinHost* out) {
    *out = WinJoinHost();
    out->shift = shift;
    out->win0 = win0;
    const uint32_t n_win = win0[n_seqs];
    std::vector<std::vector<uint32_t>> by_seq(n_seqs);
    for (size_t i = 0; i < items.size(); i++) by_seq[items[i].seq].push_back((uint32_t)i);

    // ---- split every sequence into its short list and its long entries ---------------------------
    std::vector<std::vector<uint32_t>> shorts(n_seqs), longs(n_seqs);
    size_t n_short = 0;
    for (uint32_t s = 0; s < n_seqs; s++) {
        std::vector<uint32_t>& ids = by_seq[s];
        std::stable_sort(ids.begin(), ids.end(), [&](uint32_t a, uint32_t b) {
            if (items[a].start != items[b].start) return items[a].start < items[b].start;
            return items[a].end < items[b].end;
        });
        std::vector<uint32_t> cur = ids, keep;
        for (int level = 0; level < levels && cur.size() > (size_t)long_k; level++) {
            keep.clear();
            size_t peeled = 0;
            for (size_t i = 0; i < cur.size(); i++) {
                const bool is_long = i + long_k < cur.size() && items[cur[i]].end > items[cur[i + long_k]].start;
                if (is_long) { longs[s].push_back(cur[i]); peeled++; }
                else keep.push_back(cur[i]);
            }
            cur.swap(keep);
            if (!peeled) break;
        }
        shorts[s] = std::move(cur);
        n_short += shorts[s].size();
    }
    out->n_short = (uint32_t)n_short + n_seqs;  // + one sentinel per sequence

    // directory resolution: about 4 bins per short entry, at most 2^10 per window
    uint32_t sub_bits = 0;
    while (sub_bits < 10 && sub_bits < shift && ((uint64_t)n_win << sub_bits) < 4ull * std::max<size_t>(n_short, 1)) sub_bits++;
    out->sub_bits = sub_bits;
    const uint32_t sub_shift = shift - sub_bits;

    // ---- unified entry arrays: shorts of all sequences, then the longs ---------------------------
    auto push_entry = [&](uint32_t item) {
        out->ent_start.push_back(items[item].start);
        out->ent_end.push_back(items[item].end);
        out->ent_payload.push_back(items[item].payload);
        out->ent_item.push_back(item);
    };
    out->seq_send.assign(n_seqs, 0);
    out->sdir.assign((size_t)n_win << sub_bits, 0);
    std::vector<uint32_t> pmax;
    for (uint32_t s = 0; s < n_seqs; s++) {
        const uint32_t base = (uint32_t)out->ent_start.size();
        const std::vector<uint32_t>& sh = shorts[s];
        pmax.resize(sh.size());
        uint32_t pm = 0;
        for (size_t k = 0; k < sh.size(); k++) {
            push_entry(sh[k]);
            pm = std::max(pm, items[sh[k]].end);
            pmax[k] = pm;
        }
        out->seq_send[s] = base + (uint32_t)sh.size();
        // sentinel: stops every forward scan (start >= any region end) without a bounds check
        out->ent_start.push_back(0xFFFFFFFFu);
        out->ent_end.push_back(0);
        out->ent_payload.push_back(0);
        out->ent_item.push_back(0xFFFFFFFFu);
        // sdir[(w << sub_bits) + b] = first short entry whose inclusive running max-end exceeds the bin's first base
        size_t j = 0;
        for (uint32_t w = win0[s]; w < win0[s + 1]; w++) {
            const uint64_t lo = (uint64_t)(w - win0[s]) << shift;
            for (uint32_t b = 0; b < (1u << sub_bits); b++) {
                const uint64_t x = lo + ((uint64_t)b << sub_shift);
                while (j < sh.size() && (uint64_t)pmax[j] <= x) j++;
                out->sdir[((size_t)w << sub_bits) + b] = base + (uint32_t)j;
            }
        }
    }
    // long entries: unified ids continue after the shorts; per window the ones that overlap it
    std::vector<std::vector<uint32_t>> per_win(n_win);
    for (uint32_t s = 0; s < n_seqs; s++) {
        const uint32_t nw = win0[s + 1] - win0[s];
        for (uint32_t item : longs[s]) {
            const uint32_t id = (uint32_t)out->ent_start.size();
            push_entry(item);
            const uint32_t st = items[item].start, en = items[item].end;
            const uint32_t w_first = std::min(st >> shift, nw - 1);
            const uint32_t w_last = std::min((en > st ? en - 1 : st) >> shift, nw - 1);
            for (uint32_t w = w_first; w <= w_last; w++) per_win[win0[s] + w].push_back(id);
        }
    }
    out->alw0.assign((size_t)n_win + 1, 0);
    for (uint32_t w = 0; w < n_win; w++) {
        out->alw0[w] = (uint32_t)out->alw_id.size();
        std::vector<uint32_t>& v = per_win[w];
        std::sort(v.begin(), v.end(), [&](uint32_t a, uint32_t b) {
            if (out->ent_start[a] != out->ent_start[b]) return out->ent_start[a] < out->ent_start[b];
            return a < b;
        });
        for (uint32_t id : v) {
            out->alw_start.push_back(out->ent_start[id]);
            out->alw_end.push_back(out->ent_end[id]);
            out->alw_id.push_back(id);
        }
    }
    out->alw0[n_win] = (uint32_t)out->alw_id.size();
}

}  // namespace gffx
