#include "winjoin_build.h"
#include <algorithm>

namespace gffx {

void build_win_bins(uint32_t n_seqs, const std::vector<JoinItem>& items, const std::vector<uint32_t>& win0, uint32_t shift,
                    uint32_t max_sub, uint64_t max_list, WinBinsHost* out) {
    *out = WinBinsHost();
    out->shift = shift;
    const uint32_t n_win = win0[n_seqs];
    out->n_win = n_win;
    out->win_seq.assign(n_win, 0);
    for (uint32_t s = 0; s < n_seqs; s++)
        for (uint32_t w = win0[s]; w < win0[s + 1]; w++) out->win_seq[w] = s;

    // ---- start-sorted entries per sequence --------------------------------------------------------
    std::vector<std::vector<uint32_t>> by_seq(n_seqs);
    for (size_t i = 0; i < items.size(); i++) by_seq[items[i].seq].push_back((uint32_t)i);
    std::vector<uint32_t> base(n_seqs + 1, 0);
    for (uint32_t s = 0; s < n_seqs; s++) {
        std::vector<uint32_t>& ids = by_seq[s];
        std::stable_sort(ids.begin(), ids.end(), [&](uint32_t a, uint32_t b) {
            if (items[a].start != items[b].start) return items[a].start < items[b].start;
            return items[a].end < items[b].end;
        });
        base[s] = (uint32_t)out->ent_start.size();
        for (uint32_t id : ids) {
            out->ent_start.push_back(items[id].start);
            out->ent_end.push_back(items[id].end);
            out->ent_payload.push_back(items[id].payload);
            out->ent_item.push_back(id);
        }
        out->ent_start.push_back(0xFFFFFFFFu);  // sentinel: start >= any region end stops every run
        out->ent_end.push_back(0xFFFFFFFFu);
        out->ent_payload.push_back(0);
        out->ent_item.push_back(0xFFFFFFFFu);
    }
    base[n_seqs] = (uint32_t)out->ent_start.size();

    // ---- bin resolution: about 4 bins per interval, at most 2^max_sub per window, lists within max_list ---------
    uint32_t sub = 0;
    while (sub < max_sub && sub < shift && ((uint64_t)n_win << sub) < 4ull * std::max<size_t>(items.size(), 1)) sub++;
    for (;; sub--) {
        const uint32_t bsh = shift - sub;
        uint64_t tot = 0;
        for (const JoinItem& it : items) {
            const uint32_t last = it.end > it.start ? it.end - 1 : it.start;
            tot += (uint64_t)(last >> bsh) - (it.start >> bsh) + 1;
        }
        if (tot <= max_list || sub == 0) break;
    }
    out->sub = sub;
    const uint32_t nb = 1u << sub, bsh = shift - sub;

    // ---- per window: first[] and the bin lists -----------------------------------------------------
    out->first.assign((size_t)n_win * (nb + 1), 0);
    out->boff.assign((size_t)n_win * (nb + 1), 0);
    std::vector<std::vector<uint32_t>> bins;  // bins of one sequence (all its windows)
    for (uint32_t s = 0; s < n_seqs; s++) {
        const uint32_t nw = win0[s + 1] - win0[s];
        const uint64_t n_bins = (uint64_t)nw << sub;
        const uint32_t b0 = base[s], cnt = base[s + 1] - base[s] - 1;  // without the sentinel
        bins.assign((size_t)n_bins, {});
        for (uint32_t k = 0; k < cnt; k++) {
            const uint32_t st = out->ent_start[b0 + k], en = out->ent_end[b0 + k];
            // an interval hits a region iff start < re && end > rs (tree.rs:110), which an EMPTY or inverted interval
            // (end <= start; never written by `gffx index`, but gffx_index_create takes any arrays) can satisfy too: it is
            // listed in the bin of its start only.  Intervals starting beyond the grid are reached by the runs alone.
            if ((uint64_t)(st >> bsh) >= n_bins) continue;
            const uint64_t bf = st >> bsh, bl = en > st ? std::min<uint64_t>((en - 1) >> bsh, n_bins - 1) : bf;
            for (uint64_t b = bf; b <= bl; b++) bins[(size_t)b].push_back(b0 + k);
        }
        size_t j = 0;
        for (uint32_t w = 0; w < nw; w++) {
            const size_t row = (size_t)(win0[s] + w) * (nb + 1);
            for (uint32_t b = 0; b <= nb; b++) {
                const uint64_t x = ((uint64_t)w << shift) + ((uint64_t)b << bsh);  // first base of the bin (b == nb: of the next window)
                while (j < cnt && (uint64_t)out->ent_start[b0 + j] < x) j++;
                out->first[row + b] = b0 + (uint32_t)j;
                out->boff[row + b] = (uint32_t)out->list.size();
                if (b < nb) {
                    const std::vector<uint32_t>& v = bins[((size_t)w << sub) + b];
                    out->list.insert(out->list.end(), v.begin(), v.end());
                }
            }
        }
    }
}

}  // namespace gffx
