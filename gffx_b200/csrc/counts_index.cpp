#include "counts_index.h"
#include <algorithm>
#include <numeric>

namespace gffx {

void build_counts_index(const HostIndex& H, CountsIndexHost* out) {
    CountsIndexHost& C = *out;
    C = CountsIndexHost();
    const LineTable& L = H.lines;
    const VSeqMap& V = H.vmap;
    const uint32_t n_seqs = (uint32_t)V.n_vseqs();   // spaces 0 .. n_seqs-1 are the engine's VIRTUAL sequences (host_index.h)
    const size_t nb = H.blocks.size();
    C.n_ids = (uint32_t)H.ids.size();

    // tree intervals per block
    std::vector<uint32_t> br_off(nb + 1, 0), br_list(0);
    for (size_t i = 0; i < H.roots.size(); i++)
        if (H.root_blk[i] != NONE32) br_off[H.root_blk[i] + 1]++;
    for (size_t b = 0; b < nb; b++) br_off[b + 1] += br_off[b];
    br_list.resize(br_off[nb]);
    {
        std::vector<uint32_t> fill(br_off.begin(), br_off.end() - 1);
        for (size_t i = 0; i < H.roots.size(); i++)
            if (H.root_blk[i] != NONE32) br_list[fill[H.root_blk[i]]++] = (uint32_t)i;
    }

    C.blk_space.assign(nb, NONE32);
    uint32_t n_spaces = n_seqs;
    std::vector<std::vector<uint32_t>> id_units(C.n_ids);
    std::vector<JoinItem> gap_items, irr_items;
    std::vector<uint32_t> irr_item_grp, irr_item_pos, irr_item_base, irr_item_blk;
    std::vector<uint32_t> sp_rmin(n_seqs, 0xFFFFFFFFu), sp_rmax(n_seqs, 0);   // hull of the regular blocks' tree intervals per sequence space
    std::vector<uint32_t> ord;
    C.u_line0.push_back(0);
    C.u_seg0.push_back(0);
    C.grp_off.push_back(0);

    for (size_t b = 0; b < nb; b++) {
        const uint32_t nroots = br_off[b + 1] - br_off[b];
        if (nroots == 0) continue;  // never routed
        const uint64_t l0 = H.blk_line0[b], l1 = H.blk_line0[b + 1];
        ord.clear();
        for (uint64_t i = l0; i < l1; i++)
            if (L.flags[i] & LN_D_OK) ord.push_back((uint32_t)i);
        if (ord.empty()) continue;
        // regular <=> one tree interval and every line nested inside it
        bool regular = nroots == 1;
        const RootIv& r0 = H.roots[br_list[br_off[b]]];
        if (regular)
            for (uint32_t i : ord)
                if (L.fs[i] < r0.start || L.fe[i] > r0.end) { regular = false; break; }
        // a regular block's lines live on its root's virtual sequence, moved by the root sequence's base; a private space keeps
        // the lines' own coordinates and the regions routed to it are moved back (irr_base / the root join's ent_base)
        uint32_t space, vbase = 0;
        if (regular) {
            space = V.seq[r0.seq].vseq; vbase = V.seq[r0.seq].base;
            sp_rmin[space] = std::min(sp_rmin[space], vbase + r0.start);
            sp_rmax[space] = std::max(sp_rmax[space], vbase + r0.end);
        }
        else {
            space = n_spaces++;
            C.blk_space[b] = space;
            C.n_irregular++;
            const uint32_t grp = nroots > 1 ? (uint32_t)(C.grp_off.size() - 1) : NONE32;
            for (uint32_t k = 0; k < nroots; k++) {
                const RootIv& r = H.roots[br_list[br_off[b] + k]];
                const VSeqMap::Seq& m = V.seq[r.seq];
                irr_items.push_back(JoinItem{m.vseq, m.base + r.start, m.base + r.end, space});
                irr_item_grp.push_back(grp);
                irr_item_pos.push_back(k);
                irr_item_base.push_back(m.base);
                irr_item_blk.push_back((uint32_t)b);
                if (grp != NONE32) { C.grp_seq.push_back(m.vseq); C.grp_s.push_back(m.base + r.start); C.grp_e.push_back(m.base + r.end); }
            }
            if (grp != NONE32) C.grp_off.push_back((uint32_t)C.grp_seq.size());
        }
        // units: lines grouped by ID; chrom = column 1 of the ID's first line in the block (depth.rs:148-153)
        std::stable_sort(ord.begin(), ord.end(), [&](uint32_t a, uint32_t c) { return L.id[a] < L.id[c]; });
        size_t p = 0;
        while (p < ord.size()) {
            size_t q = p;
            while (q < ord.size() && L.id[ord[q]] == L.id[ord[p]]) q++;
            const uint32_t u = (uint32_t)C.u_space.size();
            C.u_space.push_back(space);
            C.u_chrom.push_back(L.chrom[ord[p]]);  // stable sort: ord[p] is the first line in file order
            C.u_blk.push_back((uint32_t)b);
            C.u_base.push_back(vbase);
            C.u_rs.push_back(regular ? vbase + r0.start : NONE32);
            C.u_re.push_back(regular ? vbase + r0.end : NONE32);
            id_units[L.id[ord[p]]].push_back(u);
            std::sort(ord.begin() + p, ord.begin() + q, [&](uint32_t a, uint32_t c) {
                if (L.fs[a] != L.fs[c]) return L.fs[a] < L.fs[c];
                return L.fe[a] < L.fe[c];
            });
            const uint32_t seg_first = (uint32_t)C.g_s.size();
            uint32_t cs = vbase + L.fs[ord[p]], ce = vbase + L.fe[ord[p]];
            for (size_t k = p; k < q; k++) {
                const uint32_t s = vbase + L.fs[ord[k]], e = vbase + L.fe[ord[k]];
                C.l_fs.push_back(s);
                C.l_fe.push_back(e);
                if (s <= ce) ce = std::max(ce, e);
                else {
                    C.g_s.push_back(cs); C.g_e.push_back(ce);
                    cs = s; ce = e;
                }
            }
            C.g_s.push_back(cs); C.g_e.push_back(ce);
            for (uint32_t g = seg_first; g + 1 < C.g_s.size(); g++) gap_items.push_back(JoinItem{space, C.g_e[g], C.g_s[g + 1], g});
            C.u_line0.push_back((uint32_t)C.l_fs.size());
            C.u_seg0.push_back((uint32_t)C.g_s.size());
            p = q;
        }
    }
    C.n_units = (uint32_t)C.u_space.size();
    C.n_segs = (uint32_t)C.g_s.size();
    C.n_clines = (uint32_t)C.l_fs.size();
    C.n_spaces = n_spaces;

    // per-space coordinate sets
    std::vector<std::vector<uint32_t>> xs(n_spaces);
    for (uint32_t u = 0; u < C.n_units; u++) {
        auto& v = xs[C.u_space[u]];
        for (uint32_t i = C.u_line0[u]; i < C.u_line0[u + 1]; i++) { v.push_back(C.l_fs[i]); v.push_back(C.l_fe[i]); }
    }
    C.rank.assign(n_spaces, RankSeqHost{});
    C.cov.assign(n_spaces, CovSpaceHost{});
    uint32_t slot = 0;
    uint64_t word = 0;
    for (uint32_t s = 0; s < n_spaces; s++) {
        auto& v = xs[s];
        std::sort(v.begin(), v.end());
        v.erase(std::unique(v.begin(), v.end()), v.end());
        RankSeqHost& r = C.rank[s];
        r.x0 = (uint32_t)C.X.size();
        r.cnt = (uint32_t)v.size();
        r.slot0 = slot;
        slot += r.cnt + 1;
        r.canon = 1;
        r.dir_base = (uint32_t)C.dir.size();
        CovSpaceHost& cv = C.cov[s];
        cv.word0 = word;
        if (v.empty()) { r.nb = 0; r.shift = 0; C.dir.push_back(0); cv.lo = 0; cv.nbits = 0; continue; }
        const uint32_t maxv = v.back();
        uint32_t shift = 4;
        while (((uint64_t)maxv >> shift) + 1 > std::max<uint64_t>(64, 2ull * v.size())) shift++;
        r.shift = shift;
        r.nb = (maxv >> shift) + 1;
        uint32_t j = 0;
        for (uint32_t bq = 0; bq <= r.nb; bq++) {  // nb+1 entries: dir[b] = #(X < b<<shift)
            const uint64_t lo = (uint64_t)bq << shift;
            while (j < v.size() && (uint64_t)v[j] < lo) j++;
            C.dir.push_back(j);
        }
        C.X.insert(C.X.end(), v.begin(), v.end());
        cv.lo = v.front();
        uint32_t hi = v.back();
        if (s < n_seqs && sp_rmin[s] != 0xFFFFFFFFu) {   // the bitmap also spans the regular blocks' tree intervals (coverage routing)
            cv.lo = std::min(cv.lo, sp_rmin[s]);
            hi = std::max(hi, sp_rmax[s]);
        }
        cv.nbits = hi - cv.lo;
        word += (uint64_t)cv.nbits / 64 + 1;
    }
    C.n_slots = slot;
    C.n_words = word;

    auto slot_of = [&](uint32_t space, uint32_t v) {
        const auto& x = xs[space];
        return C.rank[space].slot0 + (uint32_t)(std::lower_bound(x.begin(), x.end(), v) - x.begin());
    };
    C.l_sfs.resize(C.l_fs.size());
    C.l_sfe.resize(C.l_fs.size());
    C.g_ss.resize(C.g_s.size());
    C.g_se.resize(C.g_s.size());
    for (uint32_t u = 0; u < C.n_units; u++) {
        const uint32_t sp = C.u_space[u];
        for (uint32_t i = C.u_line0[u]; i < C.u_line0[u + 1]; i++) { C.l_sfs[i] = slot_of(sp, C.l_fs[i]); C.l_sfe[i] = slot_of(sp, C.l_fe[i]); }
        for (uint32_t i = C.u_seg0[u]; i < C.u_seg0[u + 1]; i++) { C.g_ss[i] = slot_of(sp, C.g_s[i]); C.g_se[i] = slot_of(sp, C.g_e[i]); }
    }

    C.id_unit0.assign((size_t)C.n_ids + 1, 0);
    for (uint32_t i = 0; i < C.n_ids; i++) C.id_unit0[i + 1] = C.id_unit0[i] + (uint32_t)id_units[i].size();
    C.id_units.reserve(C.n_units);
    for (uint32_t i = 0; i < C.n_ids; i++) C.id_units.insert(C.id_units.end(), id_units[i].begin(), id_units[i].end());

    {   // gaps by (space, a) + the per-slot entry point
        std::vector<uint32_t> ord_g(gap_items.size());
        std::iota(ord_g.begin(), ord_g.end(), 0u);
        std::sort(ord_g.begin(), ord_g.end(), [&](uint32_t x, uint32_t y) {
            const JoinItem &p = gap_items[x], &q = gap_items[y];
            if (p.seq != q.seq) return p.seq < q.seq;
            if (p.start != q.start) return p.start < q.start;
            return p.end < q.end;
        });
        C.gap_bits.assign(((size_t)C.n_slots + 31) / 32 + 3, 0);   // spare words: depth_window_kernel realigns the bits two words at a time
        C.gs_first.assign(std::max<uint32_t>(C.n_slots, 1), 0);
        C.gsp0.assign((size_t)n_spaces + 1, 0);
        size_t k = 0;
        for (uint32_t s = 0; s < n_spaces; s++) {
            const auto& x = xs[s];
            const RankSeqHost& r = C.rank[s];
            size_t k1 = k;
            while (k1 < ord_g.size() && gap_items[ord_g[k1]].seq == s) k1++;
            const uint32_t g0 = (uint32_t)C.gap_a.size();
            for (size_t j = k; j < k1; j++) {
                const JoinItem& g = gap_items[ord_g[j]];
                C.gap_a.push_back(g.start); C.gap_b.push_back(g.end); C.gap_seg.push_back(g.payload);
            }
            C.gap_a.push_back(0xFFFFFFFFu); C.gap_b.push_back(0xFFFFFFFFu); C.gap_seg.push_back(0);
            C.gsp0[s] = g0;
            C.gsp0[s + 1] = (uint32_t)C.gap_a.size();   // past the sentinel: the next space starts here
            for (uint32_t j = g0; j + 1 < C.gap_a.size(); j++) {   // a is a coordinate of the space: mark its slot
                if (C.gap_b[j] - C.gap_a[j] >= GAP_SHORT) continue;   // only a region at least that long can reach across it: those always look
                const uint32_t slot = r.slot0 + (uint32_t)(std::lower_bound(x.begin(), x.end(), C.gap_a[j]) - x.begin());
                C.gap_bits[slot >> 5] |= 1u << (slot & 31);
            }
            {   // entry point per slot: the first gap whose a is >= the slot's coordinate (the space's sentinel when there is none)
                uint32_t j = g0;
                for (uint32_t i = 0; i < r.cnt; i++) {
                    while (C.gap_a[j] < x[i]) j++;
                    C.gs_first[r.slot0 + i] = j;
                }
                C.gs_first[r.slot0 + r.cnt] = (uint32_t)C.gap_a.size() - 1;
            }
            k = k1;
        }
    }
    {   // irregular tree intervals by (sequence, start) with the running maximum of their ends
        std::vector<uint32_t> ord_i(irr_items.size());
        std::iota(ord_i.begin(), ord_i.end(), 0u);
        std::stable_sort(ord_i.begin(), ord_i.end(), [&](uint32_t a, uint32_t b) {
            const JoinItem &p = irr_items[a], &q = irr_items[b];
            if (p.seq != q.seq) return p.seq < q.seq;
            return p.start < q.start;
        });
        const size_t m = ord_i.size();
        C.irr_seq0.assign((size_t)n_seqs + 1, 0);
        C.irr_s.resize(m); C.irr_e.resize(m); C.irr_pm.resize(m); C.irr_space.resize(m);
        C.irr_grp.resize(m); C.irr_pos.resize(m); C.irr_base.resize(m); C.irr_blk.resize(m);
        for (size_t k = 0; k < m; k++) {
            const JoinItem& it = irr_items[ord_i[k]];
            C.irr_seq0[it.seq + 1]++;
            C.irr_s[k] = it.start; C.irr_e[k] = it.end; C.irr_space[k] = it.payload;
            C.irr_pm[k] = (k && irr_items[ord_i[k - 1]].seq == it.seq) ? std::max(C.irr_pm[k - 1], it.end) : it.end;
            C.irr_grp[k] = irr_item_grp[ord_i[k]];
            C.irr_pos[k] = irr_item_pos[ord_i[k]];
            C.irr_base[k] = irr_item_base[ord_i[k]];
            C.irr_blk[k] = irr_item_blk[ord_i[k]];
        }
        for (uint32_t s = 0; s < n_seqs; s++) C.irr_seq0[s + 1] += C.irr_seq0[s];
    }
}

}  // namespace gffx
