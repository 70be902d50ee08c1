#include "host_index.h"
#include <functional>
#include <algorithm>
#include <thread>
#include <cstdlib>
#include <fcntl.h>
#include <sys/uio.h>
#include <unistd.h>

namespace gffx {

// ------------------------------------------------------------------------------------------
// StringPool
// ------------------------------------------------------------------------------------------
uint32_t StringPool::intern(std::string_view s) {
    auto it = map.find(s);
    if (it != map.end()) return it->second;
    keys.push_back(std::make_unique<std::string>(s));
    uint32_t id = (uint32_t)(keys.size() - 1);
    map.emplace(std::string_view(*keys.back()), id);
    return id;
}
int64_t StringPool::find(std::string_view s) const {
    auto it = map.find(s);
    return it == map.end() ? -1 : (int64_t)it->second;
}

// ------------------------------------------------------------------------------------------
// On-disk files
// ------------------------------------------------------------------------------------------
void load_sqs_file(const std::string& path, std::vector<std::string>* names) {
    Mapped m;
    if (!m.try_open(path)) throw Err(E_IO, "Failed to open SQS file: \"" + path + "\"");
    names->clear();
    const uint8_t* p = m.data();
    const uint8_t* end = p + m.size();
    while (p < end) {  // BufRead::lines(): '\n' terminated, one trailing '\r' stripped
        const uint8_t* nl = find_byte(p, end, '\n');
        const uint8_t* stop = nl ? nl : end;
        const uint8_t* last = stop;
        if (nl && last > p && last[-1] == '\r') last--;
        names->emplace_back((const char*)p, (size_t)(last - p));
        p = stop + 1;
    }
}

static std::vector<uint64_t> read_rix(const std::string& path) {
    Mapped m;
    if (!m.try_open(path)) throw Err(E_IO, "open " + path);
    // a JSON array of unsigned integers, compact or pretty-printed (src/utils/tree_io.rs:57-63)
    std::vector<uint64_t> v;
    const uint8_t* p = m.data();
    const uint8_t* end = p + m.size();
    auto ws = [&] { while (p < end && (*p == ' ' || *p == '\t' || *p == '\n' || *p == '\r')) p++; };
    ws();
    if (p >= end || *p != '[') throw Err(E_FORMAT, "parse json " + path);
    p++;
    ws();
    if (p < end && *p == ']') return v;
    for (;;) {
        ws();
        if (p >= end || *p < '0' || *p > '9') throw Err(E_FORMAT, "parse json " + path);
        uint64_t x = 0;
        while (p < end && *p >= '0' && *p <= '9') x = x * 10 + (uint64_t)(*p++ - '0');
        v.push_back(x);
        ws();
        if (p < end && *p == ',') { p++; continue; }
        if (p < end && *p == ']') break;
        throw Err(E_FORMAT, "parse json " + path);
    }
    return v;
}

// Flattens one bincode2-encoded IntervalTree<u32> (grammar in SURVEY.md App. B) without building
// nodes: every OptNode still owed is counted in `owed`; a present node owes two more.
static void flatten_tree(const uint8_t* p, size_t n, uint32_t seq, size_t tree_no, std::vector<RootIv>* out) {
    size_t pos = 0;
    uint64_t owed = 1;
    auto fail = [&]() -> void { throw Err(E_FORMAT, "bincode2 deserialize tree #" + std::to_string(tree_no)); };
    while (owed > 0) {
        if (pos >= n) fail();
        uint8_t tag = p[pos++];
        owed--;
        if (tag == 0) continue;
        if (tag != 1) fail();
        if (n - pos < 12) fail();
        pos += 4;  // center: only needed by the pointer tree, not by the flat layout
        uint64_t cnt;
        memcpy(&cnt, p + pos, 8);
        pos += 8;
        if (cnt > (n - pos) / 12) fail();
        for (uint64_t i = 0; i < cnt; i++) {
            uint32_t t[3];
            memcpy(t, p + pos, 12);
            pos += 12;
            out->push_back(RootIv{seq, t[0], t[1], t[2]});
        }
        owed += 2;
    }
}

void load_trees_flat(const std::string& rit_path, const std::string& rix_path, std::vector<RootIv>* out, uint32_t* n_trees) {
    Mapped rit;
    if (!rit.try_open(rit_path)) throw Err(E_IO, "open " + rit_path);
    std::vector<uint64_t> offs = read_rix(rix_path);
    *n_trees = (uint32_t)offs.size();
    if (offs.empty()) return;
    for (size_t i = 1; i < offs.size(); i++)
        if (offs[i - 1] > offs[i]) throw Err(E_FORMAT, "offsets not sorted ascending");
    if (offs.back() > rit.size()) throw Err(E_FORMAT, "last offset out of file size");
    for (size_t i = 0; i < offs.size(); i++) {
        uint64_t s = offs[i], e = i + 1 < offs.size() ? offs[i + 1] : rit.size();
        flatten_tree(rit.data() + s, (size_t)(e - s), (uint32_t)i, i, out);
    }
}

void load_gof_file(const std::string& path, std::vector<uint32_t>* fid, std::vector<uint64_t>* s, std::vector<uint64_t>* e) {
    Mapped m;
    if (!m.try_open(path)) throw Err(E_IO, "Failed to mmap " + path);
    if (m.size() % 24 != 0)
        throw Err(E_FORMAT, "Corrupted GOF (" + path + "): length " + std::to_string(m.size()) + " not multiple of 24");
    size_t n = m.size() / 24;
    fid->resize(n);
    s->resize(n);
    e->resize(n);
    for (size_t i = 0; i < n; i++) {
        const uint8_t* r = m.data() + 24 * i;
        memcpy(&(*fid)[i], r, 4);  // bytes 4..8: seq number (0 in files of older writers) — unused here
        memcpy(&(*s)[i], r + 8, 8);
        memcpy(&(*e)[i], r + 16, 8);
    }
}

// ------------------------------------------------------------------------------------------
// HostIndex
// ------------------------------------------------------------------------------------------
void HostIndex::open(const std::string& path, int threads) {
    gff_path = path;
    PhaseTimer t_all("host index (total)");
    PhaseTimer* t_files = new PhaseTimer("  index files");
    load_sqs_file(path + ".sqs", &seq_names);
    uint32_t n_trees = 0;
    load_trees_flat(path + ".rit", path + ".rix", &roots, &n_trees);
    std::vector<uint32_t> gfid;
    std::vector<uint64_t> gs, ge;
    load_gof_file(path + ".gof", &gfid, &gs, &ge);
    if (gff_map.try_open(path)) {
        gff = gff_map.data();
        gff_len = gff_map.size();
    }
    delete t_files;
    finish(gfid, gs, ge, threads);
}

void HostIndex::create(uint32_t n_seqs, const char* const* names, uint64_t n_roots, const uint32_t* rseq, const uint32_t* rs,
                       const uint32_t* re, const uint32_t* rfid, uint64_t n_gof, const uint32_t* gfid, const uint64_t* gs,
                       const uint64_t* ge, const uint8_t* text, uint64_t text_len, int threads) {
    seq_names.assign(names, names + n_seqs);
    roots.resize(n_roots);
    for (uint64_t i = 0; i < n_roots; i++) roots[i] = RootIv{rseq[i], rs[i], re[i], rfid[i]};
    // the caller's text is COPIED (include/gffx_b200.h: inputs are borrowed only for the duration of the call): the line table,
    // the ID strings of depth / coverage rows and the record emission all read it long after gffx_index_create has returned
    if (text && text_len) {
        gff_copy.assign(text, text + text_len);
        gff = gff_copy.data();
        gff_len = text_len;
    } else {
        gff = nullptr;
        gff_len = 0;
    }
    finish(std::vector<uint32_t>(gfid, gfid + n_gof), std::vector<uint64_t>(gs, gs + n_gof), std::vector<uint64_t>(ge, ge + n_gof),
           threads);
}

void HostIndex::finish(const std::vector<uint32_t>& gfid, const std::vector<uint64_t>& gs, const std::vector<uint64_t>& ge,
                       int threads) {
    seq_num.clear();
    for (size_t i = 0; i < seq_names.size(); i++) seq_num[seq_names[i]] = (uint32_t)i;
    // trees beyond .sqs would make the reference index out of bounds (intersect.rs:125); refuse them here
    for (const RootIv& r : roots)
        if (r.seq >= seq_names.size()) throw Err(E_FORMAT, "tree sequence number beyond the .sqs list");

    // reachable blocks: the last .gof record of each fid (GofMap::index, gof.rs:32-37)
    std::unordered_map<uint32_t, size_t> last;
    last.reserve(gfid.size() * 2);
    for (size_t i = 0; i < gfid.size(); i++) last[gfid[i]] = i;
    std::vector<size_t> keep;
    keep.reserve(last.size());
    for (size_t i = 0; i < gfid.size(); i++)
        if (last[gfid[i]] == i) keep.push_back(i);
    std::stable_sort(keep.begin(), keep.end(), [&](size_t a, size_t b) { return gs[a] < gs[b]; });
    blocks.resize(keep.size());
    std::unordered_map<uint32_t, uint32_t> fid_blk;
    fid_blk.reserve(keep.size() * 2);
    for (size_t k = 0; k < keep.size(); k++) {
        blocks[k] = Block{gfid[keep[k]], gs[keep[k]], ge[keep[k]]};
        fid_blk[blocks[k].fid] = (uint32_t)k;
    }
    root_blk.resize(roots.size());
    for (size_t i = 0; i < roots.size(); i++) {
        auto it = fid_blk.find(roots[i].fid);
        root_blk[i] = it == fid_blk.end() ? NONE32 : it->second;
    }
    build_lines(threads);
    build_vseqs();
    if (cache_pending) {
        cache_pending = false;
        const std::string target = line_cache_target(*this);   // decided here: the writer thread must not look at the environment later
        if (getenv("GFFX_CACHE_SYNC")) line_cache_store(*this, target);
        else if (!target.empty()) cache_writer = std::thread([this, target] { line_cache_store(*this, target); });
    }
}

// see VSeqMap (host_index.h)
void HostIndex::build_vseqs() {
    const size_t n = seq_names.size();
    vmap = VSeqMap();
    vmap.seq.assign(n, VSeqMap::Seq{0, 0, 0xFFFFFFFFu});
    if (n == 0) return;
    std::vector<uint64_t> ext(n, 0), root_ext(n, 0);
    for (const RootIv& r : roots) {
        root_ext[r.seq] = std::max<uint64_t>(root_ext[r.seq], std::max<uint64_t>(r.end, (uint64_t)r.start + 1));
        ext[r.seq] = std::max(ext[r.seq], root_ext[r.seq]);
    }
    const LineTable& L = lines;
    {
        // two passes over the line table, on the host threads when the per-thread copies of the extents are small
        const size_t T = (L.size() >= (1u << 16) && n <= (1u << 16)) ? (size_t)std::max(1, std::min(n_threads, 32)) : 1;
        std::vector<std::vector<uint64_t>> ext_t(T, std::vector<uint64_t>(T > 1 ? n : 0, 0));
        std::vector<uint32_t> blk_max(blocks.size(), 0);
        auto work = [&](size_t t) {
            uint64_t* e = T > 1 ? ext_t[t].data() : ext.data();
            for (size_t i = L.size() * t / T; i < L.size() * (t + 1) / T; i++)
                if (L.flags[i] & LN_I_OK) e[L.iseq[i]] = std::max<uint64_t>(e[L.iseq[i]], std::max(L.S[i], L.E[i]));
            // ID lines are compared with the regions routed to their block, whatever their own column 1 says (depth.rs:194-206)
            if (L.size())
                for (size_t b = blocks.size() * t / T; b < blocks.size() * (t + 1) / T; b++)
                    for (uint64_t i = blk_line0[b]; i < blk_line0[b + 1]; i++)
                        if (L.flags[i] & LN_D_OK) blk_max[b] = std::max(blk_max[b], L.fe[i]);
        };
        if (T > 1) {
            parallel_jobs(T, (int)T, work);
            for (size_t t = 0; t < T; t++)
                for (size_t c = 0; c < n; c++) ext[c] = std::max(ext[c], ext_t[t][c]);
        } else {
            work(0);
        }
        if (L.size())
            for (size_t i = 0; i < roots.size(); i++)
                if (root_blk[i] != NONE32) ext[roots[i].seq] = std::max<uint64_t>(ext[roots[i].seq], blk_max[root_blk[i]]);
    }
    std::vector<uint8_t> canon(n);
    for (size_t c = 0; c < n; c++) canon[c] = seq_num.at(seq_names[c]) == c ? 1 : 0;
    // capacity of a virtual sequence: at least the longest real sequence by its roots (a human chromosome stays alone or shares
    // with a smaller one), doubled until the whole index fits on VSEQ_MAX virtual sequences
    uint64_t cap = 1ull << 27;
    for (size_t c = 0; c < n; c++) cap = std::max(cap, root_ext[c] + 2);
    if (const char* e = getenv("GFFX_VSEQ_CAP")) cap = std::max<uint64_t>(16, strtoull(e, nullptr, 10));   // test hook
    const uint64_t CAP_MAX = 0xFFFFFFFFull;
    cap = std::min(cap, CAP_MAX);
    std::vector<uint32_t> vs(n);
    std::vector<uint64_t> base(n);
    for (;;) {
        // two packing streams (canonical names / shadowed duplicates), each greedy in sequence order
        uint32_t nv = 0;
        int64_t open_v[2] = {-1, -1};
        uint64_t fill[2] = {0, 0};
        for (size_t c = 0; c < n; c++) {
            const uint64_t len = ext[c] + 2;
            const int k = canon[c] ? 0 : 1;
            if (len > cap) { vs[c] = nv++; base[c] = 0; continue; }   // alone, on its own coordinates
            if (open_v[k] < 0 || fill[k] + len > cap) { open_v[k] = nv++; fill[k] = 0; }
            vs[c] = (uint32_t)open_v[k];
            base[c] = fill[k];
            fill[k] += len;
        }
        if (nv <= VSEQ_MAX || cap >= CAP_MAX) {
            if (nv > VSEQ_MAX)
                throw Err(E_FORMAT, "index needs " + std::to_string(nv) + " sequences with coordinates near 2^32: more than the engine's " +
                                        std::to_string(VSEQ_MAX) + " virtual sequences");
            vmap.v_extent.assign(nv, 0);
            vmap.v_canon.assign(nv, 1);
            vmap.v_members.assign(nv, 0);
            break;
        }
        cap = std::min(cap * 2, CAP_MAX);
    }
    vmap.capacity = cap;
    for (size_t c = 0; c < n; c++) vmap.v_members[vs[c]]++;
    for (size_t c = 0; c < n; c++) {
        const uint32_t v = vs[c];
        vmap.v_canon[v] = canon[c];
        if (vmap.v_members[v] == 1) {
            vmap.seq[c] = VSeqMap::Seq{v, 0, 0xFFFFFFFFu};
            vmap.v_extent[v] = (uint32_t)std::min<uint64_t>(root_ext[c], 0xFFFFFFFFull);   // the window grid follows the roots, as regions do
        } else {
            vmap.seq[c] = VSeqMap::Seq{v, (uint32_t)base[c], (uint32_t)(ext[c] + 1)};
            vmap.v_extent[v] = std::max<uint32_t>(vmap.v_extent[v], (uint32_t)(base[c] + ext[c] + 1));
        }
    }
}

namespace {

struct LocalDict {
    std::vector<std::string> items;
    std::unordered_map<std::string, uint32_t> map;
    std::string last;
    uint32_t last_id = NONE32;
    uint32_t get(const uint8_t* p, size_t n) {
        if (last_id != NONE32 && last.size() == n && memcmp(last.data(), p, n) == 0) return last_id;
        last.assign((const char*)p, n);
        auto it = map.find(last);
        if (it == map.end()) {
            last_id = (uint32_t)items.size();
            items.push_back(last);
            map.emplace(last, last_id);
        } else {
            last_id = it->second;
        }
        return last_id;
    }
};

struct Part {  // lines of a contiguous range of blocks
    LineTable t;
    std::vector<uint64_t> id_off;  // absolute offset of the ID string (d_ok lines)
    std::vector<uint32_t> id_len;
    std::vector<uint64_t> id_hash; // FNV-1a of the ID string (d_ok lines): the dictionary is built from these, in parallel
    LocalDict types, chroms;
    std::vector<uint64_t> blk_nlines;
    std::vector<uint8_t> blk_utf8;
};

// fast_id of src/commands/depth.rs:104-118: first "ID=" anywhere in column 9, value up to ; space tab
inline bool find_id(const uint8_t* a, size_t n, size_t* vs, size_t* vn) {
    for (size_t i = 0; i + 2 < n; i++) {
        if (a[i] == 'I' && a[i + 1] == 'D' && a[i + 2] == '=') {
            size_t j = i + 3;
            while (j < n && a[j] != ';' && a[j] != ' ' && a[j] != '\t') j++;
            *vs = i + 3;
            *vn = j - (i + 3);
            return true;
        }
    }
    return false;
}

}  // namespace

void HostIndex::build_lines(int threads) {
    const size_t nb = blocks.size();
    blk_utf8.assign(nb, 0);
    blk_line0.assign(nb + 1, 0);
    lines = LineTable();
    if (threads < 1) threads = 1;
    n_threads = threads;
    if (!gff || gff_len == 0 || nb == 0) return;
    if (line_cache_load(*this, threads)) { lines_source = 2; return; }
    // GFFX_LINE_SCAN=host|device picks the scanner; by default the device scans texts of 8 MB and more when the index is being
    // opened on one (below that the host threads are done before the text has crossed PCIe)
    const char* sel = getenv("GFFX_LINE_SCAN");
    const bool want_dev = scan_device >= 0 && (sel ? strcmp(sel, "device") == 0 : gff_len >= (8u << 20));
    bool done = false;
    if (want_dev) {
        RawLines R;
        if (device_scan_lines(*this, scan_device, threads, &R, &dev_text, &dev_text_cap)) {
            dev_text_device = scan_device;
            dev_text_release = device_text_release;
            lines_from_raw(R, threads);
            lines_source = 1;
            done = true;
        }
    }
    if (!done) scan_lines(threads);
    cache_pending = true;   // written by finish() once everything the file holds is final
}

// The device scanner's columns -> the line table: blocks with non-ASCII bytes re-scanned by the host scanner (it knows UTF-8;
// their lines keep their places), then the three dictionaries, numbered by first appearance as scan_lines numbers them.
void HostIndex::lines_from_raw(RawLines& R, int threads) {
    PhaseTimer t_all("    dictionaries (host)");
    PhaseTimer t_na("      non-ASCII blocks");
    const size_t nb = blocks.size();
    LineTable& L = lines;
    L = std::move(R.t);
    const size_t total = L.size();
    for (size_t b = 0; b < nb; b++) blk_line0[b + 1] = blk_line0[b] + R.blk_nlines[b];
    if (blk_line0[nb] != total) throw Err(E_STATE, "device line scan: line counts do not add up (internal)");
    for (size_t b = 0; b < nb; b++) {
        if (!block_scannable(b)) continue;
        blk_utf8[b] = 1;
        if (!R.blk_nonascii[b]) continue;
        Part P;
        scan_blocks(b, b + 1, &P);
        const size_t o = (size_t)blk_line0[b], m = P.t.size();
        if (m != R.blk_nlines[b]) throw Err(E_STATE, "device line scan: a block's line count differs from the host scanner's (internal)");
        blk_utf8[b] = P.blk_utf8[0];
        for (size_t i = 0; i < m; i++) {
            if (L.off[o + i] != P.t.off[i] || L.len[o + i] != P.t.len[i]) throw Err(E_STATE, "device line scan: a line's extent differs from the host scanner's (internal)");
            L.flags[o + i] = P.t.flags[i];
            L.S[o + i] = P.t.S[i]; L.E[o + i] = P.t.E[i]; L.iseq[o + i] = P.t.iseq[i]; L.fs[o + i] = P.t.fs[i]; L.fe[o + i] = P.t.fe[i];
            if (P.t.type_id[i] == NONE32) R.type_len[o + i] = NONE32;   // the field is where the device found it; the host decides whether it is text
            if (P.t.flags[i] & LN_D_OK) {
                R.id_off[o + i] = (uint32_t)(P.id_off[i] - P.t.off[i]);
                R.id_len[o + i] = P.id_len[i];
                R.id_hash[o + i] = P.id_hash[i];
            } else {
                R.chrom_len[o + i] = NONE32;
            }
        }
    }
    t_na.stop();
    if (R.ids_ready) {
        // the device built all three dictionaries (linescan.cu: every block ASCII, no two strings under one hash): the lines' id /
        // type_id / chrom columns are filled, the entries arrive in order of first appearance as (offset, length) into the text
        PhaseTimer t_dev("      dictionaries (from the device's entries)");
        if (L.id.size() != total || L.type_id.size() != total || L.chrom.size() != total)
            throw Err(E_STATE, "device line scan: a dictionary column of the wrong length (internal)");
        for (size_t k = 0; k < R.types_off.size(); k++)
            if (types.intern(std::string_view((const char*)gff + R.types_off[k], R.types_len[k])) != k) throw Err(E_STATE, "device line scan: type dictionary out of order (internal)");
        for (size_t k = 0; k < R.chroms_off.size(); k++)
            if (chroms.intern(std::string_view((const char*)gff + R.chroms_off[k], R.chroms_len[k])) != k) throw Err(E_STATE, "device line scan: chromosome dictionary out of order (internal)");
        ids = IdTable();
        ids.base = gff;
        ids.off = std::move(R.ids_off);
        ids.len = std::move(R.ids_len);
        return;
    }
    PhaseTimer t_tc("      type / chromosome dictionaries");
    // type / chromosome dictionaries: chunks of lines side by side, each with its own dictionary in order of appearance; the
    // chunks' dictionaries are merged in chunk order, which numbers every string by its first appearance in the file
    const size_t nparts = std::max<size_t>(1, std::min<size_t>((size_t)std::max(1, threads), total / 65536 + 1));
    struct Chunk { LocalDict types, chroms; std::vector<uint32_t> ty, ch; };
    std::vector<Chunk> chunks(nparts);
    auto run_parallel = [&](size_t n_jobs, const std::function<void(size_t)>& f) { parallel_jobs(n_jobs, (int)std::max<size_t>(1, (size_t)threads), f); };   // util.h: the thread pool
    run_parallel(nparts, [&](size_t c) {
        Chunk& C = chunks[c];
        const size_t i0 = total * c / nparts, i1 = total * (c + 1) / nparts;
        C.ty.assign(i1 - i0, NONE32);
        C.ch.assign(i1 - i0, NONE32);
        for (size_t i = i0; i < i1; i++) {
            if (R.type_len[i] != NONE32) C.ty[i - i0] = C.types.get(gff + L.off[i] + R.type_off[i], R.type_len[i]);
            if ((L.flags[i] & LN_D_OK) && R.chrom_len[i] != NONE32) C.ch[i - i0] = C.chroms.get(gff + L.off[i], R.chrom_len[i]);
        }
    });
    std::vector<std::vector<uint32_t>> tmaps(nparts), cmaps(nparts);
    for (size_t c = 0; c < nparts; c++) {
        tmaps[c].resize(chunks[c].types.items.size());
        cmaps[c].resize(chunks[c].chroms.items.size());
        for (size_t i = 0; i < tmaps[c].size(); i++) tmaps[c][i] = types.intern(chunks[c].types.items[i]);
        for (size_t i = 0; i < cmaps[c].size(); i++) cmaps[c][i] = chroms.intern(chunks[c].chroms.items[i]);
    }
    L.type_id.resize(total);
    L.chrom.resize(total);
    std::vector<uint64_t> g_off(total);
    run_parallel(nparts, [&](size_t c) {
        const size_t i0 = total * c / nparts, i1 = total * (c + 1) / nparts;
        for (size_t i = i0; i < i1; i++) {
            const uint32_t ty = chunks[c].ty[i - i0], ch = chunks[c].ch[i - i0];
            L.type_id[i] = ty == NONE32 ? NONE32 : tmaps[c][ty];
            L.chrom[i] = ch == NONE32 ? NONE32 : cmaps[c][ch];
            const bool d = L.flags[i] & LN_D_OK;
            g_off[i] = d ? L.off[i] + R.id_off[i] : 0;
            if (!d) { R.id_len[i] = 0; R.id_hash[i] = 0; }
        }
    });
    t_tc.stop();
    PhaseTimer t_id("      ID dictionary");
    number_ids(R.id_hash, g_off, R.id_len, threads);
}

// ID dictionary: the lines are cut into `nshard` shards by hash; every shard finds, for each of its lines, the FIRST line
// with the same ID (its own open-addressing table over line numbers, lines visited in file order); an ID's number is then
// the rank of its first line among all first lines = order of first appearance, as a serial intern loop would number them.
// g_hash / g_off / g_len: FNV-1a hash, absolute offset and length of the ID value of every LN_D_OK line.
void HostIndex::number_ids(const std::vector<uint64_t>& g_hash, const std::vector<uint64_t>& g_off, const std::vector<uint32_t>& g_len, int threads) {
    LineTable& L = lines;
    const size_t total = L.size();
    const size_t T = (size_t)std::max(1, std::min(threads, 64));
    const size_t nshard = T, nchunk = total >= (1u << 16) ? T : 1;
    auto run_parallel = [&](size_t n_jobs, const std::function<void(size_t)>& f) { parallel_jobs(n_jobs, (int)std::max<size_t>(1, (size_t)T), f); };   // util.h: the thread pool
    auto chunk_lo = [&](size_t c) { return total * c / nchunk; };
    auto shard_of = [&](uint64_t h) { return (size_t)((h >> 32) % nshard); };
    // the shards' line lists, in line order: chunks of lines count, then place, their lines side by side (every shard used to
    // read the whole table twice)
    std::vector<std::vector<uint32_t>> at(nchunk, std::vector<uint32_t>(nshard, 0));
    run_parallel(nchunk, [&](size_t c) {
        for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++)
            if (L.flags[i] & LN_D_OK) at[c][shard_of(g_hash[i])]++;
    });
    std::vector<std::vector<uint32_t>> list(nshard);
    for (size_t sh = 0; sh < nshard; sh++) {
        uint32_t tot = 0;
        for (size_t c = 0; c < nchunk; c++) { const uint32_t k = at[c][sh]; at[c][sh] = tot; tot += k; }
        list[sh].resize(tot);
    }
    run_parallel(nchunk, [&](size_t c) {
        for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++)
            if (L.flags[i] & LN_D_OK) { const size_t sh = shard_of(g_hash[i]); list[sh][at[c][sh]++] = (uint32_t)i; }
    });
    std::vector<uint32_t> first(total, NONE32);
    run_parallel(nshard, [&](size_t sh) {
        const std::vector<uint32_t>& mine = list[sh];
        size_t cap = 16;
        while (cap < 2 * mine.size() + 2) cap <<= 1;
        std::vector<uint32_t> tab(cap, 0);   // line number + 1
        const size_t mask = cap - 1;
        for (const uint32_t i : mine) {
            const uint64_t h = g_hash[i];
            size_t k = (size_t)h & mask;
            for (;; k = (k + 1) & mask) {
                if (tab[k] == 0) { tab[k] = i + 1; first[i] = i; break; }
                const size_t j = tab[k] - 1;
                if (g_hash[j] == h && g_len[j] == g_len[i] && memcmp(gff + g_off[j], gff + g_off[i], g_len[i]) == 0) { first[i] = (uint32_t)j; break; }
            }
        }
    });
    // an ID's number = the rank of its first line among all first lines: counted per chunk, numbered side by side
    std::vector<size_t> n_first(nchunk + 1, 0);
    run_parallel(nchunk, [&](size_t c) {
        size_t k = 0;
        for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++) k += first[i] == (uint32_t)i;
        n_first[c + 1] = k;
    });
    for (size_t c = 0; c < nchunk; c++) n_first[c + 1] += n_first[c];
    ids = IdTable();
    ids.base = gff;
    ids.off.resize(n_first[nchunk]);
    ids.len.resize(n_first[nchunk]);
    L.id.assign(total, NONE32);
    run_parallel(nchunk, [&](size_t c) {
        size_t k = n_first[c];
        for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++)
            if (first[i] == (uint32_t)i) {
                L.id[i] = (uint32_t)k;
                ids.off[k] = g_off[i];
                ids.len[k] = g_len[i];
                k++;
            }
    });
    run_parallel(nchunk, [&](size_t c) {   // first[i] < i: its number is in place since the pass above
        for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++)
            if (first[i] != NONE32 && first[i] != (uint32_t)i) L.id[i] = L.id[first[i]];
    });
}

// the host scanner proper: the lines of blocks [b0, b1) appended to *part (a Part), every rule of the reference's per-block text
// loops (intersect.rs:284-321 + :441-489, depth.rs:131-157, coverage.rs:296-323)
void HostIndex::scan_blocks(size_t b0, size_t b1, void* part) const {
    Part& P = *(Part*)part;
    P.blk_nlines.assign(b1 - b0, 0);
    P.blk_utf8.assign(b1 - b0, 0);
    for (size_t b = b0; b < b1; b++) {
        if (!block_scannable(b)) continue;
        const uint64_t bs = blocks[b].start, be = std::min<uint64_t>(blocks[b].end, gff_len);
        const bool u8ok = utf8_ok(gff + bs, (size_t)(be - bs));
        P.blk_utf8[b - b0] = u8ok;
        const bool countable = block_countable(b) && blocks[b].end <= gff_len;
        const uint8_t* p = gff + bs;
        const uint8_t* end = gff + be;
        uint64_t nl_count = 0;
        while (p < end) {
            const uint8_t* nl = find_byte(p, end, '\n');
            const uint8_t* next = nl ? nl + 1 : end;
            const uint8_t* q = nl ? nl : end;  // line body [p,q)
            if (q > p && p[0] != '#') {
                const uint8_t* tab[8];
                int nt = 0;
                const uint8_t* c = p;
                while (nt < 8) {
                    const uint8_t* t = find_byte(c, q, '\t');
                    if (!t) break;
                    tab[nt++] = t;
                    c = t + 1;
                }
                uint8_t fl = 0;
                uint32_t S = 0, E = 0, iseq = NONE32, ty = NONE32, fs = 0, fe = 0, chrom = NONE32;
                bool numok = false;
                if (nt >= 3 && utf8_ok(tab[1] + 1, (size_t)(tab[2] - tab[1] - 1)))
                    ty = P.types.get(tab[1] + 1, (size_t)(tab[2] - tab[1] - 1));
                if (nt >= 5) {
                    numok = parse_u32_strict(tab[2] + 1, (size_t)(tab[3] - tab[2] - 1), &S) &&
                            parse_u32_strict(tab[3] + 1, (size_t)(tab[4] - tab[3] - 1), &E);
                    if (numok && utf8_ok(p, (size_t)(tab[0] - p))) {
                        auto it = seq_num.find(std::string((const char*)p, (size_t)(tab[0] - p)));
                        if (it != seq_num.end()) { iseq = it->second; fl |= LN_I_OK; }
                    }
                }
                size_t ivs = 0, ivn = 0;
                if (nt >= 8 && numok && E != 0 && u8ok && countable &&
                    find_id(tab[7] + 1, (size_t)(q - tab[7] - 1), &ivs, &ivn)) {
                    uint32_t a = std::min(S, E), z = std::max(S, E);
                    fs = a == 0 ? 0 : a - 1;
                    fe = z;
                    chrom = P.chroms.get(p, (size_t)(tab[0] - p));
                    fl |= LN_D_OK;
                }
                LineTable& t = P.t;
                t.off.push_back((uint64_t)(p - gff));
                t.len.push_back((uint32_t)(next - p));
                t.blk.push_back((uint32_t)b);
                t.flags.push_back(fl);
                t.S.push_back(S); t.E.push_back(E); t.iseq.push_back(iseq); t.type_id.push_back(ty);
                t.fs.push_back(fs); t.fe.push_back(fe); t.chrom.push_back(chrom);
                if (fl & LN_D_OK) {
                    const uint8_t* ids0 = tab[7] + 1 + ivs;
                    uint64_t h = 1469598103934665603ull;
                    for (size_t z = 0; z < ivn; z++) { h ^= ids0[z]; h *= 1099511628211ull; }
                    P.id_off.push_back((uint64_t)(ids0 - gff));
                    P.id_len.push_back((uint32_t)ivn);
                    P.id_hash.push_back(h);
                } else {
                    P.id_off.push_back(0);
                    P.id_len.push_back(0);
                    P.id_hash.push_back(0);
                }
                nl_count++;
            }
            p = next;
        }
        P.blk_nlines[b - b0] = nl_count;
    }
}

void HostIndex::scan_lines(int threads) {
    PhaseTimer t_lines("  GFF line table (host scan)");
    const size_t nb = blocks.size();
    size_t nparts = std::min<size_t>((size_t)threads, std::max<size_t>(1, nb / 64));
    std::vector<Part> parts(nparts);
    // split blocks so that each part scans about the same number of bytes
    std::vector<size_t> cut(nparts + 1, nb);
    {
        uint64_t total = 0;
        for (size_t b = 0; b < nb; b++) if (block_scannable(b)) total += std::min<uint64_t>(blocks[b].end, gff_len) - blocks[b].start;
        uint64_t acc = 0;
        size_t k = 0;
        cut[0] = 0;
        for (size_t b = 0; b < nb && k + 1 < nparts; b++) {
            if (acc >= (total / nparts) * (k + 1)) cut[++k] = b;
            if (block_scannable(b)) acc += std::min<uint64_t>(blocks[b].end, gff_len) - blocks[b].start;
        }
        for (size_t j = k + 1; j < nparts; j++) cut[j] = nb;
        cut[nparts] = nb;
    }
    auto work = [&](size_t pi) { scan_blocks(cut[pi], cut[pi + 1], &parts[pi]); };
    {
        PhaseTimer t_scan("    parallel scan");
        parallel_jobs(nparts, (int)nparts, work);
    }
    PhaseTimer t_merge("    merge + ID dictionary");
    // concatenate (every part copies its own slice, in parallel), remapping the per-part type / chromosome dictionaries
    std::vector<size_t> base(nparts + 1, 0);
    for (size_t pi = 0; pi < nparts; pi++) base[pi + 1] = base[pi] + parts[pi].t.size();
    const size_t total = base[nparts];
    LineTable& L = lines;
    L.off.resize(total); L.len.resize(total); L.blk.resize(total); L.flags.resize(total);
    L.S.resize(total); L.E.resize(total); L.iseq.resize(total); L.type_id.resize(total);
    L.fs.resize(total); L.fe.resize(total); L.id.assign(total, NONE32); L.chrom.resize(total);
    std::vector<uint64_t> g_hash(total), g_off(total);
    std::vector<uint32_t> g_len(total);
    std::vector<std::vector<uint32_t>> tmaps(nparts), cmaps(nparts);
    for (size_t pi = 0; pi < nparts; pi++) {   // dictionary order = file order: parts in turn
        Part& P = parts[pi];
        tmaps[pi].resize(P.types.items.size());
        cmaps[pi].resize(P.chroms.items.size());
        for (size_t i = 0; i < tmaps[pi].size(); i++) tmaps[pi][i] = types.intern(P.types.items[i]);
        for (size_t i = 0; i < cmaps[pi].size(); i++) cmaps[pi][i] = chroms.intern(P.chroms.items[i]);
        for (size_t b = cut[pi]; b < cut[pi + 1]; b++) {
            blk_utf8[b] = P.blk_utf8[b - cut[pi]];
            blk_line0[b + 1] = P.blk_nlines[b - cut[pi]];
        }
    }
    auto copy_part = [&](size_t pi) {
        Part& P = parts[pi];
        const LineTable& t = P.t;
        const size_t o = base[pi], m = t.size();
        if (m) {
            memcpy(&L.off[o], t.off.data(), m * 8); memcpy(&L.len[o], t.len.data(), m * 4); memcpy(&L.blk[o], t.blk.data(), m * 4);
            memcpy(&L.flags[o], t.flags.data(), m); memcpy(&L.S[o], t.S.data(), m * 4); memcpy(&L.E[o], t.E.data(), m * 4);
            memcpy(&L.iseq[o], t.iseq.data(), m * 4); memcpy(&L.fs[o], t.fs.data(), m * 4); memcpy(&L.fe[o], t.fe.data(), m * 4);
            memcpy(&g_hash[o], P.id_hash.data(), m * 8); memcpy(&g_off[o], P.id_off.data(), m * 8); memcpy(&g_len[o], P.id_len.data(), m * 4);
        }
        for (size_t i = 0; i < m; i++) {
            L.type_id[o + i] = t.type_id[i] == NONE32 ? NONE32 : tmaps[pi][t.type_id[i]];
            L.chrom[o + i] = t.chrom[i] == NONE32 ? NONE32 : cmaps[pi][t.chrom[i]];
        }
        P = Part();
    };
    auto run_parallel = [&](size_t n_jobs, const std::function<void(size_t)>& f) { parallel_jobs(n_jobs, (int)std::max<size_t>(1, (size_t)threads), f); };   // util.h: the thread pool
    if (total >= (1ull << 32) - 1) throw Err(E_FORMAT, "more than 2^32-2 GFF lines inside blocks");
    run_parallel(nparts, copy_part);
    number_ids(g_hash, g_off, g_len, threads);
    for (size_t b = 0; b < nb; b++) blk_line0[b + 1] += blk_line0[b];
}

// ------------------------------------------------------------------------------------------
// BED parsing
// ------------------------------------------------------------------------------------------
// one line [line, q) of the intersect dialect (q = its '\n' or the end of the file): false = skipped, throws where the
// reference aborts (intersect.rs:215, :223-224)
bool bed_intersect_line(const uint8_t* line, const uint8_t* q, const std::unordered_map<std::string, uint32_t>& seq_num, const std::string& path,
                        std::string& scratch, uint32_t* chr, uint32_t* s, uint32_t* e) {
    if (q == line || line[0] == '#') return false;
    if (!utf8_ok(line, (size_t)(q - line))) throw Err(E_FORMAT, "invalid utf-8 sequence in BED file " + path);
    const uint8_t* f[3];
    size_t fl[3];
    int nf = 0;
    const uint8_t* c = line;
    while (nf < 3) {  // split_ascii_whitespace
        while (c < q && ascii_ws(*c)) c++;
        if (c >= q) break;
        const uint8_t* st = c;
        while (c < q && !ascii_ws(*c)) c++;
        f[nf] = st;
        fl[nf] = (size_t)(c - st);
        nf++;
    }
    if (nf < 3) return false;
    if (scratch.size() != fl[0] || memcmp(scratch.data(), f[0], fl[0]) != 0 || *chr == NONE32 - 1) {
        scratch.assign((const char*)f[0], fl[0]);
        auto it = seq_num.find(scratch);
        *chr = it == seq_num.end() ? NONE32 : it->second;
    }
    if (*chr == NONE32) return false;  // unknown chromosome: skipped BEFORE the numbers are looked at
    // lexical_core::parse::<u32> (intersect.rs:223-224): decimal digits after at most one '+' (lexical-parse-integer 1.0, parse_sign)
    if (!parse_u32_rust(std::string_view((const char*)f[1], fl[1]), s))
        throw Err(E_FORMAT, "BED " + path + ": cannot parse start '" + std::string((const char*)f[1], fl[1]) + "' as u32");
    if (!parse_u32_rust(std::string_view((const char*)f[2], fl[2]), e))
        throw Err(E_FORMAT, "BED " + path + ": cannot parse end '" + std::string((const char*)f[2], fl[2]) + "' as u32");
    return true;
}

void parse_bed_intersect(const std::string& path, const std::unordered_map<std::string, uint32_t>& seq_num, Regions* out) {
    Mapped m;
    m.open(path);
    const uint8_t* p = m.data();
    const uint8_t* end = p + m.size();
    std::string last_name;            // the previous line's chromosome name and number: BED files come in runs
    uint32_t chr = NONE32 - 1;        // "nothing looked up yet"
    while (p < end) {
        const uint8_t* nl = find_byte(p, end, '\n');
        const uint8_t* q = nl ? nl : end;
        const uint8_t* line = p;
        p = q + 1;
        uint32_t s, e;
        if (bed_intersect_line(line, q, seq_num, path, last_name, &chr, &s, &e)) {
            out->chr.push_back(chr);
            out->start.push_back(s);
            out->end.push_back(e);
        }
    }
}

bool bed_counts_line(const uint8_t* line, const uint8_t* q, const std::unordered_map<std::string, uint32_t>& seq_num,
                            std::string& scratch, uint32_t* chr, uint32_t* s, uint32_t* e) {
    if (q == line || line[0] == '#') return false;
    const uint8_t* f[3];
    size_t fl[3];
    int nf = 0;
    const uint8_t* c = line;
    while (nf < 3) {  // split on '\t' or ' ', empty fields dropped; '\n' and '\r' stay inside fields
        while (c < q && (*c == '\t' || *c == ' ')) c++;
        if (c >= q) break;
        const uint8_t* st = c;
        while (c < q && *c != '\t' && *c != ' ') c++;
        f[nf] = st;
        fl[nf] = (size_t)(c - st);
        nf++;
    }
    if (nf < 3) return false;
    if (!utf8_ok(f[0], fl[0]) || !utf8_ok(f[1], fl[1]) || !utf8_ok(f[2], fl[2])) return false;
    if (!parse_u32_rust(std::string_view((const char*)f[1], fl[1]), s)) return false;
    if (!parse_u32_rust(rtrim_unicode(std::string_view((const char*)f[2], fl[2])), e)) return false;
    if (*s >= *e) return false;
    scratch.assign((const char*)f[0], fl[0]);
    auto it = seq_num.find(scratch);
    if (it == seq_num.end()) return false;
    *chr = it->second;
    return true;
}

void parse_bed_counts(const std::string& path, const std::unordered_map<std::string, uint32_t>& seq_num, int threads, Regions* out) {
    Mapped m;
    m.open(path);
    const uint8_t* base = m.data();
    const size_t n = m.size();
    if (threads < 1) threads = 1;
    size_t nparts = std::min<size_t>((size_t)threads, std::max<size_t>(1, n >> 20));
    std::vector<size_t> cut(nparts + 1);
    cut[0] = 0;
    cut[nparts] = n;
    for (size_t k = 1; k < nparts; k++) {  // cut at line starts
        size_t c = n / nparts * k;
        const uint8_t* nl = find_byte(base + c, base + n, '\n');
        cut[k] = nl ? (size_t)(nl - base) + 1 : n;
    }
    std::vector<Regions> parts(nparts);
    auto work = [&](size_t k) {
        std::string scratch;
        const uint8_t* p = base + cut[k];
        const uint8_t* end = base + cut[k + 1];
        Regions& R = parts[k];
        while (p < end) {
            const uint8_t* nl = find_byte(p, end, '\n');
            const uint8_t* q = nl ? nl + 1 : end;  // the reference's line slice keeps its '\n'
            uint32_t c, s, e;
            if (bed_counts_line(p, q, seq_num, scratch, &c, &s, &e)) {
                R.chr.push_back(c);
                R.start.push_back(s);
                R.end.push_back(e);
            }
            p = q;
        }
    };
    parallel_jobs(nparts, (int)nparts, work);
    for (auto& R : parts) {
        out->chr.insert(out->chr.end(), R.chr.begin(), R.chr.end());
        out->start.insert(out->start.end(), R.start.begin(), R.start.end());
        out->end.insert(out->end.end(), R.end.begin(), R.end.end());
    }
}

void parse_region_arg(const std::string& region, const std::unordered_map<std::string, uint32_t>& seq_num, uint32_t* chr,
                      uint32_t* start, uint32_t* end) {
    size_t colon = region.find(':');
    if (colon == std::string::npos) throw Err(E_ARG, "Invalid region format, expected 'chr:start-end'");
    std::string_view rest = std::string_view(region).substr(colon + 1);
    size_t dash = rest.find('-');
    if (dash == std::string_view::npos) throw Err(E_ARG, "Invalid range format, expected 'start-end'");
    if (!parse_u32_rust(rest.substr(0, dash), start) || !parse_u32_rust(rest.substr(dash + 1), end))
        throw Err(E_ARG, "invalid digit found in string");
    auto it = seq_num.find(region.substr(0, colon));
    if (it == seq_num.end()) throw Err(E_ARG, "Sequence ID not found: " + region.substr(0, colon));
    if (*start >= *end)
        throw Err(E_ARG, "Region start must be less than end (" + std::to_string(*start) + " >= " + std::to_string(*end) + ")");
    *chr = it->second;
}

// ------------------------------------------------------------------------------------------
// Emission
// ------------------------------------------------------------------------------------------
void write_ranges_to(const uint8_t* text, uint64_t text_len, const uint64_t* ranges, uint64_t n, const char* out_path) {
    int fd = out_path ? ::open(out_path, O_WRONLY | O_CREAT | O_TRUNC, 0644) : 1;
    if (fd < 0) throw Err(E_IO, std::string("cannot create output file ") + out_path);
    const size_t BATCH = 1024;  // IOV_MAX
    std::vector<iovec> iov;
    iov.reserve(BATCH);
    auto flush = [&]() {
        size_t i = 0;
        while (i < iov.size()) {
            ssize_t w = ::writev(fd, iov.data() + i, (int)std::min<size_t>(iov.size() - i, BATCH));
            if (w < 0) {
                if (out_path) ::close(fd);
                throw Err(E_IO, "write failed");
            }
            size_t left = (size_t)w;
            while (i < iov.size() && left >= iov[i].iov_len) { left -= iov[i].iov_len; i++; }
            if (left > 0) {
                iov[i].iov_base = (char*)iov[i].iov_base + left;
                iov[i].iov_len -= left;
            }
        }
        iov.clear();
    };
    for (uint64_t k = 0; k < n; k++) {
        uint64_t s = ranges[2 * k], e = ranges[2 * k + 1];
        if (s >= e || e > text_len) continue;
        // coalesce with the previous slice when contiguous: fewer iovecs, same bytes
        if (!iov.empty() && (const uint8_t*)iov.back().iov_base + iov.back().iov_len == text + s) {
            iov.back().iov_len += (size_t)(e - s);
        } else {
            if (iov.size() == BATCH) flush();
            iov.push_back(iovec{(void*)(text + s), (size_t)(e - s)});
        }
    }
    flush();
    if (out_path) ::close(fd);
}

void write_text_to(const uint8_t* text, uint64_t n, const char* out_path) {
    int fd = out_path ? ::open(out_path, O_WRONLY | O_CREAT | O_TRUNC, 0644) : 1;
    if (fd < 0) throw Err(E_IO, std::string("cannot create output file ") + out_path);
    uint64_t done = 0;
    while (done < n) {
        const ssize_t w = ::write(fd, text + done, (size_t)std::min<uint64_t>(n - done, 1ull << 30));
        if (w < 0) {
            if (out_path) ::close(fd);
            throw Err(E_IO, "write failed");
        }
        done += (uint64_t)w;
    }
    if (out_path) ::close(fd);
}

}  // namespace gffx
