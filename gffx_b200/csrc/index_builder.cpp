// index_builder.cpp — writes the eight GFFx index files beside a GFF, byte-compatible with the
// reference's `gffx index` (src/index_builder/core.rs:41-242, tree layout src/utils/tree.rs:35-64,
// bincode2 framing src/utils/tree_io.rs:37-63).  The on-disk formats are frozen; this exists so that
// workloads (bench.py, the synthetic generators) can be indexed without the reference binary.
#include <algorithm>
#include <unordered_map>
#include <unordered_set>
#include "engine.h"

namespace gffx {

namespace {

struct Feat {
    uint32_t seq_off, seq_len;      // column 1 inside the text
    uint32_t start0, end0;
    uint64_t line_off;
    uint64_t id_off;                // ID value inside the text
    uint32_t id_len;
    uint64_t par_off;               // Parent value (par_len 0 = none)
    uint32_t par_len;
    uint64_t attr_off;
    uint32_t attr_len;
    bool has_attr;
};

// length of the Unicode White_Space character starting at s[i], 0 if none (Rust regex `\s`)
size_t ws_len_at(const uint8_t* s, size_t n, size_t i) {
    const uint8_t c = s[i];
    if (c == ' ' || (c >= 0x09 && c <= 0x0d)) return 1;
    if (c == 0xc2 && i + 1 < n && (s[i + 1] == 0x85 || s[i + 1] == 0xa0)) return 2;
    if (i + 2 < n) {
        const uint8_t d = s[i + 1], e = s[i + 2];
        if (c == 0xe1 && d == 0x9a && e == 0x80) return 3;
        if (c == 0xe2 && d == 0x80 && ((e >= 0x80 && e <= 0x8a) || e == 0xa8 || e == 0xa9 || e == 0xaf)) return 3;
        if (c == 0xe2 && d == 0x81 && e == 0x9f) return 3;
        if (c == 0xe3 && d == 0x80 && e == 0x80) return 3;
    }
    return 0;
}

// leftmost match of `<key>([^;\s]+)` (stop_ws) or `<key>([^;]+)` (!stop_ws) inside [s, s+n)
bool find_attr(const uint8_t* s, size_t n, const std::string& key, bool stop_ws, size_t* vs, size_t* vn) {
    const size_t kl = key.size();
    if (kl == 0 || n < kl) return false;
    for (size_t i = 0; i + kl <= n; i++) {
        if (s[i] != (uint8_t)key[0] || memcmp(s + i, key.data(), kl) != 0) continue;
        size_t j = i + kl;
        while (j < n && s[j] != ';' && !(stop_ws && ws_len_at(s, n, j))) j++;
        if (j > i + kl) { *vs = i + kl; *vn = j - (i + kl); return true; }
    }
    return false;
}

struct Iv { uint32_t start, end, fid; };

void put32(std::string& o, uint32_t v) { o.append((const char*)&v, 4); }
void put64(std::string& o, uint64_t v) { o.append((const char*)&v, 8); }

// Option<Box<Node>> of the centered tree over `ivs` (already sorted by start, stable): tag, center,
// Vec<Interval> with a u64 length, left, right.  Subsets of a start-sorted list stay start-sorted, so
// the reference's per-level stable re-sort is the identity.
void emit_tree(const std::vector<Iv>& ivs, std::string& out) {
    if (ivs.empty()) { out.push_back('\0'); return; }
    const uint32_t center = ivs[ivs.size() / 2].start;
    std::vector<Iv> left, right;
    size_t n_mid = 0;
    for (const Iv& v : ivs) n_mid += !(v.end < center) && !(v.start > center);
    out.push_back('\1');
    put32(out, center);
    put64(out, n_mid);
    for (const Iv& v : ivs) {
        if (v.end < center) left.push_back(v);
        else if (v.start > center) right.push_back(v);
        else { put32(out, v.start); put32(out, v.end); put32(out, v.fid); }
    }
    emit_tree(left, out);
    emit_tree(right, out);
}

void dump(const std::string& path, const std::string& bytes) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw Err(E_IO, "cannot create " + path);
    const size_t w = bytes.empty() ? 0 : fwrite(bytes.data(), 1, bytes.size(), f);
    if (fclose(f) != 0 || w != bytes.size()) throw Err(E_IO, "write failed: " + path);
}

}  // namespace

void build_index_files(const std::string& gff_path, const std::string& attr_key, const std::string& skip_types) {
    Mapped m;
    if (!m.try_open(gff_path)) throw Err(E_IO, "No such file or directory: " + gff_path);
    const uint8_t* d = m.data();
    const size_t len = m.size();
    if (len >= (1ull << 40)) throw Err(E_ARG, "GFF larger than 1 TiB");
    std::unordered_set<std::string> skip;
    for (size_t i = 0; i <= skip_types.size();) {
        size_t j = skip_types.find(',', i);
        if (j == std::string::npos) j = skip_types.size();
        skip.insert(skip_types.substr(i, j - i));
        i = j + 1;
    }
    const std::string id_key = "ID=", par_key = "Parent=", at_key = attr_key + "=";

    std::vector<Feat> feats;
    size_t off = 0;
    std::string scratch;
    while (off < len) {
        const uint8_t* nlp = find_byte(d + off, d + len, '\n');
        const size_t nl = nlp ? (size_t)(nlp - d) : len;
        const size_t line_off = off;
        off = nl + 1;
        if (nl == line_off || d[line_off] == '#') continue;
        if (!utf8_ok(d + line_off, nl - line_off)) throw Err(E_FORMAT, "invalid utf-8 sequence in GFF");
        std::string_view line = trim_unicode(std::string_view((const char*)d + line_off, nl - line_off));
        if (line.empty()) continue;
        const uint8_t* p = (const uint8_t*)line.data();
        const size_t n = line.size();
        size_t tabs[9];
        size_t nt = 0;
        for (size_t i = 0; i < n; i++)
            if (p[i] == '\t') { if (nt < 9) tabs[nt] = i; nt++; }
        if (nt != 8) throw Err(E_FORMAT, "Invalid GFF line (expected 9 columns): " + std::string(line));
        scratch.assign((const char*)p + tabs[1] + 1, tabs[2] - tabs[1] - 1);
        if (skip.count(scratch)) continue;
        uint32_t s1, e1;
        if (!parse_u32_rust(line.substr(tabs[2] + 1, tabs[3] - tabs[2] - 1), &s1) ||
            !parse_u32_rust(line.substr(tabs[3] + 1, tabs[4] - tabs[3] - 1), &e1))
            throw Err(E_FORMAT, "invalid digit found in string");
        if (e1 == 0) continue;
        if (s1 > e1) std::swap(s1, e1);
        Feat f{};
        f.seq_off = 0;  // filled below relative to text
        const uint64_t base = (uint64_t)(p - d);
        f.line_off = line_off;
        f.start0 = s1 == 0 ? 0 : s1 - 1;
        f.end0 = e1;
        size_t vs, vn;
        if (!find_attr(p, n, id_key, true, &vs, &vn)) throw Err(E_FORMAT, "Missing ID in feature: " + std::string(line));
        f.id_off = base + vs;
        f.id_len = (uint32_t)vn;
        if (find_attr(p, n, par_key, true, &vs, &vn)) { f.par_off = base + vs; f.par_len = (uint32_t)vn; }
        if (find_attr(p, n, at_key, false, &vs, &vn)) { f.attr_off = base + vs; f.attr_len = (uint32_t)vn; f.has_attr = true; }
        // column 1 offsets are stored relative to the line start to stay within 32 bits
        f.seq_off = (uint32_t)(base - line_off);
        f.seq_len = (uint32_t)tabs[0];
        feats.push_back(f);
    }
    auto sv = [&](uint64_t o, uint32_t l) { return std::string_view((const char*)d + o, l); };

    // ID -> index of the LAST line carrying it (core.rs:141-144)
    std::unordered_map<std::string_view, uint32_t> fmap;
    fmap.reserve(feats.size() * 2);
    for (size_t i = 0; i < feats.size(); i++) fmap[sv(feats[i].id_off, feats[i].id_len)] = (uint32_t)i;

    std::string fts, gof, prt, a2f, atn = "#attribute=" + attr_key + "\n", sqs;
    fts.reserve(feats.size() * 24);
    prt.reserve(feats.size() * 4);
    a2f.reserve(feats.size() * 4);
    std::unordered_map<std::string_view, uint32_t> seq_num, attr_num;
    std::vector<std::vector<Iv>> trees;
    bool open_root = false;
    uint32_t cur_fid = 0, cur_seq = 0, n_attr = 0;
    uint64_t cur_off = 0;
    auto put_gof = [&](uint32_t fid, uint32_t sq, uint64_t s, uint64_t e) { put32(gof, fid); put32(gof, sq); put64(gof, s); put64(gof, e); };
    for (const Feat& f : feats) {
        const std::string_view id = sv(f.id_off, f.id_len);
        const uint32_t fid = fmap[id];
        fts.append(id);
        fts.push_back('\n');
        uint32_t parent = fid;
        if (f.par_len) {
            auto it = fmap.find(sv(f.par_off, f.par_len));
            if (it != fmap.end()) parent = it->second;
        }
        put32(prt, parent);
        if (parent == fid) {  // a root: owns a tree interval and opens a .gof block
            const std::string_view seq = sv(f.line_off + f.seq_off, f.seq_len);
            auto it = seq_num.find(seq);
            uint32_t sn;
            if (it == seq_num.end()) {
                sn = (uint32_t)trees.size();
                seq_num.emplace(seq, sn);
                trees.emplace_back();
                sqs.append(seq);
                sqs.push_back('\n');
            } else {
                sn = it->second;
            }
            trees[sn].push_back(Iv{f.start0, f.end0, fid});
            if (open_root) put_gof(cur_fid, cur_seq, cur_off, f.line_off);
            open_root = true;
            cur_fid = fid; cur_seq = sn; cur_off = f.line_off;
        }
        if (f.has_attr) {
            const std::string_view v = sv(f.attr_off, f.attr_len);
            auto it = attr_num.find(v);
            uint32_t a;
            if (it == attr_num.end()) {
                a = n_attr++;
                attr_num.emplace(v, a);
                atn.append(v);
                atn.push_back('\n');
            } else {
                a = it->second;
            }
            put32(a2f, a);
        } else {
            put32(a2f, UINT32_MAX);
        }
    }
    if (open_root) put_gof(cur_fid, cur_seq, cur_off, len);

    std::string rit, rix = "[";
    for (size_t t = 0; t < trees.size(); t++) {
        if (t) rix += ",";
        rix += std::to_string(rit.size());
        std::stable_sort(trees[t].begin(), trees[t].end(), [](const Iv& a, const Iv& b) { return a.start < b.start; });
        emit_tree(trees[t], rit);
    }
    rix += "]";
    dump(gff_path + ".fts", fts);
    dump(gff_path + ".gof", gof);
    dump(gff_path + ".rit", rit);
    dump(gff_path + ".rix", rix);
    dump(gff_path + ".sqs", sqs);
    dump(gff_path + ".atn", atn);
    dump(gff_path + ".a2f", a2f);
    dump(gff_path + ".prt", prt);
}

}  // namespace gffx
