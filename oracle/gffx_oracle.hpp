// gffx_oracle.hpp — CPU restatement of GFFx v0.4.0's intersect / depth / coverage hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under gffx_b200/ may include, link or execute this file;
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs do.
//
// Parity status: "parity unpinned by the reference" — the reference (pure Rust) ships no tests,
// no golden vectors and cannot be compiled in this image (no cargo/rustc).  The restatement is
// pinned instead by (1) the hand-derived known-answer fixture of SURVEY.md App. C,
// (2) root-level counts on the reference's real `benchmark/ps.gff.{rit,rix,sqs,gof}` index files
// with `benchmark/intersect/ps/input_001.txt`, both committed under tests/golden/.
//
// Every function cites the reference file:line (relative to /root/reference/) it follows.
// The algorithms are kept as the reference has them (boxed centered interval tree, single-threaded
// root join, per-block text re-scan with a linear scan over the chromosome's regions, 100k-line
// batching in depth) so that timing this code is a fair "port" CPU baseline.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <string_view>
#include <unordered_map>
#include <unordered_set>
#include <vector>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace gffx_oracle {

static constexpr uint64_t MISSING = UINT64_MAX;  // src/index_loader/gof.rs:7

struct Error : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// ---------------------------------------------------------------------------------------------
// mmap helper (reference: memmap2::Mmap::map, e.g. src/index_loader/core.rs:14-17)
// ---------------------------------------------------------------------------------------------
struct MappedFile {
    const uint8_t* data = nullptr;
    size_t len = 0;
    int fd = -1;
    MappedFile() = default;
    MappedFile(const MappedFile&) = delete;
    MappedFile& operator=(const MappedFile&) = delete;
    explicit MappedFile(const std::string& path) { open(path); }
    void open(const std::string& path) {
        fd = ::open(path.c_str(), O_RDONLY);
        if (fd < 0) throw Error("Failed to open file: " + path);
        struct stat st;
        if (fstat(fd, &st) != 0) throw Error("fstat failed: " + path);
        len = (size_t)st.st_size;
        if (len > 0) {
            void* p = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
            if (p == MAP_FAILED) throw Error("Failed to mmap file: " + path);
            data = (const uint8_t*)p;
        }
    }
    ~MappedFile() {
        if (data) munmap((void*)data, len);
        if (fd >= 0) ::close(fd);
    }
};

// src/utils/common.rs:123-127 — `{filename}{suffix}` beside the GFF
inline std::string append_suffix(const std::string& path, const std::string& suffix) { return path + suffix; }

// ---------------------------------------------------------------------------------------------
// Small text helpers mirroring the Rust std behaviours the reference leans on
// ---------------------------------------------------------------------------------------------
inline bool valid_utf8(const uint8_t* s, size_t n) {  // std::str::from_utf8
    size_t i = 0;
    while (i < n) {
        uint8_t c = s[i];
        if (c < 0x80) { i++; continue; }
        size_t need;
        uint32_t cp;
        if ((c & 0xE0) == 0xC0) { need = 1; cp = c & 0x1F; if (c < 0xC2) return false; }
        else if ((c & 0xF0) == 0xE0) { need = 2; cp = c & 0x0F; }
        else if ((c & 0xF8) == 0xF0) { need = 3; cp = c & 0x07; if (c > 0xF4) return false; }
        else return false;
        if (n - i <= need) return false;  // truncated sequence
        for (size_t k = 1; k <= need; k++) {
            uint8_t cc = s[i + k];
            if ((cc & 0xC0) != 0x80) return false;
            cp = (cp << 6) | (cc & 0x3F);
        }
        if (need == 2 && (cp < 0x800 || (cp >= 0xD800 && cp <= 0xDFFF))) return false;
        if (need == 3 && (cp < 0x10000 || cp > 0x10FFFF)) return false;
        i += need + 1;
    }
    return true;
}

// u8::is_ascii_whitespace: space, \t, \n, \x0C, \r  (NOT \x0B)
inline bool is_ascii_ws(uint8_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\x0c' || c == '\r'; }

// Length of the Unicode White_Space char ending at s[n-1] (0 if none) — for str::trim_end / trim.
inline size_t ws_suffix_len(const uint8_t* s, size_t n) {
    if (n == 0) return 0;
    uint8_t c = s[n - 1];
    if (c == ' ' || (c >= 0x09 && c <= 0x0D)) return 1;
    if (n >= 2 && s[n - 2] == 0xC2 && (c == 0x85 || c == 0xA0)) return 2;
    if (n >= 3) {
        uint8_t a = s[n - 3], b = s[n - 2];
        if (a == 0xE1 && b == 0x9A && c == 0x80) return 3;                        // U+1680
        if (a == 0xE2 && b == 0x80 && ((c >= 0x80 && c <= 0x8A) || c == 0xA8 || c == 0xA9 || c == 0xAF)) return 3;
        if (a == 0xE2 && b == 0x81 && c == 0x9F) return 3;                        // U+205F
        if (a == 0xE3 && b == 0x80 && c == 0x80) return 3;                        // U+3000
    }
    return 0;
}
inline size_t ws_prefix_len(const uint8_t* s, size_t n) {
    if (n == 0) return 0;
    uint8_t c = s[0];
    if (c == ' ' || (c >= 0x09 && c <= 0x0D)) return 1;
    if (n >= 2 && c == 0xC2 && (s[1] == 0x85 || s[1] == 0xA0)) return 2;
    if (n >= 3) {
        uint8_t b = s[1], d = s[2];
        if (c == 0xE1 && b == 0x9A && d == 0x80) return 3;
        if (c == 0xE2 && b == 0x80 && ((d >= 0x80 && d <= 0x8A) || d == 0xA8 || d == 0xA9 || d == 0xAF)) return 3;
        if (c == 0xE2 && b == 0x81 && d == 0x9F) return 3;
        if (c == 0xE3 && b == 0x80 && d == 0x80) return 3;
    }
    return 0;
}
inline std::string_view trim_end(std::string_view s) {
    while (true) {
        size_t k = ws_suffix_len((const uint8_t*)s.data(), s.size());
        if (!k) break;
        s.remove_suffix(k);
    }
    return s;
}
inline std::string_view trim(std::string_view s) {
    s = trim_end(s);
    while (true) {
        size_t k = ws_prefix_len((const uint8_t*)s.data(), s.size());
        if (!k) break;
        s.remove_prefix(k);
    }
    return s;
}

// Rust `str::parse::<u32>()`: optional leading '+', then >=1 ASCII digits, overflow is an error.
inline bool rust_parse_u32(std::string_view s, uint32_t& out) {
    size_t i = 0;
    if (!s.empty() && s[0] == '+') i = 1;
    if (i >= s.size()) return false;
    uint64_t v = 0;
    for (; i < s.size(); i++) {
        unsigned d = (unsigned)(uint8_t)s[i] - '0';
        if (d > 9) return false;
        v = v * 10 + d;
        if (v > UINT32_MAX) return false;
    }
    out = (uint32_t)v;
    return true;
}

// src/commands/intersect.rs:526-538 parse_u32_ascii == src/commands/depth.rs:86-95 parse_u32_fast:
// non-empty, digits only, checked overflow.
inline bool parse_u32_digits(const uint8_t* s, size_t n, uint32_t& out) {
    if (n == 0) return false;
    uint64_t v = 0;
    for (size_t i = 0; i < n; i++) {
        unsigned d = (unsigned)s[i] - '0';
        if (d > 9) return false;
        v = v * 10 + d;
        if (v > UINT32_MAX) return false;
    }
    out = (uint32_t)v;
    return true;
}

// lexical_core::parse::<u32> as used at src/commands/intersect.rs:223-224: whole slice must be a decimal u32.
// lexical-parse-integer 1.0's parse_sign accepts ONE leading '+' for unsigned types in the standard format (a '-' is an
// error, so is a sign without digits) - recalled from the crate's source, which is not vendored here: the same language
// as str::parse::<u32>.  tests/test_kat_quirks.py::test_bed_dialects holds the KAT line.
inline bool lexical_parse_u32(std::string_view s, uint32_t& out) {
    return rust_parse_u32(s, out);
}

// ---------------------------------------------------------------------------------------------
// Interval tree — src/utils/tree.rs
// ---------------------------------------------------------------------------------------------
struct Interval {  // src/utils/tree.rs:6-10
    uint32_t start, end, root_fid;
};

struct Node {  // src/utils/tree.rs:18-23
    uint32_t center;
    std::vector<Interval> intervals;
    std::unique_ptr<Node> left, right;
};

struct IntervalTree {  // src/utils/tree.rs:13-15
    std::unique_ptr<Node> root;

    // src/utils/tree.rs:30-64
    static std::unique_ptr<Node> build(std::vector<Interval> ivs) {
        if (ivs.empty()) return nullptr;
        std::stable_sort(ivs.begin(), ivs.end(), [](const Interval& a, const Interval& b) { return a.start < b.start; });
        size_t mid = ivs.size() / 2;
        uint32_t center = ivs[mid].start;
        std::vector<Interval> left, right, mine;
        for (const auto& iv : ivs) {
            if (iv.end < center) left.push_back(iv);
            else if (iv.start > center) right.push_back(iv);
            else mine.push_back(iv);
        }
        auto n = std::make_unique<Node>();
        n->center = center;
        n->intervals = std::move(mine);
        n->left = build(std::move(left));
        n->right = build(std::move(right));
        return n;
    }
    static IntervalTree from_intervals(std::vector<Interval> ivs) {
        IntervalTree t;
        t.root = build(std::move(ivs));
        return t;
    }

    // src/utils/tree.rs:98-121 — predicate at :110 `iv.start < end && iv.end > start`
    static void query_rec(const Node* n, uint32_t start, uint32_t end, std::vector<const Interval*>& out) {
        if (!n) return;
        for (const auto& iv : n->intervals)
            if (iv.start < end && iv.end > start) out.push_back(&iv);
        if (start < n->center) query_rec(n->left.get(), start, end, out);
        if (end > n->center) query_rec(n->right.get(), start, end, out);
    }
    void query_interval(uint32_t start, uint32_t end, std::vector<const Interval*>& out) const {
        query_rec(root.get(), start, end, out);
    }
    static void collect(const Node* n, std::vector<Interval>& out) {
        if (!n) return;
        for (const auto& iv : n->intervals) out.push_back(iv);
        collect(n->left.get(), out);
        collect(n->right.get(), out);
    }
    std::vector<Interval> flatten() const {
        std::vector<Interval> v;
        collect(root.get(), v);
        return v;
    }
};

// bincode2 v2 default config (little-endian, fixed-width ints, u64 lengths, Option = 1 tag byte):
// the `.rit` grammar of SURVEY.md App. B, written by src/utils/tree_io.rs:37-54.
struct ByteReader {
    const uint8_t* p;
    size_t n, pos = 0;
    uint8_t u8() { need(1); return p[pos++]; }
    uint32_t u32() { need(4); uint32_t v; memcpy(&v, p + pos, 4); pos += 4; return v; }
    uint64_t u64() { need(8); uint64_t v; memcpy(&v, p + pos, 8); pos += 8; return v; }
    void need(size_t k) { if (pos + k > n) throw Error("bincode2 deserialize: unexpected end of input"); }
};
inline std::unique_ptr<Node> decode_opt_node(ByteReader& r, int depth = 0) {
    uint8_t tag = r.u8();
    if (tag == 0) return nullptr;
    if (tag != 1) throw Error("bincode2 deserialize: bad Option tag");
    if (depth > 100000) throw Error("bincode2 deserialize: tree too deep");
    auto n = std::make_unique<Node>();
    n->center = r.u32();
    uint64_t cnt = r.u64();
    if (cnt > (r.n - r.pos) / 12) throw Error("bincode2 deserialize: bad vec length");
    n->intervals.resize((size_t)cnt);
    for (auto& iv : n->intervals) { iv.start = r.u32(); iv.end = r.u32(); iv.root_fid = r.u32(); }
    n->left = decode_opt_node(r, depth + 1);
    n->right = decode_opt_node(r, depth + 1);
    return n;
}
inline void encode_opt_node(const Node* n, std::string& out) {
    if (!n) { out.push_back('\0'); return; }
    out.push_back('\1');
    auto put32 = [&](uint32_t v) { out.append((const char*)&v, 4); };
    uint64_t cnt = n->intervals.size();
    put32(n->center);
    out.append((const char*)&cnt, 8);
    for (const auto& iv : n->intervals) { put32(iv.start); put32(iv.end); put32(iv.root_fid); }
    encode_opt_node(n->left.get(), out);
    encode_opt_node(n->right.get(), out);
}

// ---------------------------------------------------------------------------------------------
// Index loaders
// ---------------------------------------------------------------------------------------------
// src/index_loader/core.rs:19-34 (BufRead::lines: split on '\n', strip one trailing '\r')
inline void load_sqs(const std::string& gff, std::vector<std::string>& id_to_name,
                     std::unordered_map<std::string, uint32_t>& name_to_id) {
    std::string p = append_suffix(gff, ".sqs");
    MappedFile f(p);
    id_to_name.clear();
    name_to_id.clear();
    size_t pos = 0;
    while (pos < f.len) {
        const uint8_t* nl = (const uint8_t*)memchr(f.data + pos, '\n', f.len - pos);
        size_t e = nl ? (size_t)(nl - f.data) : f.len;
        size_t le = e;
        if (nl && le > pos && f.data[le - 1] == '\r') le--;
        id_to_name.emplace_back((const char*)f.data + pos, le - pos);
        pos = e + 1;
    }
    for (size_t i = 0; i < id_to_name.size(); i++) name_to_id[id_to_name[i]] = (uint32_t)i;  // later dup wins
}

// serde_json Vec<u64> (compact or pretty) — src/utils/tree_index.rs:49-53
inline std::vector<uint64_t> parse_json_u64_array(const uint8_t* s, size_t n, const std::string& what) {
    std::vector<uint64_t> out;
    size_t i = 0;
    auto skipws = [&] { while (i < n && (s[i] == ' ' || s[i] == '\n' || s[i] == '\r' || s[i] == '\t')) i++; };
    skipws();
    if (i >= n || s[i] != '[') throw Error("parse json " + what);
    i++;
    skipws();
    if (i < n && s[i] == ']') { return out; }
    while (true) {
        skipws();
        if (i >= n || s[i] < '0' || s[i] > '9') throw Error("parse json " + what);
        uint64_t v = 0;
        while (i < n && s[i] >= '0' && s[i] <= '9') { v = v * 10 + (s[i] - '0'); i++; }
        out.push_back(v);
        skipws();
        if (i < n && s[i] == ',') { i++; continue; }
        if (i < n && s[i] == ']') break;
        throw Error("parse json " + what);
    }
    return out;
}

struct TreeIndexData {  // src/utils/tree_index.rs:12-16
    std::unordered_map<uint32_t, IntervalTree> chr_entries;
    std::unordered_map<std::string, uint32_t> seqid_to_num;
    std::vector<std::string> num_to_seqid;

    // src/utils/tree_index.rs:21-82
    static TreeIndexData load_tree_index(const std::string& gff) {
        TreeIndexData d;
        load_sqs(gff, d.num_to_seqid, d.seqid_to_num);
        std::string rit = append_suffix(gff, ".rit"), rix = append_suffix(gff, ".rix");
        MappedFile buf(rit);
        MappedFile jf(rix);
        std::vector<uint64_t> offsets = parse_json_u64_array(jf.data, jf.len, rix);
        if (offsets.empty()) return d;
        for (size_t i = 1; i < offsets.size(); i++)
            if (offsets[i - 1] > offsets[i]) throw Error("offsets not sorted ascending");
        if (offsets.back() > buf.len) throw Error("last offset out of file size");
        for (size_t i = 0; i < offsets.size(); i++) {
            size_t s = (size_t)offsets[i];
            size_t e = (i + 1 < offsets.size()) ? (size_t)offsets[i + 1] : buf.len;
            if (e < s || e > buf.len) throw Error("bad slice range");
            ByteReader r{buf.data + s, e - s};
            IntervalTree t;
            t.root = decode_opt_node(r);
            d.chr_entries.emplace((uint32_t)i, std::move(t));
        }
        return d;
    }
};

struct GofEntry {  // src/index_loader/gof.rs:10-15
    uint32_t feature_id, seqid_num;
    uint64_t start_offset, end_offset;
};
struct GofMap {  // src/index_loader/gof.rs:20-27
    std::vector<GofEntry> entries;
    std::unordered_map<uint32_t, std::pair<uint64_t, uint64_t>> index;  // :32-37, later duplicate wins

    // src/index_loader/gof.rs:54-84
    std::vector<std::tuple<uint32_t, uint64_t, uint64_t>> roots_to_offsets(const std::vector<uint32_t>& roots) const {
        std::vector<std::tuple<uint32_t, uint64_t, uint64_t>> out;
        out.reserve(roots.size());
        for (uint32_t r : roots) {
            auto it = index.find(r);
            if (it != index.end()) out.emplace_back(r, it->second.first, it->second.second);
            else out.emplace_back(r, MISSING, MISSING);
        }
        return out;
    }
};
// src/index_loader/gof.rs:95-128
inline GofMap load_gof(const std::string& gff) {
    std::string p = append_suffix(gff, ".gof");
    MappedFile f(p);
    const size_t REC = 24;
    if (f.len % REC != 0) throw Error("Corrupted GOF (" + p + "): length not multiple of 24");
    GofMap m;
    m.entries.reserve(f.len / REC);
    for (size_t off = 0; off < f.len; off += REC) {
        GofEntry e;
        memcpy(&e.feature_id, f.data + off, 4);
        memcpy(&e.seqid_num, f.data + off + 4, 4);
        memcpy(&e.start_offset, f.data + off + 8, 8);
        memcpy(&e.end_offset, f.data + off + 16, 8);
        m.entries.push_back(e);
    }
    for (const auto& e : m.entries) m.index[e.feature_id] = {e.start_offset, e.end_offset};
    return m;
}

// ---------------------------------------------------------------------------------------------
// intersect — src/commands/intersect.rs
// ---------------------------------------------------------------------------------------------
enum class OverlapMode { Contained = 0, ContainsRegion = 1, Overlap = 2 };  // :74-78

struct Region { uint32_t chr, start, end; };

// src/commands/intersect.rs:201-230 (BED dialect "intersect")
inline std::vector<Region> parse_bed_file(const std::string& bed, const std::unordered_map<std::string, uint32_t>& seqid_map) {
    MappedFile f(bed);
    std::vector<Region> regions;
    size_t pos = 0;
    while (pos <= f.len) {
        const uint8_t* nl = pos < f.len ? (const uint8_t*)memchr(f.data + pos, '\n', f.len - pos) : nullptr;
        size_t e = nl ? (size_t)(nl - f.data) : f.len;
        const uint8_t* line = f.data + pos;
        size_t n = e - pos;
        pos = e + 1;
        if (n == 0 || line[0] == '#') { if (!nl) break; continue; }
        if (!valid_utf8(line, n)) throw Error("invalid utf-8 sequence in BED line");  // :215
        std::string_view parts[3];
        int np = 0;
        size_t i = 0;
        while (i < n && np < 3) {  // split_ascii_whitespace
            while (i < n && is_ascii_ws(line[i])) i++;
            if (i >= n) break;
            size_t s = i;
            while (i < n && !is_ascii_ws(line[i])) i++;
            parts[np++] = std::string_view((const char*)line + s, i - s);
        }
        if (np == 3) {
            auto it = seqid_map.find(std::string(parts[0]));  // chromosome lookup BEFORE number parse (:220-222)
            if (it != seqid_map.end()) {
                uint32_t s, en;
                if (!lexical_parse_u32(parts[1], s)) throw Error("BED: invalid start '" + std::string(parts[1]) + "'");
                if (!lexical_parse_u32(parts[2], en)) throw Error("BED: invalid end '" + std::string(parts[2]) + "'");
                regions.push_back({it->second, s, en});
            }
        }
        if (!nl) break;
    }
    return regions;
}

// src/commands/intersect.rs:172-198
inline Region parse_region(const std::string& region, const std::unordered_map<std::string, uint32_t>& seqid_map) {
    size_t c = region.find(':');
    if (c == std::string::npos) throw Error("Invalid region format, expected 'chr:start-end'");
    std::string seq = region.substr(0, c), range = region.substr(c + 1);
    size_t d = range.find('-');
    if (d == std::string::npos) throw Error("Invalid range format, expected 'start-end'");
    uint32_t s, e;
    if (!rust_parse_u32(std::string_view(range).substr(0, d), s)) throw Error("invalid digit found in string");
    if (!rust_parse_u32(std::string_view(range).substr(d + 1), e)) throw Error("invalid digit found in string");
    auto it = seqid_map.find(seq);
    if (it == seqid_map.end()) throw Error("Sequence ID not found: " + seq);
    if (s >= e) throw Error("Region start must be less than end");
    return {it->second, s, e};
}

struct Pair { uint32_t root_fid, start, end; };

// src/commands/intersect.rs:105-169.  Chromosome iteration order is FxHashMap order in the
// reference (unspecified); here ascending seq number — only the multiset matters downstream.
inline std::vector<Pair> query_features(const TreeIndexData& idx, const std::vector<Region>& regions, OverlapMode mode, bool invert) {
    std::vector<std::vector<Region>> buckets(idx.seqid_to_num.size());
    for (const auto& r : regions) {
        if (r.chr >= buckets.size()) throw Error("region chromosome out of range");  // Rust would panic on b[chr]
        buckets[r.chr].push_back(r);
    }
    std::vector<Pair> results;
    std::vector<uint32_t> seqs;
    for (const auto& kv : idx.chr_entries) seqs.push_back(kv.first);
    std::sort(seqs.begin(), seqs.end());
    std::vector<const Interval*> hits;
    for (uint32_t seq : seqs) {
        if (seq >= buckets.size()) throw Error("tree seq number beyond .sqs");  // Rust: index out of bounds panic
        const auto& regs = buckets[seq];
        if (regs.empty()) continue;
        const IntervalTree& tree = idx.chr_entries.at(seq);
        for (const auto& r : regs) {
            hits.clear();
            tree.query_interval(r.start, r.end, hits);
            for (const Interval* iv : hits) {
                bool keep;
                switch (mode) {
                    case OverlapMode::Contained: keep = iv->start >= r.start && iv->end <= r.end; break;       // :146-149
                    case OverlapMode::ContainsRegion: keep = iv->start <= r.start && iv->end >= r.end; break;  // :150-153
                    default: keep = true;                                                                        // :154-157
                }
                if (invert != keep) results.push_back({iv->root_fid, iv->start, iv->end});  // :161-163
            }
        }
    }
    return results;
}

// src/commands/intersect.rs:80-102
inline bool gff_type_allowed(const uint8_t* line, size_t n, const std::unordered_set<std::string>& allow) {
    size_t off = 0;
    for (int t = 0; t < 2; t++) {
        const uint8_t* p = (const uint8_t*)memchr(line + off, '\t', n - off);
        if (!p) return false;
        off = (size_t)(p - line) + 1;
    }
    const uint8_t* p = (const uint8_t*)memchr(line + off, '\t', n - off);
    if (!p) return false;
    size_t i2 = (size_t)(p - line);
    if (!valid_utf8(line + off, i2 - off)) return false;
    return allow.count(std::string((const char*)line + off, i2 - off)) != 0;
}

using IvMap = std::unordered_map<std::string, std::vector<std::pair<uint32_t, uint32_t>>>;

// src/commands/intersect.rs:441-523
inline bool gff_line_overlaps_queries(const uint8_t* line, size_t n, const IvMap& ivmap, OverlapMode mode) {
    size_t off = 0;
    size_t tabs[5];
    for (int t = 0; t < 5; t++) {
        const uint8_t* p = (const uint8_t*)memchr(line + off, '\t', n - off);
        if (!p) return false;
        tabs[t] = (size_t)(p - line);
        off = tabs[t] + 1;
    }
    uint32_t start, end;
    if (!parse_u32_digits(line + tabs[2] + 1, tabs[3] - tabs[2] - 1, start)) return false;
    if (!parse_u32_digits(line + tabs[3] + 1, tabs[4] - tabs[3] - 1, end)) return false;
    if (!valid_utf8(line, tabs[0])) return false;
    auto it = ivmap.find(std::string((const char*)line, tabs[0]));
    if (it == ivmap.end()) return false;
    for (const auto& q : it->second) {
        uint32_t qs = q.first, qe = q.second;
        bool keep;
        switch (mode) {
            case OverlapMode::Contained: keep = start >= qs && end <= qe; break;
            case OverlapMode::ContainsRegion: keep = start <= qs && end >= qe; break;
            default:
                keep = (qs <= start && start <= qe) || (qs <= end && end <= qe) || (start <= qs && qs <= end) ||
                       (start <= qe && qe <= end);
        }
        if (keep) return true;
    }
    return false;
}

using Blocks = std::vector<std::tuple<uint32_t, uint64_t, uint64_t>>;
using Ranges = std::vector<std::pair<uint64_t, uint64_t>>;

// src/commands/intersect.rs:232-438 up to the write: returns the emitted (abs_start, abs_end)
// ranges in output order (blocks by ascending start, lines in file order).  `warn` receives the
// fids skipped for a sentinel offset (:269-272).
inline Ranges match_only_ranges(const uint8_t* gff, size_t file_len, const Blocks& blocks, const IvMap& ivmap,
                                const char* types_filter, OverlapMode mode, std::vector<uint32_t>* warn = nullptr) {
    bool have_types = types_filter != nullptr;
    std::unordered_set<std::string> allow;
    if (have_types) {  // :252-259
        std::string_view s(types_filter);
        size_t i = 0;
        while (i <= s.size()) {
            size_t j = s.find(',', i);
            if (j == std::string_view::npos) j = s.size();
            std::string_view t = trim(s.substr(i, j - i));
            if (!t.empty()) allow.insert(std::string(t));
            i = j + 1;
        }
    }
    std::vector<std::pair<uint64_t, Ranges>> parts(blocks.size());
    std::vector<char> used(blocks.size(), 0);
#pragma omp parallel for schedule(dynamic, 16)
    for (long bi = 0; bi < (long)blocks.size(); bi++) {  // rayon par_iter over blocks (:266-329)
        auto [root, start, end] = blocks[bi];
        if (start == MISSING) continue;
        size_t s = (size_t)start;
        size_t e = std::min<uint64_t>(end, file_len);
        if (s >= e || e > file_len) continue;
        const uint8_t* src = gff + s;
        size_t slen = e - s;
        Ranges matched;
        size_t pos = 0;
        while (pos < slen) {
            const uint8_t* p = (const uint8_t*)memchr(src + pos, '\n', slen - pos);
            size_t nl = p ? (size_t)(p - src) + 1 : slen;
            const uint8_t* line = src + pos;
            size_t ln = nl - pos;
            if (ln > 0 && line[ln - 1] == '\n') ln--;
            if (ln > 0 && line[0] != '#') {
                bool pass = true;
                if (have_types && !gff_type_allowed(line, ln, allow)) pass = false;
                if (pass && gff_line_overlaps_queries(line, ln, ivmap, mode))
                    matched.emplace_back(start + pos, start + nl);
            }
            pos = nl;
        }
        if (!matched.empty()) { parts[bi] = {start, std::move(matched)}; used[bi] = 1; }
    }
    if (warn)
        for (const auto& b : blocks) if (std::get<1>(b) == MISSING) warn->push_back(std::get<0>(b));
    std::vector<size_t> order;
    for (size_t i = 0; i < parts.size(); i++) if (used[i]) order.push_back(i);
    std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return parts[a].first < parts[b].first; });  // :335
    Ranges out;
    for (size_t i : order) out.insert(out.end(), parts[i].second.begin(), parts[i].second.end());
    return out;
}

// src/utils/common.rs:188-287 up to the write: merged byte ranges in output order.
inline Ranges entire_group_ranges(size_t file_len, const Blocks& blocks, std::vector<uint32_t>* warn = nullptr) {
    Ranges sorted;
    for (const auto& [fid, s, e] : blocks) {
        if (s == MISSING) { if (warn) warn->push_back(fid); continue; }
        sorted.emplace_back(s, e);
    }
    std::stable_sort(sorted.begin(), sorted.end(), [](auto& a, auto& b) { return a.first < b.first; });
    Ranges merged;
    if (!sorted.empty()) {
        uint64_t cs = sorted[0].first, ce = sorted[0].second;
        for (size_t i = 1; i < sorted.size(); i++) {
            auto [s, e] = sorted[i];
            if (s <= ce) ce = std::max(ce, e);
            else { if (cs < ce) merged.emplace_back(cs, ce); cs = s; ce = e; }
        }
        if (cs < ce) merged.emplace_back(cs, ce);
    }
    Ranges out;
    for (auto [so, eo] : merged) {
        if (so >= eo) continue;
        if (eo > file_len) continue;
        out.emplace_back(so, eo);
    }
    return out;
}

inline void write_ranges(const uint8_t* gff, const Ranges& ranges, FILE* out) {
    for (auto [s, e] : ranges) fwrite(gff + s, 1, (size_t)(e - s), out);
}

struct IntersectArgs {
    std::string input;          // -i
    std::string bed;            // -b  (exactly one of bed/region)
    std::string region;         // -r
    std::string output;         // -o ("" = stdout)
    bool contained = false, contains_region = false, overlap = false, invert = false;
    bool entire_group = false;  // -e/--entire_group
    bool has_types = false;
    std::string types;          // -T
    bool verbose = false;
};

struct IntersectResult {
    std::vector<Pair> pairs;
    Ranges ranges;
    std::vector<uint32_t> warned_fids;
};

// src/commands/intersect.rs:541-655 without the final write.
inline IntersectResult intersect_compute(const IntersectArgs& a, const TreeIndexData& idx, const GofMap& gof,
                                         const std::vector<Region>& regions, const uint8_t* gff, size_t gff_len) {
    OverlapMode mode = a.contained ? OverlapMode::Contained : a.contains_region ? OverlapMode::ContainsRegion : OverlapMode::Overlap;
    IntersectResult res;
    res.pairs = query_features(idx, regions, mode, a.invert);
    std::vector<uint32_t> roots;
    {
        std::unordered_set<uint32_t> seen;
        for (const auto& p : res.pairs) if (seen.insert(p.root_fid).second) roots.push_back(p.root_fid);
    }
    Blocks blocks = gof.roots_to_offsets(roots);
    if (!a.entire_group || a.has_types) {  // :619
        IvMap ivmap;                        // :621-633 — all regions, keyed by seq NAME
        for (const auto& r : regions)
            if (r.chr < idx.num_to_seqid.size()) {
                // num_to_seq is rebuilt from seqid_to_num in the reference: a name maps to its LAST number
                auto it = idx.seqid_to_num.find(idx.num_to_seqid[r.chr]);
                if (it != idx.seqid_to_num.end() && it->second == r.chr) ivmap[idx.num_to_seqid[r.chr]].emplace_back(r.start, r.end);
            }
        res.ranges = match_only_ranges(gff, gff_len, blocks, ivmap, a.has_types ? a.types.c_str() : nullptr, mode, &res.warned_fids);
    } else {
        res.ranges = entire_group_ranges(gff_len, blocks, &res.warned_fids);
    }
    return res;
}

inline void intersect_run(const IntersectArgs& a) {
    TreeIndexData idx = TreeIndexData::load_tree_index(a.input);
    std::vector<Region> regions;
    if (!a.bed.empty()) regions = parse_bed_file(a.bed, idx.seqid_to_num);
    else if (!a.region.empty()) regions.push_back(parse_region(a.region, idx.seqid_to_num));
    else throw Error("No region specified");
    GofMap gof = load_gof(a.input);
    MappedFile gff(a.input);
    IntersectResult r = intersect_compute(a, idx, gof, regions, gff.data, gff.len);
    for (uint32_t fid : r.warned_fids) fprintf(stderr, "[WARN] skipped fid=%u due to sentinel start offset\n", fid);
    FILE* out = a.output.empty() ? stdout : fopen(a.output.c_str(), "wb");
    if (!out) throw Error("cannot create output " + a.output);
    write_ranges(gff.data, r.ranges, out);
    if (!a.output.empty()) fclose(out); else fflush(out);
}

// ---------------------------------------------------------------------------------------------
// depth / coverage shared pieces
// ---------------------------------------------------------------------------------------------
// src/commands/depth.rs:104-118 == src/commands/coverage.rs:74-88
inline bool fast_id(const uint8_t* attrs, size_t n, std::string_view& id) {
    size_t i = 0;
    while (i + 2 < n) {
        if (attrs[i] == 'I' && attrs[i + 1] == 'D' && attrs[i + 2] == '=') {
            size_t j = i + 3;
            while (j < n && attrs[j] != ';' && attrs[j] != ' ' && attrs[j] != '\t') j++;
            id = std::string_view((const char*)attrs + i + 3, j - (i + 3));
            return true;  // sub-slice of valid UTF-8 cut at ASCII bytes is valid UTF-8
        }
        i++;
    }
    return false;
}

struct FeatLine { uint32_t start, end, id_idx; };

// The per-block line parse shared by src/commands/depth.rs:131-157 and src/commands/coverage.rs:296-323.
inline void parse_block_lines(const uint8_t* slice, size_t n, std::vector<std::string>& id_strings,
                              std::vector<std::string>& id_chrom, std::vector<FeatLine>& feats) {
    if (!valid_utf8(slice, n)) return;  // `if let Ok(text) = str::from_utf8(gff_slice)`
    std::unordered_map<std::string, uint32_t> id_to_idx;
    size_t pos = 0;
    while (pos < n) {  // split_terminator('\n')
        const uint8_t* p = (const uint8_t*)memchr(slice + pos, '\n', n - pos);
        size_t e = p ? (size_t)(p - slice) : n;
        const uint8_t* line = slice + pos;
        size_t ln = e - pos;
        pos = e + 1;
        if (ln == 0 || line[0] == '#') continue;
        size_t tabs[8];
        size_t off = 0;
        bool ok = true;
        for (int t = 0; t < 8; t++) {  // splitn(9,'\t') needs 9 pieces
            const uint8_t* q = (const uint8_t*)memchr(line + off, '\t', ln - off);
            if (!q) { ok = false; break; }
            tabs[t] = (size_t)(q - line);
            off = tabs[t] + 1;
        }
        if (!ok) continue;
        uint32_t s1, e1;
        if (!parse_u32_digits(line + tabs[2] + 1, tabs[3] - tabs[2] - 1, s1)) continue;
        if (!parse_u32_digits(line + tabs[3] + 1, tabs[4] - tabs[3] - 1, e1)) continue;
        if (e1 == 0) continue;
        if (s1 > e1) std::swap(s1, e1);
        uint32_t fs = s1 == 0 ? 0 : s1 - 1, fe = e1;
        std::string_view id;
        if (fast_id(line + tabs[7] + 1, ln - tabs[7] - 1, id)) {
            std::string key(id);
            auto it = id_to_idx.find(key);
            uint32_t idx;
            if (it == id_to_idx.end()) {
                idx = (uint32_t)id_strings.size();
                id_to_idx.emplace(key, idx);
                id_strings.push_back(key);
                id_chrom.emplace_back((const char*)line, tabs[0]);
            } else idx = it->second;
            feats.push_back({fs, fe, idx});
        }
    }
}

struct IdRow {
    std::string chrom;
    uint32_t start, end;
    uint64_t value;  // depth or breadth
    uint64_t first_block;  // tie-break for `chrom` (reference: first writer wins, order unspecified)
};
using IdMap = std::unordered_map<std::string, IdRow>;

// merge rule of src/commands/depth.rs:266-271,501-508 / src/commands/coverage.rs:431-443.
// The reference keeps the chrom of whichever partial arrives first (hash / thread order); the
// restatement keeps the chrom of the block with the lowest file offset so results are reproducible.
inline void merge_row(IdMap& g, const std::string& id, const std::string& chrom, uint32_t s, uint32_t e, uint64_t v, uint64_t blk_off) {
    auto it = g.find(id);
    if (it == g.end()) { g.emplace(id, IdRow{chrom, s, e, v, blk_off}); return; }
    IdRow& r = it->second;
    if (s < r.start) r.start = s;
    if (e > r.end) r.end = e;
    r.value += v;
    if (blk_off < r.first_block) { r.first_block = blk_off; r.chrom = chrom; }
}

// BED dialect of src/commands/depth.rs:449-495 and src/commands/coverage.rs:222-255.
// `line` includes its trailing '\n' when present.
inline bool parse_bed_line_dc(const uint8_t* line, size_t n, const std::unordered_map<std::string, uint32_t>& seqid_map, Region& out) {
    if (n == 0 || line[0] == '#') return false;
    std::string_view f[3];
    int nf = 0;
    size_t i = 0;
    while (i < n && nf < 3) {
        while (i < n && (line[i] == '\t' || line[i] == ' ')) i++;
        if (i >= n) break;
        size_t s = i;
        while (i < n && line[i] != '\t' && line[i] != ' ') i++;
        f[nf++] = std::string_view((const char*)line + s, i - s);
    }
    if (nf < 3) return false;
    for (int k = 0; k < 3; k++) if (!valid_utf8((const uint8_t*)f[k].data(), f[k].size())) return false;
    uint32_t s, e;
    if (!rust_parse_u32(f[1], s)) return false;
    if (!rust_parse_u32(trim_end(f[2]), e)) return false;
    if (s >= e) return false;
    auto it = seqid_map.find(std::string(f[0]));
    if (it == seqid_map.end()) return false;
    out = {it->second, s, e};
    return true;
}

inline std::vector<size_t> bed_line_offsets(const uint8_t* data, size_t n) {  // depth.rs:441-459
    std::vector<size_t> off;
    off.push_back(0);
    for (size_t i = 0; i < n; i++) if (data[i] == '\n') off.push_back(i + 1);
    if (off.back() != n) off.push_back(n);
    return off;
}

// ---------------------------------------------------------------------------------------------
// depth — src/commands/depth.rs
// ---------------------------------------------------------------------------------------------
struct RegionRef { uint32_t start, end; };
static constexpr size_t DEPTH_BATCH_SIZE = 100000;  // :24

inline bool overlaps(uint32_t a1, uint32_t a2, uint32_t b1, uint32_t b2) {  // :78-82
    return std::max(a1, b1) < std::min(a2, b2);
}

// src/commands/depth.rs:121-217
inline void compute_root_depth(const uint8_t* slice, size_t n, const std::vector<RegionRef>& regions, uint32_t bin_shift,
                               uint64_t blk_off, IdMap& out) {
    std::vector<std::string> id_strings, id_chrom;
    std::vector<FeatLine> feats;
    parse_block_lines(slice, n, id_strings, id_chrom, feats);
    if (feats.empty() || regions.empty()) return;
    if (bin_shift > 31) bin_shift = 31;
    uint32_t max_bin = 0;
    for (const auto& f : feats) max_bin = std::max(max_bin, (f.end - 1) >> bin_shift);
    std::vector<std::vector<uint32_t>> feat_bins((size_t)max_bin + 1);
    for (size_t i = 0; i < feats.size(); i++) {
        uint32_t b0 = feats[i].start >> bin_shift, b1 = (feats[i].end - 1) >> bin_shift;
        for (uint32_t b = b0; b <= b1; b++) feat_bins[b].push_back((uint32_t)i);
    }
    size_t n_ids = id_strings.size();
    std::vector<uint32_t> min_s(n_ids, UINT32_MAX), max_e(n_ids, 0);
    std::vector<uint64_t> depths(n_ids, 0);
    std::vector<uint32_t> cand, hit_ids;
    for (const auto& r : regions) {
        cand.clear();
        uint32_t rb0 = r.start >> bin_shift, rb1 = (r.end == 0 ? 0 : r.end - 1) >> bin_shift;
        for (uint64_t b = rb0; b <= rb1 && b < feat_bins.size(); b++) cand.insert(cand.end(), feat_bins[b].begin(), feat_bins[b].end());
        std::sort(cand.begin(), cand.end());
        cand.erase(std::unique(cand.begin(), cand.end()), cand.end());
        hit_ids.clear();
        for (uint32_t fi : cand) {
            const FeatLine& f = feats[fi];
            if (overlaps(f.start, f.end, r.start, r.end)) {
                hit_ids.push_back(f.id_idx);
                if (f.start < min_s[f.id_idx]) min_s[f.id_idx] = f.start;
                if (f.end > max_e[f.id_idx]) max_e[f.id_idx] = f.end;
            }
        }
        std::sort(hit_ids.begin(), hit_ids.end());
        hit_ids.erase(std::unique(hit_ids.begin(), hit_ids.end()), hit_ids.end());
        for (uint32_t ii : hit_ids) depths[ii]++;
    }
    for (size_t i = 0; i < n_ids; i++)
        if (depths[i] > 0) merge_row(out, id_strings[i], id_chrom[i], min_s[i] == UINT32_MAX ? 0 : min_s[i], max_e[i], depths[i], blk_off);
}

// src/commands/depth.rs:222-293
inline void compute_hit_depth(const TreeIndexData& idx, const std::vector<Region>& regions, const GofMap& gof,
                              const uint8_t* gff, size_t gff_len, uint32_t bin_shift, IdMap& global) {
    std::unordered_map<uint32_t, std::vector<RegionRef>> by_root;
    std::vector<const Interval*> hits;
    std::vector<uint32_t> seen;
    for (const auto& r : regions) {
        auto t = idx.chr_entries.find(r.chr);
        if (t == idx.chr_entries.end()) continue;
        hits.clear();
        t->second.query_interval(r.start, r.end, hits);
        seen.clear();
        for (const Interval* h : hits) {
            if (std::find(seen.begin(), seen.end(), h->root_fid) != seen.end()) continue;
            seen.push_back(h->root_fid);
            auto g = gof.index.find(h->root_fid);
            if (g == gof.index.end()) continue;
            auto [s_off, e_off] = g->second;
            if (s_off == MISSING || e_off == MISSING || e_off <= s_off) continue;
            by_root[h->root_fid].push_back({r.start, r.end});
        }
    }
    if (by_root.empty()) return;
    std::vector<std::pair<uint32_t, const std::vector<RegionRef>*>> roots;
    for (const auto& kv : by_root) roots.emplace_back(kv.first, &kv.second);
    std::vector<IdMap> partials(roots.size());
#pragma omp parallel for schedule(dynamic, 4)
    for (long i = 0; i < (long)roots.size(); i++) {  // par_bridge over roots (:256-261)
        auto [s_off, e_off] = gof.index.at(roots[i].first);
        if (e_off > gff_len) throw Error("gof block beyond GFF length");  // Rust slice index would panic
        compute_root_depth(gff + s_off, (size_t)(e_off - s_off), *roots[i].second, bin_shift, s_off, partials[i]);
    }
    for (auto& m : partials)
        for (auto& kv : m) merge_row(global, kv.first, kv.second.chrom, kv.second.start, kv.second.end, kv.second.value, kv.second.first_block);
}

// src/commands/depth.rs:429-512
inline IdMap depth_process_bed(const std::string& bed, const TreeIndexData& idx, const GofMap& gof, const uint8_t* gff,
                               size_t gff_len, uint32_t bin_shift) {
    IdMap global;
    MappedFile f(bed);
    std::vector<size_t> off = bed_line_offsets(f.data, f.len);
    size_t n_lines = off.size() - 1;
    for (size_t b0 = 0; b0 < n_lines; b0 += DEPTH_BATCH_SIZE) {
        size_t b1 = std::min(n_lines, b0 + DEPTH_BATCH_SIZE);
        std::vector<Region> regions;
        std::vector<char> ok(b1 - b0, 0);
        std::vector<Region> tmp(b1 - b0);
#pragma omp parallel for schedule(static)
        for (long i = (long)b0; i < (long)b1; i++) {  // par_iter parse (:464-495)
            if (off[i] >= off[i + 1]) continue;
            ok[i - b0] = parse_bed_line_dc(f.data + off[i], off[i + 1] - off[i], idx.seqid_to_num, tmp[i - b0]);
        }
        for (size_t i = 0; i < ok.size(); i++) if (ok[i]) regions.push_back(tmp[i]);
        compute_hit_depth(idx, regions, gof, gff, gff_len, bin_shift, global);
    }
    return global;
}

// In-memory variant (regions already parsed & validated) with the same 100k batching.
inline IdMap depth_from_regions(const std::vector<Region>& regions, const TreeIndexData& idx, const GofMap& gof,
                                const uint8_t* gff, size_t gff_len, uint32_t bin_shift) {
    IdMap global;
    for (size_t b0 = 0; b0 < regions.size(); b0 += DEPTH_BATCH_SIZE) {
        size_t b1 = std::min(regions.size(), b0 + DEPTH_BATCH_SIZE);
        std::vector<Region> batch;
        for (size_t i = b0; i < b1; i++) if (regions[i].start < regions[i].end) batch.push_back(regions[i]);
        compute_hit_depth(idx, batch, gof, gff, gff_len, bin_shift, global);
    }
    return global;
}

// src/commands/depth.rs:515-545 — row order is hash order in the reference; here as iterated.
inline void write_depth_results(const IdMap& m, FILE* out) {
    fputs("id\tchr\tstart\tend\tdepth\n", out);
    for (const auto& kv : m)
        fprintf(out, "%s\t%s\t%u\t%u\t%llu\n", kv.first.c_str(), kv.second.chrom.c_str(), kv.second.start, kv.second.end,
                (unsigned long long)kv.second.value);
}

// ---------------------------------------------------------------------------------------------
// coverage — src/commands/coverage.rs
// ---------------------------------------------------------------------------------------------
using IvVec = std::vector<std::pair<uint32_t, uint32_t>>;

inline IvVec merge_intervals(IvVec ivs) {  // :92-107
    if (ivs.empty()) return ivs;
    std::sort(ivs.begin(), ivs.end(), [](auto& a, auto& b) { return a.first < b.first; });
    IvVec out;
    uint32_t cs = ivs[0].first, ce = ivs[0].second;
    for (size_t i = 1; i < ivs.size(); i++) {
        auto [s, e] = ivs[i];
        if (s <= ce) { if (e > ce) ce = e; }
        else { out.emplace_back(cs, ce); cs = s; ce = e; }
    }
    out.emplace_back(cs, ce);
    return out;
}
inline uint64_t union_len(IvVec ivs) {  // :110-121
    if (ivs.empty()) return 0;
    std::sort(ivs.begin(), ivs.end(), [](auto& a, auto& b) { return a.first < b.first; });
    uint64_t total = 0;
    uint32_t cs = ivs[0].first, ce = ivs[0].second;
    for (size_t i = 1; i < ivs.size(); i++) {
        auto [s, e] = ivs[i];
        if (s <= ce) ce = std::max(ce, e);
        else { total += ce - cs; cs = s; ce = e; }
    }
    total += ce - cs;
    return total;
}

// src/commands/coverage.rs:277-377
inline void compute_breadth_for_root(const uint8_t* slice, size_t n, const IvVec& cov, uint64_t blk_off, IdMap& out) {
    std::vector<std::string> id_strings, id_chrom;
    std::vector<FeatLine> lines;
    parse_block_lines(slice, n, id_strings, id_chrom, lines);
    if (lines.empty() || cov.empty()) return;
    std::stable_sort(lines.begin(), lines.end(), [](const FeatLine& a, const FeatLine& b) { return a.start < b.start; });
    size_t n_ids = id_strings.size();
    std::vector<uint32_t> min_s(n_ids, UINT32_MAX), max_e(n_ids, 0);
    std::vector<IvVec> id_overlap(n_ids);
    size_t j = 0;
    for (const auto& fl : lines) {
        while (j < cov.size() && cov[j].second <= fl.start) j++;
        if (fl.start < min_s[fl.id_idx]) min_s[fl.id_idx] = fl.start;
        if (fl.end > max_e[fl.id_idx]) max_e[fl.id_idx] = fl.end;
        size_t k = j;
        while (k < cov.size() && cov[k].first < fl.end) {
            uint32_t s = std::max(fl.start, cov[k].first), e = std::min(fl.end, cov[k].second);
            if (e > s) id_overlap[fl.id_idx].emplace_back(s, e);
            if (cov[k].second <= fl.end) k++; else break;
        }
    }
    for (size_t i = 0; i < n_ids; i++) {
        uint64_t length = max_e[i] > min_s[i] ? max_e[i] - min_s[i] : 0;
        uint64_t breadth = union_len(std::move(id_overlap[i]));
        if (length > 0 || breadth > 0) merge_row(out, id_strings[i], id_chrom[i], min_s[i], max_e[i], breadth, blk_off);
    }
}

using ByRoot = std::unordered_map<uint32_t, IvVec>;

// routing loop of src/commands/coverage.rs:257-266
inline void coverage_route(const TreeIndexData& idx, const Region& r, ByRoot& by_root, std::vector<const Interval*>& hits,
                           std::vector<uint32_t>& seen) {
    auto t = idx.chr_entries.find(r.chr);
    if (t == idx.chr_entries.end()) return;
    hits.clear();
    t->second.query_interval(r.start, r.end, hits);
    seen.clear();
    for (const Interval* h : hits) {
        if (std::find(seen.begin(), seen.end(), h->root_fid) != seen.end()) continue;
        seen.push_back(h->root_fid);
        by_root[h->root_fid].emplace_back(r.start, r.end);
    }
}

// src/commands/coverage.rs:208-274
inline ByRoot collect_by_root_from_bed(const std::string& bed, const TreeIndexData& idx) {
    MappedFile f(bed);
    std::vector<size_t> off = bed_line_offsets(f.data, f.len);
    ByRoot by_root;
    std::vector<const Interval*> hits;
    std::vector<uint32_t> seen;
    for (size_t i = 0; i + 1 < off.size(); i++) {
        if (off[i] >= off[i + 1]) continue;
        Region r;
        if (!parse_bed_line_dc(f.data + off[i], off[i + 1] - off[i], idx.seqid_to_num, r)) continue;
        coverage_route(idx, r, by_root, hits, seen);
    }
    return by_root;
}

// src/commands/coverage.rs:383-450
inline IdMap finalize_compute_breadth(ByRoot& by_root, const GofMap& gof, const uint8_t* gff, size_t gff_len) {
    IdMap global;
    if (by_root.empty()) return global;
    std::vector<std::pair<uint32_t, IvVec*>> roots;
    for (auto& kv : by_root) roots.emplace_back(kv.first, &kv.second);
    std::vector<IdMap> partials(roots.size());
#pragma omp parallel for schedule(dynamic, 4)
    for (long i = 0; i < (long)roots.size(); i++) {
        IvVec cov = merge_intervals(std::move(*roots[i].second));
        auto g = gof.index.find(roots[i].first);
        if (g == gof.index.end()) continue;
        auto [s_off, e_off] = g->second;
        if (s_off == MISSING || e_off == MISSING || e_off <= s_off) continue;
        if (e_off > gff_len) throw Error("gof block beyond GFF length");
        compute_breadth_for_root(gff + s_off, (size_t)(e_off - s_off), cov, s_off, partials[i]);
    }
    for (auto& m : partials)
        for (auto& kv : m) merge_row(global, kv.first, kv.second.chrom, kv.second.start, kv.second.end, kv.second.value, kv.second.first_block);
    return global;
}

inline IdMap coverage_from_regions(const std::vector<Region>& regions, const TreeIndexData& idx, const GofMap& gof,
                                   const uint8_t* gff, size_t gff_len) {
    ByRoot by_root;
    std::vector<const Interval*> hits;
    std::vector<uint32_t> seen;
    for (const auto& r : regions) if (r.start < r.end) coverage_route(idx, r, by_root, hits, seen);
    return finalize_compute_breadth(by_root, gof, gff, gff_len);
}

// src/commands/coverage.rs:453-490
inline void write_breadth_results(const IdMap& m, FILE* out) {
    fputs("id\tchr\tstart\tend\tbreadth\tfraction\n", out);
    for (const auto& kv : m) {
        const IdRow& r = kv.second;
        uint64_t length = r.end > r.start ? r.end - r.start : 0;
        double fraction = length > 0 ? (double)r.value / (double)length : 0.0;
        fprintf(out, "%s\t%s\t%u\t%u\t%llu\t%.6f\n", kv.first.c_str(), r.chrom.c_str(), r.start, r.end,
                (unsigned long long)r.value, fraction);
    }
}

// ---------------------------------------------------------------------------------------------
// index builder (fixture writer) — src/index_builder/core.rs:41-242, src/utils/tree_io.rs:37-63
// Writes .sqs .gof .rit .rix .fts .prt (and an empty-valued .a2f/.atn pair) beside `gff`.
// ---------------------------------------------------------------------------------------------
inline bool is_regex_space_at(const uint8_t* s, size_t n, size_t i) { return ws_prefix_len(s + i, n - i) != 0; }

// first leftmost match of `KEY=([^;\s]+)` in line
inline bool regex_attr(const uint8_t* line, size_t n, const char* key, std::string& val) {
    size_t kl = strlen(key);
    for (size_t i = 0; i + kl <= n; i++) {
        if (memcmp(line + i, key, kl) != 0) continue;
        size_t j = i + kl;
        size_t st = j;
        while (j < n && line[j] != ';' && !is_regex_space_at(line, n, j)) j++;
        if (j > st) { val.assign((const char*)line + st, j - st); return true; }
    }
    return false;
}

inline void build_index(const std::string& gff, const std::string& skip_types =
                            "remark,note,comment,region,gap,assembly_gap,contig,scaffold,source") {
    std::unordered_set<std::string> skip;
    {
        size_t i = 0;
        while (i <= skip_types.size()) {
            size_t j = skip_types.find(',', i);
            if (j == std::string::npos) j = skip_types.size();
            skip.insert(skip_types.substr(i, j - i));
            i = j + 1;
        }
    }
    MappedFile f(gff);
    struct Raw { std::string seqid; uint32_t start, end; uint64_t line_offset; std::string id; bool has_parent; std::string parent; };
    std::vector<Raw> raw;
    size_t offset = 0;
    while (offset < f.len) {
        const uint8_t* p = (const uint8_t*)memchr(f.data + offset, '\n', f.len - offset);
        size_t nl = p ? (size_t)(p - f.data) : f.len;
        const uint8_t* lb = f.data + offset;
        size_t ln = nl - offset;
        uint64_t line_offset = offset;
        offset = nl + 1;
        if (ln == 0 || lb[0] == '#') continue;
        if (!valid_utf8(lb, ln)) throw Error("invalid utf-8 in GFF");
        std::string_view line = trim(std::string_view((const char*)lb, ln));
        if (line.empty()) continue;
        std::vector<std::string_view> fields;
        {
            size_t i = 0;
            while (true) {
                size_t j = line.find('\t', i);
                if (j == std::string_view::npos) { fields.push_back(line.substr(i)); break; }
                fields.push_back(line.substr(i, j - i));
                i = j + 1;
            }
        }
        if (fields.size() != 9) throw Error("Invalid GFF line (expected 9 columns): " + std::string(line));
        if (skip.count(std::string(fields[2]))) continue;
        uint32_t s1, e1;
        if (!rust_parse_u32(fields[3], s1) || !rust_parse_u32(fields[4], e1)) throw Error("invalid digit found in string");
        if (e1 == 0) continue;
        if (s1 > e1) std::swap(s1, e1);
        Raw r;
        r.seqid = std::string(fields[0]);
        r.start = s1 == 0 ? 0 : s1 - 1;
        r.end = e1;
        r.line_offset = line_offset;
        if (!regex_attr((const uint8_t*)line.data(), line.size(), "ID=", r.id)) throw Error("Missing ID in feature: " + std::string(line));
        r.has_parent = regex_attr((const uint8_t*)line.data(), line.size(), "Parent=", r.parent);
        raw.push_back(std::move(r));
    }
    std::unordered_map<std::string, uint32_t> feature_map;
    for (size_t i = 0; i < raw.size(); i++) feature_map[raw[i].id] = (uint32_t)i;  // last line with the ID wins (:141-144)

    std::string fts, gofb;
    std::vector<uint32_t> prt;
    std::vector<std::string> seq_names;
    std::unordered_map<std::string, uint32_t> seq_num;
    std::vector<std::vector<Interval>> trees_input;
    bool have_root = false;
    uint32_t cur_id = 0, cur_seq = 0;
    uint64_t cur_off = 0;
    auto put_gof = [&](uint32_t id, uint32_t sq, uint64_t s, uint64_t e) {
        gofb.append((const char*)&id, 4); gofb.append((const char*)&sq, 4);
        gofb.append((const char*)&s, 8); gofb.append((const char*)&e, 8);
    };
    for (const auto& rf : raw) {
        uint32_t fid = feature_map[rf.id];
        fts += rf.id; fts += '\n';
        uint32_t parent_id = fid;
        if (rf.has_parent) { auto it = feature_map.find(rf.parent); if (it != feature_map.end()) parent_id = it->second; }
        prt.push_back(parent_id);
        if (parent_id == fid) {
            auto it = seq_num.find(rf.seqid);
            uint32_t sn;
            if (it == seq_num.end()) { sn = (uint32_t)seq_names.size(); seq_num[rf.seqid] = sn; seq_names.push_back(rf.seqid); trees_input.emplace_back(); }
            else sn = it->second;
            trees_input[sn].push_back({rf.start, rf.end, fid});
            if (have_root) put_gof(cur_id, cur_seq, cur_off, rf.line_offset);
            have_root = true; cur_id = fid; cur_off = rf.line_offset; cur_seq = sn;
        }
    }
    if (have_root) put_gof(cur_id, cur_seq, cur_off, f.len);

    std::string rit, rix = "[";
    for (size_t i = 0; i < trees_input.size(); i++) {
        if (i) rix += ",";
        rix += std::to_string(rit.size());
        IntervalTree t = IntervalTree::from_intervals(trees_input[i]);
        encode_opt_node(t.root.get(), rit);
    }
    rix += "]";
    std::string sqs;
    for (const auto& s : seq_names) { sqs += s; sqs += '\n'; }
    auto dump = [&](const std::string& suffix, const std::string& bytes) {
        std::ofstream o(append_suffix(gff, suffix), std::ios::binary);
        if (!o) throw Error("cannot write " + gff + suffix);
        o.write(bytes.data(), (std::streamsize)bytes.size());
    };
    dump(".fts", fts);
    dump(".gof", gofb);
    dump(".rit", rit);
    dump(".rix", rix);
    dump(".sqs", sqs);
    dump(".prt", std::string((const char*)prt.data(), prt.size() * 4));
    std::vector<uint32_t> a2f(raw.size(), UINT32_MAX);
    dump(".a2f", std::string((const char*)a2f.data(), a2f.size() * 4));
    dump(".atn", "#attribute=gene_name\n");
}

}  // namespace gffx_oracle
