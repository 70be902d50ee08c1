// oracle_capi.cpp — C entry points onto the CPU oracle (gffx_oracle.hpp) for the ctypes test harness.
// TEST INFRASTRUCTURE ONLY: loaded by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs; never by the product library.
#include "gffx_oracle.hpp"
#include "reads_oracle.hpp"
#include <cstdlib>
#include <omp.h>

using namespace gffx_oracle;

namespace {
thread_local std::string g_err;
struct Handle {
    std::string gff_path;
    TreeIndexData idx;
    GofMap gof;
    MappedFile gff;
};
template <class F>
int guard(F&& f) {
    try { f(); return 0; }
    catch (const std::exception& e) { g_err = e.what(); return -1; }
    catch (...) { g_err = "unknown error"; return -1; }
}
std::vector<Region> to_regions(const uint32_t* chr, const uint32_t* s, const uint32_t* e, size_t n) {
    std::vector<Region> r(n);
    for (size_t i = 0; i < n; i++) r[i] = {chr[i], s[i], e[i]};
    return r;
}
char* dup_string(const std::string& s, uint64_t* len) {
    char* p = (char*)malloc(s.size() + 1);
    memcpy(p, s.data(), s.size());
    p[s.size()] = 0;
    if (len) *len = s.size();
    return p;
}
std::string rows_to_sorted_tsv(const IdMap& m, bool breadth) {
    std::vector<std::string> rows;
    rows.reserve(m.size());
    char buf[128];
    for (const auto& kv : m) {
        const IdRow& r = kv.second;
        std::string row = kv.first + "\t" + r.chrom + "\t";
        if (breadth) {
            uint64_t length = r.end > r.start ? r.end - r.start : 0;
            double fraction = length > 0 ? (double)r.value / (double)length : 0.0;
            snprintf(buf, sizeof buf, "%u\t%u\t%llu\t%.6f\n", r.start, r.end, (unsigned long long)r.value, fraction);
        } else {
            snprintf(buf, sizeof buf, "%u\t%u\t%llu\n", r.start, r.end, (unsigned long long)r.value);
        }
        row += buf;
        rows.push_back(std::move(row));
    }
    std::sort(rows.begin(), rows.end());
    std::string out = breadth ? "id\tchr\tstart\tend\tbreadth\tfraction\n" : "id\tchr\tstart\tend\tdepth\n";
    for (auto& r : rows) out += r;
    return out;
}
}  // namespace

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }
void orc_free(void* p) { free(p); }
void orc_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int orc_max_threads() { return omp_get_max_threads(); }

int orc_build_index(const char* gff_path) { return guard([&] { build_index(gff_path); }); }

int orc_open(const char* gff_path, void** out) {
    return guard([&] {
        auto h = std::make_unique<Handle>();
        h->gff_path = gff_path;
        h->idx = TreeIndexData::load_tree_index(gff_path);
        h->gof = load_gof(gff_path);
        // The real benchmark .gff files are git-LFS pointers; tolerate a missing GFF text so that the
        // root-level join can still be exercised on the reference's index files.
        try { h->gff.open(gff_path); } catch (...) {}
        *out = h.release();
    });
}
void orc_close(void* h) { delete (Handle*)h; }
uint32_t orc_n_seqs(void* h) { return (uint32_t)((Handle*)h)->idx.num_to_seqid.size(); }
const char* orc_seq_name(void* h, uint32_t i) { return ((Handle*)h)->idx.num_to_seqid[i].c_str(); }
uint64_t orc_n_roots(void* h) {
    uint64_t n = 0;
    for (const auto& kv : ((Handle*)h)->idx.chr_entries) n += kv.second.flatten().size();
    return n;
}

// Pin of IntervalTree::build (src/utils/tree.rs:30-64) + the bincode2 encoding (src/utils/tree_io.rs:37-54)
// against bytes the reference itself wrote: decode every tree of <gff>.rit, rebuild each tree from its flattened
// intervals with the restated constructor, encode, and return the concatenation — equal to the file iff both
// restatements reproduce the reference's.  (Ties in `start` keep their within-node order under the stable sort;
// across nodes the order cannot change node membership.)
int orc_reencode_rit(void* hv, uint8_t** out, uint64_t* n) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        std::string bytes;
        for (uint32_t i = 0; i < h->idx.num_to_seqid.size(); i++) {
            auto it = h->idx.chr_entries.find(i);
            if (it == h->idx.chr_entries.end()) { bytes.push_back('\0'); continue; }
            IntervalTree t = IntervalTree::from_intervals(it->second.flatten());
            encode_opt_node(t.root.get(), bytes);
        }
        uint8_t* p = (uint8_t*)malloc(std::max<size_t>(1, bytes.size()));
        memcpy(p, bytes.data(), bytes.size());
        *out = p; *n = bytes.size();
    });
}

// Flattened tree intervals of one sequence (for cross-checks of the product's own .rit decoder).
int orc_seq_intervals(void* hv, uint32_t seq, uint32_t** out, uint64_t* n) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        auto it = h->idx.chr_entries.find(seq);
        std::vector<Interval> v;
        if (it != h->idx.chr_entries.end()) v = it->second.flatten();
        uint32_t* p = (uint32_t*)malloc(std::max<size_t>(1, v.size()) * 12);
        for (size_t i = 0; i < v.size(); i++) { p[3 * i] = v[i].start; p[3 * i + 1] = v[i].end; p[3 * i + 2] = v[i].root_fid; }
        *out = p; *n = v.size();
    });
}

// BED dialect "intersect" (src/commands/intersect.rs:201-230)
int orc_parse_bed_intersect(void* hv, const char* bed, uint32_t** out, uint64_t* n) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        auto v = parse_bed_file(bed, h->idx.seqid_to_num);
        uint32_t* p = (uint32_t*)malloc(std::max<size_t>(1, v.size()) * 12);
        for (size_t i = 0; i < v.size(); i++) { p[3 * i] = v[i].chr; p[3 * i + 1] = v[i].start; p[3 * i + 2] = v[i].end; }
        *out = p; *n = v.size();
    });
}
// BED dialect "depth/coverage" (src/commands/depth.rs:449-495, src/commands/coverage.rs:222-255)
int orc_parse_bed_dc(void* hv, const char* bed, uint32_t** out, uint64_t* n) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        MappedFile f(bed);
        auto off = bed_line_offsets(f.data, f.len);
        std::vector<Region> v;
        for (size_t i = 0; i + 1 < off.size(); i++) {
            Region r;
            if (off[i] < off[i + 1] && parse_bed_line_dc(f.data + off[i], off[i + 1] - off[i], h->idx.seqid_to_num, r)) v.push_back(r);
        }
        uint32_t* p = (uint32_t*)malloc(std::max<size_t>(1, v.size()) * 12);
        for (size_t i = 0; i < v.size(); i++) { p[3 * i] = v[i].chr; p[3 * i + 1] = v[i].start; p[3 * i + 2] = v[i].end; }
        *out = p; *n = v.size();
    });
}

// read intervals of a SAM / BAM file (reads_oracle.hpp: process_bam's record rules, src/commands/depth.rs:335-366)
int orc_parse_reads(void* hv, const char* path, uint32_t** out, uint64_t* n) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        auto v = orc_reads::read_alignments(path, h->idx.seqid_to_num);
        uint32_t* p = (uint32_t*)malloc(std::max<size_t>(1, v.size()) * 12);
        for (size_t i = 0; i < v.size(); i++) { p[3 * i] = v[i].chr; p[3 * i + 1] = v[i].start; p[3 * i + 2] = v[i].end; }
        *out = p; *n = v.size();
    });
}

// query_features (src/commands/intersect.rs:105-169): pairs as (root_fid,start,end) triples.
int orc_query_features(void* hv, const uint32_t* chr, const uint32_t* s, const uint32_t* e, uint64_t n, int mode, int invert,
                       uint32_t** out, uint64_t* n_pairs) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        auto pairs = query_features(h->idx, to_regions(chr, s, e, n), (OverlapMode)mode, invert != 0);
        if (out) {
            uint32_t* p = (uint32_t*)malloc(std::max<size_t>(1, pairs.size()) * 12);
            memcpy(p, pairs.data(), pairs.size() * 12);
            *out = p;
        }
        *n_pairs = pairs.size();
    });
}

// Full intersect run() without the write: emitted byte ranges in output order.
int orc_intersect(void* hv, const uint32_t* chr, const uint32_t* s, const uint32_t* e, uint64_t n, int mode, int invert,
                  int entire_group, const char* types_csv, uint64_t** ranges, uint64_t* n_ranges, uint64_t* n_pairs,
                  uint64_t* n_warn) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        IntersectArgs a;
        a.contained = mode == 0; a.contains_region = mode == 1; a.overlap = mode == 2;
        a.invert = invert != 0; a.entire_group = entire_group != 0;
        a.has_types = types_csv != nullptr;
        if (types_csv) a.types = types_csv;
        auto r = intersect_compute(a, h->idx, h->gof, to_regions(chr, s, e, n), h->gff.data, h->gff.len);
        uint64_t* p = (uint64_t*)malloc(std::max<size_t>(1, r.ranges.size()) * 16);
        for (size_t i = 0; i < r.ranges.size(); i++) { p[2 * i] = r.ranges[i].first; p[2 * i + 1] = r.ranges[i].second; }
        *ranges = p; *n_ranges = r.ranges.size();
        if (n_pairs) *n_pairs = r.pairs.size();
        if (n_warn) *n_warn = r.warned_fids.size();
    });
}

// depth / coverage from already-parsed regions; result = header + rows sorted bytewise.
int orc_depth(void* hv, const uint32_t* chr, const uint32_t* s, const uint32_t* e, uint64_t n, uint32_t bin_shift, char** tsv,
              uint64_t* len) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        IdMap m = depth_from_regions(to_regions(chr, s, e, n), h->idx, h->gof, h->gff.data, h->gff.len, bin_shift);
        *tsv = dup_string(rows_to_sorted_tsv(m, false), len);
    });
}
int orc_coverage(void* hv, const uint32_t* chr, const uint32_t* s, const uint32_t* e, uint64_t n, char** tsv, uint64_t* len) {
    return guard([&] {
        Handle* h = (Handle*)hv;
        IdMap m = coverage_from_regions(to_regions(chr, s, e, n), h->idx, h->gof, h->gff.data, h->gff.len);
        *tsv = dup_string(rows_to_sorted_tsv(m, true), len);
    });
}

}  // extern "C"
