// host_index.h — host-side model of a GFFx index: the on-disk files (.sqs .rix .rit .gof, read
// unchanged) plus the line table derived from the GFF text.  This is the input of the device
// upload in engine.cu.  Reference formats: SURVEY.md App. B; src/utils/tree_io.rs:37-63,
// src/index_loader/{core.rs:19-34,gof.rs:95-128}, src/utils/tree_index.rs:36-82.
#pragma once
#include <thread>
#include <memory>
#include <unordered_map>
#include "util.h"

namespace gffx {

struct RootIv {  // one Interval<u32> of src/utils/tree.rs:6-10, tagged with the tree (seq) it sits in
    uint32_t seq, start, end, fid;
};

struct Block {  // one reachable .gof record (GofMap::index keeps the LAST record of a fid, gof.rs:32-37)
    uint32_t fid;
    uint64_t start, end;
};

// bits of LineTable::flags
enum : uint8_t {
    LN_I_OK = 1,   // columns 1..5 usable by gff_line_overlaps_queries (intersect.rs:441-489)
    LN_D_OK = 2,   // line contributes to depth/coverage (depth.rs:131-157 / coverage.rs:296-323)
};

// Struct-of-arrays over every non-empty, non-'#' line that lies inside a reachable block, in
// output order (blocks by ascending start offset, lines in file order).
struct LineTable {
    std::vector<uint64_t> off;     // absolute byte offset of the line
    std::vector<uint32_t> len;     // byte length including the '\n' when present
    std::vector<uint32_t> blk;     // index into HostIndex::blocks
    std::vector<uint8_t> flags;
    // intersect view: RAW columns 4,5 (no swap, no -1) and column 1 resolved by NAME against .sqs
    std::vector<uint32_t> S, E, iseq, type_id;
    // depth/coverage view: swapped, 0-based half-open; ID dictionary index; column-1 dictionary index
    std::vector<uint32_t> fs, fe, id, chrom;
    size_t size() const { return off.size(); }
};

struct StringPool {  // append-only dictionary: string -> dense index
    std::vector<char> bytes;
    std::vector<uint64_t> start;  // start[i]..start[i+1]
    std::unordered_map<std::string_view, uint32_t> map;  // views into `keys` (stable storage)
    std::vector<std::unique_ptr<std::string>> keys;
    uint32_t intern(std::string_view s);
    int64_t find(std::string_view s) const;
    std::string_view get(uint32_t i) const { return std::string_view(keys[i]->data(), keys[i]->size()); }
    size_t size() const { return keys.size(); }
};

// ID dictionary of the depth / coverage view: index = order of first appearance in the GFF; the strings stay where they are,
// in the GFF text (the mapped file, or the index's own copy of the buffer gffx_index_create was given)
struct IdTable {
    const uint8_t* base = nullptr;
    std::vector<uint64_t> off;
    std::vector<uint32_t> len;
    size_t size() const { return off.size(); }
    std::string_view get(uint32_t i) const { return std::string_view((const char*)base + off[i], len[i]); }
};

// Virtual sequences (the engine's coordinate system).  The partition pass and every kernel after it address at most
// a few hundred windows; the reference has no limit on the number of sequences (one tree per .rix offset,
// src/utils/tree_index.rs:36-82: benchmark/hs.gff.sqs has 583, a scaffold-level assembly thousands).  So the real
// sequences are laid side by side on a few VIRTUAL sequences: sequence c occupies [base, base + clamp] of virtual
// sequence vseq, clamp = ext + 1 where ext is the largest coordinate any region of c is ever compared with (root
// starts / ends of c, raw columns 4/5 of the lines named c, 0-based ends of the ID lines in blocks whose root is on c).
// A region (c, s, e) becomes (vseq, base + min(s, clamp), base + min(e, clamp)): the clamp sits above every coordinate
// of c and below the next sequence's base, so every comparison the reference makes (tree.rs:110, intersect.rs:145-158,
// :500-517, depth.rs:78-82, coverage.rs:349-368) has the same outcome, and regions of c can never meet anything of
// another sequence.  A sequence alone on its virtual sequence keeps its own coordinates (base 0, clamp = u32::MAX).
struct VSeqMap {
    struct Seq { uint32_t vseq, base, clamp; };
    std::vector<Seq> seq;              // per real sequence
    std::vector<uint32_t> v_extent;    // per virtual sequence: extent the window grid covers
    std::vector<uint8_t> v_canon;      // per virtual sequence: its members' names resolve to their own numbers (intersect.rs:622-625)
    std::vector<uint32_t> v_members;   // per virtual sequence: number of real sequences on it
    uint64_t capacity = 0;             // packing capacity that was used
    size_t n_vseqs() const { return v_extent.size(); }
};

// What a line scanner returns before the dictionaries are built (the device scanner, linescan.cu): the line table's plain
// columns, and per line where its type field, its column 1 and its ID value lie (offsets relative to the line; len NONE32 =
// none) with the ID's FNV-1a hash; per block its line count and whether it holds a byte >= 0x80 (such a block is re-scanned
// by the host scanner, which knows UTF-8).
struct RawLines {
    LineTable t;
    std::vector<uint32_t> type_off, type_len, chrom_len, id_off, id_len;
    std::vector<uint64_t> id_hash;
    std::vector<uint32_t> blk_nlines;
    std::vector<uint8_t> blk_nonascii;
    // the ID dictionary when the device built it (linescan.cu; every block ASCII, no hash collision): t.id is filled, these are
    // the dictionary's entries, and id_off / id_len / id_hash were not brought back
    // (the same for the type and chromosome dictionaries: t.type_id / t.chrom are filled, type_off / type_len / chrom_len stayed behind)
    bool ids_ready = false;
    std::vector<uint64_t> ids_off, types_off, chroms_off;
    std::vector<uint32_t> ids_len, types_len, chroms_len;
};

struct HostIndex {
    ~HostIndex() {
        if (cache_writer.joinable()) cache_writer.join();
        if (dev_text && dev_text_release) dev_text_release(dev_text, dev_text_device);
    }
    bool cache_pending = false;
    std::thread cache_writer;   // writes the side-car cache behind the open (the index is immutable once built)
    // the GFF text on the device, left there by the device scanner for the engine to adopt (record emission reads it)
    void* dev_text = nullptr;
    size_t dev_text_cap = 0;
    int dev_text_device = -1;
    void (*dev_text_release)(void*, int) = nullptr;
    std::string gff_path;
    std::vector<std::string> seq_names;                     // num_to_seqid
    std::unordered_map<std::string, uint32_t> seq_num;      // seqid_to_num (later duplicate wins)
    std::vector<RootIv> roots;                              // all tree intervals, grouped by seq, any order
    std::vector<Block> blocks;                              // reachable blocks sorted by start (stable)
    std::vector<uint32_t> root_blk;                         // per root: block index or NONE32 (fid absent from .gof)
    std::vector<uint8_t> blk_utf8;                          // per block: whole slice is valid UTF-8 (depth.rs:131)
    std::vector<uint64_t> blk_line0;                        // per block: first line index (blocks.size()+1 entries)
    Mapped gff_map;                                         // owned mapping when opened from a path
    std::vector<uint8_t> gff_copy;                          // owned copy when created from a caller's buffer (gffx_index_create)
    const uint8_t* gff = nullptr;
    uint64_t gff_len = 0;
    LineTable lines;
    StringPool types, chroms;
    IdTable ids;
    int n_threads = 1;   // host threads the loaders were given (the engine's load-time builders use as many)
    bool lines_from_cache = false;   // the line table came from the side-car cache (linecache.cpp)
    int lines_source = 0;            // 0 host scanner, 1 device scanner (linescan.cu), 2 side-car cache
    int scan_device = -1;            // CUDA device the line scanner may use (-1: host threads only)
    VSeqMap vmap;        // real sequence -> (virtual sequence, base, clamp); built by finish()

    // load .sqs/.rix/.rit/.gof beside `path` and (when the text is there) build the line table
    void open(const std::string& path, int threads);
    // from caller arrays (gffx_index_desc)
    void create(uint32_t n_seqs, const char* const* names, uint64_t n_roots, const uint32_t* rseq, const uint32_t* rs,
                const uint32_t* re, const uint32_t* rfid, uint64_t n_gof, const uint32_t* gfid, const uint64_t* gs,
                const uint64_t* ge, const uint8_t* text, uint64_t text_len, int threads);
    // true when block b may be scanned by the match-only path (intersect.rs:268-277)
    bool block_scannable(size_t b) const {
        const Block& k = blocks[b];
        return k.start != MISSING64 && k.start < std::min<uint64_t>(k.end, gff_len);
    }
    // true when block b is accepted by depth/coverage routing (depth.rs:243-244, coverage.rs:407)
    bool block_countable(size_t b) const {
        const Block& k = blocks[b];
        return k.start != MISSING64 && k.end != MISSING64 && k.end > k.start;
    }

   private:
    void finish(const std::vector<uint32_t>& gfid, const std::vector<uint64_t>& gs, const std::vector<uint64_t>& ge, int threads);
    void build_lines(int threads);
    void scan_lines(int threads);
    void scan_blocks(size_t b0, size_t b1, void* part) const;
    void lines_from_raw(struct RawLines& R, int threads);
    void number_ids(const std::vector<uint64_t>& g_hash, const std::vector<uint64_t>& g_off, const std::vector<uint32_t>& g_len, int threads);
    void build_vseqs();
};

// linecache.cpp: `<gff>.b200cache`, the line table of an earlier open (keyed by the GFF's size + mtime and a digest of the block
// list and sequence names); load returns false when there is no usable cache
bool line_cache_load(HostIndex& H, int threads);
std::string line_cache_target(const HostIndex& H);   // where the cache of this index goes ("": nowhere); reads the environment NOW
void line_cache_store(const HostIndex& H, const std::string& path);
uint64_t line_table_digest(const HostIndex& H);
// linescan.cu: the line table's columns scanned on CUDA device `device`; false when there is no such device
bool device_scan_lines(const HostIndex& H, int device, int threads, RawLines* out, void** d_text, size_t* d_text_cap);
void device_text_release(void* p, int device);

constexpr uint32_t VSEQ_MAX = 192;   // virtual sequences at most (each owns >= 1 of the <= 511 partition windows)

// .rit/.rix decoding: flatten every tree into RootIv records (seq = position in .rix)
void load_trees_flat(const std::string& rit_path, const std::string& rix_path, std::vector<RootIv>* out, uint32_t* n_trees);
void load_sqs_file(const std::string& path, std::vector<std::string>* names);
void load_gof_file(const std::string& path, std::vector<uint32_t>* fid, std::vector<uint64_t>* s, std::vector<uint64_t>* e);

// ---- BED parsing -------------------------------------------------------
struct Regions {
    std::vector<uint32_t> chr, start, end;
    size_t size() const { return chr.size(); }
};
// intersect dialect: src/commands/intersect.rs:201-230 — throws Err(E_FORMAT) where the reference aborts
void parse_bed_intersect(const std::string& path, const std::unordered_map<std::string, uint32_t>& seq_num, Regions* out);
// depth/coverage dialect: src/commands/depth.rs:449-495 == src/commands/coverage.rs:222-255
void parse_bed_counts(const std::string& path, const std::unordered_map<std::string, uint32_t>& seq_num, int threads, Regions* out);
// src/commands/intersect.rs:172-198
void parse_region_arg(const std::string& region, const std::unordered_map<std::string, uint32_t>& seq_num, uint32_t* chr,
                      uint32_t* start, uint32_t* end);

// Writes the byte ranges of `text` in order with writev batches (intersect.rs:385-429, common.rs:231-278)
void write_ranges_to(const uint8_t* text, uint64_t text_len, const uint64_t* ranges, uint64_t n, const char* out_path);
// the exact per-line deciders of the two BED dialects; the device parser (bedparse.cuh) hands them the
// lines it does not decide itself.  `chr` / `scratch` of the intersect form cache the previous line's name look-up
// (start with *chr = NONE32 - 1).
bool bed_intersect_line(const uint8_t* line, const uint8_t* q, const std::unordered_map<std::string, uint32_t>& seq_num, const std::string& path,
                        std::string& scratch, uint32_t* chr, uint32_t* s, uint32_t* e);
bool bed_counts_line(const uint8_t* line, const uint8_t* q, const std::unordered_map<std::string, uint32_t>& seq_num, std::string& scratch,
                     uint32_t* chr, uint32_t* s, uint32_t* e);
void write_text_to(const uint8_t* text, uint64_t n, const char* out_path);

}  // namespace gffx
