// linecache.cpp — the side-car cache of the GFF line table: `<gff>.b200cache`.
//
// The reference re-parses the text of every selected block on every run (src/commands/intersect.rs:284-321,
// src/commands/depth.rs:131-157, src/commands/coverage.rs:296-323); this engine scans the text once per index open into
// the line table (host_index.cpp: build_lines, or the device scanner linescan.cu) — and, with this file, once per GFF: the
// table, the type / chromosome / ID dictionaries and the per-block facts are written next to the GFF after the first
// open and read back by every later one.  The eight index files of `gffx index` are untouched.
//
// The cache is only ever an accelerator: it is keyed by the GFF's size and modification time and by a digest of
// everything else the table depends on (block list, sequence names), any mismatch or short read means "no cache", and a
// cache that cannot be written (read-only directory) is skipped silently.  GFFX_CACHE=0 turns it off; GFFX_CACHE_DIR puts
// the files elsewhere (tests).
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <cstdlib>
#include <thread>

#include "host_index.h"

namespace gffx {

namespace {

constexpr char MAGIC[8] = {'G', 'F', 'X', 'B', '2', '0', '0', 'C'};
constexpr uint32_t VERSION = 2;

struct Header {
    char magic[8];
    uint32_t version, n_types, n_chroms, pad;
    uint64_t gff_len, gff_mtime_ns, key, n_lines, n_blocks, n_ids, types_bytes, chroms_bytes;
};

uint64_t fnv(uint64_t h, const void* p, size_t n) {
    const uint8_t* b = (const uint8_t*)p;
    for (size_t i = 0; i < n; i++) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}

// everything besides the text that the line table is a function of
uint64_t cache_key(const HostIndex& H) {
    uint64_t h = 1469598103934665603ull;
    for (const Block& b : H.blocks) { h = fnv(h, &b.fid, 4); h = fnv(h, &b.start, 8); h = fnv(h, &b.end, 8); }
    for (const std::string& s : H.seq_names) { h = fnv(h, s.data(), s.size()); h = fnv(h, "\n", 1); }
    return h;
}

std::string cache_path(const HostIndex& H) {
    if (const char* d = getenv("GFFX_CACHE_DIR")) {
        std::string base = H.gff_path;
        const size_t s = base.rfind('/');
        if (s != std::string::npos) base = base.substr(s + 1);
        return std::string(d) + "/" + base + ".b200cache";
    }
    return H.gff_path + ".b200cache";
}

bool enabled(const HostIndex& H) {
    if (H.gff_path.empty() || !H.gff || H.gff != H.gff_map.data()) return false;   // an index made from a caller's buffer has no file to key on
    const char* e = getenv("GFFX_CACHE");
    return !(e && atoi(e) == 0);
}

bool gff_stat(const HostIndex& H, uint64_t* len, uint64_t* mtime_ns) {
    struct stat st;
    if (stat(H.gff_path.c_str(), &st) != 0) return false;
    *len = (uint64_t)st.st_size;
    *mtime_ns = (uint64_t)st.st_mtim.tv_sec * 1000000000ull + (uint64_t)st.st_mtim.tv_nsec;
    return true;
}

struct Piece { void* p; size_t n; };

template <class T> Piece piece(std::vector<T>& v) { return Piece{v.data(), v.size() * sizeof(T)}; }

std::string pool_bytes(const StringPool& P) {   // u32 length + bytes per string, in index order
    std::string s;
    for (uint32_t i = 0; i < P.size(); i++) {
        const std::string_view v = P.get(i);
        const uint32_t l = (uint32_t)v.size();
        s.append((const char*)&l, 4);
        s.append(v.data(), v.size());
    }
    return s;
}

bool pool_from(const std::string& s, uint32_t n, StringPool* P) {
    size_t o = 0;
    for (uint32_t i = 0; i < n; i++) {
        if (o + 4 > s.size()) return false;
        uint32_t l;
        memcpy(&l, s.data() + o, 4);
        o += 4;
        if (o + l > s.size()) return false;
        if (P->intern(std::string_view(s.data() + o, l)) != i) return false;
        o += l;
    }
    return o == s.size();
}

}  // namespace

bool line_cache_load(HostIndex& H, int threads) {
    if (!enabled(H)) return false;
    PhaseTimer t("  GFF line table (side-car cache)");
    uint64_t len, mtime;
    if (!gff_stat(H, &len, &mtime) || len != H.gff_len) return false;
    const int fd = ::open(cache_path(H).c_str(), O_RDONLY);
    if (fd < 0) return false;
    Header h;
    bool ok = pread(fd, &h, sizeof h, 0) == (ssize_t)sizeof h && memcmp(h.magic, MAGIC, 8) == 0 && h.version == VERSION && h.gff_len == len &&
              h.gff_mtime_ns == mtime && h.key == cache_key(H) && h.n_blocks == H.blocks.size() && h.n_lines < (1ull << 32) - 1;
    struct stat st;
    LineTable L;
    IdTable ids;
    std::vector<uint8_t> blk_utf8;
    std::vector<uint64_t> blk_line0;
    std::string tb, cb;
    std::vector<Piece> pcs;
    if (ok) {
        const size_t n = (size_t)h.n_lines, nb = (size_t)h.n_blocks;
        L.off.resize(n); L.len.resize(n); L.blk.resize(n); L.flags.resize(n); L.S.resize(n); L.E.resize(n); L.iseq.resize(n); L.type_id.resize(n);
        L.fs.resize(n); L.fe.resize(n); L.id.resize(n); L.chrom.resize(n);
        ids.off.resize((size_t)h.n_ids); ids.len.resize((size_t)h.n_ids);
        blk_utf8.resize(nb); blk_line0.resize(nb + 1);
        tb.resize((size_t)h.types_bytes); cb.resize((size_t)h.chroms_bytes);
        pcs = {piece(L.off), piece(L.len), piece(L.blk), piece(L.flags), piece(L.S), piece(L.E), piece(L.iseq), piece(L.type_id), piece(L.fs), piece(L.fe),
               piece(L.id), piece(L.chrom), piece(ids.off), piece(ids.len), piece(blk_utf8), piece(blk_line0), Piece{tb.data(), tb.size()},
               Piece{cb.data(), cb.size()}};
        size_t total = sizeof h;
        for (const Piece& p : pcs) total += p.n;
        ok = fstat(fd, &st) == 0 && (size_t)st.st_size == total;
    }
    if (ok) {   // the pieces are read side by side: the page cache serves several readers faster than one
        std::vector<size_t> at(pcs.size());
        size_t o = sizeof h;
        for (size_t i = 0; i < pcs.size(); i++) { at[i] = o; o += pcs[i].n; }
        std::vector<uint8_t> good(pcs.size(), 1);
        // the large pieces in slices of 4 MB, so that every thread has its share of the bytes
        struct Slice { size_t i, a, z; };
        std::vector<Slice> slices;
        for (size_t i = 0; i < pcs.size(); i++)
            for (size_t a = 0; a < pcs[i].n; a += 4u << 20) slices.push_back(Slice{i, a, std::min(pcs[i].n, a + (4u << 20))});
        parallel_jobs(slices.size(), std::max(1, std::min(threads, 16)), [&](size_t k) {
            const Slice& sl = slices[k];
            size_t done = sl.a;
            while (done < sl.z) {
                const ssize_t r = pread(fd, (char*)pcs[sl.i].p + done, sl.z - done, (off_t)(at[sl.i] + done));
                if (r <= 0) { good[sl.i] = 0; break; }
                done += (size_t)r;
            }
        });
        for (uint8_t g : good) ok = ok && g;
    }
    ::close(fd);
    if (!ok) return false;
    StringPool types, chroms;
    if (!pool_from(tb, h.n_types, &types) || !pool_from(cb, h.n_chroms, &chroms)) return false;
    // cheap sanity: the per-block line counts add up and every dictionary index is in range at the table's ends
    if (blk_line0.empty() || blk_line0.back() != h.n_lines) return false;
    H.lines = std::move(L);
    H.ids = std::move(ids);
    H.ids.base = H.gff;
    H.blk_utf8 = std::move(blk_utf8);
    H.blk_line0 = std::move(blk_line0);
    H.types = std::move(types);
    H.chroms = std::move(chroms);
    H.lines_from_cache = true;
    return true;
}

std::string line_cache_target(const HostIndex& H) { return enabled(H) ? cache_path(H) : std::string(); }

void line_cache_store(const HostIndex& H, const std::string& path) {
    if (path.empty()) return;
    PhaseTimer t("  GFF line table (writing the side-car cache)");
    Header h{};
    memcpy(h.magic, MAGIC, 8);
    h.version = VERSION;
    if (!gff_stat(H, &h.gff_len, &h.gff_mtime_ns) || h.gff_len != H.gff_len) return;
    h.key = cache_key(H);
    const LineTable& L = H.lines;
    h.n_lines = L.size(); h.n_blocks = H.blocks.size(); h.n_ids = H.ids.size();
    h.n_types = (uint32_t)H.types.size(); h.n_chroms = (uint32_t)H.chroms.size();
    const std::string tb = pool_bytes(H.types), cb = pool_bytes(H.chroms);
    h.types_bytes = tb.size(); h.chroms_bytes = cb.size();
    const std::string tmp = path + ".tmp" + std::to_string((long)getpid());
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return;
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    auto put = [&](const void* p, size_t n) { ok = ok && (n == 0 || fwrite(p, 1, n, f) == n); };
    put(L.off.data(), L.off.size() * 8); put(L.len.data(), L.len.size() * 4); put(L.blk.data(), L.blk.size() * 4); put(L.flags.data(), L.flags.size());
    put(L.S.data(), L.S.size() * 4); put(L.E.data(), L.E.size() * 4); put(L.iseq.data(), L.iseq.size() * 4); put(L.type_id.data(), L.type_id.size() * 4);
    put(L.fs.data(), L.fs.size() * 4); put(L.fe.data(), L.fe.size() * 4); put(L.id.data(), L.id.size() * 4); put(L.chrom.data(), L.chrom.size() * 4);
    put(H.ids.off.data(), H.ids.off.size() * 8); put(H.ids.len.data(), H.ids.len.size() * 4);
    put(H.blk_utf8.data(), H.blk_utf8.size()); put(H.blk_line0.data(), H.blk_line0.size() * 8);
    put(tb.data(), tb.size()); put(cb.data(), cb.size());
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path.c_str()) != 0) unlink(tmp.c_str());
}

// FNV-1a digest of the line table, the dictionaries and the per-block facts: equal digests = equal tables (tests compare the
// host scanner's, the device scanner's and the cache's)
uint64_t line_table_digest(const HostIndex& H) {
    const LineTable& L = H.lines;
    uint64_t h = 1469598103934665603ull;
    auto add = [&](const void* p, size_t n) { h = fnv(h, p, n); h = fnv(h, &n, sizeof n); };
    add(L.off.data(), L.off.size() * 8); add(L.len.data(), L.len.size() * 4); add(L.blk.data(), L.blk.size() * 4); add(L.flags.data(), L.flags.size());
    add(L.S.data(), L.S.size() * 4); add(L.E.data(), L.E.size() * 4); add(L.iseq.data(), L.iseq.size() * 4); add(L.type_id.data(), L.type_id.size() * 4);
    add(L.fs.data(), L.fs.size() * 4); add(L.fe.data(), L.fe.size() * 4); add(L.id.data(), L.id.size() * 4); add(L.chrom.data(), L.chrom.size() * 4);
    add(H.ids.off.data(), H.ids.off.size() * 8); add(H.ids.len.data(), H.ids.len.size() * 4);
    add(H.blk_utf8.data(), H.blk_utf8.size()); add(H.blk_line0.data(), H.blk_line0.size() * 8);
    const std::string tb = pool_bytes(H.types), cb = pool_bytes(H.chroms);
    add(tb.data(), tb.size()); add(cb.data(), cb.size());
    return h;
}

}  // namespace gffx
