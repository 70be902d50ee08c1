// reads.cpp — alignment files as a source of read intervals for depth / coverage: SAM text, BAM (BGZF), bgzip-ed or gzip-ed
// SAM.  Replaces what the reference gets from rust-htslib in process_bam (src/commands/depth.rs:297-423) and
// collect_by_root_from_bam (src/commands/coverage.rs:125-205): per record (tid, pos, reference_end), with the skips of
// depth.rs:335-366 — unmapped, tid < 0, a target name the index does not know, pos < 0, end <= pos — and the clamp to u32.
//
// htslib is not in this image and its sources are not in /root/reference (a Cargo dependency, `rust-htslib`,
// Cargo.toml:32), so the record rules below restate its published behaviour:
//   * the container: BGZF = gzip members with a 'BC' extra field holding the member size (SAM spec §4.1); each member is
//     inflated on its own (inflate.h) and its CRC-32 and ISIZE are checked as htslib's bgzf reader does;
//   * BAM records (SAM spec §4.2): refID, pos, n_cigar_op, flag, l_seq, the CIGAR words; a CIGAR of more than 65535
//     operations sits in the CG:B,I tag behind a `<l_seq>S<n>N` placeholder (bam_tag2cigar);
//   * reference_end = bam_endpos: pos + the summed lengths of the M, D, N, = and X operations, + 1 when that sum is 0;
//   * SAM text (sam_parse1): tid = the position of RNAME among the @SQ lines ('*' or an unknown name: -1), pos = POS - 1;
//     a mapped record with POS 0 loses its tid; a record without tid gets the unmapped flag.
// The format is decided by the file's CONTENT, as hts_open does (the reference sends .bam, .sam and .cram to the same
// reader, depth.rs:589).  CRAM is not decoded: its codecs and reference-based sequences are out of this engine's scope.
#include "reads.h"

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <memory>
#include <thread>

#include "inflate.h"

namespace gffx {

namespace {

inline uint32_t le16(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8); }
inline uint32_t le32(const uint8_t* p) { return (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24); }

struct Sink {   // the record rules of depth.rs:335-366 / coverage.rs:160-178
    const std::vector<int64_t>* tid2num;   // -1: the index has no sequence of that name
    Regions* out;
    ReadStats* st;
    void record(int64_t tid, int64_t pos, uint32_t flag, uint64_t rlen) {
        st->records++;
        if (flag & 4u) { st->unmapped++; return; }                       // rec.is_unmapped()
        if (tid < 0) { st->unmapped++; return; }                         // tid < 0
        if ((uint64_t)tid >= tid2num->size()) throw Err(E_FORMAT, "alignment record names a reference beyond the header's target list");
        const int64_t chr = (*tid2num)[(size_t)tid];
        if (chr < 0) { st->unknown_ref++; return; }                      // tid2num[tid] is None
        if (pos < 0) { st->skipped++; return; }
        const int64_t end = pos + (int64_t)(rlen ? rlen : 1);            // bam_endpos
        if (end <= pos) { st->skipped++; return; }
        out->chr.push_back((uint32_t)chr);
        out->start.push_back((uint32_t)std::min<int64_t>(pos, 0xFFFFFFFFll));
        out->end.push_back((uint32_t)std::min<int64_t>(end, 0xFFFFFFFFll));
    }
};

inline bool cigar_consumes_ref(uint32_t op) { return op == 0 || op == 2 || op == 3 || op == 7 || op == 8; }   // M D N = X

// ---- BAM ------------------------------------------------------------------------------------------------------------------
// header: magic, l_text, text, n_ref, (l_name, name, l_ref)*; returns the bytes it took, 0 when `n` does not hold all of it yet
size_t bam_header(const uint8_t* p, size_t n, std::vector<std::string>* names) {
    if (n < 12) return 0;
    if (memcmp(p, "BAM\1", 4) != 0) throw Err(E_FORMAT, "not a BAM stream (bad magic)");
    const uint64_t l_text = le32(p + 4);
    size_t o = 8 + l_text;
    if (o + 4 > n) return 0;
    const uint32_t n_ref = le32(p + o);
    o += 4;
    names->clear();
    for (uint32_t i = 0; i < n_ref; i++) {
        if (o + 4 > n) return 0;
        const uint32_t l_name = le32(p + o);
        if (o + 4 + (uint64_t)l_name + 4 > n) return 0;
        const char* s = (const char*)p + o + 4;
        names->emplace_back(s, strnlen(s, l_name));
        o += 4 + (size_t)l_name + 4;
    }
    return o;
}

// one record body (after block_size): the fields the engine needs
void bam_record(const uint8_t* r, size_t len, Sink& sink) {
    if (len < 32) throw Err(E_FORMAT, "truncated BAM record");
    const int32_t tid = (int32_t)le32(r), pos = (int32_t)le32(r + 4);
    const uint32_t l_name = r[8], n_cigar = le16(r + 12), flag = le16(r + 14), l_seq = le32(r + 16);
    const size_t cig_off = 32 + (size_t)l_name, aux_off = cig_off + 4ull * n_cigar + ((size_t)l_seq + 1) / 2 + l_seq;
    if (aux_off > len) throw Err(E_FORMAT, "BAM record shorter than its fields");
    const uint8_t* cig = r + cig_off;
    uint32_t nc = n_cigar;
    // a CIGAR too long for the 16-bit count: the real one is the CG:B,I tag (bam_tag2cigar's conditions)
    if (n_cigar != 0 && tid >= 0 && pos >= 0 && (le32(cig) & 15u) == 4u && (le32(cig) >> 4) == l_seq) {
        const uint8_t* a = r + aux_off;
        const uint8_t* e = r + len;
        while (a + 3 <= e) {
            const uint8_t t0 = a[0], t1 = a[1], ty = a[2];
            a += 3;
            size_t sz;
            if (ty == 'A' || ty == 'c' || ty == 'C') sz = 1;
            else if (ty == 's' || ty == 'S') sz = 2;
            else if (ty == 'i' || ty == 'I' || ty == 'f') sz = 4;
            else if (ty == 'd') sz = 8;
            else if (ty == 'Z' || ty == 'H') {
                const uint8_t* z = (const uint8_t*)memchr(a, 0, (size_t)(e - a));
                if (!z) break;
                sz = (size_t)(z - a) + 1;
            } else if (ty == 'B') {
                if (a + 5 > e) break;
                const uint8_t sub = a[0];
                const uint32_t cnt = le32(a + 1);
                const size_t es = (sub == 'c' || sub == 'C') ? 1 : (sub == 's' || sub == 'S') ? 2 : (sub == 'i' || sub == 'I' || sub == 'f') ? 4 : 0;
                if (es == 0) break;
                if (t0 == 'C' && t1 == 'G') {
                    if ((sub == 'I' || sub == 'i') && cnt >= n_cigar && cnt < (1u << 29) && a + 5 + 4ull * cnt <= e) { cig = a + 5; nc = cnt; }
                    break;
                }
                sz = 5 + es * (size_t)cnt;
            } else break;
            if (t0 == 'C' && t1 == 'G') break;   // a CG tag of another type is not a CIGAR
            if ((size_t)(e - a) < sz) break;
            a += sz;
        }
    }
    uint64_t rlen = 0;
    for (uint32_t k = 0; k < nc; k++) {
        const uint32_t c = le32(cig + 4ull * k);
        if (cigar_consumes_ref(c & 15u)) rlen += c >> 4;
    }
    sink.record(tid, pos, flag, rlen);
}

struct BgzfBlock {
    size_t data, clen;   // the raw DEFLATE bytes in the file
    uint32_t isize, crc;
};

void run_threads(int T, const std::function<void(int)>& f);

// member at p of the bytes f[0, n): its total size, the DEFLATE bytes, ISIZE and CRC.  Returns 1, 0 when this gzip member has no
// BC field (not BGZF), 2 when f ends inside the member and more of the file follows (at_eof false: read on and ask again).
int bgzf_member(const uint8_t* f, size_t n, size_t p, bool at_eof, size_t* total, BgzfBlock* b) {
    if (n - p < 18) {
        if (!at_eof) return 2;
        throw Err(E_FORMAT, "corrupt BGZF / gzip member header");
    }
    if (f[p] != 0x1f || f[p + 1] != 0x8b || f[p + 2] != 8) throw Err(E_FORMAT, "corrupt BGZF / gzip member header");
    if (!(f[p + 3] & 4)) return 0;
    const size_t xlen = le16(f + p + 10);
    if (p + 12 + xlen > n) {
        if (!at_eof) return 2;
        throw Err(E_FORMAT, "truncated BGZF header");
    }
    size_t q = p + 12, qe = q + xlen;
    long bsize = -1;
    while (q + 4 <= qe) {
        const size_t slen = le16(f + q + 2);
        if (f[q] == 'B' && f[q + 1] == 'C' && slen == 2 && q + 6 <= qe) bsize = (long)le16(f + q + 4);
        q += 4 + slen;
    }
    if (bsize < 0) return 0;
    *total = (size_t)bsize + 1;
    if (*total < 12 + xlen + 8) throw Err(E_FORMAT, "truncated BGZF block");
    if (p + *total > n) {
        if (!at_eof) return 2;
        throw Err(E_FORMAT, "truncated BGZF block");
    }
    b->data = p + 12 + xlen;
    b->clen = *total - 12 - xlen - 8;
    b->crc = le32(f + p + *total - 8);
    b->isize = le32(f + p + *total - 4);
    return 1;
}

// The file is read with pread into plain buffers, never touched through a mapping: page-cache -> buffer copies run at memory
// speed on any number of threads, where first touches of a mapping are one page fault each and the threads' faults queue up
// behind each other (r03: eight threads scanning a mapped 111 MB SAM file took as long as one).
struct FileBytes {
    int fd = -1;
    size_t size = 0;
    explicit FileBytes(const std::string& path) {
        fd = ::open(path.c_str(), O_RDONLY);
        struct stat sb;
        if (fd < 0 || fstat(fd, &sb) != 0) {
            if (fd >= 0) ::close(fd);
            throw Err(E_IO, "Failed to open file: \"" + path + "\"");
        }
        size = (size_t)sb.st_size;
    }
    ~FileBytes() { if (fd >= 0) ::close(fd); }
    void read(size_t off, uint8_t* dst, size_t len, int T) const {   // all of [off, off + len): the caller stays inside the file
        auto part = [&](size_t a, size_t z) {
            while (a < z) {
                const ssize_t r = pread(fd, dst + a, z - a, (off_t)(off + a));
                if (r <= 0) throw Err(E_IO, "read error (or the file shrank while it was read)");
                a += (size_t)r;
            }
        };
        const int W = (len >= (8u << 20)) ? std::max(1, std::min(T, 8)) : 1;
        if (W == 1) { part(0, len); return; }
        run_threads(W, [&](int w) { part(len * (size_t)w / (size_t)W, len * (size_t)(w + 1) / (size_t)W); });
    }
};

void run_threads(int T, const std::function<void(int)>& f) {   // f(0) .. f(T-1) side by side (util.h: the process's thread pool)
    parallel_jobs((size_t)std::max(1, T), T, [&](size_t t) { f((int)t); });
}

// ---- SAM text -------------------------------------------------------------------------------------------------------------
// strtol(p, &p, 0) on a field: decimal, 0x hex or 0 octal, optional sign; false when no digits were read
bool parse_int_base0(const uint8_t* p, const uint8_t* e, int64_t* out) {
    bool neg = false;
    if (p < e && (*p == '+' || *p == '-')) { neg = *p == '-'; p++; }
    int base = 10;
    if (e - p >= 2 && p[0] == '0' && (p[1] == 'x' || p[1] == 'X')) { base = 16; p += 2; }
    else if (e - p >= 1 && p[0] == '0') base = 8;
    int64_t v = 0;
    bool any = false;
    for (; p < e; p++) {
        int d;
        if (*p >= '0' && *p <= '9') d = *p - '0';
        else if (*p >= 'a' && *p <= 'f') d = *p - 'a' + 10;
        else if (*p >= 'A' && *p <= 'F') d = *p - 'A' + 10;
        else break;
        if (d >= base) break;
        v = v * base + d;
        any = true;
        if (v > (1ll << 40)) break;
    }
    *out = neg ? -v : v;
    return any;
}

struct SamState {
    std::vector<std::string> names;
    std::unordered_map<std::string, int64_t> name2tid;
    std::vector<int64_t> tid2num;
    bool header_done = false;
};

void sam_line(const uint8_t* p, const uint8_t* e, SamState& S, const std::unordered_map<std::string, uint32_t>& seq_num, Sink& sink) {
    if (e > p && e[-1] == '\r') e--;
    if (p == e) return;
    if (!S.header_done && *p == '@') {
        if (e - p >= 4 && p[1] == 'S' && p[2] == 'Q' && p[3] == '\t') {   // @SQ: the target list, tid = order of appearance
            const uint8_t* q = p + 4;
            std::string sn;
            bool have = false;
            while (q < e) {
                const uint8_t* t = find_byte(q, e, '\t');
                if (!t) t = e;
                if (t - q >= 3 && q[0] == 'S' && q[1] == 'N' && q[2] == ':') { sn.assign((const char*)q + 3, (size_t)(t - q - 3)); have = true; }
                q = t + 1;
            }
            if (!have) throw Err(E_FORMAT, "@SQ line without an SN tag");
            if (!S.name2tid.count(sn)) {
                S.name2tid.emplace(sn, (int64_t)S.names.size());
                S.names.push_back(sn);
            }
        }
        return;
    }
    if (!S.header_done) {
        S.header_done = true;
        S.tid2num.assign(S.names.size(), -1);
        for (size_t i = 0; i < S.names.size(); i++) {
            auto it = seq_num.find(S.names[i]);
            if (it != seq_num.end()) S.tid2num[i] = it->second;
        }
    }
    // QNAME FLAG RNAME POS MAPQ CIGAR RNEXT PNEXT TLEN SEQ QUAL: eleven mandatory fields
    const uint8_t* f[12];
    int nf = 0;
    f[0] = p;
    for (const uint8_t* q = p; q < e && nf < 11; q++)
        if (*q == '\t') f[++nf] = q + 1;
    if (nf < 10) throw Err(E_FORMAT, "SAM record with fewer than 11 fields");
    auto fend = [&](int k) { return k < nf ? f[k + 1] - 1 : e; };
    int64_t flag = 0, pos1 = 0;
    if (!parse_int_base0(f[1], fend(1), &flag)) throw Err(E_FORMAT, "SAM record with a non-numeric FLAG");
    {
        const uint8_t* q = f[3];
        const uint8_t* qe = fend(3);
        bool neg = false, any = false;
        if (q < qe && (*q == '-' || *q == '+')) { neg = *q == '-'; q++; }
        for (; q < qe && *q >= '0' && *q <= '9'; q++) { pos1 = pos1 * 10 + (*q - '0'); any = true; if (pos1 > (1ll << 40)) break; }
        if (!any) throw Err(E_FORMAT, "SAM record with a non-numeric POS");
        if (neg) pos1 = -pos1;
    }
    int64_t tid = -1;
    {
        const size_t l = (size_t)(fend(2) - f[2]);
        if (!(l == 1 && f[2][0] == '*')) {
            auto it = S.name2tid.find(std::string((const char*)f[2], l));
            if (it != S.name2tid.end()) tid = it->second;   // an unknown name: treated as unmapped
        }
    }
    int64_t pos = pos1 - 1;
    if (pos < 0 && tid >= 0) tid = -1;   // "mapped query cannot have zero coordinate; treated as unmapped"
    uint32_t fl = (uint32_t)flag;
    if (tid < 0) fl |= 4u;
    uint64_t rlen = 0;
    {
        const uint8_t* q = f[5];
        const uint8_t* qe = fend(5);
        if (!(qe - q == 1 && *q == '*')) {
            while (q < qe) {
                uint64_t v = 0;
                bool any = false;
                for (; q < qe && *q >= '0' && *q <= '9'; q++) { v = v * 10 + (uint64_t)(*q - '0'); any = true; }
                if (!any || q >= qe) throw Err(E_FORMAT, "malformed CIGAR in a SAM record");
                static const char OPS[] = "MIDNSHP=XB";
                const char* o = (const char*)memchr(OPS, *q, 10);
                if (!o) throw Err(E_FORMAT, "unknown CIGAR operation in a SAM record");
                if (cigar_consumes_ref((uint32_t)(o - OPS))) rlen += v;
                q++;
            }
        }
    }
    sink.tid2num = &S.tid2num;
    sink.record(tid, pos, fl, rlen);
}

void sam_text(const uint8_t* p, size_t n, bool final_chunk, SamState& S, const std::unordered_map<std::string, uint32_t>& seq_num, Sink& sink,
              size_t* consumed) {
    const uint8_t* e = p + n;
    const uint8_t* q = p;
    while (q < e) {
        const uint8_t* nl = find_byte(q, e, '\n');
        if (!nl) {
            if (!final_chunk) break;
            nl = e;
        }
        sam_line(q, nl, S, seq_num, sink);
        q = nl < e ? nl + 1 : e;
    }
    *consumed = (size_t)(q - p);
}

// The same on T threads: the header lines (and the first record, which closes the header) serially, then the rest of the text cut
// into T pieces at line ends, each piece into its own interval list and counters, appended in file order.  An error is the one of
// the earliest piece that failed, as a serial pass would have met it first.
void sam_text_threads(const uint8_t* p, size_t n, bool final_chunk, SamState& S, const std::unordered_map<std::string, uint32_t>& seq_num, Sink& sink,
                      size_t* consumed, int T) {
    size_t par_min = 4u << 20;
    if (const char* ev = getenv("GFFX_SAM_PAR_MIN")) par_min = (size_t)std::max<long long>(0, atoll(ev));   // test hook
    if (T <= 1 || n < par_min) { sam_text(p, n, final_chunk, S, seq_num, sink, consumed); return; }
    const uint8_t* e = p + n;
    const uint8_t* q = p;
    while (q < e && !S.header_done) {   // serial until the header is closed
        const uint8_t* nl = find_byte(q, e, '\n');
        if (!nl) {
            if (!final_chunk) { *consumed = (size_t)(q - p); return; }
            nl = e;
        }
        sam_line(q, nl, S, seq_num, sink);
        q = nl < e ? nl + 1 : e;
    }
    // the last complete line's end bounds the parallel part
    const uint8_t* stop = e;
    if (!final_chunk) {
        while (stop > q && stop[-1] != '\n') stop--;
    }
    std::vector<const uint8_t*> cut((size_t)T + 1, stop);
    cut[0] = q;
    for (int t = 1; t < T; t++) {
        const uint8_t* c = q + (size_t)(stop - q) * (size_t)t / (size_t)T;
        if (c < cut[(size_t)t - 1]) c = cut[(size_t)t - 1];
        const uint8_t* nl = find_byte(c, stop, '\n');
        cut[(size_t)t] = nl ? nl + 1 : stop;
    }
    std::vector<Regions> part((size_t)T);
    std::vector<ReadStats> pst((size_t)T);
    std::vector<std::exception_ptr> perr((size_t)T);
    parallel_jobs((size_t)T, T, [&](size_t t) {
        try {
            Regions mine;      // on the thread's own stack: the pieces' list headers and counters side by side in one
            ReadStats mst{};   // array share cache lines, and every record updates them
            Sink local{&S.tid2num, &mine, &mst};
            size_t used;
            if (cut[t] < cut[t + 1]) sam_text(cut[t], (size_t)(cut[t + 1] - cut[t]), true, S, seq_num, local, &used);
            part[t] = std::move(mine);
            pst[t] = mst;
        } catch (...) { perr[t] = std::current_exception(); }
    });
    for (int t = 0; t < T; t++) {
        // what a serial pass had collected before the first failing piece is of no use to the caller either: fail as it would
        if (perr[(size_t)t]) std::rethrow_exception(perr[(size_t)t]);
        Regions& r = part[(size_t)t];
        sink.out->chr.insert(sink.out->chr.end(), r.chr.begin(), r.chr.end());
        sink.out->start.insert(sink.out->start.end(), r.start.begin(), r.start.end());
        sink.out->end.insert(sink.out->end.end(), r.end.begin(), r.end.end());
        sink.st->records += pst[(size_t)t].records; sink.st->unmapped += pst[(size_t)t].unmapped;
        sink.st->unknown_ref += pst[(size_t)t].unknown_ref; sink.st->skipped += pst[(size_t)t].skipped;
    }
    *consumed = (size_t)(stop - p);
}

}  // namespace

void read_alignments(const std::string& path, const std::unordered_map<std::string, uint32_t>& seq_num, int threads, Regions* out, ReadStats* st_out) {
    ReadStats st{};
    const FileBytes file(path);
    const size_t n = file.size;
    const int T = std::max(1, threads);
    Sink sink{nullptr, out, &st};
    SamState S;
    uint8_t magic[4] = {0, 0, 0, 0};
    file.read(0, magic, std::min<size_t>(n, 4), 1);
    if (n >= 4 && memcmp(magic, "CRAM", 4) == 0)
        throw Err(E_ARG, "CRAM input is not supported by this engine (no reference-based decoder); convert to BAM or SAM: " + path);
    if (!(n >= 2 && magic[0] == 0x1f && magic[1] == 0x8b)) {   // plain SAM text, in pieces of 8 MB: an incomplete last line waits for the next piece
        st.format = 0;
        size_t piece = 8u << 20;   // small enough to stay in cache-warm, already-faulted memory from piece to piece
        if (const char* e = getenv("GFFX_BGZF_CHUNK")) piece = (size_t)std::max<long long>(1, atoll(e));   // test hook
        std::unique_ptr<uint8_t[]> text;
        size_t cap = 0, have = 0, off = 0;
        do {
            const size_t m = std::min(piece, n - off);
            if (have + m > cap) {
                std::unique_ptr<uint8_t[]> bigger(new uint8_t[cap = have + m + (have + m) / 4 + 16]);
                if (have) memcpy(bigger.get(), text.get(), have);
                text = std::move(bigger);
            }
            file.read(off, text.get() + have, m, T);
            off += m;
            size_t used = 0;
            sam_text_threads(text.get(), have + m, off >= n, S, seq_num, sink, &used, T);
            have = have + m - used;
            if (have && used) memmove(text.get(), text.get() + used, have);
        } while (off < n);
        if (st_out) *st_out = st;
        return;
    }
    // ---- gzip family: BGZF members in parallel, anything else member by member -------------------------------------------
    std::vector<uint8_t> buf;      // carry (an incomplete record / line of the previous chunk) + this chunk's inflated bytes
    size_t carry = 0;
    int kind = -1;                 // 1 BAM, 0 SAM text; decided by the first inflated bytes
    std::vector<std::string> names;
    std::vector<int64_t> tid2num;
    bool bam_header_done = false;
    auto consume_at = [&](const uint8_t* p, const size_t len, bool final_chunk) -> size_t {
        size_t o = 0;
        if (kind < 0 && (len >= 4 || final_chunk)) kind = (len >= 4 && memcmp(p, "BAM\1", 4) == 0) ? 1 : 0;
        if (kind == 1) {
            if (!bam_header_done) {
                const size_t h = bam_header(p, len, &names);
                if (h == 0) {
                    if (final_chunk) throw Err(E_FORMAT, "truncated BAM header");
                    return 0;   // nothing used: the header continues in the next chunk
                }
                tid2num.assign(names.size(), -1);
                for (size_t i = 0; i < names.size(); i++) {
                    auto it = seq_num.find(names[i]);
                    if (it != seq_num.end()) tid2num[i] = it->second;
                }
                sink.tid2num = &tid2num;
                bam_header_done = true;
                o = h;
            }
            while (o + 4 <= len) {
                const size_t bs = le32(p + o);
                if (o + 4 + bs > len) break;
                bam_record(p + o + 4, bs, sink);
                o += 4 + bs;
            }
            if (final_chunk && o != len) throw Err(E_FORMAT, "truncated BAM record at the end of the file");
        } else if (kind == 0) {
            sam_text(p, len, final_chunk, S, seq_num, sink, &o);
        }
        return o;
    };
    auto consume = [&](bool final_chunk) {   // the vector form: what is left moves to the front of `buf`
        const size_t o = consume_at(buf.data(), buf.size(), final_chunk);
        carry = buf.size() - o;
        if (carry && o) memmove(buf.data(), buf.data() + o, carry);
    };
    size_t p = 0;
    // BGZF: chunks of ~16 MB of inflated bytes; while the main thread walks one chunk's records (or lines), the workers inflate
    // the next chunk into the other buffer.  An incomplete record at a chunk's end is copied in front of the next chunk's bytes
    // (the buffers keep a head room for it).  The buffers are plain allocations, reused: no zero fill, no page faults per chunk.
    size_t CHUNK = 16u << 20;
    if (const char* e = getenv("GFFX_BGZF_CHUNK")) CHUNK = (size_t)std::max<long long>(1, atoll(e));   // test hook: many chunks, small head rooms
    const size_t HEAD = std::min<size_t>(1u << 20, CHUNK);
    struct RawBuf {
        std::unique_ptr<uint8_t[]> mem;
        size_t cap = 0, head = 0;
        void need(size_t head_room, size_t bytes) {
            if (std::max(head, head_room) + bytes + 16 > cap || head_room > head) {
                head = std::max(head, head_room);
                cap = head + bytes + bytes / 4 + 16;
                mem.reset(new uint8_t[cap]);
            }
        }
    };
    struct Chunk {
        std::unique_ptr<uint8_t[]> comp;   // the piece of the file the chunk's blocks lie in (BgzfBlock::data is relative to it)
        size_t comp_cap = 0;
        std::vector<BgzfBlock> blocks;
        std::vector<size_t> at;   // offset of every block's bytes behind the buffer's head
        size_t tot = 0;
    };
    bool plain = false;   // a gzip member without the BGZF field follows the blocks read so far
    // the next piece of the file from p on, and the whole blocks in it, until ~CHUNK inflated bytes; a block cut off by the piece's
    // end is read again with the next piece
    auto gather = [&](Chunk& c) {
        c.blocks.clear(); c.at.clear(); c.tot = 0;
        size_t piece = std::max<size_t>(std::min<size_t>(CHUNK, 8u << 20), 1u << 16);   // at least one whole block
        for (;;) {
            const size_t m = std::min(piece, n - p);
            if (m > c.comp_cap) c.comp.reset(new uint8_t[c.comp_cap = m + 16]);
            file.read(p, c.comp.get(), m, T);
            const bool at_eof = p + m >= n;
            size_t r = 0, total = 0;
            BgzfBlock b;
            int rc = 1;
            while (r < m && c.tot < CHUNK) {
                rc = bgzf_member(c.comp.get(), m, r, at_eof, &total, &b);
                if (rc != 1) break;
                c.at.push_back(c.tot);
                c.blocks.push_back(b);
                c.tot += b.isize;
                r += total;
            }
            if (rc == 0) plain = true;
            if (rc == 2 && c.blocks.empty()) { piece *= 2; continue; }   // not even one block fits: a larger piece (a member is at most 64 KB + extra fields)
            p += r;
            return !c.blocks.empty();
        }
    };
    auto inflate_chunk = [&](const Chunk& c, RawBuf& B) {
        B.need(HEAD, c.tot);
        uint8_t* dst = B.mem.get() + B.head;
        std::atomic<size_t> next{0};
        run_threads((int)std::min<size_t>((size_t)T, c.blocks.size()), [&](int) {
            InflateTables tab;
            for (size_t k; (k = next.fetch_add(1)) < c.blocks.size();) {
                const BgzfBlock& x = c.blocks[k];
                if (x.isize == 0 && x.clen <= 2) continue;   // the empty end-of-file member
                const long r = inflate_raw(c.comp.get() + x.data, x.clen, dst + c.at[k], x.isize, tab);
                if (r != (long)x.isize) throw Err(E_FORMAT, "BGZF block does not inflate to its recorded size: " + path);
                if (crc32_bytes(dst + c.at[k], x.isize) != x.crc) throw Err(E_FORMAT, "BGZF block with a CRC32 mismatch: " + path);
            }
        });
    };
    {
        Chunk ck[2];
        RawBuf B[2];
        int cur = 0;
        bool have = gather(ck[0]);
        if (have) { st.format = 1; inflate_chunk(ck[0], B[0]); }
        size_t carry_len = 0;   // bytes of the previous chunk sitting right in front of B[cur]'s head
        while (have) {
            const bool have_next = !plain && p < n && gather(ck[1 - cur]);
            struct Bg {
                std::thread t;
                std::exception_ptr err;
                ~Bg() { if (t.joinable()) t.join(); }
            } bg;
            if (have_next) {
                bg.t = std::thread([&] {
                    try { inflate_chunk(ck[1 - cur], B[1 - cur]); } catch (...) { bg.err = std::current_exception(); }
                });
            }
            const uint8_t* data = B[cur].mem.get() + B[cur].head - carry_len;
            const size_t len = carry_len + ck[cur].tot;
            const size_t used = consume_at(data, len, !have_next && !plain);
            const size_t left = len - used;
            if (bg.t.joinable()) bg.t.join();
            if (bg.err) std::rethrow_exception(bg.err);
            if (!have_next) {   // what is left goes to the member-by-member path below (or was an error already)
                buf.assign(data + used, data + len);
                carry = left;
                break;
            }
            RawBuf& N = B[1 - cur];
            if (left > N.head) {   // an incomplete record larger than the head room: a new buffer with room for it
                RawBuf big;
                big.need(left + HEAD, ck[1 - cur].tot);
                memcpy(big.mem.get() + big.head, N.mem.get() + N.head, ck[1 - cur].tot);
                N = std::move(big);
            }
            if (left) memcpy(N.mem.get() + N.head - left, data + used, left);
            carry_len = left;
            cur = 1 - cur;
        }
    }
    {
        if (plain) {   // a gzip member without the BGZF field (gzip-ed SAM): sequential, the size is found by trial
            st.format = 2;
            const size_t base = p;   // the rest of the file in memory; f[x] = file byte x
            std::unique_ptr<uint8_t[]> rest(new uint8_t[n - base + 16]);
            file.read(base, rest.get(), n - base, T);
            const uint8_t* f = rest.get() - base;
            if (n - p < 18) throw Err(E_FORMAT, "truncated gzip member");
            const size_t xoff = 10 + ((f[p + 3] & 4) ? 2 + (size_t)le16(f + p + 10) : 0);
            size_t q = p + xoff;
            if (f[p + 3] & 8) { while (q < n && f[q]) q++; q++; }
            if (f[p + 3] & 16) { while (q < n && f[q]) q++; q++; }
            if (f[p + 3] & 2) q += 2;
            if (q + 8 > n) throw Err(E_FORMAT, "truncated gzip member");
            InflateTables tab;
            size_t cap = std::max<size_t>(4 * (n - q), 1 << 20);
            for (;;) {
                buf.resize(carry + cap);
                const long r = inflate_raw(f + q, n - q - 8, buf.data() + carry, cap, tab);
                if (r == -2) { cap *= 2; continue; }
                if (r < 0) throw Err(E_FORMAT, "corrupt gzip stream: " + path);
                buf.resize(carry + (size_t)r);
                if (le32(f + n - 4) != (uint32_t)r || le32(f + n - 8) != crc32_bytes(buf.data() + carry, (size_t)r))
                    throw Err(E_FORMAT, "gzip member with a size or CRC32 mismatch: " + path);
                break;
            }
            p = n;   // one member (what gzip writes); concatenated plain members are not looked for
            consume(true);
        }
    }
    if (kind < 0) consume(true);   // nothing has been looked at yet (an empty stream)
    if (st_out) *st_out = st;
}

}  // namespace gffx
