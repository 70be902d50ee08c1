// reads_oracle.hpp — TEST INFRASTRUCTURE (CPU oracle): the read source of depth / coverage from alignment files, i.e. what
// process_bam (/root/reference/src/commands/depth.rs:297-423) and collect_by_root_from_bam
// (/root/reference/src/commands/coverage.rs:125-205) obtain from rust-htslib (Cargo.toml:32, not vendored; htslib itself is a
// C library that is absent from this image).  PARITY UNPINNED BY THE REFERENCE for this file: nothing under /root/reference
// holds a BAM/SAM fixture or an expected output, and htslib cannot be run here; the rules below restate the SAM
// specification (SAMv1 §1.4, §4.1 BGZF, §4.2 BAM) and htslib's published behaviour of sam_parse1 / bam_endpos /
// bam_tag2cigar, and the tests pin them with hand-built files whose expected triples the test generator knows.
// Deliberately independent of the product's reader (gffx_b200/csrc/reads.cpp, inflate.h): the container is read with zlib's
// gzread (any gzip member sequence, BGZF included, or plain text), whole file at once, records with std::string tools.
#pragma once
#include <zlib.h>

#include <cstdint>
#include <sstream>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace orc_reads {

struct Triple { uint32_t chr, start, end; };

inline std::string slurp(const std::string& path) {
    gzFile g = gzopen(path.c_str(), "rb");
    if (!g) throw std::runtime_error("cannot open " + path);
    std::string s;
    char buf[1 << 16];
    for (;;) {
        int k = gzread(g, buf, sizeof buf);
        if (k < 0) { gzclose(g); throw std::runtime_error("gzread failed: " + path); }
        if (k == 0) break;
        s.append(buf, (size_t)k);
    }
    gzclose(g);
    return s;
}

// depth.rs:335-366: is_unmapped -> skip; tid < 0 -> skip; tid2num[tid] None -> skip; pos < 0 -> skip; end <= pos -> skip; clamp
inline void keep(std::vector<Triple>& out, const std::vector<long>& tid2num, long tid, long long pos, unsigned flag, unsigned long long ref_len) {
    if (flag & 0x4) return;
    if (tid < 0) return;
    if ((size_t)tid >= tid2num.size()) throw std::runtime_error("tid beyond the header");
    if (tid2num[(size_t)tid] < 0) return;
    if (pos < 0) return;
    const long long end = pos + (long long)(ref_len == 0 ? 1 : ref_len);   // htslib bam_endpos: rlen 0 counts as 1
    if (end <= pos) return;
    auto clamp = [](long long v) { return (uint32_t)(v < 0 ? 0 : v > 0xFFFFFFFFll ? 0xFFFFFFFFll : v); };
    out.push_back(Triple{(uint32_t)tid2num[(size_t)tid], clamp(pos), clamp(end)});
}

inline uint32_t rd32(const std::string& s, size_t o) {
    if (o + 4 > s.size()) throw std::runtime_error("truncated BAM");
    return (uint32_t)(uint8_t)s[o] | ((uint32_t)(uint8_t)s[o + 1] << 8) | ((uint32_t)(uint8_t)s[o + 2] << 16) | ((uint32_t)(uint8_t)s[o + 3] << 24);
}

inline std::vector<Triple> from_bam(const std::string& s, const std::unordered_map<std::string, uint32_t>& seqid_to_num) {
    size_t o = 4;
    o += 4 + rd32(s, o);   // header text
    const uint32_t n_ref = rd32(s, o);
    o += 4;
    std::vector<long> tid2num;
    for (uint32_t i = 0; i < n_ref; i++) {
        const uint32_t l = rd32(s, o);
        std::string name(s.c_str() + o + 4);   // NUL-terminated
        auto it = seqid_to_num.find(name);
        tid2num.push_back(it == seqid_to_num.end() ? -1 : (long)it->second);
        o += 4 + l + 4;
    }
    std::vector<Triple> out;
    while (o < s.size()) {
        const uint32_t bs = rd32(s, o);
        const size_t r = o + 4;
        if (r + bs > s.size()) throw std::runtime_error("truncated BAM record");
        const int32_t tid = (int32_t)rd32(s, r), pos = (int32_t)rd32(s, r + 4);
        const uint32_t l_read_name = (uint8_t)s[r + 8];
        const uint32_t flag_nc = rd32(s, r + 12);   // n_cigar_op (low 16) | flag (high 16)
        uint32_t n_cigar = flag_nc & 0xFFFF;
        const uint32_t flag = flag_nc >> 16, l_seq = rd32(s, r + 16);
        size_t c = r + 32 + l_read_name;
        // htslib bam_tag2cigar: a placeholder CIGAR "<l_seq>S<x>N" with the real one in CG:B,I
        if (n_cigar > 0 && tid >= 0 && pos >= 0 && (rd32(s, c) & 0xF) == 4 && (rd32(s, c) >> 4) == l_seq) {
            size_t a = c + 4ull * n_cigar + (l_seq + 1) / 2 + l_seq, e = r + bs;
            while (a + 3 <= e) {
                const std::string tag = s.substr(a, 2);
                const char ty = s[a + 2];
                a += 3;
                size_t w = 0;
                switch (ty) {
                    case 'A': case 'c': case 'C': w = 1; break;
                    case 's': case 'S': w = 2; break;
                    case 'i': case 'I': case 'f': w = 4; break;
                    case 'd': w = 8; break;
                    case 'Z': case 'H': w = s.find('\0', a) - a + 1; break;
                    case 'B': {
                        const char sub = s[a];
                        const uint32_t cnt = rd32(s, a + 1);
                        const size_t es = (sub == 'c' || sub == 'C') ? 1 : (sub == 's' || sub == 'S') ? 2 : 4;
                        if (tag == "CG" && (sub == 'I' || sub == 'i') && cnt >= n_cigar && cnt < (1u << 29)) { c = a + 5; n_cigar = cnt; a = e; }
                        w = 5 + es * cnt;
                        break;
                    }
                    default: a = e;
                }
                if (tag == "CG") break;
                a += w;
            }
        }
        unsigned long long ref_len = 0;
        for (uint32_t k = 0; k < n_cigar; k++) {
            const uint32_t v = rd32(s, c + 4ull * k), op = v & 0xF;
            if (op == 0 || op == 2 || op == 3 || op == 7 || op == 8) ref_len += v >> 4;   // M D N = X consume the reference
        }
        keep(out, tid2num, tid, pos, flag, ref_len);
        o = r + bs;
    }
    return out;
}

inline std::vector<Triple> from_sam(const std::string& s, const std::unordered_map<std::string, uint32_t>& seqid_to_num) {
    std::vector<std::string> targets;
    std::unordered_map<std::string, long> tid_of;
    std::vector<long> tid2num;
    std::vector<Triple> out;
    std::istringstream in(s);
    std::string line;
    bool in_header = true;
    while (std::getline(in, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.empty()) continue;
        if (in_header && line[0] == '@') {
            if (line.compare(0, 4, "@SQ\t") == 0) {
                std::istringstream f(line.substr(4));
                std::string kv, sn;
                bool have = false;
                while (std::getline(f, kv, '\t'))
                    if (kv.compare(0, 3, "SN:") == 0) { sn = kv.substr(3); have = true; }
                if (!have) throw std::runtime_error("@SQ without SN");
                if (!tid_of.count(sn)) { tid_of[sn] = (long)targets.size(); targets.push_back(sn); }
            }
            continue;
        }
        if (in_header) {
            in_header = false;
            for (auto& t : targets) {
                auto it = seqid_to_num.find(t);
                tid2num.push_back(it == seqid_to_num.end() ? -1 : (long)it->second);
            }
        }
        std::vector<std::string> col;
        {
            std::istringstream f(line);
            std::string c;
            while (std::getline(f, c, '\t')) col.push_back(c);
        }
        if (col.size() < 11) throw std::runtime_error("SAM record with fewer than 11 fields");
        char* endp = nullptr;
        unsigned flag = (unsigned)strtol(col[1].c_str(), &endp, 0);   // sam_parse1: strtol(p, &p, 0)
        if (endp == col[1].c_str()) throw std::runtime_error("bad FLAG");
        long tid = -1;
        if (col[2] != "*") { auto it = tid_of.find(col[2]); if (it != tid_of.end()) tid = it->second; }
        long long pos = strtoll(col[3].c_str(), &endp, 10) - 1;
        if (endp == col[3].c_str()) throw std::runtime_error("bad POS");
        if (pos < 0 && tid >= 0) tid = -1;   // "mapped query cannot have zero coordinate; treated as unmapped"
        if (tid < 0) flag |= 0x4;
        unsigned long long ref_len = 0;
        if (col[5] != "*") {
            const char* p = col[5].c_str();
            while (*p) {
                char* q;
                unsigned long long v = strtoull(p, &q, 10);
                if (q == p || !*q) throw std::runtime_error("bad CIGAR");
                const char op = *q;
                if (!strchr("MIDNSHP=XB", op)) throw std::runtime_error("bad CIGAR op");
                if (op == 'M' || op == 'D' || op == 'N' || op == '=' || op == 'X') ref_len += v;
                p = q + 1;
            }
        }
        keep(out, tid2num, tid, pos, flag, ref_len);
    }
    return out;
}

inline std::vector<Triple> read_alignments(const std::string& path, const std::unordered_map<std::string, uint32_t>& seqid_to_num) {
    const std::string s = slurp(path);
    if (s.size() >= 4 && s.compare(0, 4, "BAM\1") == 0) return from_bam(s, seqid_to_num);
    return from_sam(s, seqid_to_num);
}

}  // namespace orc_reads
