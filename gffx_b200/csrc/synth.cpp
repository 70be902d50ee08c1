// synth.cpp — synthetic workloads of SURVEY.md §8d: a human-like GFF3 (NCBI flavour) and BED regions.
#include <algorithm>
#include <cmath>
#include "engine.h"
#include "synth_common.h"

namespace gffx {

namespace {
// GRCh38 primary assembly lengths (chr1..22, X, Y)
const uint32_t kHsLen[24] = {248956422, 242193529, 198295559, 190214555, 181538259, 170805979, 159345973, 145138636,
                                138394717, 133797422, 135086622, 133275309, 114364328, 107043718, 101991189, 90338345,
                                83257441,  80373285,  58617616,  64444167,  46709983,  50818468,  156040895, 57227415};
const char* kHsName[24] = {"chr1",  "chr2",  "chr3",  "chr4",  "chr5",  "chr6",  "chr7",  "chr8",  "chr9",  "chr10", "chr11", "chr12",
                              "chr13", "chr14", "chr15", "chr16", "chr17", "chr18", "chr19", "chr20", "chr21", "chr22", "chrX",  "chrY"};

struct Gene {
    uint32_t chr, start, end;  // 1-based closed
    uint64_t key;              // generator index
};

struct Out {
    FILE* f;
    std::string buf;
    uint64_t lines = 0;
    void line(const std::string& s) {
        buf += s;
        buf += '\n';
        lines++;
        if (buf.size() > (8u << 20)) flush();
    }
    void raw(const std::string& s) { buf += s; }
    void flush() {
        if (!buf.empty()) fwrite(buf.data(), 1, buf.size(), f);
        buf.clear();
    }
};

std::string row(const char* chr, const char* type, uint64_t s, uint64_t e, char strand, const std::string& attrs, const char* phase = ".") {
    char head[160];
    snprintf(head, sizeof head, "%s\tBestRefSeq\t%s\t%llu\t%llu\t.\t%c\t%s\t", chr, type, (unsigned long long)s, (unsigned long long)e,
             strand, phase);
    return std::string(head) + attrs;
}
}  // namespace

// Returns the number of lines written (comments included).  n_scaffolds > 0 appends that many unplaced scaffolds
// (3-150 kb, one to three genes each) after the 24 chromosomes: the shape of a scaffold-level assembly, or of the 583
// sequences of the reference's benchmark/hs.gff.sqs.
uint64_t synth_gff_file(const std::string& path, uint64_t seed, uint32_t n_genes, bool quirks, uint32_t n_scaffolds) {
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw Err(E_IO, "cannot create " + path);
    Out o{f, std::string(), 0};
    uint64_t total = 0;
    for (uint32_t l : kHsLen) total += l;
    std::vector<std::string> name_store(24 + (size_t)n_scaffolds);
    std::vector<uint32_t> kChromLen(kHsLen, kHsLen + 24);
    std::vector<const char*> kChromName(kHsName, kHsName + 24);
    for (uint32_t k = 0; k < n_scaffolds; k++) {
        char nm[32];
        snprintf(nm, sizeof nm, "scaf%05u", k);
        name_store[24 + k] = nm;
        kChromLen.push_back(3000 + (uint32_t)(rnd2(seed, 1000000 + k, 7) % 147000));
    }
    for (uint32_t k = 0; k < n_scaffolds; k++) kChromName.push_back(name_store[24 + k].c_str());
    std::vector<Gene> genes(n_genes);
    for (uint32_t g = 0; g < n_genes; g++) {
        uint64_t t = mulhi64(rnd2(seed, g, 1), total);
        uint32_t c = 0;
        while (t >= kChromLen[c]) { t -= kChromLen[c]; c++; }
        // log-normal length: median 25 kb, sigma 1.2, clamped to [300, 2.3 Mb]
        double u1 = ((rnd2(seed, g, 2) >> 11) + 1) * (1.0 / 9007199254740993.0), u2 = (rnd2(seed, g, 3) >> 11) * (1.0 / 9007199254740992.0);
        double z = std::sqrt(-2.0 * std::log(u1)) * std::cos(6.283185307179586 * u2);
        double len = std::exp(std::log(25000.0) + 1.2 * z);
        uint32_t L = (uint32_t)std::min(2300000.0, std::max(300.0, len));
        uint32_t s = (uint32_t)t + 1;
        if ((uint64_t)s + L - 1 > kChromLen[c]) s = kChromLen[c] - L + 1;
        genes[g] = Gene{c, s, s + L - 1, g};
    }
    for (uint32_t k = 0; k < n_scaffolds; k++) {
        const uint32_t ng = 1 + (rnd2(seed, 1000000 + k, 8) % 4 == 0 ? (uint32_t)(rnd2(seed, 1000000 + k, 9) % 2) + 1 : 0);
        for (uint32_t j = 0; j < ng; j++) {
            const uint32_t len = kChromLen[24 + k];
            const uint32_t L = std::max<uint32_t>(300, (uint32_t)(rnd2(seed, 1000000 + k, 20 + j) % (len / 2)));
            const uint32_t s = 1 + (uint32_t)(rnd2(seed, 1000000 + k, 30 + j) % (len - L));
            genes.push_back(Gene{24 + k, s, s + L - 1, (uint64_t)n_genes + 4ull * k + j});
        }
    }
    std::sort(genes.begin(), genes.end(), [](const Gene& a, const Gene& b) {
        if (a.chr != b.chr) return a.chr < b.chr;
        if (a.start != b.start) return a.start < b.start;
        return a.key < b.key;
    });
    o.line("##gff-version 3");
    o.line("#!gff-spec-version 1.21");
    o.line("#!processor gffx_b200 synthetic generator");
    char tmp[512];
    uint32_t cur_chr = UINT32_MAX;
    for (size_t gi = 0; gi < genes.size(); gi++) {
        const Gene& G = genes[gi];
        const char* cn = kChromName[G.chr];
        if (G.chr != cur_chr) {
            // NCBI style: a directive and a `region` line open each chromosome.  `region` is in the index
            // builder's skip list, so the line lands physically inside the previous chromosome's last block.
            for (uint32_t c = cur_chr == UINT32_MAX ? 0 : cur_chr + 1; c <= G.chr; c++) {
                snprintf(tmp, sizeof tmp, "##sequence-region %s 1 %u", kChromName[c], kChromLen[c]);
                o.line(tmp);
                snprintf(tmp, sizeof tmp, "ID=%s:1..%u;Dbxref=taxon:9606;Name=%s;gbkey=Src;genome=chromosome;mol_type=genomic DNA",
                         kChromName[c], kChromLen[c], kChromName[c] + 3);
                o.line(row(kChromName[c], "region", 1, kChromLen[c], '+', tmp));
                if (quirks && c % 3 == 1) {
                    // a chromosome-spanning root of a non-skipped type: the degenerate case for start-sorted scans
                    snprintf(tmp, sizeof tmp, "ID=span-%s;Name=whole_%s", kChromName[c], kChromName[c]);
                    o.line(row(kChromName[c], "biological_region", 1, kChromLen[c], '+', tmp));
                }
            }
            cur_chr = G.chr;
        }
        const uint64_t k = G.key;
        const char strand = (rnd2(seed, k, 4) & 1) ? '+' : '-';
        const uint32_t kind = (uint32_t)(rnd2(seed, k, 5) % 100);
        char gid[48];
        snprintf(gid, sizeof gid, "gene-G%06llu", (unsigned long long)k);
        const uint32_t glen = G.end - G.start + 1;
        if (kind < 3) {  // childless pseudogene root
            snprintf(tmp, sizeof tmp, "ID=%s;Dbxref=GeneID:%llu;Name=G%06lluP;gbkey=Gene;gene=G%06lluP;gene_biotype=pseudogene;pseudo=true",
                     gid, (unsigned long long)(100000 + k), (unsigned long long)k, (unsigned long long)k);
            o.line(row(cn, "pseudogene", G.start, G.end, strand, tmp));
            continue;
        }
        const bool lnc = kind < 5;
        if (quirks && k % 97 == 0) snprintf(tmp, sizeof tmp, "geneID=%s;ID=shadow-%llu;Name=G%06llu", gid, (unsigned long long)k, (unsigned long long)k);
        else
            snprintf(tmp, sizeof tmp, "ID=%s;Dbxref=GeneID:%llu,HGNC:HGNC:%llu;Name=G%06llu;gbkey=Gene;gene=G%06llu;gene_biotype=%s",
                     gid, (unsigned long long)(100000 + k), (unsigned long long)(k % 50000), (unsigned long long)k, (unsigned long long)k,
                     lnc ? "lncRNA" : "protein_coding");
        std::string gattrs = tmp;
        if (quirks && k % 53 == 0) {
            // reversed columns 4/5: swapped by the index builder and by depth/coverage, raw in the line filter
            snprintf(tmp, sizeof tmp, "%s\tBestRefSeq\tgene\t%u\t%u\t.\t%c\t.\t", cn, G.end, G.start, strand);
            o.line(std::string(tmp) + gattrs);
        } else {
            o.line(row(cn, "gene", G.start, G.end, strand, gattrs));
        }
        const std::string parent_gene = (quirks && k % 97 == 0) ? std::string(gid) : std::string(gid);
        const uint32_t n_tx = 1 + (uint32_t)(rnd2(seed, k, 6) % 5);
        for (uint32_t m = 0; m < n_tx; m++) {
            uint32_t n_ex = 2 + (uint32_t)(rnd2(seed, k, 10 + m) % 13);
            if (glen < 8 * n_ex) n_ex = std::max(1u, glen / 8);
            const uint32_t cells = 2 * n_ex, w = glen / cells;
            std::vector<uint32_t> p(cells);
            for (uint32_t c = 0; c < cells; c++) p[c] = G.start + c * w + (uint32_t)(rnd2(seed, k, 100 + m * 64 + c) % w);
            if (m == 0) { p[0] = G.start; p[cells - 1] = G.end; }
            char tid[64];
            snprintf(tid, sizeof tid, "rna-T%06llu.%u", (unsigned long long)k, m + 1);
            snprintf(tmp, sizeof tmp, "ID=%s;Parent=%s;Dbxref=GeneID:%llu;Name=T%06llu.%u;gbkey=%s;gene=G%06llu;product=synthetic transcript %u;transcript_id=T%06llu.%u",
                     tid, parent_gene.c_str(), (unsigned long long)(100000 + k), (unsigned long long)k, m + 1, lnc ? "ncRNA" : "mRNA",
                     (unsigned long long)k, m + 1, (unsigned long long)k, m + 1);
            o.line(row(cn, lnc ? "lnc_RNA" : "mRNA", p[0], p[cells - 1], strand, tmp));
            for (uint32_t x = 0; x < n_ex; x++) {
                if (quirks && (k + m + x) % 211 == 0) {
                    // an exon ID reused by several genes: depth/coverage merge rows by ID string across blocks
                    snprintf(tmp, sizeof tmp, "ID=exon-shared-%llu;Parent=%s;gbkey=mRNA", (unsigned long long)((k + m + x) / 211 % 7), tid);
                } else {
                    snprintf(tmp, sizeof tmp, "ID=exon-T%06llu.%u-%u;Parent=%s;Dbxref=GeneID:%llu;gbkey=%s;gene=G%06llu;transcript_id=T%06llu.%u",
                             (unsigned long long)k, m + 1, x + 1, tid, (unsigned long long)(100000 + k), lnc ? "ncRNA" : "mRNA",
                             (unsigned long long)k, (unsigned long long)k, m + 1);
                }
                std::string l = row(cn, "exon", p[2 * x], p[2 * x + 1], strand, tmp);
                if (quirks && (k + x) % 61 == 0) l += "\r";  // CRLF line: '\r' is never stripped
                o.line(l);
            }
            if (!lnc) {
                const uint32_t c0 = n_ex >= 3 ? 1 : 0, c1 = n_ex >= 3 ? n_ex - 1 : n_ex;
                for (uint32_t x = c0; x < c1; x++) {  // one ID shared by all CDS lines of the transcript
                    snprintf(tmp, sizeof tmp, "ID=cds-P%06llu.%u;Parent=%s;Dbxref=GeneID:%llu;Name=P%06llu.%u;gbkey=CDS;gene=G%06llu;protein_id=P%06llu.%u",
                             (unsigned long long)k, m + 1, tid, (unsigned long long)(100000 + k), (unsigned long long)k, m + 1,
                             (unsigned long long)k, (unsigned long long)k, m + 1);
                    o.line(row(cn, "CDS", p[2 * x], p[2 * x + 1], strand, tmp, x % 3 == 0 ? "0" : x % 3 == 1 ? "1" : "2"));
                }
            }
            if (quirks && (k + m) % 37 == 0) {
                // a child reaching outside its gene (trans-spliced style): an "irregular" unit for depth/coverage
                const uint32_t es = G.end + 1000 + (uint32_t)(rnd2(seed, k, 900 + m) % 5000);
                if ((uint64_t)es + 200 < kChromLen[G.chr]) {
                    snprintf(tmp, sizeof tmp, "ID=exon-far-T%06llu.%u;Parent=%s", (unsigned long long)k, m + 1, tid);
                    o.line(row(cn, "exon", es, es + 150, strand, tmp));
                }
            }
        }
        if (quirks && G.chr >= 24) {
            const uint32_t sc = G.chr - 24;
            if (sc % 7 == 0) {   // a child that names another sequence and lies far outside its root (App. D.4/5)
                snprintf(tmp, sizeof tmp, "ID=exon-foreign-%u;Parent=%s", sc, gid);
                o.line(row("chr1", "exon", 5000000 + 1000 * (sc % 50), 5000150 + 1000 * (sc % 50), strand, tmp));
            }
            if (sc % 401 == 5) {  // a coordinate at the top of u32: that sequence cannot share a virtual sequence
                snprintf(tmp, sizeof tmp, "ID=exon-top-%u;Parent=%s", sc, gid);
                o.line(row(cn, "exon", 4294960000ull, 4294967295ull, strand, tmp));
            }
        }
        if (quirks) {
            if (k % 29 == 0) o.line("# a comment line inside a block");
            if (k % 31 == 0) o.raw("\n");  // blank line inside a block
            if (k % 41 == 0) {             // skipped-type line with unusable coordinates
                snprintf(tmp, sizeof tmp, "ID=remark-%llu;Note=free text", (unsigned long long)k);
                o.line(std::string(cn) + "\tsynth\tremark\tabc\t12x\t.\t.\t.\t" + tmp);
            }
            if (k % 43 == 0) {  // end == 0: dropped by the indexer and by depth/coverage, raw 0/0 in the line filter
                snprintf(tmp, sizeof tmp, "ID=zero-%llu;Parent=%s", (unsigned long long)k, gid);
                o.line(std::string(cn) + "\tsynth\texon\t0\t0\t.\t+\t.\t" + tmp);
            }
            if (k % 47 == 0) {  // multi-parent child: resolves to no parent, becomes a root (App. E.4)
                snprintf(tmp, sizeof tmp, "ID=multi-%llu;Parent=%s,gene-G%06llu", (unsigned long long)k, gid, (unsigned long long)(k + 1));
                o.line(row(cn, "exon", G.start, std::min(G.end, G.start + 99), strand, tmp));
            }
            if (k % 59 == 0) {  // duplicate root ID: two tree intervals with one fid, block = last .gof record
                snprintf(tmp, sizeof tmp, "ID=gene-dup-%llu;Name=dup", (unsigned long long)(k % 3));
                o.line(row(cn, "gene", G.start, std::min(G.end, G.start + 499), strand, tmp));
            }
            if (k % 67 == 0) {  // ID terminated by a space inside column 9
                snprintf(tmp, sizeof tmp, "Note=x;ID=spaced-%llu trailing words;Parent=%s", (unsigned long long)k, gid);
                o.line(row(cn, "exon", G.start, std::min(G.end, G.start + 49), strand, tmp));
            }
        }
    }
    o.flush();
    fclose(f);
    return o.lines;
}

void synth_seq_table(const HostIndex& H, std::vector<uint32_t>* seq_extent, std::vector<uint64_t>* cum) {
    const size_t n = H.seq_names.size();
    seq_extent->assign(n, 0);
    for (const RootIv& r : H.roots) (*seq_extent)[r.seq] = std::max((*seq_extent)[r.seq], r.end);
    cum->assign(n + 1, 0);
    for (size_t i = 0; i < n; i++) (*cum)[i + 1] = (*cum)[i] + (*seq_extent)[i];
}

void synth_regions_host(const HostIndex& H, uint64_t seed, uint64_t first, uint64_t n, uint32_t length, uint32_t* chr,
                        uint32_t* start, uint32_t* end) {
    std::vector<uint32_t> ext;
    std::vector<uint64_t> cum;
    synth_seq_table(H, &ext, &cum);
    if (ext.empty() || cum.back() == 0) throw Err(E_ARG, "index has no root intervals to place regions on");
    for (uint64_t i = 0; i < n; i++) synth_region(seed, first + i, length, cum.data(), ext.data(), (uint32_t)ext.size(), chr + i, start + i, end + i);
}

}  // namespace gffx
