// capi.cpp — the extern "C" boundary declared in include/gffx_b200.h.
#include <cstdlib>
#include <thread>
#include "../../include/gffx_b200.h"
#include "engine.h"
#include "reads.h"

#include "inflate.h"
using namespace gffx;

struct gffx_index {
    std::unique_ptr<HostIndex> host_only;  // device == -1: loaders / parsers only, no compute
    std::unique_ptr<Engine> eng_;
    const HostIndex& host() const { return eng_ ? eng_->host() : *host_only; }
    Engine* engine() const {
        if (!eng_) throw Err(E_CUDA, "index was opened host-only (device -1): no GPU engine behind it, and there is no CPU fallback");
        return eng_.get();
    }
};
struct gffx_result {
    IntersectResult r;
    gffx_index* idx = nullptr;   // the handle the job ran on (a GFFX_F_ASYNC result waits on its engine)
    int wait_rc = 0;             // status of the wait an accessor had to do on its own
};
struct gffx_counts {
    gffx_index* idx;
    int kind;
    bool finished = false;
    CountRows rows;
};

namespace {
thread_local std::string g_err;

template <class F> int guard(F&& f) {
    try {
        f();
        return GFFX_OK;
    } catch (const Err& e) {
        g_err = e.what();
        return e.code;
    } catch (const std::bad_alloc&) {
        g_err = "out of host memory";
        return GFFX_E_IO;
    } catch (const std::exception& e) {
        g_err = e.what();
        return GFFX_E_ARG;
    } catch (...) {
        g_err = "unknown error";
        return GFFX_E_ARG;
    }
}

int host_threads() {
    if (const char* s = getenv("GFFX_HOST_THREADS")) {
        int v = atoi(s);
        if (v > 0) return v;
    }
    unsigned n = std::thread::hardware_concurrency();
    return n ? (int)std::min(n, 32u) : 4;
}

uint32_t* dup_u32(const std::vector<uint32_t>& v) {
    uint32_t* p = (uint32_t*)malloc(std::max<size_t>(v.size(), 1) * 4);
    if (!p) throw std::bad_alloc();
    if (!v.empty()) memcpy(p, v.data(), v.size() * 4);
    return p;
}

void need(const void* p, const char* what) {
    if (!p) throw Err(E_ARG, std::string(what) + " must not be NULL");
}

// a GFFX_F_ASYNC result whose counters have not been read yet: wait for them now (accessors call this; gffx_result_wait reports the status)
int settle(const gffx_result* cr) {
    gffx_result* r = const_cast<gffx_result*>(cr);
    if (!r || r->r.pending_slot < 0) return r ? r->wait_rc : GFFX_E_ARG;
    r->wait_rc = guard([&] { r->idx->engine()->wait(&r->r); });
    if (r->wait_rc) { IntersectResult z; z.n_queries = r->r.n_queries; r->r = std::move(z); }
    return r->wait_rc;
}
}  // namespace

extern "C" {

const char* gffx_last_error(void) { return g_err.c_str(); }
const char* gffx_version(void) { return "gffx_b200 0.1.0 (sm_100a)"; }
int gffx_device_count(void) { return cuda_device_count(); }

int gffx_index_open(const char* gff_path, int device, gffx_index** out) {
    return guard([&] {
        need(gff_path, "gff_path");
        need(out, "out");
        auto H = std::make_unique<HostIndex>();
        H->scan_device = device;   // the line table may be scanned on the device the index is opened on (linescan.cu)
        H->open(gff_path, host_threads());
        auto h = std::make_unique<gffx_index>();
        if (device == -1) h->host_only = std::move(H);
        else h->eng_ = std::make_unique<Engine>(std::move(H), device);
        *out = h.release();
    });
}

int gffx_index_create(const gffx_index_desc* d, gffx_index** out) {
    return guard([&] {
        need(d, "desc");
        need(out, "out");
        if (d->n_seqs && !d->seq_names) throw Err(E_ARG, "seq_names is NULL");
        if (d->n_roots && (!d->root_seq || !d->root_start || !d->root_end || !d->root_fid)) throw Err(E_ARG, "root arrays are NULL");
        if (d->n_gof && (!d->gof_fid || !d->gof_start || !d->gof_end)) throw Err(E_ARG, "gof arrays are NULL");
        auto H = std::make_unique<HostIndex>();
        H->scan_device = d->device;
        H->create(d->n_seqs, d->seq_names, d->n_roots, d->root_seq, d->root_start, d->root_end, d->root_fid, d->n_gof, d->gof_fid,
                  d->gof_start, d->gof_end, d->gff_text, d->gff_len, host_threads());
        auto h = std::make_unique<gffx_index>();
        if (d->device == -1) h->host_only = std::move(H);
        else h->eng_ = std::make_unique<Engine>(std::move(H), d->device);
        *out = h.release();
    });
}

void gffx_index_destroy(gffx_index* idx) { delete idx; }
void* gffx_index_cuda_stream(const gffx_index* idx) { return idx && idx->eng_ ? idx->eng_->cuda_stream() : nullptr; }

int gffx_index_build(const char* gff_path, const char* attr_key, const char* skip_types) {
    return guard([&] {
        need(gff_path, "gff_path");
        build_index_files(gff_path, attr_key ? attr_key : "gene_name",
                          skip_types ? skip_types : "remark,note,comment,region,gap,assembly_gap,contig,scaffold,source");
    });
}

uint32_t gffx_index_n_seqs(const gffx_index* idx) { return idx ? (uint32_t)idx->host().seq_names.size() : 0; }
const char* gffx_index_seq_name(const gffx_index* idx, uint32_t seq) {
    if (!idx || seq >= idx->host().seq_names.size()) return nullptr;
    return idx->host().seq_names[seq].c_str();
}
int64_t gffx_index_seq_num(const gffx_index* idx, const char* name) {
    if (!idx || !name) return -1;
    auto it = idx->host().seq_num.find(name);
    return it == idx->host().seq_num.end() ? -1 : (int64_t)it->second;
}
uint64_t gffx_index_n_roots(const gffx_index* idx) { return idx ? idx->host().roots.size() : 0; }
uint64_t gffx_index_n_blocks(const gffx_index* idx) { return idx ? idx->host().blocks.size() : 0; }
uint64_t gffx_index_n_lines(const gffx_index* idx) { return idx ? idx->host().lines.size() : 0; }
uint64_t gffx_index_n_ids(const gffx_index* idx) { return idx ? idx->host().ids.size() : 0; }
int gffx_index_lines_info(const gffx_index* idx, uint64_t* digest, int* source) {
    return guard([&] {
        need(idx, "idx");
        if (digest) *digest = line_table_digest(idx->host());
        if (source) *source = idx->host().lines_source;
    });
}
uint64_t gffx_index_device_bytes(const gffx_index* idx) { return idx && idx->eng_ ? idx->eng_->device_bytes() : 0; }

int gffx_index_seq_roots(const gffx_index* idx, uint32_t seq, const uint32_t** triples, uint64_t* n) {
    return guard([&] {
        need(idx, "idx");
        need(triples, "triples");
        need(n, "n");
        const auto& v = idx->engine()->sorted_roots(seq);
        *triples = v.data();
        *n = v.size() / 3;
    });
}

int gffx_parse_bed_intersect(const gffx_index* idx, const char* bed_path, uint32_t** chr, uint32_t** start, uint32_t** end, uint64_t* n) {
    return guard([&] {
        need(idx, "idx"); need(bed_path, "bed_path"); need(chr, "chr"); need(start, "start"); need(end, "end"); need(n, "n");
        Regions R;
        parse_bed_intersect(bed_path, idx->host().seq_num, &R);
        *chr = dup_u32(R.chr); *start = dup_u32(R.start); *end = dup_u32(R.end);
        *n = R.size();
    });
}
int gffx_parse_bed_device(gffx_index* idx, const char* bed_path, int dialect, const uint32_t** d_chr, const uint32_t** d_start,
                          const uint32_t** d_end, uint64_t* n) {
    return guard([&] {
        need(idx, "idx"); need(bed_path, "bed_path"); need(d_chr, "d_chr"); need(d_start, "d_start"); need(d_end, "d_end"); need(n, "n");
        idx->engine()->parse_bed_device(bed_path, dialect, d_chr, d_start, d_end, n);
    });
}
double gffx_bed_parse_stat(const gffx_index* idx, int what) { return idx ? const_cast<gffx_index*>(idx)->engine()->bed_parse_stat(what) : 0; }
int gffx_parse_bed_counts(const gffx_index* idx, const char* bed_path, uint32_t** chr, uint32_t** start, uint32_t** end, uint64_t* n) {
    return guard([&] {
        need(idx, "idx"); need(bed_path, "bed_path"); need(chr, "chr"); need(start, "start"); need(end, "end"); need(n, "n");
        Regions R;
        parse_bed_counts(bed_path, idx->host().seq_num, host_threads(), &R);
        *chr = dup_u32(R.chr); *start = dup_u32(R.start); *end = dup_u32(R.end);
        *n = R.size();
    });
}
int gffx_parse_reads(const gffx_index* idx, const char* path, uint32_t** chr, uint32_t** start, uint32_t** end, uint64_t* n, uint64_t stats[5]) {
    return guard([&] {
        need(idx, "idx"); need(path, "path"); need(chr, "chr"); need(start, "start"); need(end, "end"); need(n, "n");
        Regions R;
        ReadStats st;
        read_alignments(path, idx->host().seq_num, host_threads(), &R, &st);
        *chr = dup_u32(R.chr); *start = dup_u32(R.start); *end = dup_u32(R.end);
        *n = R.size();
        if (stats) { stats[0] = st.records; stats[1] = st.unmapped; stats[2] = st.unknown_ref; stats[3] = st.skipped; stats[4] = (uint64_t)st.format; }
    });
}
int gffx_parse_region(const gffx_index* idx, const char* region, uint32_t* chr, uint32_t* start, uint32_t* end) {
    return guard([&] {
        need(idx, "idx"); need(region, "region"); need(chr, "chr"); need(start, "start"); need(end, "end");
        parse_region_arg(region, idx->host().seq_num, chr, start, end);
    });
}
void gffx_free(void* p) { free(p); }

int gffx_query_features(gffx_index* idx, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, int mode,
                        int invert, uint32_t flags, gffx_result** out) {
    return guard([&] {
        need(idx, "idx"); need(out, "out");
        auto r = std::make_unique<gffx_result>();
        r->idx = idx;
        idx->engine()->intersect(chr, start, end, n, mode, invert != 0, false, nullptr, flags, true, &r->r);
        *out = r.release();
    });
}

int gffx_intersect(gffx_index* idx, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, int mode, int invert,
                   int entire_group, const char* types_csv, uint32_t flags, gffx_result** out) {
    return guard([&] {
        need(idx, "idx"); need(out, "out");
        auto r = std::make_unique<gffx_result>();
        r->idx = idx;
        idx->engine()->intersect(chr, start, end, n, mode, invert != 0, entire_group != 0, types_csv, flags, false, &r->r);
        *out = r.release();
    });
}

int gffx_shard_begin(gffx_index* idx, int mode, int invert, int entire_group, const char* types_csv, uint32_t flags, uint64_t n_expected,
                     uint32_t batches_expected) {
    return guard([&] {
        need(idx, "idx");
        idx->engine()->shard_begin(mode, invert != 0, entire_group != 0, types_csv, flags, n_expected, batches_expected);
    });
}
int gffx_shard_add(gffx_index* idx, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags,
                   uint64_t* n_pairs_total) {
    return guard([&] {
        need(idx, "idx");
        idx->engine()->shard_add(chr, start, end, n, flags, n_pairs_total);
    });
}
int gffx_shard_finish(gffx_index* idx, gffx_result** out) {
    return guard([&] {
        need(idx, "idx"); need(out, "out");
        auto r = std::make_unique<gffx_result>();
        r->idx = idx;
        idx->engine()->shard_finish(&r->r);
        *out = r.release();
    });
}
int gffx_job_buffer(gffx_index* idx, gffx_counts* counts, uint32_t k, void** dptr, uint64_t* n_elems, int* op) {
    int past = 0;
    int rc = guard([&] {
        need(idx, "idx"); need(dptr, "dptr"); need(n_elems, "n_elems"); need(op, "op");
        if (counts && counts->idx != idx) throw Err(E_ARG, "the counts job belongs to another index");
        past = idx->engine()->shard_buffer(counts ? 1 : 0, k, dptr, n_elems, op) ? 0 : 1;
    });
    return rc ? rc : past;
}
int gffx_fold(gffx_index* idx, void* dst, const void* src, uint32_t rows, uint64_t stride_elems, uint64_t n_elems, int op) {
    return guard([&] {
        need(idx, "idx"); need(dst, "dst"); need(src, "src");
        idx->engine()->fold(dst, src, rows, stride_elems, n_elems, op);
    });
}
int64_t gffx_inflate_raw(const uint8_t* in, uint64_t in_len, uint8_t* out, uint64_t out_cap) {
    if ((!in && in_len) || (!out && out_cap)) return -1;
    static thread_local gffx::InflateTables tab;
    return (int64_t)gffx::inflate_raw(in, (size_t)in_len, out, (size_t)out_cap, tab);
}
uint32_t gffx_crc32(const uint8_t* p, uint64_t n) { return (p || n == 0) ? gffx::crc32_bytes(p, (size_t)n) : 0u; }
int gffx_index_sync(gffx_index* idx) {
    return guard([&] {
        need(idx, "idx");
        idx->engine()->sync();
    });
}
int gffx_count_pairs_noalloc(gffx_index* idx, const uint32_t* d_chr, const uint32_t* d_start, const uint32_t* d_end, uint64_t n, int mode, int invert,
                             uint64_t* d_out, void* stream) {
    return guard([&] {
        need(idx, "idx"); need(d_out, "d_out");
        if (n) { need(d_chr, "d_chr"); need(d_start, "d_start"); need(d_end, "d_end"); }
        idx->engine()->count_pairs_noalloc(d_chr, d_start, d_end, n, mode, invert != 0, d_out, stream);
    });
}
uint64_t gffx_noalloc_max_regions(void) { return gffx::Engine::noalloc_max_regions(); }
uint64_t gffx_noalloc_max_roots(void) { return gffx::Engine::noalloc_max_roots(); }

int gffx_result_wait(gffx_result* r) {
    if (!r) { g_err = "result must not be NULL"; return GFFX_E_ARG; }
    return settle(r);
}
int gffx_result_pending(const gffx_result* r) { return r && r->r.pending_slot >= 0 ? 1 : 0; }
uint64_t gffx_result_n_queries(const gffx_result* r) { return r ? r->r.n_queries : 0; }
uint64_t gffx_result_n_pairs(const gffx_result* r) { return settle(r) == 0 ? r->r.n_pairs : 0; }
const uint32_t* gffx_result_pairs(const gffx_result* r) { return settle(r) == 0 && !r->r.pairs.empty() ? r->r.pairs.data() : nullptr; }
const uint32_t* gffx_result_pairs_device(const gffx_result* r) { return settle(r) == 0 ? r->r.d_pairs : nullptr; }
uint64_t gffx_result_n_roots(const gffx_result* r) { return settle(r) == 0 ? r->r.n_roots : 0; }
uint64_t gffx_result_n_ranges(const gffx_result* r) { return settle(r) == 0 ? r->r.n_ranges : 0; }
const uint64_t* gffx_result_ranges(const gffx_result* r) { return settle(r) == 0 && !r->r.ranges.empty() ? r->r.ranges.data() : nullptr; }
uint64_t gffx_result_bytes_out(const gffx_result* r) { return settle(r) == 0 ? r->r.bytes_out : 0; }
uint64_t gffx_result_n_warn(const gffx_result* r) { return settle(r) == 0 ? r->r.warn_fids.size() : 0; }
const uint32_t* gffx_result_warn_fids(const gffx_result* r) { return settle(r) == 0 && !r->r.warn_fids.empty() ? r->r.warn_fids.data() : nullptr; }
double gffx_result_kernel_ms(const gffx_result* r) { return settle(r) == 0 ? r->r.kernel_ms : 0; }
uint32_t gffx_result_launches(const gffx_result* r) { return r ? r->r.launches : 0; }
double gffx_result_stage_ms(const gffx_result* r, int stage) { return (settle(r) == 0 && stage >= 0 && stage < 12) ? r->r.stage_ms[stage] : 0; }
int gffx_result_get_info(const gffx_result* r, gffx_result_info* out) {
    if (!r || !out) return GFFX_E_ARG;
    if (int rc = settle(r)) return rc;
    const auto& R = r->r;
    out->n_queries = R.n_queries; out->n_pairs = R.n_pairs; out->n_roots = R.n_roots; out->n_ranges = R.n_ranges;
    out->bytes_out = R.bytes_out; out->n_warn = R.warn_fids.size();
    out->pairs = gffx_result_pairs(r); out->pairs_device = gffx_result_pairs_device(r);
    out->ranges = gffx_result_ranges(r); out->warn_fids = gffx_result_warn_fids(r);
    out->kernel_ms = R.kernel_ms;
    for (int k = 0; k < 12; k++) out->stage_ms[k] = R.stage_ms[k];
    out->launches = R.launches; out->reserved = 0;
    return GFFX_OK;
}
void gffx_result_free(gffx_result* r) {
    if (r && r->r.pending_slot >= 0 && r->idx) r->idx->engine()->abandon(&r->r);   // never waited for: its slot frees itself when the copy has landed
    delete r;
}

int gffx_write_ranges(const gffx_index* idx, const uint64_t* ranges, uint64_t n_ranges, const char* out_path) {
    return guard([&] {
        need(idx, "idx");
        if (n_ranges) need(ranges, "ranges");
        const HostIndex& H = idx->host();
        if (!H.gff && n_ranges) throw Err(E_STATE, "index was opened without the GFF text");
        write_ranges_to(H.gff, H.gff_len, ranges, n_ranges, out_path);
    });
}

int gffx_gather_text(gffx_index* idx, const uint64_t* ranges, uint64_t n_ranges, uint32_t flags, const uint8_t** text, uint64_t* n_bytes) {
    return guard([&] {
        need(idx, "idx"); need(text, "text"); need(n_bytes, "n_bytes");
        idx->engine()->gather_text(ranges, n_ranges, flags, text, n_bytes);
    });
}
double gffx_gather_kernel_ms(const gffx_index* idx, int part) {
    if (!idx) return 0;
    return part == 0 ? const_cast<gffx_index*>(idx)->engine()->gather_kernel_ms() : const_cast<gffx_index*>(idx)->engine()->gather_scan_ms();
}
int gffx_write_text(const uint8_t* text, uint64_t n_bytes, const char* out_path) {
    return guard([&] {
        if (n_bytes) need(text, "text");
        write_text_to(text, n_bytes, out_path);
    });
}

int gffx_counts_begin(gffx_index* idx, int kind, gffx_counts** out) {
    return guard([&] {
        need(idx, "idx"); need(out, "out");
        if (kind != 0 && kind != 1) throw Err(E_ARG, "kind must be 0 (depth) or 1 (coverage)");
        idx->engine()->counts_begin(kind);
        auto c = std::make_unique<gffx_counts>();
        c->idx = idx;
        c->kind = kind;
        *out = c.release();
    });
}
int gffx_counts_add(gffx_counts* st, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags) {
    return guard([&] {
        need(st, "state");
        if (st->finished) throw Err(E_STATE, "gffx_counts_add after gffx_counts_finish");
        st->idx->engine()->counts_add(chr, start, end, n, flags);
    });
}
int gffx_counts_finish(gffx_counts* st, uint64_t* n_rows, const uint32_t** id_idx, const uint32_t** chrom_idx, const uint32_t** start,
                       const uint32_t** end, const uint64_t** value) {
    return guard([&] {
        need(st, "state");
        if (!st->finished) {
            st->idx->engine()->counts_finish(&st->rows);
            st->finished = true;
        }
        if (n_rows) *n_rows = st->rows.id_idx.size();
        if (id_idx) *id_idx = st->rows.id_idx.data();
        if (chrom_idx) *chrom_idx = st->rows.chrom_idx.data();
        if (start) *start = st->rows.start.data();
        if (end) *end = st->rows.end.data();
        if (value) *value = st->rows.value.data();
    });
}
double gffx_counts_kernel_ms(const gffx_counts* st) { return st ? st->idx->engine()->counts_kernel_ms() : 0; }
uint32_t gffx_counts_launches(const gffx_counts* st) { return st ? st->idx->engine()->counts_launches() : 0; }
void gffx_counts_free(gffx_counts* st) { delete st; }

const char* gffx_index_id_string(const gffx_index* idx, uint32_t i, uint32_t* len) {
    if (!idx || i >= idx->host().ids.size()) return nullptr;
    std::string_view s = idx->host().ids.get(i);
    if (len) *len = (uint32_t)s.size();
    return s.data();
}
const char* gffx_index_chrom_string(const gffx_index* idx, uint32_t i, uint32_t* len) {
    if (!idx || i >= idx->host().chroms.size()) return nullptr;
    std::string_view s = idx->host().chroms.get(i);
    if (len) *len = (uint32_t)s.size();
    return s.data();
}

// write_depth_results (depth.rs:515-545) / write_breadth_results (coverage.rs:453-490); rows in ID
// dictionary order (the reference's order is hash order, i.e. unspecified)
int gffx_counts_write(gffx_counts* st, const char* out_path) {
    return guard([&] {
        need(st, "state");
        if (!st->finished) {
            st->idx->engine()->counts_finish(&st->rows);
            st->finished = true;
        }
        FILE* f = out_path ? fopen(out_path, "wb") : stdout;
        if (!f) throw Err(E_IO, std::string("cannot create output file ") + out_path);
        const HostIndex& H = st->idx->host();
        std::string buf;
        buf.reserve(1 << 22);
        buf += st->kind == 0 ? "id\tchr\tstart\tend\tdepth\n" : "id\tchr\tstart\tend\tbreadth\tfraction\n";
        char num[96];
        const CountRows& R = st->rows;
        for (size_t i = 0; i < R.id_idx.size(); i++) {
            buf += H.ids.get(R.id_idx[i]);
            buf += '\t';
            buf += H.chroms.get(R.chrom_idx[i]);
            if (st->kind == 0) {
                snprintf(num, sizeof num, "\t%u\t%u\t%llu\n", R.start[i], R.end[i], (unsigned long long)R.value[i]);
            } else {
                const uint64_t length = R.end[i] > R.start[i] ? R.end[i] - R.start[i] : 0;  // end.saturating_sub(start)
                const double fraction = length > 0 ? (double)R.value[i] / (double)length : 0.0;
                snprintf(num, sizeof num, "\t%u\t%u\t%llu\t%.6f\n", R.start[i], R.end[i], (unsigned long long)R.value[i], fraction);
            }
            buf += num;
            if (buf.size() > (1u << 22)) { fwrite(buf.data(), 1, buf.size(), f); buf.clear(); }
        }
        fwrite(buf.data(), 1, buf.size(), f);
        if (out_path) fclose(f); else fflush(f);
    });
}

// ---- command level -------------------------------------------------------------------------
int gffx_intersect_run(const gffx_intersect_args* a) {
    return guard([&] {
        need(a, "args"); need(a->input, "input");
        if ((a->bed != nullptr) == (a->region != nullptr)) throw Err(E_ARG, "exactly one of --region / --bed is required");
        gffx_index* idx = nullptr;
        if (int rc = gffx_index_open(a->input, a->device, &idx)) throw Err(rc, g_err);
        std::unique_ptr<gffx_index> hold(idx);
        Regions R;
        const uint32_t *qc = nullptr, *qs = nullptr, *qe = nullptr;
        uint64_t nq = 0;
        uint32_t qflags = 0;
        if (a->bed && !getenv("GFFX_HOST_BED")) {   // K9: the BED text is parsed on the device, the regions stay there
            idx->engine()->parse_bed_device(a->bed, 0, &qc, &qs, &qe, &nq);
            qflags = GFFX_F_DEVICE_INPUT;
        } else {
            if (a->bed) parse_bed_intersect(a->bed, idx->host().seq_num, &R);
            else {
                uint32_t c, s, e;
                parse_region_arg(a->region, idx->host().seq_num, &c, &s, &e);
                R.chr.push_back(c); R.start.push_back(s); R.end.push_back(e);
            }
            qc = R.chr.data(); qs = R.start.data(); qe = R.end.data(); nq = R.size();
        }
        if (a->verbose) fprintf(stderr, "[DEBUG] Starting query_features with %llu regions\n", (unsigned long long)nq);
        IntersectResult res;
        idx->engine()->intersect(qc, qs, qe, nq, a->mode, a->invert != 0, a->entire_group != 0, a->types, qflags, false, &res);
        for (uint32_t fid : res.warn_fids) fprintf(stderr, "[WARN] skipped fid=%u due to sentinel start offset\n", fid);
        const HostIndex& H = idx->host();
        if (!H.gff) throw Err(E_IO, std::string("Cannot open GFF: \"") + a->input + "\"");
        if (getenv("GFFX_HOST_EMIT")) {   // the reference's way: one slice per range from the mapped file
            write_ranges_to(H.gff, H.gff_len, res.ranges.data(), res.n_ranges, a->output);
        } else {                          // K8: gathered on the device, one write
            const uint8_t* text = nullptr;
            uint64_t nb = 0;
            idx->engine()->gather_text(nullptr, res.n_ranges, 0, &text, &nb);
            write_text_to(text, nb, a->output);
        }
        if (a->verbose)
            fprintf(stderr, "[INFO] gpu intersect: %llu regions, %llu pairs, %llu roots, %llu ranges, %.3f ms on device\n",
                    (unsigned long long)res.n_queries, (unsigned long long)res.n_pairs, (unsigned long long)res.n_roots,
                    (unsigned long long)res.n_ranges, res.kernel_ms);
    });
}

static int counts_run(const gffx_counts_args* a, int kind) {
    return guard([&] {
        need(a, "args"); need(a->input, "input"); need(a->source, "source");
        std::string src = a->source, ext;
        size_t dot = src.rfind('.');
        if (dot != std::string::npos) ext = src.substr(dot + 1);
        for (auto& c : ext) c = (char)tolower(c);
        const bool reads = ext == "bam" || ext == "sam" || ext == "cram";   // depth.rs:589: one reader for the three
        if (!reads && ext != "bed") throw Err(E_ARG, "Unsupported file type: \"" + src + "\". Expected .bam/.sam/.cram or .bed");
        gffx_index* idx = nullptr;
        if (int rc = gffx_index_open(a->input, a->device, &idx)) throw Err(rc, g_err);
        std::unique_ptr<gffx_index> hold(idx);
        if (!idx->host().gff) throw Err(E_IO, std::string("Cannot open GFF file: \"") + a->input + "\"");
        Regions R;
        const uint32_t *qc = nullptr, *qs = nullptr, *qe = nullptr;
        uint64_t nq = 0;
        uint32_t qflags = 0;
        if (reads) {
            ReadStats rs;
            read_alignments(src, idx->host().seq_num, a->threads > 0 ? a->threads : host_threads(), &R, &rs);
            qc = R.chr.data(); qs = R.start.data(); qe = R.end.data(); nq = R.size();
            if (a->verbose)
                fprintf(stderr, "[INFO] %llu alignment records: %llu kept, %llu unmapped, %llu on sequences the index lacks, %llu without a usable interval\n",
                        (unsigned long long)rs.records, (unsigned long long)nq, (unsigned long long)rs.unmapped, (unsigned long long)rs.unknown_ref,
                        (unsigned long long)rs.skipped);
        } else if (!getenv("GFFX_HOST_BED")) {   // K9: parsed on the device, the regions stay there
            idx->engine()->parse_bed_device(src, 1, &qc, &qs, &qe, &nq);
            qflags = GFFX_F_DEVICE_INPUT;
        } else {
            parse_bed_counts(src, idx->host().seq_num, a->threads > 0 ? a->threads : host_threads(), &R);
            qc = R.chr.data(); qs = R.start.data(); qe = R.end.data(); nq = R.size();
        }
        gffx_counts* st = nullptr;
        if (int rc = gffx_counts_begin(idx, kind, &st)) throw Err(rc, g_err);
        std::unique_ptr<gffx_counts> hold2(st);
        if (int rc = gffx_counts_add(st, qc, qs, qe, nq, qflags)) throw Err(rc, g_err);
        if (int rc = gffx_counts_write(st, a->output)) throw Err(rc, g_err);
        if (a->verbose)
            fprintf(stderr, "[INFO] gpu %s: %llu regions, %.3f ms on device\n", kind ? "coverage" : "depth", (unsigned long long)nq, gffx_counts_kernel_ms(st));
    });
}
int gffx_depth_run(const gffx_counts_args* a) { return counts_run(a, 0); }
int gffx_coverage_run(const gffx_counts_args* a) { return counts_run(a, 1); }

// ---- synthetic workloads -------------------------------------------------------------------
int gffx_synth_gff(const char* path, uint64_t seed, uint32_t n_genes, int quirks, uint64_t* n_lines) {
    return guard([&] {
        need(path, "path");
        uint64_t n = synth_gff_file(path, seed, n_genes, quirks != 0);
        if (n_lines) *n_lines = n;
    });
}
int gffx_synth_gff_scaffolds(const char* path, uint64_t seed, uint32_t n_genes, int quirks, uint32_t n_scaffolds, uint64_t* n_lines) {
    return guard([&] {
        need(path, "path");
        uint64_t n = synth_gff_file(path, seed, n_genes, quirks != 0, n_scaffolds);
        if (n_lines) *n_lines = n;
    });
}
int gffx_index_vseq(const gffx_index* idx, uint32_t seq, uint32_t* vseq, uint32_t* base, uint32_t* clamp, uint32_t* n_vseqs) {
    return guard([&] {
        need(idx, "idx");
        const VSeqMap& V = idx->host().vmap;
        if (seq >= V.seq.size()) throw Err(E_ARG, "sequence number out of range");
        if (vseq) *vseq = V.seq[seq].vseq;
        if (base) *base = V.seq[seq].base;
        if (clamp) *clamp = V.seq[seq].clamp;
        if (n_vseqs) *n_vseqs = (uint32_t)V.n_vseqs();
    });
}
int gffx_synth_regions_host(const gffx_index* idx, uint64_t seed, uint64_t first, uint64_t n, uint32_t length, uint32_t* chr,
                            uint32_t* start, uint32_t* end) {
    return guard([&] {
        need(idx, "idx");
        if (n) { need(chr, "chr"); need(start, "start"); need(end, "end"); }
        synth_regions_host(idx->host(), seed, first, n, length, chr, start, end);
    });
}
int gffx_synth_regions_device(gffx_index* idx, uint64_t seed, uint64_t first, uint64_t n, uint32_t length, uint32_t* d_chr,
                              uint32_t* d_start, uint32_t* d_end) {
    return guard([&] {
        need(idx, "idx");
        if (n) { need(d_chr, "d_chr"); need(d_start, "d_start"); need(d_end, "d_end"); }
        idx->engine()->synth_regions_device(seed, first, n, length, d_chr, d_start, d_end);
    });
}

}  // extern "C"
