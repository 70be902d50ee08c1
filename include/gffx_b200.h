/* gffx_b200.h — C ABI of the B200-native interval-overlap engine for GFFx's hot path.
 *
 * The reference (pure Rust, /root/reference) has no FFI; the path is reached through its CLI and
 * the public functions re-exported in src/lib.rs:7-22.  Each entry point below names the reference
 * item it replaces (file:line relative to the reference root).  INTEGRATION.md shows the Rust
 * `extern "C"` block + safe wrappers a maintainer adds to call them from `run()`.
 *
 * Conventions
 *   - every function returns 0 on success and a negative gffx_status on failure;
 *     gffx_last_error() then returns a thread-local, NUL-terminated message (the reference's
 *     anyhow::Error text bubbling to main, src/main.rs:28-43).  Nothing throws across the ABI.
 *   - inputs are borrowed (the library copies what it keeps); outputs are library-owned and
 *     released by the matching *_free / *_destroy call.
 *   - a handle may be used from one thread at a time; distinct handles (e.g. one per GPU) are
 *     independent.  CUDA streams are internal.
 *   - there is NO CPU fallback: every compute entry point fails with GFFX_E_CUDA when no
 *     sm_100-class device is usable.
 */
#ifndef GFFX_B200_H
#define GFFX_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    GFFX_OK = 0,
    GFFX_E_ARG = -1,    /* bad argument (NULL, bad mode, chromosome number out of range …) */
    GFFX_E_IO = -2,     /* file open / mmap / write failure                                 */
    GFFX_E_FORMAT = -3, /* corrupt index file, BED number parse error (intersect dialect)   */
    GFFX_E_CUDA = -4,   /* no usable GPU, CUDA runtime error, out of device memory          */
    GFFX_E_STATE = -5   /* call sequence error                                              */
} gffx_status;

/* OverlapMode, src/commands/intersect.rs:74-78 (CLI -c / -C / -O). */
typedef enum { GFFX_MODE_CONTAINED = 0, GFFX_MODE_CONTAINS_REGION = 1, GFFX_MODE_OVERLAP = 2 } gffx_mode;

/* flags for the compute calls */
enum {
    GFFX_F_DEVICE_INPUT = 1u << 0, /* chr/start/end are device pointers on the handle's GPU (HBM-resident queries) */
    GFFX_F_WANT_PAIRS = 1u << 1,   /* materialise the (root_fid,start,end) pairs and copy them to the host         */
    GFFX_F_NO_HOST_COPY = 1u << 2, /* leave pairs / ranges on the device (bench: kernel-only timing)               */
    GFFX_F_PAIRS_ON_DEVICE = 1u << 3, /* materialise pairs in HBM but do not copy them to the host                 */
    GFFX_F_STAGE_TIMES = 1u << 4,  /* fill gffx_result_stage_ms: a dozen CUDA events inside the call, ~40 us of it  */
    GFFX_F_ASYNC = 1u << 5         /* queue the job and return: see gffx_result_wait                                */
};

typedef struct gffx_index gffx_index;     /* device-resident index: TreeIndexData + GofMap + GFF line table */
typedef struct gffx_result gffx_result;   /* result of one intersect / query_features call                 */
typedef struct gffx_counts gffx_counts;   /* accumulation state of a depth or coverage job                 */

const char* gffx_last_error(void);
const char* gffx_version(void);
int gffx_device_count(void);

/* ------------------------------------------------------------------------------------------
 * Index: replaces TreeIndexData::load_tree_index (src/utils/tree_index.rs:21-82), load_sqs
 * (src/index_loader/core.rs:19-34), load_gof (src/index_loader/gof.rs:95-128) and the GFF mmap
 * (src/commands/intersect.rs:242-249, src/commands/depth.rs:563-566); builds the flattened
 * struct-of-arrays and uploads them to `device`.
 * ------------------------------------------------------------------------------------------ */
/* device -1 opens the index host-only (loaders, BED parsers, accessors; compute calls fail) */
int gffx_index_open(const char* gff_path, int device, gffx_index** out);

/* Same, from what a Rust caller already holds in memory (TreeIndexData, GofMap, &Mmap). */
typedef struct {
    uint32_t n_seqs;
    const char* const* seq_names;   /* .sqs order = seq number (TreeIndexData.num_to_seqid)        */
    uint64_t n_roots;               /* all tree intervals, any order                               */
    const uint32_t* root_seq;       /* seq number of each interval (tree it came from)             */
    const uint32_t* root_start;     /* Interval.start (0-based)            src/utils/tree.rs:6-10  */
    const uint32_t* root_end;       /* Interval.end   (exclusive)                                  */
    const uint32_t* root_fid;       /* Interval.root_fid                                           */
    uint64_t n_gof;                 /* .gof records in file order (last duplicate fid wins)        */
    const uint32_t* gof_fid;
    const uint64_t* gof_start;
    const uint64_t* gof_end;
    const uint8_t* gff_text;        /* the mmap'd GFF (may be NULL with gff_len 0: join-only use)  */
    uint64_t gff_len;
    int device;
} gffx_index_desc;
int gffx_index_create(const gffx_index_desc* desc, gffx_index** out);
void gffx_index_destroy(gffx_index* idx);

/* The engine's CUDA stream (a cudaStream_t) so that a caller can record its own events around calls;
 * NULL for a host-only index. */
void* gffx_index_cuda_stream(const gffx_index* idx);

/* `gffx index` (src/commands/index.rs:25-37 -> build_index, src/index_builder/core.rs:41-242): writes
 * .fts .prt .a2f .atn .sqs .gof .rit .rix beside the GFF, byte-compatible with the reference's files.
 * attr_key NULL = "gene_name", skip_types NULL = the reference's default list (src/commands/index.rs:18). */
int gffx_index_build(const char* gff_path, const char* attr_key, const char* skip_types);

uint32_t gffx_index_n_seqs(const gffx_index* idx);
const char* gffx_index_seq_name(const gffx_index* idx, uint32_t seq);
/* seqid_to_num lookup (TreeIndexData.seqid_to_num); returns -1 when unknown */
int64_t gffx_index_seq_num(const gffx_index* idx, const char* name);
uint64_t gffx_index_n_roots(const gffx_index* idx);
uint64_t gffx_index_n_blocks(const gffx_index* idx);
uint64_t gffx_index_n_lines(const gffx_index* idx);
uint64_t gffx_index_n_ids(const gffx_index* idx);
/* The GFF line table (what the reference re-parses per run: src/commands/intersect.rs:284-321,
 * src/commands/depth.rs:131-157): a digest over every column, the type / chromosome / ID dictionaries and the per-block facts,
 * and where the table came from: 0 the host scanner, 1 the device scanner, 2 the side-car cache `<gff>.b200cache` (written by
 * the first open of a GFF, keyed by its size + mtime + the block list; GFFX_CACHE=0 disables it, GFFX_LINE_SCAN=host|device
 * picks the scanner).  Equal digests = equal tables. */
int gffx_index_lines_info(const gffx_index* idx, uint64_t* digest, int* source);
uint64_t gffx_index_device_bytes(const gffx_index* idx);
/* Where sequence `seq` sits in the engine's coordinate system (diagnostic).  TreeIndexData holds one tree per sequence
 * with no limit on their number (src/utils/tree_index.rs:36-82); the engine lays the sequences side by side on at most
 * a few hundred VIRTUAL sequences: a region (seq, s, e) is processed as (vseq, base + min(s, clamp), base + min(e, clamp)).
 * clamp exceeds every coordinate of the sequence a region is ever compared with, so results are unchanged; pairs and
 * depth / coverage rows come back in the sequence's own coordinates.  Any out pointer may be NULL. */
int gffx_index_vseq(const gffx_index* idx, uint32_t seq, uint32_t* vseq, uint32_t* base, uint32_t* clamp, uint32_t* n_vseqs);
/* flattened tree intervals of one sequence as (start,end,root_fid) triples, sorted by start */
int gffx_index_seq_roots(const gffx_index* idx, uint32_t seq, const uint32_t** triples, uint64_t* n);

/* ------------------------------------------------------------------------------------------
 * BED parsing on the host: the three dialects of the reference.
 *   intersect: parse_bed_file, src/commands/intersect.rs:201-230 (number errors abort: GFFX_E_FORMAT)
 *   counts   : src/commands/depth.rs:449-495 == src/commands/coverage.rs:222-255 (bad lines dropped)
 * Output arrays are library-owned; free with gffx_free().
 * ------------------------------------------------------------------------------------------ */
int gffx_parse_bed_intersect(const gffx_index* idx, const char* bed_path, uint32_t** chr, uint32_t** start, uint32_t** end,
                             uint64_t* n);
int gffx_parse_bed_counts(const gffx_index* idx, const char* bed_path, uint32_t** chr, uint32_t** start, uint32_t** end,
                          uint64_t* n);
/* The same two dialects parsed on the device (bedparse.cuh): the file's text goes to HBM, one thread per
 * line decides the plain lines, the host's exact per-line functions the rest (non-ASCII, signs, overlong numbers...), so
 * the result — and dialect 0's abort on a bad number — is that of the host parsers above.  dialect: 0 intersect, 1 counts.
 * The arrays are DEVICE pointers owned by the handle until its next parse: pass them to gffx_intersect / gffx_counts_add
 * with GFFX_F_DEVICE_INPUT.  gffx_bed_parse_stat: 0 kernel ms, 1 launches, 2 lines left to the host, 3 bytes of text. */
int gffx_parse_bed_device(gffx_index* idx, const char* bed_path, int dialect, const uint32_t** d_chr, const uint32_t** d_start,
                          const uint32_t** d_end, uint64_t* n);
double gffx_bed_parse_stat(const gffx_index* idx, int what);
/* Alignment files as the read source of depth / coverage: SAM text, BAM, bgzip-ed or gzip-ed SAM (decided by content, as
 * hts_open does).  Replaces what process_bam (src/commands/depth.rs:297-423) and collect_by_root_from_bam
 * (src/commands/coverage.rs:125-205) get from rust-htslib: one (sequence number, pos, reference_end) triple per record, with
 * the skips of depth.rs:335-366 (unmapped, no target, a target name the index lacks, pos < 0, end <= pos), clamped to u32, in
 * file order.  BGZF blocks are inflated by the library's own decoder on all host threads, CRC-32 checked.  CRAM: GFFX_E_ARG.
 * stats (may be NULL): [0] records seen, [1] unmapped / no target, [2] unknown target name, [3] pos < 0 or end <= pos,
 * [4] container (0 text, 1 BGZF, 2 gzip).  Output arrays are library-owned; free with gffx_free(). */
int gffx_parse_reads(const gffx_index* idx, const char* path, uint32_t** chr, uint32_t** start, uint32_t** end, uint64_t* n,
                     uint64_t stats[5]);
/* parse_region, src/commands/intersect.rs:172-198 */
/* The reader's own DEFLATE decoder and CRC-32 (inflate.h; the reference gets both from rust-htslib -> htslib -> zlib), exported
 * for tests and tools: one raw DEFLATE stream (a BGZF block's payload) -> out.  Returns the bytes written, -1 for a malformed
 * stream, -2 when out_cap is too small. */
int64_t gffx_inflate_raw(const uint8_t* in, uint64_t in_len, uint8_t* out, uint64_t out_cap);
uint32_t gffx_crc32(const uint8_t* p, uint64_t n);
int gffx_parse_region(const gffx_index* idx, const char* region, uint32_t* chr, uint32_t* start, uint32_t* end);
void gffx_free(void* p);

/* ------------------------------------------------------------------------------------------
 * Root-level join: replaces query_features (src/commands/intersect.rs:105-169) and the
 * IntervalTree::query_interval calls under it (src/utils/tree.rs:98-121).
 * A pair names the root only — (iv.root_fid, iv.start, iv.end), intersect.rs:162 — so the result is the multiset
 * "root r, once per region that hits r".  Pairs come back grouped by root, roots in (sequence, start, end) order; the
 * reference's order (chromosomes in hash-map order, regions in BED order) is not reproduced, nothing downstream of
 * query_features depends on it (intersect.rs:598-617 only dedups the roots).
 * ------------------------------------------------------------------------------------------ */
int gffx_query_features(gffx_index* idx, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n,
                        int mode, int invert, uint32_t flags, gffx_result** out);

/* ------------------------------------------------------------------------------------------
 * Full intersect: replaces the body of run() after BED parsing (src/commands/intersect.rs:586-654):
 * query_features, root dedup + GofMap::roots_to_offsets (src/index_loader/gof.rs:54-84), then
 * write_gff_match_only_by_coords (src/commands/intersect.rs:232-438) when !entire_group || types,
 * else write_gff_output's sort+merge (src/utils/common.rs:199-229).  The result holds the byte
 * ranges [start,end) of the GFF to write, in output order.
 * `types_csv` = the -T value or NULL.
 * ------------------------------------------------------------------------------------------ */
int gffx_intersect(gffx_index* idx, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, int mode,
                   int invert, int entire_group, const char* types_csv, uint32_t flags, gffx_result** out);

/* Asynchronous jobs.  The reference's run() is one synchronous pass (src/commands/intersect.rs:586-654); a caller that has
 * several passes to run (a mode sweep over resident BED sets, the chunks of a huge BED, the ranks of a sharded job) need not
 * pay a host round trip between them: with GFFX_F_ASYNC (together with GFFX_F_DEVICE_INPUT | GFFX_F_NO_HOST_COPY, on
 * gffx_query_features / gffx_intersect / gffx_shard_begin) the call returns as soon as the job is QUEUED on the engine's
 * stream, the 64 bytes of counters travel to a page-locked slot behind it, and the next job can be queued at once.
 * gffx_result_wait blocks until the counters are there (and reports the job's status, e.g. a bad sequence number); every
 * gffx_result_* accessor waits by itself when needed.  Queued jobs rotate over three LANES (three sets of the per-job
 * device state, three streams), so up to three of them run side by side on the GPU: the region arrays of a GFFX_F_ASYNC job must be
 * complete when the call is made (not merely queued on the engine's stream) and stay untouched until the wait.  Device
 * outputs (ranges for gffx_gather_text, pairs) stay valid until the handle's next job on the same lane, i.e. at least until the
 * next job is queued.  Between gffx_shard_begin and gffx_shard_finish of a queued job gffx_index_cuda_stream reports the
 * job's lane.  If the pairs outgrow the buffer
 * sized from earlier jobs, gffx_result_n_pairs is exact but gffx_result_pairs_device is NULL (run that pass synchronously).
 * At most 32 results may be pending per handle; gffx_result_free on a pending result abandons it.
 * Host arrays work too (r02): without GFFX_F_DEVICE_INPUT the regions are copied from the caller's arrays behind the call —
 * they must stay valid and unchanged until the wait (page-locked arrays make the copy asynchronous) — and without
 * GFFX_F_NO_HOST_COPY the host copy of the ranges follows the job into the result's page-locked buffer, trimmed at the wait;
 * with two jobs in flight one job's ranges travel to the host while the next job's regions travel to the device.  Pairs are
 * never copied to the host behind a queued job (GFFX_F_WANT_PAIRS is refused). */
int gffx_result_wait(gffx_result* r);
int gffx_result_pending(const gffx_result* r);   /* 1 while the counters of a GFFX_F_ASYNC job have not been read */
uint64_t gffx_result_n_queries(const gffx_result* r);
uint64_t gffx_result_n_pairs(const gffx_result* r);
/* n_pairs (root_fid, start, end) u32 triples on the host; NULL unless GFFX_F_WANT_PAIRS */
const uint32_t* gffx_result_pairs(const gffx_result* r);
const uint32_t* gffx_result_pairs_device(const gffx_result* r);
uint64_t gffx_result_n_roots(const gffx_result* r);   /* distinct selected root fids */
uint64_t gffx_result_n_ranges(const gffx_result* r);
/* n_ranges (byte_start, byte_end) u64 pairs on the host, output order */
const uint64_t* gffx_result_ranges(const gffx_result* r);
uint64_t gffx_result_bytes_out(const gffx_result* r);
/* fids skipped for a sentinel offset (the reference prints "[WARN] skipped fid=…", intersect.rs:269-272) */
uint64_t gffx_result_n_warn(const gffx_result* r);
const uint32_t* gffx_result_warn_fids(const gffx_result* r);
/* device time of the call's kernels in ms (CUDA events on the engine's stream) and launch count */
double gffx_result_kernel_ms(const gffx_result* r);
uint32_t gffx_result_launches(const gffx_result* r);
/* per-stage device ms (all 0 unless the call carried GFFX_F_STAGE_TIMES; gffx_result_kernel_ms is always there): 0 root join (hits per root) with the region scatter onto the line rank space fused in,
 * 3 root/block select + merge, 4 (unused since the scatter rides on stage 0), 5 rank scans, 6 line filter + compaction,
 * 7 pair offsets scan + pair expansion, 8 the window partition pass (totals + offsets + scatter);
 * stages that did not run report 0 */
double gffx_result_stage_ms(const gffx_result* r, int stage);
/* everything above in one call (the accessors stay): what run() reads after the join, src/commands/intersect.rs:596-654 */
typedef struct gffx_result_info {
    uint64_t n_queries, n_pairs, n_roots, n_ranges, bytes_out, n_warn;
    const uint32_t* pairs;        /* as gffx_result_pairs */
    const uint32_t* pairs_device; /* as gffx_result_pairs_device */
    const uint64_t* ranges;       /* as gffx_result_ranges */
    const uint32_t* warn_fids;    /* as gffx_result_warn_fids */
    double kernel_ms;
    double stage_ms[12];
    uint32_t launches;
    uint32_t reserved;
} gffx_result_info;
int gffx_result_get_info(const gffx_result* r, gffx_result_info* out);
void gffx_result_free(gffx_result* r);

/* Write the selected byte ranges of the GFF (write_vectored loops of intersect.rs:385-429 and
 * common.rs:231-278). out_path NULL = stdout. */
int gffx_write_ranges(const gffx_index* idx, const uint64_t* ranges, uint64_t n_ranges, const char* out_path);

/* Record emission on the device: the same bytes the write_vectored loops above produce — the
 * concatenation of text[start,end) over the ranges, skipping start >= end and end > file length — gathered from the
 * HBM-resident GFF text (uploaded on first use) into ONE buffer, so the host issues a single write.
 * ranges: n_ranges x (start,end) u64 on the host, or NULL = the ranges of this handle's last finished intersect, still
 * on the device (n_ranges must then be that result's gffx_result_n_ranges).
 * flags & GFFX_F_NO_HOST_COPY: *text is a device pointer, else a page-locked host copy; both stay valid until the
 * handle's next gather.  gffx_gather_kernel_ms: device time of the gather kernel alone / of its offsets scan. */
int gffx_gather_text(gffx_index* idx, const uint64_t* ranges, uint64_t n_ranges, uint32_t flags, const uint8_t** text,
                     uint64_t* n_bytes);
double gffx_gather_kernel_ms(const gffx_index* idx, int part /* 0 gather kernel, 1 offsets scan */);
/* one write(2) loop over a gathered buffer; out_path NULL = stdout */
int gffx_write_text(const uint8_t* text, uint64_t n_bytes, const char* out_path);

/* ------------------------------------------------------------------------------------------
 * depth / coverage: replace compute_hit_depth + compute_root_depth (src/commands/depth.rs:121-293)
 * and collect_by_root_* + finalize_compute_breadth (src/commands/coverage.rs:208-450).
 * kind: 0 = depth, 1 = coverage (breadth).  *_add may be called once per batch of regions
 * (the reference batches 100k BED lines, depth.rs:24,462); regions with start >= end are ignored
 * as both BED dialects drop them before the engine is reached.
 * ------------------------------------------------------------------------------------------ */
int gffx_counts_begin(gffx_index* idx, int kind, gffx_counts** out);
int gffx_counts_add(gffx_counts* st, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n,
                    uint32_t flags);
/* rows: parallel arrays, library-owned until gffx_counts_free; order = ID dictionary order */
int gffx_counts_finish(gffx_counts* st, uint64_t* n_rows, const uint32_t** id_idx, const uint32_t** chrom_idx,
                       const uint32_t** start, const uint32_t** end, const uint64_t** value);
double gffx_counts_kernel_ms(const gffx_counts* st);
uint32_t gffx_counts_launches(const gffx_counts* st);
void gffx_counts_free(gffx_counts* st);
const char* gffx_index_id_string(const gffx_index* idx, uint32_t id_idx, uint32_t* len);
const char* gffx_index_chrom_string(const gffx_index* idx, uint32_t chrom_idx, uint32_t* len);
/* write_depth_results (depth.rs:515-545) / write_breadth_results (coverage.rs:453-490) */
int gffx_counts_write(gffx_counts* st, const char* out_path);

/* ------------------------------------------------------------------------------------------
 * Batched / sharded jobs.  The reference runs one process over one BED (intersect.rs:586-654) and batches only in
 * depth (100k lines, depth.rs:24,462).  Here the same job can take its regions in any number of batches and can be
 * spread over several GPUs (one handle per GPU, each fed its own share of the regions):
 *     gffx_shard_begin -> gffx_shard_add ... -> [combine] -> gffx_shard_finish
 * [combine], multi-GPU only: for every accumulator k reported by gffx_job_buffer, exchange the shards' copies and
 * reduce them element-wise with the reported op (gffx_fold is the reduction kernel; gffx_b200/shard.py drives it over
 * torch.distributed: all-to-all of 1/W slices, fold, all-gather).  Every shard then finishes with the job's global
 * state and produces the complete result.  The pairs of a shard's own regions are materialised (GFFX_F_WANT_PAIRS /
 * GFFX_F_PAIRS_ON_DEVICE) only when the job announces its size: n_expected regions in at most batches_expected adds
 * (the bucketed batches are then kept on the device); n_expected = 0 is a streaming job that only counts them.
 * ------------------------------------------------------------------------------------------ */
typedef enum {
    GFFX_RED_MAX_U8 = 0, GFFX_RED_MAX_U64 = 1, GFFX_RED_MIN_U64 = 2, GFFX_RED_SUM_U32 = 3, GFFX_RED_SUM_I32 = 4, GFFX_RED_OR_U64 = 5
} gffx_reduce_op;
int gffx_shard_begin(gffx_index* idx, int mode, int invert, int entire_group, const char* types_csv, uint32_t flags,
                     uint64_t n_expected, uint32_t batches_expected);
/* n_pairs_total (may be NULL): pairs counted so far by this shard.  The arrays may be reused when the call returns — except in a
 * GFFX_F_ASYNC job, where host arrays are copied behind the call (n_pairs_total must be NULL) and stay the job's until the wait. */
int gffx_shard_add(gffx_index* idx, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags,
                   uint64_t* n_pairs_total);
int gffx_shard_finish(gffx_index* idx, gffx_result** out);
/* accumulator k of the open intersect job (counts == NULL) or of a depth / coverage job: device pointer, element count,
 * gffx_reduce_op.  Returns 1 when k is past the last accumulator, 0 when filled in, negative on error. */
int gffx_job_buffer(gffx_index* idx, gffx_counts* counts, uint32_t k, void** dptr, uint64_t* n_elems, int* op);
/* dst[i] = op over r < rows of src[r * stride_elems + i] for i < n_elems; device pointers on the handle's GPU,
 * asynchronous on the engine's stream (gffx_index_cuda_stream) */
int gffx_fold(gffx_index* idx, void* dst, const void* src, uint32_t rows, uint64_t stride_elems, uint64_t n_elems, int op);
/* wait for the engine's stream */
int gffx_index_sync(gffx_index* idx);

/* ------------------------------------------------------------------------------------------
 * Allocation-free entry point.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  The join-pair count of one device-resident batch — the number gffx_query_features reports as n_pairs, from
 * the same partition and join kernels — with every intermediate in a fixed static device array inside the library
 * (10.6 MB; every GPU that loads the library holds it until exit).  The call allocates and frees nothing, on the device
 * or the host.
 *   d_chr / d_start / d_end : n sequence numbers / starts / ends, device memory of the handle's GPU
 *   mode 0 overlap, 1 contained (-c), 2 contains-region (-C); invert: -I
 *   d_out   : two uint64 in device memory (8-byte aligned): [0] pairs, [1] non-zero when a region's sequence number lies
 *             beyond the index's sequence list (the count is then that of the other regions)
 *   stream  : a cudaStream_t, or NULL for the handle's stream (gffx_index_cuda_stream); the call is asynchronous on it
 * Limits: n <= gffx_noalloc_max_regions() (1 048 576) and an index of at most gffx_noalloc_max_roots() (262 144) root
 * intervals; beyond either the call returns GFFX_E_ARG with a message that says so.  Every call on a GPU shares the one
 * array: calls that overlap on the GPU (two streams, two handles on one device) race and must be serialised by the caller. */
int gffx_count_pairs_noalloc(gffx_index* idx, const uint32_t* d_chr, const uint32_t* d_start, const uint32_t* d_end, uint64_t n,
                             int mode, int invert, uint64_t* d_out, void* stream);
uint64_t gffx_noalloc_max_regions(void);
uint64_t gffx_noalloc_max_roots(void);

/* ------------------------------------------------------------------------------------------
 * Command level: the reference's run() functions (src/commands/intersect.rs:541-655,
 * src/commands/depth.rs:548-635, src/commands/coverage.rs:493-582) with the GPU engine inside.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
    const char* input;       /* -i */
    const char* output;      /* -o, NULL = stdout */
    const char* bed;         /* -b  | exactly one  */
    const char* region;      /* -r  | of the two   */
    int mode;                /* gffx_mode */
    int invert;              /* -I */
    int entire_group;        /* -e */
    const char* types;       /* -T or NULL */
    int verbose;             /* -v */
    int device;
} gffx_intersect_args;
int gffx_intersect_run(const gffx_intersect_args* args);

typedef struct {
    const char* input;   /* -i */
    const char* source;  /* -s: .bed, or .bam / .sam (gffx_parse_reads); .cram: GFFX_E_ARG (not decoded) */
    const char* output;  /* -o, NULL = stdout */
    uint32_t bin_shift;  /* accepted, cannot change results (depth.rs:163-191) */
    int threads;         /* -t: host parse threads */
    int verbose;
    int device;
} gffx_counts_args;
int gffx_depth_run(const gffx_counts_args* args);
int gffx_coverage_run(const gffx_counts_args* args);

/* ------------------------------------------------------------------------------------------
 * Synthetic workload generators, counter-based splitmix64 so any shard can be
 * generated independently, on the host or straight into HBM.
 * ------------------------------------------------------------------------------------------ */
/* human-like GFF3 text written to `path`; returns number of lines through *n_lines */
int gffx_synth_gff(const char* path, uint64_t seed, uint32_t n_genes, int quirks, uint64_t* n_lines);
/* the same with n_scaffolds unplaced scaffolds (3-150 kb, 1-3 genes each) behind the 24 chromosomes: the shape of a
 * scaffold-level assembly (the reference's benchmark/hs.gff.sqs lists 583 sequences) */
int gffx_synth_gff_scaffolds(const char* path, uint64_t seed, uint32_t n_genes, int quirks, uint32_t n_scaffolds, uint64_t* n_lines);
/* regions for sequences of the index (placement proportional to sequence extent); host arrays */
int gffx_synth_regions_host(const gffx_index* idx, uint64_t seed, uint64_t first, uint64_t n, uint32_t length,
                            uint32_t* chr, uint32_t* start, uint32_t* end);
/* same, generated by a kernel into device arrays d_chr/d_start/d_end on the handle's GPU */
int gffx_synth_regions_device(gffx_index* idx, uint64_t seed, uint64_t first, uint64_t n, uint32_t length, uint32_t* d_chr,
                              uint32_t* d_start, uint32_t* d_end);

#ifdef __cplusplus
}
#endif
#endif
