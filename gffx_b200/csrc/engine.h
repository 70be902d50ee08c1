// engine.h — the device-resident index and the GPU pipelines behind the C ABI (include/gffx_b200.h).
// Host-only declarations: CUDA types stay inside engine.cu.
#pragma once
#include <memory>
#include "host_index.h"

namespace gffx {

// Page-locked host memory from a process-wide grow-only pool (engine.cu): result arrays land in it by DMA at PCIe
// speed (a pageable std::vector costs a zero-fill, page faults and a staged copy: ~5x slower for a 40 MB range list).
template <class T> class HostBlob {
   public:
    HostBlob() = default;
    HostBlob(const HostBlob&) = delete;
    HostBlob& operator=(const HostBlob&) = delete;
    HostBlob(HostBlob&& o) noexcept { *this = std::move(o); }
    HostBlob& operator=(HostBlob&& o) noexcept {
        if (this != &o) { release(); p_ = o.p_; cap_ = o.cap_; n_ = o.n_; o.p_ = nullptr; o.cap_ = 0; o.n_ = 0; }
        return *this;
    }
    ~HostBlob() { release(); }
    void resize(size_t n);   // contents are NOT preserved
    void release();
    T* data() { return (T*)p_; }
    const T* data() const { return (const T*)p_; }
    size_t size() const { return n_; }
    bool empty() const { return n_ == 0; }
    T& operator[](size_t i) { return ((T*)p_)[i]; }
    const T& operator[](size_t i) const { return ((const T*)p_)[i]; }
   private:
    void* p_ = nullptr;
    size_t cap_ = 0, n_ = 0;
};
void* pinned_acquire(size_t bytes, size_t* cap);
void pinned_release(void* p, size_t cap);
template <class T> void HostBlob<T>::resize(size_t n) {
    if (n * sizeof(T) > cap_) { release(); p_ = pinned_acquire(n * sizeof(T), &cap_); }
    n_ = n;
}
template <class T> void HostBlob<T>::release() {
    if (p_) pinned_release(p_, cap_);
    p_ = nullptr; cap_ = 0; n_ = 0;
}

struct PendingInfo {   // what turns a job's counter block into result fields (kept with the result while it is pending)
    bool pairs_dev = false, join_only = false, match_only = true, nothing = false, no_host = false, want_pairs_host = false;
    unsigned long long pair_cap = 0;
    const uint32_t* d_pairs = nullptr;   // the pair buffer of the job's lane
    int lane = 0;
    bool host_in = false;                // asynchronous job whose regions were host arrays: they are read until the job's copies have run
    uint64_t ranges_cap = 0;             // asynchronous job with a host copy: ranges in flight to IntersectResult::ranges (an upper bound)
};

struct IntersectResult {
    int pending_slot = -1;   // >= 0: an asynchronous job whose counters have not been read yet (Engine::wait)
    PendingInfo pend;
    uint64_t n_queries = 0;
    uint64_t n_pairs = 0;
    HostBlob<uint32_t> pairs;         // host copy (WANT_PAIRS)
    const uint32_t* d_pairs = nullptr;  // device copy (PAIRS_ON_DEVICE / WANT_PAIRS), owned by the engine until the next call
    uint64_t n_roots = 0;
    uint64_t n_ranges = 0;
    HostBlob<uint64_t> ranges;        // host copy unless NO_HOST_COPY
    uint64_t bytes_out = 0;
    std::vector<uint32_t> warn_fids;
    double kernel_ms = 0;
    double stage_ms[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t launches = 0;
};



struct CountRows {  // parallel arrays in page-locked memory
    HostBlob<uint32_t> id_idx, chrom_idx, start, end;
    HostBlob<uint64_t> value;
};

// how the shards' copies of an accumulator combine (gffx_reduce_op of the C ABI)
enum ReduceOp { RED_MAX_U8 = 0, RED_MAX_U64 = 1, RED_MIN_U64 = 2, RED_SUM_U32 = 3, RED_SUM_I32 = 4, RED_OR_U64 = 5 };

class EngineImpl;

class Engine {
   public:
    // takes ownership of a fully built HostIndex and uploads it to `device`
    Engine(std::unique_ptr<HostIndex> host, int device);
    ~Engine();
    const HostIndex& host() const;
    int device() const;
    uint64_t device_bytes() const;
    void* cuda_stream() const;
    // (start,end,root_fid) triples of one sequence's tree intervals, sorted by start
    const std::vector<uint32_t>& sorted_roots(uint32_t seq) const;

    // mode: 0 contained, 1 contains-region, 2 overlap.  join_only: stop after the root-level join.
    void intersect(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, int mode, bool invert,
                   bool entire_group, const char* types_csv, uint32_t flags, bool join_only, IntersectResult* out);

    // The same pipeline in three steps, for jobs whose regions arrive in batches or are spread over several GPUs: begin,
    // any number of adds (accumulate), [combine the accumulators of the shards: shard_buffer + fold], finish.
    void shard_begin(int mode, bool invert, bool entire_group, const char* types_csv, uint32_t flags, uint64_t n_expected,
                     uint32_t batches_expected);
    void shard_add(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags, uint64_t* n_pairs_total);
    void shard_finish(IntersectResult* out);
    // GFFX_F_ASYNC jobs: block until the result's counters have arrived and fill its fields (no-op otherwise); abandon() is
    // for a result that is dropped without having been waited for
    void wait(IntersectResult* out);
    void abandon(IntersectResult* out);
    // accumulator k of the running intersect job (which = 0) or depth / coverage job (which = 1): device pointer, element
    // count and how shards combine it (ReduceOp).  false when k is past the last one.
    bool shard_buffer(int which, uint32_t k, void** dptr, uint64_t* n_elems, int* op);
    // dst[i] = op over r < rows of src[r * stride + i], i < n_elems (device pointers on this engine's GPU, engine stream)
    void fold(void* dst, const void* src, uint32_t rows, uint64_t stride_elems, uint64_t n_elems, int op);
    void sync();
    // join-pair count of a device-resident batch with NO allocation in the call: every intermediate lives in a fixed static
    // device array (engine.cu: gffx_noalloc_scratch).  d_out[0] = pairs, d_out[1] != 0: a sequence number out of range.
    // Asynchronous on `stream` (NULL: the engine's stream); overlapping calls on one GPU share the array and race.
    void count_pairs_noalloc(const uint32_t* d_chr, const uint32_t* d_start, const uint32_t* d_end, uint64_t n, int mode, bool invert,
                             uint64_t* d_out, void* stream);
    static uint64_t noalloc_max_regions();
    static uint64_t noalloc_max_roots();

    // depth (kind 0) / coverage (kind 1)
    void counts_begin(int kind);
    void counts_add(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags);
    void counts_finish(CountRows* rows);
    double counts_kernel_ms() const;
    uint32_t counts_launches() const;

    // record emission on the device (gather.cuh): the text of the byte ranges — `ranges` (host, n_ranges x (start,end)) or,
    // when NULL, the ranges of the last finished intersect still on the device — gathered from the HBM-resident GFF.
    // flags & 4 (NO_HOST_COPY): *text is a device pointer; else a page-locked host copy.  Both live until the next call.
    void gather_text(const uint64_t* ranges, uint64_t n_ranges, uint32_t flags, const uint8_t** text, uint64_t* n_bytes);
    double gather_kernel_ms() const;   // gather_text_kernel alone (CUDA events)
    double gather_scan_ms() const;     // the offsets scan before it
    uint32_t gather_launches() const;

    // BED text parsed on the device (bedparse.cuh): dialect 0 = intersect (parse_bed_file), 1 = depth / coverage.  The arrays are
    // device pointers owned by the engine until its next parse; pass them on with GFFX_F_DEVICE_INPUT.
    void parse_bed_device(const std::string& path, int dialect, const uint32_t** d_chr, const uint32_t** d_start, const uint32_t** d_end,
                          uint64_t* n);
    // 0 kernel ms, 1 launches, 2 lines decided by the host's per-line functions, 3 bytes of text
    double bed_parse_stat(int what) const;

    void synth_regions_device(uint64_t seed, uint64_t first, uint64_t n, uint32_t length, uint32_t* d_chr, uint32_t* d_start,
                              uint32_t* d_end);

   private:
    std::unique_ptr<EngineImpl> impl_;
};

int cuda_device_count();

// index_builder.cpp
void build_index_files(const std::string& gff_path, const std::string& attr_key, const std::string& skip_types);

// synthetic workloads (synth.cpp)
uint64_t synth_gff_file(const std::string& path, uint64_t seed, uint32_t n_genes, bool quirks, uint32_t n_scaffolds = 0);
void synth_regions_host(const HostIndex& H, uint64_t seed, uint64_t first, uint64_t n, uint32_t length, uint32_t* chr,
                        uint32_t* start, uint32_t* end);
// per-sequence extent table shared by the host and device region generators
void synth_seq_table(const HostIndex& H, std::vector<uint32_t>* seq_extent, std::vector<uint64_t>* cum);

}  // namespace gffx
