// engine_impl.cuh — device state of one index (shared by engine.cu and counts.cu)
#pragma once
#include <cuda_runtime.h>
#include "engine.h"
#include "filter.cuh"
#include "partition.cuh"
#include "bedparse.cuh"
#include "gather.cuh"
#include "winjoin.cuh"
#include "winjoin_build.h"
#include "rootjoin.cuh"

namespace gffx {

#define GFFX_CUDA(call)                                                                                              \
    do {                                                                                                             \
        cudaError_t _e = (call);                                                                                     \
        if (_e != cudaSuccess)                                                                                       \
            throw Err(E_CUDA, std::string("CUDA error: ") + cudaGetErrorString(_e) + " at " + __FILE__ + ":" +       \
                                  std::to_string(__LINE__) + " (" #call ")");                                        \
    } while (0)

// grow-only device buffer
struct DBuf {
    void* p = nullptr;
    size_t cap = 0;
    size_t* total = nullptr;  // accounting
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
    void ensure(size_t bytes) {
        if (bytes <= cap) return;
        if (p) { cudaFree(p); if (total) *total -= cap; }
        p = nullptr;
        size_t want = bytes + bytes / 8 + 256;
        GFFX_CUDA(cudaMalloc(&p, want));
        cap = want;
        if (total) *total += cap;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

template <class T> inline void upload(DBuf& b, const std::vector<T>& v, cudaStream_t st) {
    b.ensure(std::max<size_t>(v.size(), 1) * sizeof(T));
    if (!v.empty()) GFFX_CUDA(cudaMemcpyAsync(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, st));
}

struct CountsState;  // counts.cu

struct IxChunk {
    uint64_t off = 0, n = 0;  // position of the batch in the job's bucketed arrays (retained jobs), regions in it
};
// an asynchronous job's read-back target: page-locked bytes + the job's own pair of events
constexpr size_t ASYNC_SLOT_BYTES = 8192;
constexpr size_t ASYNC_WARN_OFF = 64;        // [0,64) counters, [64, 64 + 4 * ASYNC_WARN_CAP) head of the warn list, [4096, 8192) -T allow bytes
constexpr uint32_t ASYNC_WARN_CAP = 1000;
constexpr size_t ASYNC_ALLOW_OFF = 4096, ASYNC_ALLOW_CAP = 4096;
struct AsyncSlot {
    uint8_t* h = nullptr;
    cudaEvent_t ev_start = nullptr, ev_end = nullptr;
    int state = 0;   // 0 free, 1 owned by a pending result, 2 abandoned (its result was freed before the copy landed)
};

struct IxJob {  // state of one intersect job between ix_begin and ix_finish
    bool active = false;
    bool async = false;
    int slot = -1;
    int mode = 2;
    bool invert = false, entire_group = false, join_only = false, match_only = true, run_filter = false, have_types = false;
    bool pairs_dev = false, want_pairs_host = false, no_host = false, nothing = false, sealed = false;
    bool host_in = false;        // a queued job whose regions are copied from host arrays behind the call
    bool stage_events = false;   // GFFX_F_STAGE_TIMES: per-stage CUDA events (stage_ms); otherwise only the call's whole span is timed
    uint64_t n_total = 0, retained = 0;
    unsigned long long pair_cap = 0;
    uint32_t launches = 0;
    std::vector<IxChunk> chunks;
};

class EngineImpl {
   public:
    std::unique_ptr<HostIndex> H;
    int dev = 0;
    cudaStream_t st = nullptr, st_copy = nullptr;  // st: every kernel; st_copy: host -> device chunks of the regions
    size_t dev_bytes = 0;
    int sm_count = 148;
    unsigned char* noalloc_scratch = nullptr;   // device address of engine.cu's static scratch array on this engine's GPU

    std::vector<std::vector<uint32_t>> seq_sorted_roots;  // per seq: (start,end,fid) triples sorted by start (accessor)

    // ---- window-addressed root join (winjoin.cuh): what intersect uses -------------------------
    DBuf d_wj_win0, d_wj_base, d_wj_winseq, d_wj_ent4, d_wj_list4, d_wj_fid, d_wj_first, d_wj_boff, d_wj_rend, d_wj_blk, d_wj_hit, d_wj_cnt, d_wj_off;
    WinBins WC{};
    uint32_t wj_n_ent = 0;
    int wj_force = 0;                           // GFFX_WJ_KERNEL: 0 = by job size, 1 = region-parallel rounds, 2 = root-parallel
    uint32_t wj_tail_bins = WR_TAIL_BINS;       // GFFX_WR_TAIL_BINS: regions longer than this many fine bins take the tail path

    // ---- blocks ------------------------------------------------------------------------------
    DBuf d_bstart, d_bend, d_blk_sel;
    uint32_t n_blocks = 0;

    // ---- lines + filter rank space -----------------------------------------------------------
    DBuf d_ln_flags, d_ln_blk, d_ln_type, d_ln_iseq, d_ln_S, d_ln_E, d_ln_iS, d_ln_iE, d_ln_off, d_ln_len;
    DBuf d_fX, d_fdir, d_fseqs;
    DBuf d_wr_xw0, d_wr_xlen, d_wr_xoff;   // rootjoin.cuh: per window its part of X (RankSlices)
    RankSlices XS{};
    bool wj_slices = true;                 // GFFX_WR_SLICE=0: the root-parallel kernel looks ranks up in global memory
    uint32_t f_slots = 0;
    RankSpace FR{};
    LineArrays LA{};
    // per-line match flags of a sealed (sharded) job live behind the per-root hit flags in d_wj_hit
    uint64_t match_off() const { return ((uint64_t)wj_n_ent + 255u) & ~255ull; }
    uint64_t flag_bytes() const { return match_off() + std::max<uint64_t>(LA.n, 1); }
    uint8_t* match_flags() const { return d_wj_hit.as<uint8_t>() + match_off(); }
    DBuf d_A, d_B, d_Ms, d_Me, d_PMX, d_SMN, d_NS, d_NE, d_keep, d_type_allow;

    // ---- window grid + partition scratch (partition.cuh) --------------------------------------
    DBuf d_seqmap, d_win0, d_poff, d_ptot, d_pcur, d_bse;
    std::vector<uint32_t> h_win0;
    WinGrid WG{};

    // ---- per-call scratch --------------------------------------------------------------------
    DBuf d_q[3], d_tile_count, d_tile_off, d_scan_scratch, d_pairs, d_ranges, d_small, d_warn;
    void* h_small = nullptr;  // pinned, 256 B
    cudaEvent_t ev[12];  // [0..7] stage boundaries of one intersect call, [8],[9] around the pair expansion (K1b), [10] end of the partition pass

    IxJob job;
    // ---- job lanes ---------------------------------------------------------------------------
    // Two sets of the per-job device state (lane_bufs(): accumulators, bucketed regions, scan arrays, outputs) and two streams, so
    // that two queued GFFX_F_ASYNC jobs run side by side: the partition pass of one fills the issue slots the join of the other
    // leaves idle (profiles/exp_two_handles.py: -19 %).  The members ARE the active lane's buffers; lane_activate swaps them with
    // the other lane's saved copies, so the kernels' call sites do not know about lanes.  Synchronous jobs, depth / coverage, the
    // BED parser and the gather run on lane 0 (`st` outside a job is lane 0's stream, the one gffx_index_cuda_stream reports).
    struct Lane {
        std::vector<DBuf> bufs;   // the lane's buffers while it is NOT the active one
        cudaStream_t st = nullptr;
        uint32_t epoch = 0, epoch8 = 0;
        bool ready = false;
        cudaEvent_t ev_done = nullptr;   // end of the lane's last asynchronous job: its staged regions may be overwritten after it
    };
    static constexpr int MAX_LANES = 4;
    Lane lanes[MAX_LANES];
    ScanChain chains[MAX_LANES];   // the lanes' single-pass scan records (scan.cuh); chain() = the active lane's
    DBuf d_chain_flags;
    ScanChain* chain() { return chains[lane_active].flags ? &chains[lane_active] : nullptr; }
    int lane_active = 0, lane_next = 0, lane_last = 0, n_lanes = 3;   // three lanes: 1.15 -> 1.07 ms per four-pass step over two; a fourth adds nothing (r03g)
    uint32_t epoch_start = 0;
    std::vector<DBuf*> lane_bufs();
    void lane_activate(int k);
    void alloc_job_buffers();
    DBuf& lane_buf(int lane, DBuf* member);   // `member` as lane `lane` sees it
    std::vector<AsyncSlot> slots = std::vector<AsyncSlot>(32);
    uint32_t epoch = 0, epoch8 = 0;         // filter.cuh: the rank-space arrays are tagged per job, not cleared
    unsigned long long pairs_hwm = 0;       // most pairs any job of this engine produced: sizes the pair buffer of the next one
    DBuf d_blk_ent0, d_blk_ent, d_orphans;  // per block the tree intervals that name it; intervals whose fid has no .gof record
    uint32_t n_orphans = 0;
    int slot_acquire();
    void ix_wait(IntersectResult* out);
    void ix_abandon(IntersectResult* out);
    void read_counters(IntersectResult* out, const uint8_t* h, const uint32_t* warn_head);
    std::vector<cudaEvent_t> ev_chunk, ev_copy;  // 4 timing events per batch of a job; one copy-done event per chunk

    // ---- counts (depth / coverage) -----------------------------------------------------------
    CountsState* counts = nullptr;  // owned; freed by counts_release()

    EngineImpl(std::unique_ptr<HostIndex> host, int device);
    ~EngineImpl();
    void build_windows();
    void build_join();
    void build_blocks();
    void build_lines();
    void intersect(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, int mode, bool invert,
                   bool entire_group, const char* types_csv, uint32_t flags, bool join_only, IntersectResult* out);
    void ix_begin(int mode, bool invert, bool entire_group, const char* types_csv, uint32_t flags, bool join_only, uint64_t n_retained,
                  uint32_t n_chunks);
    void ix_add(const uint32_t* dc, const uint32_t* ds, const uint32_t* de, uint64_t n);
    void ix_finish(IntersectResult* out);
    void ix_seal();
    uint32_t ix_scans(FilterScans* P);
    uint32_t expand_pairs(unsigned long long cap);
    // K8 (gather.cuh): the text of `n_ranges` byte ranges gathered on the device; ranges == nullptr: those of the last intersect
    void gather_text(const uint64_t* ranges, uint64_t n_ranges, uint32_t flags, const uint8_t** text, uint64_t* n_bytes);
    DBuf d_text, d_text_off, d_text_out, d_text_ranges;
    bool text_resident = false;
    uint64_t last_n_ranges = 0;      // ranges of the last finished intersect still in d_ranges
    HostBlob<uint8_t> h_text_out;    // page-locked copy of the last gathered text
    // K9 (bedparse.cuh / bedparse.cu): BED text -> device-resident regions
    DBuf d_bed_text, d_bed_tiles, d_bed_tfirst, d_bed_lstart, d_bed_status, d_bed_line[3], d_bed_flag, d_bed_pos, d_bed_out[3];
    DBuf d_bed_nhash, d_bed_nid, d_bed_noff, d_bed_nlen, d_bed_nblob;
    BedNames bed_names{};
    bool bed_names_built = false;
    cudaEvent_t ev_bed[4] = {nullptr, nullptr, nullptr, nullptr};
    double bed_ms = 0;            // device time of the parse kernels
    uint32_t bed_launches = 0;
    uint64_t bed_host_lines = 0;  // lines the kernel left to the host's exact per-line functions
    uint64_t bed_bytes = 0;
    double gather_ms = 0, gather_scan_ms = 0;   // gather_text_kernel alone; its offsets scan
    cudaEvent_t ev_gather[4] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t gather_launches = 0;  // K1b: offsets scan + wj_expand_kernel into d_pairs
    // stage the regions on the device; returns device pointers (either the caller's or the staging buffers)
    void stage_queries(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags,
                       const uint32_t** dc, const uint32_t** ds, const uint32_t** de);
    // K0: bucket n (< 2^31) device-resident regions by window; returns the bucketed arrays (engine-owned
    // scratch, valid until the next call) and the number of kernels launched
    uint32_t partition_into(const uint32_t* dc, const uint32_t* ds, const uint32_t* de, uint64_t n, uint2* ose, uint32_t* poff);
};

// counts.cu
void counts_release(EngineImpl* E);
void counts_begin_impl(EngineImpl* E, int kind);
void counts_add_impl(EngineImpl* E, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags);
void counts_finish_impl(EngineImpl* E, CountRows* rows);
double counts_ms_impl(const EngineImpl* E);
bool counts_buffer_impl(EngineImpl* E, uint32_t k, void** dptr, uint64_t* n_elems, int* op);
uint32_t counts_launches_impl(const EngineImpl* E);

}  // namespace gffx
