// engine.cu — index upload and the intersect pipeline (join -> block select -> line filter / merge).
#include <algorithm>
#include <numeric>
#include <mutex>
#include <atomic>
#include <functional>
#include <thread>
#include "engine_impl.cuh"
#include "synth_common.h"

namespace gffx {

constexpr uint32_t WARN_CAP = 1u << 16;

static unsigned char* noalloc_scratch_address();   // device address of the allocation-free entry point's static scratch (below)

// The join kernel's dynamic shared memory on top of its static 26 KB exceeds the 48 KB default.  Function attributes belong to
// the device (context), so every engine sets them for its own device when it is built (one process may hold an engine per GPU).
template <int MODE, bool INV> static void wj_set_attrs() {
    GFFX_CUDA(cudaFuncSetAttribute(wj_count_kernel<MODE, INV, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WJ_DYN_SMEM));
    GFFX_CUDA(cudaFuncSetAttribute(wj_count_kernel<MODE, INV, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WJ_DYN_SMEM));
    GFFX_CUDA(cudaFuncSetAttribute(wj_count_kernel<MODE, INV, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WJ_DYN_SMEM));
    GFFX_CUDA(cudaFuncSetAttribute(wj_count_kernel<MODE, INV, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WJ_DYN_SMEM));
    GFFX_CUDA(cudaFuncSetAttribute(wj_root_kernel<MODE, INV, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WJ_DYN_SMEM));
    GFFX_CUDA(cudaFuncSetAttribute(wj_root_kernel<MODE, INV, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WJ_DYN_SMEM));
    GFFX_CUDA(cudaFuncSetAttribute(wj_root_kernel<MODE, INV, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(WJ_DYN_SMEM + WR_SLICE_SMEM)));
}
static void set_kernel_attributes() {
    wj_set_attrs<0, false>(); wj_set_attrs<0, true>(); wj_set_attrs<1, false>(); wj_set_attrs<1, true>(); wj_set_attrs<2, false>();
    GFFX_CUDA(cudaFuncSetAttribute(part_scatter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PART_SCATTER_SMEM));
    GFFX_CUDA(cudaFuncSetAttribute(part_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PART_COUNT_SMEM));
}

// small kernels -------------------------------------------------------------------------------
struct TileCountLoad {
    const uint32_t* p;
    __device__ uint64_t operator()(uint64_t i) const { return (uint64_t)p[i]; }
};
struct MarkLoad {  // next-set scan input: the cell's own index when it carries THIS job's mark (filter.cuh: epochs)
    const uint8_t* m;
    uint32_t epoch8;
    __device__ uint32_t operator()(uint64_t i) const { return m[i] == epoch8 ? (uint32_t)i : 0xFFFFFFFFu; }
};

// the per-job accumulators that do need zeros — the 256-byte counter block, the per-root hit flags and hit counters — in ONE launch
// (they were four memset nodes; the rank-space arrays are never cleared: filter.cuh, epochs)
__global__ void clear_job_kernel(uint4* __restrict__ small, uint4* __restrict__ hit, uint32_t n_hit16, uint4* __restrict__ cnt, uint32_t n_cnt16) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint4 z = make_uint4(0, 0, 0, 0);
    if (i < 16) small[i] = z;
    if (i < n_hit16) hit[i] = z;
    for (uint32_t k = i; k < n_cnt16; k += gridDim.x * blockDim.x) cnt[k] = z;
}

// the finished range list -> the result's page-locked host buffer (device-accessible: unified addressing), exactly as many as the
// job's counter says: a queued job's host copy without the host knowing the count (a cudaMemcpyAsync would need it)
__global__ void ranges_to_host_kernel(const ulonglong2* __restrict__ src, ulonglong2* __restrict__ dst_host, const unsigned long long* __restrict__ n_ptr,
                                      unsigned long long cap) {
    const unsigned long long n = min(*n_ptr, cap);
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x)
        dst_host[i] = src[i];
}

__global__ void fill_u64_kernel(unsigned long long* p, uint64_t n, unsigned long long v) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ------------------------------------------------------------------------------------------------
EngineImpl::EngineImpl(std::unique_ptr<HostIndex> host, int device) : H(std::move(host)), dev(device) {
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        throw Err(E_CUDA, std::string("no usable CUDA device (") + cudaGetErrorString(e) + "); this engine has no CPU fallback");
    if (device < 0 || device >= ndev) throw Err(E_ARG, "device ordinal out of range");
    GFFX_CUDA(cudaSetDevice(dev));
    cudaDeviceProp prop;
    GFFX_CUDA(cudaGetDeviceProperties(&prop, dev));
    if (prop.major < 10) throw Err(E_CUDA, std::string("device ") + prop.name + " is not sm_100-class; the kernels are built for sm_100a only");
    sm_count = prop.multiProcessorCount;
    GFFX_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    GFFX_CUDA(cudaStreamCreateWithFlags(&st_copy, cudaStreamNonBlocking));
    lanes[0].st = st;
    lanes[0].ready = true;
    if (const char* e = getenv("GFFX_LANES")) n_lanes = std::min((int)MAX_LANES, std::max(1, atoi(e)));   // A/B hook: 1 = every job on lane 0
    if (const char* e = getenv("GFFX_WJ_KERNEL")) wj_force = !strcmp(e, "rounds") ? 1 : !strcmp(e, "roots") ? 2 : 0;
    if (const char* e = getenv("GFFX_WR_SLICE")) wj_slices = atoi(e) != 0;
    if (const char* e = getenv("GFFX_WR_TAIL_BINS")) wj_tail_bins = (uint32_t)std::min(1 << 20, std::max(0, atoi(e)));
    for (auto& x : ev) GFFX_CUDA(cudaEventCreate(&x));
    GFFX_CUDA(cudaMallocHost(&h_small, 256));
    set_kernel_attributes();
    if (!getenv("GFFX_SCAN_TWO_PASS")) {   // A/B hook: the reduce-then-scan form everywhere
        d_chain_flags.total = &dev_bytes;
        d_chain_flags.ensure((size_t)MAX_LANES * SCAN_CHAIN_TILES * 4);
        GFFX_CUDA(cudaMemsetAsync(d_chain_flags.p, 0, (size_t)MAX_LANES * SCAN_CHAIN_TILES * 4, st));
        for (int k = 0; k < MAX_LANES; k++) { chains[k].flags = d_chain_flags.as<uint32_t>() + (size_t)k * SCAN_CHAIN_TILES; chains[k].cap_tiles = SCAN_CHAIN_TILES; }
    }
    noalloc_scratch = noalloc_scratch_address();   // once per engine = per device, this device current
    for (DBuf* b : {&d_bstart, &d_bend, &d_blk_sel,
                    &d_ln_flags, &d_ln_blk, &d_ln_type, &d_ln_iseq, &d_ln_S, &d_ln_E, &d_ln_iS, &d_ln_iE, &d_ln_off, &d_ln_len,
                    &d_fX, &d_fdir, &d_fseqs, &d_wr_xw0, &d_wr_xlen, &d_wr_xoff, &d_A, &d_B, &d_Ms, &d_Me, &d_PMX, &d_SMN, &d_NS, &d_NE, &d_keep, &d_type_allow,
                    &d_wj_win0, &d_wj_base, &d_seqmap, &d_blk_ent0, &d_blk_ent, &d_orphans, &d_wj_winseq, &d_wj_ent4, &d_wj_list4, &d_wj_fid, &d_wj_first, &d_wj_boff, &d_wj_rend, &d_wj_blk, &d_wj_hit, &d_wj_cnt, &d_wj_off, &d_text, &d_text_off, &d_text_out, &d_text_ranges, &d_bed_text, &d_bed_tiles, &d_bed_tfirst, &d_bed_lstart, &d_bed_status,
                    &d_bed_line[0], &d_bed_line[1], &d_bed_line[2], &d_bed_flag, &d_bed_pos, &d_bed_out[0], &d_bed_out[1], &d_bed_out[2],
                    &d_bed_nhash, &d_bed_nid, &d_bed_noff, &d_bed_nlen, &d_bed_nblob,
                    &d_win0, &d_poff, &d_ptot, &d_pcur, &d_bse,
                    &d_q[0], &d_q[1], &d_q[2], &d_tile_count, &d_tile_off, &d_scan_scratch, &d_pairs, &d_ranges, &d_small, &d_warn})
        b->total = &dev_bytes;
    {
        PhaseTimer t("engine: windows + join index");
        build_windows();
        build_join();
        build_blocks();
    }
    {
        PhaseTimer t("engine: lines + rank space");
        build_lines();
        // roots hit | lines matched: one allocation, so that a sharded job has ONE flag array to exchange (match_flags())
        d_wj_hit.ensure(flag_bytes());
        GFFX_CUDA(cudaMemsetAsync(d_wj_hit.p, 0, flag_bytes(), st));
    }
    d_small.ensure(256);
    d_warn.ensure(WARN_CAP * 4);
    if (H->dev_text && H->dev_text_device == dev) {   // the device line scanner left the GFF text in HBM: record emission adopts it
        d_text.p = H->dev_text;
        d_text.cap = H->dev_text_cap;
        dev_bytes += d_text.cap;
        text_resident = true;
        H->dev_text = nullptr;
    }
    GFFX_CUDA(cudaStreamSynchronize(st));
}

EngineImpl::~EngineImpl() {
    cudaSetDevice(dev);
#ifdef WR_STATS
    {
        unsigned long long h[8];
        cudaDeviceSynchronize();
        cudaMemcpyFromSymbol(h, wr_stats, sizeof h);
        fprintf(stderr, "WR_STATS roots_tested %llu iters %llu nn %llu segs %llu n_roots %llu n1 %llu n_norm %llu maxlen_sum %llu\n", h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
    }
#endif
    for (Lane& l : lanes)
        if (l.st) cudaStreamSynchronize(l.st);
    for (int k = 0; k < MAX_LANES; k++)   // the active lane's buffers are the members, released below
        if (k != lane_active)
            for (DBuf& b : lanes[k].bufs) b.release();
    for (DBuf* b : {&d_bstart, &d_bend, &d_blk_sel,
                    &d_ln_flags, &d_ln_blk, &d_ln_type, &d_ln_iseq, &d_ln_S, &d_ln_E, &d_ln_iS, &d_ln_iE, &d_ln_off, &d_ln_len,
                    &d_fX, &d_fdir, &d_fseqs, &d_wr_xw0, &d_wr_xlen, &d_wr_xoff, &d_A, &d_B, &d_Ms, &d_Me, &d_PMX, &d_SMN, &d_NS, &d_NE, &d_keep, &d_type_allow,
                    &d_wj_win0, &d_wj_base, &d_seqmap, &d_blk_ent0, &d_blk_ent, &d_orphans, &d_wj_winseq, &d_wj_ent4, &d_wj_list4, &d_wj_fid, &d_wj_first, &d_wj_boff, &d_wj_rend, &d_wj_blk, &d_wj_hit, &d_wj_cnt, &d_wj_off, &d_text, &d_text_off, &d_text_out, &d_text_ranges, &d_bed_text, &d_bed_tiles, &d_bed_tfirst, &d_bed_lstart, &d_bed_status,
                    &d_bed_line[0], &d_bed_line[1], &d_bed_line[2], &d_bed_flag, &d_bed_pos, &d_bed_out[0], &d_bed_out[1], &d_bed_out[2],
                    &d_bed_nhash, &d_bed_nid, &d_bed_noff, &d_bed_nlen, &d_bed_nblob,
                    &d_win0, &d_poff, &d_ptot, &d_pcur, &d_bse,
                    &d_q[0], &d_q[1], &d_q[2], &d_tile_count, &d_tile_off, &d_scan_scratch, &d_pairs, &d_ranges, &d_small, &d_warn})
        b->release();
    d_chain_flags.release();
    counts_release(this);
    if (h_small) cudaFreeHost(h_small);
    for (auto& S : slots) {
        if (S.h) cudaFreeHost(S.h);
        if (S.ev_start) cudaEventDestroy(S.ev_start);
        if (S.ev_end) cudaEventDestroy(S.ev_end);
    }
    for (auto& x : ev) cudaEventDestroy(x);
    for (auto& x : ev_chunk) cudaEventDestroy(x);
    for (auto& x : ev_copy) cudaEventDestroy(x);
    for (auto& x : ev_gather) if (x) cudaEventDestroy(x);
    for (auto& x : ev_bed) if (x) cudaEventDestroy(x);
    if (st_copy) cudaStreamDestroy(st_copy);
    for (Lane& l : lanes) {
        if (l.ev_done) cudaEventDestroy(l.ev_done);
        if (l.st) cudaStreamDestroy(l.st);
    }
}

// ---- job lanes (engine_impl.cuh) ------------------------------------------------------------------------------------
std::vector<DBuf*> EngineImpl::lane_bufs() {
    return {&d_wj_hit, &d_wj_cnt, &d_wj_off, &d_blk_sel, &d_A, &d_B, &d_Ms, &d_Me, &d_PMX, &d_SMN, &d_NS, &d_NE, &d_keep, &d_type_allow,
            &d_poff, &d_ptot, &d_pcur, &d_bse, &d_tile_count, &d_tile_off, &d_scan_scratch, &d_pairs, &d_ranges, &d_small, &d_warn,
            &d_q[0], &d_q[1], &d_q[2]};   // a queued job's staged regions stay put while the next job's cross PCIe
}

DBuf& EngineImpl::lane_buf(int lane, DBuf* member) {
    if (lane == lane_active) return *member;
    const std::vector<DBuf*> m = lane_bufs();
    for (size_t i = 0; i < m.size(); i++)
        if (m[i] == member && i < lanes[lane].bufs.size()) return lanes[lane].bufs[i];
    throw Err(E_STATE, "lane buffer (internal)");
}

// what a lane needs before its first job: the buffers the index sizes once (the others grow on demand)
void EngineImpl::alloc_job_buffers() {
    d_poff.ensure(((size_t)WG.n_win + 2) * 4);
    d_ptot.ensure((size_t)PART_NB_MAX * 4);
    d_pcur.ensure((size_t)PART_NB_MAX * 4);
    GFFX_CUDA(cudaMemsetAsync(d_ptot.p, 0, (size_t)PART_NB_MAX * 4, st));
    d_wj_cnt.ensure(std::max<uint32_t>(wj_n_ent, 1) * 8ull);
    d_wj_off.ensure(((uint64_t)wj_n_ent + 1) * 8);
    d_blk_sel.ensure(std::max<uint32_t>(n_blocks, 1));
    d_ranges.ensure(std::max<uint32_t>(n_blocks, 1) * 16ull);
    const size_t ns = std::max<uint32_t>(f_slots, 1);
    d_A.ensure(ns * 8); d_B.ensure(ns * 8); d_PMX.ensure(ns * 8); d_SMN.ensure((ns + 1) * 8);
    d_Ms.ensure(ns * 2); d_Me.ensure(ns * 2); d_NS.ensure(ns * 8); d_NE.ensure(ns * 8);
    GFFX_CUDA(cudaMemsetAsync(d_A.p, 0, ns * 8, st));
    GFFX_CUDA(cudaMemsetAsync(d_B.p, 0xFF, ns * 8, st));
    GFFX_CUDA(cudaMemsetAsync(d_Ms.p, 0, ns * 2, st));
    GFFX_CUDA(cudaMemsetAsync(d_Me.p, 0, ns * 2, st));
    epoch = epoch_start; epoch8 = epoch_start % 256u;
    d_keep.ensure(std::max<size_t>(LA.n, 1));
    d_type_allow.ensure(std::max<size_t>(H->types.size(), 1));
    d_wj_hit.ensure(flag_bytes());
    GFFX_CUDA(cudaMemsetAsync(d_wj_hit.p, 0, flag_bytes(), st));
    d_small.ensure(256);
    d_warn.ensure(WARN_CAP * 4);
}

void EngineImpl::lane_activate(int k) {
    if (k == lane_active) return;
    const std::vector<DBuf*> m = lane_bufs();
    Lane& cur = lanes[lane_active];
    Lane& nxt = lanes[k];
    cur.bufs.resize(m.size());
    if (nxt.bufs.size() != m.size()) {
        nxt.bufs.assign(m.size(), DBuf());
        for (DBuf& b : nxt.bufs) b.total = &dev_bytes;
    }
    for (size_t i = 0; i < m.size(); i++) { cur.bufs[i] = *m[i]; *m[i] = nxt.bufs[i]; }
    cur.epoch = epoch; cur.epoch8 = epoch8;
    epoch = nxt.epoch; epoch8 = nxt.epoch8;
    if (!nxt.st) GFFX_CUDA(cudaStreamCreateWithFlags(&nxt.st, cudaStreamNonBlocking));
    st = nxt.st;
    lane_active = k;
    if (!nxt.ready) {
        alloc_job_buffers();
        nxt.ready = true;
    }
}

// Window grid of the partition pass (partition.cuh): windows of 2^shift bases over each VIRTUAL sequence's extent (VSeqMap,
// host_index.h: the largest root end of a sequence that has a virtual sequence to itself, the laid-out length of a shared
// one), the finest grid with at most PART_MAX_WINDOWS windows.  Windows are a locality device only: a region that starts
// beyond the extent falls into the virtual sequence's last window.
void EngineImpl::build_windows() {
    const VSeqMap& V = H->vmap;
    const size_t n_seqs = H->seq_names.size(), n_v = V.n_vseqs();
    uint64_t max_win = PART_MAX_WINDOWS;
    if (const char* e = getenv("GFFX_MAX_WINDOWS")) max_win = std::min<uint64_t>(PART_MAX_WINDOWS, std::max<uint64_t>(n_v, strtoull(e, nullptr, 10)));
    uint32_t shift = 12;
    for (;; shift++) {
        uint64_t tot = 0;
        for (size_t v = 0; v < n_v; v++) tot += ((uint64_t)V.v_extent[v] >> shift) + 1;
        if (tot <= max_win || shift == 31) break;
    }
    h_win0.assign(n_v + 1, 0);
    for (size_t v = 0; v < n_v; v++) h_win0[v + 1] = h_win0[v] + (uint32_t)(((uint64_t)V.v_extent[v] >> shift) + 1);
    if (h_win0[n_v] > PART_MAX_WINDOWS) throw Err(E_ARG, "window grid overflow (internal): more virtual sequences than windows");
    std::vector<uint4> seqmap(std::max<size_t>(n_seqs, 1), make_uint4(0, 0xFFFFFFFFu, 0, 0));
    for (size_t c = 0; c < n_seqs; c++) {
        const VSeqMap::Seq& m = V.seq[c];
        const uint32_t nw = h_win0[m.vseq + 1] - h_win0[m.vseq];
        seqmap[c] = make_uint4(m.base, m.clamp, h_win0[m.vseq] | (nw << 16), m.vseq);
    }
    upload(d_win0, h_win0, st);
    upload(d_seqmap, seqmap, st);
    WG.win0 = d_win0.as<uint32_t>();
    WG.seqmap = d_seqmap.as<uint4>();
    WG.n_seqs = (uint32_t)n_seqs;
    WG.n_vseqs = (uint32_t)n_v;
    WG.shift = shift;
    WG.n_win = h_win0[n_v];
    d_poff.ensure(((size_t)WG.n_win + 2) * 4);
    d_ptot.ensure((size_t)PART_NB_MAX * 4);   // window totals: zero between partition passes (part_offsets_kernel clears them)
    d_pcur.ensure((size_t)PART_NB_MAX * 4);
    GFFX_CUDA(cudaMemsetAsync(d_ptot.p, 0, (size_t)PART_NB_MAX * 4, st));
    GFFX_CUDA(cudaStreamSynchronize(st));
}

// K0 with caller-chosen outputs: `ose` takes the n bucketed regions, `poff` the n_win + 2
// window offsets.  The window totals and cursors are scratch that the NEXT call on the stream reuses (stream order keeps that safe).
uint32_t EngineImpl::partition_into(const uint32_t* dc, const uint32_t* ds, const uint32_t* de, uint64_t n, uint2* ose, uint32_t* poff) {
    if (n == 0) return 0;
    if (n >= (1ull << 31)) throw Err(E_ARG, "more than 2^31-1 regions in one call: split the batch");
    const uint32_t nb = WG.n_win + 1;
    const unsigned tiles = (unsigned)((n + PART_TILE - 1) / PART_TILE);
    part_count_kernel<<<std::min(tiles, (unsigned)(sm_count * PART_COUNT_CTAS)), PART_THREADS, PART_COUNT_SMEM, st>>>(WG, dc, ds, (uint32_t)n, d_ptot.as<uint32_t>());
    part_offsets_kernel<<<1, PART_NB_MAX, 0, st>>>(d_ptot.as<uint32_t>(), nb, (uint32_t)n, poff, d_pcur.as<uint32_t>());
    const unsigned stiles = (unsigned)((n + PART_STILE - 1) / PART_STILE);
    part_scatter_kernel<<<std::min(stiles, (unsigned)(sm_count * PART_SCATTER_CTAS)), PART_THREADS, PART_SCATTER_SMEM, st>>>(WG, dc, ds, de, (uint32_t)n, d_pcur.as<uint32_t>(), ose);
    return 3;
}

// Join index over the root intervals (winjoin_build.h: start-sorted records + per-window bin lists).
void EngineImpl::build_join() {
    const size_t n_seqs = H->seq_names.size();
    const VSeqMap& V = H->vmap;
    const uint32_t n_v = (uint32_t)V.n_vseqs();
    const auto& R = H->roots;
    std::vector<JoinItem> items(R.size());   // in the engine's (virtual) coordinates
    for (size_t i = 0; i < R.size(); i++) {
        const VSeqMap::Seq& m = V.seq[R[i].seq];
        items[i] = JoinItem{m.vseq, m.base + R[i].start, m.base + R[i].end, R[i].fid};
    }
    // accessor copy: per sequence (start,end,fid) sorted by (start,end)
    seq_sorted_roots.assign(n_seqs, {});
    {
        std::vector<std::vector<uint32_t>> by_seq(n_seqs);
        for (size_t i = 0; i < R.size(); i++) by_seq[R[i].seq].push_back((uint32_t)i);
        for (size_t s = 0; s < n_seqs; s++) {
            auto& ids = by_seq[s];
            std::stable_sort(ids.begin(), ids.end(), [&](uint32_t a, uint32_t b) {
                if (R[a].start != R[b].start) return R[a].start < R[b].start;
                return R[a].end < R[b].end;
            });
            auto& flat = seq_sorted_roots[s];
            flat.reserve(ids.size() * 3);
            for (uint32_t i : ids) { flat.push_back(R[i].start); flat.push_back(R[i].end); flat.push_back(R[i].fid); }
        }
    }
    // the window-addressed form of the same intervals (winjoin.cuh), used by intersect: on the PARTITION's grid, so that
    // the bucketed regions of one window find their whole index slice in one contiguous piece (copied to shared memory)
    const std::vector<uint32_t>& jwin0 = h_win0;
    const uint32_t jshift = WG.shift;
    WinBinsHost bh;
    uint32_t max_sub = 7;  // GFFX_WJ_SUB=0..9; 2^7 bins per window (~65 kb at human scale): 6, 7 and 8 measure the same on the bench
    if (const char* e = getenv("GFFX_WJ_SUB")) max_sub = (uint32_t)std::min(9, std::max(0, atoi(e)));
    build_win_bins(n_v, items, jwin0, jshift, max_sub, 1ull << 27, &bh);
    std::vector<uint4> ent4(bh.n_entries()), list4(std::max<size_t>(bh.list.size(), 1));
    for (size_t i = 0; i < ent4.size(); i++) ent4[i] = make_uint4(bh.ent_start[i], bh.ent_end[i], bh.ent_payload[i], (uint32_t)i);
    for (size_t i = 0; i < bh.list.size(); i++) list4[i] = ent4[bh.list[i]];
    std::vector<uint32_t> wblk(bh.n_entries()), wbase(bh.n_entries(), 0);
    for (size_t k = 0; k < wblk.size(); k++) {
        wblk[k] = bh.ent_item[k] == 0xFFFFFFFFu ? 0xFFFFFFFFu : H->root_blk[bh.ent_item[k]];
        if (bh.ent_item[k] != 0xFFFFFFFFu) wbase[k] = V.seq[R[bh.ent_item[k]].seq].base;
    }
    upload(d_wj_win0, jwin0, st);
    upload(d_wj_winseq, bh.win_seq, st);
    upload(d_wj_ent4, ent4, st);
    upload(d_wj_list4, list4, st);
    upload(d_wj_fid, bh.ent_payload, st);
    upload(d_wj_first, bh.first, st);
    upload(d_wj_boff, bh.boff, st);
    upload(d_wj_blk, wblk, st);
    upload(d_wj_base, wbase, st);
    {   // rootjoin.cuh: per window, where its root run ends — the first interval of the sequence that starts at or beyond the
        // window's end + the longest region the kernel's normal part admits (no such region reaches it); the sequence's
        // sentinel for the last window, whose regions may start anywhere beyond the grid
        const uint32_t fsh = jshift - WR_FB;
        const uint64_t taillen = std::min<uint64_t>((uint64_t)wj_tail_bins << fsh, 0xFFFFFFFFull);
        std::vector<uint32_t> rend(bh.n_win, 0);
        size_t e0 = 0;   // first entry of the current sequence
        for (uint32_t v = 0; v < n_v; v++) {
            size_t sent = e0;
            while (bh.ent_item[sent] != 0xFFFFFFFFu) sent++;
            for (uint32_t w = jwin0[v]; w < jwin0[v + 1]; w++) {
                if (w + 1 == jwin0[v + 1]) { rend[w] = (uint32_t)sent; continue; }
                const uint64_t lim = ((uint64_t)(w - jwin0[v] + 1) << jshift) + taillen;
                rend[w] = (uint32_t)(std::lower_bound(bh.ent_start.begin() + (long)e0, bh.ent_start.begin() + (long)sent, lim,
                                                      [](uint32_t a, uint64_t b) { return (uint64_t)a < b; }) - bh.ent_start.begin());
            }
            e0 = sent + 1;
        }
        upload(d_wj_rend, rend, st);
        GFFX_CUDA(cudaStreamSynchronize(st));
    }
    {   // block -> the tree intervals that name it (wj_select_blocks_kernel gathers); intervals without a .gof record
        const uint32_t nb = (uint32_t)H->blocks.size();
        std::vector<uint32_t> off(nb + 1, 0), ent, orph;
        for (size_t k = 0; k < wblk.size(); k++)
            if (bh.ent_item[k] != 0xFFFFFFFFu) { if (wblk[k] != 0xFFFFFFFFu) off[wblk[k] + 1]++; else orph.push_back((uint32_t)k); }
        for (uint32_t b = 0; b < nb; b++) off[b + 1] += off[b];
        ent.resize(off[nb]);
        std::vector<uint32_t> fill(off.begin(), off.end() - 1);
        for (size_t k = 0; k < wblk.size(); k++)
            if (bh.ent_item[k] != 0xFFFFFFFFu && wblk[k] != 0xFFFFFFFFu) ent[fill[wblk[k]]++] = (uint32_t)k;
        n_orphans = (uint32_t)orph.size();
        upload(d_blk_ent0, off, st);
        upload(d_blk_ent, ent, st);
        upload(d_orphans, orph, st);
        GFFX_CUDA(cudaStreamSynchronize(st));
    }
    wj_n_ent = (uint32_t)bh.n_entries();
    d_wj_hit.ensure(std::max<uint32_t>(wj_n_ent, 1));   // grown to flag_bytes() once the line table is known (constructor)
    d_wj_cnt.ensure(std::max<uint32_t>(wj_n_ent, 1) * 8ull);
    d_wj_off.ensure(((uint64_t)wj_n_ent + 1) * 8);
    GFFX_CUDA(cudaStreamSynchronize(st));
    WC = WinBins{};
    WC.win_seq = d_wj_winseq.as<uint32_t>();
    WC.win0 = d_wj_win0.as<uint32_t>();
    WC.ent4 = d_wj_ent4.as<uint4>();
    WC.ent_fid = d_wj_fid.as<uint32_t>();
    WC.list4 = d_wj_list4.as<uint4>();
    WC.first = d_wj_first.as<uint32_t>();
    WC.boff = d_wj_boff.as<uint32_t>();
    WC.run_end = d_wj_rend.as<uint32_t>();
    WC.shift = bh.shift;
    WC.sub = bh.sub;
    WC.n_win = bh.n_win;
    WC.n_ent = wj_n_ent;
}

void EngineImpl::build_blocks() {
    n_blocks = (uint32_t)H->blocks.size();
    std::vector<uint64_t> bs(n_blocks), be(n_blocks);
    for (uint32_t b = 0; b < n_blocks; b++) { bs[b] = H->blocks[b].start; be[b] = H->blocks[b].end; }
    upload(d_bstart, bs, st);
    upload(d_bend, be, st);
    d_blk_sel.ensure(std::max<uint32_t>(n_blocks, 1));
    d_ranges.ensure(std::max<uint32_t>(n_blocks, 1) * 16ull);
    GFFX_CUDA(cudaStreamSynchronize(st));
}

// a piece of coordinates, ascending: three counting passes over 11-bit digits (a piece is a few ten thousand values: the digit
// tables and both buffers stay in cache), std::sort below a few hundred values
static void radix_sort_u32(uint32_t* v, size_t n) {
    if (n < 512) { std::sort(v, v + n); return; }
    static thread_local std::vector<uint32_t> tmp;
    if (tmp.size() < n) tmp.resize(n);
    uint32_t* a = v;
    uint32_t* b = tmp.data();
    uint32_t cnt[3][2048];
    memset(cnt, 0, sizeof cnt);
    for (size_t i = 0; i < n; i++) { const uint32_t x = a[i]; cnt[0][x & 2047u]++; cnt[1][(x >> 11) & 2047u]++; cnt[2][x >> 22]++; }
    for (int p = 0; p < 3; p++) {
        uint32_t* c = cnt[p];
        if (p == 2 && c[a[0] >> 22] == n) break;   // the values share their top digit: two passes were enough
        uint32_t run = 0;
        for (int d = 0; d < 2048; d++) { const uint32_t k = c[d]; c[d] = run; run += k; }
        const int sh = 11 * p;
        for (size_t i = 0; i < n; i++) { const uint32_t x = a[i]; b[c[(x >> sh) & 2047u]++] = x; }
        std::swap(a, b);
    }
    if (a != v) memcpy(v, a, n * 4);
}

// Lines + the filter rank space (see filter.cuh).
void EngineImpl::build_lines() {
    const LineTable& L = H->lines;
    const VSeqMap& V = H->vmap;
    // the rank space lives on the VIRTUAL sequences: a line's sequence, start and end move with its sequence's base
    const size_t n = L.size(), n_seqs = V.n_vseqs();
    std::vector<uint32_t> vS(n, 0), vE(n, 0), vseq(n, 0);
    // per sequence: sorted distinct S/E of usable lines (host threads: sequences are sorted side by side, longest first)
    const size_t T = (size_t)std::max(1, H->n_threads);
    auto run_parallel = [&](size_t n_jobs, const std::function<void(size_t)>& f) { parallel_jobs(n_jobs, (int)T, f); };   // util.h: the thread pool
    // the columns that do not depend on the rank space cross PCIe from a helper thread while the host builds it (pageable
    // copies occupy their calling thread)
    d_ln_flags.ensure(std::max<size_t>(n, 1)); d_ln_blk.ensure(std::max<size_t>(n, 1) * 4); d_ln_type.ensure(std::max<size_t>(n, 1) * 4);
    d_ln_off.ensure(std::max<size_t>(n, 1) * 8); d_ln_len.ensure(std::max<size_t>(n, 1) * 4);
    d_ln_iseq.ensure(std::max<size_t>(n, 1) * 4); d_ln_S.ensure(std::max<size_t>(n, 1) * 4); d_ln_E.ensure(std::max<size_t>(n, 1) * 4);
    cudaError_t up_err = cudaSuccess;
    std::atomic<int> coords{0};   // 1: the lines' virtual coordinates are final (the helper uploads them too), -1: given up
    struct Joiner {
        std::thread t;
        std::atomic<int>* coords;
        ~Joiner() {
            int none = 0;
            coords->compare_exchange_strong(none, -1);   // leaving early (an exception): the helper must not wait for them
            if (t.joinable()) t.join();
        }
    } early{std::thread(), &coords};
    early.t = std::thread([&] {
        auto put = [&](DBuf& b, const void* p, size_t bytes) {
            if (up_err == cudaSuccess && bytes) up_err = cudaMemcpyAsync(b.p, p, bytes, cudaMemcpyHostToDevice, st_copy);
        };
        up_err = cudaSetDevice(dev);
        put(d_ln_flags, L.flags.data(), n); put(d_ln_blk, L.blk.data(), n * 4); put(d_ln_type, L.type_id.data(), n * 4);
        put(d_ln_off, L.off.data(), n * 8); put(d_ln_len, L.len.data(), n * 4);
        while (coords.load() == 0) std::this_thread::yield();
        if (coords.load() == 1) { put(d_ln_iseq, vseq.data(), n * 4); put(d_ln_S, vS.data(), n * 4); put(d_ln_E, vE.data(), n * 4); }
        if (up_err == cudaSuccess) up_err = cudaStreamSynchronize(st_copy);
    });
    const size_t n_chunks = n >= (1u << 16) ? T : 1;   // the line table in T pieces of consecutive lines
    auto chunk_lo = [&](size_t c) { return n * c / n_chunks; };
    std::vector<std::vector<uint32_t>> chunk_cnt(n_chunks, std::vector<uint32_t>(n_seqs, 0));   // values a chunk adds to each sequence's set
    {
        PhaseTimer t("  rank space: virtual coordinates");
        run_parallel(n_chunks, [&](size_t c) {
            std::vector<uint32_t>& cnt = chunk_cnt[c];
            for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++)
                if (L.flags[i] & LN_I_OK) {
                    const VSeqMap::Seq& m = V.seq[L.iseq[i]];
                    vseq[i] = m.vseq; vS[i] = m.base + L.S[i]; vE[i] = m.base + L.E[i];
                    cnt[m.vseq] += 2;
                }
        });
    }
    coords.store(1);
    std::vector<std::vector<uint32_t>> xs(n_seqs);
    {
        PhaseTimer t("  rank space: coordinate sets (sort)");
        // every chunk writes its values straight to its own place in the sequences' arrays
        for (size_t s = 0; s < n_seqs; s++) {
            size_t tot = 0;
            for (size_t c = 0; c < n_chunks; c++) { const uint32_t k = chunk_cnt[c][s]; chunk_cnt[c][s] = (uint32_t)tot; tot += k; }
            xs[s].resize(tot);
        }
        run_parallel(n_chunks, [&](size_t c) {
            std::vector<uint32_t>& at = chunk_cnt[c];
            for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++)
                if (L.flags[i] & LN_I_OK) {
                    uint32_t* v = xs[vseq[i]].data() + at[vseq[i]];
                    v[0] = vS[i]; v[1] = vE[i];
                    at[vseq[i]] += 2;
                }
        });
        // a sequence's array is n_chunks pieces (chunk c's values end at chunk_cnt[c][s] now), each nearly sorted already when the
        // file is: the pieces are sorted side by side, then merged pairwise in log2(n_chunks) rounds — a long sequence no longer
        // occupies one thread for the whole sort
        struct Job { uint32_t s, lo, mid, hi; };
        auto run_jobs = [&](const std::vector<Job>& jobs, bool merge) {
            std::atomic<size_t> next{0};
            run_parallel(std::min(T, std::max<size_t>(jobs.size(), 1)), [&](size_t) {
                for (size_t k; (k = next.fetch_add(1)) < jobs.size();) {
                    uint32_t* v = xs[jobs[k].s].data();
                    if (merge) std::inplace_merge(v + jobs[k].lo, v + jobs[k].mid, v + jobs[k].hi);
                    else radix_sort_u32(v + jobs[k].lo, jobs[k].hi - jobs[k].lo);
                }
            });
        };
        std::vector<std::vector<uint32_t>> cut(n_seqs);   // piece boundaries of every sequence
        std::vector<Job> jobs;
        for (size_t s = 0; s < n_seqs; s++) {
            cut[s].push_back(0);
            for (size_t c = 0; c < n_chunks; c++)
                if (chunk_cnt[c][s] != cut[s].back()) cut[s].push_back(chunk_cnt[c][s]);
            for (size_t k = 0; k + 1 < cut[s].size(); k++) jobs.push_back(Job{(uint32_t)s, cut[s][k], 0, cut[s][k + 1]});
        }
        std::sort(jobs.begin(), jobs.end(), [](const Job& a, const Job& b) { return a.hi - a.lo > b.hi - b.lo; });
        run_jobs(jobs, false);
        for (;;) {
            jobs.clear();
            for (size_t s = 0; s < n_seqs; s++) {
                std::vector<uint32_t>& c = cut[s];
                if (c.size() <= 2) continue;
                std::vector<uint32_t> nc;
                for (size_t k = 0; k + 2 < c.size(); k += 2) { jobs.push_back(Job{(uint32_t)s, c[k], c[k + 1], c[k + 2]}); nc.push_back(c[k]); }
                if (c.size() % 2 == 0) nc.push_back(c[c.size() - 2]);   // an odd piece waits for the next round
                nc.push_back(c.back());
                c.swap(nc);
            }
            if (jobs.empty()) break;
            run_jobs(jobs, true);
        }
        {
            std::atomic<size_t> next{0};
            run_parallel(std::min(T, n_seqs), [&](size_t) {
                for (size_t k; (k = next.fetch_add(1)) < n_seqs;) {
                    auto& v = xs[k];
                    v.erase(std::unique(v.begin(), v.end()), v.end());
                }
            });
        }
    }
    PhaseTimer* t_xd = new PhaseTimer("  rank space: X + directories");
    std::vector<RankSeq> seqs(n_seqs);
    std::vector<uint32_t> X, dir;
    uint32_t slot = 0;
    {
        size_t nx = 0, nd = 0;
        for (size_t s = 0; s < n_seqs; s++) {   // layout first, so that the sequences (and pieces of the long ones) fill side by side
            auto& v = xs[s];
            RankSeq& r = seqs[s];
            r.x0 = (uint32_t)nx;
            r.cnt = (uint32_t)v.size();
            r.slot0 = slot;
            slot += r.cnt + 1;
            // a name listed twice in .sqs resolves to its LAST number (intersect.rs:622-625): regions carrying
            // the earlier number never reach the line filter (such sequences share virtual sequences among themselves)
            r.canon = V.v_canon[s];
            r.dir_base = (uint32_t)nd;
            if (v.empty()) { r.nb = 0; r.shift = 0; nd += 1; continue; }
            uint32_t maxv = v.back(), shift = 4;
            while (((uint64_t)maxv >> shift) + 1 > std::max<uint64_t>(64, 2ull * v.size())) shift++;
            r.shift = shift;
            r.nb = (maxv >> shift) + 1;
            nd += (size_t)r.nb + 1;   // nb+1 entries: dir[b] = #(X < b<<shift)
            nx += v.size();
        }
        X.assign(nx + 8, 0u);   // rank_lb (filter.cuh) reads up to 4 entries past a bin
        dir.assign(nd, 0u);
        struct Piece { uint32_t s, b0, b1; };
        std::vector<Piece> pieces;
        constexpr uint32_t PIECE = 1u << 16;   // directory cells per task
        for (size_t s = 0; s < n_seqs; s++)
            for (uint32_t b0 = 0; b0 <= seqs[s].nb && !xs[s].empty(); b0 += PIECE) pieces.push_back(Piece{(uint32_t)s, b0, std::min(seqs[s].nb + 1, b0 + PIECE)});
        std::atomic<size_t> next{0};
        run_parallel(std::min(T, std::max<size_t>(pieces.size(), 1)), [&](size_t) {
            for (size_t k; (k = next.fetch_add(1)) < pieces.size();) {
                const Piece pc = pieces[k];
                const auto& v = xs[pc.s];
                const RankSeq& r = seqs[pc.s];
                uint32_t* d = dir.data() + r.dir_base;
                size_t j = (size_t)(std::lower_bound(v.begin(), v.end(), (uint64_t)pc.b0 << r.shift, [](uint32_t a, uint64_t c) { return (uint64_t)a < c; }) - v.begin());
                const size_t j0 = j;
                for (uint32_t b = pc.b0; b < pc.b1; b++) {
                    const uint64_t lo = (uint64_t)b << r.shift;
                    while (j < v.size() && (uint64_t)v[j] < lo) j++;
                    d[b] = (uint32_t)j;
                }
                // the piece's values of X: from its first cell's rank up to the next piece's first cell's (the last piece: the rest)
                const uint64_t lim = (uint64_t)pc.b1 << r.shift;
                while (j < v.size() && (uint64_t)v[j] < lim) j++;
                if (j > j0) memcpy(X.data() + r.x0 + j0, v.data() + j0, (j - j0) * 4);
            }
        });
    }
    f_slots = slot;
    delete t_xd;
    {   // rootjoin.cuh: per window of the partition grid, its part of X and the fine-bin offsets into it
        PhaseTimer t_xo("  rank space: window slices");
        const uint32_t n_win = WG.n_win, fsh = WG.shift - WR_FB;
        std::vector<uint32_t> xw0(std::max<uint32_t>(n_win, 1), 0), xlen(std::max<uint32_t>(n_win, 1), 0xFFFFFFFFu);
        std::vector<uint16_t> xoff((size_t)std::max<uint32_t>(n_win, 1) * WR_XOFF_STRIDE, 0);
        std::vector<uint32_t> win_v(n_win, 0);
        for (size_t v = 0; v < n_seqs; v++)
            for (uint32_t w = h_win0[v]; w < h_win0[v + 1]; w++) win_v[w] = (uint32_t)v;
        std::atomic<uint32_t> next{0};
        run_parallel(std::min<size_t>(T, std::max<uint32_t>(n_win, 1)), [&](size_t) {
            for (uint32_t w; (w = next.fetch_add(1)) < n_win;) {
                const uint32_t v = win_v[w];
                const auto& xv = xs[v];
                const uint64_t lo = (uint64_t)(w - h_win0[v]) << WG.shift;
                const bool last = w + 1 == h_win0[v + 1];
                auto below_in = [&](uint64_t c, size_t a, size_t b) {
                    return (size_t)(std::lower_bound(xv.begin() + (long)a, xv.begin() + (long)b, c, [](uint32_t x, uint64_t y) { return (uint64_t)x < y; }) - xv.begin());
                };
                const size_t i0 = below_in(lo, 0, xv.size()), i1 = last ? xv.size() : below_in(lo + (1ull << WG.shift), i0, xv.size());
                xw0[w] = seqs[v].x0 + (uint32_t)i0;
                if (i1 - i0 > WR_SLICE_CAP - 4 || i1 - i0 > 0xFFFFu) continue;   // keeps the global look-up
                xlen[w] = (uint32_t)(i1 - i0);
                uint16_t* row = xoff.data() + (size_t)w * WR_XOFF_STRIDE;
                size_t j = i0;   // the fine bins ascend: one walk over the window's part
                for (uint32_t b = 0; b < WR_NBF; b++) {
                    const uint64_t c = lo + ((uint64_t)b << fsh);
                    while (j < i1 && (uint64_t)xv[j] < c) j++;
                    row[b] = (uint16_t)(j - i0);
                }
                row[WR_NBF] = (uint16_t)(i1 - i0);
            }
        });
        upload(d_wr_xw0, xw0, st);
        upload(d_wr_xlen, xlen, st);
        upload(d_wr_xoff, xoff, st);
        GFFX_CUDA(cudaStreamSynchronize(st));
        XS.xw0 = d_wr_xw0.as<uint32_t>();
        XS.xlen = d_wr_xlen.as<uint32_t>();
        XS.xoff = d_wr_xoff.as<uint16_t>();
    }
    std::vector<uint32_t> iS(n, 0), iE(n, 0);
    PhaseTimer* t_rank = new PhaseTimer("  rank space: line ranks");
    run_parallel(n_chunks, [&](size_t c) {
        // lb through the sequence's directory: the cell of the value's bin bounds the search to the bin's few values
        auto rank_of = [&](const RankSeq& r, uint32_t v) {
            const uint32_t b = v >> r.shift;   // a line's coordinate is in the set: b < nb
            const uint32_t* x = X.data() + r.x0;
            const uint32_t lo = dir[r.dir_base + b], hi = dir[r.dir_base + b + 1];
            return r.slot0 + (uint32_t)(std::lower_bound(x + lo, x + hi, v) - x);
        };
        for (size_t i = chunk_lo(c); i < chunk_lo(c + 1); i++)
            if (L.flags[i] & LN_I_OK) {
                const RankSeq& r = seqs[vseq[i]];
                iS[i] = rank_of(r, vS[i]);
                iE[i] = rank_of(r, vE[i]);
            }
    });
    delete t_rank;
    PhaseTimer t_up("  line table upload");
    upload(d_ln_iS, iS, st);
    upload(d_ln_iE, iE, st);
    upload(d_fX, X, st);
    upload(d_fdir, dir, st);
    upload(d_fseqs, seqs, st);
    GFFX_CUDA(cudaStreamSynchronize(st));
    early.t.join();
    GFFX_CUDA(up_err);
    t_up.stop();
    PhaseTimer t_al("  rank space: accumulators (alloc + clear)");
    const size_t ns = std::max<uint32_t>(f_slots, 1);
    d_A.ensure(ns * 8); d_B.ensure(ns * 8); d_PMX.ensure(ns * 8); d_SMN.ensure((ns + 1) * 8);
    d_Ms.ensure(ns * 2); d_Me.ensure(ns * 2); d_NS.ensure(ns * 8); d_NE.ensure(ns * 8);
    // cleared ONCE: jobs tag their keys and marks with an epoch instead (filter.cuh)
    GFFX_CUDA(cudaMemsetAsync(d_A.p, 0, ns * 8, st));
    GFFX_CUDA(cudaMemsetAsync(d_B.p, 0xFF, ns * 8, st));
    GFFX_CUDA(cudaMemsetAsync(d_Ms.p, 0, ns * 2, st));
    GFFX_CUDA(cudaMemsetAsync(d_Me.p, 0, ns * 2, st));
    epoch_start = 0;
    if (const char* e = getenv("GFFX_EPOCH_START"))   // test hook: start close to the wrap of both counters
        epoch_start = (uint32_t)std::min<uint64_t>(strtoull(e, nullptr, 10), EPOCH_MAX);
    epoch = epoch_start; epoch8 = epoch_start % 256u;
    d_keep.ensure(std::max<size_t>(n, 1));
    d_type_allow.ensure(std::max<size_t>(H->types.size(), 1));
    GFFX_CUDA(cudaStreamSynchronize(st));
    FR.X = d_fX.as<uint32_t>();
    FR.dir = d_fdir.as<uint32_t>();
    FR.seqs = d_fseqs.as<RankSeq>();
    FR.n_seqs = (uint32_t)n_seqs;
    LA.flags = d_ln_flags.as<uint8_t>();
    LA.blk = d_ln_blk.as<uint32_t>();
    LA.type_id = d_ln_type.as<uint32_t>();
    LA.iseq = d_ln_iseq.as<uint32_t>();
    LA.S = d_ln_S.as<uint32_t>();
    LA.E = d_ln_E.as<uint32_t>();
    LA.iS = d_ln_iS.as<uint32_t>();
    LA.iE = d_ln_iE.as<uint32_t>();
    LA.off = d_ln_off.as<uint64_t>();
    LA.len = d_ln_len.as<uint32_t>();
    LA.n = n;
}

// ---- pinned host pool ---------------------------------------------------------------------------
namespace {
struct PinnedPool {
    std::mutex mu;
    std::vector<std::pair<void*, size_t>> free_list;
    size_t pooled = 0;
    static constexpr size_t MAX_POOLED = 4ull << 30;
    ~PinnedPool() {
        // the CUDA context may already be gone at process exit: leave the blocks to the driver
    }
};
PinnedPool& pinned_pool() {
    static PinnedPool* P = new PinnedPool();  // never destroyed: results may outlive static destruction order
    return *P;
}
}  // namespace

void* pinned_acquire(size_t bytes, size_t* cap) {
    PinnedPool& P = pinned_pool();
    {
        std::lock_guard<std::mutex> g(P.mu);
        size_t best = SIZE_MAX;
        for (size_t i = 0; i < P.free_list.size(); i++)
            if (P.free_list[i].second >= bytes && (best == SIZE_MAX || P.free_list[i].second < P.free_list[best].second)) best = i;
        if (best != SIZE_MAX && P.free_list[best].second <= 4 * std::max<size_t>(bytes, 1 << 20)) {
            auto b = P.free_list[best];
            P.free_list.erase(P.free_list.begin() + (long)best);
            P.pooled -= b.second;
            *cap = b.second;
            return b.first;
        }
    }
    size_t want = std::max<size_t>(bytes, 1 << 16);
    want = (want + (want >> 3) + 0xFFFF) & ~(size_t)0xFFFF;  // 12 % slack: a slightly larger next result reuses the block
    void* p = nullptr;
    cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocPortable);
    if (e != cudaSuccess) {
        cudaGetLastError();
        throw Err(E_CUDA, std::string("cudaHostAlloc failed: ") + cudaGetErrorString(e));
    }
    *cap = want;
    return p;
}

void pinned_release(void* p, size_t cap) {
    PinnedPool& P = pinned_pool();
    {
        std::lock_guard<std::mutex> g(P.mu);
        if (P.pooled + cap <= PinnedPool::MAX_POOLED && P.free_list.size() < 64) {
            P.free_list.emplace_back(p, cap);
            P.pooled += cap;
            return;
        }
    }
    cudaFreeHost(p);
}

void EngineImpl::stage_queries(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags,
                               const uint32_t** dc, const uint32_t** ds, const uint32_t** de) {
    if (flags & 1u) {  // GFFX_F_DEVICE_INPUT
        *dc = chr; *ds = start; *de = end;
        return;
    }
    const uint32_t* src[3] = {chr, start, end};
    for (int k = 0; k < 3; k++) {
        d_q[k].ensure(std::max<uint64_t>(n, 1) * 4);
        if (n) GFFX_CUDA(cudaMemcpyAsync(d_q[k].p, src[k], n * 4, cudaMemcpyHostToDevice, st));
    }
    *dc = d_q[0].as<uint32_t>(); *ds = d_q[1].as<uint32_t>(); *de = d_q[2].as<uint32_t>();
}

template <int MODE, bool INV, bool FILT, bool SMALL>
static void launch_wj2(EngineImpl* E, const uint32_t* poff, const uint2* se, uint64_t n, unsigned tiles, unsigned long long* n_pairs, uint32_t* err) {
    typedef unsigned long long ull;
    FilterArrays F{E->d_A.as<ull>(), E->d_B.as<ull>(), E->d_Ms.as<uint8_t>(), E->d_Me.as<uint8_t>(), E->d_small.as<uint32_t>() + 48, E->epoch, E->epoch8};
    const dim3 grid(tiles, SMALL ? 16u : 1u);   // SMALL: 16 CTAs share a tile's small segments
    wj_count_kernel<MODE, INV, FILT, SMALL><<<grid, WJ_THREADS, WJ_DYN_SMEM, E->st>>>(E->WC, poff, se, n, E->d_wj_hit.as<uint8_t>(),
                                                                                       E->d_wj_cnt.as<ull>(), n_pairs, err, E->FR, F);
}
template <int MODE, bool INV, bool FILT>
static void launch_wj(EngineImpl* E, const uint32_t* poff, const uint2* se, uint64_t n, unsigned tiles, unsigned long long* n_pairs, uint32_t* err) {
    typedef unsigned long long ull;
    // fewer than ~64 regions per window on average: most (tile, window) segments are small -> the region-parallel kernel's SMALL form;
    // otherwise the root-parallel kernel (rootjoin.cuh).  GFFX_WJ_KERNEL=rounds|roots forces one of them (tests, A/B runs).
    const bool small = n < 64ull * std::max<uint32_t>(E->WC.n_win, 1);
    if (E->wj_force == 1 || (E->wj_force == 0 && small)) {
        if (small) launch_wj2<MODE, INV, FILT, true>(E, poff, se, n, tiles, n_pairs, err);
        else launch_wj2<MODE, INV, FILT, false>(E, poff, se, n, tiles, n_pairs, err);
        return;
    }
    FilterArrays F{E->d_A.as<ull>(), E->d_B.as<ull>(), E->d_Ms.as<uint8_t>(), E->d_Me.as<uint8_t>(), E->d_small.as<uint32_t>() + 48, E->epoch, E->epoch8};
    if (FILT && E->wj_slices)   // the window's part of the rank space in shared memory (rootjoin.cuh); GFFX_WR_SLICE=0: global look-ups
        wj_root_kernel<MODE, INV, FILT, FILT><<<tiles, WJ_THREADS, WJ_DYN_SMEM + WR_SLICE_SMEM, E->st>>>(E->WC, poff, se, n, E->d_wj_hit.as<uint8_t>(), E->d_wj_cnt.as<ull>(),
                                                                                                       n_pairs, err, E->FR, F, E->wj_tail_bins, E->XS);
    else
        wj_root_kernel<MODE, INV, FILT, false><<<tiles, WJ_THREADS, WJ_DYN_SMEM, E->st>>>(E->WC, poff, se, n, E->d_wj_hit.as<uint8_t>(), E->d_wj_cnt.as<ull>(), n_pairs, err,
                                                                                          E->FR, F, E->wj_tail_bins, E->XS);
}
// `filt`: the line filter's rank-space scatter (K3a) rides on the same pass
static void launch_wj_mode(EngineImpl* E, int mode, bool invert, bool filt, const uint32_t* poff, const uint2* se, uint64_t n, unsigned tiles,
                           unsigned long long* n_pairs, uint32_t* err) {
    switch ((mode * 2 + (invert ? 1 : 0)) * 2 + (filt ? 1 : 0)) {
        case 0: launch_wj<0, false, false>(E, poff, se, n, tiles, n_pairs, err); break;
        case 1: launch_wj<0, false, true>(E, poff, se, n, tiles, n_pairs, err); break;
        case 2: launch_wj<0, true, false>(E, poff, se, n, tiles, n_pairs, err); break;
        case 3: launch_wj<0, true, true>(E, poff, se, n, tiles, n_pairs, err); break;
        case 4: launch_wj<1, false, false>(E, poff, se, n, tiles, n_pairs, err); break;
        case 5: launch_wj<1, false, true>(E, poff, se, n, tiles, n_pairs, err); break;
        case 6: launch_wj<1, true, false>(E, poff, se, n, tiles, n_pairs, err); break;
        case 7: launch_wj<1, true, true>(E, poff, se, n, tiles, n_pairs, err); break;
        case 8: launch_wj<2, false, false>(E, poff, se, n, tiles, n_pairs, err); break;
        default: launch_wj<2, false, true>(E, poff, se, n, tiles, n_pairs, err); break;
    }
}

// K1b: per-root counters -> first-pair offsets -> the pair array (winjoin.cuh); nothing is written when the total exceeds `cap`
uint32_t EngineImpl::expand_pairs(unsigned long long cap) {
    typedef unsigned long long ull;
    d_scan_scratch.ensure(scan_scratch_elems((uint64_t)wj_n_ent + 1) * 8);
    uint32_t launches = device_scan<ull, OpSum, WjCntLoad, PtrStore<ull>, false, false>(
        WjCntLoad{d_wj_cnt.as<ull>(), wj_n_ent}, PtrStore<ull>{d_wj_off.as<ull>()}, (uint64_t)wj_n_ent + 1, OpSum(), 0ull,
        d_scan_scratch.as<ull>(), nullptr, st, nullptr, chain());
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    wj_expand_kernel<<<(unsigned)sms * 8, WJ_EXP_THREADS, 0, st>>>(d_wj_off.as<ull>(), WC.ent4, d_wj_base.as<uint32_t>(), wj_n_ent, d_pairs.as<uint32_t>(), cap);
    return launches + 1;
}

// ------------------------------------------------------------------------------------------------
// One intersect JOB = ix_begin, any number of ix_add (each a batch of device-resident regions: K0 partition, K1 join,
// K3a scatter onto the line rank space — all of them accumulate), ix_finish (root -> block selection, rank-space scans,
// line filter, ordered compaction, read-back).  gffx_intersect is one job whose batches are the chunks of the host ->
// device copy (the copy of chunk i+1 overlaps the kernels of chunk i); the shard API (gffx_shard_*) exposes the three
// steps so that the accumulators of several GPUs can be combined before ix_finish.
void EngineImpl::ix_begin(int mode, bool invert, bool entire_group, const char* types_csv, uint32_t flags, bool join_only,
                          uint64_t n_retained, uint32_t n_chunks) {
    if (mode < 0 || mode > 2) throw Err(E_ARG, "mode must be 0 (contained), 1 (contains-region) or 2 (overlap)");
    GFFX_CUDA(cudaSetDevice(dev));
    if ((flags & 32u) && n_lanes > 1) {   // queued jobs alternate between the lanes; everything else runs on lane 0
        lane_activate(lane_next);
        lane_next = (lane_next + 1) % n_lanes;
    } else {
        lane_activate(0);
    }
    IxJob& jb = job;
    jb = IxJob();
    jb.mode = mode; jb.invert = invert; jb.entire_group = entire_group; jb.join_only = join_only;
    jb.want_pairs_host = flags & 2u;
    jb.pairs_dev = (flags & 8u) || jb.want_pairs_host;
    jb.no_host = flags & 4u;
    jb.stage_events = flags & 16u;
    jb.match_only = !entire_group || types_csv != nullptr;  // intersect.rs:619
    jb.nothing = (mode == 2 && invert) || wj_n_ent == 0;    // -I with overlap selects nothing (intersect.rs:161)
    jb.run_filter = !jb.nothing && !join_only && jb.match_only && LA.n > 0;
    jb.have_types = types_csv != nullptr;
    jb.retained = n_retained;
    jb.active = true;
    while (ev_chunk.size() < 4 * (size_t)std::max<uint32_t>(n_chunks, 1)) {
        cudaEvent_t e;
        GFFX_CUDA(cudaEventCreate(&e));
        ev_chunk.push_back(e);
    }
    if (n_retained) {  // the bucketed form of every batch is kept (a too-small pair buffer re-runs the join)
        d_bse.ensure(n_retained * 8 + 32);
    }
    d_poff.ensure((size_t)std::max<uint32_t>(n_chunks, 1) * (WG.n_win + 2) * 4);
    jb.async = flags & 32u;   // GFFX_F_ASYNC: the counters are read back into a slot of the engine's ring, the caller waits later
    if (jb.async) {
        if (jb.want_pairs_host) throw Err(E_ARG, "GFFX_F_ASYNC does not copy pairs to the host: use GFFX_F_PAIRS_ON_DEVICE, not GFFX_F_WANT_PAIRS");
        if (jb.stage_events) throw Err(E_ARG, "GFFX_F_ASYNC and GFFX_F_STAGE_TIMES exclude each other");
        jb.slot = slot_acquire();
    }
    GFFX_CUDA(cudaEventRecord(jb.async ? slots[jb.slot].ev_start : ev[0], st));
    {
        const uint32_t n_hit16 = (wj_n_ent + 15u) / 16u, n_cnt16 = (wj_n_ent + 1u) / 2u;   // both buffers are padded to whole 16-byte words
        const uint32_t threads = std::max<uint32_t>(std::max(n_hit16, 16u), std::min<uint32_t>(n_cnt16, 148u * 1024u));
        clear_job_kernel<<<(threads + 255) / 256, 256, 0, st>>>(d_small.as<uint4>(), d_wj_hit.as<uint4>(), n_hit16, d_wj_cnt.as<uint4>(), n_cnt16);
        jb.launches++;
    }
    if (jb.run_filter) {
        if (jb.have_types) {  // -T: comma-split, trimmed, non-empty tokens (intersect.rs:252-259)
            const size_t nt = std::max<size_t>(H->types.size(), 1);
            std::vector<uint8_t> allow_local;
            uint8_t* allow;
            if (jb.async) {   // the slot's page-locked bytes: a pageable source would synchronise the stream
                if (nt > ASYNC_ALLOW_CAP) throw Err(E_ARG, "GFFX_F_ASYNC with -T supports up to 4096 feature types");
                allow = slots[jb.slot].h + ASYNC_ALLOW_OFF;
                memset(allow, 0, nt);
            } else {
                allow_local.assign(nt, 0);
                allow = allow_local.data();
            }
            std::string_view sv(types_csv);
            size_t i = 0;
            while (i <= sv.size()) {
                size_t j = sv.find(',', i);
                if (j == std::string_view::npos) j = sv.size();
                std::string_view t = trim_unicode(sv.substr(i, j - i));
                if (!t.empty()) {
                    int64_t id = H->types.find(t);
                    if (id >= 0) allow[(size_t)id] = 1;
                }
                i = j + 1;
            }
            GFFX_CUDA(cudaMemcpyAsync(d_type_allow.p, allow, nt, cudaMemcpyHostToDevice, st));
            if (!jb.async) GFFX_CUDA(cudaStreamSynchronize(st));  // `allow_local` is a stack-local
        }
        // a new epoch instead of clearing the rank-space arrays (filter.cuh); a counter that wraps clears its arrays once
        const uint64_t ns = f_slots;
        if (epoch == EPOCH_MAX) {
            GFFX_CUDA(cudaMemsetAsync(d_A.p, 0, ns * 8, st));
            GFFX_CUDA(cudaMemsetAsync(d_B.p, 0xFF, ns * 8, st));
            epoch = 0;
        }
        epoch++;
        if (mode == 2) {
            if (epoch8 == 255u) {
                GFFX_CUDA(cudaMemsetAsync(d_Ms.p, 0, ns * 2, st));
                GFFX_CUDA(cudaMemsetAsync(d_Me.p, 0, ns * 2, st));
                epoch8 = 0;
            }
            epoch8++;
        }
    }
}

// regions already on the device; asynchronous on the engine's stream
void EngineImpl::ix_add(const uint32_t* dc, const uint32_t* ds, const uint32_t* de, uint64_t n) {
    IxJob& jb = job;
    if (!jb.active) throw Err(E_STATE, "intersect batch without a job");
    if (n == 0) return;
    if (jb.n_total + n >= (1ull << 40)) throw Err(E_ARG, "too many regions in one job");
    const uint32_t slot = (uint32_t)jb.chunks.size();
    const bool timed = jb.stage_events && 4 * (size_t)slot + 3 < ev_chunk.size();
    IxChunk ck;
    ck.n = n;
    ck.off = jb.retained ? jb.n_total : 0;
    if (jb.retained) {
        if (jb.n_total + n > jb.retained) throw Err(E_STATE, "more regions than announced at the job's begin");
        if ((size_t)(slot + 1) * (WG.n_win + 2) * 4 > d_poff.cap) throw Err(E_STATE, "more batches than announced at the job's begin");
    } else {
        d_bse.ensure(n * 8 + 32);
    }
    uint2* ose = d_bse.as<uint2>() + ck.off;
    uint32_t* poff = d_poff.as<uint32_t>() + (jb.retained ? (size_t)slot * (WG.n_win + 2) : 0);
    jb.n_total += n;
    jb.chunks.push_back(ck);
    if (jb.nothing) return;
    unsigned long long* small = d_small.as<unsigned long long>();
    uint32_t* small32 = d_small.as<uint32_t>();
    if (timed) GFFX_CUDA(cudaEventRecord(ev_chunk[4 * slot], st));
    jb.launches += partition_into(dc, ds, de, n, ose, poff);
    if (timed) GFFX_CUDA(cudaEventRecord(ev_chunk[4 * slot + 1], st));
    if (jb.pairs_dev && jb.pair_cap == 0) {
        // grow-only; a larger result re-runs the expansion (or, for an asynchronous job, reports its pairs as not materialised)
        d_pairs.ensure(std::max<uint64_t>({2 * std::max<uint64_t>(jb.retained, n), pairs_hwm + pairs_hwm / 8, (uint64_t)1024}) * 12);
        jb.pair_cap = d_pairs.cap / 12;
    }
    const unsigned tiles = (unsigned)((n + WJ_TILE - 1) / WJ_TILE);
    launch_wj_mode(this, jb.mode, jb.invert, jb.run_filter, poff, ose, n, tiles, small, small32 + 12);   // K1 (+ K3a when the line filter runs)
    jb.launches++;
    if (timed) GFFX_CUDA(cudaEventRecord(ev_chunk[4 * slot + 2], st));
    if (timed) GFFX_CUDA(cudaEventRecord(ev_chunk[4 * slot + 3], st));
}

// the rank-space scans of the line filter (filter.cuh): PMX / SMN / NS / NE from the scattered region keys
uint32_t EngineImpl::ix_scans(FilterScans* Pout) {
    typedef unsigned long long ull;
    const IxJob& jb = job;
    const int mode = jb.mode;
    const uint64_t ns = f_slots;
    uint32_t launches = 0;
    d_scan_scratch.ensure(scan_scratch_elems(LA.n / LINE_THREADS + 1) * 8 + scan_scratch_elems(2ull * f_slots + 2) * 8);
    FilterArrays F{d_A.as<ull>(), d_B.as<ull>(), d_Ms.as<uint8_t>(), d_Me.as<uint8_t>(), d_small.as<uint32_t>() + 48, epoch, epoch8};
    ull* scr = d_scan_scratch.as<ull>();
    *Pout = FilterScans{d_PMX.as<ull>(), d_SMN.as<ull>(), d_NS.as<uint32_t>(), d_NE.as<uint32_t>(), F.any_inv, epoch};
    if (mode == 0 || mode == 2)
        launches += device_scan<ull, OpMax, PtrLoad<ull>, PtrStore<ull>, false, true, 16>(PtrLoad<ull>{F.A}, PtrStore<ull>{d_PMX.as<ull>()}, ns,
                                                                                     OpMax(), 0ull, scr, nullptr, st, nullptr, chain());
    if (mode == 1) {
        launches += device_scan<ull, OpMin, PtrLoad<ull>, PtrStore<ull>, true, true, 16>(PtrLoad<ull>{F.B}, PtrStore<ull>{d_SMN.as<ull>()}, ns,
                                                                                    OpMin(), ~0ull, scr, nullptr, st, nullptr, chain());
        fill_u64_kernel<<<1, 32, 0, st>>>(d_SMN.as<ull>() + ns, 1, ~0ull);  // SMN[iS+1] may index one past the last slot
        launches++;
    }
    if (mode == 2) {
        launches += device_scan<uint32_t, OpMin, MarkLoad, PtrStore<uint32_t>, true, true, 16>(
            MarkLoad{F.Ms, epoch8}, PtrStore<uint32_t>{d_NS.as<uint32_t>()}, 2 * ns, OpMin(), 0xFFFFFFFFu, (uint32_t*)scr, nullptr, st, nullptr, chain());
        launches += device_scan<uint32_t, OpMin, MarkLoad, PtrStore<uint32_t>, true, true, 16>(
            MarkLoad{F.Me, epoch8}, PtrStore<uint32_t>{d_NE.as<uint32_t>()}, 2 * ns, OpMin(), 0xFFFFFFFFu, (uint32_t*)scr, nullptr, st, F.any_inv, chain());
    }
    return launches;
}

// No more batches: the shard-local half of the finish step.  Afterwards the job's state that other shards can add to is
// two byte-flag arrays — which roots were hit, which lines match some region — both combined by MAX.
void EngineImpl::ix_seal() {
    IxJob& jb = job;
    if (!jb.active) throw Err(E_STATE, "no intersect job is open");
    if (jb.sealed) return;
    GFFX_CUDA(cudaSetDevice(dev));
    jb.sealed = true;
    if (!jb.run_filter) return;
    if (jb.nothing || jb.n_total == 0) {
        GFFX_CUDA(cudaMemsetAsync(match_flags(), 0, std::max<uint64_t>(LA.n, 1), st));
        return;
    }
    FilterScans P{};
    jb.launches += ix_scans(&P);
    const unsigned tiles = (unsigned)((LA.n + LINE_TILE - 1) / LINE_TILE);
    if (jb.mode == 0) match_lines_kernel<0><<<tiles, LINE_THREADS, 0, st>>>(LA, P, match_flags());
    else if (jb.mode == 1) match_lines_kernel<1><<<tiles, LINE_THREADS, 0, st>>>(LA, P, match_flags());
    else match_lines_kernel<2><<<tiles, LINE_THREADS, 0, st>>>(LA, P, match_flags());
    jb.launches++;
}

void EngineImpl::ix_finish(IntersectResult* out) {
    IxJob& jb = job;
    if (!jb.active) throw Err(E_STATE, "intersect finish without a job");
    jb.active = false;
    GFFX_CUDA(cudaSetDevice(dev));
    *out = IntersectResult();
    out->n_queries = jb.n_total;
    const int mode = jb.mode;
    // a sealed job may hold other shards' flags even when this shard added no regions
    const bool nothing = jb.nothing || (jb.n_total == 0 && !jb.sealed);
    uint32_t launches = jb.launches;
    unsigned long long* small = d_small.as<unsigned long long>();  // [0] n_pairs [1..3] merge counts [4] bytes_out [5] n_emit (u64); u32 view: [12] err [13] n_warn
    uint32_t* small32 = d_small.as<uint32_t>();
    if (jb.stage_events) GFFX_CUDA(cudaEventRecord(ev[3], st));

    // ---- root -> block selection, merged ranges (-e) --------------------------------------------
    if (!nothing) {
        d_scan_scratch.ensure(scan_scratch_elems(LA.n / LINE_THREADS + 1) * 8 + scan_scratch_elems(2ull * f_slots + 2) * 8);
        // every block looks at the hit flags of the tree intervals that name it (no flag array to clear first); the intervals
        // whose fid has no .gof record go to the warn list; the selected blocks are counted here unless the merge does it
        const bool merge = n_blocks && !jb.match_only && !jb.join_only;
        const uint32_t n_thr = std::max<uint32_t>(std::max(n_blocks, n_orphans), 1);
        wj_select_blocks_kernel<<<(n_thr + 255) / 256, 256, 0, st>>>(d_wj_hit.as<uint8_t>(), d_blk_ent0.as<uint32_t>(), d_blk_ent.as<uint32_t>(), n_blocks,
                                                                     d_bstart.as<uint64_t>(), d_orphans.as<uint32_t>(), n_orphans, WC.ent_fid,
                                                                     d_blk_sel.as<uint8_t>(), merge ? nullptr : small + 2, d_warn.as<uint32_t>(),
                                                                     small32 + 13, WARN_CAP);
        launches++;
        if (merge) {
            merge_blocks_kernel<<<1, 1024, 0, st>>>(d_bstart.as<uint64_t>(), d_bend.as<uint64_t>(), d_blk_sel.as<uint8_t>(), n_blocks,
                                                   H->gff_len, true, d_ranges.as<ulonglong2>(), small + 1);
            launches++;
        }
    }
    if (jb.stage_events) GFFX_CUDA(cudaEventRecord(ev[4], st));

    // ---- line filter: rank-space scans, per-line decision, ordered compaction --------------------
    if (jb.run_filter && !nothing) {
        FilterScans P{};
        if (!jb.sealed) launches += ix_scans(&P);
        if (jb.stage_events) GFFX_CUDA(cudaEventRecord(ev[6], st));
        const uint64_t n_emit_tiles = (LA.n + LINE_TILE - 1) / LINE_TILE;
        d_tile_count.ensure(n_emit_tiles * 4);
        const uint8_t* sel = d_blk_sel.as<uint8_t>();
        const uint8_t* ta = d_type_allow.as<uint8_t>();
        const bool ht = jb.have_types;
        if (jb.sealed)  // the per-line match flags exist already (and may hold the other shards' matches too)
            keep_lines_kernel<<<(unsigned)n_emit_tiles, LINE_THREADS, 0, st>>>(LA, match_flags(), sel, ta, ht, d_keep.as<uint8_t>(), d_tile_count.as<uint32_t>());
        else if (mode == 0) eval_lines_kernel<0><<<(unsigned)n_emit_tiles, LINE_THREADS, 0, st>>>(LA, P, sel, ta, ht, d_keep.as<uint8_t>(), d_tile_count.as<uint32_t>());
        else if (mode == 1) eval_lines_kernel<1><<<(unsigned)n_emit_tiles, LINE_THREADS, 0, st>>>(LA, P, sel, ta, ht, d_keep.as<uint8_t>(), d_tile_count.as<uint32_t>());
        else eval_lines_kernel<2><<<(unsigned)n_emit_tiles, LINE_THREADS, 0, st>>>(LA, P, sel, ta, ht, d_keep.as<uint8_t>(), d_tile_count.as<uint32_t>());
        launches++;
        // the compacted list can never exceed the number of lines
        d_ranges.ensure(std::max<uint64_t>(LA.n, n_blocks) * 16);
        const unsigned long long* tile_off = nullptr;
        static const bool force_scan = getenv("GFFX_EMIT_SCAN") != nullptr;   // test hook for the large-table path
        if (n_emit_tiles > LINE_FOLD_TILES || force_scan) {   // a very large line table: offsets by a scan pass instead
            d_tile_off.ensure(n_emit_tiles * 8);
            d_scan_scratch.ensure(scan_scratch_elems(n_emit_tiles) * 8);
            launches += device_scan<uint64_t, OpSum, TileCountLoad, PtrStore<uint64_t>, false, false>(
                TileCountLoad{d_tile_count.as<uint32_t>()}, PtrStore<uint64_t>{d_tile_off.as<uint64_t>()}, n_emit_tiles, OpSum(), 0ull,
                d_scan_scratch.as<uint64_t>(), nullptr, st);
            tile_off = d_tile_off.as<unsigned long long>();
        }
        emit_ranges_kernel<<<(unsigned)n_emit_tiles, LINE_THREADS, 0, st>>>(LA, d_keep.as<uint8_t>(), d_tile_count.as<uint32_t>(), tile_off,
                                                                          d_ranges.as<ulonglong2>(), small + 4, small + 5);
        launches++;
    } else {
        if (jb.stage_events) GFFX_CUDA(cudaEventRecord(ev[6], st));
    }
    if (jb.stage_events) GFFX_CUDA(cudaEventRecord(ev[7], st));

    // ---- pairs: the run-length expansion of the per-root hit counters (K1b) ----------------------
    if (jb.stage_events) GFFX_CUDA(cudaEventRecord(ev[8], st));
    if (jb.pairs_dev && !nothing && jb.n_total) launches += expand_pairs(jb.pair_cap);
    GFFX_CUDA(cudaEventRecord(ev[9], st));

    // ---- results -----------------------------------------------------------------------------
    out->launches = launches;
    out->pend = PendingInfo{jb.pairs_dev, jb.join_only, jb.match_only, nothing, jb.no_host, jb.want_pairs_host, jb.pair_cap, d_pairs.as<uint32_t>(), lane_active};
    if (jb.async) {   // queue the read-back of the counters (and of the head of the warn list) into the job's slot; gffx_result_wait reads it
        AsyncSlot& S = slots[jb.slot];
        GFFX_CUDA(cudaMemcpyAsync(S.h, d_small.p, 64, cudaMemcpyDeviceToHost, st));
        GFFX_CUDA(cudaMemcpyAsync(S.h + ASYNC_WARN_OFF, d_warn.p, ASYNC_WARN_CAP * 4, cudaMemcpyDeviceToHost, st));
        if (!jb.no_host && !nothing && !jb.join_only) {
            // the host copy of the ranges rides behind the job: their number is not known on the host yet, so a kernel stores the
            // list (as long as the job's counter says) into the result's page-locked buffer, sized for the upper bound (every
            // line / every block); wait() trims it.  The next queued job's regions cross PCIe the other way meanwhile.
            const uint64_t cap = jb.match_only ? LA.n : n_blocks;
            if (cap) {
                out->ranges.resize(cap * 2);
                // a few dozen CTAs keep the link busy (16 to 296 measure the same); the fewer, the fewer thread slots the other lanes' kernels lose
                ranges_to_host_kernel<<<(unsigned)std::max(1, sm_count / 4), 256, 0, st>>>(d_ranges.as<ulonglong2>(), reinterpret_cast<ulonglong2*>(out->ranges.data()),
                                                                            small + (jb.match_only ? 5 : 1), cap);
                out->launches++;
                out->pend.ranges_cap = cap;
            }
        }
        GFFX_CUDA(cudaEventRecord(S.ev_end, st));
        {
            Lane& ln = lanes[lane_active];
            if (!ln.ev_done) GFFX_CUDA(cudaEventCreateWithFlags(&ln.ev_done, cudaEventDisableTiming));
            GFFX_CUDA(cudaEventRecord(ln.ev_done, st));
        }
        {
            cudaError_t le = cudaGetLastError();
            if (le != cudaSuccess) { S.state = 0; throw Err(E_CUDA, std::string("kernel launch failed: ") + cudaGetErrorString(le)); }
        }
        out->pending_slot = jb.slot;
        out->pend.host_in = jb.host_in;
        lane_activate(0);   // outside a job the engine's stream and buffers are lane 0's
        return;
    }
    GFFX_CUDA(cudaMemcpyAsync(h_small, d_small.p, 64, cudaMemcpyDeviceToHost, st));
    GFFX_CUDA(cudaStreamSynchronize(st));
    {
        cudaError_t le = cudaGetLastError();
        if (le != cudaSuccess) throw Err(E_CUDA, std::string("kernel launch failed: ") + cudaGetErrorString(le));
    }
    read_counters(out, (const uint8_t*)h_small, nullptr);
    if (out->pend.pairs_dev && out->n_pairs > out->pend.pair_cap) {
        // the pair buffer was too small for this result: grow it to the exact size and expand again (the counters are intact)
        d_pairs.ensure(out->n_pairs * 12);
        out->launches += expand_pairs(d_pairs.cap / 12);
        GFFX_CUDA(cudaStreamSynchronize(st));
        GFFX_CUDA(cudaGetLastError());
        out->d_pairs = out->n_pairs ? d_pairs.as<uint32_t>() : nullptr;
    }
    if (out->n_ranges && !jb.no_host) {
        out->ranges.resize(out->n_ranges * 2);
        GFFX_CUDA(cudaMemcpyAsync(out->ranges.data(), d_ranges.p, out->n_ranges * 16, cudaMemcpyDeviceToHost, st));
    }
    if (jb.want_pairs_host && out->n_pairs && !jb.no_host) {
        out->pairs.resize(out->n_pairs * 3);
        GFFX_CUDA(cudaMemcpyAsync(out->pairs.data(), d_pairs.p, out->n_pairs * 12, cudaMemcpyDeviceToHost, st));
    }
    GFFX_CUDA(cudaStreamSynchronize(st));
    // per-stage device time: the batches' partition / join / scatter intervals add up; the rest is measured once
    float ms;
    for (size_t k = 0; jb.stage_events && k < jb.chunks.size() && 4 * k + 3 < ev_chunk.size() && !jb.nothing; k++) {
        GFFX_CUDA(cudaEventElapsedTime(&ms, ev_chunk[4 * k], ev_chunk[4 * k + 1]));
        out->stage_ms[8] += ms;
        GFFX_CUDA(cudaEventElapsedTime(&ms, ev_chunk[4 * k + 1], ev_chunk[4 * k + 2]));
        out->stage_ms[0] += ms;
        GFFX_CUDA(cudaEventElapsedTime(&ms, ev_chunk[4 * k + 2], ev_chunk[4 * k + 3]));
        out->stage_ms[4] += ms;
    }
    if (jb.stage_events) {
    GFFX_CUDA(cudaEventElapsedTime(&ms, ev[8], ev[9]));
    out->stage_ms[7] = ms;  // offsets scan + pair expansion
    GFFX_CUDA(cudaEventElapsedTime(&ms, ev[3], ev[4]));
    out->stage_ms[3] = ms;
    GFFX_CUDA(cudaEventElapsedTime(&ms, ev[4], ev[6]));
    out->stage_ms[5] = ms;
    GFFX_CUDA(cudaEventElapsedTime(&ms, ev[6], ev[7]));
    out->stage_ms[6] = ms;
    }
    GFFX_CUDA(cudaEventElapsedTime(&ms, ev[0], ev[9]));
    out->kernel_ms = ms;
}

// the 64-byte counter block of a finished job -> the result's fields.  u64 view: [0] n_pairs [1] n_ranges (-e merge) [2] blocks selected
// [3] bytes (-e merge) [4] bytes_out [5] n_emit; u32 view: [12] region with a bad sequence number [13] n_warn.
// warn_head: the first ASYNC_WARN_CAP entries of the device warn list (asynchronous jobs), else it is read from the device here.
void EngineImpl::read_counters(IntersectResult* out, const uint8_t* h, const uint32_t* warn_head) {
    const uint64_t* hs = (const uint64_t*)h;
    const uint32_t* hs32 = (const uint32_t*)h;
    const PendingInfo& P = out->pend;
    if (hs32[12]) throw Err(E_ARG, "region with a chromosome number beyond the index's sequence list");
    out->n_pairs = hs[0];
    if (out->n_pairs > pairs_hwm) pairs_hwm = out->n_pairs;
    out->d_pairs = (P.pairs_dev && out->n_pairs && out->n_pairs <= P.pair_cap) ? P.d_pairs : nullptr;
    uint32_t n_warn = std::min<uint32_t>(hs32[13], WARN_CAP);
    if (n_warn) {
        std::vector<uint32_t> w(n_warn);
        if (warn_head) {
            n_warn = std::min<uint32_t>(n_warn, ASYNC_WARN_CAP);   // an asynchronous job keeps the head of the list only
            w.assign(warn_head, warn_head + n_warn);
        } else {
            GFFX_CUDA(cudaMemcpy(w.data(), d_warn.p, n_warn * 4ull, cudaMemcpyDeviceToHost));
        }
        std::sort(w.begin(), w.end());
        w.erase(std::unique(w.begin(), w.end()), w.end());
        out->warn_fids = std::move(w);
    }
    out->n_roots = hs[2] + out->warn_fids.size();
    if (!P.join_only && !P.nothing) {
        if (P.match_only) { out->n_ranges = hs[5]; out->bytes_out = hs[4]; }
        else { out->n_ranges = hs[1]; out->bytes_out = hs[3]; }
    }
    last_n_ranges = out->n_ranges;
    lane_last = P.lane;
}

// ---- asynchronous jobs: a ring of page-locked slots, each with its own pair of events -------------------------------------
int EngineImpl::slot_acquire() {
    for (int pass = 0; pass < 2; pass++)
        for (size_t k = 0; k < slots.size(); k++) {
            AsyncSlot& S = slots[k];
            if (S.state == 2 && cudaEventQuery(S.ev_end) == cudaSuccess) S.state = 0;   // an abandoned result whose copy has landed
            if (S.state == 0) {
                if (!S.h) {
                    GFFX_CUDA(cudaMallocHost((void**)&S.h, ASYNC_SLOT_BYTES));
                    GFFX_CUDA(cudaEventCreate(&S.ev_start));
                    GFFX_CUDA(cudaEventCreate(&S.ev_end));
                }
                S.state = 1;
                return (int)k;
            }
        }
    throw Err(E_STATE, "too many asynchronous results pending on this index (" + std::to_string(slots.size()) + "): wait for or free some");
}

void EngineImpl::ix_wait(IntersectResult* out) {
    if (out->pending_slot < 0) return;
    GFFX_CUDA(cudaSetDevice(dev));
    AsyncSlot& S = slots[(size_t)out->pending_slot];
    out->pending_slot = -1;
    cudaError_t e = cudaEventSynchronize(S.ev_end);
    if (e != cudaSuccess) { S.state = 0; throw Err(E_CUDA, std::string("asynchronous intersect failed: ") + cudaGetErrorString(e)); }
    float ms = 0;
    cudaEventElapsedTime(&ms, S.ev_start, S.ev_end);
    out->kernel_ms = ms;
    S.state = 0;   // the slot's bytes are read below, before any other call on this handle can take it (one thread per handle)
    const uint64_t cap = out->pend.ranges_cap;
    out->pend.ranges_cap = 0;
    try {
        read_counters(out, S.h, (const uint32_t*)(S.h + ASYNC_WARN_OFF));
    } catch (...) {
        out->ranges.release();
        throw;
    }
    if (cap) {   // the copy has landed (it preceded the slot's event): keep the ranges that exist
        if (out->n_ranges > cap) throw Err(E_STATE, "more ranges than the list's upper bound (internal)");
        if (out->n_ranges) out->ranges.resize(out->n_ranges * 2);   // shrinks in place
        else out->ranges.release();
    }
}

void EngineImpl::ix_abandon(IntersectResult* out) {
    if (out->pending_slot < 0) return;
    if (out->pend.ranges_cap || out->pend.host_in) {   // the result's page-locked buffer goes back to the pool with the result, the caller's region
        // arrays are free to go after this call: let the copies into / out of them finish first
        cudaSetDevice(dev);
        cudaEventSynchronize(slots[(size_t)out->pending_slot].ev_end);
        out->pend.ranges_cap = 0;
    }
    slots[(size_t)out->pending_slot].state = 2;
    out->pending_slot = -1;
}

void EngineImpl::intersect(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, int mode, bool invert,
                           bool entire_group, const char* types_csv, uint32_t flags, bool join_only, IntersectResult* out) {
    if (n && (!chr || !start || !end)) throw Err(E_ARG, "NULL region arrays");
    if (n >= (1ull << 31)) throw Err(E_ARG, "more than 2^31-1 regions in one call: split the batch");
    GFFX_CUDA(cudaSetDevice(dev));
    const bool dev_in = flags & 1u;  // GFFX_F_DEVICE_INPUT
    // GFFX_F_ASYNC with host regions: the arrays must stay valid (and should be page-locked) until the result has been waited for
    // host input: the copy is cut into chunks so that chunk i+1 crosses PCIe while chunk i is partitioned and joined
    uint32_t n_chunks = 1;
    if (!dev_in && n >= (1u << 21)) n_chunks = (uint32_t)std::min<uint64_t>(8, n >> 20);
    if (!dev_in && (flags & 32u)) {   // queued: the NEXT job's copy hides this job's kernels; fewer, larger copies cost less per byte
        static const int async_chunks = getenv("GFFX_ASYNC_CHUNKS") ? std::max(1, atoi(getenv("GFFX_ASYNC_CHUNKS"))) : 2;
        n_chunks = std::min<uint32_t>(n_chunks, (uint32_t)async_chunks);
    }
    uint64_t chunk = (n + n_chunks - 1) / std::max<uint32_t>(n_chunks, 1);
    chunk = (chunk + PART_TILE - 1) / PART_TILE * PART_TILE;
    if (chunk == 0) chunk = PART_TILE;
    n_chunks = (uint32_t)std::max<uint64_t>(1, (n + chunk - 1) / chunk);
    ix_begin(mode, invert, entire_group, types_csv, flags, join_only, std::max<uint64_t>(n, 1), n_chunks);
    if (dev_in) {
        ix_add(chr, start, end, n);
    } else if (n) {
        const uint32_t* src[3] = {chr, start, end};
        for (int k = 0; k < 3; k++) d_q[k].ensure(n * 4);
        // the lane's previous queued job may still be reading its staged regions
        if (lanes[lane_active].ev_done) GFFX_CUDA(cudaStreamWaitEvent(st_copy, lanes[lane_active].ev_done, 0));
        while (ev_copy.size() < n_chunks) {
            cudaEvent_t e;
            GFFX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            ev_copy.push_back(e);
        }
        for (uint32_t c = 0; c < n_chunks; c++) {
            const uint64_t o = (uint64_t)c * chunk, m = std::min<uint64_t>(chunk, n - o);
            for (int k = 0; k < 3; k++)
                GFFX_CUDA(cudaMemcpyAsync(d_q[k].as<uint32_t>() + o, src[k] + o, m * 4, cudaMemcpyHostToDevice, st_copy));
            GFFX_CUDA(cudaEventRecord(ev_copy[c], st_copy));
        }
        for (uint32_t c = 0; c < n_chunks; c++) {
            const uint64_t o = (uint64_t)c * chunk, m = std::min<uint64_t>(chunk, n - o);
            GFFX_CUDA(cudaStreamWaitEvent(st, ev_copy[c], 0));
            ix_add(d_q[0].as<uint32_t>() + o, d_q[1].as<uint32_t>() + o, d_q[2].as<uint32_t>() + o, m);
        }
    }
    ix_finish(out);
    if (!dev_in && out->pending_slot >= 0) out->pend.host_in = true;
}

// ------------------------------------------------------------------------------------------------
// synthetic regions straight into HBM (same function of (seed, index) as the host generator)
__global__ void synth_regions_kernel(uint64_t seed, uint64_t first, uint64_t n, uint32_t length, const uint64_t* __restrict__ cum,
                                     const uint32_t* __restrict__ extent, uint32_t n_seq, uint32_t* __restrict__ chr,
                                     uint32_t* __restrict__ start, uint32_t* __restrict__ end) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t c, s, e;
    synth_region(seed, first + i, length, cum, extent, n_seq, &c, &s, &e);
    chr[i] = c; start[i] = s; end[i] = e;
}

void Engine::synth_regions_device(uint64_t seed, uint64_t first, uint64_t n, uint32_t length, uint32_t* d_chr, uint32_t* d_start,
                                  uint32_t* d_end) {
    EngineImpl* E = impl_.get();
    GFFX_CUDA(cudaSetDevice(E->dev));
    std::vector<uint32_t> ext;
    std::vector<uint64_t> cum;
    synth_seq_table(*E->H, &ext, &cum);
    if (ext.empty() || cum.back() == 0) throw Err(E_ARG, "index has no root intervals to place regions on");
    DBuf dcum, dext;
    upload(dcum, cum, E->st);
    upload(dext, ext, E->st);
    if (n) synth_regions_kernel<<<(unsigned)((n + 255) / 256), 256, 0, E->st>>>(seed, first, n, length, dcum.as<uint64_t>(), dext.as<uint32_t>(),
                                                                             (uint32_t)ext.size(), d_chr, d_start, d_end);
    GFFX_CUDA(cudaStreamSynchronize(E->st));
    GFFX_CUDA(cudaGetLastError());
    dcum.release();
    dext.release();
}

// ------------------------------------------------------------------------------------------------
// Engine facade
// ------------------------------------------------------------------------------------------------
Engine::Engine(std::unique_ptr<HostIndex> host, int device) : impl_(new EngineImpl(std::move(host), device)) {}
Engine::~Engine() = default;
const HostIndex& Engine::host() const { return *impl_->H; }
int Engine::device() const { return impl_->dev; }
uint64_t Engine::device_bytes() const { return impl_->dev_bytes; }
void* Engine::cuda_stream() const { return (void*)impl_->st; }
const std::vector<uint32_t>& Engine::sorted_roots(uint32_t seq) const {
    if (seq >= impl_->seq_sorted_roots.size()) throw Err(E_ARG, "sequence number out of range");
    return impl_->seq_sorted_roots[seq];
}
int cuda_device_count() {
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}
// ---- shards ---------------------------------------------------------------------------------------
void Engine::shard_begin(int mode, bool invert, bool entire_group, const char* types_csv, uint32_t flags, uint64_t n_expected,
                         uint32_t batches_expected) {
    if ((flags & (2u | 8u)) && n_expected == 0)
        throw Err(E_ARG, "a batched intersect job materialises its pairs only when the number of regions is announced at begin (n_expected)");
    if (n_expected >= (1ull << 31)) throw Err(E_ARG, "n_expected must stay below 2^31 (announce 0 for a streaming job)");
    impl_->ix_begin(mode, invert, entire_group, types_csv, flags, false, n_expected, std::max<uint32_t>(batches_expected, 1));
}
void Engine::shard_add(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags, uint64_t* n_pairs_total) {
    EngineImpl* E = impl_.get();
    if (n && (!chr || !start || !end)) throw Err(E_ARG, "NULL region arrays");
    if (!E->job.active) throw Err(E_STATE, "gffx_shard_add without gffx_shard_begin");
    // a queued job copies host regions behind the call: no count can be handed back, the arrays stay the caller's until the wait
    if (E->job.async && !(flags & 1u) && n_pairs_total) throw Err(E_ARG, "a GFFX_F_ASYNC job cannot report pair counts per batch (n_pairs_total must be NULL)");
    if (E->job.async && !(flags & 1u)) E->job.host_in = true;
    GFFX_CUDA(cudaSetDevice(E->dev));
    const uint32_t *dc, *ds, *de;
    E->stage_queries(chr, start, end, n, flags, &dc, &ds, &de);
    E->ix_add(dc, ds, de, n);
    if (((flags & 1u) || E->job.async) && !n_pairs_total) return;  // device input (or a queued job), no count wanted: stays asynchronous (gffx.h: arrays live until finish / sync / wait)
    // the caller's arrays are free again when this returns
    GFFX_CUDA(cudaMemcpyAsync(E->h_small, E->d_small.p, 64, cudaMemcpyDeviceToHost, E->st));
    GFFX_CUDA(cudaStreamSynchronize(E->st));
    GFFX_CUDA(cudaGetLastError());
    if (((const uint32_t*)E->h_small)[12]) throw Err(E_ARG, "region with a chromosome number beyond the index's sequence list");
    if (n_pairs_total) *n_pairs_total = *(const uint64_t*)E->h_small;
}
void Engine::shard_finish(IntersectResult* out) { impl_->ix_finish(out); }
void Engine::wait(IntersectResult* out) { impl_->ix_wait(out); }
void Engine::abandon(IntersectResult* out) { impl_->ix_abandon(out); }
// ---- the allocation-free entry point (gffx_count_pairs_noalloc) ---------------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA
// graph capture, tight serving loops).  The join-pair count of ONE device-resident batch — K0 partition + K1 join, the same
// kernels as gffx_query_features — with every intermediate (the bucketed regions, the window totals / cursors / offsets, the
// per-root hit counters and flags) in this fixed static device array instead of the engine's grow-on-demand buffers.  Every
// GPU that loads the library holds the array (10.6 MB) until exit.  All calls on a GPU share it and it keeps the last call's
// contents: what must start at zero is cleared on the stream at the head of each call, and two calls that overlap on the GPU
// (different streams, or two indexes on one device) race — serialise them.
constexpr uint64_t NOALLOC_MAX_REGIONS = 1ull << 20;   // regions per call
constexpr uint32_t NOALLOC_MAX_ROOTS = 1u << 18;       // root intervals of the index
constexpr size_t NA_SE = 0;                                                                   // uint2[MAX_REGIONS + 4]
constexpr size_t NA_CNT = NA_SE + (NOALLOC_MAX_REGIONS + 4) * 8;                              // u64[MAX_ROOTS + 2]
constexpr size_t NA_HIT = NA_CNT + ((size_t)NOALLOC_MAX_ROOTS + 2) * 8;                       // u8[MAX_ROOTS + 16]
constexpr size_t NA_TOT = NA_HIT + (size_t)NOALLOC_MAX_ROOTS + 16;                            // u32[PART_NB_MAX]
constexpr size_t NA_CUR = NA_TOT + (size_t)PART_NB_MAX * 4;                                   // u32[PART_NB_MAX]
constexpr size_t NA_POFF = NA_CUR + (size_t)PART_NB_MAX * 4;                                  // u32[PART_NB_MAX + 4]
constexpr size_t NOALLOC_BYTES = (NA_POFF + ((size_t)PART_NB_MAX + 4) * 4 + 255) & ~(size_t)255;
static_assert(NA_CNT % 16 == 0 && NA_HIT % 16 == 0 && NA_TOT % 16 == 0, "aligned parts");
static __device__ __align__(256) unsigned char gffx_noalloc_scratch[NOALLOC_BYTES];

// In host code the array's name is a stand-in, not a device address: the address is asked for once per device, with that
// device current (the engine's constructor), and only that pointer is handed to kernels.
static unsigned char* noalloc_scratch_address() {
    unsigned char* ptr = nullptr;
    GFFX_CUDA(cudaGetSymbolAddress((void**)&ptr, gffx_noalloc_scratch));
    if (!ptr) throw Err(E_CUDA, "cudaGetSymbolAddress returned no address for the allocation-free scratch array");
    return ptr;
}

template <int MODE, bool INV>
static void launch_wj_noalloc(EngineImpl* E, unsigned char* S, uint64_t n, unsigned long long* n_pairs, uint32_t* err, cudaStream_t s) {
    typedef unsigned long long ull;
    const uint32_t* poff = reinterpret_cast<const uint32_t*>(S + NA_POFF);
    const uint2* se = reinterpret_cast<const uint2*>(S + NA_SE);
    uint8_t* hit = S + NA_HIT;
    ull* cnt = reinterpret_cast<ull*>(S + NA_CNT);
    const unsigned tiles = (unsigned)((n + WJ_TILE - 1) / WJ_TILE);
    const FilterArrays F{};   // FILT = false: never read
    const bool small = n < 64ull * std::max<uint32_t>(E->WC.n_win, 1);   // the choice of launch_wj
    if (E->wj_force == 1 || (E->wj_force == 0 && small)) {
        if (small) wj_count_kernel<MODE, INV, false, true><<<dim3(tiles, 16u), WJ_THREADS, WJ_DYN_SMEM, s>>>(E->WC, poff, se, n, hit, cnt, n_pairs, err, E->FR, F);
        else wj_count_kernel<MODE, INV, false, false><<<dim3(tiles, 1u), WJ_THREADS, WJ_DYN_SMEM, s>>>(E->WC, poff, se, n, hit, cnt, n_pairs, err, E->FR, F);
    } else {
        wj_root_kernel<MODE, INV, false, false><<<tiles, WJ_THREADS, WJ_DYN_SMEM, s>>>(E->WC, poff, se, n, hit, cnt, n_pairs, err, E->FR, F, E->wj_tail_bins, E->XS);
    }
}

uint64_t Engine::noalloc_max_regions() { return NOALLOC_MAX_REGIONS; }
uint64_t Engine::noalloc_max_roots() { return NOALLOC_MAX_ROOTS; }

void Engine::count_pairs_noalloc(const uint32_t* d_chr, const uint32_t* d_start, const uint32_t* d_end, uint64_t n, int mode, bool invert,
                                 uint64_t* d_out, void* stream) {
    EngineImpl* E = impl_.get();
    if (mode < 0 || mode > 2) throw Err(E_ARG, "mode must be 0 (overlap), 1 (contained) or 2 (contains-region)");
    if (n > NOALLOC_MAX_REGIONS)
        throw Err(E_ARG, "gffx_count_pairs_noalloc: " + std::to_string(n) + " regions exceed the fixed scratch array's " + std::to_string(NOALLOC_MAX_REGIONS) +
                             " per call: split the batch or use gffx_query_features");
    if (E->wj_n_ent > NOALLOC_MAX_ROOTS)
        throw Err(E_ARG, "gffx_count_pairs_noalloc: the index has " + std::to_string(E->wj_n_ent) + " root intervals, the fixed scratch array holds counters for " +
                             std::to_string(NOALLOC_MAX_ROOTS) + ": use gffx_query_features");
    if (!E->noalloc_scratch) throw Err(E_STATE, "the allocation-free scratch array has no device address on this GPU");
    GFFX_CUDA(cudaSetDevice(E->dev));
    cudaStream_t s = stream ? (cudaStream_t)stream : E->st;
    unsigned char* S = E->noalloc_scratch;
    GFFX_CUDA(cudaMemsetAsync(d_out, 0, 16, s));
    if (n == 0 || (mode == 2 && invert) || E->wj_n_ent == 0) return;   // -I with overlap selects nothing (intersect.rs:161)
    // the array keeps the previous call's contents: the window totals and the per-root counters / flags must start at zero
    GFFX_CUDA(cudaMemsetAsync(S + NA_TOT, 0, (size_t)PART_NB_MAX * 4, s));
    GFFX_CUDA(cudaMemsetAsync(S + NA_CNT, 0, ((size_t)E->wj_n_ent + 2) * 8, s));
    GFFX_CUDA(cudaMemsetAsync(S + NA_HIT, 0, (size_t)E->wj_n_ent + 16, s));
    uint32_t* tot = reinterpret_cast<uint32_t*>(S + NA_TOT);
    uint32_t* cur = reinterpret_cast<uint32_t*>(S + NA_CUR);
    uint32_t* poff = reinterpret_cast<uint32_t*>(S + NA_POFF);
    const unsigned tiles = (unsigned)((n + PART_TILE - 1) / PART_TILE), stiles = (unsigned)((n + PART_STILE - 1) / PART_STILE);
    part_count_kernel<<<std::min(tiles, (unsigned)(E->sm_count * PART_COUNT_CTAS)), PART_THREADS, PART_COUNT_SMEM, s>>>(E->WG, d_chr, d_start, (uint32_t)n, tot);
    part_offsets_kernel<<<1, PART_NB_MAX, 0, s>>>(tot, E->WG.n_win + 1, (uint32_t)n, poff, cur);
    part_scatter_kernel<<<std::min(stiles, (unsigned)(E->sm_count * PART_SCATTER_CTAS)), PART_THREADS, PART_SCATTER_SMEM, s>>>(
        E->WG, d_chr, d_start, d_end, (uint32_t)n, cur, reinterpret_cast<uint2*>(S + NA_SE));
    unsigned long long* n_pairs = reinterpret_cast<unsigned long long*>(d_out);
    uint32_t* err = reinterpret_cast<uint32_t*>(d_out + 1);   // the kernels store 1 into the low word
    switch (mode * 2 + (invert ? 1 : 0)) {
        case 0: launch_wj_noalloc<0, false>(E, S, n, n_pairs, err, s); break;
        case 1: launch_wj_noalloc<0, true>(E, S, n, n_pairs, err, s); break;
        case 2: launch_wj_noalloc<1, false>(E, S, n, n_pairs, err, s); break;
        case 3: launch_wj_noalloc<1, true>(E, S, n, n_pairs, err, s); break;
        default: launch_wj_noalloc<2, false>(E, S, n, n_pairs, err, s); break;
    }
    GFFX_CUDA(cudaGetLastError());
}

void Engine::sync() {
    GFFX_CUDA(cudaSetDevice(impl_->dev));
    for (const EngineImpl::Lane& l : impl_->lanes)
        if (l.st) GFFX_CUDA(cudaStreamSynchronize(l.st));
}
bool Engine::shard_buffer(int which, uint32_t k, void** dptr, uint64_t* n_elems, int* op) {
    EngineImpl* E = impl_.get();
    if (which == 1) return counts_buffer_impl(E, k, dptr, n_elems, op);
    E->ix_seal();
    const IxJob& jb = E->job;
    struct B { void* p; uint64_t n; int op; };
    std::vector<B> v;
    if (!jb.nothing) {
        // byte flags, MAX: which roots were hit (-> selected blocks, warn list) and, behind them in the same allocation, which
        // lines match some region (filter.cuh); the padding between the two stays 0
        v.push_back({E->d_wj_hit.p, jb.run_filter ? E->flag_bytes() : (uint64_t)E->wj_n_ent, RED_MAX_U8});
    }
    if (k >= v.size()) return false;
    *dptr = v[k].p; *n_elems = v[k].n; *op = v[k].op;
    return true;
}

template <class T, class F>
__global__ void fold_kernel(T* __restrict__ dst, const T* __restrict__ src, uint32_t rows, uint64_t stride, uint64_t n, F f) {
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) {
        T a = src[i];
        for (uint32_t r = 1; r < rows; r++) a = f(a, src[(uint64_t)r * stride + i]);
        dst[i] = a;
    }
}
struct FMax { template <class T> __device__ T operator()(T a, T b) const { return a > b ? a : b; } };
struct FMin { template <class T> __device__ T operator()(T a, T b) const { return a < b ? a : b; } };
struct FSum { template <class T> __device__ T operator()(T a, T b) const { return a + b; } };
struct FOr { template <class T> __device__ T operator()(T a, T b) const { return a | b; } };

void Engine::fold(void* dst, const void* src, uint32_t rows, uint64_t stride, uint64_t n, int op) {
    EngineImpl* E = impl_.get();
    if (n == 0 || rows == 0) return;
    GFFX_CUDA(cudaSetDevice(E->dev));
    const unsigned blocks = (unsigned)std::min<uint64_t>((n + 255) / 256, (uint64_t)E->sm_count * 16);
    typedef unsigned long long ull;
    switch (op) {
        case RED_MAX_U8: fold_kernel<<<blocks, 256, 0, E->st>>>((uint8_t*)dst, (const uint8_t*)src, rows, stride, n, FMax()); break;
        case RED_MAX_U64: fold_kernel<<<blocks, 256, 0, E->st>>>((ull*)dst, (const ull*)src, rows, stride, n, FMax()); break;
        case RED_MIN_U64: fold_kernel<<<blocks, 256, 0, E->st>>>((ull*)dst, (const ull*)src, rows, stride, n, FMin()); break;
        case RED_SUM_U32: fold_kernel<<<blocks, 256, 0, E->st>>>((uint32_t*)dst, (const uint32_t*)src, rows, stride, n, FSum()); break;
        case RED_SUM_I32: fold_kernel<<<blocks, 256, 0, E->st>>>((int*)dst, (const int*)src, rows, stride, n, FSum()); break;
        case RED_OR_U64: fold_kernel<<<blocks, 256, 0, E->st>>>((ull*)dst, (const ull*)src, rows, stride, n, FOr()); break;
        default: throw Err(E_ARG, "unknown reduce op");
    }
    GFFX_CUDA(cudaGetLastError());
}

// ------------------------------------------------------------------------------------------------
// K8: record emission on the device (gather.cuh)
void EngineImpl::gather_text(const uint64_t* ranges, uint64_t n_ranges, uint32_t flags, const uint8_t** text, uint64_t* n_bytes) {
    typedef unsigned long long ull;
    GFFX_CUDA(cudaSetDevice(dev));
    *text = nullptr; *n_bytes = 0;
    gather_ms = 0; gather_launches = 0;
    if (!H->gff) throw Err(E_STATE, "index was opened without the GFF text");
    if (!text_resident) {   // once per index: the GFF text goes to HBM (+ 8 bytes of padding for the copy's look-ahead word)
        d_text.ensure(H->gff_len + 16);
        GFFX_CUDA(cudaMemsetAsync((uint8_t*)d_text.p + H->gff_len, 0, 16, st));
        if (H->gff_len) GFFX_CUDA(cudaMemcpyAsync(d_text.p, H->gff, H->gff_len, cudaMemcpyHostToDevice, st));
        text_resident = true;
    }
    const ulonglong2* dr;
    if (ranges) {
        d_text_ranges.ensure(std::max<uint64_t>(n_ranges, 1) * 16);
        if (n_ranges) GFFX_CUDA(cudaMemcpyAsync(d_text_ranges.p, ranges, n_ranges * 16, cudaMemcpyHostToDevice, st));
        dr = d_text_ranges.as<ulonglong2>();
    } else {
        if (n_ranges != last_n_ranges) throw Err(E_STATE, "the ranges of that result are no longer on the device: pass them explicitly");
        dr = lane_buf(lane_last, &d_ranges).as<ulonglong2>();
    }
    if (n_ranges == 0) return;
    if (!ev_gather[0])
        for (auto& e : ev_gather) GFFX_CUDA(cudaEventCreate(&e));
    GFFX_CUDA(cudaEventRecord(ev_gather[0], st));
    d_text_off.ensure((n_ranges + 1) * 8);
    d_scan_scratch.ensure(scan_scratch_elems(n_ranges + 1) * 8);
    gather_launches += device_scan<ull, OpSum, RangeLenLoad, PtrStore<ull>, false, false>(
        RangeLenLoad{dr, n_ranges, H->gff_len}, PtrStore<ull>{d_text_off.as<ull>()}, n_ranges + 1, OpSum(), 0ull, d_scan_scratch.as<ull>(),
        (ull*)((uint8_t*)d_small.p + 128), st);
    GFFX_CUDA(cudaEventRecord(ev_gather[1], st));
    GFFX_CUDA(cudaMemcpyAsync((uint8_t*)h_small + 128, (uint8_t*)d_small.p + 128, 8, cudaMemcpyDeviceToHost, st));
    GFFX_CUDA(cudaStreamSynchronize(st));
    const uint64_t total = *(const uint64_t*)((const uint8_t*)h_small + 128);
    *n_bytes = total;
    if (total == 0) return;
    d_text_out.ensure(total + 16);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    GFFX_CUDA(cudaEventRecord(ev_gather[2], st));
    gather_text_kernel<<<(unsigned)sms * 8, GATHER_THREADS, 0, st>>>(d_text.as<uint8_t>(), dr, d_text_off.as<ull>(), n_ranges, d_text_out.as<uint8_t>());
    gather_launches++;
    GFFX_CUDA(cudaEventRecord(ev_gather[3], st));
    if (flags & 4u) {
        *text = d_text_out.as<uint8_t>();
    } else {
        h_text_out.resize(total);
        GFFX_CUDA(cudaMemcpyAsync(h_text_out.data(), d_text_out.p, total, cudaMemcpyDeviceToHost, st));
        *text = h_text_out.data();
    }
    GFFX_CUDA(cudaStreamSynchronize(st));
    GFFX_CUDA(cudaGetLastError());
    float ms0, ms1;   // the offsets scan and the gather kernel, without the host round trip for the output size between them
    GFFX_CUDA(cudaEventElapsedTime(&ms0, ev_gather[0], ev_gather[1]));
    GFFX_CUDA(cudaEventElapsedTime(&ms1, ev_gather[2], ev_gather[3]));
    gather_scan_ms = ms0;
    gather_ms = ms1;
}
void Engine::gather_text(const uint64_t* ranges, uint64_t n_ranges, uint32_t flags, const uint8_t** text, uint64_t* n_bytes) {
    impl_->gather_text(ranges, n_ranges, flags, text, n_bytes);
}
double Engine::gather_kernel_ms() const { return impl_->gather_ms; }
double Engine::gather_scan_ms() const { return impl_->gather_scan_ms; }
uint32_t Engine::gather_launches() const { return impl_->gather_launches; }

void bed_parse_device_impl(EngineImpl* E, const std::string& path, int dialect, const uint32_t** dc, const uint32_t** ds, const uint32_t** de,
                           uint64_t* n_out);   // bedparse.cu
void Engine::parse_bed_device(const std::string& path, int dialect, const uint32_t** d_chr, const uint32_t** d_start, const uint32_t** d_end,
                              uint64_t* n) {
    bed_parse_device_impl(impl_.get(), path, dialect, d_chr, d_start, d_end, n);
}
double Engine::bed_parse_stat(int what) const {
    switch (what) {
        case 0: return impl_->bed_ms;
        case 1: return (double)impl_->bed_launches;
        case 2: return (double)impl_->bed_host_lines;
        case 3: return (double)impl_->bed_bytes;
        default: return 0;
    }
}

void Engine::counts_begin(int kind) { counts_begin_impl(impl_.get(), kind); }
void Engine::counts_add(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags) {
    counts_add_impl(impl_.get(), chr, start, end, n, flags);
}
void Engine::counts_finish(CountRows* rows) { counts_finish_impl(impl_.get(), rows); }
double Engine::counts_kernel_ms() const { return counts_ms_impl(impl_.get()); }
uint32_t Engine::counts_launches() const { return counts_launches_impl(impl_.get()); }
void Engine::intersect(const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, int mode, bool invert,
                       bool entire_group, const char* types_csv, uint32_t flags, bool join_only, IntersectResult* out) {
    impl_->intersect(chr, start, end, n, mode, invert, entire_group, types_csv, flags, join_only, out);
}

}  // namespace gffx
