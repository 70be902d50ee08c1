// counts.cu — depth / coverage pipelines (K6 / K7): index upload, per-batch scatter, finish.
#include <algorithm>
#include "counts.cuh"
#include "counts_index.h"
#include "engine_impl.cuh"

namespace gffx {

struct CountsState {
    // ---- index (built lazily by the first counts_begin, kept for the engine's lifetime) ----
    bool built = false;
    CountsIndexHost host;  // scalars are read after the upload; the big vectors are released
    DBuf d_rank, d_cov, d_X, d_dir;
    DBuf d_u_space, d_u_chrom, d_u_blk, d_u_base, d_irr_base, d_u_line0, d_u_seg0, d_l_fs, d_l_fe, d_l_sfs, d_l_sfe, d_g_s, d_g_e, d_g_ss, d_g_se;
    DBuf d_id_unit0, d_id_units, d_irr_grp, d_irr_pos, d_grp_off, d_grp_seq, d_grp_s, d_grp_e, d_blk_space;
    DBuf d_irr_seq0, d_irr_s, d_irr_e, d_irr_pm, d_irr_space;
    DBuf d_cxw0, d_cxlen, d_cxoff;   // depth_window_kernel: the windows' parts of the counts coordinate set (RankSlices)
    RankSlices CXS{};
    DBuf d_gsp0, d_gap_bits, d_gs_first, d_gap_a, d_gap_b, d_gap_seg, d_irr_blk, d_u_rs, d_u_re;
    // ---- accumulators of the running job ----
    DBuf d_Hs, d_He, d_G, d_S, d_E, d_bits, d_diff, d_P, d_routed, d_scratch, d_err;
    DBuf r_chrom, r_start, r_end, r_value, r_flag, r_pos, o_id, o_chrom, o_start, o_end, o_value;
    int kind = -1;
    bool active = false;
    bool exchanged = false;   // a sharded coverage job handed its accumulators out (counts_buffer_impl)
    uint64_t n_regions = 0;
    double ms = 0;
    uint32_t launches = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;

    std::vector<DBuf*> all() {
        return {&d_rank, &d_cov, &d_X, &d_dir, &d_u_space, &d_u_chrom, &d_u_blk, &d_u_base, &d_irr_base, &d_u_line0, &d_u_seg0, &d_l_fs, &d_l_fe, &d_l_sfs,
                &d_l_sfe, &d_g_s, &d_g_e, &d_g_ss, &d_g_se, &d_irr_seq0, &d_irr_s, &d_irr_e, &d_irr_pm, &d_irr_space, &d_cxw0, &d_cxlen, &d_cxoff, &d_gsp0, &d_gap_bits, &d_gs_first, &d_gap_a, &d_gap_b, &d_gap_seg, &d_irr_blk, &d_u_rs, &d_u_re, &d_id_unit0, &d_id_units, &d_irr_grp, &d_irr_pos, &d_grp_off, &d_grp_seq,
                &d_grp_s, &d_grp_e, &d_blk_space, &d_Hs, &d_He, &d_G, &d_S, &d_E, &d_bits, &d_diff, &d_P, &d_routed, &d_scratch, &d_err,
                &r_chrom, &r_start, &r_end, &r_value, &r_flag, &r_pos, &o_id, &o_chrom, &o_start, &o_end, &o_value};
    }
};

void counts_release(EngineImpl* E) {
    CountsState* C = E->counts;
    if (!C) return;
    for (DBuf* b : C->all()) b->release();
    if (C->ev0) cudaEventDestroy(C->ev0);
    if (C->ev1) cudaEventDestroy(C->ev1);
    delete C;
    E->counts = nullptr;
}

static void build_counts(EngineImpl* E, CountsState* C) {
    for (DBuf* b : C->all()) b->total = &E->dev_bytes;
    CountsIndexHost& h = C->host;
    build_counts_index(*E->H, &h);
    cudaStream_t st = E->st;
    upload(C->d_rank, h.rank, st);
    upload(C->d_cov, h.cov, st);
    {   // per window of the partition grid: its part of the sequence space's coordinate set + fine-bin offsets (rootjoin.cuh: RankSlices)
        const uint32_t n_win = E->WG.n_win, shift = E->WG.shift, fsh = shift - DW_FB;
        const size_t n_v = E->H->vmap.n_vseqs();
        std::vector<uint32_t> xw0(std::max<uint32_t>(n_win, 1), 0), xlen(std::max<uint32_t>(n_win, 1), 0xFFFFFFFFu);
        std::vector<uint16_t> xoff((size_t)std::max<uint32_t>(n_win, 1) * DW_XOFF_STRIDE, 0);
        for (size_t v = 0; v < n_v; v++) {
            const uint32_t* xv = h.X.data() + h.rank[v].x0;
            const size_t cnt = h.rank[v].cnt;
            auto below = [&](uint64_t c) { return (size_t)(std::lower_bound(xv, xv + cnt, c, [](uint32_t a, uint64_t b) { return (uint64_t)a < b; }) - xv); };
            for (uint32_t w = E->h_win0[v]; w < E->h_win0[v + 1]; w++) {
                const uint64_t lo = (uint64_t)(w - E->h_win0[v]) << shift;
                const bool last = w + 1 == E->h_win0[v + 1];
                const size_t i0 = below(lo), i1 = last ? cnt : below(lo + (1ull << shift));
                xw0[w] = h.rank[v].x0 + (uint32_t)i0;
                if (i1 - i0 > WR_SLICE_CAP - 4 || i1 - i0 > 0xFFFFu) continue;   // depth_window_kernel falls back to the global look-ups there
                xlen[w] = (uint32_t)(i1 - i0);
                uint16_t* row = xoff.data() + (size_t)w * DW_XOFF_STRIDE;
                for (uint32_t b = 0; b < DW_NBF; b++) row[b] = (uint16_t)(std::min(below(lo + ((uint64_t)b << fsh)), i1) - i0);
                row[DW_NBF] = (uint16_t)(i1 - i0);
            }
        }
        upload(C->d_cxw0, xw0, st);
        upload(C->d_cxlen, xlen, st);
        upload(C->d_cxoff, xoff, st);
        GFFX_CUDA(cudaStreamSynchronize(st));
        C->CXS = RankSlices{C->d_cxw0.as<uint32_t>(), C->d_cxlen.as<uint32_t>(), C->d_cxoff.as<uint16_t>()};
    }
    h.X.insert(h.X.end(), 8, 0u);   // rank_lb (filter.cuh) reads up to 4 entries past a bin
    upload(C->d_X, h.X, st);
    upload(C->d_dir, h.dir, st);
    upload(C->d_u_space, h.u_space, st);
    upload(C->d_u_chrom, h.u_chrom, st);
    upload(C->d_u_blk, h.u_blk, st);
    upload(C->d_u_base, h.u_base, st);
    upload(C->d_irr_base, h.irr_base, st);
    upload(C->d_u_line0, h.u_line0, st);
    upload(C->d_u_seg0, h.u_seg0, st);
    upload(C->d_l_fs, h.l_fs, st);
    upload(C->d_l_fe, h.l_fe, st);
    upload(C->d_l_sfs, h.l_sfs, st);
    upload(C->d_l_sfe, h.l_sfe, st);
    upload(C->d_g_s, h.g_s, st);
    upload(C->d_g_e, h.g_e, st);
    upload(C->d_g_ss, h.g_ss, st);
    upload(C->d_g_se, h.g_se, st);
    upload(C->d_id_unit0, h.id_unit0, st);
    upload(C->d_id_units, h.id_units, st);
    upload(C->d_irr_grp, h.irr_grp, st);
    upload(C->d_irr_pos, h.irr_pos, st);
    upload(C->d_grp_off, h.grp_off, st);
    upload(C->d_grp_seq, h.grp_seq, st);
    upload(C->d_grp_s, h.grp_s, st);
    upload(C->d_grp_e, h.grp_e, st);
    upload(C->d_blk_space, h.blk_space, st);
    upload(C->d_gsp0, h.gsp0, st);
    upload(C->d_gap_bits, h.gap_bits, st);
    upload(C->d_gs_first, h.gs_first, st);
    upload(C->d_gap_a, h.gap_a, st);
    upload(C->d_gap_b, h.gap_b, st);
    upload(C->d_gap_seg, h.gap_seg, st);
    upload(C->d_irr_blk, h.irr_blk, st);
    upload(C->d_u_rs, h.u_rs, st);
    upload(C->d_u_re, h.u_re, st);
    upload(C->d_irr_seq0, h.irr_seq0, st);
    upload(C->d_irr_s, h.irr_s, st);
    upload(C->d_irr_e, h.irr_e, st);
    upload(C->d_irr_pm, h.irr_pm, st);
    upload(C->d_irr_space, h.irr_space, st);
    C->d_err.ensure(256);
    GFFX_CUDA(cudaFuncSetAttribute(depth_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DW_SMEM));
    GFFX_CUDA(cudaStreamSynchronize(st));
    GFFX_CUDA(cudaEventCreate(&C->ev0));
    GFFX_CUDA(cudaEventCreate(&C->ev1));
    // keep the scalars, drop the bulk
    for (auto* v : {&h.X, &h.dir, &h.u_space, &h.u_chrom, &h.u_blk, &h.u_base, &h.irr_base, &h.u_line0, &h.u_seg0, &h.l_fs, &h.l_fe, &h.l_sfs, &h.l_sfe, &h.g_s,
                    &h.g_e, &h.g_ss, &h.g_se, &h.id_unit0, &h.id_units, &h.irr_grp, &h.irr_pos, &h.blk_space, &h.irr_seq0, &h.irr_s, &h.irr_e, &h.irr_pm, &h.irr_space, &h.gsp0, &h.gap_bits, &h.gs_first, &h.gap_a, &h.gap_b, &h.gap_seg, &h.irr_blk, &h.u_rs, &h.u_re})
        std::vector<uint32_t>().swap(*v);
    C->built = true;
}

void counts_begin_impl(EngineImpl* E, int kind) {
    GFFX_CUDA(cudaSetDevice(E->dev));
    if (!E->H->gff) throw Err(E_STATE, "depth/coverage need the GFF text (the index was opened without it)");
    if (!E->counts) E->counts = new CountsState();
    CountsState* C = E->counts;
    if (!C->built) build_counts(E, C);
    const CountsIndexHost& h = C->host;
    cudaStream_t st = E->st;
    C->kind = kind;
    C->ms = 0;
    C->launches = 0;
    C->n_regions = 0;
    GFFX_CUDA(cudaMemsetAsync(C->d_err.p, 0, 256, st));
    if (kind == 0) {
        const size_t ns = std::max<size_t>(h.n_slots, 1), ng = std::max<size_t>(h.n_segs, 1);
        C->d_Hs.ensure(ns * 4); C->d_He.ensure(ns * 4); C->d_S.ensure(ns * 4); C->d_E.ensure(ns * 4);
        C->d_G.ensure(ng * 4);
        GFFX_CUDA(cudaMemsetAsync(C->d_Hs.p, 0, ns * 4, st));
        GFFX_CUDA(cudaMemsetAsync(C->d_He.p, 0, ns * 4, st));
        GFFX_CUDA(cudaMemsetAsync(C->d_G.p, 0, ng * 4, st));
    } else {
        const size_t nw = std::max<uint64_t>(h.n_words, 1);
        C->d_bits.ensure((nw + 8) * 8); C->d_diff.ensure(nw * 4); C->d_P.ensure((nw / 8 + 2) * 4);
        C->d_routed.ensure(std::max<uint32_t>(E->n_blocks, 1));
        GFFX_CUDA(cudaMemsetAsync(C->d_bits.p, 0, (nw + 8) * 8, st));   // + the tail of the last group of eight words
        GFFX_CUDA(cudaMemsetAsync(C->d_diff.p, 0, nw * 4, st));
        GFFX_CUDA(cudaMemsetAsync(C->d_routed.p, 0, std::max<uint32_t>(E->n_blocks, 1), st));
    }
    GFFX_CUDA(cudaStreamSynchronize(st));
    C->active = true;
    C->exchanged = false;
}

static CountsDev make_dev(EngineImpl* E, CountsState* C) {
    CountsDev D{};
    D.R.X = C->d_X.as<uint32_t>();
    D.R.dir = C->d_dir.as<uint32_t>();
    D.R.seqs = C->d_rank.as<RankSeq>();
    D.R.n_seqs = C->host.n_spaces;
    D.cov = C->d_cov.as<CovSpace>();
    D.gsp0 = C->d_gsp0.as<uint32_t>();
    D.gap_bits = C->d_gap_bits.as<uint32_t>();
    D.gs_first = C->d_gs_first.as<uint32_t>();
    D.gap_a = C->d_gap_a.as<uint32_t>();
    D.gap_b = C->d_gap_b.as<uint32_t>();
    D.gap_seg = C->d_gap_seg.as<uint32_t>();
    D.irr_blk = C->d_irr_blk.as<uint32_t>();
    D.irr_seq0 = C->d_irr_seq0.as<uint32_t>(); D.irr_s = C->d_irr_s.as<uint32_t>(); D.irr_e = C->d_irr_e.as<uint32_t>();
    D.irr_pm = C->d_irr_pm.as<uint32_t>(); D.irr_space = C->d_irr_space.as<uint32_t>();
    D.irr_grp = C->d_irr_grp.as<uint32_t>();
    D.irr_pos = C->d_irr_pos.as<uint32_t>();
    D.irr_base = C->d_irr_base.as<uint32_t>();
    D.grp_off = C->d_grp_off.as<uint32_t>();
    D.grp_seq = C->d_grp_seq.as<uint32_t>();
    D.grp_s = C->d_grp_s.as<uint32_t>();
    D.grp_e = C->d_grp_e.as<uint32_t>();
    D.n_seqs = (uint32_t)E->H->vmap.n_vseqs();   // the bucketed regions carry virtual sequence numbers (0xFFFFFFFF: unknown sequence)
    D.Hs = C->d_Hs.as<uint32_t>();
    D.He = C->d_He.as<uint32_t>();
    D.G = C->d_G.as<uint32_t>();
    D.bits = C->d_bits.as<unsigned long long>();
    D.diff = C->d_diff.as<int>();
    D.blk_routed = C->d_routed.as<uint8_t>();
    D.diff_used = C->d_err.as<uint32_t>() + 4;
    return D;
}

// one batch of device-resident regions: K0 partition + the scatter kernel; asynchronous on the engine's stream
static void counts_add_device(EngineImpl* E, CountsState* C, const CountsDev& D, const uint32_t* dc, const uint32_t* ds, const uint32_t* de,
                              uint64_t n, uint2* ose) {
    C->launches += E->partition_into(dc, ds, de, n, ose, E->d_poff.as<uint32_t>());
    if (C->kind == 0) {   // a CTA per run of a window's reads, shared-memory histograms (counts.cuh: depth_window_kernel)
        const DepthWin W{E->WC.win_seq, E->WC.win0, E->d_poff.as<uint32_t>(), E->WC.shift, E->WC.n_win, C->CXS, C->host.n_irregular != 0};
        const unsigned grid = (unsigned)(n / DW_RUN) + E->WC.n_win + 1;
        depth_window_kernel<<<grid, DW_THREADS, DW_SMEM, E->st>>>(D, W, ose, C->d_err.as<uint32_t>());
    } else {
        coverage_scatter_kernel<<<(unsigned)((n + CV_TILE - 1) / CV_TILE), CV_THREADS, 0, E->st>>>(D, E->WC.win_seq, E->WC.win0, E->WC.shift, E->d_poff.as<uint32_t>(), E->WC.n_win, ose, n,
                                                                                                  C->d_err.as<uint32_t>());
    }
    C->launches += 1;
}

void counts_add_impl(EngineImpl* E, const uint32_t* chr, const uint32_t* start, const uint32_t* end, uint64_t n, uint32_t flags) {
    CountsState* C = E->counts;
    if (!C || !C->active) throw Err(E_STATE, "gffx_counts_add without gffx_counts_begin");
    if (n == 0) return;
    if (!chr || !start || !end) throw Err(E_ARG, "NULL region arrays");
    if (n >= (1ull << 31)) throw Err(E_ARG, "more than 2^31-1 regions in one call: split the batch");
    if (C->n_regions + n >= (1ull << 32)) throw Err(E_ARG, "more than 2^32-1 regions in one depth/coverage job (32-bit histogram slots)");
    GFFX_CUDA(cudaSetDevice(E->dev));
    const CountsDev D = make_dev(E, C);
    const bool dev_in = flags & 1u;  // GFFX_F_DEVICE_INPUT
    // host input: chunks of the copy cross PCIe while the previous chunk is partitioned and scattered
    uint64_t chunk = n;
    if (!dev_in && n >= (1u << 22)) chunk = ((n + 7) / 8 + PART_TILE - 1) / PART_TILE * PART_TILE;
    const uint32_t n_chunks = (uint32_t)((n + chunk - 1) / chunk);
    E->d_bse.ensure(chunk * 8 + 32);
    GFFX_CUDA(cudaEventRecord(C->ev0, E->st));
    if (dev_in) {
        counts_add_device(E, C, D, chr, start, end, n, E->d_bse.as<uint2>());
    } else {
        const uint32_t* src[3] = {chr, start, end};
        for (int k = 0; k < 3; k++) E->d_q[k].ensure(n * 4);
        // a queued intersect job of this lane may still be reading its staged regions
        if (E->lanes[E->lane_active].ev_done) GFFX_CUDA(cudaStreamWaitEvent(E->st_copy, E->lanes[E->lane_active].ev_done, 0));
        while (E->ev_copy.size() < n_chunks) {
            cudaEvent_t e;
            GFFX_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            E->ev_copy.push_back(e);
        }
        for (uint32_t c = 0; c < n_chunks; c++) {
            const uint64_t o = (uint64_t)c * chunk, m = std::min<uint64_t>(chunk, n - o);
            for (int k = 0; k < 3; k++)
                GFFX_CUDA(cudaMemcpyAsync(E->d_q[k].as<uint32_t>() + o, src[k] + o, m * 4, cudaMemcpyHostToDevice, E->st_copy));
            GFFX_CUDA(cudaEventRecord(E->ev_copy[c], E->st_copy));
        }
        for (uint32_t c = 0; c < n_chunks; c++) {
            const uint64_t o = (uint64_t)c * chunk, m = std::min<uint64_t>(chunk, n - o);
            GFFX_CUDA(cudaStreamWaitEvent(E->st, E->ev_copy[c], 0));
            // the bucketed form of a chunk is scratch: the next chunk's partition overwrites it (stream order)
            counts_add_device(E, C, D, E->d_q[0].as<uint32_t>() + o, E->d_q[1].as<uint32_t>() + o, E->d_q[2].as<uint32_t>() + o, m,
                              E->d_bse.as<uint2>());
        }
    }
    GFFX_CUDA(cudaEventRecord(C->ev1, E->st));
    GFFX_CUDA(cudaMemcpyAsync(E->h_small, C->d_err.p, 4, cudaMemcpyDeviceToHost, E->st));
    GFFX_CUDA(cudaStreamSynchronize(E->st));
    GFFX_CUDA(cudaGetLastError());
    float ms;
    GFFX_CUDA(cudaEventElapsedTime(&ms, C->ev0, C->ev1));
    C->ms += ms;
    C->n_regions += n;
    if (*(uint32_t*)E->h_small) throw Err(E_ARG, "region with a chromosome number beyond the index's sequence list");
}

void counts_finish_impl(EngineImpl* E, CountRows* rows) {
    CountsState* C = E->counts;
    if (!C || !C->active) throw Err(E_STATE, "gffx_counts_finish without gffx_counts_begin");
    GFFX_CUDA(cudaSetDevice(E->dev));
    const CountsIndexHost& h = C->host;
    cudaStream_t st = E->st;
    *rows = CountRows();
    C->active = false;
    const uint32_t n_ids = h.n_ids;
    if (n_ids == 0 || h.n_units == 0) return;
    typedef unsigned long long ull;
    const uint64_t scan_n = std::max<uint64_t>({(uint64_t)h.n_slots, h.n_words, (uint64_t)n_ids, 1});
    C->d_scratch.ensure(scan_scratch_elems(scan_n) * 8);
    C->r_chrom.ensure(n_ids * 4ull); C->r_start.ensure(n_ids * 4ull); C->r_end.ensure(n_ids * 4ull); C->r_value.ensure(n_ids * 8ull);
    C->r_flag.ensure(n_ids); C->r_pos.ensure(n_ids * 4ull);
    GFFX_CUDA(cudaEventRecord(C->ev0, st));
    CountsFin F{};
    F.id_unit0 = C->d_id_unit0.as<uint32_t>(); F.id_units = C->d_id_units.as<uint32_t>();
    F.u_space = C->d_u_space.as<uint32_t>(); F.u_chrom = C->d_u_chrom.as<uint32_t>(); F.u_blk = C->d_u_blk.as<uint32_t>(); F.u_base = C->d_u_base.as<uint32_t>(); F.u_rs = C->d_u_rs.as<uint32_t>(); F.u_re = C->d_u_re.as<uint32_t>();
    F.u_line0 = C->d_u_line0.as<uint32_t>(); F.u_seg0 = C->d_u_seg0.as<uint32_t>();
    F.l_fs = C->d_l_fs.as<uint32_t>(); F.l_fe = C->d_l_fe.as<uint32_t>(); F.l_sfs = C->d_l_sfs.as<uint32_t>(); F.l_sfe = C->d_l_sfe.as<uint32_t>();
    F.g_s = C->d_g_s.as<uint32_t>(); F.g_e = C->d_g_e.as<uint32_t>(); F.g_ss = C->d_g_ss.as<uint32_t>(); F.g_se = C->d_g_se.as<uint32_t>();
    F.cov = C->d_cov.as<CovSpace>();
    F.n_ids = n_ids;
    if (C->kind == 0) {
        const uint64_t ns = h.n_slots;
        C->launches += device_scan<uint32_t, OpSum, PtrLoad<uint32_t>, PtrStore<uint32_t>, false, true>(
            PtrLoad<uint32_t>{C->d_Hs.as<uint32_t>()}, PtrStore<uint32_t>{C->d_S.as<uint32_t>()}, ns, OpSum(), 0u, C->d_scratch.as<uint32_t>(), nullptr, st, nullptr, E->chain());
        C->launches += device_scan<uint32_t, OpSum, PtrLoad<uint32_t>, PtrStore<uint32_t>, false, true>(
            PtrLoad<uint32_t>{C->d_He.as<uint32_t>()}, PtrStore<uint32_t>{C->d_E.as<uint32_t>()}, ns, OpSum(), 0u, C->d_scratch.as<uint32_t>(), nullptr, st, nullptr, E->chain());
        F.S = C->d_S.as<uint32_t>(); F.E = C->d_E.as<uint32_t>(); F.G = C->d_G.as<uint32_t>();
    } else {
        const uint64_t nw = h.n_words;
        // the difference array is only ever touched by regions longer than COV_DIRECT_WORDS words: while the job's flag says none
        // came, its scan returns at once (a sharded job's arrays were combined with other shards': always scanned)
        C->launches += device_scan<int, OpSum, DiffLoad, CoverStore, false, true>(DiffLoad{C->d_diff.as<int>()}, CoverStore{C->d_bits.as<ull>()}, nw,
                                                                                OpSum(), 0, C->d_scratch.as<int>(), nullptr, st,
                                                                                C->exchanged ? nullptr : C->d_err.as<uint32_t>() + 4, E->chain());
        C->launches += device_scan<uint32_t, OpSum, PopcLoad8, PtrStore<uint32_t>, false, false>(
            PopcLoad8{C->d_bits.as<ull>()}, PtrStore<uint32_t>{C->d_P.as<uint32_t>()}, (nw + 7) / 8, OpSum(), 0u, C->d_scratch.as<uint32_t>(), nullptr, st, nullptr, E->chain(),
            1024 /* 64 bytes per element: the bitmap is read once instead of twice */);
        F.bits = C->d_bits.as<ull>(); F.P = C->d_P.as<uint32_t>(); F.blk_routed = C->d_routed.as<uint8_t>();
    }
    RowArrays R{C->r_chrom.as<uint32_t>(), C->r_start.as<uint32_t>(), C->r_end.as<uint32_t>(), C->r_value.as<ull>(), C->r_flag.as<uint8_t>()};
    const unsigned nb = (n_ids + 255) / 256;
    if (C->kind == 0) counts_rows_kernel<0><<<nb, 256, 0, st>>>(F, R);
    else counts_rows_kernel<1><<<nb, 256, 0, st>>>(F, R);
    C->launches++;
    uint32_t* d_total = C->d_err.as<uint32_t>() + 8;
    C->launches += device_scan<uint32_t, OpSum, FlagLoad, PtrStore<uint32_t>, false, false>(FlagLoad{R.flag}, PtrStore<uint32_t>{C->r_pos.as<uint32_t>()}, n_ids,
                                                                                         OpSum(), 0u, C->d_scratch.as<uint32_t>(), d_total, st, nullptr, E->chain());
    GFFX_CUDA(cudaMemcpyAsync(E->h_small, d_total, 4, cudaMemcpyDeviceToHost, st));
    GFFX_CUDA(cudaStreamSynchronize(st));
    GFFX_CUDA(cudaGetLastError());
    const uint32_t n_rows = *(uint32_t*)E->h_small;
    if (n_rows) {
        C->o_id.ensure(n_rows * 4ull); C->o_chrom.ensure(n_rows * 4ull); C->o_start.ensure(n_rows * 4ull); C->o_end.ensure(n_rows * 4ull);
        C->o_value.ensure(n_rows * 8ull);
        counts_compact_kernel<<<nb, 256, 0, st>>>(R, C->r_pos.as<uint32_t>(), n_ids, C->o_id.as<uint32_t>(), C->o_chrom.as<uint32_t>(),
                                                  C->o_start.as<uint32_t>(), C->o_end.as<uint32_t>(), C->o_value.as<ull>());
        C->launches++;
    }
    GFFX_CUDA(cudaEventRecord(C->ev1, st));
    if (n_rows) {
        rows->id_idx.resize(n_rows); rows->chrom_idx.resize(n_rows); rows->start.resize(n_rows); rows->end.resize(n_rows); rows->value.resize(n_rows);
        GFFX_CUDA(cudaMemcpyAsync(rows->id_idx.data(), C->o_id.p, n_rows * 4ull, cudaMemcpyDeviceToHost, st));
        GFFX_CUDA(cudaMemcpyAsync(rows->chrom_idx.data(), C->o_chrom.p, n_rows * 4ull, cudaMemcpyDeviceToHost, st));
        GFFX_CUDA(cudaMemcpyAsync(rows->start.data(), C->o_start.p, n_rows * 4ull, cudaMemcpyDeviceToHost, st));
        GFFX_CUDA(cudaMemcpyAsync(rows->end.data(), C->o_end.p, n_rows * 4ull, cudaMemcpyDeviceToHost, st));
        GFFX_CUDA(cudaMemcpyAsync(rows->value.data(), C->o_value.p, n_rows * 8ull, cudaMemcpyDeviceToHost, st));
    }
    GFFX_CUDA(cudaStreamSynchronize(st));
    GFFX_CUDA(cudaGetLastError());
    float ms;
    GFFX_CUDA(cudaEventElapsedTime(&ms, C->ev0, C->ev1));
    C->ms += ms;
}

// accumulators of the running depth / coverage job and how shards combine them (engine.h ReduceOp)
bool counts_buffer_impl(EngineImpl* E, uint32_t k, void** dptr, uint64_t* n_elems, int* op) {
    CountsState* C = E->counts;
    if (!C || !C->active) throw Err(E_STATE, "no depth / coverage job is open");
    const CountsIndexHost& h = C->host;
    struct B { void* p; uint64_t n; int op; };
    B v[3];
    if (C->kind == 0) {
        v[0] = {C->d_Hs.p, h.n_slots, RED_SUM_U32};   // region starts per rank slot
        v[1] = {C->d_He.p, h.n_slots, RED_SUM_U32};   // region ends per rank slot
        v[2] = {C->d_G.p, h.n_segs, RED_SUM_U32};     // regions lying inside the gap after each segment
    } else {
        v[0] = {C->d_bits.p, h.n_words, RED_OR_U64};  // covered bases (partial words)
        v[1] = {C->d_diff.p, h.n_words, RED_SUM_I32}; // fully covered words, difference array
        C->exchanged = true;                          // other shards' long regions arrive through it
        v[2] = {C->d_routed.p, E->n_blocks, RED_MAX_U8};
    }
    if (k >= 3) return false;
    *dptr = v[k].p; *n_elems = v[k].n; *op = v[k].op;
    return true;
}

double counts_ms_impl(const EngineImpl* E) { return E->counts ? E->counts->ms : 0.0; }
uint32_t counts_launches_impl(const EngineImpl* E) { return E->counts ? E->counts->launches : 0; }

}  // namespace gffx
