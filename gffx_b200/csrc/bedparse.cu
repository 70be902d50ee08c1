// bedparse.cu — host driver of K9 (bedparse.cuh): a BED file to device-resident (chr, start, end) arrays.
#include <algorithm>
#include "bedparse.cuh"
#include "engine_impl.cuh"

namespace gffx {

// sequence names -> numbers as H->seq_num resolves them (a name listed twice in .sqs maps to its LAST number)
static void build_bed_names(EngineImpl* E) {
    const auto& M = E->H->seq_num;
    uint32_t size = 16;
    while (size < 4 * M.size() + 4) size <<= 1;
    std::vector<uint64_t> hash(size, 0);
    std::vector<uint32_t> id(size, 0), off(size, 0), len(size, 0);
    std::vector<uint8_t> blob;
    for (const auto& kv : M) {
        const uint64_t h = bed_fnv1a((const uint8_t*)kv.first.data(), (uint32_t)kv.first.size());
        uint32_t k = (uint32_t)h & (size - 1);
        while (hash[k]) k = (k + 1) & (size - 1);
        hash[k] = h; id[k] = kv.second; off[k] = (uint32_t)blob.size(); len[k] = (uint32_t)kv.first.size();
        blob.insert(blob.end(), kv.first.begin(), kv.first.end());
    }
    if (blob.empty()) blob.push_back(0);
    upload(E->d_bed_nhash, hash, E->st);
    upload(E->d_bed_nid, id, E->st);
    upload(E->d_bed_noff, off, E->st);
    upload(E->d_bed_nlen, len, E->st);
    upload(E->d_bed_nblob, blob, E->st);
    GFFX_CUDA(cudaStreamSynchronize(E->st));
    E->bed_names = BedNames{E->d_bed_nhash.as<uint64_t>(), E->d_bed_nid.as<uint32_t>(), E->d_bed_noff.as<uint32_t>(),
                            E->d_bed_nlen.as<uint32_t>(), E->d_bed_nblob.as<uint8_t>(), size - 1};
    E->bed_names_built = true;
}

void bed_parse_device_impl(EngineImpl* E, const std::string& path, int dialect, const uint32_t** dc, const uint32_t** ds, const uint32_t** de,
                           uint64_t* n_out) {
    if (dialect != 0 && dialect != 1) throw Err(E_ARG, "dialect must be 0 (intersect) or 1 (depth / coverage)");
    GFFX_CUDA(cudaSetDevice(E->dev));
    cudaStream_t st = E->st;
    *dc = *ds = *de = nullptr;
    *n_out = 0;
    E->bed_ms = 0; E->bed_launches = 0; E->bed_host_lines = 0; E->bed_bytes = 0;
    Mapped m;
    m.open(path);
    const uint64_t n = m.size();
    if (n == 0) return;
    if (!E->bed_names_built) build_bed_names(E);
    if (!E->ev_bed[0])
        for (auto& e : E->ev_bed) GFFX_CUDA(cudaEventCreate(&e));
    E->bed_bytes = n;
    E->d_bed_text.ensure(n + 32);
    GFFX_CUDA(cudaMemcpyAsync(E->d_bed_text.p, m.data(), n, cudaMemcpyHostToDevice, st));
    const uint8_t* text = E->d_bed_text.as<uint8_t>();
    GFFX_CUDA(cudaEventRecord(E->ev_bed[0], st));

    // ---- line starts -----------------------------------------------------------------------------------------
    const uint64_t tiles = (n + BED_TILE - 1) / BED_TILE;
    if (tiles >= (1ull << 31)) throw Err(E_ARG, "BED file too large for one pass");
    E->d_bed_tiles.ensure((tiles + 1) * 4);
    E->d_bed_tfirst.ensure((tiles + 1) * 4);
    E->d_scan_scratch.ensure(scan_scratch_elems(tiles) * 4 + 64);
    uint32_t* small32 = E->d_small.as<uint32_t>() + 40;   // [40] newline total, [41] flagged lines, [42] kept lines
    uint32_t* h32 = (uint32_t*)E->h_small + 40;
    GFFX_CUDA(cudaMemsetAsync(small32, 0, 16, st));
    bed_count_nl_kernel<<<(unsigned)tiles, BED_THREADS, 0, st>>>(text, n, E->d_bed_tiles.as<uint32_t>());
    E->bed_launches += 1 + device_scan<uint32_t, OpSum, PtrLoad<uint32_t>, PtrStore<uint32_t>, false, false>(
                               PtrLoad<uint32_t>{E->d_bed_tiles.as<uint32_t>()}, PtrStore<uint32_t>{E->d_bed_tfirst.as<uint32_t>()}, tiles, OpSum(),
                               0u, E->d_scan_scratch.as<uint32_t>(), small32, st);
    GFFX_CUDA(cudaMemcpyAsync(h32, small32, 4, cudaMemcpyDeviceToHost, st));
    GFFX_CUDA(cudaStreamSynchronize(st));
    const uint64_t n_lines = (uint64_t)h32[0] + 1;
    if (n_lines >= (1ull << 31)) throw Err(E_ARG, "more than 2^31-1 lines in one BED file: split it");
    E->d_bed_lstart.ensure((n_lines + 1) * 8);
    E->d_bed_status.ensure(n_lines);
    for (int k = 0; k < 3; k++) E->d_bed_line[k].ensure(n_lines * 4);
    E->d_bed_flag.ensure(n_lines * 4);
    E->d_bed_pos.ensure(n_lines * 4);
    bed_line_starts_kernel<<<(unsigned)tiles, BED_THREADS, 0, st>>>(text, n, E->d_bed_tfirst.as<uint32_t>(), E->d_bed_lstart.as<uint64_t>());

    // ---- one thread per line ---------------------------------------------------------------------------------
    const unsigned lblocks = (unsigned)((n_lines + BED_THREADS - 1) / BED_THREADS);
    uint8_t* status = E->d_bed_status.as<uint8_t>();
    uint32_t *lc = E->d_bed_line[0].as<uint32_t>(), *ls = E->d_bed_line[1].as<uint32_t>(), *le = E->d_bed_line[2].as<uint32_t>();
    if (dialect == 0)
        bed_parse_lines_kernel<0><<<lblocks, BED_THREADS, 0, st>>>(text, n, E->d_bed_lstart.as<uint64_t>(), n_lines, E->bed_names, status, lc, ls, le,
                                                                    small32 + 1, E->d_bed_flag.as<uint32_t>());
    else
        bed_parse_lines_kernel<1><<<lblocks, BED_THREADS, 0, st>>>(text, n, E->d_bed_lstart.as<uint64_t>(), n_lines, E->bed_names, status, lc, ls, le,
                                                                    small32 + 1, E->d_bed_flag.as<uint32_t>());
    E->bed_launches += 2;
    GFFX_CUDA(cudaEventRecord(E->ev_bed[1], st));
    GFFX_CUDA(cudaMemcpyAsync(h32 + 1, small32 + 1, 4, cudaMemcpyDeviceToHost, st));
    GFFX_CUDA(cudaStreamSynchronize(st));
    const uint32_t n_flag = h32[1];

    // ---- lines the kernel left to the host: the exact per-line functions, in file order -----------------------------
    if (n_flag) {
        std::vector<uint32_t> lines(n_flag);
        GFFX_CUDA(cudaMemcpy(lines.data(), E->d_bed_flag.p, (size_t)n_flag * 4, cudaMemcpyDeviceToHost));
        std::sort(lines.begin(), lines.end());
        // their byte offsets: a handful of 8-byte reads (runs of flagged lines are rare; a file of them still works, slowly)
        std::vector<uint64_t> offs(n_flag);
        for (uint32_t i = 0; i < n_flag; i++)
            GFFX_CUDA(cudaMemcpyAsync(&offs[i], E->d_bed_lstart.as<uint64_t>() + lines[i], 8, cudaMemcpyDeviceToHost, st));
        GFFX_CUDA(cudaStreamSynchronize(st));
        std::vector<uint8_t> pst(n_flag);
        std::vector<uint32_t> pc(n_flag, 0), ps(n_flag, 0), pe(n_flag, 0);
        std::string scratch;
        uint32_t cached_chr = NONE32 - 1;
        const uint8_t* base = m.data();
        for (uint32_t i = 0; i < n_flag; i++) {
            const uint8_t* line = base + offs[i];
            const uint8_t* nl = find_byte(line, base + n, '\n');
            uint32_t c = 0, s = 0, e = 0;
            bool keep;
            if (dialect == 0) {
                keep = bed_intersect_line(line, nl ? nl : base + n, E->H->seq_num, path, scratch, &cached_chr, &s, &e);   // may throw: the reference aborts
                c = cached_chr;
            } else {
                keep = bed_counts_line(line, nl ? nl + 1 : base + n, E->H->seq_num, scratch, &c, &s, &e);   // this dialect's slice keeps its '\n'
            }
            pst[i] = keep ? BED_KEEP : BED_SKIP;
            pc[i] = c; ps[i] = s; pe[i] = e;
        }
        DBuf pl, pt, dcx, dsx, dex;
        upload(pl, lines, st); upload(pt, pst, st); upload(dcx, pc, st); upload(dsx, ps, st); upload(dex, pe, st);
        bed_patch_kernel<<<(n_flag + BED_THREADS - 1) / BED_THREADS, BED_THREADS, 0, st>>>(pl.as<uint32_t>(), pt.as<uint8_t>(), dcx.as<uint32_t>(),
                                                                                           dsx.as<uint32_t>(), dex.as<uint32_t>(), n_flag, status, lc, ls, le);
        GFFX_CUDA(cudaStreamSynchronize(st));
        pl.release(); pt.release(); dcx.release(); dsx.release(); dex.release();
        E->bed_launches++;
        E->bed_host_lines = n_flag;
    }

    // ---- kept lines -> dense arrays in file order ------------------------------------------------------------------
    GFFX_CUDA(cudaEventRecord(E->ev_bed[2], st));
    E->d_scan_scratch.ensure(scan_scratch_elems(n_lines) * 4 + 64);
    E->bed_launches += device_scan<uint32_t, OpSum, BedKeepLoad, PtrStore<uint32_t>, false, false>(
        BedKeepLoad{status}, PtrStore<uint32_t>{E->d_bed_pos.as<uint32_t>()}, n_lines, OpSum(), 0u, E->d_scan_scratch.as<uint32_t>(), small32 + 2, st);
    GFFX_CUDA(cudaMemcpyAsync(h32 + 2, small32 + 2, 4, cudaMemcpyDeviceToHost, st));
    for (int k = 0; k < 3; k++) E->d_bed_out[k].ensure(n_lines * 4 + 16);   // kept <= lines: no second round trip for the size
    bed_compact_kernel<<<lblocks, BED_THREADS, 0, st>>>(status, E->d_bed_pos.as<uint32_t>(), lc, ls, le, n_lines, E->d_bed_out[0].as<uint32_t>(),
                                                        E->d_bed_out[1].as<uint32_t>(), E->d_bed_out[2].as<uint32_t>());
    E->bed_launches++;
    GFFX_CUDA(cudaEventRecord(E->ev_bed[3], st));
    GFFX_CUDA(cudaStreamSynchronize(st));
    GFFX_CUDA(cudaGetLastError());
    float a, b;
    GFFX_CUDA(cudaEventElapsedTime(&a, E->ev_bed[0], E->ev_bed[1]));
    GFFX_CUDA(cudaEventElapsedTime(&b, E->ev_bed[2], E->ev_bed[3]));
    E->bed_ms = (double)a + b;   // kernels only: the host round trips for sizes and the host-decided lines are outside
    *n_out = h32[2];
    *dc = E->d_bed_out[0].as<uint32_t>();
    *ds = E->d_bed_out[1].as<uint32_t>();
    *de = E->d_bed_out[2].as<uint32_t>();
}

}  // namespace gffx
