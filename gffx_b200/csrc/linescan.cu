// linescan.cu — the GFF line table built on the device.
//
// What it replaces: the reference re-parses the text of every block it touches on every run — split on '\n', skip '#',
// split on '\t', parse columns 4/5, look for "ID=" (src/commands/intersect.rs:284-321 + :441-489, src/commands/depth.rs:131-157 +
// fast_id :104-118, src/commands/coverage.rs:296-323).  This engine does that once per index, into the line table; the host
// scanner (host_index.cpp: scan_lines) spends ~155 ms of a 0.27 s open on it for the 416 MB bench GFF.  Here the text goes to
// HBM (where record emission, gather.cuh, wants it anyway) and three kernels produce the same columns:
//   ls_count_kernel   a warp per block: lines of the block (non-empty, not starting with '#'), "has a byte >= 0x80"
//   device_scan       first line number of every block
//   ls_starts_kernel  a warp per block again: the byte offset of each of its lines, in order
//   ls_parse_kernel   a thread per line: tabs, columns 4/5 (digits only, checked u32), column 1 against the sequence names
//                     (FNV-1a table, verified byte by byte), the type field, the first "ID=" of column 9 and its FNV-1a hash
// Every rule is the host scanner's; what needs more than bytes — UTF-8 validity — is left to it: a block with a byte >= 0x80 is
// flagged and the host re-scans just that block (its lines keep their places: counting lines is pure byte work).  The
// dictionaries (type, chromosome, ID: numbered by first appearance) are built on the host from the (offset, length, hash)
// triples the kernels return (host_index.cpp: lines_from_raw).
#include <fcntl.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <functional>
#include <string>
#include <thread>

#include "engine_impl.cuh"

namespace gffx {

namespace {

constexpr int LS_THREADS = 256;

struct LsBlocks {
    const uint64_t* bs;      // block start
    const uint64_t* be;      // min(block end, text length); == bs for a block that is not scanned
    const uint8_t* countable;
    uint32_t n;
};

// line starts inside [bs, be): p == bs or text[p-1] == '\n'; a line counts when its first byte is neither '\n' nor '#'
__device__ __forceinline__ bool ls_is_line(const uint8_t* __restrict__ text, uint64_t bs, uint64_t p) {
    const uint8_t c = text[p];
    return (p == bs || text[p - 1] == 10) && c != 10 && c != '#';
}

__global__ void __launch_bounds__(LS_THREADS) ls_count_kernel(const uint8_t* __restrict__ text, LsBlocks B, uint32_t* __restrict__ nlines, uint8_t* __restrict__ nonascii) {
    const uint32_t b = (blockIdx.x * LS_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (b >= B.n) return;
    const uint64_t bs = B.bs[b], be = B.be[b];
    uint32_t cnt = 0, hi = 0;
    for (uint64_t p = bs + lane; p < be; p += 32) {
        cnt += ls_is_line(text, bs, p) ? 1u : 0u;
        hi |= text[p] & 0x80u;
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    hi = __reduce_or_sync(0xffffffffu, hi);
    if (lane == 0) { nlines[b] = cnt; nonascii[b] = hi ? 1 : 0; }
}

struct U32Load {
    const uint32_t* p;
    __device__ unsigned long long operator()(uint64_t i) const { return p[i]; }
};

__global__ void __launch_bounds__(LS_THREADS) ls_starts_kernel(const uint8_t* __restrict__ text, LsBlocks B, const unsigned long long* __restrict__ line0,
                                                               unsigned long long* __restrict__ off, uint32_t* __restrict__ blk) {
    const uint32_t b = (blockIdx.x * LS_THREADS + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    if (b >= B.n) return;
    const uint64_t bs = B.bs[b], be = B.be[b];
    unsigned long long k = line0[b];
    for (uint64_t p0 = bs; p0 < be; p0 += 32) {   // uniform trip count
        const uint64_t p = p0 + lane;
        const bool is = p < be && ls_is_line(text, bs, p);
        const uint32_t m = __ballot_sync(0xffffffffu, is);
        if (is) {
            const unsigned long long o = k + __popc(m & ((1u << lane) - 1u));
            off[o] = p;
            blk[o] = b;
        }
        k += __popc(m);
    }
}

struct LsNames {   // open-addressing table over the sequence names (the LAST number of a name listed twice, as seq_num resolves it)
    const uint64_t* hash;   // 0 = empty
    const uint32_t* id;
    const uint32_t* name_off;
    const uint32_t* name_len;
    const uint8_t* blob;
    uint32_t mask;
};

__host__ __device__ inline uint64_t ls_fnv(const uint8_t* s, uint32_t n) {
    uint64_t h = 1469598103934665603ull;
    for (uint32_t i = 0; i < n; i++) { h ^= s[i]; h *= 1099511628211ull; }
    return h;
}

__device__ __forceinline__ uint32_t ls_lookup(const LsNames& T, const uint8_t* s, uint32_t n) {
    uint64_t h = ls_fnv(s, n);
    if (h == 0) h = 1;
    for (uint32_t k = (uint32_t)h & T.mask;; k = (k + 1) & T.mask) {
        const uint64_t hk = T.hash[k];
        if (hk == 0) return 0xFFFFFFFFu;
        if (hk == h && T.name_len[k] == n) {
            const uint8_t* b = T.blob + T.name_off[k];
            bool eq = true;
            for (uint32_t i = 0; i < n; i++) eq = eq && b[i] == s[i];
            if (eq) return T.id[k];
        }
    }
}

// parse_u32_strict (util.cpp): digits only, at least one, checked against u32::MAX; *out is written on success only
__device__ __forceinline__ bool ls_u32(const uint8_t* s, uint64_t n, uint32_t* out) {
    if (n == 0) return false;
    unsigned long long acc = 0;
    for (uint64_t i = 0; i < n; i++) {
        const uint32_t d = (uint32_t)s[i] - 48u;
        if (d > 9u) return false;
        acc = acc * 10 + d;
        if (acc > 0xFFFFFFFFull) return false;
    }
    *out = (uint32_t)acc;
    return true;
}

struct LsOut {
    uint32_t* len;
    uint8_t* flags;
    uint32_t *S, *E, *iseq, *fs, *fe;
    uint32_t *type_off, *type_len, *chrom_len, *id_off, *id_len;
    unsigned long long* id_hash;
};

// one thread per line; the rules are those of host_index.cpp: scan_lines for a block whose bytes are all ASCII
__global__ void __launch_bounds__(LS_THREADS) ls_parse_kernel(const uint8_t* __restrict__ text, LsBlocks B, const unsigned long long* __restrict__ off,
                                                              const uint32_t* __restrict__ blk, uint64_t n_lines, LsNames T, LsOut O) {
    const uint64_t i = (uint64_t)blockIdx.x * LS_THREADS + threadIdx.x;
    if (i >= n_lines) return;
    const uint64_t p = off[i];
    const uint32_t b = blk[i];
    const uint64_t be = B.be[b];
    uint64_t tab[8];
    int nt = 0;
    uint64_t q = p;
    for (; q < be; q++) {
        const uint8_t c = text[q];
        if (c == 10) break;
        if (c == 9 && nt < 8) tab[nt++] = q;
    }
    const uint64_t next = q < be ? q + 1 : be;
    uint8_t fl = 0;
    uint32_t S = 0, E = 0, iseq = 0xFFFFFFFFu, fs = 0, fe = 0;
    uint32_t type_off = 0, type_len = 0xFFFFFFFFu, chrom_len = 0xFFFFFFFFu, id_off = 0, id_len = 0;
    unsigned long long id_hash = 0;
    bool numok = false;
    if (nt >= 3) { type_off = (uint32_t)(tab[1] + 1 - p); type_len = (uint32_t)(tab[2] - tab[1] - 1); }
    if (nt >= 5) {
        numok = ls_u32(text + tab[2] + 1, tab[3] - tab[2] - 1, &S) && ls_u32(text + tab[3] + 1, tab[4] - tab[3] - 1, &E);
        if (numok) {
            iseq = ls_lookup(T, text + p, (uint32_t)(tab[0] - p));
            if (iseq != 0xFFFFFFFFu) fl |= LN_I_OK;
        }
    }
    if (nt >= 8 && numok && E != 0 && B.countable[b]) {
        // fast_id (depth.rs:104-118): the first "ID=" anywhere in column 9, its value up to ';', ' ' or '\t'
        const uint64_t a = tab[7] + 1;
        for (uint64_t k = a; k + 2 < q; k++)
            if (text[k] == 'I' && text[k + 1] == 'D' && text[k + 2] == '=') {
                uint64_t j = k + 3;
                while (j < q && text[j] != ';' && text[j] != ' ' && text[j] != 9) j++;
                const uint32_t lo = min(S, E), hi = max(S, E);
                fs = lo == 0 ? 0 : lo - 1;
                fe = hi;
                chrom_len = (uint32_t)(tab[0] - p);
                fl |= LN_D_OK;
                id_off = (uint32_t)(k + 3 - p);
                id_len = (uint32_t)(j - (k + 3));
                id_hash = ls_fnv(text + k + 3, id_len);
                break;
            }
    }
    O.len[i] = (uint32_t)(next - p);
    O.flags[i] = fl;
    O.S[i] = S; O.E[i] = E; O.iseq[i] = (fl & LN_I_OK) ? iseq : 0xFFFFFFFFu; O.fs[i] = fs; O.fe[i] = fe;
    O.type_off[i] = type_off; O.type_len[i] = type_len; O.chrom_len[i] = chrom_len; O.id_off[i] = id_off; O.id_len[i] = id_len;
    O.id_hash[i] = id_hash;
}

// ---- the dictionaries on the device (r03i: IDs, r03k: types and chromosomes) ---------------------------------------------------
// A string's number is the rank of its FIRST line among all first lines (host_index.cpp: number_ids; the type / chromosome pools
// number the same way).  Open-addressing table keyed by the string's FNV-1a hash, value = the smallest line number carrying it
// (atomicMin); every line then looks its first line up and compares its bytes with that line's — two different strings under
// one 64-bit hash, or anything else that does not compare equal, raises `bad` and the host builds the dictionaries instead.
// First lines are numbered by a scan.  FIELD: 0 the ID value (lines that count for depth / coverage), 1 the type column (every
// line that has one), 2 column 1 of the lines that count.
struct LsDict {
    const uint8_t* text;
    const unsigned long long* off;   // line offsets
    const uint8_t* flags;
    const uint32_t* id_off;
    const uint32_t* id_len;
    const unsigned long long* id_hash;
    const uint32_t* type_off;
    const uint32_t* type_len;
    const uint32_t* chrom_len;
    unsigned long long* keys;        // table: 0 = empty
    uint32_t* vals;                  // table: smallest line number of the key
    uint32_t mask;
    uint32_t* first;                 // per line: its string's first line, NONE32 without one
    uint64_t n;
};
template <int FIELD> __device__ __forceinline__ bool ls_field(const LsDict& D, uint64_t i, uint32_t* fo, uint32_t* fl) {
    if (FIELD == 0) { *fo = D.id_off[i]; *fl = D.id_len[i]; return D.flags[i] & LN_D_OK; }
    if (FIELD == 1) { *fo = D.type_off[i]; *fl = D.type_len[i]; return *fl != 0xFFFFFFFFu; }
    *fo = 0; *fl = D.chrom_len[i];
    return (D.flags[i] & LN_D_OK) && *fl != 0xFFFFFFFFu;
}
template <int FIELD> __device__ __forceinline__ unsigned long long ls_field_key(const LsDict& D, uint64_t i, uint32_t fo, uint32_t fl) {
    const unsigned long long h = FIELD == 0 ? D.id_hash[i] : ls_fnv(D.text + D.off[i] + fo, fl);
    return h ? h : 1ull;
}

template <int FIELD> __global__ void __launch_bounds__(256) ls_dict_insert_kernel(LsDict D) {
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    uint32_t fo, fl;
    if (i >= D.n || !ls_field<FIELD>(D, i, &fo, &fl)) return;
    const unsigned long long h = ls_field_key<FIELD>(D, i, fo, fl);
    for (uint32_t s = (uint32_t)h & D.mask;; s = (s + 1) & D.mask) {
        unsigned long long prev = D.keys[s];   // a few strings on millions of lines (types): most probes find their key already there
        if (prev == 0ull) prev = atomicCAS(D.keys + s, 0ull, h);
        if (prev == 0ull || prev == h) {
            if (D.vals[s] > (uint32_t)i) atomicMin(D.vals + s, (uint32_t)i);
            return;
        }
    }
}
template <int FIELD> __global__ void __launch_bounds__(256) ls_dict_first_kernel(LsDict D, uint8_t* __restrict__ is_first, uint32_t* __restrict__ bad) {
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= D.n) return;
    uint32_t f = 0xFFFFFFFFu, fo, fl;
    if (ls_field<FIELD>(D, i, &fo, &fl)) {
        const unsigned long long h = ls_field_key<FIELD>(D, i, fo, fl);
        uint32_t s = (uint32_t)h & D.mask;
        while (D.keys[s] != h) s = (s + 1) & D.mask;   // present: inserted by the kernel before
        f = D.vals[s];
        if (f != (uint32_t)i) {   // the same bytes as the first line's string?
            uint32_t go, gl;
            bool same = f < i && ls_field<FIELD>(D, f, &go, &gl) && gl == fl;
            if (same) {
                const uint8_t* a = D.text + D.off[i] + fo;
                const uint8_t* b = D.text + D.off[f] + go;
                for (uint32_t k = 0; k < fl; k++) same = same && a[k] == b[k];
            }
            if (!same) *bad = 1;
        }
    }
    D.first[i] = f;
    is_first[i] = f == (uint32_t)i ? 1 : 0;
}
struct U8Load {
    const uint8_t* p;
    __device__ uint32_t operator()(uint64_t i) const { return p[i]; }
};
// rank[i] = first lines before line i.  Line i's number = rank[first[i]]; a first line also files its string's place in the text
template <int FIELD>
__global__ void __launch_bounds__(256) ls_dict_number_kernel(LsDict D, const uint32_t* __restrict__ rank, uint32_t* __restrict__ id,
                                                             unsigned long long* __restrict__ ent_off, uint32_t* __restrict__ ent_len) {
    const uint64_t i = (uint64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= D.n) return;
    const uint32_t f = D.first[i];
    if (f == 0xFFFFFFFFu) { id[i] = 0xFFFFFFFFu; return; }
    const uint32_t r = rank[f];
    id[i] = r;
    if (f == (uint32_t)i) {
        uint32_t fo, fl;
        ls_field<FIELD>(D, i, &fo, &fl);
        ent_off[r] = D.off[i] + fo;
        ent_len[r] = fl;
    }
}

template <class T> void download(std::vector<T>& v, const DBuf& b, size_t n, cudaStream_t st) {
    v.resize(n);
    if (n) GFFX_CUDA(cudaMemcpyAsync(v.data(), b.p, n * sizeof(T), cudaMemcpyDeviceToHost, st));
}

}  // namespace

// The scanner proper.  Returns false (and leaves *out untouched) when no usable device is there; throws on CUDA errors.
// On success the text stays on the device: *d_text / *d_text_cap are handed to the engine (record emission reads it).
bool device_scan_lines(const HostIndex& H, int device, int threads, RawLines* out, void** d_text, size_t* d_text_cap) {
    int ndev = 0;
    if (device < 0 || cudaGetDeviceCount(&ndev) != cudaSuccess || device >= ndev) { cudaGetLastError(); return false; }
    PhaseTimer t_all("  GFF line table (device scan)");
    GFFX_CUDA(cudaSetDevice(device));
    cudaStream_t st;
    GFFX_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
    const size_t nb = H.blocks.size();
    const uint64_t n = H.gff_len;
    DBuf text, bs, be, cnt, nl, na, l0, scratch, off, blk, tot, obuf;
    std::vector<DBuf*> temps = {&bs, &be, &cnt, &nl, &na, &l0, &scratch, &tot, &obuf};
    DBuf o_len, o_flags, o_S, o_E, o_iseq, o_fs, o_fe, o_toff, o_tlen, o_clen, o_ioff, o_ilen, o_ihash, t_hash, t_id, t_off, t_len, t_blob;
    for (DBuf* b : {&t_hash, &t_id, &t_off, &t_len, &t_blob}) temps.push_back(b);   // the o_* columns are views into obuf
    struct Cleanup {
        std::vector<DBuf*>& v; cudaStream_t st; DBuf* text; bool keep_text = false;
        ~Cleanup() { for (DBuf* b : v) b->release(); if (!keep_text) text->release(); cudaStreamDestroy(st); }
    } cleanup{temps, st, &text};
    // ---- the text: page-locked staging buffers filled by the host threads, copied as they fill -------------------------------
    {
        PhaseTimer t("    text -> HBM");
        {
            PhaseTimer ta("      device allocation");
            text.ensure(n + 32);
            GFFX_CUDA(cudaMemsetAsync((uint8_t*)text.p + n, 0, 32, st));
            GFFX_CUDA(cudaStreamSynchronize(st));
        }
        // W workers, each with its own stream and two small page-locked buffers, take the text's chunks round robin: pread
        // (page cache -> staging, no faulting of the mapping's pages into this process) and the copy of the previous chunk overlap
        // within a worker, the workers' preads run side by side.  Small buffers: page-locking memory costs ~0.25 ms per MB.
        constexpr size_t CH = 2u << 20;
        const int W = std::max(1, std::min(threads, 6));
        const bool from_file = H.gff == H.gff_map.data() && !H.gff_path.empty();
        const uint64_t n_chunks = (n + CH - 1) / CH;
        std::vector<cudaError_t> werr((size_t)W, cudaSuccess);
        // ONE page-locked block for all the workers' buffers, from the library's pool (engine.cu: later results reuse it):
        // a cudaMallocHost per buffer cost 3 ms apiece under the driver's lock, more than the copies themselves
        size_t stage_cap = 0;
        uint8_t* const stage_all = (uint8_t*)pinned_acquire((size_t)W * 2 * CH, &stage_cap);
        struct StageGuard { void* p; size_t cap; ~StageGuard() { pinned_release(p, cap); } } stage_guard{stage_all, stage_cap};
        parallel_jobs((size_t)W, W, [&](size_t w) {
            {
                cudaError_t& e = werr[w];
                uint8_t* stage[2] = {stage_all + (size_t)(2 * w) * CH, stage_all + (size_t)(2 * w + 1) * CH};
                cudaEvent_t ev[2] = {nullptr, nullptr};
                cudaStream_t ws = nullptr;
                const int fd = from_file ? ::open(H.gff_path.c_str(), O_RDONLY) : -1;
                e = cudaSetDevice(device);
                if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ws, cudaStreamNonBlocking);
                for (int k = 0; k < 2 && e == cudaSuccess; k++) e = cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming);
                int k = 0;
                PhaseTimer tw(w == 0 ? "      worker 0: chunks" : "");
                if (w) tw.on = false;
                for (uint64_t c = (uint64_t)w; c < n_chunks && e == cudaSuccess; c += (uint64_t)W, k ^= 1) {
                    const uint64_t o = c * CH;
                    const size_t m = (size_t)std::min<uint64_t>(CH, n - o);
                    e = cudaEventSynchronize(ev[k]);   // the copy that last used this buffer
                    if (e != cudaSuccess) break;
                    size_t done = 0;
                    while (fd >= 0 && done < m) {
                        const ssize_t r = pread(fd, stage[k] + done, m - done, (off_t)(o + done));
                        if (r <= 0) break;
                        done += (size_t)r;
                    }
                    if (done != m) memcpy(stage[k], H.gff + o, m);
                    e = cudaMemcpyAsync((uint8_t*)text.p + o, stage[k], m, cudaMemcpyHostToDevice, ws);
                    if (e == cudaSuccess) e = cudaEventRecord(ev[k], ws);
                }
                if (ws) { const cudaError_t e2 = cudaStreamSynchronize(ws); if (e == cudaSuccess) e = e2; }
                tw.stop();
                for (int j = 0; j < 2; j++)
                    if (ev[j]) cudaEventDestroy(ev[j]);
                if (ws) cudaStreamDestroy(ws);
                if (fd >= 0) ::close(fd);
            }
        });
        for (cudaError_t e : werr) GFFX_CUDA(e);
    }
    PhaseTimer t_k("    kernels + read-back");
    PhaseTimer* t_pre = new PhaseTimer("      blocks, names, counts");
    // ---- blocks and sequence names ---------------------------------------------------------------------------------------------
    std::vector<uint64_t> h_bs(nb), h_be(nb);
    std::vector<uint8_t> h_countable(nb);
    for (size_t b = 0; b < nb; b++) {
        const bool scan = H.block_scannable(b);
        h_bs[b] = scan ? H.blocks[b].start : 0;
        h_be[b] = scan ? std::min<uint64_t>(H.blocks[b].end, n) : 0;
        h_countable[b] = H.block_countable(b) && H.blocks[b].end <= n;
    }
    upload(bs, h_bs, st); upload(be, h_be, st); upload(cnt, h_countable, st);
    size_t cap = 16;
    while (cap < 2 * H.seq_num.size() + 2) cap <<= 1;
    std::vector<uint64_t> th(cap, 0);
    std::vector<uint32_t> tid(cap, 0), toff(cap, 0), tlen(cap, 0);
    std::vector<uint8_t> blob;
    for (const auto& kv : H.seq_num) {
        uint64_t h = ls_fnv((const uint8_t*)kv.first.data(), (uint32_t)kv.first.size());
        if (h == 0) h = 1;
        size_t k = (size_t)h & (cap - 1);
        while (th[k]) k = (k + 1) & (cap - 1);
        th[k] = h; tid[k] = kv.second; toff[k] = (uint32_t)blob.size(); tlen[k] = (uint32_t)kv.first.size();
        blob.insert(blob.end(), kv.first.begin(), kv.first.end());
    }
    blob.push_back(0);
    upload(t_hash, th, st); upload(t_id, tid, st); upload(t_off, toff, st); upload(t_len, tlen, st); upload(t_blob, blob, st);
    const LsNames T{t_hash.as<uint64_t>(), t_id.as<uint32_t>(), t_off.as<uint32_t>(), t_len.as<uint32_t>(), t_blob.as<uint8_t>(), (uint32_t)cap - 1};
    const LsBlocks B{bs.as<uint64_t>(), be.as<uint64_t>(), cnt.as<uint8_t>(), (uint32_t)nb};
    // ---- lines per block, their first line numbers, their offsets ------------------------------------------------------------------
    typedef unsigned long long ull;
    nl.ensure(nb * 4 + 4); na.ensure(nb + 1); l0.ensure((nb + 1) * 8); tot.ensure(8);
    const unsigned warps_grid = (unsigned)((nb * 32 + LS_THREADS - 1) / LS_THREADS);
    ls_count_kernel<<<warps_grid, LS_THREADS, 0, st>>>(text.as<uint8_t>(), B, nl.as<uint32_t>(), na.as<uint8_t>());
    scratch.ensure(scan_scratch_elems(nb) * 8);
    device_scan<ull, OpSum, U32Load, PtrStore<ull>, false, false>(U32Load{nl.as<uint32_t>()}, PtrStore<ull>{l0.as<ull>()}, nb, OpSum(), 0ull, scratch.as<ull>(),
                                                                tot.as<ull>(), st);
    ull n_lines = 0;
    GFFX_CUDA(cudaMemcpyAsync(&n_lines, tot.p, 8, cudaMemcpyDeviceToHost, st));
    GFFX_CUDA(cudaStreamSynchronize(st));
    GFFX_CUDA(cudaGetLastError());
    if (n_lines >= (1ull << 32) - 1) throw Err(E_FORMAT, "more than 2^32-2 GFF lines inside blocks");
    delete t_pre;
    RawLines R;
    const size_t L = (size_t)n_lines;
    PhaseTimer* t_al = new PhaseTimer("      output buffers");
    // one allocation for the 15 output columns (a cudaMalloc each cost 2 ms apiece next to a 400 MB text allocation)
    {
        const size_t Lc = std::max<size_t>(L, 1);
        struct Col { DBuf* b; size_t width; };
        const Col cols[] = {{&off, 8}, {&o_ihash, 8}, {&blk, 4}, {&o_len, 4}, {&o_S, 4}, {&o_E, 4}, {&o_iseq, 4}, {&o_fs, 4}, {&o_fe, 4},
                            {&o_toff, 4}, {&o_tlen, 4}, {&o_clen, 4}, {&o_ioff, 4}, {&o_ilen, 4}, {&o_flags, 1}};
        size_t total = 0;
        for (const Col& c : cols) total += (Lc * c.width + 255) & ~(size_t)255;
        obuf.ensure(total);
        size_t o = 0;
        for (const Col& c : cols) {   // views into obuf: not owners (they are not in `temps`)
            c.b->p = (uint8_t*)obuf.p + o;
            c.b->cap = 0;
            o += (Lc * c.width + 255) & ~(size_t)255;
        }
    }
    delete t_al;
    if (L) {
        ls_starts_kernel<<<warps_grid, LS_THREADS, 0, st>>>(text.as<uint8_t>(), B, l0.as<ull>(), off.as<ull>(), blk.as<uint32_t>());
        const LsOut O{o_len.as<uint32_t>(), o_flags.as<uint8_t>(), o_S.as<uint32_t>(), o_E.as<uint32_t>(), o_iseq.as<uint32_t>(), o_fs.as<uint32_t>(),
                      o_fe.as<uint32_t>(), o_toff.as<uint32_t>(), o_tlen.as<uint32_t>(), o_clen.as<uint32_t>(), o_ioff.as<uint32_t>(), o_ilen.as<uint32_t>(),
                      o_ihash.as<ull>()};
        ls_parse_kernel<<<(unsigned)((L + LS_THREADS - 1) / LS_THREADS), LS_THREADS, 0, st>>>(text.as<uint8_t>(), B, off.as<ull>(), blk.as<uint32_t>(), L, T, O);
    }
    if (getenv("GFFX_TIMING")) { PhaseTimer tk("      (kernels, synchronised)"); GFFX_CUDA(cudaStreamSynchronize(st)); }
    // ---- read-back ---------------------------------------------------------------------------------------------------------------------
    PhaseTimer t_rb("      read-back");
    static_assert(sizeof(ull) == sizeof(uint64_t), "u64");
    GFFX_CUDA(cudaStreamSynchronize(st));   // the kernels
    GFFX_CUDA(cudaGetLastError());
    // ---- the ID dictionary, on the device when every block is ASCII (a block with other bytes is re-scanned by the host, which
    // may see its ID fields differently) ----------------------------------------------------------------------------------------
    DBuf ibuf;   // one allocation for everything below (the i_* are views into it)
    temps.push_back(&ibuf);
    DBuf i_keys, i_vals, i_first, i_isfirst, i_rank, i_bad, i_id[3], i_entoff[3], i_entlen[3];
    bool dev_ids = false;
    uint32_t n_ent[3] = {0, 0, 0};   // IDs, types, chromosomes
    {
        const char* e = getenv("GFFX_ID_DICT");   // "host": the host builds the dictionaries (A/B runs, tests)
        download(R.blk_nonascii, na, nb, st);
        GFFX_CUDA(cudaStreamSynchronize(st));
        bool ascii = true;
        for (uint8_t x : R.blk_nonascii) ascii = ascii && !x;
        if (L && ascii && !(e && !strcmp(e, "host"))) {
            PhaseTimer t_ids("      dictionaries (device)");
            size_t cap = 1024;
            while (cap < 2 * L) cap <<= 1;
            {
                struct Part { DBuf* b; size_t bytes; };
                const Part parts[] = {{&i_keys, cap * 8}, {&i_vals, cap * 4}, {&i_entoff[0], L * 8}, {&i_entoff[1], L * 8}, {&i_entoff[2], L * 8}, {&i_first, L * 4},
                                      {&i_rank, L * 4 + 4}, {&i_id[0], L * 4}, {&i_id[1], L * 4}, {&i_id[2], L * 4}, {&i_entlen[0], L * 4}, {&i_entlen[1], L * 4},
                                      {&i_entlen[2], L * 4}, {&i_isfirst, L}, {&i_bad, 16}};
                size_t total = 0;
                for (const Part& q : parts) total += (q.bytes + 255) & ~(size_t)255;
                ibuf.ensure(total);
                size_t o = 0;
                for (const Part& q : parts) { q.b->p = (uint8_t*)ibuf.p + o; q.b->cap = 0; o += (q.bytes + 255) & ~(size_t)255; }
            }
            GFFX_CUDA(cudaMemsetAsync(i_bad.p, 0, 16, st));
            scratch.ensure(scan_scratch_elems(L) * 4 + 64);
            const LsDict D{text.as<uint8_t>(), off.as<ull>(), o_flags.as<uint8_t>(), o_ioff.as<uint32_t>(), o_ilen.as<uint32_t>(), o_ihash.as<ull>(),
                           o_toff.as<uint32_t>(), o_tlen.as<uint32_t>(), o_clen.as<uint32_t>(),
                           i_keys.as<ull>(), i_vals.as<uint32_t>(), (uint32_t)cap - 1, i_first.as<uint32_t>(), (uint64_t)L};
            const unsigned g = (unsigned)((L + 255) / 256);
            uint32_t* bad = i_bad.as<uint32_t>();
            auto run = [&](int f, auto ins, auto fst, auto num) {
                GFFX_CUDA(cudaMemsetAsync(i_keys.p, 0, cap * 8, st));
                GFFX_CUDA(cudaMemsetAsync(i_vals.p, 0xFF, cap * 4, st));
                ins<<<g, 256, 0, st>>>(D);
                fst<<<g, 256, 0, st>>>(D, i_isfirst.as<uint8_t>(), bad);
                device_scan<uint32_t, OpSum, U8Load, PtrStore<uint32_t>, false, false>(U8Load{i_isfirst.as<uint8_t>()}, PtrStore<uint32_t>{i_rank.as<uint32_t>()}, L,
                                                                                      OpSum(), 0u, scratch.as<uint32_t>(), bad + 1 + f, st);
                num<<<g, 256, 0, st>>>(D, i_rank.as<uint32_t>(), i_id[f].as<uint32_t>(), i_entoff[f].as<ull>(), i_entlen[f].as<uint32_t>());
            };
            run(0, ls_dict_insert_kernel<0>, ls_dict_first_kernel<0>, ls_dict_number_kernel<0>);
            run(1, ls_dict_insert_kernel<1>, ls_dict_first_kernel<1>, ls_dict_number_kernel<1>);
            run(2, ls_dict_insert_kernel<2>, ls_dict_first_kernel<2>, ls_dict_number_kernel<2>);
            uint32_t res[4] = {1, 0, 0, 0};
            GFFX_CUDA(cudaMemcpyAsync(res, i_bad.p, 16, cudaMemcpyDeviceToHost, st));
            GFFX_CUDA(cudaStreamSynchronize(st));
            GFFX_CUDA(cudaGetLastError());
            dev_ids = res[0] == 0;   // otherwise: two strings under one hash — the host's dictionaries compare bytes on every probe
            for (int f = 0; f < 3; f++) n_ent[f] = res[1 + f];
        }
    }
    {
        // pageable destinations: a copy occupies its calling thread (driver staging) and the fresh vectors fault their pages in
        // as they are written, so the columns come back on several threads / streams side by side
        std::vector<std::function<void(cudaStream_t)>> jobs;
        auto col = [&](auto& v, const DBuf& b, size_t cnt) { jobs.push_back([&v, &b, cnt](cudaStream_t s) { download(v, b, cnt, s); }); };
        col(R.t.off, off, L); col(R.t.blk, blk, L); col(R.t.len, o_len, L); col(R.t.flags, o_flags, L);
        if (dev_ids) {
            col(R.t.id, i_id[0], L); col(R.ids_off, i_entoff[0], n_ent[0]); col(R.ids_len, i_entlen[0], n_ent[0]);
            col(R.t.type_id, i_id[1], L); col(R.types_off, i_entoff[1], n_ent[1]); col(R.types_len, i_entlen[1], n_ent[1]);
            col(R.t.chrom, i_id[2], L); col(R.chroms_off, i_entoff[2], n_ent[2]); col(R.chroms_len, i_entlen[2], n_ent[2]);
        } else {
            col(R.id_hash, o_ihash, L); col(R.id_off, o_ioff, L); col(R.id_len, o_ilen, L);
            col(R.type_off, o_toff, L); col(R.type_len, o_tlen, L); col(R.chrom_len, o_clen, L);
        }
        col(R.t.S, o_S, L); col(R.t.E, o_E, L); col(R.t.iseq, o_iseq, L); col(R.t.fs, o_fs, L); col(R.t.fe, o_fe, L);
        col(R.blk_nlines, nl, nb);
        const int W = std::max(1, std::min(threads, 6));
        std::vector<std::string> werr((size_t)W);
        std::atomic<size_t> next{0};
        parallel_jobs((size_t)W, W, [&](size_t w) {
            {
                cudaStream_t ws = nullptr;
                try {
                    GFFX_CUDA(cudaSetDevice(device));
                    GFFX_CUDA(cudaStreamCreateWithFlags(&ws, cudaStreamNonBlocking));
                    for (size_t k; (k = next.fetch_add(1)) < jobs.size();) {
                        jobs[k](ws);
                        GFFX_CUDA(cudaStreamSynchronize(ws));
                    }
                } catch (const std::exception& e) { werr[w] = e.what(); }
                if (ws) cudaStreamDestroy(ws);
            }
        });
        for (const std::string& e : werr)
            if (!e.empty()) throw Err(E_CUDA, e);
    }
    R.ids_ready = dev_ids;
    *out = std::move(R);
    *d_text = text.p;
    *d_text_cap = text.cap;
    cleanup.keep_text = true;
    return true;
}

}  // namespace gffx
namespace gffx {
void device_text_release(void* p, int device) {
    if (!p) return;
    int cur = 0;
    cudaGetDevice(&cur);
    cudaSetDevice(device);
    cudaFree(p);
    cudaSetDevice(cur);
}
}  // namespace gffx
