// bedparse.cuh — K9: BED text -> (chr, start, end) regions on the device.
//
// Replaces the single-threaded text loops of parse_bed_file (src/commands/intersect.rs:201-230, dialect 0) and of
// process_bed / collect_by_root_from_bed (src/commands/depth.rs:449-495, src/commands/coverage.rs:222-255, dialect 1);
// the two dialects differ in field separators, evaluation order and error behaviour.
//
//   bed_count_nl_kernel      newlines per 4 KB tile                      -> device_scan -> first line number of every tile
//   bed_line_starts_kernel   byte offset of every line start (u64)
//   bed_parse_lines_kernel   one thread per line: PLAIN lines (ASCII only, digit-only numbers of <= 10 digits that fit a
//                            u32) are decided on the spot with the reference's rules; anything else — non-ASCII bytes,
//                            signs, overlong or overflowing numbers, non-digits where the dialect would abort or needs
//                            Rust's str::parse / trim_end corner cases — is flagged and decided by the host's exact
//                            per-line functions (host_index.cpp), in file order, so the abort message of dialect 0 is
//                            that of the first offending line
//   device_scan + bed_compact_kernel   kept lines -> dense chr[] / start[] / end[] in file order
//
// The sequence names live in a small open-addressing table (FNV-1a 64 of the name, verified byte by byte).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include "scan.cuh"

namespace gffx {

constexpr int BED_THREADS = 256;
constexpr int BED_TILE = BED_THREADS * 16;   // bytes per CTA of the newline passes
enum : uint8_t { BED_SKIP = 0, BED_KEEP = 1, BED_HOST = 2 };

struct BedNames {
    const uint64_t* hash;      // table_size entries, 0 = empty
    const uint32_t* id;        // sequence number (the LAST number of a name listed twice, as seq_num resolves it)
    const uint32_t* name_off;  // into `blob`
    const uint32_t* name_len;
    const uint8_t* blob;
    uint32_t mask;             // table_size - 1
};

__host__ __device__ inline uint64_t bed_fnv1a(const uint8_t* s, uint32_t n) {
    uint64_t h = 1469598103934665603ull;
    for (uint32_t i = 0; i < n; i++) { h ^= s[i]; h *= 1099511628211ull; }
    return h ? h : 1ull;
}

__device__ __forceinline__ uint32_t bed_lookup(const BedNames& T, const uint8_t* s, uint32_t n) {
    const uint64_t h = bed_fnv1a(s, n);
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

// newlines that start another line: '\n' at position p with p + 1 < n
__device__ __forceinline__ uint32_t bed_nl_mask(const uint8_t* __restrict__ text, uint64_t n, uint64_t p0) {
    uint32_t m = 0;
    if (p0 + 16 < n) {   // all 16 positions have a successor
        const uint4 v = *reinterpret_cast<const uint4*>(text + p0);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int k = 0; k < 4; k++)
#pragma unroll
            for (int b = 0; b < 4; b++)
                if (((w[k] >> (8 * b)) & 0xFFu) == 10u) m |= 1u << (4 * k + b);
    } else {
        for (int i = 0; i < 16; i++)
            if (p0 + i + 1 < n && text[p0 + i] == 10) m |= 1u << i;
    }
    return m;
}

static __global__ void __launch_bounds__(BED_THREADS) bed_count_nl_kernel(const uint8_t* __restrict__ text, uint64_t n, uint32_t* __restrict__ tile_counts) {
    __shared__ uint32_t sm[BED_THREADS / 32];
    const uint64_t p0 = ((uint64_t)blockIdx.x * BED_THREADS + threadIdx.x) * 16;
    uint32_t c = p0 < n ? (uint32_t)__popc(bed_nl_mask(text, n, p0)) : 0u;
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
        for (int k = 0; k < BED_THREADS / 32; k++) t += sm[k];
        tile_counts[blockIdx.x] = t;
    }
}

// line 0 starts at byte 0; line tile_first[tile] + 1 + (rank inside the tile) starts after that newline
static __global__ void __launch_bounds__(BED_THREADS)
    bed_line_starts_kernel(const uint8_t* __restrict__ text, uint64_t n, const uint32_t* __restrict__ tile_first, uint64_t* __restrict__ line_start) {
    __shared__ uint32_t sm[BED_THREADS / 32 + 1];
    const uint64_t p0 = ((uint64_t)blockIdx.x * BED_THREADS + threadIdx.x) * 16;
    const uint32_t m = p0 < n ? bed_nl_mask(text, n, p0) : 0u;
    const uint32_t c = (uint32_t)__popc(m);
    uint32_t tot;
    const uint32_t ex = block_scan_excl(c, OpSum(), 0u, sm, &tot);
    uint64_t k = (uint64_t)tile_first[blockIdx.x] + ex + 1;
    for (uint32_t mm = m; mm; mm &= mm - 1) line_start[k++] = p0 + (uint32_t)__ffs(mm);   // byte after the newline
    if (blockIdx.x == 0 && threadIdx.x == 0) line_start[0] = 0;
}

__device__ __forceinline__ bool bed_ws0(uint8_t c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == 0x0c; }   // u8::is_ascii_whitespace
__device__ __forceinline__ bool bed_ws1(uint8_t c) { return c == ' ' || c == '\t'; }

// digits only, 1..10 of them, value <= u32::MAX: the plain case every number parser of the reference agrees on
__device__ __forceinline__ bool bed_plain_u32(const uint8_t* s, uint32_t n, uint32_t* out) {
    if (n == 0 || n > 10) return false;
    uint64_t v = 0;
    for (uint32_t i = 0; i < n; i++) {
        const uint32_t d = (uint32_t)s[i] - 48u;
        if (d > 9u) return false;
        v = v * 10 + d;
    }
    if (v > 0xFFFFFFFFull) return false;
    *out = (uint32_t)v;
    return true;
}

// one line [L, L + len) without its '\n': status and, for a kept line, its region.  L may point into shared or global memory
// (the caller instantiates both, so every byte load is of a known address space)
template <int DIALECT>
__device__ __forceinline__ uint8_t bed_parse_one(const uint8_t* __restrict__ L, uint32_t len, const BedNames& T, uint32_t* pc, uint32_t* ps,
                                                 uint32_t* pe) {
    uint8_t st = BED_SKIP;
    uint32_t c = 0, s = 0, e = 0;
    if (len != 0 && L[0] != '#') {
        bool ascii = true;
        for (uint32_t j = 0; j < len; j++) ascii = ascii && L[j] < 0x80;
        uint32_t f0[3] = {0, 0, 0}, f1[3] = {0, 0, 0};
        int nf = 0;
        uint32_t i = 0;
        while (nf < 3) {   // the first three fields: split_ascii_whitespace (dialect 0) / split on '\t' or ' ', empty fields dropped (1)
            while (i < len && (DIALECT == 0 ? bed_ws0(L[i]) : bed_ws1(L[i]))) i++;
            if (i >= len) break;
            f0[nf] = i;
            while (i < len && !(DIALECT == 0 ? bed_ws0(L[i]) : bed_ws1(L[i]))) i++;
            f1[nf] = i;
            nf++;
        }
        if (!ascii) {
            st = BED_HOST;   // UTF-8 validity (dialect 0 aborts on bad UTF-8) and Unicode trimming are the host's business
        } else if (nf == 3) {
            if (DIALECT == 0) {
                c = bed_lookup(T, L + f0[0], f1[0] - f0[0]);
                if (c != 0xFFFFFFFFu)   // unknown chromosome: skipped BEFORE the numbers are looked at
                    st = (bed_plain_u32(L + f0[1], f1[1] - f0[1], &s) && bed_plain_u32(L + f0[2], f1[2] - f0[2], &e)) ? BED_KEEP : BED_HOST;
            } else {
                // the reference's line slice of this dialect keeps its '\n', which can only end the line's last field; field 3
                // is trim_end()ed anyway (ASCII white space here, anything else was sent to the host above), so leaving the
                // '\n' out of [p, q) decides every line the same way
                uint32_t e1 = f1[2];
                while (e1 > f0[2] && (L[e1 - 1] == '\r' || L[e1 - 1] == '\n' || L[e1 - 1] == 0x0b || L[e1 - 1] == 0x0c)) e1--;
                if (bed_plain_u32(L + f0[1], f1[1] - f0[1], &s) && bed_plain_u32(L + f0[2], e1 - f0[2], &e)) {
                    if (s < e) {
                        c = bed_lookup(T, L + f0[0], f1[0] - f0[0]);
                        if (c != 0xFFFFFFFFu) st = BED_KEEP;
                    }
                } else {
                    st = BED_HOST;   // '+', overlong digits, junk: Rust's str::parse decides (a failure just drops the line)
                }
            }
        }
    }
    *pc = c; *ps = s; *pe = e;
    return st;
}

constexpr uint32_t BED_STAGE = 16384;   // bytes of text a CTA stages in shared memory (256 lines of a plain BED are ~6 KB)

// DIALECT 0: intersect (parse_bed_file), 1: depth / coverage.  text must be readable up to n + 16 (the staging copy is 16-byte wide).
template <int DIALECT>
static __global__ void __launch_bounds__(BED_THREADS)
    bed_parse_lines_kernel(const uint8_t* __restrict__ text, uint64_t n, const uint64_t* __restrict__ line_start, uint64_t n_lines, BedNames T,
                           uint8_t* __restrict__ status, uint32_t* __restrict__ oc, uint32_t* __restrict__ os, uint32_t* __restrict__ oe,
                           uint32_t* __restrict__ n_flag, uint32_t* __restrict__ flag_list) {
    __shared__ __align__(16) uint8_t buf[BED_STAGE + 16];
    const uint64_t k0 = (uint64_t)blockIdx.x * BED_THREADS, k = k0 + threadIdx.x;
    const uint64_t k1 = min(k0 + BED_THREADS, n_lines);
    // the CTA's lines are one contiguous span of the text: copied once with 16-byte loads, parsed from shared memory
    const uint64_t p_lo = line_start[k0], p_hi = k1 < n_lines ? line_start[k1] : n;
    const uint64_t a0 = p_lo & ~15ull;
    const bool staged = p_hi - a0 <= BED_STAGE;   // uniform
    if (staged)
        for (uint64_t o = (uint64_t)threadIdx.x * 16; a0 + o < p_hi; o += (uint64_t)BED_THREADS * 16)
            *reinterpret_cast<uint4*>(buf + o) = *reinterpret_cast<const uint4*>(text + a0 + o);
    __syncthreads();
    if (k >= n_lines) return;
    const uint64_t p = k == k0 ? p_lo : line_start[k];
    // q: the line's '\n', or the end of the text for an unterminated last line
    const uint64_t q = k + 1 < n_lines ? line_start[k + 1] - 1 : (text[n - 1] == 10 ? n - 1 : n);
    uint8_t st;
    uint32_t c = 0, s = 0, e = 0;
    const uint32_t len = (uint32_t)min(q - p, (uint64_t)0x7FFFFFFF);
    if (q - p > 4096) st = BED_HOST;   // a line this long is not a plain BED line
    else if (staged) st = bed_parse_one<DIALECT>(buf + (p - a0), len, T, &c, &s, &e);
    else st = bed_parse_one<DIALECT>(text + p, len, T, &c, &s, &e);
    status[k] = st;
    oc[k] = c; os[k] = s; oe[k] = e;
    if (st == BED_HOST) flag_list[atomicAdd(n_flag, 1u)] = (uint32_t)k;
}

struct BedKeepLoad {
    const uint8_t* status;
    __device__ uint32_t operator()(uint64_t i) const { return status[i] == BED_KEEP ? 1u : 0u; }
};

static __global__ void __launch_bounds__(BED_THREADS)
    bed_patch_kernel(const uint32_t* __restrict__ line, const uint8_t* __restrict__ st, const uint32_t* __restrict__ c, const uint32_t* __restrict__ s,
                     const uint32_t* __restrict__ e, uint32_t m, uint8_t* __restrict__ status, uint32_t* __restrict__ oc, uint32_t* __restrict__ os,
                     uint32_t* __restrict__ oe) {
    const uint32_t i = blockIdx.x * BED_THREADS + threadIdx.x;
    if (i >= m) return;
    const uint32_t k = line[i];
    status[k] = st[i]; oc[k] = c[i]; os[k] = s[i]; oe[k] = e[i];
}

static __global__ void __launch_bounds__(BED_THREADS)
    bed_compact_kernel(const uint8_t* __restrict__ status, const uint32_t* __restrict__ pos, const uint32_t* __restrict__ c, const uint32_t* __restrict__ s,
                       const uint32_t* __restrict__ e, uint64_t n_lines, uint32_t* __restrict__ oc, uint32_t* __restrict__ os, uint32_t* __restrict__ oe) {
    const uint64_t k = (uint64_t)blockIdx.x * BED_THREADS + threadIdx.x;
    if (k >= n_lines || status[k] != BED_KEEP) return;
    const uint32_t o = pos[k];
    oc[o] = c[k]; os[o] = s[k]; oe[o] = e[k];
}

}  // namespace gffx
