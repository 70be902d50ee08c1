// gather.cuh — K8: record emission on the device.
//
// The reference writes the selected records slice by slice from the mmap'd GFF: write_gff_match_only_by_coords copies
// every emitted (abs_start, abs_end) line range (src/commands/intersect.rs:385-429), write_gff_output the merged block
// ranges (src/utils/common.rs:231-278).  Here the GFF text is resident in HBM and the ranges of a finished intersect
// are still on the device, so the output text is gathered there: an exclusive scan of the range lengths gives every
// range its place in the output, and gather_text_kernel copies OUTPUT-centric — a warp owns a chunk of output bytes,
// finds its first range by binary search over the scanned offsets and walks forward, fusing source-contiguous ranges
// (consecutive lines of a block) into one run — so one 400 MB range and a million 150-byte ranges load the grid the
// same way.  Algorithmic bytes: 2 x bytes_out (every output byte read once, written once).
//
// Copy primitive: destination words are 4-byte aligned stores of consecutive lanes (128 B per warp instruction); the
// source is read as aligned words too and realigned with a funnel shift, the upper word coming from the next lane's
// load (one extra load for lane 31 only).  The text buffer is padded by 8 bytes so the look-ahead word exists.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gffx {

constexpr int GATHER_THREADS = 256;
constexpr uint32_t GATHER_CHUNK = 16384;   // output bytes per warp and step

// length of range k as the host writer takes it (host_index.cpp: write_ranges_to skips s >= e and e > text_len)
struct RangeLenLoad {
    const ulonglong2* ranges;
    uint64_t n, text_len;
    __device__ unsigned long long operator()(uint64_t i) const {
        if (i >= n) return 0ull;
        const ulonglong2 r = ranges[i];
        return (r.x < r.y && r.y <= text_len) ? r.y - r.x : 0ull;
    }
};

// warp-cooperative copy of nbytes from src to dst (any alignment), all 32 lanes call it with the same arguments
__device__ __forceinline__ void warp_copy_bytes(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, uint64_t nbytes, uint32_t lane) {
    // head: up to the first 4-byte boundary of dst
    const uint32_t head = (uint32_t)min((uint64_t)((4u - (uint32_t)(reinterpret_cast<uintptr_t>(dst) & 3u)) & 3u), nbytes);
    if (lane < head) dst[lane] = __ldg(src + lane);
    dst += head; src += head; nbytes -= head;
    const uint64_t nwords = nbytes >> 2;
    const uint32_t mis = (uint32_t)(reinterpret_cast<uintptr_t>(src) & 3u);
    const uint32_t* sw = reinterpret_cast<const uint32_t*>(src - mis);   // aligned word holding src[0]
    uint32_t* dw = reinterpret_cast<uint32_t*>(dst);
    constexpr int U = 8;   // independent loads in flight per lane: 1 KB per warp and step
    for (uint64_t w0 = 0; w0 < nwords; w0 += 32 * U) {
        uint32_t lo[U], hi[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint64_t w = w0 + 32u * u + lane;
            lo[u] = w < nwords ? __ldg(sw + w) : 0u;
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint64_t w = w0 + 32u * u + lane;
            hi[u] = __shfl_down_sync(0xffffffffu, lo[u], 1);
            // lane 31's upper word is lane 0's word of the next group (or, at the end of the step or the run, one more load)
            if (u + 1 < U) {
                const uint32_t nxt = __shfl_sync(0xffffffffu, lo[u + 1 < U ? u + 1 : u], 0);
                if (lane == 31) hi[u] = nxt;
            }
            if (mis && w < nwords && (w + 1 == nwords || (lane == 31 && u + 1 == U))) hi[u] = __ldg(sw + w + 1);
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint64_t w = w0 + 32u * u + lane;
            if (w < nwords) __stcs(dw + w, __funnelshift_r(lo[u], hi[u], mis * 8u));
        }
    }
    // tail: the last nbytes & 3 bytes
    const uint32_t tail = (uint32_t)(nbytes & 3u);
    if (lane < tail) dst[4 * nwords + lane] = __ldg(src + 4 * nwords + lane);
}

// off: n_ranges + 1 ascending output offsets (off[n_ranges] = bytes_out)
static __global__ void __launch_bounds__(GATHER_THREADS)
    gather_text_kernel(const uint8_t* __restrict__ text, const ulonglong2* __restrict__ ranges, const unsigned long long* __restrict__ off,
                       uint64_t n_ranges, uint8_t* __restrict__ out) {
    const unsigned long long total = off[n_ranges];
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t n_warps = (uint64_t)gridDim.x * (GATHER_THREADS / 32);
    for (uint64_t chunk = (uint64_t)blockIdx.x * (GATHER_THREADS / 32) + (threadIdx.x >> 5); chunk * GATHER_CHUNK < total; chunk += n_warps) {
        const unsigned long long c0 = chunk * GATHER_CHUNK, c1 = min(c0 + GATHER_CHUNK, total);
        uint64_t lo = 0, hi = n_ranges;   // off[lo] <= c0 < off[hi]
        while (hi - lo > 1) {
            const uint64_t mid = lo + ((hi - lo) >> 1);
            if (off[mid] <= c0) lo = mid;
            else hi = mid;
        }
        uint64_t k = lo;   // off[k] <= c0 < off[k + 1]: range k is not empty
        unsigned long long o = c0;   // next output byte
        while (o < c1) {
            const unsigned long long ok = off[k];
            const ulonglong2 rk = ranges[k];
            const unsigned long long src0 = rk.x + (o - ok);   // source byte of output byte o
            // extend over the following ranges while they continue the source (and hold bytes): lanes look at k+1+lane
            unsigned long long run_end_out = min((unsigned long long)off[k + 1], c1);
            unsigned long long src_next = rk.y;   // source position the next range must start at
            uint64_t kk = k + 1;
            while (run_end_out < c1) {
                const uint64_t j = kk + lane;
                bool cont = false;
                unsigned long long oj1 = 0, ej = 0;
                if (j < n_ranges) {
                    const ulonglong2 rj = ranges[j];
                    oj1 = off[j + 1];
                    ej = rj.y;
                    // contiguous with ITS predecessor: its start equals the previous range's end; empty ranges (no bytes) break a run
                    const unsigned long long prev_end = j == kk ? src_next : ranges[j - 1].y;
                    cont = oj1 > off[j] && rj.x == prev_end;
                }
                const uint32_t bal = __ballot_sync(0xffffffffu, cont);
                const uint32_t ncont = bal == 0xffffffffu ? 32u : (uint32_t)__ffs(~bal) - 1u;   // leading ranges that continue the run
                if (ncont) {
                    run_end_out = min((unsigned long long)__shfl_sync(0xffffffffu, oj1, ncont - 1), c1);
                    src_next = __shfl_sync(0xffffffffu, ej, ncont - 1);
                    kk += ncont;
                }
                if (ncont < 32u) break;
            }
            warp_copy_bytes(out + o, text + src0, run_end_out - o, lane);
            o = run_end_out;
            if (o >= c1) break;
            // next range with bytes: the first j >= kk with off[j + 1] > o
            k = kk;
            while (true) {
                const uint64_t j = k + lane;
                const bool has = j < n_ranges && off[j + 1] > o;
                const uint32_t bal = __ballot_sync(0xffffffffu, has);
                if (bal) { k += (uint32_t)__ffs(bal) - 1u; break; }
                k += 32;
            }
        }
    }
}

}  // namespace gffx
