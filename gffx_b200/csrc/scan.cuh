// scan.cuh — device-wide scans used by the join (tile offsets), the rank-space post-passes
// (prefix-max, suffix-min, next-set) and the order-preserving compactions.
//
// Scheme: reduce-then-scan (tile reduce -> scan of tile aggregates, recursive above one tile of
// aggregates -> tile rescan with carry-in); up to SCAN_FOLD_TILES tiles the middle step is folded
// into the rescan (two launches).  Warp shuffles inside a warp, one
// shared-memory pass across the warps of a CTA.  The reduce-then-scan kernels have no spin-waits between CTAs; the
// single-pass form (scan_chain_kernel, r03h) waits only for CTAs of lower index, which the hardware has started earlier.  Every array these run over is small next to the query stream (<= a few
// hundred MB); the streaming kernels do their own block scans in registers/shared memory.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>

namespace gffx {

constexpr int SCAN_THREADS = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;
constexpr uint64_t SCAN_FOLD_TILES = 8192;   // up to here (16.8 M elements) the rescan kernel folds the tile totals itself

struct OpSum {
    template <class T> __host__ __device__ T operator()(T a, T b) const { return a + b; }
};
struct OpMax {
    template <class T> __host__ __device__ T operator()(T a, T b) const { return a > b ? a : b; }
};
struct OpMin {
    template <class T> __host__ __device__ T operator()(T a, T b) const { return a < b ? a : b; }
};

template <class T> __device__ __forceinline__ T shfl_up_any(T v, int d) {
    if constexpr (sizeof(T) == 8) {
        unsigned long long x = (unsigned long long)v;
        x = __shfl_up_sync(0xffffffffu, x, d);
        return (T)x;
    } else {
        return (T)__shfl_up_sync(0xffffffffu, v, d);
    }
}
template <class T> __device__ __forceinline__ T shfl_idx_any(T v, int l) {
    if constexpr (sizeof(T) == 8) {
        unsigned long long x = (unsigned long long)v;
        x = __shfl_sync(0xffffffffu, x, l);
        return (T)x;
    } else {
        return (T)__shfl_sync(0xffffffffu, v, l);
    }
}

// inclusive scan across the warp
template <class T, class Op> __device__ __forceinline__ T warp_scan_incl(T v, Op op) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = shfl_up_any(v, d);
        if (lane >= d) v = op(o, v);
    }
    return v;
}

// CTA-wide exclusive scan of one value per thread; returns the exclusive prefix and writes the CTA total.
// `smem` must hold blockDim.x/32 + 1 elements of T.  Contains two __syncthreads().
template <class T, class Op> __device__ __forceinline__ T block_scan_excl(T v, Op op, T identity, T* smem, T* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    T incl = warp_scan_incl(v, op);
    if (lane == 31) smem[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        T w = lane < nwarps ? smem[lane] : identity;
        T wi = warp_scan_incl(w, op);
        if (lane < nwarps) smem[lane] = wi;
        if (lane == nwarps - 1) smem[nwarps] = wi;
    }
    __syncthreads();
    T excl = shfl_up_any(incl, 1);
    if (lane == 0) excl = identity;
    if (warp > 0) excl = op(smem[warp - 1], excl);
    *total = smem[nwarps];
    return excl;
}

// ---- device-wide scan -----------------------------------------------------------------------
// Load(i) -> T gives element i; REVERSE scans from the high end (suffix scan).  Element j of the
// scan order is array index (REVERSE ? n-1-j : j).
template <class T, class Op, class Load, bool REVERSE, int ITEMS = SCAN_ITEMS>
__global__ void __launch_bounds__(SCAN_THREADS) scan_reduce_kernel(Load load, uint64_t n, Op op, T identity, T* __restrict__ agg,
                                                                   const uint32_t* __restrict__ gate) {
    __shared__ T sm[SCAN_THREADS / 32 + 1];
    if (gate && *gate == 0) return;   // the whole scan is switched off by a device-side flag (uniform)
    const uint64_t base = (uint64_t)blockIdx.x * (SCAN_THREADS * ITEMS);
    T acc = identity;
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        uint64_t j = base + (uint64_t)k * SCAN_THREADS + threadIdx.x;
        if (j < n) acc = op(acc, load(REVERSE ? n - 1 - j : j));
    }
    // order inside the tile does not matter for a reduction (ops are commutative)
    T tot;
    block_scan_excl(acc, op, identity, sm, &tot);
    if (threadIdx.x == 0) agg[blockIdx.x] = tot;
}

// single CTA: in-place exclusive scan of `agg[0..m)`; writes the grand total to agg[m]
template <class T, class Op>
__global__ void __launch_bounds__(1024) scan_agg_kernel(T* __restrict__ agg, uint32_t m, Op op, T identity, const uint32_t* __restrict__ gate) {
    __shared__ T sm[1024 / 32 + 1];
    if (gate && *gate == 0) return;
    T carry = identity;
    for (uint32_t base = 0; base < m; base += 1024) {
        uint32_t i = base + threadIdx.x;
        T v = i < m ? agg[i] : identity;
        T tot;
        T ex = block_scan_excl(v, op, identity, sm, &tot);
        if (i < m) agg[i] = op(carry, ex);
        carry = op(carry, tot);
        __syncthreads();
    }
    if (threadIdx.x == 0) agg[m] = carry;
}

template <class T, class Op, class Load, class Store, bool REVERSE, bool INCLUSIVE, int ITEMS = SCAN_ITEMS>
__global__ void __launch_bounds__(SCAN_THREADS)
    scan_apply_kernel(Load load, Store store, uint64_t n, Op op, T identity, const T* __restrict__ agg, const uint32_t* __restrict__ gate,
                      bool fold, T* __restrict__ total_out) {
    __shared__ T sm[SCAN_THREADS / 32 + 1];
    if (gate && *gate == 0) return;
    // fold: agg holds the tiles' raw totals and this CTA combines its predecessors' itself (a few loads per thread while the
    // tile count is small) - the single-CTA pass over the aggregates and its launch are skipped
    T carry = identity;
    if (fold) {
        for (uint32_t j = threadIdx.x; j < blockIdx.x; j += SCAN_THREADS) carry = op(carry, agg[j]);
        T ctot;
        block_scan_excl(carry, op, identity, sm, &ctot);
        carry = ctot;
        __syncthreads();   // sm is used again below
    } else {
        carry = agg[blockIdx.x];
    }
    // striped arrangement: element k * 32 + lane of the warp's 256 -> every load and store of a warp is one contiguous run
    // (a blocked arrangement, 64 bytes per thread, costs four times the L1 wavefronts); the price is one warp scan per k
    // instead of one per thread, a few dozen shuffles that nothing here waits for
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint64_t wbase = (uint64_t)blockIdx.x * (SCAN_THREADS * ITEMS) + (uint64_t)warp * (32 * ITEMS) + lane;
    T v[ITEMS];
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        const uint64_t j = wbase + 32u * k;
        v[k] = j < n ? load(REVERSE ? n - 1 - j : j) : identity;
    }
    T run = identity;   // this warp's elements before row k, combined
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        const T inc = warp_scan_incl(v[k], op);
        const T row = shfl_idx_any(inc, 31);
        T ex = shfl_up_any(inc, 1);
        if (lane == 0) ex = identity;
        v[k] = op(run, INCLUSIVE ? inc : ex);   // prefix inside the warp's 256 elements
        run = op(run, row);
    }
    // exclusive scan of the warps' totals
    if (lane == 0) sm[warp] = run;
    __syncthreads();
    T wpre = identity, tot = identity;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) {
        const T x = sm[w];
        if (w < (int)warp) wpre = op(wpre, x);
        tot = op(tot, x);
    }
    const T pre = op(carry, wpre);
    if (fold && total_out && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0) *total_out = op(carry, tot);
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        const uint64_t j = wbase + 32u * k;
        if (j < n) store(REVERSE ? n - 1 - j : j, op(pre, v[k]));
    }
}

// ---- single-pass scan (r03h): a tile's carry-in comes from its predecessors' published totals ("decoupled look-back") ----------
// One launch instead of reduce + rescan, the input read once: a CTA scans its tile, publishes the tile's total, and its first warp
// walks back over the predecessors' records — 32 at a time — adding totals until it meets a tile whose INCLUSIVE prefix is already
// known; then it publishes its own inclusive prefix.  A record is (flag, total, inclusive); the flag carries the launch's epoch, so
// records of earlier launches read as "not there yet" and nothing is cleared between launches.  A CTA waits only for CTAs with a
// lower blockIdx, which were dispatched before it: they are resident or finished, so the wait ends.
struct ScanChain {            // host-side handle: one per stream of scans (records of two scans in flight must not mix)
    uint32_t* flags = nullptr;   // cap_tiles words, zero when allocated
    uint32_t cap_tiles = 0;
    uint32_t epoch = 0;          // bumped per launch; flag = epoch << 2 | state (1: total published, 2: inclusive prefix published)
};
constexpr uint32_t SCAN_CHAIN_TILES = 1u << 16;   // records per chain (134 M elements); larger scans take the two-pass form

template <class T> __device__ __forceinline__ T ld_volatile_any(const T* p) { return *reinterpret_cast<const volatile T*>(p); }

template <class T, class Op, class Load, class Store, bool REVERSE, bool INCLUSIVE>
__global__ void __launch_bounds__(SCAN_THREADS)
    scan_chain_kernel(Load load, Store store, uint64_t n, Op op, T identity, uint32_t* __restrict__ flags, T* __restrict__ agg, T* __restrict__ incl,
                      uint32_t epoch, const uint32_t* __restrict__ gate, T* __restrict__ total_out) {
    __shared__ T sm[SCAN_THREADS / 32 + 1];
    __shared__ T s_carry;
    if (gate && *gate == 0) return;
    const uint32_t tile = blockIdx.x, lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint64_t wbase = (uint64_t)tile * SCAN_TILE + (uint64_t)warp * (32 * SCAN_ITEMS) + lane;
    T v[SCAN_ITEMS];
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const uint64_t j = wbase + 32u * k;
        v[k] = j < n ? load(REVERSE ? n - 1 - j : j) : identity;
    }
    T run = identity;   // this warp's elements before row k, combined (striped arrangement: see scan_apply_kernel)
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const T inc = warp_scan_incl(v[k], op);
        const T row = shfl_idx_any(inc, 31);
        T ex = shfl_up_any(inc, 1);
        if (lane == 0) ex = identity;
        v[k] = op(run, INCLUSIVE ? inc : ex);
        run = op(run, row);
    }
    if (lane == 0) sm[warp] = run;
    __syncthreads();
    T wpre = identity, tot = identity;
#pragma unroll
    for (int w = 0; w < SCAN_THREADS / 32; w++) {
        const T x = sm[w];
        if (w < (int)warp) wpre = op(wpre, x);
        tot = op(tot, x);
    }
    if (warp == 0) {
        T carry = identity;
        if (tile == 0) {
            if (lane == 0) {
                incl[0] = tot;
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(flags) = (epoch << 2) | 2u;
            }
        } else {
            if (lane == 0) {
                agg[tile] = tot;
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(flags + tile) = (epoch << 2) | 1u;
            }
            long long j = (long long)tile - 1;
            for (;;) {
                const long long idx = j - (long long)lane;
                uint32_t f = (epoch << 2) | 2u;   // before the first tile: an inclusive prefix of nothing
                if (idx >= 0) {
                    do { f = ld_volatile_any(flags + idx); } while ((f >> 2) != epoch);
                }
                __threadfence();
                const uint32_t have = __ballot_sync(0xffffffffu, (f & 3u) == 2u);
                const uint32_t first = have ? (uint32_t)__ffs((int)have) - 1u : 32u;   // the nearest predecessor with a known prefix
                T x = identity;
                if (idx >= 0 && lane <= first) x = lane == first ? ld_volatile_any(incl + idx) : ld_volatile_any(agg + idx);
#pragma unroll
                for (int d = 16; d >= 1; d >>= 1) {
                    T o;
                    if constexpr (sizeof(T) == 8) o = (T)__shfl_xor_sync(0xffffffffu, (unsigned long long)x, d);
                    else o = (T)__shfl_xor_sync(0xffffffffu, x, d);
                    x = op(x, o);
                }
                carry = op(carry, x);
                if (have) break;
                j -= 32;
            }
            if (lane == 0) {
                incl[tile] = op(carry, tot);
                __threadfence();
                *reinterpret_cast<volatile uint32_t*>(flags + tile) = (epoch << 2) | 2u;
            }
        }
        if (lane == 0) {
            s_carry = carry;
            if (total_out && tile == gridDim.x - 1) *total_out = op(carry, tot);
        }
    }
    __syncthreads();
    const T pre = op(s_carry, wpre);
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; k++) {
        const uint64_t j = wbase + 32u * k;
        if (j < n) store(REVERSE ? n - 1 - j : j, op(pre, v[k]));
    }
}

struct ScanScratch {  // caller-provided aggregate storage: needs scan_scratch_elems(n) elements of T
    void* ptr;
};
inline uint64_t scan_scratch_elems(uint64_t n) {
    uint64_t tot = 0;
    while (true) {
        uint64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
        if (tiles == 0) tiles = 1;
        tot += tiles + 1;
        if (tiles <= 1024 * 64) break;  // one CTA finishes the rest
        n = tiles;
    }
    return 2 * tot + 16;   // the single-pass form keeps two values per tile (total, inclusive prefix)
}

template <class T> struct PtrLoad {
    const T* p;
    __device__ T operator()(uint64_t i) const { return p[i]; }
};
template <class T> struct PtrStore {
    T* p;
    __device__ void operator()(uint64_t i, T v) const { p[i] = v; }
};

// Scans n elements; `scratch` holds scan_scratch_elems(n) Ts.  If total_out != nullptr the grand
// total (device pointer) is written there.  gate != nullptr: a device-side word; when it is 0 the kernels return at once
// and nothing is written (the consumer checks the same word).  Returns the number of kernels launched.
// ITEMS: elements per thread of the reduce / rescan kernels (16 for the big rank-space scans: half the CTAs, -7 % of their time)
template <class T, class Op, class Load, class Store, bool REVERSE, bool INCLUSIVE, int ITEMS = SCAN_ITEMS>
int device_scan(Load load, Store store, uint64_t n, Op op, T identity, T* scratch, T* total_out, cudaStream_t st, const uint32_t* gate = nullptr,
                ScanChain* chain = nullptr, uint64_t chain_from_tiles = 8192) {   // chain_from_tiles: lower it when a tile's load is heavy (a long tile hides the look-back)
    int launches = 0;
    uint64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
    if (tiles == 0) tiles = 1;
    // One launch where it pays (measured, r03h): a scan of a few dozen tiles (the carry arrives within two look-back steps; the
    // join's offsets scan: -5 us) and a scan of many waves of tiles (the look-back hides behind the following tiles; coverage's
    // 48 M-word scans: -45 us each).  A scan of one or two waves — the rank-space scans, 1300-2600 tiles all resident at once — waits
    // for the prefix to travel 32 tiles per step through the whole grid and is slower than reduce + rescan (56 against 35 us).
    if (chain && chain->flags && tiles > 1 && tiles <= chain->cap_tiles && (tiles <= 64 || tiles >= chain_from_tiles)) {
        if (++chain->epoch >= (1u << 30)) {   // the epoch is about to run out of bits: start over from cleared records
            cudaMemsetAsync(chain->flags, 0, (size_t)chain->cap_tiles * 4, st);
            chain->epoch = 1;
        }
        scan_chain_kernel<T, Op, Load, Store, REVERSE, INCLUSIVE><<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(load, store, n, op, identity, chain->flags, scratch,
                                                                                                            scratch + tiles + 1, chain->epoch, gate, total_out);
        return 1;
    }
    if (ITEMS != SCAN_ITEMS) tiles = std::max<uint64_t>(1, (n + (uint64_t)SCAN_THREADS * ITEMS - 1) / ((uint64_t)SCAN_THREADS * ITEMS));
    T* agg = scratch;
    scan_reduce_kernel<T, Op, Load, REVERSE, ITEMS><<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(load, n, op, identity, agg, gate);
    launches++;
    if (tiles <= SCAN_FOLD_TILES) {   // two launches: the rescan folds the raw tile totals itself and writes the grand total
        scan_apply_kernel<T, Op, Load, Store, REVERSE, INCLUSIVE, ITEMS><<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(load, store, n, op, identity, agg, gate,
                                                                                                            true, total_out);
        return launches + 1;
    }
    if (tiles <= 1024 * 64) {
        scan_agg_kernel<T, Op><<<1, 1024, 0, st>>>(agg, (uint32_t)tiles, op, identity, gate);
        launches++;
    } else {
        // scan the aggregates with the same machinery (exclusive, forward), in place
        T* next = agg + tiles + 1;
        launches += device_scan<T, Op, PtrLoad<T>, PtrStore<T>, false, false>(PtrLoad<T>{agg}, PtrStore<T>{agg}, tiles, op, identity,
                                                                             next, agg + tiles, st, gate);
    }
    scan_apply_kernel<T, Op, Load, Store, REVERSE, INCLUSIVE, ITEMS><<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(load, store, n, op, identity, agg, gate,
                                                                                                        false, nullptr);
    launches++;
    if (total_out) {
        cudaMemcpyAsync(total_out, agg + tiles, sizeof(T), cudaMemcpyDeviceToDevice, st);
    }
    return launches;
}

}  // namespace gffx
