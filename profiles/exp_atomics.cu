// Micro-benchmark behind DESIGN.md §4: how fast can a B200 apply scattered L2 updates of the kinds the kernels issue?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o profiles/exp_atomics profiles/exp_atomics.cu && profiles/exp_atomics
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint64_t mix(uint64_t x) { x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31); }
// KIND 0: atomicMax u64; 1: atomicAdd u32; 2: byte store; 3: atomicMax u64 + byte store; 4: plain 8-byte load (sum)
template <int KIND> __global__ void k(unsigned long long* a, uint32_t* h, uint8_t* m, uint64_t n, uint64_t slots, uint32_t window, unsigned long long* sink) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // regions of a tile fall into one window of `window` slots (as after the partition pass), windows taken in tile order
    const uint32_t tile = (uint32_t)(i >> 12), w0 = ((uint32_t)(mix(tile) >> 40) % (uint32_t)(slots / window)) * window;   // window is a power of two
    const uint64_t s = w0 + ((uint32_t)mix(i) & (window - 1));
    if (KIND == 0 || KIND == 3) atomicMax(a + s, (unsigned long long)i);
    if (KIND == 1) atomicAdd(h + s, 1u);
    if (KIND == 2 || KIND == 3) m[2 * s] = 1;
    if (KIND == 4) { unsigned long long v = a[s]; if (v == 0xdeadbeefull) *sink = v; }
}
template <int KIND> float run(unsigned long long* a, uint32_t* h, uint8_t* m, uint64_t n, uint64_t slots, uint32_t window, unsigned long long* sink) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    for (int w = 0; w < 2; w++) k<KIND><<<(unsigned)((n + 255) / 256), 256>>>(a, h, m, n, slots, window, sink);
    cudaEventRecord(e0);
    for (int r = 0; r < 5; r++) k<KIND><<<(unsigned)((n + 255) / 256), 256>>>(a, h, m, n, slots, window, sink);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms / 5;
}
int main() {
    const uint64_t slots = 1u << 22, n = 10000000;   // 4 M slots (the bench index has 2.7 M)
    unsigned long long *a, *sink; uint32_t* h; uint8_t* m;
    cudaMalloc(&a, slots * 8); cudaMalloc(&h, slots * 4); cudaMalloc(&m, slots * 2); cudaMalloc(&sink, 8);
    cudaMemset(a, 0, slots * 8); cudaMemset(h, 0, slots * 4); cudaMemset(m, 0, slots * 2);
    for (uint32_t window : {8192u, 1u << 22}) {
        printf("10 M updates over 4 M slots, %u slots per tile's window:\n", window);
        printf("  atomicMax u64          %.1f us\n", run<0>(a, h, m, n, slots, window, sink) * 1e3);
        printf("  atomicAdd u32          %.1f us\n", run<1>(a, h, m, n, slots, window, sink) * 1e3);
        printf("  byte store             %.1f us\n", run<2>(a, h, m, n, slots, window, sink) * 1e3);
        printf("  atomicMax u64 + byte   %.1f us\n", run<3>(a, h, m, n, slots, window, sink) * 1e3);
        printf("  8-byte load            %.1f us\n", run<4>(a, h, m, n, slots, window, sink) * 1e3);
    }
    return 0;
}
