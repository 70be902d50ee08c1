// bulkcopy.cuh — the bulk asynchronous copy (cp.async.bulk, the TMA's 1-D form) and its mbarrier, as used by the kernels that
// stage a contiguous piece of global memory in shared memory while their warps do something else (rootjoin.cuh: the window's
// rank slice; counts.cuh: the same for the counts coordinate set; partition.cuh: the next tile of regions).
//   thread 0:   mbar_init(&bar, 1) once, then per copy group  mbar_expect_tx(&bar, bytes); bulk_g2s(dst, src, bytes, &bar) ...
//   everybody:  mbar_wait(&bar, phase); phase ^= 1
// A buffer that generic loads have read is handed back to the copy engine behind `fence.proxy.async.shared::cta`.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace gffx {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {   // 16-byte aligned, bytes % 16 == 0
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
                 "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok = 0;
    while (!ok)
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace gffx
