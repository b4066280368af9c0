// lr_tma.cuh -- bulk asynchronous copy global -> shared through the TMA unit (cp.async.bulk, SASS UBLKCP) completed on an mbarrier:
// one thread programs the copy, the copy engine moves the bytes, every thread of the block waits on the barrier's phase.
// Used to stage the majorant super-grid of the DDA walks in shared memory (lr_wavefront.cuh, k_wv_track_dda).
#pragma once
#include <cstdint>

namespace lr {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

// `bar`: 8-byte aligned shared word.  Call from ONE thread, then __syncthreads() before anybody arrives / waits.
__device__ __forceinline__ void tma_bar_init(uint64_t *bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// one thread: announce `bytes` of transactions and arrive
__device__ __forceinline__ void tma_bar_expect(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// one thread: copy `bytes` (multiple of 16; src and dst 16-byte aligned) from global to shared; completion is counted on `bar`
__device__ __forceinline__ void tma_load_1d(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// every thread: wait until the phase with parity `phase` of `bar` has completed
__device__ __forceinline__ void tma_bar_wait(uint64_t *bar, uint32_t phase) {
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(smem_u32(bar)), "r"(phase) : "memory");
    } while (!done);
}

} // namespace lr
