// Bulk asynchronous copies (TMA, cp.async.bulk) global -> shared memory with mbarrier completion, sm_90+/sm_100a.
// Every block kernel of the pipeline streams contiguous 16 KB blocks (2048 residues) - a job's limb block, a key row block,
// a block's twiddle table - so one elected thread requests whole blocks one stage ahead and the warps never hold registers
// for operands in flight.
#pragma once
#include "device.cuh"

namespace lmn {

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(u64 *bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// make freshly initialised barriers visible to the async proxy (call by the initialising thread, before a CTA barrier)
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(u64 *bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, u64 *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(u64 *bar, unsigned parity) {
    asm volatile("{\n\t.reg .pred p;\n\tLMN_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra LMN_DONE;\n\tbra LMN_WAIT;\n\tLMN_DONE:\n\t}"
                 ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// Before a CTA barrier after which a bulk copy will overwrite shared memory this thread has read: the reads must be complete
// and ordered before the async proxy's writes.  (Measured in k_mac_rd: without it a copy occasionally lands before a queued
// LDS has executed.)  Loads that nothing has consumed yet must additionally be made a register dependency by the caller.
__device__ __forceinline__ void release_smem_to_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

}  // namespace lmn
