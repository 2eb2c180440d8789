// cg_ptx.cuh -- thin inline-PTX wrappers for the sm_100a async machinery used here:
// mbarrier (transaction barriers) and the TMA bulk-copy engine (cp.async.bulk).
#pragma once
#include <stdint.h>

namespace cg {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

// make barrier inits visible to the async proxy (TMA) before first use
__device__ __forceinline__ void mbar_init_fence() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}

// TMA bulk copy global -> shared, completion counted in bytes on `bar`.
// dst/src 16-byte aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes,
                                         uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// The same with 32-bit shared-memory addresses the caller keeps in registers.
__device__ __forceinline__ void mbar_arrive_expect_tx_s(uint32_t bar_s, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_wait_s(uint32_t bar_s, uint32_t parity) {
    // the loop lives in the PTX block (labels are local to it): test, leave, or try again
    asm volatile(
        "{\n\t.reg .pred p;\n"
        "CG_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra CG_DONE;\n\t"
        "bra CG_WAIT;\n"
        "CG_DONE:\n\t}"
        ::"r"(bar_s), "r"(parity)
        : "memory");
}

__device__ __forceinline__ void bulk_g2s_s(uint32_t dst_s, const void* gmem_src, uint32_t bytes, uint32_t bar_s) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s),
        "l"(gmem_src), "r"(bytes), "r"(bar_s)
        : "memory");
}

}  // namespace cg
