// Global -> shared staging with the TMA bulk-copy engine (cp.async.bulk + mbarrier complete_tx; SASS: UBLKCP).
// One thread issues a copy of a whole tile; every thread then waits on the tile's mbarrier.  Phases are counted by the
// caller (`n_done` = how many loads were already completed on that barrier) so that the host emulator can implement
// the same calls with a monotonically increasing counter.
// Requirements of cp.async.bulk: 16-byte aligned source and destination, size a multiple of 16 bytes.
#pragma once
#include "gqp_common.cuh"

#ifndef GQP_EMU
__device__ __forceinline__ uint32_t gqp_smem_addr(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(gqp_smem_addr(bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
// issue by ONE thread: arm the barrier with the byte count, then start the bulk copy
__device__ __forceinline__ void bulk_load(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    uint32_t b = gqp_smem_addr(bar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     gqp_smem_addr(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(b)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t n_done) {
    uint32_t b = gqp_smem_addr(bar), parity = n_done & 1u, ok;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok)
            : "r"(b), "r"(parity)
            : "memory");
    } while (!ok);
}
#else
inline void mbar_init(uint64_t *bar) { std::atomic_ref<uint64_t>(*bar).store(0); }
inline void bulk_load(void *smem_dst, const void *gsrc, uint32_t bytes, uint64_t *bar) {
    std::memcpy(smem_dst, gsrc, bytes);
    std::atomic_ref<uint64_t>(*bar).fetch_add(1, std::memory_order_release);
}
inline void mbar_wait(uint64_t *bar, uint32_t n_done) {
    while (std::atomic_ref<uint64_t>(*bar).load(std::memory_order_acquire) <= n_done) std::this_thread::yield();
}
#endif
