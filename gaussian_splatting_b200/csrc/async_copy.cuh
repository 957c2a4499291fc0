// async_copy.cuh — mbarrier + bulk-copy (TMA engine, `cp.async.bulk`, SASS UBLKCP)
// helpers used by the render kernels to stage per-tile Gaussian batches in shared memory.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gsr {

#ifdef GSR_HOST_EMUL
}  // namespace gsr
#include "async_copy_emul.cuh"      // tests/host_emul/cuda_emul: the CPU emulation's stand-ins (not part of the product)
namespace gsr {
#else

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(arrivals) : "memory");
}

// make the barrier initialisation visible to the async proxy
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
                 : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// A wait that lasts longer than this can only be a lost copy / a lost predecessor: trap (launch failure reported to
// the host) instead of hanging the GPU.  The bound is wall time, not a poll count, so profiler replay, time-slicing
// or a debugger stretching a healthy kernel do not trip it.
constexpr unsigned long long SPIN_LIMIT_NS = 4000000000ull;

__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    unsigned long long t0 = 0;
    for (uint32_t spins = 1; !mbar_try_wait(bar, parity); ++spins) {
        if ((spins & 0xfffu) == 0) {
            const unsigned long long now = global_timer_ns();
            if (t0 == 0) t0 = now;
            else if (now - t0 > SPIN_LIMIT_NS) __trap();
        }
    }
}

// global -> shared bulk copy completing on an mbarrier. bytes % 16 == 0, both
// addresses 16 B aligned.
__device__ __forceinline__ void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(dst_smem), "l"(src), "r"(bytes), "r"(bar)
        : "memory");
}

// shared -> global bulk copy (bulk async-group completion). bytes % 16 == 0, 16 B aligned.
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src_smem, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_smem), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// wait until the bulk stores of this thread have finished READING shared memory
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// order generic-proxy writes to shared memory before async-proxy (bulk copy) reads of it
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
#endif  // GSR_HOST_EMUL

}  // namespace gsr
