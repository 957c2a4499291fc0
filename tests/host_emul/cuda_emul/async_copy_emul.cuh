// async_copy_emul.cuh — CPU stand-ins for csrc/async_copy.cuh (mbarrier + bulk copies), test infrastructure.
// Included by the product header only when the emulated library is built (-DGSR_HOST_EMUL, tests/host_emul on the
// include path); the product build never sees this file.
#pragma once
#include <stdint.h>
#include <string.h>

namespace gsr {

// CPU emulation (tests/host_emul): shared memory is static storage; a "shared address" is the 32-bit offset from
// an anchor in the same image, bulk copies complete synchronously at issue, mbarrier waits are therefore no-ops.
inline char g_smem_anchor;
inline uint32_t smem_u32(const void* p) {
    return (uint32_t)(int32_t)(reinterpret_cast<intptr_t>(p) - reinterpret_cast<intptr_t>(&g_smem_anchor));
}
inline char* smem_ptr(uint32_t h) {
    return reinterpret_cast<char*>(reinterpret_cast<intptr_t>(&g_smem_anchor) + (intptr_t)(int32_t)h);
}
// mbarrier state kept in the 8 bytes of the barrier object: a phase completes when all expected arrivals have
// happened and the transaction count is back to zero (it may be transiently negative, as on the hardware).
struct EmulMbar { int32_t tx; int16_t pending; uint8_t count; uint8_t phase; };
static_assert(sizeof(EmulMbar) == 8, "fits the mbarrier object");
inline EmulMbar* mbar_obj(uint32_t bar) { return reinterpret_cast<EmulMbar*>(smem_ptr(bar)); }
inline void mbar_check(EmulMbar* m) {
    if (m->pending == 0 && m->tx == 0) { m->phase ^= 1; m->pending = m->count; }
}
inline void mbar_init(uint32_t bar, uint32_t arrivals) { *mbar_obj(bar) = EmulMbar{0, (int16_t)arrivals, (uint8_t)arrivals, 0}; }
inline void mbar_fence_init() {}
inline void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    EmulMbar* m = mbar_obj(bar);
    m->tx += (int32_t)bytes;
    m->pending -= 1;
    mbar_check(m);
}
inline bool mbar_try_wait(uint32_t bar, uint32_t parity) { return mbar_obj(bar)->phase != (uint8_t)parity; }
inline void mbar_wait(uint32_t bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) emu::spin();      // let the fibers that issue the copies run
}
inline void bulk_g2s(uint32_t dst_smem, const void* src, uint32_t bytes, uint32_t bar) {
    memcpy(smem_ptr(dst_smem), src, bytes);
    EmulMbar* m = mbar_obj(bar);
    m->tx -= (int32_t)bytes;
    mbar_check(m);
}
inline void bulk_s2g(void* dst, uint32_t src_smem, uint32_t bytes) { memcpy(dst, smem_ptr(src_smem), bytes); }
inline void bulk_commit() {}
inline void bulk_wait_read0() {}
inline void fence_proxy_async_smem() {}

}  // namespace gsr
