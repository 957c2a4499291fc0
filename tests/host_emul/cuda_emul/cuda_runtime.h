// cuda_runtime.h — stand-in for the CUDA runtime header when the product's .cu files are compiled as plain C++
// by g++ for the CPU emulation (tests/host_emul; -DGSR_HOST_EMUL, this directory first on the include path).
// TEST INFRASTRUCTURE: lets `-m "not gpu"` tests execute the real kernel sources (block scans, warp shuffles,
// shared memory, atomics) in the GPU-less authoring container.  A block's threads run as cooperatively scheduled
// fibers on one OS thread (emul.cpp): __syncthreads() and the *_sync warp primitives are scheduling points, blocks
// run one after another, so `__shared__` maps to `static`.  Nothing here is a performance model.
#pragma once
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <functional>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __launch_bounds__(...)
#define __shared__ static
#define __grid_constant__
#define __restrict__
#define __align__(n) __attribute__((aligned(n)))

struct uint2 { uint32_t x, y; };
struct uint3 { uint32_t x, y, z; };
struct alignas(16) uint4 { uint32_t x, y, z, w; };
struct int2 { int x, y; };
struct alignas(16) ulonglong2 { unsigned long long x, y; };
struct alignas(8) float2 { float x, y; };
struct float3 { float x, y, z; };
struct alignas(16) float4 { float x, y, z, w; };
struct dim3 {
    unsigned x, y, z;
    dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
static inline float2 make_float2(float x, float y) { return {x, y}; }
static inline float3 make_float3(float x, float y, float z) { return {x, y, z}; }
static inline float4 make_float4(float x, float y, float z, float w) { return {x, y, z, w}; }
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return {x, y}; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return {x, y, z, w}; }
static inline int2 make_int2(int x, int y) { return {x, y}; }

typedef void* cudaStream_t;
typedef int cudaError_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice, cudaMemcpyDefault };
static inline const char* cudaGetErrorString(cudaError_t) { return "emulation"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaMemsetAsync(void* p, int v, size_t n, cudaStream_t) { memset(p, v, n); return cudaSuccess; }
static inline cudaError_t cudaMemcpyAsync(void* d, const void* s, size_t n, cudaMemcpyKind, cudaStream_t) {
    memcpy(d, s, n);
    return cudaSuccess;
}

// events and side streams: everything is synchronous, so these only have to exist
typedef void* cudaEvent_t;
enum { cudaEventDisableTiming = 2, cudaStreamNonBlocking = 1 };
enum cudaFuncAttribute { cudaFuncAttributePreferredSharedMemoryCarveout = 9, cudaFuncAttributeMaxDynamicSharedMemorySize = 8 };
enum { cudaSharedmemCarveoutMaxShared = 100 };
static inline cudaError_t cudaEventCreate(cudaEvent_t* e) { *e = (void*)1; return cudaSuccess; }
static inline cudaError_t cudaEventCreateWithFlags(cudaEvent_t* e, unsigned) { *e = (void*)1; return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t = nullptr) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float* ms, cudaEvent_t, cudaEvent_t) { *ms = 0.f; return cudaSuccess; }
static inline cudaError_t cudaStreamWaitEvent(cudaStream_t, cudaEvent_t, unsigned = 0) { return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t* s, unsigned) { *s = (void*)2; return cudaSuccess; }
static inline cudaError_t cudaGetDevice(int* d) { *d = 0; return cudaSuccess; }
// a __device__ array is ordinary static storage here, so its "device address" is its address
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void** p, T& symbol) { *p = (void*)&symbol; return cudaSuccess; }
template <typename F> static inline cudaError_t cudaFuncSetAttribute(F, cudaFuncAttribute, int) { return cudaSuccess; }

namespace emu {
struct Idx { unsigned x, y, z; };
// per-fiber / per-block coordinates, updated by the scheduler on every switch
extern Idx g_threadIdx, g_blockIdx, g_blockDim, g_gridDim;
void launch(dim3 grid, dim3 block, const std::function<void()>& body);
void syncthreads();
int syncthreads_count(int pred);
uint32_t warp_exchange(uint32_t mask, uint32_t value, int src_lane_of_me(int lane, int arg), int arg);
uint32_t warp_shfl(uint32_t mask, uint32_t value, int src_lane);
uint32_t warp_ballot(uint32_t mask, int pred);
void warp_sync(uint32_t mask);
unsigned lane_id();
uint32_t warp_alive();   // lanes of the caller's warp that have not exited
void spin();     // a polling loop's body: give the other fibers of the block a turn
}  // namespace emu

#define threadIdx emu::g_threadIdx
#define blockIdx emu::g_blockIdx
#define blockDim emu::g_blockDim
#define gridDim emu::g_gridDim

static inline void __syncthreads() { emu::syncthreads(); }
static inline int __syncthreads_count(int pred) { return emu::syncthreads_count(pred); }
static inline void __syncwarp(uint32_t mask = 0xffffffffu) { emu::warp_sync(mask); }
static inline uint32_t __ballot_sync(uint32_t mask, int pred) { return emu::warp_ballot(mask, pred); }
static inline int __any_sync(uint32_t mask, int pred) { return emu::warp_ballot(mask, pred) != 0; }
static inline int __all_sync(uint32_t mask, int pred) { return emu::warp_ballot(mask, !pred) == 0; }
static inline uint32_t __activemask() { return 0xffffffffu; }

// shuffle of any trivially copyable type up to 8 bytes, 32 bits at a time
template <typename T> static inline T emu_shfl_from(uint32_t mask, T v, int src_lane) {
    static_assert(sizeof(T) <= 8, "shuffle operand too wide");
    uint32_t w[2] = {0, 0};
    memcpy(w, &v, sizeof(T));
    w[0] = emu::warp_shfl(mask, w[0], src_lane);
    if (sizeof(T) > 4) w[1] = emu::warp_shfl(mask, w[1], src_lane);
    T r;
    memcpy(&r, w, sizeof(T));
    return r;
}
template <typename T> static inline T __shfl_sync(uint32_t mask, T v, int src, int width = 32) {
    const int lane = (int)emu::lane_id();
    return emu_shfl_from(mask, v, (lane & ~(width - 1)) | (src & (width - 1)));
}
template <typename T> static inline T __shfl_xor_sync(uint32_t mask, T v, int delta, int width = 32) {
    (void)width;
    return emu_shfl_from(mask, v, (int)emu::lane_id() ^ delta);
}
// redux.sync: warp-wide integer reductions (all lanes named in the mask take part; exited lanes contribute nothing)
static inline uint32_t __reduce_or_sync(uint32_t mask, uint32_t v) {
    for (int o = 16; o > 0; o >>= 1) v |= emu_shfl_from(mask, v, (int)emu::lane_id() ^ o);
    return v;
}
static inline uint32_t __reduce_max_sync(uint32_t mask, uint32_t v) {
    for (int o = 16; o > 0; o >>= 1) { const uint32_t w = emu_shfl_from(mask, v, (int)emu::lane_id() ^ o); v = w > v ? w : v; }
    return v;
}
template <typename T> static inline T __shfl_up_sync(uint32_t mask, T v, unsigned delta, int width = 32) {
    const int lane = (int)emu::lane_id();
    const int src = lane - (int)delta;
    return emu_shfl_from(mask, v, (src < (lane & ~(width - 1))) ? lane : src);
}
template <typename T> static inline T __shfl_down_sync(uint32_t mask, T v, unsigned delta, int width = 32) {
    const int lane = (int)emu::lane_id();
    const int src = lane + (int)delta;
    return emu_shfl_from(mask, v, (src > (lane | (width - 1))) ? lane : src);
}
// lanes of `mask` holding the same value as the caller
static inline uint32_t __match_any_sync(uint32_t mask, uint32_t v) {
    uint32_t m = 0;
    const uint32_t live = mask & emu::warp_alive();
    for (int l = 0; l < 32; l++) {
        const uint32_t other = emu::warp_shfl(mask, v, l);
        if (((live >> l) & 1u) && other == v) m |= 1u << l;
    }
    return m;
}
template <typename T> static inline uint32_t emu_bits(T v) { uint32_t b; static_assert(sizeof(T) == 4, "32-bit"); memcpy(&b, &v, 4); return b; }
template <typename T> static inline T emu_from_bits(uint32_t b) { T v; memcpy(&v, &b, 4); return v; }

// fibers of one block never run concurrently: plain read-modify-write is atomic
template <typename T> static inline T atomicAdd(T* p, T v) { T o = *p; *p = o + v; return o; }
static inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) { auto o = *p; *p = o + v; return o; }
static inline float atomicAdd(float* p, float v) { float o = *p; *p = o + v; return o; }
template <typename T> static inline T atomicMax(T* p, T v) { T o = *p; if (v > o) *p = v; return o; }
template <typename T> static inline T atomicMin(T* p, T v) { T o = *p; if (v < o) *p = v; return o; }
template <typename T> static inline T atomicOr(T* p, T v) { T o = *p; *p = o | v; return o; }
template <typename T> static inline T atomicExch(T* p, T v) { T o = *p; *p = v; return o; }
template <typename T> static inline T atomicCAS(T* p, T cmp, T v) { T o = *p; if (o == cmp) *p = v; return o; }
static inline void __threadfence() {}
static inline void __threadfence_block() {}
static inline void __trap() { __builtin_trap(); }
static inline void __nanosleep(unsigned) {}

static inline int __popc(uint32_t v) { return __builtin_popcount(v); }
static inline int __popcll(unsigned long long v) { return __builtin_popcountll(v); }
static inline int __ffs(int v) { return __builtin_ffs(v); }
static inline uint32_t __umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((unsigned long long)a * b) >> 32); }
static inline int __clz(int v) { return v ? __builtin_clz((unsigned)v) : 32; }
static inline uint32_t __brev(uint32_t v) { uint32_t r = 0; for (int i = 0; i < 32; i++) r |= ((v >> i) & 1u) << (31 - i); return r; }
static inline float __uint_as_float(uint32_t v) { return emu_from_bits<float>(v); }
static inline uint32_t __float_as_uint(float v) { return emu_bits(v); }
static inline float __int_as_float(int v) { return emu_from_bits<float>((uint32_t)v); }
static inline int __float_as_int(float v) { return (int)emu_bits(v); }
static inline float __fmaf_rn(float a, float b, float c) { return fmaf(a, b, c); }
static inline float __fmul_rn(float a, float b) { volatile float r = a * b; return r; }
static inline float __fadd_rn(float a, float b) { volatile float r = a + b; return r; }
static inline float __fdividef(float a, float b) { return a / b; }
static inline float __frcp_rn(float a) { return 1.0f / a; }
static inline float rsqrtf(float a) { return 1.0f / sqrtf(a); }
static inline float __saturatef(float a) { return a < 0.f ? 0.f : (a > 1.f ? 1.f : a); }
template <typename T> static inline T __ldg(const T* p) { return *p; }
using std::max;
using std::min;
