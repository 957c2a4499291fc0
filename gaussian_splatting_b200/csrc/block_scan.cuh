// block_scan.cuh — warp / block prefix-sum helpers shared by binning.cu and densify.cu.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gsr {

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t n = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += n;
    }
    return v;
}

// Exclusive scan over the 256 threads of a block. s_w: 8 words of shared memory.
// All threads must call. Returns exclusive prefix; *total gets the block total.
__device__ __forceinline__ uint32_t block_excl_scan_256(uint32_t v, uint32_t* s_w, uint32_t* total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = warp_incl_scan(v, lane);
    __syncthreads();                       // protect s_w reuse
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    uint32_t wp = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t x = s_w[w];
        if (w < warp) wp += x;
        tot += x;
    }
    if (total) *total = tot;
    return wp + inc - v;
}

}  // namespace gsr
