// selftest.cu — checks of the CUDA emulation itself (compiled only by tests/test_cuda_emul_cpu.py).
#include <cuda_runtime.h>
#include <stdint.h>

// Every warp primitive once; results are checked on the host against closed forms.
__global__ void primitives_kernel(uint32_t* out, int n) {
    __shared__ uint32_t s_sum[8];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t gid = blockIdx.x * blockDim.x + tid;
    if ((int)gid >= n) return;                                   // exited threads leave the barriers
    uint32_t v = gid;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);      // warp sum of gid
    const uint32_t up = __shfl_up_sync(0xffffffffu, gid, 1);                        // lane 0 keeps its own
    const uint32_t bal = __ballot_sync(0xffffffffu, gid & 1u);
    const uint32_t first = __shfl_sync(0xffffffffu, gid, 0);
    const uint32_t match = __match_any_sync(0xffffffffu, gid % 3u);
    if (lane == 0) s_sum[warp] = v;
    const int odd = __syncthreads_count(gid & 1u);
    uint32_t block_sum = 0;
    for (uint32_t w = 0; w < (blockDim.x >> 5); w++) block_sum += s_sum[w];
    uint32_t* o = out + 8 * (size_t)gid;
    o[0] = v; o[1] = up; o[2] = bal; o[3] = first; o[4] = match; o[5] = (uint32_t)odd; o[6] = block_sum;
    o[7] = __any_sync(0xffffffffu, gid == 5) | (__all_sync(0xffffffffu, gid < 1000000u) << 1);
    atomicAdd(out + 8 * (size_t)n, 1u);
}

// A reversal through shared memory: correct with the barrier, schedule-dependent without it.
template <bool BARRIER>
__global__ void reverse_kernel(const uint32_t* in, uint32_t* out) {
    __shared__ uint32_t s[256];
    s[threadIdx.x] = in[threadIdx.x];
    if (BARRIER) __syncthreads();
    out[threadIdx.x] = s[255 - threadIdx.x];
    __syncthreads();
    s[threadIdx.x] = 0;                                          // leave the static storage clean for the next launch
}

extern "C" void run_primitives(uint32_t* out, int n, int block) {
    emu::launch(dim3((n + block - 1) / block), dim3(block), [&]() { primitives_kernel(out, n); });
}
extern "C" void run_reverse(const uint32_t* in, uint32_t* out, int barrier) {
    if (barrier) emu::launch(dim3(1), dim3(256), [&]() { reverse_kernel<true>(in, out); });
    else emu::launch(dim3(1), dim3(256), [&]() { reverse_kernel<false>(in, out); });
}
