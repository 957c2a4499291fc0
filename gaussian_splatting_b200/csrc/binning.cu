// binning.cu — kernel group 2 of the path: depth ordering, tile-count scan, instance
// emission, stable tile ordering, tile ranges.
//
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::
// SortPairs<u64,u32> + identifyTileRanges (DGR/cuda_rasterizer/rasterizer_impl.cu:78-164,
// 309-350).  The reference sorts R (tile|depth) 64-bit keys in 6 radix passes.  Here the
// same total order (tile, depth bits, Gaussian id) is produced in two cheaper steps:
//   1. stable LSD radix sort of the P Gaussians by depth bits (4 passes over P items);
//   2. instances are emitted in that depth order and then stably ordered by tile id
//      only: ceil(getHigherMsb(T)/8) passes over R items (2 at every benchmark size).
// Stability makes the result identical to the reference's stable sort of
// (tile<<32 | depth) keys with emission order (id, tile) as the tie-break.
#include "common.cuh"
#include "block_scan.cuh"

namespace gsr {

// ---------------------------------------------------------------- radix sort (u64 items)
__global__ void __launch_bounds__(SORT_THREADS)
radix_hist_kernel(const uint64_t* __restrict__ in, uint32_t n, int shift,
                  uint32_t* __restrict__ hist, uint32_t nb) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t base = blockIdx.x * SORT_CHUNK;
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t idx = base + i * SORT_THREADS + threadIdx.x;
        if (idx < n) atomicAdd(&h[(uint32_t)(in[idx] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[threadIdx.x * nb + blockIdx.x] = h[threadIdx.x];
}

// One block per digit: exclusive scan of its row of per-block counts, in place.
__global__ void __launch_bounds__(256)
radix_rowscan_kernel(uint32_t* __restrict__ hist, uint32_t nb, uint32_t* __restrict__ digit_tot) {
    __shared__ uint32_t s_w[8];
    uint32_t* row = hist + (size_t)blockIdx.x * nb;
    const uint32_t seg = (nb + 255) / 256;
    const uint32_t b0 = threadIdx.x * seg;
    const uint32_t b1 = min(nb, b0 + seg);
    uint32_t sum = 0;
    for (uint32_t b = b0; b < b1; b++) sum += row[b];
    uint32_t total;
    uint32_t run = block_excl_scan_256(sum, s_w, &total);
    for (uint32_t b = b0; b < b1; b++) {
        uint32_t c = row[b];
        row[b] = run;
        run += c;
    }
    if (threadIdx.x == 0) digit_tot[blockIdx.x] = total;
}

// Stable scatter of one 8-bit digit.  SPLIT: final pass, write low word to out_lo
// (Gaussian id) and high word to out_hi (tile id) instead of the u64 item.
template <bool SPLIT>
__global__ void __launch_bounds__(SORT_THREADS)
radix_scatter_kernel(const uint64_t* __restrict__ in, uint32_t n, int shift,
                     const uint32_t* __restrict__ hist, uint32_t nb,
                     const uint32_t* __restrict__ digit_tot,
                     uint64_t* __restrict__ out, uint32_t* __restrict__ out_lo,
                     uint32_t* __restrict__ out_hi) {
    __shared__ uint32_t wh[8][256];
    __shared__ uint64_t stage[SORT_CHUNK];
    __shared__ uint32_t s_gbase[256];
    __shared__ uint32_t s_dstart[256];
    __shared__ uint32_t s_w[8];

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const uint32_t block_base = blockIdx.x * SORT_CHUNK;
    const uint32_t n_block = min((uint32_t)SORT_CHUNK, n - block_base);

    uint64_t items[SORT_ITEMS];
    uint32_t ranks[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t idx = warp * (32 * SORT_ITEMS) + i * 32 + lane;
        items[i] = idx < n_block ? in[block_base + idx] : ~0ull;
    }
#pragma unroll
    for (int w = 0; w < 8; w++) wh[w][t] = 0;
    __syncthreads();

#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t idx = warp * (32 * SORT_ITEMS) + i * 32 + lane;
        bool valid = idx < n_block;
        uint32_t d = valid ? ((uint32_t)(items[i] >> shift) & 255u) : 256u;
        uint32_t mask = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(mask) - 1;
        uint32_t r = __popc(mask & ((1u << lane) - 1u));
        uint32_t old = 0;
        if (lane == leader && valid) {
            old = wh[warp][d];
            wh[warp][d] = old + __popc(mask);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        ranks[i] = old + r;
        __syncwarp();
    }
    __syncthreads();

    // per digit (thread t owns digit t): exclusive prefix over warps, block count
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        uint32_t c = wh[w][t];
        wh[w][t] = run;
        run += c;
    }
    uint32_t dstart = block_excl_scan_256(run, s_w, nullptr);
    uint32_t dbase = block_excl_scan_256(digit_tot[t], s_w, nullptr);
    s_dstart[t] = dstart;
    s_gbase[t] = dbase + hist[(size_t)t * nb + blockIdx.x] - dstart;
    __syncthreads();

#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t idx = warp * (32 * SORT_ITEMS) + i * 32 + lane;
        if (idx < n_block) {
            uint32_t d = (uint32_t)(items[i] >> shift) & 255u;
            stage[s_dstart[d] + wh[warp][d] + ranks[i]] = items[i];
        }
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        uint32_t idx = i * SORT_THREADS + t;
        if (idx < n_block) {
            uint64_t it = stage[idx];
            uint32_t d = (uint32_t)(it >> shift) & 255u;
            uint32_t pos = s_gbase[d] + idx;
            if (SPLIT) {
                out_lo[pos] = (uint32_t)it;
                out_hi[pos] = (uint32_t)(it >> 32);
            } else {
                out[pos] = it;
            }
        }
    }
}

// ---------------------------------------------------------------- onesweep radix sort
// One read + one write of the items per digit pass: digit histograms of ALL passes come from a
// single read of the input, and each pass is one kernel whose CTAs (4096-item tiles, taken in
// ticket order) rank their items, publish their per-digit counts and obtain the count of all
// earlier tiles by decoupled look-back over `status` (aggregate / inclusive-prefix flags in the
// two top bits of each word) instead of a separate histogram + scan pass.  Stable.
constexpr int OS_THREADS = 512;
constexpr int OS_WARPS = OS_THREADS / 32;
constexpr int OS_ITEMS = 8;
constexpr int OS_TILE = OS_THREADS * OS_ITEMS;          // 4096 == SORT_CHUNK
static_assert(OS_TILE == SORT_CHUNK, "scratch sizing assumes 4096-item tiles");
constexpr uint32_t OS_FLAG_AGG = 1u << 30, OS_FLAG_PREFIX = 1u << 31, OS_COUNT_MASK = (1u << 30) - 1u;
#ifndef GSR_OS_MAX_N
#define GSR_OS_MAX_N (1u << 30)                          // look-back counts must fit in 30 bits
#endif
constexpr uint32_t OS_MAX_N = GSR_OS_MAX_N;             // (tests lower it to reach the 3-kernel pass at small sizes)

// Histograms of `passes` consecutive 8-bit digits starting at bit `shift`, one read of the input.
template <int PASSES>
__global__ void __launch_bounds__(256)
radix_hist_multi_kernel(const uint64_t* __restrict__ in, uint32_t n, int shift, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t h[PASSES][256];
#pragma unroll
    for (int p = 0; p < PASSES; p++) h[p][threadIdx.x] = 0;
    __syncthreads();
    const uint32_t stride = gridDim.x * blockDim.x * 2;
    for (uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) * 2; i < n; i += stride) {
        uint64_t a, b = 0;
        const bool two = i + 1 < n;
        if (two) {
            const ulonglong2 v = *reinterpret_cast<const ulonglong2*>(in + i);
            a = v.x; b = v.y;
        } else {
            a = in[i];
        }
#pragma unroll
        for (int p = 0; p < PASSES; p++) {
            atomicAdd(&h[p][(uint32_t)(a >> (shift + 8 * p)) & 255u], 1u);
            if (two) atomicAdd(&h[p][(uint32_t)(b >> (shift + 8 * p)) & 255u], 1u);
        }
    }
    __syncthreads();
#pragma unroll
    for (int p = 0; p < PASSES; p++) {
        const uint32_t c = h[p][threadIdx.x];
        if (c) atomicAdd(&ghist[p * 256 + threadIdx.x], c);
    }
}

// Exclusive scan of each pass's 256 digit counts (block p = pass p).
__global__ void __launch_bounds__(256)
radix_bases_kernel(const uint32_t* __restrict__ ghist, uint32_t* __restrict__ gbase) {
    __shared__ uint32_t s_w[8];
    const uint32_t c = ghist[blockIdx.x * 256 + threadIdx.x];
    gbase[blockIdx.x * 256 + threadIdx.x] = block_excl_scan_256(c, s_w, nullptr);
}

#ifndef OS_MIN_BLOCKS
#define OS_MIN_BLOCKS 3
#endif
template <bool SPLIT>
__global__ void __launch_bounds__(OS_THREADS, OS_MIN_BLOCKS)
onesweep_kernel(const uint64_t* __restrict__ in, uint32_t n, int shift, const uint32_t* __restrict__ gbase,
                uint32_t* status, uint32_t* ticket,
                uint64_t* __restrict__ out, uint32_t* __restrict__ out_lo, uint32_t* __restrict__ out_hi) {
    __shared__ uint16_t wh[OS_WARPS][256];           // per-warp digit counts (<= 256), then prefixes (<= 4096)
    __shared__ uint64_t stage[OS_TILE];
    __shared__ uint32_t s_gbase[256];
    __shared__ uint32_t s_dstart[256];
    __shared__ uint32_t s_w[8];
    __shared__ uint32_t s_tile;

    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    if (t == 0) s_tile = atomicAdd(ticket, 1u);     // tiles start in ticket order: look-back never waits on an unscheduled CTA
#pragma unroll
    for (int w = 0; w < OS_WARPS / 2; w++) wh[w * 2 + (t >> 8)][t & 255] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t block_base = tile * OS_TILE;
    const uint32_t n_block = min((uint32_t)OS_TILE, n - block_base);

    uint64_t items[OS_ITEMS];
    uint32_t ranks[OS_ITEMS / 2];
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t idx = warp * (32 * OS_ITEMS) + i * 32 + lane;
        items[i] = idx < n_block ? in[block_base + idx] : ~0ull;
    }
    // Lanes holding the same digit, by 8 ballots per item.  (MATCH.ANY costs one round per distinct
    // value in the warp -- ~30 with random 8-bit digits -- and was the bottleneck of this kernel.)
    // Ranks are < 512 and kept two per register to stay within OS_MIN_BLOCKS CTAs per SM.
#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t idx = warp * (32 * OS_ITEMS) + i * 32 + lane;
        const bool valid = idx < n_block;
        const uint32_t d = (uint32_t)(items[i] >> shift) & 255u;
        uint32_t mask = __ballot_sync(0xffffffffu, valid);
#pragma unroll
        for (int b = 0; b < 8; b++) {
            const bool bit = (d >> b) & 1u;
            const uint32_t v = __ballot_sync(0xffffffffu, bit);
            mask &= bit ? v : ~v;
        }
        const uint32_t lower = mask & ((1u << lane) - 1u);
        const uint32_t old = valid ? wh[warp][d] : 0u;     // every lane of the group reads the same counter
        __syncwarp();
        if (valid && lower == 0) wh[warp][d] = (uint16_t)(old + __popc(mask));   // group leader
        __syncwarp();
        const uint32_t r = old + __popc(lower);
        if (i & 1) ranks[i >> 1] |= r << 16; else ranks[i >> 1] = r;
    }
    __syncthreads();

    // thread d < 256 owns digit d: exclusive prefix over warps, tile count, publish, look back
    uint32_t cnt = 0;
    volatile uint32_t* vstatus = status;
    if (t < 256) {
#pragma unroll
        for (int w = 0; w < OS_WARPS; w++) {
            const uint32_t c = wh[w][t];
            wh[w][t] = (uint16_t)cnt;
            cnt += c;
        }
        vstatus[(size_t)tile * 256 + t] = cnt | (tile == 0 ? OS_FLAG_PREFIX : OS_FLAG_AGG);
    }
    // start of each digit's run inside the tile (all threads take part in the barriers)
    uint32_t dstart = 0;
    {
        const int w8 = t >> 5;
        uint32_t inc = warp_incl_scan(cnt, lane);
        if (t < 256 && lane == 31) s_w[w8] = inc;
        __syncthreads();
        if (t < 256) {
            uint32_t wp = 0;
#pragma unroll
            for (int w = 0; w < 8; w++) if (w < w8) wp += s_w[w];
            dstart = wp + inc - cnt;
        }
    }
    if (t < 256) {
        uint32_t excl = 0;
        if (tile > 0) {
            int look = (int)tile - 1;
            uint32_t spins = 0;
            while (true) {
                const uint32_t v = vstatus[(size_t)look * 256 + t];
                if ((v >> 30) == 0) {
                    if (++spins > (1u << 24)) __trap();       // a lost predecessor must not hang the GPU
                    continue;
                }
                excl += v & OS_COUNT_MASK;
                if (v & OS_FLAG_PREFIX) break;
                --look;
            }
            vstatus[(size_t)tile * 256 + t] = ((excl + cnt) & OS_COUNT_MASK) | OS_FLAG_PREFIX;
        }
        s_dstart[t] = dstart;
        s_gbase[t] = gbase[t] + excl - dstart;
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t idx = warp * (32 * OS_ITEMS) + i * 32 + lane;
        if (idx < n_block) {
            const uint32_t d = (uint32_t)(items[i] >> shift) & 255u;
            stage[s_dstart[d] + wh[warp][d] + ((ranks[i >> 1] >> ((i & 1) * 16)) & 0xffffu)] = items[i];
        }
    }
    __syncthreads();

#pragma unroll
    for (int i = 0; i < OS_ITEMS; i++) {
        const uint32_t idx = i * OS_THREADS + t;
        if (idx < n_block) {
            const uint64_t it = stage[idx];
            const uint32_t d = (uint32_t)(it >> shift) & 255u;
            const uint32_t pos = s_gbase[d] + idx;
            if (SPLIT) {
                out_lo[pos] = (uint32_t)it;
                out_hi[pos] = (uint32_t)(it >> 32);
            } else {
                out[pos] = it;
            }
        }
    }
}

// Stable LSD sort of n u64 items on `passes` 8-bit digits from bit `shift` (a -> b -> a ...).
// split_last: the final pass writes low words to out_lo and high words to out_hi instead.
// Returns through *result the buffer holding the sorted items (unused with split_last).
static int radix_sort_reset(uint32_t n, int passes, Scratch s, cudaStream_t st) {
    const uint32_t nb = div_up(n, OS_TILE);
    GSR_CUDA_TRY(cudaMemsetAsync(s.sort_ctl, 0, (SORT_CTL_STATUS + (size_t)passes * 256 * nb) * sizeof(uint32_t), st));
    return GSR_OK;
}

// have_hist: the digit histograms were already accumulated into the (reset) control block by the
// producer of the items (instance emission), so the input is not read an extra time.
static int radix_sort_onesweep(uint64_t* a, uint64_t* b, uint32_t n, int shift, int passes, Scratch s,
                               bool split_last, uint32_t* out_lo, uint32_t* out_hi, uint64_t** result,
                               bool have_hist, cudaStream_t st) {
    const uint32_t nb = div_up(n, OS_TILE);
    uint32_t* ctl = s.sort_ctl;
    if (!have_hist) {
    int rc = radix_sort_reset(n, passes, s, st);
    if (rc) return rc;
    const uint32_t hist_blocks = min(div_up(n, 512u), 148u * 8u);
    switch (passes) {
        case 1: GSR_LAUNCH(radix_hist_multi_kernel<1>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
        case 2: GSR_LAUNCH(radix_hist_multi_kernel<2>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
        case 3: GSR_LAUNCH(radix_hist_multi_kernel<3>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
        default: GSR_LAUNCH(radix_hist_multi_kernel<4>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
    }
    }
    GSR_LAUNCH(radix_bases_kernel, passes, 256, 0, st, ctl + SORT_CTL_HIST, ctl + SORT_CTL_BASE);
    for (int pass = 0; pass < passes; pass++) {
        uint32_t* status = ctl + SORT_CTL_STATUS + (size_t)pass * 256 * nb;
        const bool split = split_last && pass == passes - 1;
        if (split)
            GSR_LAUNCH(onesweep_kernel<true>, nb, OS_THREADS, 0, st, a, n, shift + 8 * pass, ctl + SORT_CTL_BASE + pass * 256,
                                                            status, ctl + SORT_CTL_TICKET + pass, nullptr, out_lo, out_hi);
        else
            GSR_LAUNCH(onesweep_kernel<false>, nb, OS_THREADS, 0, st, a, n, shift + 8 * pass, ctl + SORT_CTL_BASE + pass * 256,
                                                             status, ctl + SORT_CTL_TICKET + pass, b, nullptr, nullptr);
        uint64_t* tmp = a; a = b; b = tmp;
    }
    if (result) *result = a;
    return GSR_OK;
}

// One LSD pass src -> dst on digit `shift`. split => final pass with split outputs.
static int radix_pass(const uint64_t* src, uint64_t* dst, uint32_t n, int shift, Scratch s,
                      bool split, uint32_t* out_lo, uint32_t* out_hi, cudaStream_t st) {
    const uint32_t nb = div_up(n, SORT_CHUNK);
    GSR_LAUNCH(radix_hist_kernel, nb, SORT_THREADS, 0, st, src, n, shift, s.hist, nb);
    GSR_LAUNCH(radix_rowscan_kernel, 256, 256, 0, st, s.hist, nb, s.digit_tot);
    if (split)
        GSR_LAUNCH(radix_scatter_kernel<true>, nb, SORT_THREADS, 0, st, src, n, shift, s.hist, nb, s.digit_tot,
                                                                nullptr, out_lo, out_hi);
    else
        GSR_LAUNCH(radix_scatter_kernel<false>, nb, SORT_THREADS, 0, st, src, n, shift, s.hist, nb, s.digit_tot,
                                                                 dst, nullptr, nullptr);
    return GSR_OK;
}

// ---------------------------------------------------------------- tile-count scan
__device__ __forceinline__ uint32_t rect_tiles(uint2 r) {
    uint32_t w = (r.y & 0xffffu) - (r.x & 0xffffu);
    uint32_t h = (r.y >> 16) - (r.x >> 16);
    return w * h;
}

// Phase 1: tiles touched by the Gaussian at each depth rank -> offsets[] (raw counts)
// and the per-block sums.
__global__ void __launch_bounds__(SCAN_THREADS)
scan_count_kernel(const uint64_t* __restrict__ sorted, const uint2* __restrict__ rects, uint32_t P,
                  uint32_t* __restrict__ offsets, uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t s_w[8];
    const uint32_t base = blockIdx.x * SCAN_CHUNK;
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        uint32_t idx = base + i * SCAN_THREADS + threadIdx.x;
        if (idx < P) {
            uint32_t c = rect_tiles(rects[(uint32_t)sorted[idx]]);
            offsets[idx] = c;
            sum += c;
        }
    }
    uint32_t total;
    block_excl_scan_256(sum, s_w, &total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// Phase 2: one block; exclusive scan of the block sums in place, total -> counters.
__global__ void __launch_bounds__(256)
scan_blocksums_kernel(uint32_t* __restrict__ block_sums, uint32_t nb, int32_t* __restrict__ counters) {
    __shared__ uint32_t s_w[8];
    unsigned long long carry = 0;
    for (uint32_t b0 = 0; b0 < nb; b0 += 256) {
        uint32_t b = b0 + threadIdx.x;
        uint32_t v = b < nb ? block_sums[b] : 0u;
        uint32_t total;
        uint32_t ex = block_excl_scan_256(v, s_w, &total);
        if (b < nb) block_sums[b] = (uint32_t)carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) {
        counters[0] = carry > 0x7fffffffull ? 0x7fffffff : (int32_t)carry;
        counters[1] = carry > 0x7fffffffull ? 1 : 0;   // reference: int num_rendered (rasterizer_impl.cu:312)
    }
}

// Phase 3: inclusive scan inside each block + block prefix, in place.
__global__ void __launch_bounds__(SCAN_THREADS)
scan_apply_kernel(uint32_t* __restrict__ offsets, uint32_t P, const uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t s_w[8];
    const uint32_t base = blockIdx.x * SCAN_CHUNK + threadIdx.x * SCAN_ITEMS;
    uint32_t v[SCAN_ITEMS];
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        v[i] = (base + i < P) ? offsets[base + i] : 0u;
        sum += v[i];
    }
    uint32_t run = block_excl_scan_256(sum, s_w, nullptr) + block_sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        run += v[i];
        if (base + i < P) offsets[base + i] = run;
    }
}

// ---------------------------------------------------------------- instance emission
// duplicateWithKeys (rasterizer_impl.cu:78-126) in depth order; item = tile<<32 | id.
// Warp-cooperative: a warp owns 32 consecutive depth ranks, whose instances occupy one contiguous
// range of the output; lane j writes instance first + j, first + j + 32, ... (coalesced 8 B stores)
// after locating its Gaussian among the warp's 32 inclusive offsets with a 5-step shuffle search.
// The 8-bit digit histograms of the tile ids (what the following radix sort needs) are accumulated
// on the way, so the instance list is not read again for them.
constexpr int EMIT_THREADS = 256;
constexpr int EMIT_ROUNDS = 4;                     // 1024 Gaussians per CTA

template <int PASSES>
__global__ void __launch_bounds__(EMIT_THREADS)
emit_instances_kernel(const uint64_t* __restrict__ sorted, const uint2* __restrict__ rects,
                      const uint32_t* __restrict__ offsets, uint32_t P, uint32_t grid_x,
                      uint64_t* __restrict__ inst, uint32_t* __restrict__ ghist) {
    __shared__ uint32_t s_hist[PASSES][256];
#pragma unroll
    for (int p = 0; p < PASSES; p++) s_hist[p][threadIdx.x] = 0;
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t last_incl = offsets[P - 1];
    for (int round = 0; round < EMIT_ROUNDS; round++) {
        const uint32_t i = (blockIdx.x * EMIT_ROUNDS + round) * EMIT_THREADS + threadIdx.x;
        uint32_t g = 0, x0 = 0, y0 = 0, w = 1, incl = last_incl, cnt = 0;
        if (i < P) {
            const uint64_t it = sorted[i];
            incl = offsets[i];
            if ((uint32_t)(it >> 32) != 0xFFFFFFFFu) {          // not culled
                g = (uint32_t)it;
                const uint2 r = rects[g];
                x0 = r.x & 0xffffu; y0 = r.x >> 16;
                w = (r.y & 0xffffu) - x0;
                cnt = w * ((r.y >> 16) - y0);
            }
        }
        const uint32_t excl = incl - cnt;
        const uint32_t first = __shfl_sync(0xffffffffu, excl, 0);
        const uint32_t total = __shfl_sync(0xffffffffu, incl, 31) - first;
        const uint32_t xy = x0 | (y0 << 16);
        for (uint32_t j = lane; j < ((total + 31u) & ~31u); j += 32) {      // uniform trip count: shuffles inside
            const uint32_t target = first + j;
            // L = number of lanes whose inclusive offset is <= target = owner of instance `target`
            uint32_t L = 0;
#pragma unroll
            for (int step = 16; step > 0; step >>= 1) {
                const uint32_t probe = __shfl_sync(0xffffffffu, incl, (L + step - 1) & 31);
                if (L + step <= 32 && probe <= target) L += step;
            }
            const uint32_t src = L & 31;
            const uint32_t o_excl = __shfl_sync(0xffffffffu, excl, src);
            const uint32_t o_w = __shfl_sync(0xffffffffu, w, src);
            const uint32_t o_xy = __shfl_sync(0xffffffffu, xy, src);
            const uint32_t o_g = __shfl_sync(0xffffffffu, g, src);
            if (j < total) {
                const uint32_t local = target - o_excl;
                const uint32_t ry = local / o_w, rx = local - ry * o_w;
                const uint32_t tile = ((o_xy >> 16) + ry) * grid_x + (o_xy & 0xffffu) + rx;
                inst[target] = ((uint64_t)tile << 32) | o_g;
#pragma unroll
                for (int p = 0; p < PASSES; p++) atomicAdd(&s_hist[p][(tile >> (8 * p)) & 255u], 1u);
            }
        }
    }
    __syncthreads();
#pragma unroll
    for (int p = 0; p < PASSES; p++) {
        const uint32_t c = s_hist[p][threadIdx.x];
        if (c) atomicAdd(&ghist[p * 256 + threadIdx.x], c);
    }
}

// identifyTileRanges (rasterizer_impl.cu:132-164) on the sorted tile ids; 4 instances per thread.
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t* __restrict__ tile_ids, uint32_t L, uint2* __restrict__ ranges) {
    const uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= L) return;
    uint32_t v[4];
    if (i0 + 3 < L) {
        const uint4 q = *reinterpret_cast<const uint4*>(tile_ids + i0);
        v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = i0 + k < L ? tile_ids[i0 + k] : 0u;
    }
    uint32_t prev = i0 > 0 ? tile_ids[i0 - 1] : 0u;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t idx = i0 + k;
        if (idx >= L) break;
        const uint32_t cur = v[k];
        if (idx == 0)
            ranges[cur].x = 0;
        else if (cur != prev) {
            ranges[prev].y = idx;
            ranges[cur].x = idx;
        }
        if (idx == L - 1) ranges[cur].y = L;
        prev = cur;
    }
}

// Test hook: rebuild the reference's sorted keys  tile<<32 | depth bits.
__global__ void rebuild_keys_kernel(const uint2* __restrict__ ranges, uint32_t T,
                                    const uint32_t* __restrict__ point_list,
                                    const GsRecord* __restrict__ rec, uint64_t* __restrict__ keys) {
    uint32_t tile = blockIdx.x;
    if (tile >= T) return;
    uint2 rg = ranges[tile];
    for (uint32_t i = rg.x + threadIdx.x; i < rg.y; i += blockDim.x)
        keys[i] = ((uint64_t)tile << 32) | __float_as_uint(rec[point_list[i]].q3.z);
}

// ---------------------------------------------------------------- launchers
int launch_depth_order_and_scan(const GsrParams& p, GeomState g, Scratch s, int32_t* num_rendered_host,
                                cudaStream_t st) {
    const uint32_t P = (uint32_t)p.P;
    // depth_a holds depth_bits<<32 | id (written by preprocess). 4 stable LSD passes.
    uint64_t* a = s.depth_a;
    uint64_t* b = s.depth_b;
    {
        StageTimer tm(GSR_STAGE_DEPTH_SORT, st);
        if (P < OS_MAX_N) {
            int rc = radix_sort_onesweep(a, b, P, 32, 4, s, false, nullptr, nullptr, nullptr, false, st);
            if (rc) return rc;
        } else {
            for (int pass = 0; pass < 4; pass++) {
                radix_pass(a, b, P, 32 + 8 * pass, s, false, nullptr, nullptr, st);
                uint64_t* tmp = a; a = b; b = tmp;
            }
        }
    }
    int rc = check_stage("depth sort", p.debug, st);
    if (rc) return rc;
    // a == s.depth_a again (even number of passes)
    {
        StageTimer tm(GSR_STAGE_SCAN, st);
        const uint32_t nb = div_up(P, SCAN_CHUNK);
        GSR_LAUNCH(scan_count_kernel, nb, SCAN_THREADS, 0, st, a, s.rects, P, s.offsets, s.scan_blocks);
        GSR_LAUNCH(scan_blocksums_kernel, 1, 256, 0, st, s.scan_blocks, nb, s.counters);
        GSR_LAUNCH(scan_apply_kernel, nb, SCAN_THREADS, 0, st, s.offsets, P, s.scan_blocks);
    }
    rc = check_stage("tile-count scan", p.debug, st);
    if (rc) return rc;
    if (num_rendered_host)
        GSR_CUDA_TRY(cudaMemcpyAsync(num_rendered_host, s.counters, 2 * sizeof(int32_t),
                                     cudaMemcpyDeviceToHost, st));
    return GSR_OK;
}

int launch_binning(const GsrParams& p, GeomState g, ImgState im, BinState b, Scratch s, int64_t R,
                   cudaStream_t st) {
    const uint32_t P = (uint32_t)p.P;
    const uint32_t gx = div_up(p.W, TILE_X), gy = div_up(p.H, TILE_Y);
    const uint32_t T = gx * gy;
    GSR_CUDA_TRY(cudaMemsetAsync(im.ranges, 0, (size_t)T * sizeof(uint2), st));   // rasterizer_impl.cu:342
    if (R <= 0) return GSR_OK;
    const uint32_t n = (uint32_t)R;
    // reference sorts bits [0, 32+getHigherMsb(T)) (rasterizer_impl.cu:332-340); the low 32
    // (depth) are already in order, so only the tile bits remain.
    const int bits = (int)higher_msb(T);
    const int passes = (bits + 7) / 8;
    const bool fused_hist = n < OS_MAX_N;
    int rc = GSR_OK;
    {
        StageTimer tm(GSR_STAGE_EMIT, st);
        const uint32_t eb = div_up(P, EMIT_THREADS * EMIT_ROUNDS);
        uint32_t* gh = s.sort_ctl + SORT_CTL_HIST;
        if (fused_hist) {
            rc = radix_sort_reset(n, passes, s, st);
            if (rc) return rc;
        }
        switch (fused_hist ? passes : 0) {
            case 1: GSR_LAUNCH(emit_instances_kernel<1>, eb, EMIT_THREADS, 0, st, s.depth_a, s.rects, s.offsets, P, gx, s.inst_a, gh); break;
            case 2: GSR_LAUNCH(emit_instances_kernel<2>, eb, EMIT_THREADS, 0, st, s.depth_a, s.rects, s.offsets, P, gx, s.inst_a, gh); break;
            case 3: GSR_LAUNCH(emit_instances_kernel<3>, eb, EMIT_THREADS, 0, st, s.depth_a, s.rects, s.offsets, P, gx, s.inst_a, gh); break;
            default: GSR_LAUNCH(emit_instances_kernel<4>, eb, EMIT_THREADS, 0, st, s.depth_a, s.rects, s.offsets, P, gx, s.inst_a,
                                s.hist /* unused scratch */); break;
        }
    }
    rc = check_stage("emit instances", p.debug, st);
    if (rc) return rc;
    uint64_t* a = s.inst_a;
    uint64_t* c = s.inst_b;
    {
        StageTimer tm(GSR_STAGE_TILE_SORT, st);
        if (fused_hist) {
            rc = radix_sort_onesweep(a, c, n, 32, passes, s, true, b.point_list, s.tile_ids, nullptr, true, st);
            if (rc) return rc;
        } else
        for (int pass = 0; pass < passes; pass++) {
            const bool last = pass == passes - 1;
            radix_pass(a, c, n, 32 + 8 * pass, s, last, b.point_list, s.tile_ids, st);
            uint64_t* tmp = a; a = c; c = tmp;
        }
    }
    rc = check_stage("tile sort", p.debug, st);
    if (rc) return rc;
    {
        StageTimer tm(GSR_STAGE_TILE_RANGES, st);
        GSR_LAUNCH(tile_ranges_kernel, div_up(n, 1024), 256, 0, st, s.tile_ids, n, im.ranges);
    }
    return check_stage("tile ranges", p.debug, st);
}

int launch_rebuild_keys(const GsrParams& p, GeomState g, ImgState im, BinState b, int64_t R,
                        uint64_t* keys_out, cudaStream_t st) {
    const uint32_t gx = div_up(p.W, TILE_X), gy = div_up(p.H, TILE_Y);
    if (R <= 0) return GSR_OK;
    GSR_LAUNCH(rebuild_keys_kernel, gx * gy, 128, 0, st, im.ranges, gx * gy, b.point_list, g.rec, keys_out);
    return check_stage("rebuild keys", true, st);
}

}  // namespace gsr
