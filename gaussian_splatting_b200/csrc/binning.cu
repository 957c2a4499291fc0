// binning.cu — kernel group 2 of the path: depth ordering and the per-tile, depth-ordered Gaussian lists.
//
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + cub::DeviceRadixSort::
// SortPairs<u64,u32> + identifyTileRanges (DGR/cuda_rasterizer/rasterizer_impl.cu:78-164,
// 309-350).  The reference materialises R (tile|depth, id) pairs and sorts them in 6 radix
// passes.  The same total order (tile, depth bits, Gaussian id) is produced here without ever
// sorting the R instances:
//   1. stable LSD radix sort of the P Gaussians by depth bits (4 onesweep passes over P items);
//   2. the depth-ordered Gaussians are cut into chunks; every (chunk, tile) count comes from a
//      2-D difference grid (4 shared-memory adds per Gaussian + a prefix sum), a prefix over
//      the chunks of each tile gives every chunk its first slot in every tile list, a scan
//      over the tiles gives the tile ranges and num_rendered;
//   3. one warp per chunk walks its Gaussians in depth order and appends each to the lists of
//      the tiles of its rectangle (lanes = tiles of one Gaussian, running slots in shared memory).
// A tile's list is filled in depth order (chunks in order, Gaussians in order inside a chunk);
// ties in depth keep Gaussian-id order because the depth sort is stable: the result equals the
// reference's stable sort of (tile<<32 | depth) keys emitted in (id, tile) order, bit for bit.
#include "common.cuh"
#include "block_scan.cuh"
#include "async_copy.cuh"

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

// Stable scatter of one 8-bit digit.
__global__ void __launch_bounds__(SORT_THREADS)
radix_scatter_kernel(const uint64_t* __restrict__ in, uint32_t n, int shift,
                     const uint32_t* __restrict__ hist, uint32_t nb,
                     const uint32_t* __restrict__ digit_tot,
                     uint64_t* __restrict__ out) {
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
            out[s_gbase[d] + idx] = it;
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
__global__ void __launch_bounds__(OS_THREADS, OS_MIN_BLOCKS)
onesweep_kernel(const uint64_t* __restrict__ in, uint32_t n, int shift, const uint32_t* __restrict__ gbase,
                uint32_t* status, uint32_t* ticket, uint64_t* __restrict__ out) {
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
            unsigned long long t0 = 0;
            while (true) {
                const uint32_t v = vstatus[(size_t)look * 256 + t];
                if ((v >> 30) == 0) {
#ifndef GSR_HOST_EMUL
                    if ((++spins & 0xfffu) == 0) {             // a lost predecessor must not hang the GPU (wall-time bound)
                        const unsigned long long now = global_timer_ns();
                        if (t0 == 0) t0 = now;
                        else if (now - t0 > SPIN_LIMIT_NS) __trap();
                    }
#else
                    if (++spins > (1u << 24)) __trap();
                    (void)t0;
#endif
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
            out[s_gbase[d] + idx] = it;
        }
    }
}

// Stable LSD sort of n u64 items on `passes` 8-bit digits from bit `shift` (a -> b -> a ...).
static int radix_sort_onesweep(uint64_t* a, uint64_t* b, uint32_t n, int shift, int passes, Scratch s, cudaStream_t st) {
    const uint32_t nb = div_up(n, OS_TILE);
    uint32_t* ctl = s.sort_ctl;
    GSR_CUDA_TRY(cudaMemsetAsync(ctl, 0, (SORT_CTL_STATUS + (size_t)passes * 256 * nb) * sizeof(uint32_t), st));
    const uint32_t hist_blocks = min(div_up(n, 512u), 148u * 8u);
    switch (passes) {
        case 1: GSR_LAUNCH(radix_hist_multi_kernel<1>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
        case 2: GSR_LAUNCH(radix_hist_multi_kernel<2>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
        case 3: GSR_LAUNCH(radix_hist_multi_kernel<3>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
        default: GSR_LAUNCH(radix_hist_multi_kernel<4>, hist_blocks, 256, 0, st, a, n, shift, ctl + SORT_CTL_HIST); break;
    }
    GSR_LAUNCH(radix_bases_kernel, passes, 256, 0, st, ctl + SORT_CTL_HIST, ctl + SORT_CTL_BASE);
    for (int pass = 0; pass < passes; pass++) {
        uint32_t* status = ctl + SORT_CTL_STATUS + (size_t)pass * 256 * nb;
        GSR_LAUNCH(onesweep_kernel, nb, OS_THREADS, 0, st, a, n, shift + 8 * pass, ctl + SORT_CTL_BASE + pass * 256,
                                                    status, ctl + SORT_CTL_TICKET + pass, b);
        uint64_t* tmp = a; a = b; b = tmp;
    }
    return GSR_OK;
}

// One LSD pass src -> dst on digit `shift` (3 kernels; only for n >= 2^30, where look-back counts do not fit).
static int radix_pass(const uint64_t* src, uint64_t* dst, uint32_t n, int shift, Scratch s, cudaStream_t st) {
    const uint32_t nb = div_up(n, SORT_CHUNK);
    GSR_LAUNCH(radix_hist_kernel, nb, SORT_THREADS, 0, st, src, n, shift, s.hist, nb);
    GSR_LAUNCH(radix_rowscan_kernel, 256, 256, 0, st, s.hist, nb, s.digit_tot);
    GSR_LAUNCH(radix_scatter_kernel, nb, SORT_THREADS, 0, st, src, n, shift, s.hist, nb, s.digit_tot, dst);
    return GSR_OK;
}

// ---------------------------------------------------------------- per-tile lists
// Unit of work = (chunk of consecutive depth ranks, band of tile rows).  The band keeps a unit's
// tile counters in shared memory (<= TB_MAX_CELLS tiles); at 1080p the whole image is one band.
// Chunks cover the V visible Gaussians (they sort before the culled ones, whose keys are all-ones;
// V is counted by the preprocess kernel) and are cut so that every chunk holds about the same number
// of list-filling steps (a step appends one Gaussian to up to 32 tiles): near Gaussians cover hundreds
// of tiles, far ones a few, so equal Gaussian counts would leave the first chunks 6x the work.
// A culled Gaussian's rectangle is empty, so a stray one inside a chunk adds nothing.

// Shared-memory word access by shared-window address (the compiler otherwise rebuilds the window base in every
// step of the list-filling loop), and the fence between two steps of a warp.  On the device a warp runs the
// filling loop converged (every branch in it is warp-uniform) and the volatile accesses stay in program order, so
// one step's store precedes the next step's load in the same warp's instruction stream; the CPU emulation runs the
// lanes as fibers and needs the scheduling point.
#ifndef GSR_HOST_EMUL
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void step_fence() {}
#else
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { return *reinterpret_cast<uint32_t*>(smem_ptr(a)); }
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) { *reinterpret_cast<uint32_t*>(smem_ptr(a)) = v; }
__device__ __forceinline__ void step_fence() { __syncwarp(); }
#endif

#ifndef GSR_TB_FILL_WARPS
#define GSR_TB_FILL_WARPS 8
#endif
constexpr int TB_FILL_WARPS = GSR_TB_FILL_WARPS;       // warps (= stripes of tile rows) per CTA of the list-filling kernel

// Part of rectangle r inside tile rows [r0, r1).
__device__ __forceinline__ void clip_rect(uint2 r, uint32_t r0, uint32_t r1, uint32_t& x0, uint32_t& x1,
                                          uint32_t& ya, uint32_t& yb) {
    x0 = r.x & 0xffffu; x1 = r.y & 0xffffu;
    ya = max(r.x >> 16, r0); yb = min(r.y >> 16, r1);
}

// ceil(2^31 / w): (2 l) * magic >> 32 == l / w while l * w < 2^31 (here l < 2^20 tiles of a band, w <= 2^11 columns)
__device__ __forceinline__ uint32_t div_magic(uint32_t w) { return 0x7FFFFFFFu / (w ? w : 1u) + 1u; }

// Step 2a.  The Gaussians in depth order as 16-byte records {rect.x, rect.y, div_magic(width), id} (what the two
// following kernels stream instead of gathering by id) and, per weight block of pl.wblock consecutive depth ranks,
// the number of list-filling steps they need.  One CTA per weight block.
__global__ void __launch_bounds__(256)
tile_weights_kernel(const uint64_t* __restrict__ sorted, const uint2* __restrict__ rects,
                    const int32_t* __restrict__ counters, uint32_t wblock,
                    uint4* __restrict__ srects, uint32_t* __restrict__ weights) {
    __shared__ uint32_t s_sum;
    if (threadIdx.x == 0) s_sum = 0;
    __syncthreads();
    const uint32_t V = (uint32_t)counters[CNT_VISIBLE];
    const uint32_t k0 = blockIdx.x * wblock, k1 = min(V, k0 + wblock);
    uint32_t steps = 0;
    for (uint32_t k = k0 + threadIdx.x; k < k1; k += 256) {
        const uint32_t id = (uint32_t)sorted[k];
        const uint2 r = rects[id];
        const uint32_t w = (r.y & 0xffffu) - (r.x & 0xffffu);
        srects[k] = make_uint4(r.x, r.y, div_magic(w), id);
        steps += (w * ((r.y >> 16) - (r.x >> 16)) + 31u) >> 5;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) steps += __shfl_xor_sync(0xffffffffu, steps, o);
    if ((threadIdx.x & 31u) == 0 && steps) atomicAdd(&s_sum, steps);
    __syncthreads();
    if (threadIdx.x == 0) weights[blockIdx.x] = s_sum;
}

// Step 2b.  One CTA: exclusive scan of the (<= TB_MAX_WBLOCKS) weights in shared memory; chunk c starts where the
// running weight reaches c / n_chunks of the total, interpolated linearly inside a weight block (the cut only
// balances work: any monotone bounds with bounds[0] = 0 and bounds[n_chunks] = V give the same lists).
constexpr int TB_SCAN_THREADS = 1024;
__global__ void __launch_bounds__(TB_SCAN_THREADS)
tile_bounds_kernel(const uint32_t* __restrict__ weights, uint32_t n_wblocks, uint32_t wblock,
                   const int32_t* __restrict__ counters, uint32_t n_chunks, uint32_t* __restrict__ bounds) {
    __shared__ uint32_t s_e[TB_MAX_WBLOCKS + 1];     // exclusive prefixes, s_e[n_wblocks] = total
    __shared__ uint32_t s_w[32];
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    constexpr uint32_t PER = TB_MAX_WBLOCKS / TB_SCAN_THREADS;
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (uint32_t j = 0; j < PER; j++) {
        const uint32_t i = t * PER + j;
        v[j] = i < n_wblocks ? weights[i] : 0u;
        sum += v[j];
    }
    const uint32_t inc = warp_incl_scan(sum, lane);
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = s_w[lane];
        s_w[lane] = warp_incl_scan(w, lane) - w;
    }
    __syncthreads();
    uint32_t run = s_w[warp] + inc - sum;
#pragma unroll
    for (uint32_t j = 0; j < PER; j++) {
        s_e[t * PER + j] = run;
        run += v[j];
    }
    if (t == TB_SCAN_THREADS - 1) s_e[TB_MAX_WBLOCKS] = run;      // total (entries past n_wblocks hold the total too)
    __syncthreads();
    const unsigned long long total = s_e[TB_MAX_WBLOCKS];
    const uint32_t V = (uint32_t)counters[CNT_VISIBLE];
    for (uint32_t c = t; c <= n_chunks; c += TB_SCAN_THREADS) {
        unsigned long long rank = V;
        if (c < n_chunks) {
            const uint32_t target = (uint32_t)(total * c / n_chunks);
            uint32_t lo = 0, hi = n_wblocks;                       // last block with prefix <= target
            while (hi - lo > 1) {
                const uint32_t mid = (lo + hi) >> 1;
                if (s_e[mid] <= target) lo = mid; else hi = mid;
            }
            const uint32_t wb = s_e[lo + 1] - s_e[lo];
            rank = (unsigned long long)lo * wblock + (wb ? (unsigned long long)(target - s_e[lo]) * wblock / wb : 0ull);
        }
        bounds[c] = rank < V ? (uint32_t)rank : V;
    }
}

// Step 2c.  count[chunk][tile] = Gaussians of the chunk whose rectangle covers the tile: +-1 at the four
// corners of every rectangle in a difference grid (order-free shared-memory adds), then a 2-D prefix sum.
__global__ void __launch_bounds__(TB_COUNT_THREADS)
tile_count_kernel(const uint4* __restrict__ srects, const uint32_t* __restrict__ bounds, TileBinPlan pl,
                  uint32_t* __restrict__ M) {
    __shared__ int32_t D[TB_COUNT_CELLS];
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5, chunk = blockIdx.x, band = blockIdx.y;
    const uint32_t r0 = band * pl.band_rows, r1 = min(pl.gy, r0 + pl.band_rows);
    const uint32_t rows = r1 - r0, S = pl.stride, gx = pl.gx;
    const uint32_t k0 = bounds[chunk], k1 = bounds[chunk + 1];
    for (uint32_t i = t; i < (rows + 1) * S; i += TB_COUNT_THREADS) D[i] = 0;
    __syncthreads();
    for (uint32_t k = k0 + t; k < k1; k += TB_COUNT_THREADS) {
        uint32_t x0, x1, ya, yb;
        const uint4 g = srects[k];
        clip_rect(make_uint2(g.x, g.y), r0, r1, x0, x1, ya, yb);
        if (x1 > x0 && yb > ya) {
            const uint32_t a = (ya - r0) * S, b = (yb - r0) * S;
            atomicAdd(&D[a + x0], 1);
            atomicAdd(&D[a + x1], -1);
            atomicAdd(&D[b + x0], -1);
            atomicAdd(&D[b + x1], 1);
        }
    }
    __syncthreads();
    // along x: one warp per row, 32 columns at a time
    for (uint32_t r = warp; r < rows; r += TB_COUNT_THREADS / 32) {
        uint32_t carry = 0;
        for (uint32_t x = 0; x < gx; x += 32) {
            const uint32_t v = x + lane < gx ? (uint32_t)D[r * S + x + lane] : 0u;
            const uint32_t inc = warp_incl_scan(v, lane) + carry;
            if (x + lane < gx) D[r * S + x + lane] = (int32_t)inc;
            carry = __shfl_sync(0xffffffffu, inc, 31);
        }
    }
    __syncthreads();
    uint32_t* out = M + (size_t)chunk * pl.T + (size_t)r0 * gx;
    for (uint32_t x = t; x < gx; x += TB_COUNT_THREADS) {            // along y, fused with the store
        int32_t run = 0;
        for (uint32_t r = 0; r < rows; r++) { run += D[r * S + x]; out[r * gx + x] = (uint32_t)run; }
    }
}

// Step 2d.  Per tile: exclusive prefix of the counts over the chunks, in place, and the tile's total.
// 32 tiles x TB_CP_GROUPS chunk groups per CTA; each thread sums its group, the group totals are combined
// through shared memory, a second sweep writes the prefixes.
constexpr int TB_CP_GROUPS = 32;
__global__ void __launch_bounds__(32 * TB_CP_GROUPS)
tile_chunk_prefix_kernel(uint32_t* __restrict__ M, TileBinPlan pl, uint32_t* __restrict__ totals) {
    __shared__ uint32_t s_tot[TB_CP_GROUPS][32];
    const uint32_t lane = threadIdx.x & 31u, grp = threadIdx.x >> 5;
    const uint32_t t = blockIdx.x * 32 + lane, T = pl.T;
    const uint32_t c0 = pl.n_chunks * grp / TB_CP_GROUPS, c1 = pl.n_chunks * (grp + 1) / TB_CP_GROUPS;
    uint32_t sum = 0;
    if (t < T) {
#pragma unroll 8
        for (uint32_t c = c0; c < c1; c++) sum += M[(size_t)c * T + t];
    }
    s_tot[grp][lane] = sum;
    __syncthreads();
    uint32_t run = 0;
    for (uint32_t g = 0; g < grp; g++) run += s_tot[g][lane];
    if (t >= T) return;
    for (uint32_t c = c0; c < c1; c += 8) {
        uint32_t v[8];
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = c + j < c1 ? M[(size_t)(c + j) * T + t] : 0u;
#pragma unroll
        for (int j = 0; j < 8; j++)
            if (c + j < c1) { M[(size_t)(c + j) * T + t] = run; run += v[j]; }
    }
    if (grp == TB_CP_GROUPS - 1) totals[t] = run;
}

// Step 2e.  One CTA: exclusive scan of the tile totals -> ranges (identifyTileRanges, rasterizer_impl.cu:
// 132-164: [start, end) per tile, (0, 0) for a tile nobody touches -- the reference memsets and never writes it).
// Totals pass through registers in slabs of 8 per thread, loaded and stored coalesced through shared memory.
//
// Also: tile_start[T + 1] (the exclusive prefix of every tile, empty ones included) and, per band, the tile rows at
// which the list-filling kernel cuts the band into TB_FILL_WARPS stripes of about equal instance counts (rows near
// the image centre are denser: equal row counts would leave most warps of a CTA waiting for one).
__global__ void __launch_bounds__(TB_SCAN_THREADS)
tile_ranges_kernel(const uint32_t* __restrict__ totals, TileBinPlan pl, uint2* __restrict__ ranges,
                   uint32_t* __restrict__ tile_start, uint32_t* __restrict__ stripes) {
    const uint32_t T = pl.T;
    constexpr uint32_t PER = 8, SLAB = PER * TB_SCAN_THREADS;
    __shared__ uint32_t s_v[SLAB + SLAB / 32];       // padded: thread t reads 8 consecutive words conflict-free
    __shared__ uint32_t s_w[32];
    const uint32_t t = threadIdx.x, lane = t & 31u, warp = t >> 5;
    uint32_t carry = 0;
    for (uint32_t base = 0; base < T; base += SLAB) {
        for (uint32_t j = 0; j < PER; j++) {
            const uint32_t i = j * TB_SCAN_THREADS + t;
            s_v[i + (i >> 5)] = base + i < T ? totals[base + i] : 0u;
        }
        __syncthreads();
        uint32_t v[PER], sum = 0;
#pragma unroll
        for (uint32_t j = 0; j < PER; j++) {
            const uint32_t i = t * PER + j;
            v[j] = s_v[i + (i >> 5)];
            sum += v[j];
        }
        const uint32_t inc = warp_incl_scan(sum, lane);
        if (lane == 31) s_w[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const uint32_t w = s_w[lane];
            s_w[lane] = warp_incl_scan(w, lane) - w;
        }
        __syncthreads();
        uint32_t run = carry + s_w[warp] + inc - sum;
#pragma unroll
        for (uint32_t j = 0; j < PER; j++) {
            const uint32_t i = t * PER + j;
            s_v[i + (i >> 5)] = run;                 // exclusive prefix = start of the tile's list
            run += v[j];
        }
        __syncthreads();
        if (t == TB_SCAN_THREADS - 1) s_w[0] = run;  // slab total + carry
        for (uint32_t j = 0; j < PER; j++) {
            const uint32_t i = j * TB_SCAN_THREADS + t;
            if (base + i < T) {
                const uint32_t start = s_v[i + (i >> 5)], c = totals[base + i];
                ranges[base + i] = c ? make_uint2(start, start + c) : make_uint2(0u, 0u);
                tile_start[base + i] = start;
            }
        }
        __syncthreads();
        carry = s_w[0];
        __syncthreads();
    }
    if (t == 0) tile_start[T] = carry;
    __syncthreads();                     // tile_start (written by this CTA) is visible to all of its threads
    // stripes[band][w], w = 0..TB_FILL_WARPS: first tile row of stripe w (relative cuts at equal instance counts)
    for (uint32_t i = t; i < pl.n_bands * (TB_FILL_WARPS + 1); i += TB_SCAN_THREADS) {
        const uint32_t band = i / (TB_FILL_WARPS + 1), w = i % (TB_FILL_WARPS + 1);
        const uint32_t r0 = band * pl.band_rows, r1 = min(pl.gy, r0 + pl.band_rows);
        const uint32_t lo_v = tile_start[r0 * pl.gx], hi_v = tile_start[r1 * pl.gx];
        const uint32_t target = lo_v + (uint32_t)((unsigned long long)(hi_v - lo_v) * w / TB_FILL_WARPS);
        uint32_t lo = r0, hi = r1;       // first row whose start reaches the target
        if (w == 0) hi = r0;
        else if (w == TB_FILL_WARPS) lo = r1;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (tile_start[mid * pl.gx] >= target) hi = mid; else lo = mid + 1;
        }
        stripes[i] = lo;
    }
}

// Step 3.  One CTA of TB_FILL_WARPS warps per unit.  run[tile] = next free slot of the tile's list for this chunk.
// Every warp owns a stripe of the band's tile rows (cut by tile_ranges_kernel at equal instance counts) and walks
// ALL Gaussians of the chunk in depth order, 32 at a time: each lane clips one rectangle to the stripe and the ones
// that reach it are compacted (in order) into a parameter list in shared memory; the warp then takes them one after
// the other, lanes = tiles of the clipped rectangle (row-major): slot = run[tile]++, point_list[slot] = id.
// Within one Gaussian all tiles differ, a tile belongs to one warp, and a warp takes Gaussians in depth order: every
// list is filled in order without any ranking.  Lanes beyond the rectangle do the same read-modify-write on a
// scratch word of their own (no branch in the step).  The warps of a CTA never wait for each other after the
// counters are set up.
#ifndef GSR_TB_FILL_BLOCKS
#define GSR_TB_FILL_BLOCKS 6            // CTAs per SM: 37 KB of shared memory each
#endif
// L2 eviction priorities of the kernel's three streams (bit 0: point_list stores evict-last, bit 1: the Gaussians'
// rectangles evict-first, bit 2: the count matrix and list starts evict-first).  point_list is written in partial
// sectors that are completed later by other CTAs: the longer they stay in L2 the fewer of them go to DRAM half-written.
// Measured (B200): policy 3 takes the kernel from 0.283 to 0.276 ms at C3 and from 1.37 to 1.13 ms at C4 (6M, 4K);
// bit 2 adds nothing.
#ifndef GSR_TB_FILL_POLICY
#define GSR_TB_FILL_POLICY 3
#endif
#if GSR_TB_FILL_POLICY && !defined(GSR_HOST_EMUL)
__device__ __forceinline__ uint64_t l2_evict_last() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ uint64_t l2_evict_first() {
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void stg_u32_hint(uint32_t* a, uint32_t v, uint64_t pol) {
    asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(a), "r"(v), "l"(pol) : "memory");
}
__device__ __forceinline__ uint4 ldg_u4_hint(const uint4* a, uint64_t pol) {
    uint4 v;
    asm volatile("ld.global.nc.L2::cache_hint.v4.u32 {%0, %1, %2, %3}, [%4], %5;"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(a), "l"(pol));
    return v;
}
__device__ __forceinline__ uint32_t ldg_u32_hint(const uint32_t* a, uint64_t pol) {
    uint32_t v;
    asm volatile("ld.global.nc.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(a), "l"(pol));
    return v;
}
#define TB_FILL_HINTS 1
#else
#define TB_FILL_HINTS 0
#endif
__global__ void __launch_bounds__(32 * TB_FILL_WARPS, GSR_TB_FILL_BLOCKS)
tile_fill_kernel(const uint4* __restrict__ srects, const uint32_t* __restrict__ bounds, TileBinPlan pl,
                 const uint32_t* __restrict__ M, const uint32_t* __restrict__ tile_start,
                 const uint32_t* __restrict__ stripes, uint32_t* __restrict__ point_list, uint32_t band0) {
    __shared__ uint32_t run[TB_MAX_CELLS + 32 * TB_FILL_WARPS];  // tile counters, then one scratch word per thread
    __shared__ __align__(16) uint4 s_par[TB_FILL_WARPS][32];     // {id, first cell, n | (gx - w) << 21, div_magic(w)}
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, chunk = blockIdx.x, band = blockIdx.y + band0;
    const uint32_t r0 = band * pl.band_rows, r1 = min(pl.gy, r0 + pl.band_rows);
    const uint32_t rows = r1 - r0, gx = pl.gx, cells = rows * gx;
    const uint32_t k0 = bounds[chunk], k1 = bounds[chunk + 1];
    if (k0 >= k1) return;
    {
        const uint32_t* m = M + (size_t)chunk * pl.T + (size_t)r0 * gx;
        const uint32_t* ts = tile_start + (size_t)r0 * gx;
#if TB_FILL_HINTS && (GSR_TB_FILL_POLICY & 4)
        const uint64_t once = l2_evict_first();
        for (uint32_t i = threadIdx.x; i < cells; i += 32 * TB_FILL_WARPS)
            run[i] = ts[i] + ldg_u32_hint(m + i, once);
#else
        for (uint32_t i = threadIdx.x; i < cells; i += 32 * TB_FILL_WARPS) run[i] = ts[i] + m[i];
#endif
    }
#if TB_FILL_HINTS && (GSR_TB_FILL_POLICY & 1)
    const uint64_t keep = l2_evict_last();
#endif
#if TB_FILL_HINTS && (GSR_TB_FILL_POLICY & 2)
    const uint64_t stream = l2_evict_first();
    auto load_rect = [&](uint32_t i) { return ldg_u4_hint(srects + i, stream); };
#else
    auto load_rect = [&](uint32_t i) { return srects[i]; };
#endif
    __syncthreads();
    const uint32_t s0 = stripes[band * (TB_FILL_WARPS + 1) + warp], s1 = stripes[band * (TB_FILL_WARPS + 1) + warp + 1];
    if (s0 >= s1) return;                // an empty stripe (one very dense row next to it)
    uint4* par = s_par[warp];
    uint32_t run_s = smem_u32(run);                           // shared-window address: the steps use 32-bit addressing
    uint32_t lane2 = lane + lane;
    const uint32_t dummy = TB_MAX_CELLS + threadIdx.x;
    const uint32_t lt = (1u << lane) - 1u;
#ifndef GSR_HOST_EMUL
    asm volatile("" : "+r"(run_s), "+r"(lane2));              // keep both in registers (ptxas otherwise rebuilds them per step)
#endif
    // one slot per (Gaussian, tile): lanes = tiles l, l + 32, ... of the clipped rectangle
    auto append = [&](const uint4& q, uint32_t n, uint32_t skip, uint32_t l, uint32_t l2) {
        const uint32_t ry = __umulhi(l2, q.w);
        const uint32_t cell = l < n ? q.y + l + ry * skip : dummy;
        const uint32_t a = run_s + 4u * cell;
        const uint32_t slot = lds_u32(a);
        sts_u32(a, slot + 1u);
#if TB_FILL_HINTS && (GSR_TB_FILL_POLICY & 1)
        if (l < n) stg_u32_hint(point_list + slot, q.x, keep);
#elif !defined(GSR_TB_EXPERIMENT_NO_STORE)
        if (l < n) point_list[slot] = q.x;
#else   // measurement variant only (profiles/build_variants.py): how much of the kernel is the scattered store
        if (l < n && q.x == 0xFFFFFFFFu) point_list[slot] = q.x;
#endif
        step_fence();                    // the next Gaussian's lanes may read the counters this one wrote
    };
    // two batches of 32 Gaussians are in flight ahead of the one being walked (a batch's steps are too short to
    // cover the latency of one load)
    uint4 next = k0 + lane < k1 ? load_rect(k0 + lane) : make_uint4(0u, 0u, 0u, 0u);
    uint4 next2 = k0 + 32 + lane < k1 ? load_rect(k0 + 32 + lane) : make_uint4(0u, 0u, 0u, 0u);
    for (uint32_t k = k0; k < k1; k += 32) {
        const uint4 g = next;
        next = next2;
        const uint32_t kn = k + 64 + lane;
        next2 = kn < k1 ? load_rect(kn) : make_uint4(0u, 0u, 0u, 0u);
        uint32_t x0, x1, ya, yb;
        clip_rect(make_uint2(g.x, g.y), s0, s1, x0, x1, ya, yb);
        const bool hit = x1 > x0 && yb > ya;
        const uint32_t bits = __ballot_sync(0xffffffffu, hit);
        if (bits == 0) continue;
        if (hit) {
            const uint32_t w = x1 - x0;
            par[__popc(bits & lt)] = make_uint4(g.w, (ya - r0) * gx + x0, (w * (yb - ya)) | ((gx - w) << 21), g.z);
        }
        __syncwarp();
        const uint32_t cnt = __popc(bits);
        for (uint32_t e = 0; e < cnt; e++) {
            const uint4 q = par[e];
            const uint32_t n = q.z & 0x1fffffu, skip = q.z >> 21;
            append(q, n, skip, lane, lane2);
            if (n > 32)                  // uniform and rare: a clipped rectangle of more than 32 tiles
                for (uint32_t l = lane + 32; l - lane < n; l += 32) append(q, n, skip, l, l + l);
        }
        __syncwarp();                    // everybody has read this batch's parameters
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
// Stage 1 of the forward after preprocess: num_rendered to the host, depth order, per-tile counts and ranges.
int launch_depth_order_and_count(const GsrParams& p, GeomState g, ImgState im, Scratch s, int32_t* num_rendered_host,
                                 void* num_rendered_event, cudaStream_t st) {
    const uint32_t P = (uint32_t)p.P;
    // {num_rendered (64-bit), V} were accumulated by the preprocess kernel: the host can size `bin` while the
    // device is still ordering (the reference blocks on this copy after its scan, rasterizer_impl.cu:313)
    if (num_rendered_host)
        GSR_CUDA_TRY(cudaMemcpyAsync(num_rendered_host, s.counters, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    if (num_rendered_event) GSR_CUDA_TRY(cudaEventRecord((cudaEvent_t)num_rendered_event, st));
    // depth_a holds depth_bits<<32 | id (written by preprocess). 4 stable LSD passes.
    uint64_t* a = s.depth_a;
    uint64_t* b = s.depth_b;
    {
        StageTimer tm(GSR_STAGE_DEPTH_SORT, st);
        if (P < OS_MAX_N) {
            int rc = radix_sort_onesweep(a, b, P, 32, 4, s, st);
            if (rc) return rc;
        } else {
            for (int pass = 0; pass < 4; pass++) {
                radix_pass(a, b, P, 32 + 8 * pass, s, st);
                uint64_t* tmp = a; a = b; b = tmp;
            }
        }
    }
    int rc = check_stage("depth sort", p.debug, st);
    if (rc) return rc;
    // a == s.depth_a again (even number of passes)
    const TileBinPlan pl = tile_bin_plan(p.P, p.W, p.H);
    {
        StageTimer tm(GSR_STAGE_TILE_COUNT, st);
        GSR_LAUNCH(tile_weights_kernel, pl.n_wblocks, 256, 0, st, a, s.rects, s.counters, pl.wblock, s.srects, s.tile_weights);
        GSR_LAUNCH(tile_bounds_kernel, 1, TB_SCAN_THREADS, 0, st, s.tile_weights, pl.n_wblocks, pl.wblock, s.counters,
                                                      pl.n_chunks, s.tile_bounds);
        GSR_LAUNCH(tile_count_kernel, dim3(pl.n_chunks, pl.n_bands), TB_COUNT_THREADS, 0, st, s.srects, s.tile_bounds, pl,
                                                      s.tile_chunk);
        GSR_LAUNCH(tile_chunk_prefix_kernel, div_up(pl.T, 32), 32 * TB_CP_GROUPS, 0, st, s.tile_chunk, pl, s.tile_total);
        GSR_LAUNCH(tile_ranges_kernel, 1, TB_SCAN_THREADS, 0, st, s.tile_total, pl, im.ranges, s.tile_start, s.tile_stripes);
    }
    return check_stage("tile counts", p.debug, st);
}

// Stage 2 of the forward before the blend: fill point_list.
int launch_tile_lists(const GsrParams& p, GeomState g, ImgState im, BinState b, Scratch s, int64_t R, cudaStream_t st) {
    if (R <= 0) return GSR_OK;
    const TileBinPlan pl = tile_bin_plan(p.P, p.W, p.H);
    {
        StageTimer tm(GSR_STAGE_TILE_FILL, st);
        // One launch per band of tile rows, in sequence: the lists being written at any time then belong to one band
        // (a contiguous part of point_list that fits L2 up to 4K), so partially written sectors are completed in L2.
        // With all bands in flight at once the same stores were read-modify-writes of DRAM sectors: 2.7 ms instead of
        // 0.7 ms of compute at C4 (6M Gaussians, 3840x2160, 334 MB of point_list).
        for (uint32_t band = 0; band < pl.n_bands; band++)
            GSR_LAUNCH(tile_fill_kernel, dim3(pl.n_chunks, 1), 32 * TB_FILL_WARPS, 0, st, s.srects,
                                                         s.tile_bounds, pl, s.tile_chunk, s.tile_start, s.tile_stripes, b.point_list, band);
    }
    return check_stage("tile lists", p.debug, st);
}

int launch_rebuild_keys(const GsrParams& p, GeomState g, ImgState im, BinState b, int64_t R,
                        uint64_t* keys_out, cudaStream_t st) {
    const uint32_t gx = div_up(p.W, TILE_X), gy = div_up(p.H, TILE_Y);
    if (R <= 0) return GSR_OK;
    GSR_LAUNCH(rebuild_keys_kernel, gx * gy, 128, 0, st, im.ranges, gx * gy, b.point_list, g.rec, keys_out);
    return check_stage("rebuild keys", true, st);
}

}  // namespace gsr
