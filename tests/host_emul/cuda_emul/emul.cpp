// (Switching: a fiber is entered once through makecontext / swapcontext; every later switch is a
// _setjmp / _longjmp pair, which unlike swapcontext makes no signal-mask system call — build without _FORTIFY_SOURCE,
// whose longjmp check rejects jumps between stacks.)
// emul.cpp — cooperative fiber scheduler behind tests/host_emul/cuda_emul/cuda_runtime.h (TEST INFRASTRUCTURE).
// One block at a time; its threads are ucontext fibers resumed round-robin.  A fiber runs until it finishes or
// blocks in a barrier (block-wide: __syncthreads; warp-wide: every *_sync primitive).  Threads that have exited no
// longer count for barriers (CUDA semantics).  If no fiber can make progress the emulation aborts (a barrier that
// not all live threads reach is a bug in the kernel on the GPU too).
#include <setjmp.h>
#include <stdio.h>
#include <stdlib.h>
#include <ucontext.h>

#include <vector>

#include "cuda_runtime.h"

namespace emu {

Idx g_threadIdx, g_blockIdx, g_blockDim, g_gridDim;

namespace {
constexpr size_t STACK = 256 * 1024;

struct Warp {
    uint32_t alive = 0;            // lanes that have not exited
    uint32_t arrived = 0;          // lanes waiting in the current warp barrier
    uint32_t gen = 0;
    uint32_t slot[32];
    uint32_t ballot = 0, ballot_out = 0;
};

struct Fiber {
    ucontext_t ctx;
    jmp_buf jb;
    char* stack = nullptr;
    bool done = false, started = false;
    Idx tid;
    unsigned linear = 0;
};

struct Block {
    std::vector<Fiber> fibers;
    std::vector<Warp> warps;
    ucontext_t sched;
    jmp_buf sched_jb;
    unsigned cur = 0, alive = 0;
    unsigned bar_arrived = 0, bar_gen = 0;
    int bar_count = 0, bar_count_out = 0;
    unsigned long long progress = 0;
    const std::function<void()>* body = nullptr;
};

Block* g_blk = nullptr;

// GSR_EMUL_SCHED = "reverse" | "random[:seed]": order in which the runnable fibers of a block are resumed in each
// scheduling pass (default: thread order).  A result that changes with the order means the kernel lacks a barrier
// between a write and another thread's read — a race on the GPU.
int g_sched_mode = -1;
uint64_t g_rng = 0x9E3779B97F4A7C15ull;
void sched_init() {
    if (g_sched_mode >= 0) return;
    const char* e = getenv("GSR_EMUL_SCHED");
    g_sched_mode = 0;
    if (e && !strncmp(e, "reverse", 7)) g_sched_mode = 1;
    if (e && !strncmp(e, "random", 6)) {
        g_sched_mode = 2;
        if (e[6] == ':') g_rng ^= strtoull(e + 7, nullptr, 10) * 0xD1342543DE82EF95ull;
    }
}
uint64_t rng() { g_rng ^= g_rng << 13; g_rng ^= g_rng >> 7; g_rng ^= g_rng << 17; return g_rng; }

void yield() {
    Block* b = g_blk;
    if (!_setjmp(b->fibers[b->cur].jb)) _longjmp(b->sched_jb, 1);
}

void trampoline() {
    Block* b = g_blk;
    (*b->body)();
    Fiber& f = b->fibers[b->cur];
    f.done = true;
    b->alive--;
    b->warps[f.linear >> 5].alive &= ~(1u << (f.linear & 31));
    b->progress++;
    _longjmp(b->sched_jb, 1);
}

// Warp barrier over the live lanes of `mask`; `last` runs once, by the lane that completes the barrier.
template <typename F> void warp_barrier(uint32_t mask, F last) {
    Block* b = g_blk;
    const unsigned lin = b->fibers[b->cur].linear;
    Warp& w = b->warps[lin >> 5];
    const uint32_t me = 1u << (lin & 31);
    const uint32_t gen = w.gen;
    w.arrived |= me;
    b->progress++;
    for (;;) {
        const uint32_t need = mask & w.alive;
        if (w.gen != gen) return;
        if ((w.arrived & need) == need) {
            last(w);
            w.arrived = 0;
            w.gen++;
            b->progress++;
            return;
        }
        yield();
    }
}
}  // namespace

void spin() { yield(); }

uint32_t warp_alive() { return g_blk->warps[g_blk->fibers[g_blk->cur].linear >> 5].alive; }

unsigned lane_id() { return g_blk->fibers[g_blk->cur].linear & 31u; }

void warp_sync(uint32_t mask) { warp_barrier(mask, [](Warp&) {}); }

uint32_t warp_shfl(uint32_t mask, uint32_t value, int src_lane) {
    Block* b = g_blk;
    const unsigned lin = b->fibers[b->cur].linear;
    Warp& w = b->warps[lin >> 5];
    w.slot[lin & 31] = value;
    warp_barrier(mask, [](Warp&) {});
    const uint32_t r = w.slot[src_lane & 31];
    warp_barrier(mask, [](Warp&) {});          // nobody overwrites a slot before every lane has read
    return r;
}

uint32_t warp_ballot(uint32_t mask, int pred) {
    Block* b = g_blk;
    const unsigned lin = b->fibers[b->cur].linear;
    Warp& w = b->warps[lin >> 5];
    if (pred) w.ballot |= 1u << (lin & 31);
    warp_barrier(mask, [](Warp& x) { x.ballot_out = x.ballot; x.ballot = 0; });
    const uint32_t r = w.ballot_out & mask;
    warp_barrier(mask, [](Warp&) {});
    return r;
}

void syncthreads() { (void)syncthreads_count(0); }

int syncthreads_count(int pred) {
    Block* b = g_blk;
    const unsigned gen = b->bar_gen;
    b->bar_arrived++;
    if (pred) b->bar_count++;
    b->progress++;
    for (;;) {
        if (b->bar_gen != gen) return b->bar_count_out;
        if (b->bar_arrived >= b->alive) {
            b->bar_count_out = b->bar_count;
            b->bar_count = 0;
            b->bar_arrived = 0;
            b->bar_gen++;
            b->progress++;
            return b->bar_count_out;
        }
        yield();
    }
}

void launch(dim3 grid, dim3 block, const std::function<void()>& body) {
    const unsigned n = block.x * block.y * block.z;
    sched_init();
    std::vector<unsigned> order(n);
    for (unsigned t = 0; t < n; t++) order[t] = g_sched_mode == 1 ? n - 1 - t : t;
    Block blk;
    blk.body = &body;
    blk.fibers.resize(n);
    for (auto& f : blk.fibers) f.stack = (char*)malloc(STACK);
    g_blockDim = {block.x, block.y, block.z};
    g_gridDim = {grid.x, grid.y, grid.z};
    Block* outer = g_blk;
    g_blk = &blk;
    for (unsigned bz = 0; bz < grid.z; bz++)
        for (unsigned by = 0; by < grid.y; by++)
            for (unsigned bx = 0; bx < grid.x; bx++) {
                g_blockIdx = {bx, by, bz};
                blk.warps.assign((n + 31) / 32, Warp());
                blk.alive = n;
                blk.bar_arrived = blk.bar_gen = 0;
                blk.bar_count = 0;
                for (unsigned t = 0; t < n; t++) {
                    Fiber& f = blk.fibers[t];
                    f.done = false;
                    f.started = false;
                    f.linear = t;
                    f.tid = {t % block.x, (t / block.x) % block.y, t / (block.x * block.y)};
                    blk.warps[t >> 5].alive |= 1u << (t & 31);
                    getcontext(&f.ctx);
                    f.ctx.uc_stack.ss_sp = f.stack;
                    f.ctx.uc_stack.ss_size = STACK;
                    f.ctx.uc_link = nullptr;
                    makecontext(&f.ctx, trampoline, 0);
                }
                while (blk.alive) {
                    const unsigned long long before = blk.progress;
                    if (g_sched_mode == 2)
                        for (unsigned i = n - 1; i > 0; i--) std::swap(order[i], order[rng() % (i + 1)]);
                    for (unsigned k = 0; k < n; k++) {
                        const unsigned t = order[k];
                        Fiber& f = blk.fibers[t];
                        if (f.done) continue;
                        blk.cur = t;
                        g_threadIdx = f.tid;
                        if (!_setjmp(blk.sched_jb)) {
                            if (f.started) _longjmp(f.jb, 1);
                            f.started = true;
                            setcontext(&f.ctx);              // first entry: switch to the fiber's own stack
                        }
                    }
                    if (blk.progress == before) {
                        fprintf(stderr, "cuda_emul: deadlock in block (%u,%u,%u): %u live threads wait at barriers "
                                        "that not all of them reach\n", bx, by, bz, blk.alive);
                        abort();
                    }
                }
            }
    g_blk = outer;
    for (auto& f : blk.fibers) free(f.stack);
}

}  // namespace emu
