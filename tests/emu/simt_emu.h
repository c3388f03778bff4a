// simt_emu.h — a small CPU SIMT emulator: compiles the repo's CUDA kernel headers with g++ and runs them
// block by block.  TEST INFRASTRUCTURE ONLY: it lets `pytest -m "not gpu"` execute the product kernel SOURCE
// against the oracle in the build container; it is never part of the product path and never a fallback.
//
// Model: one OS thread per warp, 32 ucontext fibers per warp (one per lane), round-robin between sync points.
// Warp intrinsics (full mask only) are barriers over the 32 lanes that exchange values through a double-buffered
// array; __syncthreads() is a std::barrier over the warps of the block.  Blocks run one after another, so
// `__shared__` variables are plain statics.  Atomics are __atomic builtins (warps are real threads).
#pragma once
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
#include <stdio.h>
#include <ucontext.h>
#include <algorithm>
#include <barrier>
#include <functional>
#include <memory>
#include <thread>
#include <vector>

#define DV_EMU 1
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline __attribute__((always_inline))
#define __shared__ static
#define __align__(n) __attribute__((aligned(n)))
#define __launch_bounds__(...)

struct uint2 { uint32_t x, y; };
struct __attribute__((aligned(16))) uint4 { uint32_t x, y, z, w; };
static inline uint2 make_uint2(uint32_t x, uint32_t y) { return uint2{ x, y }; }
static inline uint4 make_uint4(uint32_t x, uint32_t y, uint32_t z, uint32_t w) { return uint4{ x, y, z, w }; }
using std::min;
using std::max;

namespace emu {
struct Warp {
    ucontext_t sched;
    ucontext_t ctx[32];
    std::unique_ptr<char[]> stack[32];
    bool done[32] = {};
    int n_done = 0;
    int warp_id = 0;
    uint32_t phase = 0;                 // completed warp-level barriers
    int arrived = 0;
    unsigned long long xchg[2][32];
    uint32_t blk_phase = 0;             // completed block-level barriers
    int blk_arrived = 0;
};
inline thread_local Warp* tw = nullptr;
inline thread_local int tlane = 0;
inline thread_local unsigned t_block = 0;
inline unsigned g_block_dim = 1, g_grid_dim = 1;
inline std::barrier<>* g_block_bar = nullptr;
inline const std::function<void()>* g_body = nullptr;

inline void yield() { Warp* w = tw; swapcontext(&w->ctx[tlane], &w->sched); }

inline int warp_arrive_prepare() { return (int)(tw->phase & 1u); }
inline void warp_barrier() {
    Warp* w = tw;
    const uint32_t my = w->phase;
    if (++w->arrived + w->n_done >= 32) { w->arrived = 0; w->phase = my + 1; return; }
    while (w->phase == my) yield();
}
inline void block_barrier() {
    Warp* w = tw;
    const uint32_t my = w->blk_phase;
    if (++w->blk_arrived + w->n_done >= 32) {
        w->blk_arrived = 0;
        g_block_bar->arrive_and_wait();          // the warp's OS thread meets the other warps
        w->blk_phase = my + 1;
        return;
    }
    while (w->blk_phase == my) yield();
}
inline void lane_entry() {
    (*g_body)();
    Warp* w = tw;
    w->done[tlane] = true;
    w->n_done++;
    // a lane that exits counts as arrived for every later barrier
    if (w->arrived && w->arrived + w->n_done >= 32) { w->arrived = 0; w->phase++; }
    if (w->blk_arrived && w->blk_arrived + w->n_done >= 32) { w->blk_arrived = 0; g_block_bar->arrive_and_wait(); w->blk_phase++; }
    swapcontext(&w->ctx[tlane], &w->sched);
}
inline void run_warp(int warp_id, unsigned block) {
    Warp* w = new Warp();
    w->warp_id = warp_id;
    tw = w;
    t_block = block;
    const size_t STK = 256 * 1024;
    for (int l = 0; l < 32; l++) {
        w->stack[l].reset(new char[STK]);
        getcontext(&w->ctx[l]);
        w->ctx[l].uc_stack.ss_sp = w->stack[l].get();
        w->ctx[l].uc_stack.ss_size = STK;
        w->ctx[l].uc_link = &w->sched;
        makecontext(&w->ctx[l], (void (*)())lane_entry, 0);
    }
    while (w->n_done < 32)
        for (int l = 0; l < 32; l++)
            if (!w->done[l]) { tlane = l; swapcontext(&w->sched, &w->ctx[l]); }
    delete w;
    tw = nullptr;
}
// launch<<<grid, block>>>(body): blocks one after another, warps of a block as OS threads
inline void launch(unsigned grid, unsigned block_threads, const std::function<void()>& body) {
    g_block_dim = block_threads; g_grid_dim = grid; g_body = &body;
    const int n_warps = (int)(block_threads / 32);
    for (unsigned b = 0; b < grid; b++) {
        std::barrier<> bar(n_warps);
        g_block_bar = &bar;
        std::vector<std::thread> th;
        for (int wi = 0; wi < n_warps; wi++) th.emplace_back(run_warp, wi, b);
        for (auto& t : th) t.join();
    }
}
}  // namespace emu

struct emu_tid_x { operator unsigned() const { return (unsigned)(emu::tw->warp_id * 32 + emu::tlane); } };
struct emu_bid_x { operator unsigned() const { return emu::t_block; } };
struct emu_bdim_x { operator unsigned() const { return emu::g_block_dim; } };
struct emu_gdim_x { operator unsigned() const { return emu::g_grid_dim; } };
inline struct { emu_tid_x x; } threadIdx;
inline struct { emu_bid_x x; } blockIdx;
inline struct { emu_bdim_x x; } blockDim;
inline struct { emu_gdim_x x; } gridDim;

#include "simt_emu_intrinsics.h"
