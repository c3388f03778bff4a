// sort.cuh — stable LSD radix sort of (u64 key, u32 value) pairs (kernel stage K2: index builds).
// 8 bits per pass: per-block digit histogram -> per-digit scan over blocks -> stable scatter in which
// every warp owns a contiguous sub-tile and ranks its 32 items per step with __match_any_sync.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RS_THREADS 256
#define RS_WARPS 8
#define RS_ITEMS_PER_WARP 1024
#define RS_TILE (RS_WARPS * RS_ITEMS_PER_WARP)      // 8192 items per block

static __global__ void __launch_bounds__(RS_THREADS) k_rs_hist(const unsigned long long* __restrict__ keys, int64_t n, int shift,
                                                         uint32_t* __restrict__ hist /*[256][nb]*/, int64_t nb) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
    for (int i = threadIdx.x; i < RS_TILE; i += RS_THREADS) {
        const int64_t g = base + i;
        if (g < n) atomicAdd(&h[(uint32_t)(keys[g] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(int64_t)threadIdx.x * nb + blockIdx.x] = h[threadIdx.x];
}

// block d scans hist[d][0..nb) in place (exclusive) and adds the base of digit d
static __global__ void __launch_bounds__(1024) k_rs_scan(uint32_t* __restrict__ hist, int64_t nb, uint32_t* __restrict__ totals /*[256]*/, int phase) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry;
    const int d = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* row = hist + (int64_t)d * nb;
    if (phase == 0) {                                   // within-digit exclusive scan, total to totals[d]
        if (threadIdx.x == 0) carry = 0;
        __syncthreads();
        for (int64_t b0 = 0; b0 < nb; b0 += 1024) {
            const int64_t i = b0 + threadIdx.x;
            const uint32_t v = i < nb ? row[i] : 0u;
            uint32_t x = v;
            for (int s = 1; s < 32; s <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, s); if (lane >= s) x += y; }
            if (lane == 31) warp_tot[warp] = x;
            __syncthreads();
            uint32_t before = carry;
            for (int w = 0; w < warp; w++) before += warp_tot[w];
            if (i < nb) row[i] = before + x - v;
            __syncthreads();
            if (threadIdx.x == 1023) carry = before + x;
            __syncthreads();
        }
        if (threadIdx.x == 0) totals[d] = carry;
    } else {                                            // add the digit base (exclusive scan of totals)
        uint32_t base = 0;
        for (int j = 0; j < d; j++) base += totals[j];
        for (int64_t i = threadIdx.x; i < nb; i += 1024) row[i] += base;
    }
}

static __global__ void __launch_bounds__(RS_THREADS) k_rs_scatter(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                            unsigned long long* __restrict__ okeys, uint32_t* __restrict__ ovals, int64_t n, int shift,
                                                            const uint32_t* __restrict__ hist, int64_t nb) {
    __shared__ uint32_t cnt[RS_WARPS][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&cnt[0][0])[i] = 0;
    __syncthreads();
    const int64_t wbase = (int64_t)blockIdx.x * RS_TILE + (int64_t)warp * RS_ITEMS_PER_WARP;
    // phase 1: per-warp digit counts
    for (int s = 0; s < RS_ITEMS_PER_WARP; s += 32) {
        const int64_t g = wbase + s + lane;
        if (g < n) atomicAdd(&cnt[warp][(uint32_t)(keys[g] >> shift) & 255u], 1u);
    }
    __syncthreads();
    // phase 2: cnt[w][d] -> global offset of warp w's first item with digit d
    {
        const int d = threadIdx.x;
        uint32_t run = hist[(int64_t)d * nb + blockIdx.x];
        for (int w = 0; w < RS_WARPS; w++) { const uint32_t c = cnt[w][d]; cnt[w][d] = run; run += c; }
    }
    __syncthreads();
    // phase 3: stable scatter, 32 items per step
    for (int s = 0; s < RS_ITEMS_PER_WARP; s += 32) {
        const int64_t g = wbase + s + lane;
        const bool v = g < n;
        const unsigned long long k = v ? keys[g] : 0ULL;
        const uint32_t d = v ? ((uint32_t)(k >> shift) & 255u) : 256u + lane;      // invalid lanes never match
        const uint32_t peers = __match_any_sync(0xffffffffu, d);
        const int rank = __popc(peers & ((1u << lane) - 1u));
        uint32_t off = 0;
        if (v) off = cnt[warp][d];
        __syncwarp();
        if (v && rank == 0) cnt[warp][d] = off + __popc(peers);
        __syncwarp();
        if (v) { okeys[off + rank] = k; ovals[off + rank] = vals[g]; }
    }
}

// Sorts n pairs by bits [begin_bit, end_bit) of the key.  keys/vals and tkeys/tvals are ping-pong
// buffers; returns which pair of buffers holds the result (0 = keys/vals, 1 = tkeys/tvals).
// hist needs 256 * ceil(n / RS_TILE) + 256 uint32.
static inline int dv_radix_sort_pairs(unsigned long long* keys, uint32_t* vals, unsigned long long* tkeys, uint32_t* tvals,
                                      int64_t n, int begin_bit, int end_bit, uint32_t* hist, cudaStream_t st, int64_t* launches) {
    if (n <= 1) return 0;
    const int64_t nb = (n + RS_TILE - 1) / RS_TILE;
    uint32_t* totals = hist + 256 * nb;
    int cur = 0;
    for (int shift = begin_bit; shift < end_bit; shift += 8) {
        unsigned long long* ik = cur ? tkeys : keys; uint32_t* iv = cur ? tvals : vals;
        unsigned long long* ok = cur ? keys : tkeys; uint32_t* ov = cur ? vals : tvals;
        k_rs_hist<<<(unsigned)nb, RS_THREADS, 0, st>>>(ik, n, shift, hist, nb);
        k_rs_scan<<<256, 1024, 0, st>>>(hist, nb, totals, 0);
        k_rs_scan<<<256, 1024, 0, st>>>(hist, nb, totals, 1);
        k_rs_scatter<<<(unsigned)nb, RS_THREADS, 0, st>>>(ik, iv, ok, ov, n, shift, hist, nb);
        if (launches) *launches += 4;
        cur ^= 1;
    }
    return cur;
}
