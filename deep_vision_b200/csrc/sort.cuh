// sort.cuh — stable LSD radix sort of (u64 key, u32 value) pairs (kernel stage K2: index builds).
// 8 bits per pass: per-block digit histogram -> per-digit scan over blocks -> stable scatter staged through
// shared memory (block-local digit sort, then contiguous digit runs to HBM).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define RS_THREADS 256
#define RS_WARPS 8
#define RS_PER_LANE 16                              // items per lane, held in registers across the phases
#define RS_ITEMS_PER_WARP (32 * RS_PER_LANE)
#define RS_TILE (RS_WARPS * RS_ITEMS_PER_WARP)      // 4096 items per block

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

// Same histogram from ONE BYTE per item: the scatter of the previous pass wrote the digit every item will have in this
// pass next to its new position, so passes after the first read n bytes instead of the 8 n bytes of the keys.
// A CTA counts RS_HT consecutive tiles (their loads issued together, one histogram each): a 4 KB tile per CTA left the kernel
// bound by CTA turnover (0.24 ms per pass for 0.44 GB).
#define RS_HT 4
static __global__ void __launch_bounds__(RS_THREADS) k_rs_hist8(const uint8_t* __restrict__ digits, int64_t n, uint32_t* __restrict__ hist /*[256][nb]*/, int64_t nb) {
    __shared__ uint32_t h[RS_HT][256];
#pragma unroll
    for (int t = 0; t < RS_HT; t++) h[t][threadIdx.x] = 0;
    __syncthreads();
    const int64_t tile0 = (int64_t)blockIdx.x * RS_HT;
    uint4 v[RS_HT];
#pragma unroll
    for (int t = 0; t < RS_HT; t++) {                                       // RS_TILE is a multiple of 16; digits is 16-byte aligned
        const int64_t g = (tile0 + t) * RS_TILE + (int64_t)threadIdx.x * 16;
        v[t] = (tile0 + t < nb && g + 16 <= n) ? *reinterpret_cast<const uint4*>(digits + g) : make_uint4(0u, 0u, 0u, 0u);
    }
#pragma unroll
    for (int t = 0; t < RS_HT; t++) {
        if (tile0 + t >= nb) break;
        const int64_t g = (tile0 + t) * RS_TILE + (int64_t)threadIdx.x * 16;
        if (g + 16 <= n) {
            const uint32_t w[4] = { v[t].x, v[t].y, v[t].z, v[t].w };
#pragma unroll
            for (int q = 0; q < 4; q++)
#pragma unroll
                for (int b = 0; b < 4; b++) atomicAdd(&h[t][(w[q] >> (8 * b)) & 255u], 1u);
        } else {
            for (int64_t i = g; i < n && i < g + 16; i++) atomicAdd(&h[t][digits[i]], 1u);
        }
    }
    __syncthreads();
#pragma unroll
    for (int t = 0; t < RS_HT; t++) if (tile0 + t < nb) hist[(int64_t)threadIdx.x * nb + tile0 + t] = h[t][threadIdx.x];
}
static inline size_t dv_radix_hist_words(int64_t n) {                      // uint32 words the sort needs behind `hist`
    const int64_t nb = (n + RS_TILE - 1) / RS_TILE;
    return (size_t)(256 * nb + 512) + 4 + (size_t)((n + 15) / 16) * 4;      // histogram matrix + totals + digit bases | 16-byte aligned digit bytes
}

// block d scans hist[d][0..nb) in place (exclusive, within the digit) and leaves the digit's total in totals[d].
// Four consecutive tiles per thread: the row of a 440 M item sort has 107 k entries, and a step of the loop is three barriers long.
static __global__ void __launch_bounds__(1024) k_rs_scan(uint32_t* __restrict__ hist, int64_t nb, uint32_t* __restrict__ totals /*[256]*/) {
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t carry;
    const int d = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* row = hist + (int64_t)d * nb;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int64_t b0 = 0; b0 < nb; b0 += 4096) {
        const int64_t i = b0 + (int64_t)threadIdx.x * 4;
        uint32_t v[4];
#pragma unroll
        for (int k = 0; k < 4; k++) v[k] = i + k < nb ? row[i + k] : 0u;
        const uint32_t mine = v[0] + v[1] + v[2] + v[3];
        uint32_t x = mine;
        for (int s = 1; s < 32; s <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, s); if (lane >= s) x += y; }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        uint32_t before = carry;
        for (int w = 0; w < warp; w++) before += warp_tot[w];
        uint32_t run = before + x - mine;
#pragma unroll
        for (int k = 0; k < 4; k++) { if (i + k < nb) row[i + k] = run; run += v[k]; }
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) totals[d] = carry;
}

// exclusive scan of the 256 digit totals: the global start of every digit.  The scatter adds it to the per-tile offsets itself, so
// no kernel rewrites the 256 x nb histogram matrix a second time.
static __global__ void __launch_bounds__(256) k_rs_base(const uint32_t* __restrict__ totals, uint32_t* __restrict__ dbase) {
    __shared__ uint32_t t[256];
    t[threadIdx.x] = totals[threadIdx.x];
    __syncthreads();
    uint32_t b = 0;
    for (int j = 0; j < (int)threadIdx.x; j++) b += t[j];
    dbase[threadIdx.x] = b;
}
static inline uint32_t* dv_radix_dbase_ptr(uint32_t* hist, int64_t n) { return hist + 256 * ((n + RS_TILE - 1) / RS_TILE) + 256; }

// Stable scatter through shared memory: every lane loads its 16 keys once (independent loads, kept in registers),
// the block ranks its 4096 items by digit (warps own contiguous sub-tiles, 32 items per step ranked with
// __match_any_sync), places them digit-sorted in shared memory and then streams them out, so every digit leaves the
// block as ONE contiguous run instead of 32-byte fragments: DRAM traffic = one read + one write of the data per pass.
struct RsSmem {
    unsigned long long key[RS_TILE];
    uint32_t val[RS_TILE];
    uint32_t cnt[RS_WARPS][256];     // per warp and digit: count, then block-local position of the warp's next item
    uint32_t delta[256];             // global offset of the block's digit run minus its block-local start
    uint32_t wsum[RS_WARPS];
};
// The tile's items are in registers (item j of a lane = tile position warp * 512 + j * 32 + lane); everything after the load.
static __device__ __forceinline__ void rs_scatter_body(unsigned long long (&k)[RS_PER_LANE], uint32_t (&v)[RS_PER_LANE], RsSmem& sm,
                                                       unsigned long long* __restrict__ okeys, uint32_t* __restrict__ ovals, int64_t n, int shift,
                                                       const uint32_t* __restrict__ hist, int64_t nb, uint8_t* __restrict__ next_digit, int next_shift) {
    const uint32_t* __restrict__ dbase = hist + 256 * nb + 256;               // dv_radix_dbase_ptr
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t bbase = (int64_t)blockIdx.x * RS_TILE;
    const int64_t wbase = bbase + (int64_t)warp * RS_ITEMS_PER_WARP;
    for (int i = threadIdx.x; i < RS_WARPS * 256; i += RS_THREADS) (&sm.cnt[0][0])[i] = 0;
    __syncthreads();
    // phase 1: per-warp digit counts
#pragma unroll
    for (int j = 0; j < RS_PER_LANE; j++)
        if (wbase + j * 32 + lane < n) atomicAdd(&sm.cnt[warp][(uint32_t)(k[j] >> shift) & 255u], 1u);
    __syncthreads();
    // phase 2: thread d: block-local start of digit d (exclusive scan of the digit totals), per-warp positions, delta
    {
        const int d = threadIdx.x;
        uint32_t tot = 0;
        for (int w = 0; w < RS_WARPS; w++) tot += sm.cnt[w][d];
        uint32_t x = tot;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) sm.wsum[warp] = x;
        __syncthreads();
        uint32_t before = 0;
        for (int w = 0; w < warp; w++) before += sm.wsum[w];
        const uint32_t lstart = before + x - tot;
        uint32_t run = lstart;
        for (int w = 0; w < RS_WARPS; w++) { const uint32_t c = sm.cnt[w][d]; sm.cnt[w][d] = run; run += c; }
        sm.delta[d] = hist[(int64_t)d * nb + blockIdx.x] + dbase[d] - lstart;
    }
    __syncthreads();
    // phase 3: stable placement into shared memory, 32 items per step.  The lanes holding the same digit are found
    // with 8 ballots (one per digit bit) instead of MATCH.ANY: on uniformly distributed digits the match instruction
    // was 39 % of the kernel's stall samples (its latency grows with the number of distinct values in the warp).
#pragma unroll
    for (int j0 = 0; j0 < RS_PER_LANE; j0 += 4) {
        uint32_t dd[4], peers[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const bool ok = wbase + (j0 + u) * 32 + lane < n;
            dd[u] = (uint32_t)(k[j0 + u] >> shift) & 255u;
            uint32_t m = __ballot_sync(0xffffffffu, ok);
#pragma unroll
            for (int b = 0; b < 8; b++) {                 // lanes whose digit bit b equals mine: and, setp, vote, @!p not (+ the and below)
                uint32_t same;
                asm volatile("{\n\t.reg .pred p;\n\tand.b32 %0, %1, %2;\n\tsetp.ne.u32 p, %0, 0;\n\t"
                             "vote.sync.ballot.b32 %0, p, 0xffffffff;\n\t@!p not.b32 %0, %0;\n\t}"
                             : "=r"(same) : "r"(dd[u]), "r"(1u << b));
                m &= same;
            }
            peers[u] = m;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const bool ok = wbase + (j0 + u) * 32 + lane < n;
            const int rank = __popc(peers[u] & ((1u << lane) - 1u));
            uint32_t off = 0;
            if (ok) off = sm.cnt[warp][dd[u]];
            __syncwarp();
            if (ok && rank == 0) sm.cnt[warp][dd[u]] = off + __popc(peers[u]);
            __syncwarp();
            if (ok) { sm.key[off + rank] = k[j0 + u]; sm.val[off + rank] = v[j0 + u]; }
        }
    }
    __syncthreads();
    // phase 4: stream the digit-sorted tile out; item i of digit d goes to delta[d] + i
    const int cnt_tile = (int)min((int64_t)RS_TILE, n - bbase);
    for (int i = threadIdx.x; i < cnt_tile; i += RS_THREADS) {
        const unsigned long long kk = sm.key[i];
        const uint32_t pos = sm.delta[(uint32_t)(kk >> shift) & 255u] + (uint32_t)i;
        okeys[pos] = kk; ovals[pos] = sm.val[i];
        if (next_digit) next_digit[pos] = (uint8_t)(kk >> next_shift);       // the digit of the NEXT pass, for its histogram
    }
}
static __global__ void __launch_bounds__(RS_THREADS, 3) k_rs_scatter(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                               unsigned long long* __restrict__ okeys, uint32_t* __restrict__ ovals, int64_t n, int shift,
                                                               const uint32_t* __restrict__ hist, int64_t nb, uint8_t* __restrict__ next_digit, int next_shift) {
    extern __shared__ __align__(16) unsigned char rs_raw[];
    RsSmem& sm = *reinterpret_cast<RsSmem*>(rs_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t wbase = (int64_t)blockIdx.x * RS_TILE + (int64_t)warp * RS_ITEMS_PER_WARP;
    unsigned long long k[RS_PER_LANE];
    uint32_t v[RS_PER_LANE];
#pragma unroll
    for (int j = 0; j < RS_PER_LANE; j++) { const int64_t g = wbase + j * 32 + lane; k[j] = g < n ? keys[g] : 0ULL; }
#pragma unroll
    for (int j = 0; j < RS_PER_LANE; j++) { const int64_t g = wbase + j * 32 + lane; v[j] = g < n ? vals[g] : 0u; }
    rs_scatter_body(k, v, sm, okeys, ovals, n, shift, hist, nb, next_digit, next_shift);
}

// Sorts n pairs by bits [begin_bit, end_bit) of the key.  keys/vals and tkeys/tvals are ping-pong
// buffers; returns which pair of buffers holds the result (0 = keys/vals, 1 = tkeys/tvals).
// hist needs dv_radix_hist_words(n) uint32 and 16-byte alignment.
static inline uint32_t* dv_radix_totals_ptr(uint32_t* hist, int64_t n) { return hist + 256 * ((n + RS_TILE - 1) / RS_TILE); }
static inline uint8_t* dv_radix_digits_ptr(uint32_t* hist, int64_t n) {
    const int64_t nb = (n + RS_TILE - 1) / RS_TILE;
    return reinterpret_cast<uint8_t*>(hist + (((size_t)(256 * nb + 512) + 3) & ~size_t(3)));    // 16-byte aligned (hist is)
}
// first_digits_ready: an earlier pass (the caller's own) already wrote the digit bytes of the pass at begin_bit
static inline int dv_radix_sort_pairs(unsigned long long* keys, uint32_t* vals, unsigned long long* tkeys, uint32_t* tvals,
                                      int64_t n, int begin_bit, int end_bit, uint32_t* hist, cudaStream_t st, int64_t* launches,
                                      bool first_digits_ready = false) {
    if (n <= 1) return 0;
    cudaFuncSetAttribute(k_rs_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem));
    const int64_t nb = (n + RS_TILE - 1) / RS_TILE;
    uint32_t* totals = dv_radix_totals_ptr(hist, n);
    uint8_t* digits = dv_radix_digits_ptr(hist, n);
    int cur = 0;
    for (int shift = begin_bit; shift < end_bit; shift += 8) {
        unsigned long long* ik = cur ? tkeys : keys; uint32_t* iv = cur ? tvals : vals;
        unsigned long long* ok = cur ? keys : tkeys; uint32_t* ov = cur ? vals : tvals;
        const bool more = shift + 8 < end_bit;
        if (shift == begin_bit && !first_digits_ready) k_rs_hist<<<(unsigned)nb, RS_THREADS, 0, st>>>(ik, n, shift, hist, nb);
        else k_rs_hist8<<<(unsigned)((nb + RS_HT - 1) / RS_HT), RS_THREADS, 0, st>>>(digits, n, hist, nb);
        k_rs_scan<<<256, 1024, 0, st>>>(hist, nb, totals);
        k_rs_base<<<1, 256, 0, st>>>(totals, dv_radix_dbase_ptr(hist, n));
        k_rs_scatter<<<(unsigned)nb, RS_THREADS, sizeof(RsSmem), st>>>(ik, iv, ok, ov, n, shift, hist, nb, more ? digits : nullptr, shift + 8);
        if (launches) *launches += 4;
        cur ^= 1;
    }
    return cur;
}
