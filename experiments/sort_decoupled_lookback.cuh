// MEASURED AND NOT ADOPTED (round 1, session 10): single-kernel radix passes with decoupled look-back (Onesweep style).
// One up-front read of the keys gives the global digit histograms of all passes (k_rs_ghist, 0.90 ms at 440 M postings);
// tiles take tickets, publish (aggregate | inclusive prefix) status words and walk back over their predecessors, four
// status loads per round trip; a spin limit reports a stuck look-back through an error word instead of hanging.
// Bit-exact on tests/test_index.py, test_gate.py, test_parallel.py.  Config-3 build (1 M chunks, 440 M postings):
// 21.65 ms vs 22.0 ms for the adopted per-pass histogram + scan + scatter: the four k_rs_hist + eight k_rs_scan launches
// (4.2 ms) disappear, but every scatter pass grows from ~2.6 to 3.3-3.6 ms (2.9 us per tile: the CTA sits at the phase-2
// barrier while its 256 digit threads wait for the predecessors' words), so the net gain is 1.6 % for spin-waits in
// production code.  A second variant advanced the look-back one predecessor per poll BETWEEN the ranking steps of phase 3
// (state machine per digit thread, no blocking before the final barrier): 21.45 ms; ncu: 3.0-3.3 ms and 1.88 G warp
// instructions per scatter pass against ~2.65 ms / 1.73 G without look-back — with 444 tiles in flight and a tile living
// ~12 us, the chain back to the first inclusive prefix costs more L2 round trips than the histogram kernel it replaces.
// What would make it pay: larger tiles per CTA (fewer tiles in flight per unit of work) or a ninth warp that does a
// 4-wide look-back while the other eight rank (needs <= 72 registers to keep 3 CTAs per SM).
// sort.cuh — stable LSD radix sort of (u64 key, u32 value) pairs (kernel stage K2: index builds).
// 8 bits per pass.  ONE read of the keys up front gives the global digit histogram of every pass (k_rs_ghist; digit
// counts do not depend on the order of the items); a pass is then a single kernel: tiles take tickets in order,
// rank their 4096 items by digit in shared memory, publish their digit counts and find the counts of all earlier
// tiles by decoupled look-back over the published words (aggregate -> inclusive prefix), and stream contiguous digit
// runs to HBM.  No per-pass histogram kernel and no scan kernels: a pass reads and writes the data exactly once.
#pragma once
#include <cuda_runtime.h>
#include <algorithm>
#include <stdint.h>

#define RS_THREADS 256
#define RS_WARPS 8
#define RS_PER_LANE 16                              // items per lane, held in registers across the phases
#define RS_ITEMS_PER_WARP (32 * RS_PER_LANE)
#define RS_TILE (RS_WARPS * RS_ITEMS_PER_WARP)      // 4096 items per block

#define RS_MAX_PASSES 8
#define RS_FLAG_AGG 0x4000000000000000ULL           // status word = flag (2 bits) | count (62 bits)
#define RS_FLAG_INC 0x8000000000000000ULL
#define RS_VALUE_MASK 0x3FFFFFFFFFFFFFFFULL
#define RS_SPIN_LIMIT (1u << 22)

// auxiliary words behind the tile-status matrix: [ghist 8 x 256][gbase 8 x 256][tickets 8][err 1]
static inline size_t dv_radix_hist_words(int64_t n) {
    const int64_t nb = (n + RS_TILE - 1) / RS_TILE;
    return (size_t)(512 * (nb + 1)) + 2 * RS_MAX_PASSES * 256 + 16;        // 64-bit status words first (8-byte aligned)
}

// global digit histogram of every pass in one read of the keys
static __global__ void __launch_bounds__(RS_THREADS) k_rs_ghist(const unsigned long long* __restrict__ keys, int64_t n, int begin_bit, int npass,
                                                          uint32_t* __restrict__ ghist /*[npass][256]*/) {
    __shared__ uint32_t h[RS_MAX_PASSES][256];
    for (int i = threadIdx.x; i < RS_MAX_PASSES * 256; i += RS_THREADS) (&h[0][0])[i] = 0;
    __syncthreads();
    for (int64_t g = (int64_t)blockIdx.x * RS_THREADS + threadIdx.x; g < n; g += (int64_t)gridDim.x * RS_THREADS) {
        const unsigned long long k = keys[g] >> begin_bit;
        for (int p = 0; p < npass; p++) atomicAdd(&h[p][(uint32_t)(k >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < npass * 256; i += RS_THREADS) { const uint32_t c = (&h[0][0])[i]; if (c) atomicAdd(&ghist[i], c); }
}
// block p: gbase[p][d] = number of items whose digit of pass p is < d
static __global__ void __launch_bounds__(256) k_rs_gscan(const uint32_t* __restrict__ ghist, uint32_t* __restrict__ gbase) {
    __shared__ uint32_t wsum[8];
    const int d = threadIdx.x, lane = d & 31, warp = d >> 5;
    const uint32_t v = ghist[blockIdx.x * 256 + d];
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[warp] = x;
    __syncthreads();
    uint32_t before = 0;
    for (int w = 0; w < warp; w++) before += wsum[w];
    gbase[blockIdx.x * 256 + d] = before + x - v;
}

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
    uint32_t tile;
};
static __global__ void __launch_bounds__(RS_THREADS, 3) k_rs_scatter(const unsigned long long* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                               unsigned long long* __restrict__ okeys, uint32_t* __restrict__ ovals, int64_t n, int shift,
                                                               unsigned long long* status /*[tiles][256], zeroed*/, uint32_t* ticket /*zeroed*/,
                                                               const uint32_t* __restrict__ gbase /*[256]*/, uint32_t* err) {
    extern __shared__ __align__(16) unsigned char rs_raw[];
    RsSmem& sm = *reinterpret_cast<RsSmem*>(rs_raw);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // tiles are handed out in ticket order, so every tile a look-back waits for is already running (or done)
    if (threadIdx.x == 0) sm.tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = sm.tile;
    const int64_t bbase = (int64_t)tile * RS_TILE;
    const int64_t wbase = bbase + (int64_t)warp * RS_ITEMS_PER_WARP;
    unsigned long long k[RS_PER_LANE];
    uint32_t v[RS_PER_LANE];
#pragma unroll
    for (int j = 0; j < RS_PER_LANE; j++) { const int64_t g = wbase + j * 32 + lane; k[j] = g < n ? keys[g] : 0ULL; }
#pragma unroll
    for (int j = 0; j < RS_PER_LANE; j++) { const int64_t g = wbase + j * 32 + lane; v[j] = g < n ? vals[g] : 0u; }
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
        // publish this tile's count of digit d at once, so later tiles never wait for more than the counting phase
        volatile unsigned long long* stat = status;
        stat[(size_t)tile * 256 + d] = (tile == 0 ? RS_FLAG_INC : RS_FLAG_AGG) | (unsigned long long)tot;
        uint32_t x = tot;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) sm.wsum[warp] = x;
        __syncthreads();
        uint32_t before = 0;
        for (int w = 0; w < warp; w++) before += sm.wsum[w];
        const uint32_t lstart = before + x - tot;
        uint32_t run = lstart;
        for (int w = 0; w < RS_WARPS; w++) { const uint32_t c = sm.cnt[w][d]; sm.cnt[w][d] = run; run += c; }
        // decoupled look-back: items with digit d in all earlier tiles = aggregates back to the first inclusive prefix
        unsigned long long excl = 0;
        if (tile != 0) {
            int64_t j = (int64_t)tile - 1;
            uint32_t spins = 0;
            bool done = false;
            while (!done) {                                              // four predecessors per round trip
                unsigned long long v[4];
#pragma unroll
                for (int u = 0; u < 4; u++) v[u] = j - u >= 0 ? stat[(size_t)(j - u) * 256 + d] : RS_FLAG_INC;
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    if (done) break;
                    if ((v[u] >> 62) == 0ULL) {                          // not published yet: poll again from here
                        if (++spins > RS_SPIN_LIMIT) { *err = 1u; done = true; }   // never expected; reported instead of hanging
                        break;
                    }
                    excl += v[u] & RS_VALUE_MASK;
                    j--;
                    if (v[u] & RS_FLAG_INC) done = true;
                }
            }
            stat[(size_t)tile * 256 + d] = RS_FLAG_INC | (excl + tot);
        }
        sm.delta[d] = gbase[d] + (uint32_t)excl - lstart;
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
    }
}

// Sorts n pairs by bits [begin_bit, end_bit) of the key.  keys/vals and tkeys/tvals are ping-pong
// buffers; returns which pair of buffers holds the result (0 = keys/vals, 1 = tkeys/tvals).
// hist needs dv_radix_hist_words(n) uint32 (8-byte aligned); n < 2^32.  After the stream has been synchronised, dv_radix_sort_failed(hist, n)
// tells whether a look-back gave up (never expected; results are then invalid).
static inline uint32_t* dv_radix_err_ptr(uint32_t* hist, int64_t n) {
    const int64_t nb = (n + RS_TILE - 1) / RS_TILE;
    return hist + 512 * (nb + 1) + 2 * RS_MAX_PASSES * 256 + RS_MAX_PASSES;
}
static inline int dv_radix_sort_pairs(unsigned long long* keys, uint32_t* vals, unsigned long long* tkeys, uint32_t* tvals,
                                      int64_t n, int begin_bit, int end_bit, uint32_t* hist, cudaStream_t st, int64_t* launches) {
    if (n <= 1) return 0;
    cudaFuncSetAttribute(k_rs_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem));
    const int64_t nb = (n + RS_TILE - 1) / RS_TILE;
    const int npass = (end_bit - begin_bit + 7) / 8;
    unsigned long long* status = reinterpret_cast<unsigned long long*>(hist);
    uint32_t* ghist = hist + 512 * (nb + 1);
    uint32_t* gbase = ghist + RS_MAX_PASSES * 256;
    uint32_t* tickets = gbase + RS_MAX_PASSES * 256;
    cudaMemsetAsync(ghist, 0, sizeof(uint32_t) * (2 * RS_MAX_PASSES * 256 + 16), st);
    k_rs_ghist<<<(unsigned)std::min<int64_t>((n + RS_THREADS * 16 - 1) / (RS_THREADS * 16), 148 * 8), RS_THREADS, 0, st>>>(keys, n, begin_bit, npass, ghist);
    k_rs_gscan<<<npass, 256, 0, st>>>(ghist, gbase);
    if (launches) *launches += 2;
    int cur = 0;
    for (int p = 0; p < npass; p++) {
        unsigned long long* ik = cur ? tkeys : keys; uint32_t* iv = cur ? tvals : vals;
        unsigned long long* ok = cur ? keys : tkeys; uint32_t* ov = cur ? vals : tvals;
        cudaMemsetAsync(status, 0, sizeof(unsigned long long) * 256 * (size_t)nb, st);
        k_rs_scatter<<<(unsigned)nb, RS_THREADS, sizeof(RsSmem), st>>>(ik, iv, ok, ov, n, begin_bit + 8 * p, status, tickets + p, gbase + p * 256,
                                                                       tickets + RS_MAX_PASSES);
        if (launches) *launches += 1;
        cur ^= 1;
    }
    return cur;
}
