// index.cu — positional inverted index over chunk symbols + indexed scoring with top-k (kernel stages K2-K4, bulk a7).
// replaces, at corpus scale (BASELINE configs 3 and 5): the scoring loop and top-k of select_reference_material_context
//   (reference web/server.py:32073-32087):  score = sum(lowered.count(term) for term in terms);
//   sorted(positive, key=(-score, index))[:k].   Results are identical to the scan scorer of chunk.cu (exact counts).
//
// Build (dvb_chunk_index_build):
//   K1  symbolise every chunk (chunk.cu k_symbolize: decode, CPython lower(), class map) -> u16 symbols, resident
//   K2a every adjacent same-class symbol pair (CJK-CJK or ASCII-ASCII) becomes a posting
//         key = pair id (30 bits), payload = (chunk, position of the pair's first symbol)
//       (+ the symbol that follows the pair, a free trigram filter at query time); emitted chunk by chunk, position by position
//   K2b stable LSD radix sort by the 30 key bits only (sort.cuh; 4 passes): postings of one key stay ordered by
//       (chunk, position) because that is the emission order
//   K2c segmented reduce: run heads of equal keys -> distinct key list + CSR offsets (the postings "document frequency")
// Query (dvb_chunk_index_topk), per batch of queries with a dense [batch][n_chunks] int32 score tile in HBM:
//   K3a plan   per term: binary-search every bigram of the term in the key list, keep the RAREST one (offset j)
//   K3b score  one warp per term walks that posting list 32 entries at a time: candidate start = pos - j, the term is
//              verified against the resident symbols (exact), matches are counted greedily left to right per chunk
//              (non-overlapping == str.count) and added to score[query][chunk]
//   K4   top-k the warp-level top-k of chunk.cu under (score desc, global index asc), positives only
// Query terms come from the reference's own term extractor (web/server.py:32037-32051): every term is >= 2 symbols
// of one class, so each term has at least one indexed pair.
#include "common.cuh"
#include "sort.cuh"
#include "index.cuh"
#include <stdlib.h>

__device__ __forceinline__ uint32_t ix_code(uint32_t s) { return s >= 0x8000u ? 20992u + (s & 0x7FFFu) : s; }     // 15-bit symbol code
__device__ __forceinline__ bool ix_same_class(uint32_t a, uint32_t b) { return a != 0u && b != 0u && ((a ^ b) & 0x8000u) == 0u; }
__device__ __forceinline__ uint32_t ix_pair_key(uint32_t a, uint32_t b) { return (ix_code(a) << 15) | ix_code(b); }
// A pair's posting list is sub-ordered by the low bits of the symbol AFTER the pair (sort key = pair << 2 | next & 3):
// a term that continues past the pair only streams the quarter of the list its own next symbol selects.
#define IX_SUB_BITS 2
#define IX_SUB_MASK 3u

// ---- K2a: count, then emit postings (one warp per chunk)
__global__ void __launch_bounds__(256) k_ix_count(const uint16_t* __restrict__ sym, const int64_t* __restrict__ sym_off, const int32_t* __restrict__ sym_cnt,
                                                   int64_t n_chunks, int32_t* __restrict__ pair_cnt, int32_t* __restrict__ err) {
    const int lane = threadIdx.x & 31;
    for (int64_t c = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); c < n_chunks; c += (int64_t)gridDim.x * 8) {
        const uint16_t* s = sym + sym_off[c]; const int32_t m = sym_cnt[c];
        int cnt = 0;
        for (int32_t i = lane; i + 1 < m; i += 32) cnt += ix_same_class(s[i], s[i + 1]) ? 1 : 0;
        for (int d = 16; d > 0; d >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, d);
        if (lane == 0) { pair_cnt[c] = cnt; if (m > 65535) atomicExch(err, 1); }
    }
}
__global__ void __launch_bounds__(256) k_ix_emit(const uint16_t* __restrict__ sym, const int64_t* __restrict__ sym_off, const int32_t* __restrict__ sym_cnt,
                                                  int64_t n_chunks, const int64_t* __restrict__ post_off, unsigned long long* __restrict__ pk, uint32_t* __restrict__ pv,
                                                  uint8_t* __restrict__ digit0) {
    const int lane = threadIdx.x & 31;
    for (int64_t c = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); c < n_chunks; c += (int64_t)gridDim.x * 8) {
        const uint16_t* s = sym + sym_off[c]; const int32_t m = sym_cnt[c];
        int64_t w = post_off[c];
        for (int32_t b0 = 0; b0 + 1 < m; b0 += 32) {
            const int32_t i = b0 + lane;
            uint32_t a = 0, b = 0, nx = 0;
            if (i + 1 < m) { a = s[i]; b = s[i + 1]; if (i + 2 < m) nx = s[i + 2]; }
            const bool v = ix_same_class(a, b);
            const uint32_t bal = __ballot_sync(0xffffffffu, v);
            if (v) {
                const int64_t o = w + __popc(bal & ((1u << lane) - 1u));
                const uint32_t k32 = (ix_pair_key(a, b) << IX_SUB_BITS) | (nx & IX_SUB_MASK);
                pk[o] = ((unsigned long long)c << 32) | k32;
                pv[o] = (uint32_t)i | (nx << 16);                  // position | the symbol AFTER the pair (0 = none / separator)
                digit0[o] = (uint8_t)k32;                          // the first radix pass builds its histogram from these bytes, not from the keys
            }
            w += __popc(bal);
        }
    }
}
// ---- K2a + first radix pass fused: the postings are never written in emission order.
// Block b owns the postings with emission index [b * RS_TILE, (b + 1) * RS_TILE): it finds the chunks that produce them
// (binary search in the scanned per-chunk counts), re-walks their symbols — a chunk that straddles a block border is walked
// by both neighbours, 16 steps for a 512-character chunk — and hands every posting of its range to `sink(tile position, key, value)`.
template <typename Sink>
__device__ __forceinline__ void ix_block_postings(const uint16_t* __restrict__ sym, const int64_t* __restrict__ sym_off, const int32_t* __restrict__ sym_cnt,
                                                  int64_t n_chunks, const int64_t* __restrict__ post_off, int64_t n_post, Sink sink) {
    __shared__ int64_t s_c0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t P0 = (int64_t)blockIdx.x * RS_TILE, P1 = min(n_post, P0 + RS_TILE);
    if (threadIdx.x == 0) {                                      // last chunk whose first posting index is <= P0
        int64_t lo = 0, hi = n_chunks - 1;
        while (lo < hi) { const int64_t mid = (lo + hi + 1) >> 1; if (post_off[mid] <= P0) lo = mid; else hi = mid - 1; }
        s_c0 = lo;
    }
    __syncthreads();
    for (int64_t c = s_c0 + warp; c < n_chunks; c += RS_WARPS) {
        int64_t w = post_off[c];
        if (w >= P1) break;
        const int64_t wend = post_off[c + 1];
        if (wend <= P0) continue;
        const uint16_t* s = sym + sym_off[c]; const int32_t m = sym_cnt[c];
        for (int32_t b0 = 0; b0 + 1 < m && w < P1; b0 += 32) {
            const int32_t i = b0 + lane;
            uint32_t a = 0, b = 0, nx = 0;
            if (i + 1 < m) { a = s[i]; b = s[i + 1]; if (i + 2 < m) nx = s[i + 2]; }
            const bool v = ix_same_class(a, b);
            const uint32_t bal = __ballot_sync(0xffffffffu, v);
            if (v) {
                const int64_t o = w + __popc(bal & ((1u << lane) - 1u));
                if (o >= P0 && o < P1)
                    sink((int)(o - P0), ((unsigned long long)c << 32) | (ix_pair_key(a, b) << IX_SUB_BITS) | (nx & IX_SUB_MASK), (uint32_t)i | (nx << 16));
            }
            w += __popc(bal);
        }
    }
}
struct IxHistSink { uint32_t* h; __device__ __forceinline__ void operator()(int, unsigned long long key, uint32_t) const { atomicAdd(&h[(uint32_t)key & 255u], 1u); } };
struct IxStageSink { unsigned long long* k; uint32_t* v; __device__ __forceinline__ void operator()(int pos, unsigned long long key, uint32_t val) const { k[pos] = key; v[pos] = val; } };

static __global__ void __launch_bounds__(RS_THREADS) k_ix_hist1(const uint16_t* __restrict__ sym, const int64_t* __restrict__ sym_off, const int32_t* __restrict__ sym_cnt,
                                                                int64_t n_chunks, const int64_t* __restrict__ post_off, int64_t n_post,
                                                                uint32_t* __restrict__ hist /*[256][nb]*/, int64_t nb) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    ix_block_postings(sym, sym_off, sym_cnt, n_chunks, post_off, n_post, IxHistSink{ h });
    __syncthreads();
    hist[(int64_t)threadIdx.x * nb + blockIdx.x] = h[threadIdx.x];
}
static __global__ void __launch_bounds__(RS_THREADS, 3) k_ix_scatter1(const uint16_t* __restrict__ sym, const int64_t* __restrict__ sym_off, const int32_t* __restrict__ sym_cnt,
                                                                      int64_t n_chunks, const int64_t* __restrict__ post_off, int64_t n_post,
                                                                      unsigned long long* __restrict__ okeys, uint32_t* __restrict__ ovals,
                                                                      const uint32_t* __restrict__ hist, int64_t nb, uint8_t* __restrict__ next_digit) {
    extern __shared__ __align__(16) unsigned char rs_raw[];
    RsSmem& sm = *reinterpret_cast<RsSmem*>(rs_raw);
    ix_block_postings(sym, sym_off, sym_cnt, n_chunks, post_off, n_post, IxStageSink{ sm.key, sm.val });       // emission order inside the tile
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int cnt_tile = (int)min((int64_t)RS_TILE, n_post - (int64_t)blockIdx.x * RS_TILE);
    unsigned long long k[RS_PER_LANE];
    uint32_t v[RS_PER_LANE];
#pragma unroll
    for (int j = 0; j < RS_PER_LANE; j++) { const int t = warp * RS_ITEMS_PER_WARP + j * 32 + lane; k[j] = t < cnt_tile ? sm.key[t] : 0ULL; v[j] = t < cnt_tile ? sm.val[t] : 0u; }
    rs_scatter_body(k, v, sm, okeys, ovals, n_post, 0, hist, nb, next_digit, 8);      // its first barrier orders these reads before the placement writes
}

// exclusive scan of int32 counts into int64 offsets (single CTA; n is the number of chunks / blocks)
#define IX_SCAN_PER 8
static __global__ void __launch_bounds__(1024) k_ix_scan(const int32_t* __restrict__ cnt, int64_t* __restrict__ off, int64_t n) {
    __shared__ long long warp_tot[32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base = 0; base < n; base += 1024 * IX_SCAN_PER) {
        const int64_t i0 = base + (int64_t)threadIdx.x * IX_SCAN_PER;
        int32_t c[IX_SCAN_PER];
        long long v = 0;
#pragma unroll
        for (int j = 0; j < IX_SCAN_PER; j++) { c[j] = i0 + j < n ? cnt[i0 + j] : 0; v += c[j]; }
        long long x = v;
        for (int d = 1; d < 32; d <<= 1) { long long y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        long long before = carry;
        for (int w = 0; w < warp; w++) before += warp_tot[w];
        long long run = before + x - v;
#pragma unroll
        for (int j = 0; j < IX_SCAN_PER; j++) { if (i0 + j < n) off[i0 + j] = run; run += c[j]; }
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[n] = carry;
}

// Large n: block sums -> single-CTA scan of the sums -> per-block local scan + base (three short parallel launches
// instead of one long single-CTA loop).  tmp_sum: int32[nb], tmp_base: int64[nb + 1], nb = ceil(n / IX_SCAN_TILE).
#define IX_SCAN_TILE 8192
static __global__ void __launch_bounds__(256) k_ix_scan_sum(const int32_t* __restrict__ cnt, int64_t n, int32_t* __restrict__ bsum) {
    __shared__ int s_sum;
    if (threadIdx.x == 0) s_sum = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * IX_SCAN_TILE;
    int v = 0;
    for (int i = threadIdx.x; i < IX_SCAN_TILE; i += 256) if (base + i < n) v += cnt[base + i];
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    if ((threadIdx.x & 31) == 0) atomicAdd(&s_sum, v);
    __syncthreads();
    if (threadIdx.x == 0) bsum[blockIdx.x] = s_sum;
}
static __global__ void __launch_bounds__(256) k_ix_scan_apply(const int32_t* __restrict__ cnt, int64_t n, const int64_t* __restrict__ bbase,
                                                              int64_t* __restrict__ off, int64_t nb) {
    __shared__ long long warp_tot[8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t i0 = (int64_t)blockIdx.x * IX_SCAN_TILE + (int64_t)threadIdx.x * 32;      // 32 consecutive counts per thread
    int32_t c[32];
    long long v = 0;
#pragma unroll
    for (int j = 0; j < 32; j++) { c[j] = i0 + j < n ? cnt[i0 + j] : 0; v += c[j]; }
    long long x = v;
    for (int d = 1; d < 32; d <<= 1) { long long y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
    if (lane == 31) warp_tot[warp] = x;
    __syncthreads();
    long long run = bbase[blockIdx.x] + x - v;
    for (int w = 0; w < warp; w++) run += warp_tot[w];
#pragma unroll
    for (int j = 0; j < 32; j++) { if (i0 + j < n) off[i0 + j] = run; run += c[j]; }
    if (blockIdx.x == 0 && threadIdx.x == 0) off[n] = bbase[nb];
}

// ---- K2c: run heads of equal keys.  Blocks of IX_HEAD_TILE postings: count heads, scan the block counts, write.
#define IX_HEAD_TILE 4096
__global__ void __launch_bounds__(256) k_ix_head_count(const unsigned long long* __restrict__ pk, int64_t n, int32_t* __restrict__ blk_cnt) {
    __shared__ int s_cnt;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * IX_HEAD_TILE;
    int c = 0;
    for (int i = threadIdx.x; i < IX_HEAD_TILE; i += 256) {
        const int64_t g = base + i;
        if (g < n && (g == 0 || (uint32_t)pk[g] != (uint32_t)pk[g - 1])) c++;
    }
    for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(&s_cnt, c);
    __syncthreads();
    if (threadIdx.x == 0) blk_cnt[blockIdx.x] = s_cnt;
}
// Also writes the resident form of the sorted postings: the chunk id alone (the key is implied by the key's offset range).
__global__ void __launch_bounds__(256) k_ix_head_write(const unsigned long long* __restrict__ pk, int64_t n, const int64_t* __restrict__ blk_off,
                                                        uint32_t* __restrict__ ukeys, int64_t* __restrict__ uoff, uint32_t* __restrict__ pc) {
    __shared__ int s_warp[8];
    const int64_t base = (int64_t)blockIdx.x * IX_HEAD_TILE;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int64_t w = blk_off[blockIdx.x];
    for (int i0 = 0; i0 < IX_HEAD_TILE; i0 += 256) {                 // ordered: 256 consecutive postings per step
        const int64_t g = base + i0 + threadIdx.x;
        const unsigned long long kg = g < n ? pk[g] : 0ull;
        const bool head = g < n && (g == 0 || (uint32_t)kg != (uint32_t)pk[g - 1]);
        if (g < n) pc[g] = (uint32_t)(kg >> 32);
        const uint32_t bal = __ballot_sync(0xffffffffu, head);
        if (lane == 0) s_warp[warp] = __popc(bal);
        __syncthreads();
        int before = 0, total = 0;
        for (int k = 0; k < 8; k++) { if (k < warp) before += s_warp[k]; total += s_warp[k]; }
        if (head) {
            const int64_t o = w + before + __popc(bal & ((1u << lane) - 1u));
            ukeys[o] = (uint32_t)kg; uoff[o] = g;
        }
        w += total;
        __syncthreads();
    }
}

extern "C" void dvb_chunk_index_free(dvb_index* ix);

// ---- K3a: per term, the rarest indexed pair
__global__ void __launch_bounds__(256) k_ix_plan(const uint16_t* __restrict__ term_sym, const int32_t* __restrict__ term_off, int32_t n_terms,
                                                  const uint32_t* __restrict__ ukeys, const int64_t* __restrict__ uoff, int64_t n_keys,
                                                  int64_t* __restrict__ plan_lo, int64_t* __restrict__ plan_hi, int32_t* __restrict__ plan_j, int32_t* __restrict__ err) {
    for (int32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n_terms; t += gridDim.x * blockDim.x) {
        const uint16_t* ts = term_sym + term_off[t];
        const int32_t L = term_off[t + 1] - term_off[t];
        int64_t best_lo = 0, best_hi = 0; int32_t best_j = -1; int64_t best_df = -1;
        bool usable = false, bad = L < 2;
        for (int32_t j = 0; j < L; j++) bad |= ts[j] == 0;
        // L >= 3: only pairs that are followed by a term symbol (one chunk-ordered sub-list); L == 2: the pair's four sub-lists
        for (int32_t j = 0; j + 1 < L && (j + 2 < L || L == 2) && !bad; j++) {
            if (!ix_same_class(ts[j], ts[j + 1])) continue;
            usable = true;
            const bool nxt = j + 2 < L, covered = j == 0 && L <= 3;
            const uint32_t k30 = ix_pair_key(ts[j], ts[j + 1]) << IX_SUB_BITS;
            const uint64_t key_lo = nxt ? (k30 | (ts[j + 2] & IX_SUB_MASK)) : k30;
            const uint64_t key_hi = nxt ? key_lo + 1 : (uint64_t)k30 + IX_SUB_MASK + 1;      // exclusive
            int64_t lo = 0, hi = n_keys;
            while (lo < hi) { const int64_t m = (lo + hi) >> 1; if (ukeys[m] < key_lo) lo = m + 1; else hi = m; }
            int64_t lo2 = lo, hi2 = n_keys;
            while (lo2 < hi2) { const int64_t m = (lo2 + hi2) >> 1; if (ukeys[m] < key_hi) lo2 = m + 1; else hi2 = m; }
            const int64_t a = n_keys > 0 ? uoff[lo] : 0, b = n_keys > 0 ? uoff[lo2] : 0;
            // cost in bytes: every posting of the pair streams its 4-byte position word; the chunk id (a 32-byte sector) is
            // only fetched for postings whose look-ahead symbol matches (~1/12 of a sub-list), and a candidate the pair and
            // the look-ahead symbol do not decide costs two more random sectors (chunk offset + symbols).  A 2-symbol term
            // needs the chunk of every posting.
            const int64_t cost = nxt ? (b - a) * 4 + (b - a) * (32 + (covered ? 0 : 64)) / 12 : (b - a) * 8;
            if (best_df < 0 || cost < best_df) { best_df = cost; best_lo = a; best_hi = b; best_j = j; }
            if (b == a) break;                                    // a pair that occurs nowhere: the term occurs nowhere
        }
        if (bad || !usable) atomicExch(err, 1);                   // not a term of the reference's extractor
        plan_lo[t] = best_lo; plan_hi[t] = best_hi; plan_j[t] = best_j < 0 ? 0 : best_j;
    }
}

// ---- K3b: posting lists are cut into work items of <= IX_ITEM postings, one warp each, so that a frequent pair
// (10^5..10^6 postings) does not serialise on one warp.  An item owns the CHUNKS that start inside its range: it skips the
// leading postings of a chunk the previous item started, and runs past its end to finish its last chunk (the greedy
// non-overlapping count is per chunk, so chunks must not be split).
#define IX_ITEM 2048
__global__ void __launch_bounds__(256) k_ix_items(const int64_t* __restrict__ plan_lo, const int64_t* __restrict__ plan_hi, int32_t n_terms, int32_t* __restrict__ n_items) {
    for (int32_t t = blockIdx.x * blockDim.x + threadIdx.x; t < n_terms; t += gridDim.x * blockDim.x)
        n_items[t] = (int32_t)((plan_hi[t] - plan_lo[t] + IX_ITEM - 1) / IX_ITEM);
}
__global__ void __launch_bounds__(256) k_ix_score(IxSeg ix, const uint16_t* __restrict__ term_sym, const int32_t* __restrict__ term_off,
                                                   const int32_t* __restrict__ query_term_off, int32_t q0, int32_t q1,
                                                   const int64_t* __restrict__ plan_lo, const int64_t* __restrict__ plan_hi, const int32_t* __restrict__ plan_j,
                                                   const int64_t* __restrict__ item_off, int32_t* __restrict__ score /* [q1 - q0][n_chunks] */) {
    const int lane = threadIdx.x & 31;
    const int32_t t_begin = query_term_off[q0], t_end = query_term_off[q1];
    const int64_t w_begin = item_off[t_begin], w_end = item_off[t_end];
    for (int64_t w = w_begin + (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); w < w_end; w += (int64_t)gridDim.x * 8) {
        int32_t ta = t_begin, tb = t_end;                          // term of item w: last t with item_off[t] <= w
        while (tb - ta > 1) { const int32_t m = (ta + tb) >> 1; if (item_off[m] <= w) ta = m; else tb = m; }
        const int32_t t = ta;
        int32_t a = q0, b = q1;                                    // query of term t: last q with query_term_off[q] <= t
        while (b - a > 1) { const int32_t m = (a + b) >> 1; if (query_term_off[m] <= t) a = m; else b = m; }
        int32_t* row = score + (int64_t)(a - q0) * ix.n_chunks;
        const uint16_t* ts = term_sym + term_off[t];
        const int32_t L = term_off[t + 1] - term_off[t], j = plan_j[t];
        const int64_t lo = plan_lo[t], hi = plan_hi[t];
        const int64_t r_lo = lo + (w - item_off[t]) * IX_ITEM, r_hi = min(hi, r_lo + IX_ITEM);
        const bool use_next = j + 2 < L, covered = j == 0 && L <= 3;
        // L == 2 walks all four sub-lists of the pair (each chunk-ordered, not the whole range): occurrences are counted
        // additively.  "xy" cannot overlap itself; "xx" is counted per RUN of x at the run's last pair (look-ahead != x):
        // a run of r symbols holds floor(r / 2) non-overlapping occurrences, which is what the greedy scan would count.
        const bool pair_only = L == 2, same2 = pair_only && ts[0] == ts[1];
        const uint32_t next_sym = use_next ? ts[j + 2] : 0u;
        // A term can overlap itself only if it has a border (a proper prefix that is also a suffix: "aaa", "abab").  Without
        // one — and for 2-symbol terms, whose "xx" case is counted per run — matches simply add up.
        bool bordered = false;
        if (!pair_only) {
            for (int32_t d = 1 + lane; d < L; d += 32) {
                bool eq = true;
                for (int32_t q = 0; q + d < L && eq; q++) eq = ts[q] == ts[q + d];
                bordered |= eq;
            }
            bordered = __any_sync(0xffffffffu, bordered);
        }
        const bool additive = pair_only || !bordered;
        // Only a greedy (bordered) term needs whole chunks per item: its items own the chunks that START inside them.  An
        // additive term counts exactly the postings of its own range.
        const int64_t lim = additive ? r_hi : hi;
        const int64_t skip_chunk = !additive && r_lo > lo ? (int64_t)ix.pc[r_lo - 1] : -1;   // owned by the previous item
        const int64_t last_chunk = !additive ? (int64_t)ix.pc[r_hi - 1] : -1;               // ours to finish
        if (!additive && r_lo > lo && skip_chunk == last_chunk) continue;   // the whole item lies inside a chunk the previous item owns
        int64_t cur = -1; int32_t cnt = 0, last_end = 0;
        bool more = true, leading = !additive && r_lo > lo;                     // still inside the run of postings the previous item owns
        // One step = 32 postings (position words v of postings p0 + lane).
        auto step = [&](const int64_t p0, const uint32_t v) {
            const int64_t p = p0 + lane;
            int64_t chunk = -2; int32_t start = 0, wgt = 0; bool beyond = false;
            // The chunk id is needed by every lane only at the item's borders (ownership of the first / last chunk) and by
            // 2-symbol terms; inside the item a term with a look-ahead symbol fetches it for the lanes that pass the filter.
            const bool border = leading || (!additive && p0 + 32 > r_hi) || pair_only;
            const bool cand = p < lim && (int32_t)(v & 0xFFFFu) >= j && (pair_only || (v >> 16) == next_sym);
            if (p < lim && (border || cand)) chunk = (int64_t)ix.pc[p];
            bool skip = false;
            if (leading) {                                         // only the LEADING postings of skip_chunk are skipped (a pair's
                const uint32_t other = __ballot_sync(0xffffffffu, p < lim && chunk != skip_chunk);   // sub-lists restart the chunk order)
                skip = other ? lane < __ffs(other) - 1 : true;
                if (other) leading = false;
            }
            if (p < lim) {
                start = (int32_t)(v & 0xFFFFu) - j;
                beyond = !additive && p >= r_hi && chunk != last_chunk;          // first posting of a chunk the next item owns
                if (!beyond && !skip && cand) {
                    if (pair_only) {
                        if (!same2) wgt = 1;
                        else if ((v >> 16) != ts[0]) {
                            const uint16_t* s = ix.sym + ix.sym_off[chunk];
                            int32_t q = start, r = 2;
                            while (q > 0 && s[q - 1] == ts[0]) { q--; r++; }
                            wgt = r >> 1;
                        }
                    } else {
                        // the pair and the look-ahead symbol are already matched: only the rest of the term is read back,
                        // and the chunk length only when the term reaches past the look-ahead symbol
                        if (covered) wgt = 1;
                        else if (j + 3 >= L || start + L <= ix.sym_cnt[chunk]) {
                            const uint16_t* s = ix.sym + ix.sym_off[chunk] + start;
                            bool match = true;
                            for (int32_t d = 0; d < j && match; d++) match = s[d] == ts[d];
                            for (int32_t d = j + 3; d < L && match; d++) match = s[d] == ts[d];
                            wgt = match ? 1 : 0;
                        }
                    }
                }
            }
            const bool match = wgt > 0;
            // postings are ordered by chunk: everything after the first `beyond` posting is beyond too
            const uint32_t bey = __ballot_sync(0xffffffffu, beyond);
            uint32_t bal = __ballot_sync(0xffffffffu, match);
            if (bey) { bal &= (1u << (__ffs(bey) - 1)) - 1u; more = false; }
            if (additive) {
                // occurrences cannot overlap: the count of a chunk is the number of its matching postings, whatever the
                // order — one atomic per (chunk, step) from the first lane of each group of equal chunks
                if ((bal >> lane) & 1u) {
                    if (same2) atomicAdd(&row[chunk], wgt);
                    else {
                        const uint32_t grp = __match_any_sync(bal, (uint32_t)chunk);
                        if (__ffs(grp) - 1 == lane) atomicAdd(&row[chunk], __popc(grp));
                    }
                }
                bal = 0u;
            }
            if (bal) {
                // bordered term: occurrences CAN overlap, but rarely do.  If no match of this step starts inside the previous
                // match of the same chunk (the one before it in the step, or the carried one), greedy counts them all.
                const uint32_t pm = bal & ((1u << lane) - 1u);
                const int srcp = pm ? 31 - __clz(pm) : 0;
                const int64_t pch = __shfl_sync(0xffffffffu, chunk, srcp);
                const int32_t pst = __shfl_sync(0xffffffffu, start, srcp);
                bool conflict = false;
                if ((bal >> lane) & 1u) conflict = pm ? (pch == chunk && start - pst < L) : (chunk == cur && start < last_end);
                if (!__any_sync(0xffffffffu, conflict)) {
                    if (cnt > 0 && lane == 0) atomicAdd(&row[cur], cnt);
                    cnt = 0;
                    if ((bal >> lane) & 1u) {
                        const uint32_t grp = __match_any_sync(bal, (uint32_t)chunk);
                        if (__ffs(grp) - 1 == lane) atomicAdd(&row[chunk], __popc(grp));
                    }
                    const int lastl = 31 - __clz(bal);
                    cur = __shfl_sync(0xffffffffu, chunk, lastl);
                    last_end = __shfl_sync(0xffffffffu, start, lastl) + L;
                    bal = 0u;
                }
            }
            while (bal) {                                          // greedy, left to right: non-overlapping == str.count
                const int src = __ffs(bal) - 1; bal &= bal - 1;
                const int64_t cb = __shfl_sync(0xffffffffu, chunk, src);
                const int32_t sb = __shfl_sync(0xffffffffu, start, src);
                const int32_t wb = __shfl_sync(0xffffffffu, wgt, src);
                if (cb != cur) {
                    if (cnt > 0 && lane == 0) atomicAdd(&row[cur], cnt);
                    cur = cb; cnt = 0; last_end = -2147483647;
                }
                if (pair_only) cnt += wb;
                else if (sb >= last_end) { cnt++; last_end = sb + L; }
            }
        };
        int64_t p0 = r_lo;
        while (p0 < lim && more) {
            // Interior of the item, term with a look-ahead symbol: 128 postings per iteration, and nothing but the four
            // position words is touched unless one of them passes the look-ahead filter (about 1 posting in 12 does on
            // a Zipf corpus; on rarer symbols whole iterations are skipped).
            if (!leading && use_next && p0 + 128 <= r_hi) {
                uint32_t v4[4]; bool c4[4]; bool any = false;
#pragma unroll
                for (int u = 0; u < 4; u++) v4[u] = ix.pv[p0 + 32 * u + lane];
#pragma unroll
                for (int u = 0; u < 4; u++) { c4[u] = (v4[u] >> 16) == next_sym && (int32_t)(v4[u] & 0xFFFFu) >= j; any |= c4[u]; }
                if (__any_sync(0xffffffffu, any)) {
                    if (additive) {
                        // the four sub-steps are independent: their dependent loads (chunk id -> symbol offset -> symbols)
                        // are issued side by side instead of one chain after the other
                        uint32_t ch[4]; int64_t so[4]; int32_t sc[4];
#pragma unroll
                        for (int u = 0; u < 4; u++) ch[u] = c4[u] ? ix.pc[p0 + 32 * u + lane] : 0u;
                        if (!covered) {
#pragma unroll
                            for (int u = 0; u < 4; u++) { so[u] = c4[u] ? ix.sym_off[ch[u]] : 0; sc[u] = c4[u] && j + 3 < L ? ix.sym_cnt[ch[u]] : 0x7fffffff; }
#pragma unroll
                            for (int u = 0; u < 4; u++) if (c4[u]) {
                                const int32_t start = (int32_t)(v4[u] & 0xFFFFu) - j;
                                bool m = start + L <= sc[u];
                                const uint16_t* sp = ix.sym + so[u] + start;
                                for (int32_t d = 0; d < j && m; d++) m = sp[d] == ts[d];
                                for (int32_t d = j + 3; d < L && m; d++) m = sp[d] == ts[d];
                                c4[u] = m;
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 4; u++) {
                            const uint32_t bal = __ballot_sync(0xffffffffu, c4[u]);
                            if (c4[u]) {
                                const uint32_t grp = __match_any_sync(bal, ch[u]);
                                if (__ffs(grp) - 1 == lane) atomicAdd(&row[ch[u]], __popc(grp));
                            }
                        }
                    } else {
#pragma unroll
                        for (int u = 0; u < 4; u++) if (__any_sync(0xffffffffu, c4[u])) step(p0 + 32 * u, v4[u]);
                    }
                }
                p0 += 128;
                continue;
            }
            if (pair_only && !same2 && p0 + 128 <= r_hi) {
                // "xy" with x != y: every posting of the pair is one occurrence; only the chunk ids are read
                uint32_t ch[4];
#pragma unroll
                for (int u = 0; u < 4; u++) ch[u] = ix.pc[p0 + 32 * u + lane];
#pragma unroll
                for (int u = 0; u < 4; u++) {
                    const uint32_t grp = __match_any_sync(0xffffffffu, ch[u]);
                    if (__ffs(grp) - 1 == lane) atomicAdd(&row[ch[u]], __popc(grp));
                }
                p0 += 128;
                continue;
            }
            step(p0, p0 + lane < lim ? ix.pv[p0 + lane] : 0u);
            p0 += 32;
        }
        if (cnt > 0 && lane == 0) atomicAdd(&row[cur], cnt);
    }
}

// ------------------------------------------------------------------ host side
void ix_free_seg(IxSeg& sg) {
    void* ptrs[] = { sg.sym, sg.sym_off, sg.sym_cnt, sg.pc, sg.pv, sg.ukeys, sg.uoff };
    for (void* p : ptrs) if (p) cudaFree(p);
    sg = IxSeg();
}

extern "C" void dvb_chunk_index_free(dvb_index* ix) {
    if (!ix) return;
    cudaSetDevice(ix->device);
    for (IxSeg& sg : ix->segs) ix_free_seg(sg);
    delete ix;
}

extern "C" int32_t dvb_chunk_index_stats(const dvb_index* ix, int64_t* out6) {
    if (!ix || !out6) return DVB_E_INVALID;
    out6[0] = ix->n_chunks; out6[1] = ix->n_bytes; out6[2] = ix->n_post; out6[3] = ix->n_keys;
    out6[4] = 2 * ix->n_bytes + 12 * ix->n_chunks + 8 * ix->n_post + 12 * ix->n_keys + 80 * (int64_t)ix->segs.size();   // resident device bytes
    out6[5] = ix->device;
    return DVB_OK;
}

static int64_t ix_seg_bytes() {                       // test hook: small segments exercise the multi-segment path on small corpora
    const char* e = getenv("DVB_IX_SEG_BYTES");
    if (e && *e) { const long long v = atoll(e); if (v > 0) return (int64_t)std::min<long long>(v, IX_SEG_BYTES); }
    return IX_SEG_BYTES;
}

// Builds one segment from chunks [0, n_chunks) whose rebased byte offsets are hoff[0 .. n_chunks] (hoff[0] == 0) and whose
// text starts at `text` (host or device according to mem).  On failure the partially built segment is released.
static int32_t ix_build_segment(dvb_ctx* ctx, dvb_scope& sc, cudaStream_t st, int32_t mem, const uint8_t* text, const int64_t* hoff, int64_t n_chunks, IxSeg* sg) {
    const int64_t nb = hoff[n_chunks];
    sg->n_chunks = n_chunks; sg->n_bytes = nb;
    uint8_t* d_text = nullptr; int32_t* d_pc = nullptr; int64_t* d_po = nullptr; int64_t* d_scan = nullptr;
    unsigned long long* s_pk = nullptr; unsigned long long* t_pk = nullptr; uint32_t* t_pv = nullptr; uint32_t* d_hist = nullptr; int32_t* d_bc = nullptr; int64_t* d_bo = nullptr;
    auto drop = [&]() { void* t[] = { d_text, d_pc, d_po, d_scan, s_pk, t_pk, t_pv, d_hist, d_bc, d_bo }; for (void* p : t) if (p) cudaFreeAsync(p, st); };
#define IX_TRY(expr) do { cudaError_t e2__ = (expr); if (e2__ != cudaSuccess) { \
        dvb_set_err(ctx, "CUDA error %s at %s", cudaGetErrorString(e2__), #expr); drop(); cudaStreamSynchronize(st); ix_free_seg(*sg); cudaGetLastError(); \
        return e2__ == cudaErrorMemoryAllocation ? DVB_E_NOMEM : DVB_E_CUDA; } } while (0)
    IX_TRY(cudaMallocAsync((void**)&sg->sym, 2 * (size_t)nb + 64, st));
    IX_TRY(cudaMallocAsync((void**)&sg->sym_off, 8 * (size_t)(n_chunks + 1), st));
    IX_TRY(cudaMallocAsync((void**)&sg->sym_cnt, 4 * (size_t)n_chunks + 4, st));
    IX_TRY(cudaMemcpyAsync(sg->sym_off, hoff, 8 * (size_t)(n_chunks + 1), cudaMemcpyHostToDevice, st));
    const uint8_t* d_utf8 = text;
    if (mem == DVB_MEM_HOST) {
        IX_TRY(cudaMallocAsync((void**)&d_text, (size_t)nb + 64, st));
        if ((size_t)nb >= DVB_H2D_MIN && dvb_is_pageable(text)) IX_TRY(dvb_h2d_staged(sc.l, ctx->device, d_text, text, (size_t)nb, st));
        else IX_TRY(cudaMemcpyAsync(d_text, text, (size_t)nb, cudaMemcpyHostToDevice, st));
        d_utf8 = d_text;
    }
    // K1: symbols (sym_off doubles as the rebased byte offsets the symboliser reads)
    dv_launch_symbolize(ctx, st, d_utf8, sg->sym_off, n_chunks, sg->sym, sg->sym_cnt);
    // K2a: postings
    IX_TRY(cudaMallocAsync((void**)&d_pc, 4 * (size_t)n_chunks + 4, st));
    IX_TRY(cudaMallocAsync((void**)&d_po, 8 * (size_t)(n_chunks + 1), st));
    const int grid = (int)std::min<int64_t>((n_chunks + 7) / 8, (int64_t)ctx->num_sm * 16);
    IX_TRY(cudaMemsetAsync(d_po, 0, 8, st));                   // d_po[0] doubles as the error flag until the scan overwrites it
    k_ix_count<<<grid, 256, 0, st>>>(sg->sym, sg->sym_off, sg->sym_cnt, n_chunks, d_pc, reinterpret_cast<int32_t*>(d_po));
    int32_t too_long = 0;
    IX_TRY(cudaMemcpyAsync(&too_long, d_po, 4, cudaMemcpyDeviceToHost, st));
    {                                                    // exclusive scan of the per-chunk posting counts
        const int64_t nsb = (n_chunks + IX_SCAN_TILE - 1) / IX_SCAN_TILE;
        IX_TRY(cudaMallocAsync((void**)&d_scan, 8 * (size_t)(nsb + 1) + 4 * (size_t)nsb + 16, st));
        int64_t* d_sbase = d_scan; int32_t* d_ssum = reinterpret_cast<int32_t*>(d_scan + nsb + 1);
        k_ix_scan_sum<<<(unsigned)nsb, 256, 0, st>>>(d_pc, n_chunks, d_ssum);
        k_ix_scan<<<1, 1024, 0, st>>>(d_ssum, d_sbase, nsb);
        k_ix_scan_apply<<<(unsigned)nsb, 256, 0, st>>>(d_pc, n_chunks, d_sbase, d_po, nsb);
    }
    int64_t n_post = 0;
    IX_TRY(cudaMemcpyAsync(&n_post, d_po + n_chunks, 8, cudaMemcpyDeviceToHost, st));
    IX_TRY(cudaStreamSynchronize(st));
    if (too_long) {
        drop(); cudaStreamSynchronize(st); ix_free_seg(*sg);
        dvb_set_err(ctx, "dvb_chunk_index_build: a chunk has more than 65535 characters%s%s"); return DVB_E_INVALID;
    }
    ctx->launches += 4;
    sg->n_post = n_post;
    if (d_text) { cudaFreeAsync(d_text, st); d_text = nullptr; }
    if (n_post > 0) {
        IX_TRY(cudaMallocAsync((void**)&s_pk, 8 * (size_t)n_post, st)); IX_TRY(cudaMallocAsync((void**)&sg->pv, 4 * (size_t)n_post, st));
        IX_TRY(cudaMallocAsync((void**)&t_pk, 8 * (size_t)n_post, st)); IX_TRY(cudaMallocAsync((void**)&t_pv, 4 * (size_t)n_post, st));
        IX_TRY(cudaMallocAsync((void**)&d_hist, 4 * dv_radix_hist_words(n_post), st));
        // K2a + K2b: the first radix pass generates the postings from the symbols (no emission-order copy in HBM), the other
        // three move 12-byte postings; stable, so postings of one key stay ordered by (chunk, position)
        int64_t l = 0;
        int which;
        // Measured at 1 M chunks (profiles/README.md): generating the postings inside the first pass (DVB_IX_FUSED=1) costs more
        // than it saves — 23.96 ms against 20.40 ms: the symbols are walked twice (histogram, scatter) by one warp per chunk
        // with ~9 chunks per 8-warp block — so the default emits once and sorts; the fused kernels stay as the A/B arm.
        static const bool fused = [] { const char* e = getenv("DVB_IX_FUSED"); return e && e[0] == '1'; }();
        if (!fused) {
            k_ix_emit<<<grid, 256, 0, st>>>(sg->sym, sg->sym_off, sg->sym_cnt, n_chunks, d_po, s_pk, sg->pv, dv_radix_digits_ptr(d_hist, n_post));
            ctx->launches += 1;
            which = dv_radix_sort_pairs(s_pk, sg->pv, t_pk, t_pv, n_post, 0, 32, d_hist, st, &l, /*first_digits_ready=*/true);
        } else {
            const int64_t nrb = (n_post + RS_TILE - 1) / RS_TILE;
            cudaFuncSetAttribute(k_ix_scatter1, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(RsSmem));
            k_ix_hist1<<<(unsigned)nrb, RS_THREADS, 0, st>>>(sg->sym, sg->sym_off, sg->sym_cnt, n_chunks, d_po, n_post, d_hist, nrb);
            k_rs_scan<<<256, 1024, 0, st>>>(d_hist, nrb, dv_radix_totals_ptr(d_hist, n_post));
            k_rs_base<<<1, 256, 0, st>>>(dv_radix_totals_ptr(d_hist, n_post), dv_radix_dbase_ptr(d_hist, n_post));
            k_ix_scatter1<<<(unsigned)nrb, RS_THREADS, sizeof(RsSmem), st>>>(sg->sym, sg->sym_off, sg->sym_cnt, n_chunks, d_po, n_post, t_pk, t_pv,
                                                                               d_hist, nrb, dv_radix_digits_ptr(d_hist, n_post));
            ctx->launches += 4;
            // passes 2..4 read t_*, so the ping-pong roles are swapped for the call: result 0 = t_*, 1 = s_*
            which = 1 - dv_radix_sort_pairs(t_pk, t_pv, s_pk, sg->pv, n_post, 8, 32, d_hist, st, &l, /*first_digits_ready=*/true);
        }
        ctx->launches += l;
        if (which) { std::swap(s_pk, t_pk); std::swap(sg->pv, t_pv); }
        cudaFreeAsync(t_pk, st); t_pk = nullptr; cudaFreeAsync(t_pv, st); t_pv = nullptr;      // the pool hands these bytes to pc below
        IX_TRY(cudaMallocAsync((void**)&sg->pc, 4 * (size_t)n_post + 64, st));
        // K2c: segment heads
        const int64_t nhb = (n_post + IX_HEAD_TILE - 1) / IX_HEAD_TILE;
        IX_TRY(cudaMallocAsync((void**)&d_bc, 4 * (size_t)nhb + 4, st)); IX_TRY(cudaMallocAsync((void**)&d_bo, 8 * (size_t)(nhb + 1), st));
        k_ix_head_count<<<(unsigned)nhb, 256, 0, st>>>(s_pk, n_post, d_bc);
        k_ix_scan<<<1, 1024, 0, st>>>(d_bc, d_bo, nhb);
        int64_t n_keys = 0;
        IX_TRY(cudaMemcpyAsync(&n_keys, d_bo + nhb, 8, cudaMemcpyDeviceToHost, st));
        IX_TRY(cudaStreamSynchronize(st));
        sg->n_keys = n_keys;
        IX_TRY(cudaMallocAsync((void**)&sg->ukeys, 4 * (size_t)n_keys + 4, st)); IX_TRY(cudaMallocAsync((void**)&sg->uoff, 8 * (size_t)(n_keys + 1), st));
        k_ix_head_write<<<(unsigned)nhb, 256, 0, st>>>(s_pk, n_post, d_bo, sg->ukeys, sg->uoff, sg->pc);
        IX_TRY(cudaMemcpyAsync(sg->uoff + n_keys, &n_post, 8, cudaMemcpyHostToDevice, st));
        IX_TRY(cudaStreamSynchronize(st));                     // n_post is a stack variable
        ctx->launches += 3;
    }
    drop();
#undef IX_TRY
    return DVB_OK;
}

extern "C" int32_t dvb_chunk_index_build(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_chunks, dvb_index** out) {
    if (!ctx || !out || n_chunks < 0 || !off || n_chunks >= (1LL << 32)) { dvb_set_err(ctx, "dvb_chunk_index_build: invalid argument%s%s"); return DVB_E_INVALID; }
    *out = nullptr;
    dvb_scope sc(ctx);
    cudaStream_t st = sc.stream();
    dvb_index* ix = new dvb_index();
    ix->device = ctx->device; ix->n_chunks = n_chunks;
    std::vector<int64_t> hoff((size_t)n_chunks + 1);
    if (mem == DVB_MEM_HOST) memcpy(hoff.data(), off, 8 * (size_t)(n_chunks + 1));
    else {
        cudaError_t e = cudaMemcpyAsync(hoff.data(), off, 8 * (size_t)(n_chunks + 1), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { dvb_set_err(ctx, "CUDA error %s at %s", cudaGetErrorString(e), "copy of the chunk offsets"); delete ix; cudaGetLastError(); return DVB_E_CUDA; }
    }
    const int64_t o0 = hoff[0];
    for (int64_t c = 0; c <= n_chunks; c++) hoff[c] -= o0;
    for (int64_t c = 0; c < n_chunks; c++) if (hoff[c + 1] < hoff[c] || hoff[c + 1] - hoff[c] > 0x7fffffffLL / 2) {
        dvb_set_err(ctx, "dvb_chunk_index_build: bad offsets%s%s"); delete ix; return DVB_E_INVALID; }
    ix->n_bytes = hoff[n_chunks];
    const bool timing = ctx->timing.load() != 0;
    if (timing) cudaEventRecord(sc.l->ev0, st);
    // cut into segments of <= seg_bytes of text (a single chunk larger than that is its own segment)
    const int64_t seg_bytes = ix_seg_bytes();
    std::vector<int64_t> seg_off;
    if (n_chunks == 0) ix->segs.push_back(IxSeg());              // an empty index answers every query with "no positive chunk"
    for (int64_t c0 = 0; c0 < n_chunks; ) {
        // largest c1 with hoff[c1] - hoff[c0] <= seg_bytes (at least one chunk)
        int64_t c1 = std::upper_bound(hoff.begin() + c0, hoff.end(), hoff[c0] + seg_bytes) - hoff.begin() - 1;
        c1 = std::max(c1, c0 + 1);
        seg_off.assign(hoff.begin() + c0, hoff.begin() + c1 + 1);
        const int64_t b0 = seg_off[0];
        for (int64_t& v : seg_off) v -= b0;
        IxSeg sg; sg.chunk_base = c0;
        const int32_t rc = ix_build_segment(ctx, sc, st, mem, utf8 + o0 + b0, seg_off.data(), c1 - c0, &sg);
        if (rc != DVB_OK) { dvb_chunk_index_free(ix); return rc; }
        ix->n_post += sg.n_post; ix->n_keys += sg.n_keys;
        ix->segs.push_back(sg);
        c0 = c1;
    }
    if (timing) cudaEventRecord(sc.l->ev1, st);
    cudaError_t e = cudaStreamSynchronize(st);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { dvb_set_err(ctx, "CUDA error %s at %s", cudaGetErrorString(e), "dvb_chunk_index_build"); dvb_chunk_index_free(ix); cudaGetLastError(); return DVB_E_CUDA; }
    if (timing) { float ms = 0.f; cudaEventElapsedTime(&ms, sc.l->ev0, sc.l->ev1); ctx->last_ms = ms; }
    *out = ix;
    return DVB_OK;
}

extern "C" int32_t dvb_chunk_index_topk(dvb_ctx* ctx, const dvb_index* ix, int32_t mem, const uint16_t* term_sym, const int32_t* term_off, int32_t n_terms,
                                         const int32_t* query_term_off, int32_t nq, int64_t index_base, int32_t k,
                                         int64_t* out_idx, int32_t* out_score, int32_t* out_n) {
    if (!ctx || !ix || nq < 0 || n_terms < 0 || k < 1 || k > 32 || !term_off || !query_term_off || !out_idx || !out_score || !out_n ||
        ix->device != ctx->device) { dvb_set_err(ctx, "dvb_chunk_index_topk: invalid argument (1 <= k <= 32, index of this device)%s%s"); return DVB_E_INVALID; }
    if (nq == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    int64_t N = 0;                                               // largest segment: sizes the dense score tile
    for (const IxSeg& sg : ix->segs) N = std::max(N, sg.n_chunks);
    // terms: sizes come from the offsets
    int32_t n_sym = 0;
    if (mem == DVB_MEM_HOST) n_sym = term_off[n_terms];
    else { DVB_CUDA(ctx, cudaMemcpyAsync(&n_sym, term_off + n_terms, 4, cudaMemcpyDeviceToHost, io.st)); DVB_CUDA(ctx, cudaStreamSynchronize(io.st)); }
    std::vector<int32_t> h_qto((size_t)nq + 1);
    if (mem == DVB_MEM_HOST) memcpy(h_qto.data(), query_term_off, 4 * (size_t)(nq + 1));
    else { DVB_CUDA(ctx, cudaMemcpyAsync(h_qto.data(), query_term_off, 4 * (size_t)(nq + 1), cudaMemcpyDeviceToHost, io.st)); DVB_CUDA(ctx, cudaStreamSynchronize(io.st)); }
    if (h_qto[0] != 0 || h_qto[nq] != n_terms) { dvb_set_err(ctx, "dvb_chunk_index_topk: query_term_off must span the terms%s%s"); return DVB_E_INVALID; }
    // batch of queries sharing one dense score tile (<= 512 MB)
    const int64_t budget = 512LL << 20;
    const int32_t B = (int32_t)std::max<int64_t>(1, std::min<int64_t>(nq, N > 0 ? budget / (4 * N) : nq));
    const size_t n_part = (size_t)B + (size_t)ctx->num_sm * 8;     // nb * dv_topk_splits(., nb) <= max(nb, 8 * num_sm) for every nb <= B
    size_t need = dvb_pad(4 * (size_t)B * (size_t)std::max<int64_t>(N, 1)) + dvb_pad(8 * (size_t)n_terms + 16) * 3 + dvb_pad(4 * (size_t)n_terms + 4) * 2
                + dvb_pad(4 * n_part * k) + dvb_pad(8 * n_part * k) + 4096;
    if (mem == DVB_MEM_HOST) need += dvb_pad(2 * (size_t)n_sym + 2) + dvb_pad(4 * (size_t)n_terms + 4) + dvb_pad(4 * (size_t)nq + 4)
                                   + dvb_pad(8 * (size_t)nq * k) + dvb_pad(4 * (size_t)nq * k) + dvb_pad(4 * (size_t)nq);
    int32_t rc = sc.reserve(need); if (rc) return rc;
    const uint16_t* d_ts = io.in(term_sym, (size_t)n_sym); const int32_t* d_to = io.in(term_off, (size_t)n_terms + 1);
    const int32_t* d_qto = io.in(query_term_off, (size_t)nq + 1);
    int64_t* d_oidx = io.out(out_idx, (size_t)nq * k); int32_t* d_osc = io.out(out_score, (size_t)nq * k); int32_t* d_on = io.out(out_n, (size_t)nq);
    int64_t* d_lo = io.tmp<int64_t>((size_t)n_terms + 1); int64_t* d_hi = io.tmp<int64_t>((size_t)n_terms + 1); int32_t* d_j = io.tmp<int32_t>((size_t)n_terms + 1);
    int32_t* d_err = io.tmp<int32_t>(1);
    int32_t* d_ni = io.tmp<int32_t>((size_t)n_terms + 1); int64_t* d_io = io.tmp<int64_t>((size_t)n_terms + 2);
    int32_t* d_score = io.tmp<int32_t>((size_t)B * (size_t)std::max<int64_t>(N, 1));
    int32_t* d_ps = io.tmp<int32_t>(n_part * k); int64_t* d_pi = io.tmp<int64_t>(n_part * k);
    rc = io.check(ctx); if (rc) return rc;
    DVB_CUDA(ctx, cudaMemsetAsync(d_err, 0, 4, io.st));
    const bool timing = ctx->timing.load() != 0;
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev0, io.st));
    // segment by segment: the running top-k of a query (out arrays) is one more candidate list of the next segment's merge
    bool first = true;
    for (const IxSeg& sg : ix->segs) {
        const int64_t Ns = sg.n_chunks;
        if (n_terms > 0) {
            k_ix_plan<<<std::min((n_terms + 255) / 256, ctx->num_sm * 8), 256, 0, io.st>>>(d_ts, d_to, n_terms, sg.ukeys, sg.uoff, sg.n_keys, d_lo, d_hi, d_j, d_err);
            k_ix_items<<<std::min((n_terms + 255) / 256, ctx->num_sm * 8), 256, 0, io.st>>>(d_lo, d_hi, n_terms, d_ni);
            k_ix_scan<<<1, 1024, 0, io.st>>>(d_ni, d_io, n_terms);
            ctx->launches += 3;
        }
        for (int32_t q0 = 0; q0 < nq; q0 += B) {
            const int32_t q1 = std::min(nq, q0 + B), nb = q1 - q0;
            if (Ns > 0) {
                DVB_CUDA(ctx, cudaMemsetAsync(d_score, 0, 4 * (size_t)nb * (size_t)Ns, io.st));
                const int32_t nt = h_qto[q1] - h_qto[q0];
                if (nt > 0) {
                    k_ix_score<<<ctx->num_sm * 16, 256, 0, io.st>>>(sg, d_ts, d_to, d_qto, q0, q1, d_lo, d_hi, d_j, d_io, d_score);
                    ctx->launches += 1;
                }
            }
            dv_topk_device(ctx, io.st, d_score, nullptr, Ns, nb, index_base + sg.chunk_base, k, dv_topk_splits(ctx, Ns, nb), d_ps, d_pi,
                           d_oidx + (size_t)q0 * k, d_osc + (size_t)q0 * k, d_on + q0, /*merge_prev=*/!first);
        }
        first = false;
    }
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev1, io.st));
    int32_t h_err = 0;
    DVB_CUDA(ctx, cudaMemcpyAsync(&h_err, d_err, 4, cudaMemcpyDeviceToHost, io.st));
    rc = io.finish(ctx);
    if (rc == DVB_OK && timing) { float ms = 0.f; cudaEventElapsedTime(&ms, sc.l->ev0, sc.l->ev1); ctx->last_ms = ms; }
    if (rc == DVB_OK && h_err) { dvb_set_err(ctx, "dvb_chunk_index_topk: a term has < 2 symbols, a separator, or no same-class pair%s%s"); return DVB_E_INVALID; }
    return rc;
}
