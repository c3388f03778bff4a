// chunk.cu — chunk scoring and warp-level top-k.
// replaces: the scoring loop and top-k of select_reference_material_context
//           (reference web/server.py:32073-32087):
//               score = sum(lowered.count(term) for term in terms)      # non-overlapping substring count
//               sorted(positive, key=(-score, index))[:max_chunks]
//
// dvb_symbolize     text -> dense lowered SYMBOLS (u16: 0 separator, 1..20992 CJK code, 0x8000|code ASCII).
//                   Query terms only ever contain ASCII-class / CJK symbols (web/server.py:32039), so
//                   str.count(term) on the lowered chunk equals a substring count on this stream.
// dvb_chunk_scores  exact scan scorer: one warp per (query, chunk); greedy left-to-right
//                   non-overlapping matching == CPython's str.count.  Meant for API-sized inputs
//                   (one document, one query); the inverted-index path (index.cu) serves bulk sizes.
// dvb_topk          warp-level top-k (k <= 32) under the total order (score desc, index asc), also used
//                   to merge per-shard / per-GPU candidate lists.
#include "common.cuh"
#include "text.cuh"

// ------------------------------------------------------------------ symbolise
// returns the number of symbols written (warp-uniform)
__device__ int dv_symbolize_warp(const uint8_t* text, int32_t n, dv_warp_buf* wb, uint16_t* out) {
    const int lane = threadIdx.x & 31;
    if (n <= 0) return 0;
    const int32_t a = (int32_t)((uintptr_t)text & 15u);
    uint32_t* R = wb->raw + lane * 5;
    uint32_t carry_word = 0;
    int base = 0;
    bool serial = false;
    for (int32_t tile = -a; tile < n; tile += DV_TILE) {
        uint32_t m16 = dv_tile_masks(text, n, tile, lane, carry_word, R);
        const int cnt = __popc(m16);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += y; }
        int widx = base + incl - cnt;
        bool c4 = false;
        while (m16) {
            const int t = __ffs(m16) - 1; m16 &= m16 - 1;
            out[widx++] = (uint16_t)dv_decode_sym(R, t, c4);
        }
        if (__any_sync(0xffffffffu, c4)) { serial = true; break; }
        base += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (!serial) return base;
    // exact serial path (U+0130 -> 'i' + separator)
    int k = 0;
    if (lane == 0) {
        int32_t i = 0;
        while (i < n) {
            uint32_t b = text[i], cp; int len;
            if (b < 0x80u) { cp = b; len = 1; }
            else if (b < 0xE0u) { cp = b & 0x1Fu; len = 2; }
            else if (b < 0xF0u) { cp = b & 0x0Fu; len = 3; }
            else { cp = b & 0x07u; len = 4; }
            for (int j = 1; j < len && i + j < n; j++) cp = (cp << 6) | (text[i + j] & 0x3Fu);
            i += len;
            uint32_t s = 0;
            if (cp < 0x80u) s = dv_ascii_sym(cp);
            else if (cp - 0x4E00u < 20992u) s = cp - 0x4E00u + 1u;
            else if (cp == 0x212Au) s = DV_SYM_ASCII | 11u;
            else if (cp == 0x0130u) { out[k++] = (uint16_t)(DV_SYM_ASCII | 9u); s = 0; }
            out[k++] = (uint16_t)s;
        }
    }
    return __shfl_sync(0xffffffffu, k, 0);
}

// one warp per text; symbols of text t go to sym[off[t] - off[0] ...] (a text never has more symbols than bytes)
__global__ void __launch_bounds__(256) k_symbolize(const uint8_t* __restrict__ utf8, const int64_t* __restrict__ off, int64_t n_texts,
                                                    uint16_t* __restrict__ sym, int32_t* __restrict__ sym_cnt) {
    __shared__ dv_warp_buf wbuf[8];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t base0 = off[0];
    for (int64_t t = (int64_t)blockIdx.x * 8 + warp; t < n_texts; t += (int64_t)gridDim.x * 8) {
        const int64_t b = off[t], e = off[t + 1];
        const int c = dv_symbolize_warp(utf8 + b, (int32_t)(e - b), &wbuf[warp], sym + (b - base0));
        if (lane == 0) sym_cnt[t] = c;
        __syncwarp();
    }
}

// ------------------------------------------------------------------ scan scorer
// grid.x strides over (query, chunk) pairs, one warp each.
__global__ void __launch_bounds__(256) k_count_terms(const uint16_t* __restrict__ sym, const int64_t* __restrict__ off, const int32_t* __restrict__ sym_cnt,
                                                      int64_t n_chunks, const uint16_t* __restrict__ term_sym, const int32_t* __restrict__ term_off,
                                                      const int32_t* __restrict__ query_term_off, int32_t nq, int32_t* __restrict__ score) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t base0 = off[0];
    const int64_t n_pairs = (int64_t)nq * n_chunks;
    for (int64_t pr = (int64_t)blockIdx.x * 8 + warp; pr < n_pairs; pr += (int64_t)gridDim.x * 8) {
        const int32_t q = (int32_t)(pr / n_chunks);
        const int64_t c = pr - (int64_t)q * n_chunks;
        const uint16_t* s = sym + (off[c] - base0);
        const int32_t n = sym_cnt[c];
        int32_t total = 0;
        for (int32_t t = query_term_off[q]; t < query_term_off[q + 1]; t++) {
            const uint16_t* p = term_sym + term_off[t];
            const int32_t L = term_off[t + 1] - term_off[t];
            if (L <= 0 || L > n) continue;
            const uint16_t first = p[0];
            int32_t next_ok = 0;
            for (int32_t b0 = 0; b0 + L <= n; b0 += 32) {
                const int32_t i = b0 + lane;
                bool m = (i + L <= n) && (s[i] == first);
                for (int32_t j = 1; m && j < L; j++) m = (s[i + j] == p[j]);
                uint32_t mask = __ballot_sync(0xffffffffu, m);
                while (mask) {                                   // greedy left-to-right == str.count
                    const int32_t pos = b0 + __ffs(mask) - 1; mask &= mask - 1;
                    if (pos >= next_ok) { total++; next_ok = pos + L; }
                }
            }
        }
        if (lane == 0) score[pr] = total;
    }
}

// ------------------------------------------------------------------ warp-level top-k
struct TopEntry { int32_t score; int64_t idx; };
__device__ __forceinline__ bool top_better(int32_t sa, int64_t ia, int32_t sb, int64_t ib) {
    return sa > sb || (sa == sb && ia < ib);
}
// lane j holds the j-th best entry (descending).  Insert candidates held one per lane.
__device__ __forceinline__ void warp_topk_insert(int32_t& es, int64_t& ei, int32_t cs, int64_t ci, bool valid, int k, int lane) {
    const int32_t ks = __shfl_sync(0xffffffffu, es, k - 1);
    const int64_t ki = __shfl_sync(0xffffffffu, ei, k - 1);
    uint32_t pend = __ballot_sync(0xffffffffu, valid && cs > 0 && top_better(cs, ci, ks, ki));
    while (pend) {
        const int src = __ffs(pend) - 1; pend &= pend - 1;
        const int32_t s = __shfl_sync(0xffffffffu, cs, src);
        const int64_t i = __shfl_sync(0xffffffffu, ci, src);
        const uint32_t ahead = __ballot_sync(0xffffffffu, lane < k && top_better(es, ei, s, i));
        const int pos = __popc(ahead);
        const int32_t us = __shfl_up_sync(0xffffffffu, es, 1);
        const int64_t ui = __shfl_up_sync(0xffffffffu, ei, 1);
        if (pos < k) {
            if (lane > pos) { es = us; ei = ui; }
            else if (lane == pos) { es = s; ei = i; }
        }
    }
}

#define TOPK_WARPS 8
// grid = (n_split, nq).  CTA (sp, q) scans candidates [sp*per, (sp+1)*per) of query q and writes k entries
// (score 0 = empty) to part[(q*n_split + sp)*k ...].  idx == nullptr means index = index_base + position.
__global__ void __launch_bounds__(TOPK_WARPS * 32) k_topk_partial(const int32_t* __restrict__ score, const int64_t* __restrict__ idx, int64_t n,
                                                                   int64_t index_base, int32_t k, int64_t per, int32_t nq,
                                                                   int32_t* __restrict__ part_score, int64_t* __restrict__ part_idx) {
    __shared__ int32_t s_s[TOPK_WARPS][32];
    __shared__ int64_t s_i[TOPK_WARPS][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int sp = blockIdx.x, n_split = gridDim.x;
    const int64_t lo = (int64_t)sp * per, hi = min(n, lo + per);
    for (int q = blockIdx.y; q < nq; q += gridDim.y) {
    const int32_t* sc = score + (int64_t)q * n;
    const int64_t* ix = idx ? idx + (int64_t)q * n : nullptr;
    int32_t es = 0; int64_t ei = 0x7fffffffffffffffLL;
    if (ix) {
        for (int64_t b0 = lo + (int64_t)warp * 32; b0 < hi; b0 += TOPK_WARPS * 32) {
            const int64_t i = b0 + lane;
            const bool v = i < hi;
            const int32_t cs = v ? sc[i] : 0;
            const int64_t ci = v ? ix[i] : 0;
            if (__any_sync(0xffffffffu, cs > 0)) warp_topk_insert(es, ei, cs, ci, v, k, lane);
        }
    } else {
        // dense score rows are almost all zero: stream them as 16-byte vectors (two per lane and step), insert rarely
        const int64_t v0 = min(hi, lo + (int64_t)((4 - ((reinterpret_cast<uintptr_t>(sc + lo) >> 2) & 3)) & 3));   // first 16-byte aligned item
        const int64_t nvec = (hi - v0) >> 2, v1 = v0 + 4 * nvec;
        if (warp == 0) {                                          // <= 3 + 3 unaligned items at the two ends
            const int64_t i = lane < 3 ? lo + lane : v1 + (lane - 3);
            const bool v = lane < 3 ? i < v0 : (lane < 6 && i < hi);
            const int32_t cs = v ? sc[i] : 0;
            if (__any_sync(0xffffffffu, cs > 0)) warp_topk_insert(es, ei, cs, index_base + i, v, k, lane);
        }
        const int4* vec = reinterpret_cast<const int4*>(sc + v0);
        for (int64_t g0 = (int64_t)warp * 64; g0 < nvec; g0 += TOPK_WARPS * 64) {
            const int64_t ga = g0 + lane, gb = g0 + 32 + lane;
            int4 a = make_int4(0, 0, 0, 0), b = make_int4(0, 0, 0, 0);
            if (ga < nvec) a = vec[ga];
            if (gb < nvec) b = vec[gb];
            if (__any_sync(0xffffffffu, (a.x | a.y | a.z | a.w | b.x | b.y | b.z | b.w) != 0)) {
                const int64_t ia = index_base + v0 + 4 * ga, ib = index_base + v0 + 4 * gb;
                if (__any_sync(0xffffffffu, a.x > 0)) warp_topk_insert(es, ei, a.x, ia, true, k, lane);
                if (__any_sync(0xffffffffu, a.y > 0)) warp_topk_insert(es, ei, a.y, ia + 1, true, k, lane);
                if (__any_sync(0xffffffffu, a.z > 0)) warp_topk_insert(es, ei, a.z, ia + 2, true, k, lane);
                if (__any_sync(0xffffffffu, a.w > 0)) warp_topk_insert(es, ei, a.w, ia + 3, true, k, lane);
                if (__any_sync(0xffffffffu, b.x > 0)) warp_topk_insert(es, ei, b.x, ib, true, k, lane);
                if (__any_sync(0xffffffffu, b.y > 0)) warp_topk_insert(es, ei, b.y, ib + 1, true, k, lane);
                if (__any_sync(0xffffffffu, b.z > 0)) warp_topk_insert(es, ei, b.z, ib + 2, true, k, lane);
                if (__any_sync(0xffffffffu, b.w > 0)) warp_topk_insert(es, ei, b.w, ib + 3, true, k, lane);
            }
        }
    }
    s_s[warp][lane] = es; s_i[warp][lane] = ei;
    __syncthreads();
    if (warp == 0) {
        for (int w = 1; w < TOPK_WARPS; w++) {
            const int32_t cs = s_s[w][lane]; const int64_t ci = s_i[w][lane];
            warp_topk_insert(es, ei, cs, ci, lane < k, k, lane);
        }
        if (lane < k) {
            part_score[((int64_t)q * n_split + sp) * k + lane] = es;
            part_idx[((int64_t)q * n_split + sp) * k + lane] = ei;
        }
    }
    __syncthreads();
    }
}

// one warp per query merges m candidate lists of k entries; writes k entries + the number of valid ones
// prev != 0: out_score/out_idx already hold a query's running top-k (an earlier index segment); it is merged as one more list.
__global__ void __launch_bounds__(256) k_topk_final(const int32_t* __restrict__ part_score, const int64_t* __restrict__ part_idx, int32_t m, int32_t k, int32_t nq,
                                                     int32_t* out_score, int64_t* out_idx, int32_t* __restrict__ out_n, int32_t prev) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int q = blockIdx.x * 8 + warp; q < nq; q += gridDim.x * 8) {
        int32_t es = 0; int64_t ei = 0x7fffffffffffffffLL;
        if (prev && lane < k) {                                   // already sorted, positives first
            es = out_score[(int64_t)q * k + lane];
            if (es > 0) ei = out_idx[(int64_t)q * k + lane];
        }
        const int64_t tot = (int64_t)m * k;
        for (int64_t b0 = 0; b0 < tot; b0 += 32) {
            const int64_t i = b0 + lane; const bool v = i < tot;
            const int32_t cs = v ? part_score[(int64_t)q * tot + i] : 0;
            const int64_t ci = v ? part_idx[(int64_t)q * tot + i] : 0;
            if (__any_sync(0xffffffffu, cs > 0)) warp_topk_insert(es, ei, cs, ci, v, k, lane);
        }
        if (lane < k) { out_score[(int64_t)q * k + lane] = es; out_idx[(int64_t)q * k + lane] = es > 0 ? ei : -1; }
        const uint32_t valid = __ballot_sync(0xffffffffu, lane < k && es > 0);
        if (lane == 0) out_n[q] = __popc(valid);
    }
}

// ------------------------------------------------------------------ entry points
extern "C" int32_t dvb_symbolize(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_texts,
                                  uint16_t* out_sym, int32_t* out_cnt) {
    if (!ctx || n_texts < 0 || !off || !out_sym || !out_cnt) { dvb_set_err(ctx, "dvb_symbolize: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_texts == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    int64_t ends[2];
    if (mem == DVB_MEM_HOST) { ends[0] = off[0]; ends[1] = off[n_texts]; }
    else {
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[0], off, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[1], off + n_texts, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    }
    const int64_t total = ends[1] - ends[0];
    if (total < 0) { dvb_set_err(ctx, "dvb_symbolize: bad offsets%s%s"); return DVB_E_INVALID; }
    const uint8_t* d_utf8 = mem == DVB_MEM_HOST ? io.in(utf8 + ends[0], (size_t)total + 16) - ends[0] : utf8;
    const int64_t* d_off = io.in(off, (size_t)n_texts + 1);
    uint16_t* d_sym = io.out(out_sym, (size_t)total);
    int32_t* d_cnt = io.out(out_cnt, (size_t)n_texts);
    int32_t rc = io.check(ctx); if (rc) return rc;
    const int grid = (int)std::min<int64_t>((n_texts + 7) / 8, (int64_t)ctx->num_sm * 8);
    k_symbolize<<<grid, 256, 0, io.st>>>(d_utf8, d_off, n_texts, d_sym, d_cnt);
    ctx->launches += 1;
    return io.finish(ctx);
}

extern "C" int32_t dvb_chunk_scores(dvb_ctx* ctx, int32_t mem, const uint8_t* chunk_utf8, const int64_t* chunk_off, int64_t n_chunks,
                                     const uint16_t* term_sym, const int32_t* term_off, int32_t n_terms,
                                     const int32_t* query_term_off, int32_t nq, int32_t* out_score) {
    if (!ctx || n_chunks < 0 || nq < 0 || !chunk_off || !term_off || !query_term_off || !out_score) { dvb_set_err(ctx, "dvb_chunk_scores: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_chunks == 0 || nq == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    int64_t ends[2]; int32_t n_tsym = 0;
    if (mem == DVB_MEM_HOST) { ends[0] = chunk_off[0]; ends[1] = chunk_off[n_chunks]; n_tsym = term_off[n_terms]; }
    else {
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[0], chunk_off, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[1], chunk_off + n_chunks, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(&n_tsym, term_off + n_terms, 4, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    }
    const int64_t total = ends[1] - ends[0];
    const uint8_t* d_utf8 = mem == DVB_MEM_HOST ? io.in(chunk_utf8 + ends[0], (size_t)total + 16) - ends[0] : chunk_utf8;
    const int64_t* d_off = io.in(chunk_off, (size_t)n_chunks + 1);
    const uint16_t* d_tsym = io.in(term_sym, (size_t)n_tsym);
    const int32_t* d_toff = io.in(term_off, (size_t)n_terms + 1);
    const int32_t* d_qoff = io.in(query_term_off, (size_t)nq + 1);
    int32_t* d_score = io.out(out_score, (size_t)nq * (size_t)n_chunks);
    uint16_t* d_sym = io.tmp<uint16_t>((size_t)total + 8);
    int32_t* d_cnt = io.tmp<int32_t>((size_t)n_chunks);
    int32_t rc = io.check(ctx); if (rc) return rc;
    const int grid1 = (int)std::min<int64_t>((n_chunks + 7) / 8, (int64_t)ctx->num_sm * 8);
    k_symbolize<<<grid1, 256, 0, io.st>>>(d_utf8, d_off, n_chunks, d_sym, d_cnt);
    const int64_t pairs = (int64_t)nq * n_chunks;
    const int grid2 = (int)std::min<int64_t>((pairs + 7) / 8, (int64_t)ctx->num_sm * 16);
    k_count_terms<<<grid2, 256, 0, io.st>>>(d_sym, d_off, d_cnt, n_chunks, d_tsym, d_toff, d_qoff, nq, d_score);
    ctx->launches += 2;
    return io.finish(ctx);
}

// host-side launchers shared with index.cu (device pointers, caller's stream)
void dv_launch_symbolize(dvb_ctx* ctx, cudaStream_t st, const uint8_t* d_utf8, const int64_t* d_off, int64_t n_texts, uint16_t* d_sym, int32_t* d_cnt) {
    const int grid = (int)std::min<int64_t>((n_texts + 7) / 8, (int64_t)ctx->num_sm * 8);
    k_symbolize<<<grid, 256, 0, st>>>(d_utf8, d_off, n_texts, d_sym, d_cnt);
    ctx->launches += 1;
}
int64_t dv_topk_splits(dvb_ctx* ctx, int64_t n, int32_t nq) {
    return std::max<int64_t>(1, std::min<int64_t>((n + 8191) / 8192, std::max<int64_t>(1, (int64_t)ctx->num_sm * 8 / std::max(1, nq))));
}
// d_ps / d_pi: nq * n_split * k scratch entries each
void dv_topk_device(dvb_ctx* ctx, cudaStream_t st, const int32_t* d_score, const int64_t* d_idx, int64_t n, int32_t nq, int64_t index_base, int32_t k,
                    int64_t n_split, int32_t* d_ps, int64_t* d_pi, int64_t* d_oidx, int32_t* d_oscore, int32_t* d_on, bool merge_prev) {
    const int64_t per = (n + n_split - 1) / n_split;
    dim3 g1((unsigned)n_split, (unsigned)std::min(nq, 65535));
    k_topk_partial<<<g1, TOPK_WARPS * 32, 0, st>>>(d_score, d_idx, n, index_base, k, per > 0 ? per : 1, nq, d_ps, d_pi);
    const int grid2 = std::min((nq + 7) / 8, ctx->num_sm * 8);
    k_topk_final<<<grid2, 256, 0, st>>>(d_ps, d_pi, (int32_t)n_split, k, nq, d_oscore, d_oidx, d_on, merge_prev ? 1 : 0);
    ctx->launches += 2;
}

extern "C" int32_t dvb_topk(dvb_ctx* ctx, int32_t mem, const int32_t* score, const int64_t* idx, int64_t n, int32_t nq,
                             int64_t index_base, int32_t k, int64_t* out_idx, int32_t* out_score, int32_t* out_n) {
    if (!ctx || n < 0 || nq < 0 || k < 1 || k > 32 || !out_idx || !out_score || !out_n || (n > 0 && !score)) { dvb_set_err(ctx, "dvb_topk: invalid argument (1 <= k <= 32)%s%s"); return DVB_E_INVALID; }
    if (nq == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    const int32_t* d_score = io.in(score, (size_t)nq * (size_t)n);
    const int64_t* d_idx = io.in(idx, (size_t)nq * (size_t)n);
    int64_t* d_oidx = io.out(out_idx, (size_t)nq * k);
    int32_t* d_oscore = io.out(out_score, (size_t)nq * k);
    int32_t* d_on = io.out(out_n, (size_t)nq);
    int64_t n_split = std::max<int64_t>(1, std::min<int64_t>((n + 8191) / 8192, std::max<int64_t>(1, (int64_t)ctx->num_sm * 4 / std::max(1, nq))));
    const int64_t per = n_split ? (n + n_split - 1) / n_split : 0;
    int32_t* d_ps = io.tmp<int32_t>((size_t)nq * n_split * k);
    int64_t* d_pi = io.tmp<int64_t>((size_t)nq * n_split * k);
    int32_t rc = io.check(ctx); if (rc) return rc;
    dim3 g1((unsigned)n_split, (unsigned)std::min(nq, 65535));
    k_topk_partial<<<g1, TOPK_WARPS * 32, 0, io.st>>>(d_score, d_idx, n, index_base, k, per > 0 ? per : 1, nq, d_ps, d_pi);
    const int grid2 = std::min((nq + 7) / 8, ctx->num_sm * 8);
    k_topk_final<<<grid2, 256, 0, io.st>>>(d_ps, d_pi, (int32_t)n_split, k, nq, d_oscore, d_oidx, d_on, 0);
    ctx->launches += 2;
    return io.finish(ctx);
}

// ------------------------------------------------------------------ dvb_topk_static: the same two kernels, no allocation
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, serving loops, harnesses that forbid it).  The partial lists live in ONE file-scope device array
// that every GPU loading the library holds until exit.  Host code may not pass the array's name to a kernel (it is a
// host-side stand-in, not a device address): the address is resolved once per device with cudaGetSymbolAddress while
// that device is current (dv_topk_static_setup), and only that pointer reaches the kernels.
static __device__ __align__(256) unsigned char g_topk_static_scratch[DVB_TOPK_STATIC_SCRATCH_BYTES];
#define DV_TOPK_STATIC_MAX_DEVICES 64
static std::mutex g_topk_static_mu;
static unsigned char* g_topk_static_ptr[DV_TOPK_STATIC_MAX_DEVICES] = { nullptr };

// `device` must be current.  Called from dvb_create; the entry point repeats it if a ctx of that device was never created.
cudaError_t dv_topk_static_setup(int device, unsigned char** out) {
    if (device < 0 || device >= DV_TOPK_STATIC_MAX_DEVICES) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> g(g_topk_static_mu);
    if (!g_topk_static_ptr[device]) {
        unsigned char* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&ptr, g_topk_static_scratch);
        if (e != cudaSuccess) return e;
        g_topk_static_ptr[device] = ptr;
    }
    if (out) *out = g_topk_static_ptr[device];
    return cudaSuccess;
}

extern "C" int32_t dvb_topk_static(dvb_ctx* ctx, void* stream, const int32_t* d_score, const int64_t* d_idx, int64_t n, int32_t nq,
                                    int64_t index_base, int32_t k, int64_t* d_out_idx, int32_t* d_out_score, int32_t* d_out_n) {
    if (!ctx || n < 0 || nq < 0 || k < 1 || k > 32 || !d_out_idx || !d_out_score || !d_out_n || (n > 0 && !d_score)) {
        dvb_set_err(ctx, "dvb_topk_static: invalid argument (1 <= k <= 32)%s%s"); return DVB_E_INVALID; }
    if (nq == 0) return DVB_OK;
    // one partial entry = 4 B score + 8 B index; the split count shrinks to what the array holds, a single split must fit
    const int64_t entry_cap = (int64_t)DVB_TOPK_STATIC_SCRATCH_BYTES / 12;
    if ((int64_t)nq * k > entry_cap) {
        char msg[160];
        snprintf(msg, sizeof(msg), "nq * k = %lld partial entries exceed the static scratch (%lld entries of 12 bytes, %d bytes)",
                 (long long)nq * k, (long long)entry_cap, (int)DVB_TOPK_STATIC_SCRATCH_BYTES);
        dvb_set_err(ctx, "dvb_topk_static: %s%s", msg); return DVB_E_CAPACITY;
    }
    int dev = -1;
    DVB_CUDA(ctx, cudaGetDevice(&dev));
    if (dev != ctx->device) DVB_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char* scratch = nullptr;
    DVB_CUDA(ctx, dv_topk_static_setup(ctx->device, &scratch));
    int64_t n_split = std::max<int64_t>(1, std::min<int64_t>((n + 8191) / 8192, std::max<int64_t>(1, (int64_t)ctx->num_sm * 4 / std::max(1, nq))));
    n_split = std::min<int64_t>(n_split, entry_cap / ((int64_t)nq * k));
    const int64_t entries = (int64_t)nq * n_split * k;
    int64_t* d_pi = reinterpret_cast<int64_t*>(scratch);                                  // indices first (8-byte aligned), then scores
    int32_t* d_ps = reinterpret_cast<int32_t*>(scratch + 8 * (size_t)entries);
    const int64_t per = (n + n_split - 1) / n_split;
    cudaStream_t st = (cudaStream_t)stream;
    dim3 g1((unsigned)n_split, (unsigned)std::min(nq, 65535));
    // every (query, split) CTA writes all k entries of its list before k_topk_final reads them: nothing has to start at zero
    k_topk_partial<<<g1, TOPK_WARPS * 32, 0, st>>>(d_score, d_idx, n, index_base, k, per > 0 ? per : 1, nq, d_ps, d_pi);
    const int grid2 = std::min((nq + 7) / 8, ctx->num_sm * 8);
    k_topk_final<<<grid2, 256, 0, st>>>(d_ps, d_pi, (int32_t)n_split, k, nq, d_out_score, d_out_idx, d_out_n, 0);
    ctx->launches += 2;
    DVB_CUDA(ctx, cudaGetLastError());
    return DVB_OK;
}

// ================================================================== a4: chunk builder
// replaces: _build_reference_material_chunks   (reference web/server.py:31932-31970)
// Offsets are CODE-POINT indices (Python str indexing).  k_cp_index builds cp -> byte offset per
// document; k_chunk_walk does the (inherently sequential) window walk with one warp per document,
// the rfind("\n") and strip() scans running 32 code points per step.
__device__ __forceinline__ bool dv_is_space_cp(uint32_t c) {         // str.isspace() / re \s  (29 code points)
    return (c - 9u < 5u) || (c - 0x1Cu < 5u) || c == 0x85u || c == 0xA0u || c == 0x1680u || (c - 0x2000u < 11u) ||
           c == 0x2028u || c == 0x2029u || c == 0x202Fu || c == 0x205Fu || c == 0x3000u;
}
__device__ __forceinline__ uint32_t dv_decode_at(const uint8_t* t, int64_t i, int64_t n) {
    uint32_t b = t[i], cp; int len;
    if (b < 0x80u) return b;
    if (b < 0xE0u) { cp = b & 0x1Fu; len = 2; } else if (b < 0xF0u) { cp = b & 0x0Fu; len = 3; } else { cp = b & 0x07u; len = 4; }
    for (int j = 1; j < len && i + j < n; j++) cp = (cp << 6) | (t[i + j] & 0x3Fu);
    return cp;
}

// one CTA per document (grid-stride).  cp_byte[off[d] + k] = byte offset (within the document) of code point k.
__global__ void __launch_bounds__(256) k_cp_index(const uint8_t* __restrict__ utf8, const int64_t* __restrict__ off, int64_t n_docs,
                                                   uint32_t* __restrict__ cp_byte, int32_t* __restrict__ n_cp) {
    __shared__ int warp_tot[8];
    __shared__ int carry;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t base0 = off[0];
    for (int64_t d = blockIdx.x; d < n_docs; d += gridDim.x) {
        const uint8_t* t = utf8 + off[d];
        const int64_t n = off[d + 1] - off[d];
        uint32_t* out = cp_byte + (off[d] - base0);
        if (threadIdx.x == 0) carry = 0;
        __syncthreads();
        for (int64_t tile = 0; tile < n; tile += 256 * 16) {
            const int64_t p0 = tile + (int64_t)threadIdx.x * 16;
            uint32_t starts = 0;
            for (int j = 0; j < 16; j++) if (p0 + j < n && (t[p0 + j] & 0xC0u) != 0x80u) starts |= 1u << j;
            const int cnt = __popc(starts);
            int incl = cnt;
            for (int s = 1; s < 32; s <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, s); if (lane >= s) incl += y; }
            if (lane == 31) warp_tot[warp] = incl;
            __syncthreads();
            int before = carry;
            for (int w = 0; w < warp; w++) before += warp_tot[w];
            int k = before + incl - cnt;
            while (starts) { const int j = __ffs(starts) - 1; starts &= starts - 1; out[k++] = (uint32_t)(p0 + j); }
            __syncthreads();
            if (threadIdx.x == 255) carry = before + incl;
            __syncthreads();
        }
        if (threadIdx.x == 0) n_cp[d] = carry;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(128) k_chunk_walk(const uint8_t* __restrict__ utf8, const int64_t* __restrict__ off, int64_t n_docs,
                                                     const uint32_t* __restrict__ cp_byte, const int32_t* __restrict__ n_cp_arr,
                                                     int32_t chunk_size, int32_t overlap_cfg, const int64_t* __restrict__ cap_off,
                                                     int32_t* __restrict__ out_count, int32_t* __restrict__ o_start, int32_t* __restrict__ o_end,
                                                     int32_t* __restrict__ o_tstart, int32_t* __restrict__ o_tend,
                                                     int64_t* __restrict__ o_tb0, int64_t* __restrict__ o_tb1) {
    const int lane = threadIdx.x & 31;
    const int64_t base0 = off[0];
    for (int64_t d = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5); d < n_docs; d += (int64_t)gridDim.x * 4) {
        const uint8_t* t = utf8 + off[d];
        const int64_t nb = off[d + 1] - off[d];
        const uint32_t* cb = cp_byte + (off[d] - base0);
        const int32_t n = n_cp_arr[d];
        const int64_t seg = cap_off[d], cap = cap_off[d + 1] - cap_off[d];
        const int32_t third = chunk_size / 3 > 0 ? chunk_size / 3 : 0;
        const int32_t overlap = overlap_cfg < third ? overlap_cfg : third;
        const int32_t half = chunk_size / 2 > 200 ? chunk_size / 2 : 200;
        int32_t start = 0, cnt = 0;
        while (start < n) {
            const int32_t hard_end = (int64_t)start + chunk_size < n ? start + chunk_size : n;
            int32_t end = hard_end;
            if (hard_end < n) {
                const int32_t floor_ = start + half;
                int32_t found = -1;
                for (int32_t hi = hard_end; hi > floor_ && found < 0; hi -= 32) {          // rfind("\n", floor_, hard_end)
                    const int32_t i = hi - 1 - lane;
                    const bool nl = i >= floor_ && t[cb[i]] == (uint8_t)'\n';
                    const uint32_t m = __ballot_sync(0xffffffffu, nl);
                    if (m) found = hi - 1 - (__ffs(m) - 1);
                }
                if (found > start) end = found + 1 < n ? found + 1 : n;
            }
            // strip(): first / last non-whitespace code point of [start, end)
            int32_t ts = end, te = start;
            for (int32_t lo = start; lo < end && ts == end; lo += 32) {
                const int32_t i = lo + lane;
                const bool nw = i < end && !dv_is_space_cp(dv_decode_at(t, cb[i], nb));
                const uint32_t m = __ballot_sync(0xffffffffu, nw);
                if (m) ts = lo + __ffs(m) - 1;
            }
            if (ts < end) {
                for (int32_t hi = end; hi > ts; hi -= 32) {
                    const int32_t i = hi - 1 - lane;
                    const bool nw = i >= ts && !dv_is_space_cp(dv_decode_at(t, cb[i], nb));
                    const uint32_t m = __ballot_sync(0xffffffffu, nw);
                    if (m) { te = hi - (__ffs(m) - 1); break; }
                }
                if (lane == 0 && cnt < cap) {
                    o_start[seg + cnt] = start; o_end[seg + cnt] = end; o_tstart[seg + cnt] = ts; o_tend[seg + cnt] = te;
                    o_tb0[seg + cnt] = cb[ts]; o_tb1[seg + cnt] = te < n ? (int64_t)cb[te] : nb;
                }
                cnt++;
            }
            if (end >= n) break;
            int32_t next = end - overlap > start + 1 ? end - overlap : start + 1;
            if (next <= start) next = end;
            start = next;
        }
        if (lane == 0) out_count[d] = cnt;
    }
}

extern "C" int32_t dvb_chunk_build(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_docs,
                                    int32_t chunk_size, int32_t overlap, const int64_t* cap_off,
                                    int32_t* out_count, int32_t* out_start, int32_t* out_end, int32_t* out_tstart, int32_t* out_tend,
                                    int64_t* out_tbyte0, int64_t* out_tbyte1) {
    if (!ctx || n_docs < 0 || !off || !cap_off || chunk_size < 1 || overlap < 0 || !out_count || !out_start || !out_end || !out_tstart || !out_tend ||
        !out_tbyte0 || !out_tbyte1) { dvb_set_err(ctx, "dvb_chunk_build: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_docs == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    int64_t ends[3];
    if (mem == DVB_MEM_HOST) { ends[0] = off[0]; ends[1] = off[n_docs]; ends[2] = cap_off[n_docs]; }
    else {
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[0], off, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[1], off + n_docs, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[2], cap_off + n_docs, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    }
    const int64_t total = ends[1] - ends[0], cap = ends[2];
    if (total < 0 || total > 0xFFFFFFF0LL || cap < 0) { dvb_set_err(ctx, "dvb_chunk_build: bad offsets (documents are limited to 4 GiB per call)%s%s"); return DVB_E_INVALID; }
    const uint8_t* d_utf8 = mem == DVB_MEM_HOST ? io.in(utf8 + ends[0], (size_t)total + 16) - ends[0] : utf8;
    const int64_t* d_off = io.in(off, (size_t)n_docs + 1);
    const int64_t* d_cap = io.in(cap_off, (size_t)n_docs + 1);
    int32_t* d_count = io.out(out_count, (size_t)n_docs);
    int32_t* d_s = io.out(out_start, (size_t)cap); int32_t* d_e = io.out(out_end, (size_t)cap);
    int32_t* d_ts = io.out(out_tstart, (size_t)cap); int32_t* d_te = io.out(out_tend, (size_t)cap);
    int64_t* d_b0 = io.out(out_tbyte0, (size_t)cap); int64_t* d_b1 = io.out(out_tbyte1, (size_t)cap);
    uint32_t* d_cb = io.tmp<uint32_t>((size_t)total + 1);
    int32_t* d_ncp = io.tmp<int32_t>((size_t)n_docs);
    int32_t rc = io.check(ctx); if (rc) return rc;
    const int g1 = (int)std::min<int64_t>(n_docs, (int64_t)ctx->num_sm * 8);
    k_cp_index<<<g1, 256, 0, io.st>>>(d_utf8, d_off, n_docs, d_cb, d_ncp);
    const int g2 = (int)std::min<int64_t>((n_docs + 3) / 4, (int64_t)ctx->num_sm * 8);
    k_chunk_walk<<<g2, 128, 0, io.st>>>(d_utf8, d_off, n_docs, d_cb, d_ncp, chunk_size, overlap, d_cap, d_count, d_s, d_e, d_ts, d_te, d_b0, d_b1);
    ctx->launches += 2;
    return io.finish(ctx);
}
