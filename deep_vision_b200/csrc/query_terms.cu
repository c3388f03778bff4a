// query_terms.cu — dvb_query_terms: the term extractor of the reference-material query path on the device.
// replaces: _tokenize_reference_material_query(query_text)   (reference web/server.py:32037-32051)
//   lower(); findall [a-z0-9_]{2,} | [一-鿿]{2,}; ordered de-dup; a pure-CJK term of >= 6 characters also contributes its
//   4-grams, then its 3-grams (de-dup against everything so far); the first 80 terms.
// Output is the symbol CSR dvb_chunk_index_topk / dvb_chunk_scores read (include/dvb200.h), so a batch of query texts goes
// from UTF-8 to top-k without the interpreter in between.
//   K1  k_symbolize (chunk.cu): decode, CPython lower(), class map -> u16 symbols, 0 = separator.  A term is a maximal run
//       of non-zero symbols of ONE class (the regex alternation cannot join classes) of length >= 2.
//   K2  k_qt_select: one warp per query walks the runs in order and keeps a term list of (start, length) references into
//       the query's own symbols — every term is a substring of it; "not in terms" is a warp-parallel comparison of the
//       candidate with the <= 80 kept terms.
//   K3  scans of the per-query term / symbol counts, then k_qt_write copies the substrings into the CSR.
#include "common.cuh"

#define QT_MAX_TERMS 80

// candidate [cs, cs + cl) of the query's symbols equals a kept term?  (warp-parallel over the kept terms)
__device__ __forceinline__ bool qt_seen(const uint16_t* __restrict__ s, const uint32_t* terms, int nt, int cs, int cl, int lane) {
    bool dup = false;
    for (int t = lane; t < nt; t += 32) {
        const uint32_t e = terms[t];
        const int ts = (int)(e & 0xFFFFu), tl = (int)(e >> 16);
        if (tl == cl) {
            bool same = true;
            for (int j = 0; same && j < cl; j++) same = s[ts + j] == s[cs + j];
            dup = dup || same;
        }
    }
    return __any_sync(0xffffffffu, dup);
}

// terms[q][0..80): start | length << 16; cnt2[q] = (n_terms, n_symbols)
__global__ void __launch_bounds__(256) k_qt_select(const uint16_t* __restrict__ sym, const int64_t* __restrict__ off, const int32_t* __restrict__ sym_cnt,
                                                    int64_t nq, int32_t max_terms, uint32_t* __restrict__ terms, int32_t* __restrict__ n_terms,
                                                    int32_t* __restrict__ n_syms, int32_t* __restrict__ err) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t base0 = off[0];
    for (int64_t q = (int64_t)blockIdx.x * 8 + warp; q < nq; q += (int64_t)gridDim.x * 8) {
        const uint16_t* s = sym + (off[q] - base0);
        const int m = sym_cnt[q];
        uint32_t* tq = terms + q * QT_MAX_TERMS;
        int nt = 0, ns = 0;
        if (m > 65535) { if (lane == 0) atomicExch(err, 1); }
        else {
            int i = 0;
            while (i < m && nt < max_terms) {
                // next run: 32 symbols per step, first symbol of a class, then the end of the run
                int start = -1, cls = 0;
                while (i < m) {
                    const int p = i + lane;
                    const uint32_t v = p < m ? s[p] : 0u;
                    const uint32_t nz = __ballot_sync(0xffffffffu, v != 0u);
                    if (nz) { const int l0 = __ffs(nz) - 1; start = i + l0; cls = (int)(__shfl_sync(0xffffffffu, v, l0) >> 15); break; }
                    i += 32;
                }
                if (start < 0) break;
                int end = start;
                for (;;) {
                    const int p = end + lane;
                    const uint32_t v = p < m ? s[p] : 0u;
                    const uint32_t brk = __ballot_sync(0xffffffffu, v == 0u || (int)(v >> 15) != cls);
                    if (brk) { end += __ffs(brk) - 1; break; }
                    end += 32;
                }
                const int len = end - start;
                i = end;
                if (len < 2) continue;
                if (!qt_seen(s, tq, nt, start, len, lane)) {
                    if (lane == 0) tq[nt] = (uint32_t)start | ((uint32_t)len << 16);
                    __syncwarp();
                    nt++; ns += len;
                }
                if (cls == 0 && len >= 6)                                   // CJK (ASCII symbols carry bit 15)
                    for (int size = 4; size >= 3 && nt < max_terms; size--)
                        for (int k = 0; k + size <= len && nt < max_terms; k++)
                            if (!qt_seen(s, tq, nt, start + k, size, lane)) {
                                if (lane == 0) tq[nt] = (uint32_t)(start + k) | ((uint32_t)size << 16);
                                __syncwarp();
                                nt++; ns += size;
                            }
            }
        }
        if (lane == 0) { n_terms[q] = nt; n_syms[q] = ns; }
    }
}

// exclusive scan of int32 counts into int32 offsets (single CTA), total in off[n]
static __global__ void __launch_bounds__(1024) k_qt_scan(const int32_t* __restrict__ cnt, int32_t* __restrict__ off, int64_t n, int32_t* __restrict__ overflow) {
    __shared__ long long warp_tot[32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        const long long v = i < n ? cnt[i] : 0;
        long long x = v;
        for (int d = 1; d < 32; d <<= 1) { const long long y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        long long before = carry;
        for (int w = 0; w < warp; w++) before += warp_tot[w];
        if (i < n) off[i] = (int32_t)(before + x - v);
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) { off[n] = (int32_t)carry; if (carry > 0x7fffffffLL) *overflow = 1; }
}

__global__ void __launch_bounds__(256) k_qt_write(const uint16_t* __restrict__ sym, const int64_t* __restrict__ off, int64_t nq, const uint32_t* __restrict__ terms,
                                                   const int32_t* __restrict__ qto, const int32_t* __restrict__ qso, int64_t sym_cap,
                                                   uint16_t* __restrict__ out_sym, int32_t* __restrict__ out_term_off) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t base0 = off[0];
    for (int64_t q = (int64_t)blockIdx.x * 8 + warp; q < nq; q += (int64_t)gridDim.x * 8) {
        const uint16_t* s = sym + (off[q] - base0);
        const uint32_t* tq = terms + q * QT_MAX_TERMS;
        const int t0 = qto[q], nt = qto[q + 1] - t0;
        int w = qso[q];
        for (int t = 0; t < nt; t++) {
            const uint32_t e = tq[t];
            const int ts = (int)(e & 0xFFFFu), tl = (int)(e >> 16);
            if (lane == 0) out_term_off[t0 + t] = w;
            if ((int64_t)w + tl <= sym_cap) for (int j = lane; j < tl; j += 32) out_sym[w + j] = s[ts + j];
            w += tl;
        }
        if (q == nq - 1 && lane == 0) out_term_off[t0 + nt] = w;
    }
}

extern "C" int32_t dvb_query_terms(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t nq, int32_t max_terms,
                                    uint16_t* out_term_sym, int64_t sym_cap, int32_t* out_term_off, int32_t* out_query_term_off, int64_t* out_totals2) {
    if (!ctx || nq < 0 || !off || !out_term_sym || !out_term_off || !out_query_term_off || !out_totals2 || max_terms < 1 || max_terms > QT_MAX_TERMS || sym_cap < 0) {
        dvb_set_err(ctx, "dvb_query_terms: invalid argument (1 <= max_terms <= 80)%s%s"); return DVB_E_INVALID; }
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    if (nq == 0) {
        const int32_t zero = 0;
        out_totals2[0] = out_totals2[1] = 0;
        if (mem == DVB_MEM_HOST) { out_term_off[0] = 0; out_query_term_off[0] = 0; return DVB_OK; }
        DVB_CUDA(ctx, cudaMemcpyAsync(out_term_off, &zero, 4, cudaMemcpyHostToDevice, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(out_query_term_off, &zero, 4, cudaMemcpyHostToDevice, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
        return DVB_OK;
    }
    int64_t ends[2];
    if (mem == DVB_MEM_HOST) { ends[0] = off[0]; ends[1] = off[nq]; }
    else {
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[0], off, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(&ends[1], off + nq, 8, cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    }
    const int64_t total = ends[1] - ends[0];
    if (total < 0) { dvb_set_err(ctx, "dvb_query_terms: bad offsets%s%s"); return DVB_E_INVALID; }
    const uint8_t* d_utf8 = mem == DVB_MEM_HOST ? io.in(utf8 + ends[0], (size_t)total + 16) - ends[0] : utf8;
    const int64_t* d_off = io.in(off, (size_t)nq + 1);
    uint16_t* d_sym = io.tmp<uint16_t>((size_t)total + 8);
    int32_t* d_cnt = io.tmp<int32_t>((size_t)nq);
    uint32_t* d_terms = io.tmp<uint32_t>((size_t)nq * QT_MAX_TERMS);
    int32_t* d_nt = io.tmp<int32_t>((size_t)nq); int32_t* d_ns = io.tmp<int32_t>((size_t)nq);
    int32_t* d_qso = io.tmp<int32_t>((size_t)nq + 1);
    int32_t* d_err = io.tmp<int32_t>(2);
    uint16_t* d_osym = io.out(out_term_sym, (size_t)sym_cap);
    int32_t* d_otoff = io.out(out_term_off, (size_t)nq * (size_t)max_terms + 1);
    int32_t* d_oqto = io.out(out_query_term_off, (size_t)nq + 1);
    int32_t rc = io.check(ctx); if (rc) return rc;
    DVB_CUDA(ctx, cudaMemsetAsync(d_err, 0, 8, io.st));
    const bool timing = ctx->timing.load() != 0;
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev0, io.st));
    dv_launch_symbolize(ctx, io.st, d_utf8, d_off, nq, d_sym, d_cnt);
    const int grid = (int)std::min<int64_t>((nq + 7) / 8, (int64_t)ctx->num_sm * 8);
    k_qt_select<<<grid, 256, 0, io.st>>>(d_sym, d_off, d_cnt, nq, max_terms, d_terms, d_nt, d_ns, d_err);
    k_qt_scan<<<1, 1024, 0, io.st>>>(d_nt, d_oqto, nq, d_err + 1);
    k_qt_scan<<<1, 1024, 0, io.st>>>(d_ns, d_qso, nq, d_err + 1);
    k_qt_write<<<grid, 256, 0, io.st>>>(d_sym, d_off, nq, d_terms, d_oqto, d_qso, sym_cap, d_osym, d_otoff);
    ctx->launches += 4;
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev1, io.st));
    int32_t h_err[2] = { 0, 0 }, h_tot[2] = { 0, 0 };
    DVB_CUDA(ctx, cudaMemcpyAsync(h_err, d_err, 8, cudaMemcpyDeviceToHost, io.st));
    DVB_CUDA(ctx, cudaMemcpyAsync(&h_tot[0], d_oqto + nq, 4, cudaMemcpyDeviceToHost, io.st));
    DVB_CUDA(ctx, cudaMemcpyAsync(&h_tot[1], d_qso + nq, 4, cudaMemcpyDeviceToHost, io.st));
    DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    if (h_err[0]) { dvb_set_err(ctx, "dvb_query_terms: a query has more than 65535 characters%s%s"); return DVB_E_INVALID; }
    if (h_err[1]) { dvb_set_err(ctx, "dvb_query_terms: more than 2^31 term symbols in one batch%s%s"); return DVB_E_CAPACITY; }
    if ((int64_t)h_tot[1] > sym_cap) {
        char msg[96]; snprintf(msg, sizeof(msg), "%d term symbols, out_term_sym holds %lld", h_tot[1], (long long)sym_cap);
        dvb_set_err(ctx, "dvb_query_terms: %s%s", msg); return DVB_E_CAPACITY;
    }
    out_totals2[0] = h_tot[0]; out_totals2[1] = h_tot[1];
    rc = io.finish(ctx);
    if (rc == DVB_OK && timing) { float ms = 0.f; cudaEventElapsedTime(&ms, sc.l->ev0, sc.l->ev1); ctx->last_ms = ms; }
    return rc;
}
