// tokenize.cu — dvb_tokenize_v3: K1 (decode + key generation) and K2 (per-text sort-unique).
// replaces: _tokenize_similarity_text_v3  (reference web/server.py:24722-24732)
#include "common.cuh"
#include "text.cuh"

// ------------------------------------------------------------------ K1: emit keys per text
struct EmitToGlobal {
    unsigned long long* out; int32_t cap; int32_t n; int lane;
    __device__ __forceinline__ void restart() { n = 0; }
    __device__ __forceinline__ void push(bool v, uint64_t key) {
        const uint32_t m = __ballot_sync(0xffffffffu, v);
        if (!m) return;
        const int r = __popc(m & ((1u << lane) - 1u));
        if (v && n + r < cap) out[n + r] = key;
        n += __popc(m);
    }
    __device__ __forceinline__ void bigram(bool v, uint32_t id) { push(v, (uint64_t)id); }
    __device__ __forceinline__ bool need_token(uint32_t) const { return true; }
    __device__ __forceinline__ void token(bool v, uint64_t key, uint32_t) { push(v, key); }
};

// one warp per text; raw keys (with duplicates) go to scratch[seg_off[t] .. ), count to raw_cnt[t]
__global__ void __launch_bounds__(256) k_emit_keys(const uint8_t* __restrict__ utf8, const int64_t* __restrict__ off, int64_t n_texts,
                                                    const int64_t* __restrict__ seg_off, unsigned long long* __restrict__ scratch,
                                                    int32_t* __restrict__ raw_cnt) {
    __shared__ dv_warp_buf wbuf[8];
    __shared__ unsigned long long ktab[DV_KTAB];
    dv_fill_ktab(ktab);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t t = (int64_t)blockIdx.x * 8 + warp; t < n_texts; t += (int64_t)gridDim.x * 8) {
        const int64_t b = off[t], e = off[t + 1];
        EmitToGlobal emit{ scratch + seg_off[t], (int32_t)(seg_off[t + 1] - seg_off[t]), 0, lane };
        dv_tokenize_v3_warp(utf8 + b, (int32_t)(e - b), &wbuf[warp], ktab, emit);
        if (lane == 0) raw_cnt[t] = emit.n < emit.cap ? emit.n : emit.cap;
        __syncwarp();
    }
}

// ------------------------------------------------------------------ K2: sort + unique per text
#define K2_SMEM_KEYS 4096

__device__ void bitonic_sort_u64(unsigned long long* a, int n_pow2) {   // whole CTA cooperates; a may be smem or global
    for (int k = 2; k <= n_pow2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
                int ixj = i ^ j;
                if (ixj > i) {
                    unsigned long long x = a[i], y = a[ixj];
                    bool up = ((i & k) == 0);
                    if ((x > y) == up) { a[i] = y; a[ixj] = x; }
                }
            }
            __syncthreads();
        }
    }
}

// CTA per text.  Sorts the raw keys (padded with ~0 to a power of two), removes duplicates in
// place, writes the distinct count.  Segments longer than K2_SMEM_KEYS are sorted in global memory
// (the scratch segment is padded to a power of two by the host-side layout).
__global__ void __launch_bounds__(256) k_sort_unique(const int64_t* __restrict__ seg_off, unsigned long long* __restrict__ scratch,
                                                      const int32_t* __restrict__ raw_cnt, int32_t* __restrict__ uniq_cnt, int64_t n_texts) {
    __shared__ unsigned long long skeys[K2_SMEM_KEYS];
    __shared__ int s_total;
    for (int64_t t = blockIdx.x; t < n_texts; t += gridDim.x) {
        const int m = raw_cnt[t];
        unsigned long long* seg = scratch + seg_off[t];
        if (m == 0) { if (threadIdx.x == 0) uniq_cnt[t] = 0; continue; }
        int p2 = 1; while (p2 < m) p2 <<= 1;
        unsigned long long* a;
        if (p2 <= K2_SMEM_KEYS) {
            a = skeys;
            for (int i = threadIdx.x; i < p2; i += blockDim.x) a[i] = i < m ? seg[i] : ~0ULL;
        } else {
            a = seg;                                                   // capacity is >= p2 by layout
            for (int i = m + threadIdx.x; i < p2; i += blockDim.x) a[i] = ~0ULL;
        }
        __syncthreads();
        bitonic_sort_u64(a, p2);
        // unique: sequential chunks with a running count (distinct keys stay ordered)
        if (threadIdx.x == 0) s_total = 0;
        __syncthreads();
        for (int base = 0; base < m; base += blockDim.x) {
            const int i = base + threadIdx.x;
            unsigned long long v = i < m ? a[i] : 0ULL;
            const bool head = (i < m) && (i == 0 || a[i - 1] != v);
            // block-wide exclusive scan of head flags via warp ballots
            __shared__ int warp_sums[8];
            const uint32_t bal = __ballot_sync(0xffffffffu, head);
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            if (lane == 0) warp_sums[warp] = __popc(bal);
            __syncthreads();
            int before = s_total;
            for (int w = 0; w < warp; w++) before += warp_sums[w];
            const int pos = before + __popc(bal & ((1u << lane) - 1u));
            __syncthreads();                                           // all reads of a[i-1] done before writes
            if (head) seg[pos] = v;                                    // pos <= i: never overwrites unread input when a==seg
            __syncthreads();
            if (threadIdx.x == 0) { int s = 0; for (int w = 0; w < 8; w++) s += warp_sums[w]; s_total += s; }
            __syncthreads();
        }
        if (threadIdx.x == 0) uniq_cnt[t] = s_total;
        __syncthreads();
    }
}

// exclusive scan of int32 counts into int64 offsets (single CTA; n up to a few million is fine)
__global__ void __launch_bounds__(1024) k_scan_counts(const int32_t* __restrict__ cnt, int64_t* __restrict__ off, int64_t n) {
    __shared__ long long warp_tot[32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int64_t base = 0; base < n; base += 1024) {
        const int64_t i = base + threadIdx.x;
        long long v = i < n ? cnt[i] : 0, x = v;
        for (int d = 1; d < 32; d <<= 1) { long long y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        long long before = carry;
        for (int w = 0; w < warp; w++) before += warp_tot[w];
        if (i < n) off[i] = before + x - v;
        __syncthreads();
        if (threadIdx.x == 1023) carry = before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[n] = carry;
}

__global__ void __launch_bounds__(256) k_gather_keys(const int64_t* __restrict__ seg_off, const unsigned long long* __restrict__ scratch,
                                                      const int64_t* __restrict__ out_off, unsigned long long* __restrict__ out,
                                                      int64_t out_cap, int64_t n_texts) {
    for (int64_t t = blockIdx.x; t < n_texts; t += gridDim.x) {
        const int64_t o = out_off[t], m = out_off[t + 1] - o;
        const unsigned long long* seg = scratch + seg_off[t];
        for (int64_t i = threadIdx.x; i < m; i += blockDim.x) if (o + i < out_cap) out[o + i] = seg[i];
    }
}

// ------------------------------------------------------------------ entry point
extern "C" int32_t dvb_tokenize_v3(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_texts,
                                    uint64_t* out_keys, int64_t key_cap, int64_t* out_off) {
    if (!ctx || n_texts < 0 || !off || !out_off || (key_cap > 0 && !out_keys)) { dvb_set_err(ctx, "dvb_tokenize_v3: invalid argument%s%s"); return DVB_E_INVALID; }
    dvb_scope sc(ctx);
    cudaStream_t st = sc.stream();
    if (n_texts == 0) { int64_t z = 0; if (mem == DVB_MEM_HOST) out_off[0] = 0; else DVB_CUDA(ctx, cudaMemcpyAsync(out_off, &z, 8, cudaMemcpyHostToDevice, st)); DVB_CUDA(ctx, cudaStreamSynchronize(st)); return DVB_OK; }
    // offsets are needed on the host to lay out the scratch segments
    std::vector<int64_t> hoff((size_t)n_texts + 1);
    if (mem == DVB_MEM_HOST) memcpy(hoff.data(), off, sizeof(int64_t) * (size_t)(n_texts + 1));
    else { DVB_CUDA(ctx, cudaMemcpyAsync(hoff.data(), off, sizeof(int64_t) * (size_t)(n_texts + 1), cudaMemcpyDeviceToHost, st)); DVB_CUDA(ctx, cudaStreamSynchronize(st)); }
    const int64_t total_bytes = hoff[n_texts] - hoff[0];
    std::vector<int64_t> hseg((size_t)n_texts + 1);
    hseg[0] = 0;
    for (int64_t t = 0; t < n_texts; t++) {
        int64_t nb = hoff[t + 1] - hoff[t];
        if (nb < 0 || nb > 0x7FFF0000LL) { dvb_set_err(ctx, "dvb_tokenize_v3: bad offsets%s%s"); return DVB_E_INVALID; }
        int64_t cap = nb / 2 + 1, p2 = 1; while (p2 < cap) p2 <<= 1;      // power of two for the in-place global sort
        hseg[t + 1] = hseg[t] + (cap <= K2_SMEM_KEYS ? cap : p2);
    }
    const int64_t scratch_n = hseg[n_texts];
    size_t need = dvb_pad((size_t)total_bytes) + dvb_pad(8 * (size_t)(n_texts + 1)) * 3 + dvb_pad(8 * (size_t)scratch_n)
                + dvb_pad(4 * (size_t)n_texts) * 2 + dvb_pad(8 * (size_t)(key_cap > 0 ? key_cap : 1)) + 4096;
    int32_t rc = sc.reserve(need); if (rc) return rc;
    cudaError_t cerr = cudaSuccess;
    uint8_t* d_utf8 = sc.alloc<uint8_t>((size_t)total_bytes + 64);
    int64_t* d_off = sc.alloc<int64_t>((size_t)n_texts + 1);
    int64_t* d_seg = sc.alloc<int64_t>((size_t)n_texts + 1);
    int64_t* d_ooff = sc.alloc<int64_t>((size_t)n_texts + 1);
    unsigned long long* d_scratch = sc.alloc<unsigned long long>((size_t)scratch_n + 1);
    int32_t* d_raw = sc.alloc<int32_t>((size_t)n_texts);
    int32_t* d_uniq = sc.alloc<int32_t>((size_t)n_texts);
    unsigned long long* d_out = sc.alloc<unsigned long long>((size_t)(key_cap > 0 ? key_cap : 1));
    if (!d_utf8 || !d_off || !d_seg || !d_ooff || !d_scratch || !d_raw || !d_uniq || !d_out) { dvb_set_err(ctx, "workspace layout error%s%s"); return DVB_E_NOMEM; }
    const uint8_t* k_utf8 = dvb_in(mem, utf8 + (mem == DVB_MEM_HOST ? hoff[0] : 0), d_utf8, (size_t)total_bytes, st, &cerr);
    const int64_t* k_off = off;
    if (mem == DVB_MEM_HOST) {
        // device copy of offsets rebased to the staged buffer
        std::vector<int64_t> reb((size_t)n_texts + 1);
        for (int64_t t = 0; t <= n_texts; t++) reb[t] = hoff[t] - hoff[0];
        DVB_CUDA(ctx, cudaMemcpyAsync(d_off, reb.data(), 8 * (size_t)(n_texts + 1), cudaMemcpyHostToDevice, st));
        DVB_CUDA(ctx, cudaStreamSynchronize(st));      // reb goes out of scope
        k_off = d_off;
    }
    DVB_CUDA(ctx, cudaMemcpyAsync(d_seg, hseg.data(), 8 * (size_t)(n_texts + 1), cudaMemcpyHostToDevice, st));
    DVB_CUDA(ctx, cerr);
    const int grid1 = (int)std::min<int64_t>((n_texts + 7) / 8, (int64_t)ctx->num_sm * 8);
    k_emit_keys<<<grid1, 256, 0, st>>>(k_utf8, k_off, n_texts, d_seg, d_scratch, d_raw);
    const int grid2 = (int)std::min<int64_t>(n_texts, (int64_t)ctx->num_sm * 8);
    k_sort_unique<<<grid2, 256, 0, st>>>(d_seg, d_scratch, d_raw, d_uniq, n_texts);
    int64_t* k_ooff = mem == DVB_MEM_DEVICE ? out_off : d_ooff;
    unsigned long long* k_out = mem == DVB_MEM_DEVICE ? (unsigned long long*)out_keys : d_out;
    k_scan_counts<<<1, 1024, 0, st>>>(d_uniq, k_ooff, n_texts);
    k_gather_keys<<<grid2, 256, 0, st>>>(d_seg, d_scratch, k_ooff, k_out, key_cap, n_texts);
    ctx->launches += 4;
    DVB_CUDA(ctx, cudaGetLastError());
    int64_t total = 0;
    DVB_CUDA(ctx, cudaMemcpyAsync(&total, k_ooff + n_texts, 8, cudaMemcpyDeviceToHost, st));
    if (mem == DVB_MEM_HOST) DVB_CUDA(ctx, cudaMemcpyAsync(out_off, d_ooff, 8 * (size_t)(n_texts + 1), cudaMemcpyDeviceToHost, st));
    DVB_CUDA(ctx, cudaStreamSynchronize(st));
    if (total > key_cap) { dvb_set_err(ctx, "dvb_tokenize_v3: key_cap too small%s%s"); return DVB_E_CAPACITY; }
    if (mem == DVB_MEM_HOST && total > 0) {
        DVB_CUDA(ctx, cudaMemcpyAsync(out_keys, d_out, 8 * (size_t)total, cudaMemcpyDeviceToHost, st));
        DVB_CUDA(ctx, cudaStreamSynchronize(st));
    }
    return DVB_OK;
}
