// annotate.cu — text features for the per-turn evidence annotation (rows a11-a13 of SURVEY.md §8).
// replaces the string scans of: evaluate_answer_depth (reference web/server.py:21221-21401), _score_rich_option_answer
//   (20318-20367), evaluate_answer_quality (20370-20426) and _infer_legacy_log_capture_contract (21419-21466):
//   ~100 literal substring tests (`kw in text`), exact matches (`text in [...]`), `any(c.isdigit())`, `len(text.strip())`
//   per question / answer / option, and `option in answer` containment.
//
// dvb_text_features  one warp per text: Python str.strip() bounds, stripped length in code points, any str.isdigit()
//                    character, and a 64-bit mask with bit g set iff some pattern of group g occurs in the text
//                    (substring), or — for `whole` patterns — equals the stripped text.  Matching runs on UTF-8 bytes
//                    (UTF-8 is self-synchronising, so byte-substring == code-point-substring for valid UTF-8); pattern
//                    candidates are found through a hash of the FIRST CHARACTER of each pattern, so a text position
//                    costs one table probe unless a pattern starts with that character.
// dvb_contains       one warp per (needle, haystack): raw needle in stripped haystack, stripped needle in stripped
//                    haystack, stripped needle == stripped haystack.
// The pattern lists are the reference's literals; they cross the ABI as data (nothing is baked into the kernels).
#include "common.cuh"
#include "../../include/dv_annot_tables.h"

__constant__ uint32_t c_digit_lo[DV_DIGIT_RANGES_N], c_digit_hi[DV_DIGIT_RANGES_N];
__constant__ uint32_t c_space_lo[DV_SPACE_RANGES_N], c_space_hi[DV_SPACE_RANGES_N];

struct PatDev {
    const uint8_t* bytes; const int32_t* off; const uint8_t* group; const uint8_t* whole; const int32_t* order;
    const uint32_t* ht_key; const uint32_t* ht_range;   // ht_range = start << 16 | count, into `order`
    uint32_t ht_mask; int32_t np;
};

__device__ __forceinline__ int an_char_len(uint32_t b) { return b < 0x80u ? 1 : (b < 0xE0u ? 2 : (b < 0xF0u ? 3 : 4)); }
__device__ __forceinline__ uint32_t an_decode(const uint8_t* t, int64_t i, int64_t n, int& len) {
    const uint32_t b = t[i];
    len = an_char_len(b);
    uint32_t cp = len == 1 ? b : (len == 2 ? (b & 0x1Fu) : (len == 3 ? (b & 0x0Fu) : (b & 0x07u)));
    for (int k = 1; k < len && i + k < n; k++) cp = (cp << 6) | (t[i + k] & 0x3Fu);
    return cp;
}
__device__ __forceinline__ bool an_is_space(uint32_t cp) {
    if (cp > 0x3000u) return false;
    if (cp < 0x80u) return (cp - 9u) <= 4u || (cp - 28u) <= 4u;
#pragma unroll
    for (int r = 0; r < DV_SPACE_RANGES_N; r++) if (cp >= c_space_lo[r] && cp <= c_space_hi[r]) return true;
    return false;
}
__device__ __forceinline__ bool an_is_digit(uint32_t cp) {
    if (cp < 0x80u) return cp - '0' < 10u;
    if (cp - 0x3400u < 0x6C00u) return false;                       // CJK ideographs: no digits
    int lo = 0, hi = DV_DIGIT_RANGES_N - 1;
    while (lo <= hi) {
        const int m = (lo + hi) >> 1;
        if (cp < c_digit_lo[m]) hi = m - 1; else if (cp > c_digit_hi[m]) lo = m + 1; else return true;
    }
    return false;
}
// first character of a byte string packed little-endian into a u32 (the hash key of a pattern / text position)
__host__ __device__ __forceinline__ uint32_t an_first_key(const uint8_t* t, int64_t avail) {
    const uint32_t b = t[0];
    int len = b < 0x80u ? 1 : (b < 0xE0u ? 2 : (b < 0xF0u ? 3 : 4));
    if (len > avail) len = (int)avail;
    uint32_t k = 0;
    for (int i = 0; i < len; i++) k |= (uint32_t)t[i] << (8 * i);
    return k;
}
__host__ __device__ __forceinline__ uint32_t an_hash(uint32_t key) { uint32_t x = key * 0x9E3779B1u; return x ^ (x >> 15); }

// Python str.strip() bounds of text[0..n): [b0, b1) in bytes (lane 0 computes, result broadcast)
__device__ __forceinline__ void an_strip(const uint8_t* t, int64_t n, int64_t& b0, int64_t& b1) {
    int64_t s = 0, e = n;
    if ((threadIdx.x & 31) == 0) {
        while (s < n) { int len; const uint32_t cp = an_decode(t, s, n, len); if (!an_is_space(cp)) break; s += len; }
        while (e > s) {
            int64_t c = e - 1;
            while (c > s && (t[c] & 0xC0u) == 0x80u) c--;
            int len; const uint32_t cp = an_decode(t, c, n, len);
            if (!an_is_space(cp)) break;
            e = c;
        }
    }
    b0 = __shfl_sync(0xffffffffu, s, 0); b1 = __shfl_sync(0xffffffffu, e, 0);
}

__global__ void __launch_bounds__(256) k_text_features(const uint8_t* __restrict__ utf8, const int64_t* __restrict__ off, int64_t n_texts, PatDev P,
                                                        unsigned long long* __restrict__ out_mask, int32_t* __restrict__ out_len,
                                                        uint8_t* __restrict__ out_flags, int32_t* __restrict__ out_b0, int32_t* __restrict__ out_b1) {
    const int lane = threadIdx.x & 31;
    for (int64_t ti = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); ti < n_texts; ti += (int64_t)gridDim.x * 8) {
        const uint8_t* t = utf8 + off[ti];
        const int64_t n = off[ti + 1] - off[ti];
        int64_t b0, b1;
        an_strip(t, n, b0, b1);
        unsigned long long mask = 0; int cnt = 0; bool digit = false;
        for (int64_t i = b0 + lane; i < b1; i += 32) {
            const uint32_t b = t[i];
            if ((b & 0xC0u) == 0x80u) continue;                     // continuation byte: not a character start
            cnt++;
            if (b >= 0x80u || b - '0' < 10u) { int len; digit |= an_is_digit(an_decode(t, i, b1, len)); }
            if (P.np == 0) continue;
            const uint32_t key = an_first_key(t + i, b1 - i);
            uint32_t h = an_hash(key) & P.ht_mask;
            for (;;) {
                const uint32_t rg = __ldg(&P.ht_range[h]);
                if (rg == 0u) break;                                 // empty slot: no pattern starts with this character
                if (__ldg(&P.ht_key[h]) == key) {
                    const uint32_t st = rg >> 16, c = rg & 0xFFFFu;
                    for (uint32_t q = 0; q < c; q++) {
                        const int32_t p = __ldg(&P.order[st + q]);
                        const int32_t po = __ldg(&P.off[p]), pl = __ldg(&P.off[p + 1]) - po;
                        const bool whole = __ldg(&P.whole[p]) != 0;
                        if (i + pl > b1 || (whole && (i != b0 || pl != b1 - b0))) continue;
                        bool eq = true;
                        for (int32_t k = 0; k < pl && eq; k++) eq = t[i + k] == __ldg(&P.bytes[po + k]);
                        if (eq) mask |= 1ULL << __ldg(&P.group[p]);
                    }
                    break;
                }
                h = (h + 1) & P.ht_mask;
            }
        }
        for (int s = 16; s > 0; s >>= 1) {
            mask |= __shfl_xor_sync(0xffffffffu, mask, s);
            cnt += __shfl_xor_sync(0xffffffffu, cnt, s);
            digit |= __shfl_xor_sync(0xffffffffu, (int)digit, s) != 0;
        }
        if (lane == 0) {
            out_mask[ti] = mask; out_len[ti] = cnt; out_flags[ti] = (digit ? 1 : 0) | (b1 > b0 ? 2 : 0);
            if (out_b0) { out_b0[ti] = (int32_t)b0; out_b1[ti] = (int32_t)b1; }
        }
    }
}

// does needle[0..m) occur in hay[h0..h1)?  (lanes stride the start positions; warp-uniform result)
__device__ __forceinline__ bool an_find(const uint8_t* hay, int64_t h0, int64_t h1, const uint8_t* nd, int64_t m) {
    if (m == 0) return true;
    bool found = false;
    for (int64_t base = h0; base + m <= h1 && !found; base += 32) {
        const int64_t i = base + (threadIdx.x & 31);
        bool eq = i + m <= h1;
        for (int64_t k = 0; k < m && eq; k++) eq = hay[i + k] == nd[k];
        found = __any_sync(0xffffffffu, eq);
    }
    return found;
}

__global__ void __launch_bounds__(256) k_contains(const uint8_t* __restrict__ hay_utf8, const int64_t* __restrict__ hay_off,
                                                   const uint8_t* __restrict__ nd_utf8, const int64_t* __restrict__ nd_off,
                                                   const int32_t* __restrict__ nd_hay, int64_t nn, uint8_t* __restrict__ out) {
    for (int64_t j = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); j < nn; j += (int64_t)gridDim.x * 8) {
        const int64_t hi = nd_hay[j];
        const uint8_t* hay = hay_utf8 + hay_off[hi]; const int64_t hn = hay_off[hi + 1] - hay_off[hi];
        const uint8_t* nd = nd_utf8 + nd_off[j]; const int64_t m = nd_off[j + 1] - nd_off[j];
        int64_t h0, h1, s0, s1;
        an_strip(hay, hn, h0, h1);
        an_strip(nd, m, s0, s1);
        const bool raw_in = an_find(hay, h0, h1, nd, m);
        const bool same = s0 == 0 && s1 == m;
        const bool strip_in = s1 > s0 && (same ? raw_in : an_find(hay, h0, h1, nd + s0, s1 - s0));
        bool equal = strip_in && (s1 - s0) == (h1 - h0);             // contained and same length == equal
        if ((threadIdx.x & 31) == 0) out[j] = (raw_in ? 1 : 0) | (strip_in ? 2 : 0) | (equal ? 4 : 0);
    }
}

// ------------------------------------------------------------------ host side
static int32_t an_upload_tables(dvb_ctx* ctx) {
    static std::mutex mu; static uint64_t done = 0;
    std::lock_guard<std::mutex> g(mu);
    if (ctx->device < 64 && (done >> ctx->device) & 1ULL) return DVB_OK;
    DVB_CUDA(ctx, cudaMemcpyToSymbol(c_digit_lo, dv_digit_lo, sizeof(dv_digit_lo)));
    DVB_CUDA(ctx, cudaMemcpyToSymbol(c_digit_hi, dv_digit_hi, sizeof(dv_digit_hi)));
    DVB_CUDA(ctx, cudaMemcpyToSymbol(c_space_lo, dv_space_lo, sizeof(dv_space_lo)));
    DVB_CUDA(ctx, cudaMemcpyToSymbol(c_space_hi, dv_space_hi, sizeof(dv_space_hi)));
    if (ctx->device < 64) done |= 1ULL << ctx->device;
    return DVB_OK;
}

extern "C" int32_t dvb_text_features(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_texts,
                                      const uint8_t* pat_utf8, const int32_t* pat_off, const uint8_t* pat_group, const uint8_t* pat_whole, int32_t n_pat,
                                      uint64_t* out_mask, int32_t* out_len, uint8_t* out_flags, int32_t* out_b0, int32_t* out_b1) {
    if (!ctx || n_texts < 0 || !off || n_pat < 0 || n_pat > 4096 || (n_pat > 0 && (!pat_utf8 || !pat_off || !pat_group || !pat_whole)) ||
        !out_mask || !out_len || !out_flags || ((out_b0 == nullptr) != (out_b1 == nullptr))) {
        dvb_set_err(ctx, "dvb_text_features: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_texts == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    int32_t rc = an_upload_tables(ctx); if (rc) return rc;
    // ---- pattern tables (always HOST pointers: they are the caller's small literal lists)
    std::vector<int32_t> order((size_t)n_pat);
    std::vector<uint32_t> keys((size_t)n_pat);
    for (int32_t p = 0; p < n_pat; p++) {
        const int32_t pl = pat_off[p + 1] - pat_off[p];
        if (pl <= 0 || pl > 255 || pat_group[p] > 63) { dvb_set_err(ctx, "dvb_text_features: patterns must be 1..255 bytes, groups 0..63%s%s"); return DVB_E_INVALID; }
        order[p] = p; keys[p] = an_first_key(pat_utf8 + pat_off[p], pl);
    }
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return keys[a] < keys[b]; });
    uint32_t hsz = 64; while (hsz < (uint32_t)n_pat * 4u) hsz <<= 1;
    std::vector<uint32_t> ht_key(hsz, 0u), ht_range(hsz, 0u);
    for (int32_t i = 0; i < n_pat;) {
        int32_t j = i; while (j < n_pat && keys[order[j]] == keys[order[i]]) j++;
        uint32_t h = an_hash(keys[order[i]]) & (hsz - 1);
        while (ht_range[h] != 0u) h = (h + 1) & (hsz - 1);
        ht_key[h] = keys[order[i]]; ht_range[h] = ((uint32_t)i << 16) | (uint32_t)(j - i);
        i = j;
    }
    const size_t pbytes = n_pat ? (size_t)pat_off[n_pat] : 0;
    uint8_t* d_pb = io.tmp<uint8_t>(pbytes + 16); int32_t* d_po = io.tmp<int32_t>((size_t)n_pat + 1);
    uint8_t* d_pg = io.tmp<uint8_t>((size_t)n_pat + 1); uint8_t* d_pw = io.tmp<uint8_t>((size_t)n_pat + 1);
    int32_t* d_or = io.tmp<int32_t>((size_t)n_pat + 1); uint32_t* d_hk = io.tmp<uint32_t>(hsz); uint32_t* d_hr = io.tmp<uint32_t>(hsz);
    // ---- texts
    std::vector<int64_t> hoff;
    int64_t o0 = 0, total = 0;
    if (mem == DVB_MEM_HOST) { o0 = off[0]; total = off[n_texts] - o0; }
    const uint8_t* d_utf8 = mem == DVB_MEM_HOST ? io.in(utf8 + o0, (size_t)total + 16) - o0 : utf8;
    const int64_t* d_off = io.in(off, (size_t)n_texts + 1);
    unsigned long long* d_mask = (unsigned long long*)io.out(out_mask, (size_t)n_texts);
    int32_t* d_len = io.out(out_len, (size_t)n_texts); uint8_t* d_fl = io.out(out_flags, (size_t)n_texts);
    int32_t* d_b0 = io.out(out_b0, (size_t)n_texts); int32_t* d_b1 = io.out(out_b1, (size_t)n_texts);
    rc = io.check(ctx); if (rc) return rc;
    if (n_pat) {
        DVB_CUDA(ctx, cudaMemcpyAsync(d_pb, pat_utf8, pbytes, cudaMemcpyHostToDevice, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(d_po, pat_off, 4 * (size_t)(n_pat + 1), cudaMemcpyHostToDevice, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(d_pg, pat_group, (size_t)n_pat, cudaMemcpyHostToDevice, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(d_pw, pat_whole, (size_t)n_pat, cudaMemcpyHostToDevice, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(d_or, order.data(), 4 * (size_t)n_pat, cudaMemcpyHostToDevice, io.st));
    }
    DVB_CUDA(ctx, cudaMemcpyAsync(d_hk, ht_key.data(), 4 * (size_t)hsz, cudaMemcpyHostToDevice, io.st));
    DVB_CUDA(ctx, cudaMemcpyAsync(d_hr, ht_range.data(), 4 * (size_t)hsz, cudaMemcpyHostToDevice, io.st));
    PatDev P{ d_pb, d_po, d_pg, d_pw, d_or, d_hk, d_hr, hsz - 1, n_pat };
    const int grid = (int)std::min<int64_t>((n_texts + 7) / 8, (int64_t)ctx->num_sm * 8);
    const bool timing = ctx->timing.load() != 0;
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev0, io.st));
    k_text_features<<<grid, 256, 0, io.st>>>(d_utf8, d_off, n_texts, P, d_mask, d_len, d_fl, d_b0, d_b1);
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev1, io.st));
    ctx->launches += 1;
    rc = io.finish(ctx);                                  // synchronises: the host-side tables stay alive until here
    if (rc == DVB_OK && timing) { float ms = 0.f; cudaEventElapsedTime(&ms, sc.l->ev0, sc.l->ev1); ctx->last_ms = ms; }
    return rc;
}

extern "C" int32_t dvb_contains(dvb_ctx* ctx, int32_t mem, const uint8_t* hay_utf8, const int64_t* hay_off, int64_t n_hay,
                                 const uint8_t* needle_utf8, const int64_t* needle_off, const int32_t* needle_hay, int64_t n_needles,
                                 uint8_t* out_bits) {
    if (!ctx || n_hay < 0 || n_needles < 0 || !hay_off || !needle_off || !needle_hay || !out_bits) {
        dvb_set_err(ctx, "dvb_contains: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_needles == 0) return DVB_OK;
    if (n_hay == 0) { dvb_set_err(ctx, "dvb_contains: needles without haystacks%s%s"); return DVB_E_INVALID; }
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    int32_t rc = an_upload_tables(ctx); if (rc) return rc;
    int64_t h0 = 0, ht = 0, n0 = 0, nt = 0;
    if (mem == DVB_MEM_HOST) {
        h0 = hay_off[0]; ht = hay_off[n_hay] - h0; n0 = needle_off[0]; nt = needle_off[n_needles] - n0;
        for (int64_t j = 0; j < n_needles; j++) if (needle_hay[j] < 0 || needle_hay[j] >= n_hay) {
            dvb_set_err(ctx, "dvb_contains: needle_hay out of range%s%s"); return DVB_E_INVALID; }
    }
    const uint8_t* d_hu = mem == DVB_MEM_HOST ? io.in(hay_utf8 + h0, (size_t)ht + 16) - h0 : hay_utf8;
    const uint8_t* d_nu = mem == DVB_MEM_HOST ? io.in(needle_utf8 + n0, (size_t)nt + 16) - n0 : needle_utf8;
    const int64_t* d_ho = io.in(hay_off, (size_t)n_hay + 1); const int64_t* d_no = io.in(needle_off, (size_t)n_needles + 1);
    const int32_t* d_nh = io.in(needle_hay, (size_t)n_needles);
    uint8_t* d_out = io.out(out_bits, (size_t)n_needles);
    rc = io.check(ctx); if (rc) return rc;
    const int grid = (int)std::min<int64_t>((n_needles + 7) / 8, (int64_t)ctx->num_sm * 8);
    k_contains<<<grid, 256, 0, io.st>>>(d_hu, d_ho, d_nu, d_no, d_nh, n_needles, d_out);
    ctx->launches += 1;
    return io.finish(ctx);
}

// ================================================================== a11 / a12 decision stage
// replaces, per interview-log turn, the arithmetic and control flow of: _infer_legacy_log_capture_contract (reference
//   web/server.py:21419-21466), the contract resolution of backfill_session_interview_log_evidence_annotations (21484-21520),
//   _score_rich_option_answer (20318-20367), evaluate_answer_depth (21221-21401), evaluate_answer_quality (20370-20426) and
//   classify_answer_evidence_class_v3 (22552-22565) — over the feature bits dvb_text_features / dvb_contains produced.
// One thread per turn.  fp64 sums run in the reference's order (Python sum() left to right from 0), no FMA contraction;
// round(x, 2) is exact decimal rounding (half to even on the exact value of the double), as CPython's float.__round__.
// Pattern-group bit numbers are part of the ABI (include/dvb200.h, DVB_AN_G_*).
#define AN_G_VAGUE 0
#define AN_G_INFORMATIVE 1
#define AN_G_SEMICOLON 2
#define AN_G_CONTRA_L 3
#define AN_G_CONTRA_R 9
#define AN_G_SCENARIO 15
#define AN_G_GENERIC 16
#define AN_G_OPT_PUNCT 17
#define AN_G_CAUSE 18
#define AN_G_CONTEXT 19
#define AN_G_ROLE 20
#define AN_G_PROCESS 21
#define AN_G_UNITS 22
#define AN_G_Q_RICH 24
#define AN_G_Q_EVIDENCE 25

__device__ __forceinline__ bool an_bit(unsigned long long m, int g) { return ((m >> g) & 1ULL) != 0ULL; }

// Python round(x, 2) for finite |x| < 2^40
__device__ __forceinline__ double an_round2(double x) {
    double k = floor(__dmul_rn(x, 100.0));
    if (__fma_rn(x, 100.0, -k) < 0.0) k -= 1.0;                         // the product rounded up across an integer
    else if (__fma_rn(x, 100.0, -(k + 1.0)) >= 0.0) k += 1.0;
    const double e = __fma_rn(x, 100.0, -(k + 0.5));                    // sign of (exact x * 100) - (k + 1/2); zero only on an exact tie
    const bool odd = fmod(fabs(k), 2.0) != 0.0;
    if (e > 0.0 || (e == 0.0 && odd)) k += 1.0;
    const double r = __ddiv_rn(k, 100.0);
    return r == 0.0 ? copysign(0.0, x) : r;
}

struct AnTurnIn {
    const unsigned long long* q_mask; const unsigned long long* a_mask; const int32_t* a_len; const uint8_t* a_flags;
    const int32_t* opt_off; const unsigned long long* opt_mask; const int32_t* opt_len; const uint8_t* opt_flags; const uint8_t* opt_bits;
    const uint32_t* turn_flags; const int32_t* follow_up_round; const int32_t* formal_prior; const double* sensitivity;
};

__global__ void __launch_bounds__(256) k_annotate_turns(AnTurnIn in, int64_t n, uint32_t hard_mask, int32_t overwrite_contract,
                                                         uint32_t* __restrict__ out_bits, double* __restrict__ out_fscore, double* __restrict__ out_quality) {
    const double W[8] = { 0.4, 0.5, 0.8, 0.3, 0.1, 0.2, 0.2, 0.6 };              // _SIGNAL_WEIGHTS.get(s, 0.1) by signal id
    const double PEN[8] = { 0.2, 0.25, 0.3, 0.2, 0.05, 0.1, 0.1, 0.25 };          // negative_penalty.get(s, 0.05)
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long qm = in.q_mask[t], am = in.a_mask[t];
        const int alen = in.a_len[t];
        const bool adigit = (in.a_flags[t] & 1u) != 0u;
        const uint32_t tf = in.turn_flags[t];
        const bool options_truthy = tf & 1u, multi = tf & 2u, is_fu = tf & 4u, log_hard = tf & 8u, log_rr = tf & 16u,
                   dim_exact = tf & 32u, dim_lower = tf & 64u, fu_contract = tf & (1u << 13);
        const uint32_t raw_mode = (tf >> 8) & 3u, raw_intent = (tf >> 10) & 3u;
        const int o0 = in.opt_off[t], o1 = in.opt_off[t + 1];
        // ---- inferred contract (21419-21466; build_question_capture_contract 19669 with no missing aspects)
        uint32_t imode, iintent; bool irr;
        if (fu_contract || log_hard) { imode = 2; irr = true; iintent = 3; }
        else if (in.formal_prior[t] >= 2 || dim_lower) { imode = 2; irr = true; iintent = 2; }
        else { imode = 1; irr = false; iintent = 1; }
        const bool asks = an_bit(qm, AN_G_Q_EVIDENCE);
        const unsigned long long marker = (1ULL << AN_G_CAUSE) | (1ULL << AN_G_CONTEXT) | (1ULL << AN_G_ROLE) | (1ULL << AN_G_PROCESS);
        bool opts_rich = false;
        for (int j = o0; j < o1; j++) opts_rich = opts_rich || (in.opt_len[j] >= 10 && (in.opt_mask[j] & marker) != 0ULL);
        if (asks || opts_rich) {
            imode = 2; irr = true;
            if (fu_contract || asks) iintent = 3; else if (iintent == 1) iintent = 2;
        }
        // ---- contract of the log (21484-21520)
        const bool explicit_c = raw_mode != 0u && raw_intent != 0u;
        const bool keep = explicit_c && !overwrite_contract;
        const uint32_t mode = keep ? raw_mode : imode;
        bool rr = log_rr || mode == 2u;
        if (overwrite_contract || !explicit_c) rr = irr || mode == 2u;
        const uint32_t intent = keep ? raw_intent : iintent;
        // ---- rich option (20318-20367)
        bool rich = false, exact = false;
        {
            bool any_cand = false, any_match = false;
            int best_len = -1; unsigned long long best_mask = 0; bool best_digit = false;
            for (int j = o0; j < o1; j++) {
                if (in.opt_len[j] <= 0) continue;
                any_cand = true;
                const uint32_t b = in.opt_bits[j];
                if (b & 4u) exact = true;
                if ((b & 2u) && in.opt_len[j] > best_len) {                  // max(matched, key=len): the first of the longest
                    any_match = true; best_len = in.opt_len[j]; best_mask = in.opt_mask[j]; best_digit = (in.opt_flags[j] & 1u) != 0u;
                }
            }
            if (alen == 0 || !any_cand) exact = false;
            else if (exact || any_match) {
                int plen; unsigned long long pmask; bool pdigit;
                if (exact) { plen = alen; pmask = am; pdigit = adigit; } else { plen = best_len; pmask = best_mask; pdigit = best_digit; }
                int score = (exact ? 1 : 0) + (plen >= 10 ? 1 : 0) + (plen >= 16 ? 1 : 0)
                          + (an_bit(pmask, AN_G_OPT_PUNCT) ? 1 : 0) + (an_bit(pmask, AN_G_CAUSE) ? 1 : 0) + (an_bit(pmask, AN_G_CONTEXT) ? 1 : 0)
                          + (an_bit(pmask, AN_G_ROLE) ? 1 : 0) + (an_bit(pmask, AN_G_PROCESS) ? 1 : 0) + (pdigit ? 1 : 0) + (an_bit(qm, AN_G_Q_RICH) ? 1 : 0);
                if (dim_lower && an_bit(pmask, AN_G_UNITS)) score += 1;
                rich = score >= 4 || (exact && score >= 3 && plen >= 12);
            }
            if (alen == 0 || !any_cand || !(exact || any_match)) { rich = false; exact = false; }
        }
        // ---- depth (21221-21401)
        const double sens = in.sensitivity[t];
        const bool needs_rationale = rr || mode == 2u;
        const bool pick_only = mode == 1u && !needs_rationale;
        int short_thr = (int)__dadd_rn(20.0, __dmul_rn(sens, 20.0));
        if (options_truthy) short_thr = min(short_thr, pick_only ? (multi ? 10 : 6) : (multi ? 18 : 12));
        uint32_t sig = 0, err = 0;
        if (alen < short_thr && !rich) sig |= 1u << 0;
        if (an_bit(am, AN_G_VAGUE)) sig |= 1u << 1;
        if (an_bit(am, AN_G_GENERIC)) sig |= 1u << 2;
        if (options_truthy) {
            const bool informative = alen >= 10 || rich || an_bit(am, AN_G_INFORMATIVE);
            if (!pick_only && exact && alen < 16 && !informative) sig |= 1u << 3;
            if (rich) sig |= 1u << 4;
        }
        if (dim_exact && !adigit && alen < 60) sig |= 1u << 5;
        if (multi && options_truthy && !an_bit(am, AN_G_SEMICOLON) && (o1 - o0) >= 3 && !rich) {
            int selected = 0;
            for (int j = o0; j < o1; j++) {
                if (!(in.opt_flags[j] & 2u)) err = 1u;                           // `opt in answer` with a non-str option raises in the reference
                selected += (in.opt_bits[j] & 1u) ? 1 : 0;
            }
            if (selected <= 1 && alen < 30) sig |= 1u << 6;
        }
        bool contra = false;
#pragma unroll
        for (int i = 0; i < 6; i++) contra = contra || (an_bit(am, AN_G_CONTRA_L + i) && an_bit(am, AN_G_CONTRA_R + i));
        if (contra) sig |= 1u << 7;
        const bool s_detailed = alen > 80, s_multi = an_bit(am, AN_G_SEMICOLON) && alen > 40, s_quant = adigit && alen > 30;
        double fs = 0.0;
#pragma unroll
        for (int i = 0; i < 8; i++) if (sig & (1u << i)) fs = __dadd_rn(fs, W[i]);
        fs = __dmul_rn(fs, sens);
        double suff = 0.0;
        if (s_detailed) suff = __dadd_rn(suff, 0.5);
        if (s_multi) suff = __dadd_rn(suff, 0.3);
        if (s_quant) suff = __dadd_rn(suff, 0.2);
        if (rich) suff = __dadd_rn(suff, intent == 3u ? 0.25 : 0.18);
        fs = __dadd_rn(fs, -suff);
        const bool any_suff = s_detailed || s_multi || s_quant || rich;
        const bool hard = (sig & hard_mask) != 0u;
        const bool needs = fs >= 0.4, suggest = !needs && fs >= 0.15 && !any_suff;
        // ---- quality (20370-20426)
        double q = 0.5;
        if (alen >= 120) q = __dadd_rn(q, 0.2); else if (alen >= 80) q = __dadd_rn(q, 0.1);
        if (adigit) q = __dadd_rn(q, 0.1);
        if (an_bit(am, AN_G_SCENARIO)) q = __dadd_rn(q, 0.1);
        if (sig & (1u << 4)) q = __dadd_rn(q, 0.12);
        double pen = 0.0;
#pragma unroll
        for (int i = 0; i < 8; i++) if (sig & (1u << i)) pen = __dadd_rn(pen, PEN[i]);
        q = __dadd_rn(q, -pen);
        if (is_fu && in.follow_up_round[t] >= 1 && !hard) q = __dadd_rn(q, 0.05);
        q = fmax(0.0, fmin(1.0, q));
        const bool high_quality = q >= 0.75;
        const double q2 = an_round2(q);
        // ---- evidence class (22552-22565) on the ROUNDED quality, clamped again
        const double qc = fmax(0.0, fmin(1.0, q2));
        uint32_t cls;
        if (sig & (1u << 4)) cls = 1;
        else if (sig & ((1u << 1) | (1u << 2) | (1u << 7))) cls = (intent >= 2u) ? 3 : 2;
        else if (sig & ((1u << 0) | (1u << 3) | (1u << 5) | (1u << 6))) cls = (intent == 3u) ? 3 : 2;
        else if (qc < 0.35 && intent >= 2u) cls = 2;
        else cls = 0;
        out_bits[t] = sig | (needs ? 1u << 8 : 0u) | (suggest ? 1u << 9 : 0u) | (hard ? 1u << 10 : 0u) | (adigit ? 1u << 11 : 0u) | (rr ? 1u << 12 : 0u)
                    | (mode << 13) | (intent << 15) | (cls << 17) | (err << 20) | (high_quality ? 1u << 21 : 0u) | (alen >= 80 ? 1u << 22 : 0u);
        out_fscore[t] = an_round2(fs);
        out_quality[t] = q2;
    }
}

extern "C" int32_t dvb_annotate_turns(dvb_ctx* ctx, int32_t mem, int64_t n_turns,
                                       const uint64_t* q_mask, const uint64_t* a_mask, const int32_t* a_len, const uint8_t* a_flags,
                                       const int32_t* opt_off, const uint64_t* opt_mask, const int32_t* opt_len, const uint8_t* opt_flags, const uint8_t* opt_bits,
                                       const uint32_t* turn_flags, const int32_t* follow_up_round, const int32_t* formal_prior, const double* sensitivity,
                                       uint32_t hard_signal_mask, int32_t overwrite_contract,
                                       uint32_t* out_bits, double* out_follow_up_score, double* out_quality_score) {
    if (!ctx || n_turns < 0 || !q_mask || !a_mask || !a_len || !a_flags || !opt_off || !turn_flags || !follow_up_round || !formal_prior || !sensitivity ||
        !out_bits || !out_follow_up_score || !out_quality_score) { dvb_set_err(ctx, "dvb_annotate_turns: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_turns == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    int32_t n_opt = 0;
    if (mem == DVB_MEM_HOST) n_opt = opt_off[n_turns];
    else { DVB_CUDA(ctx, cudaMemcpyAsync(&n_opt, opt_off + n_turns, 4, cudaMemcpyDeviceToHost, io.st)); DVB_CUDA(ctx, cudaStreamSynchronize(io.st)); }
    if (n_opt < 0 || (n_opt > 0 && (!opt_mask || !opt_len || !opt_flags || !opt_bits))) { dvb_set_err(ctx, "dvb_annotate_turns: bad option arrays%s%s"); return DVB_E_INVALID; }
    AnTurnIn in;
    in.q_mask = (const unsigned long long*)io.in(q_mask, (size_t)n_turns); in.a_mask = (const unsigned long long*)io.in(a_mask, (size_t)n_turns);
    in.a_len = io.in(a_len, (size_t)n_turns); in.a_flags = io.in(a_flags, (size_t)n_turns);
    in.opt_off = io.in(opt_off, (size_t)n_turns + 1);
    in.opt_mask = (const unsigned long long*)io.in(opt_mask, (size_t)n_opt); in.opt_len = io.in(opt_len, (size_t)n_opt);
    in.opt_flags = io.in(opt_flags, (size_t)n_opt); in.opt_bits = io.in(opt_bits, (size_t)n_opt);
    in.turn_flags = io.in(turn_flags, (size_t)n_turns); in.follow_up_round = io.in(follow_up_round, (size_t)n_turns);
    in.formal_prior = io.in(formal_prior, (size_t)n_turns); in.sensitivity = io.in(sensitivity, (size_t)n_turns);
    uint32_t* d_bits = io.out(out_bits, (size_t)n_turns); double* d_fs = io.out(out_follow_up_score, (size_t)n_turns); double* d_q = io.out(out_quality_score, (size_t)n_turns);
    int32_t rc = io.check(ctx); if (rc) return rc;
    const int grid = (int)std::min<int64_t>((n_turns + 255) / 256, (int64_t)ctx->num_sm * 8);
    k_annotate_turns<<<grid, 256, 0, io.st>>>(in, n_turns, hard_signal_mask, overwrite_contract, d_bits, d_fs, d_q);
    ctx->launches += 1;
    return io.finish(ctx);
}
