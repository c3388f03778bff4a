// gate.cu — question-similarity gate and solution-signature Jaccard.
// replaces: normalize_interview_question_text / question_token_set / is_similar_interview_question /
//           find_similar_interview_question_log      (reference web/server.py:20462-20501)
//           _solution_signature_ngrams / _solution_jaccard_similarity   (web/server.py:36798-36813)
//
//  k_question_keys  one thread per question: decode, strip, full CPython lower() (table from
//                   include/dv_unicode_tables.h, final-sigma rule included), the ordered synonym
//                   replaces, punctuation/whitespace removal, the stop-word replaces in the CALLER's
//                   order, then the sorted distinct bigram keys over ALL code points.
//  history index    (key, history id) pairs radix-sorted by key (sort.cuh): the postings of a bigram
//                   are one contiguous, id-ordered range.
//  k_gate_match     one CTA per candidate: postings of its bigrams are accumulated into packed 16-bit
//                   shared-memory counters (one per history question), then one pass over the counters
//                   evaluates  inter / max(1, |L|+|R|-inter) >= threshold  in fp64 exactly as the
//                   reference and keeps the LARGEST matching history index (reversed(logs) scan).
#include <type_traits>
#include <chrono>
#include "common.cuh"
#include "sort.cuh"
#include "../../include/dv_unicode_tables.h"

struct GateUnicode {            // device copies of the CPython tables
    const uint32_t* lower_src; const uint32_t* lower_dst;      // [N], [N][3]
    const uint32_t* cased_lo; const uint32_t* cased_hi; const uint32_t* cign_lo; const uint32_t* cign_hi;
    int n_lower, n_cased, n_cign;
};
struct GateTablesDev {          // synonym / stop-word tables as code-point CSR, caller order
    const uint32_t* syn_src; const int32_t* syn_src_off; const uint32_t* syn_dst; const int32_t* syn_dst_off; int32_t n_syn;
    const uint32_t* stop; const int32_t* stop_off; int32_t n_stop;
    int32_t growth;             // worst-case length factor of the synonym replaces (gate_tables)
};

__device__ __forceinline__ bool gate_in_ranges(uint32_t cp, const uint32_t* lo, const uint32_t* hi, int n) {
    int a = 0, b = n - 1;
    while (a <= b) { const int m = (a + b) >> 1; if (cp < lo[m]) b = m - 1; else if (cp > hi[m]) a = m + 1; else return true; }
    return false;
}
__device__ __forceinline__ bool gate_is_space(uint32_t c) {
    return (c - 9u < 5u) || (c - 0x1Cu < 5u) || c == 0x85u || c == 0xA0u || c == 0x1680u || (c - 0x2000u < 11u) ||
           c == 0x2028u || c == 0x2029u || c == 0x202Fu || c == 0x205Fu || c == 0x3000u;
}
__device__ __forceinline__ bool gate_is_punct(uint32_t c) {          // web/server.py:20466 character class besides \s
    switch (c) {
        case 0xFF0C: case 0x3002: case 0xFF1F: case 0xFF01: case 0x3001: case 0xFF1B: case 0xFF1A:
        case ',': case '.': case '?': case '!': case ';': case ':': case 0xFF08: case 0xFF09: case '(': case ')':
        case 0x3010: case 0x3011: case '[': case ']': case '"': case '\'': case 0x201C: case 0x201D: case 0x2018: case 0x2019:
            return true;
        default: return false;
    }
}
// CPython str.lower() of in[0..n) -> out; returns new length (<= n + #U+0130)
// Scratch views.  Questions whose capacity fits GQ_CAP code points and that stay inside the BMP (no 4-byte sequence in the
// text, no astral character in a synonym replacement) keep their two buffers in shared memory as 16-BIT code points, element i
// of thread t at half-word i * 128 + t; the others use the per-question regions of the global scratch (32-bit).  256 bytes per
// thread instead of 512: the kernel is bound by the latency of its serial per-thread loops, and the shared scratch is what caps
// the warps per SM (14 -> 28).  The helpers below are written once against operator[] / operator+.
#define GQ_THREADS 128
#define GATE_SLICES 4
#define GQ_CAP 64
#define GQ_SMEM_BYTES (2 * GQ_CAP * GQ_THREADS * 2)
struct GqShared {
    uint16_t* p;
    __device__ __forceinline__ uint16_t& operator[](int64_t i) const { return p[i * GQ_THREADS]; }
    __device__ __forceinline__ GqShared operator+(int64_t k) const { return GqShared{ p + k * GQ_THREADS }; }
};
struct GqGlobal {
    uint32_t* p;
    __device__ __forceinline__ uint32_t& operator[](int64_t i) const { return p[i]; }
    __device__ __forceinline__ GqGlobal operator+(int64_t k) const { return GqGlobal{ p + k }; }
};
__device__ __forceinline__ uint32_t gq_bit(uint32_t cp) { return (cp * 0x9E3779B1u) >> 26; }
template <class In, class Out>
__device__ int64_t gate_lower(const In in, int64_t n, const Out out, const GateUnicode& U) {
    int64_t k = 0;
    for (int64_t i = 0; i < n; i++) {
        const uint32_t cp = in[i];
        if (cp < 0x80u) { out[k++] = (cp - 'A' < 26u) ? cp + 32u : cp; continue; }
        if (cp >= 0x3000u && cp < 0xA000u) { out[k++] = cp; continue; }         // CJK block: nothing is cased
        if (cp == 0x3A3u) {                                                      // Final_Sigma context
            int64_t j; uint32_t c = 0;
            for (j = i - 1; j >= 0; j--) { c = in[j]; if (!gate_in_ranges(c, U.cign_lo, U.cign_hi, U.n_cign)) break; }
            bool fin = (j >= 0 && gate_in_ranges(c, U.cased_lo, U.cased_hi, U.n_cased));
            if (fin) {
                for (j = i + 1; j < n; j++) { c = in[j]; if (!gate_in_ranges(c, U.cign_lo, U.cign_hi, U.n_cign)) break; }
                fin = (j == n || !gate_in_ranges(c, U.cased_lo, U.cased_hi, U.n_cased));
            }
            out[k++] = fin ? 0x3C2u : 0x3C3u;
            continue;
        }
        int a = 0, b = U.n_lower - 1, hit = -1;
        while (a <= b) { const int m = (a + b) >> 1; const uint32_t s = U.lower_src[m]; if (cp < s) b = m - 1; else if (cp > s) a = m + 1; else { hit = m; break; } }
        if (hit < 0) { out[k++] = cp; continue; }
        for (int t = 0; t < 3; t++) { const uint32_t dcp = U.lower_dst[hit * 3 + t]; if (!dcp) break; out[k++] = dcp; }
    }
    return k;
}
// str.replace(src, dst): left-to-right, non-overlapping
template <class In, class Out>
__device__ int64_t gate_replace(const In in, int64_t n, const uint32_t* src, int32_t ns, const uint32_t* dst, int32_t nd, const Out out) {
    if (ns <= 0) { for (int64_t i = 0; i < n; i++) out[i] = in[i]; return n; }
    int64_t i = 0, k = 0;
    while (i < n) {
        bool m = (i + ns <= n) && in[i] == src[0];
        for (int32_t j = 1; m && j < ns; j++) m = in[i + j] == src[j];
        if (m) { for (int32_t j = 0; j < nd; j++) out[k++] = dst[j]; i += ns; }
        else out[k++] = in[i++];
    }
    return k;
}
// 64-bit key arrays: plain global memory, or split over the two shared-memory buffers of a question (high word | low word)
struct GqKeysGlobal {
    unsigned long long* p;
    __device__ __forceinline__ unsigned long long get(int64_t i) const { return p[i]; }
    __device__ __forceinline__ void set(int64_t i, unsigned long long v) const { p[i] = v; }
};
struct GqKeysShared {            // key = first code point << 21 | second: one 16-bit code point per buffer
    GqShared hi, lo;
    __device__ __forceinline__ unsigned long long get(int64_t i) const { return ((unsigned long long)hi[i] << 21) | lo[i]; }
    __device__ __forceinline__ void set(int64_t i, unsigned long long v) const { hi[i] = (uint16_t)(v >> 21); lo[i] = (uint16_t)(v & 0x1FFFFFu); }
};
template <class K>
__device__ void gate_heapsort_v(const K a, int64_t n) {
    for (int64_t start = n / 2 - 1; start >= 0; start--) {
        int64_t root = start;
        for (;;) {
            int64_t c = 2 * root + 1; if (c >= n) break;
            unsigned long long vc = a.get(c);
            if (c + 1 < n) { const unsigned long long v2 = a.get(c + 1); if (vc < v2) { c++; vc = v2; } }
            const unsigned long long vr = a.get(root);
            if (vr >= vc) break;
            a.set(root, vc); a.set(c, vr); root = c;
        }
    }
    for (int64_t end = n - 1; end > 0; end--) {
        const unsigned long long t0 = a.get(0); a.set(0, a.get(end)); a.set(end, t0);
        int64_t root = 0;
        for (;;) {
            int64_t c = 2 * root + 1; if (c >= end) break;
            unsigned long long vc = a.get(c);
            if (c + 1 < end) { const unsigned long long v2 = a.get(c + 1); if (vc < v2) { c++; vc = v2; } }
            const unsigned long long vr = a.get(root);
            if (vr >= vc) break;
            a.set(root, vc); a.set(c, vr); root = c;
        }
    }
}
__device__ void gate_heapsort(unsigned long long* a, int64_t n) { gate_heapsort_v(GqKeysGlobal{ a }, n); }

// one thread per question.  Scratch: two code-point buffers of cap_off[s+1]-cap_off[s] entries each;
// keys of question s are written at keys[cap_off[s] ...], their count to key_cnt[s]; the normalised text
// (optional) to norm[cap_off[s] ...] / norm_len[s].
template <class Buf>
__device__ __forceinline__ void gate_question(const uint8_t* t, int64_t nb, const Buf a, const Buf b, const GateUnicode& U, const GateTablesDev& T,
                                              unsigned long long* ks, int32_t* key_cnt_s, uint32_t* norm_s, int32_t* norm_len_s, bool /*keys_in_scratch: implied by Buf*/) {
    // decode
    int64_t m = 0;
    for (int64_t i = 0; i < nb;) {
        uint32_t c = t[i], cp; int len;
        if (c < 0x80u) { cp = c; len = 1; } else if (c < 0xE0u) { cp = c & 0x1Fu; len = 2; } else if (c < 0xF0u) { cp = c & 0x0Fu; len = 3; } else { cp = c & 0x07u; len = 4; }
        for (int j = 1; j < len && i + j < nb; j++) cp = (cp << 6) | (t[i + j] & 0x3Fu);
        i += len; a[m++] = cp;
    }
    // strip().lower()
    int64_t lo = 0, hi = m;
    while (lo < hi && gate_is_space(a[lo])) lo++;
    while (hi > lo && gate_is_space(a[hi - 1])) hi--;
    m = gate_lower(a + lo, hi - lo, b, U);
    Buf cur = b, oth = a;
    // `present`: a 64-bit set (one hashed bit per code point) that always covers the characters of the current text.  A
    // replace whose source has a character outside the set cannot match and is skipped without touching the buffers
    // (most of the 9 + 17 patterns never occur in a given question); replacements add their characters, deletions
    // leave the set a superset.
    unsigned long long present = 0ULL;
    for (int64_t i = 0; i < m; i++) present |= 1ULL << gq_bit(cur[i]);
    for (int32_t y = 0; y < T.n_syn; y++) {
        const uint32_t* src = T.syn_src + T.syn_src_off[y]; const int32_t ns = T.syn_src_off[y + 1] - T.syn_src_off[y];
        const uint32_t* dst = T.syn_dst + T.syn_dst_off[y]; const int32_t nd = T.syn_dst_off[y + 1] - T.syn_dst_off[y];
        unsigned long long need = 0ULL;
        for (int32_t j = 0; j < ns; j++) need |= 1ULL << gq_bit(src[j]);
        if (ns <= 0 || (present & need) != need) continue;                 // replace("", x) is not a rule the tables carry: copy = no-op
        m = gate_replace(cur, m, src, ns, dst, nd, oth);
        for (int32_t j = 0; j < nd; j++) present |= 1ULL << gq_bit(dst[j]);
        const Buf tmp = cur; cur = oth; oth = tmp;
    }
    { int64_t k = 0; for (int64_t i = 0; i < m; i++) { const uint32_t c = cur[i]; if (!gate_is_space(c) && !gate_is_punct(c)) cur[k++] = c; } m = k; }
    for (int32_t y = 0; y < T.n_stop; y++) {
        const uint32_t* src = T.stop + T.stop_off[y]; const int32_t ns = T.stop_off[y + 1] - T.stop_off[y];
        unsigned long long need = 0ULL;
        for (int32_t j = 0; j < ns; j++) need |= 1ULL << gq_bit(src[j]);
        if (ns <= 0 || (present & need) != need) continue;
        m = gate_replace(cur, m, src, ns, (const uint32_t*)nullptr, 0, oth);
        const Buf tmp = cur; cur = oth; oth = tmp;
    }
    if (norm_s) { for (int64_t i = 0; i < m; i++) norm_s[i] = cur[i]; *norm_len_s = (int32_t)m; }
    // bigram keys, sorted distinct
    int64_t nk = 0;
    if (m == 1) { ks[0] = (1ULL << 42) | cur[0]; nk = 1; }
    else if (m > 1) {
        nk = m - 1;
        if constexpr (std::is_same<Buf, GqShared>::value) {
            // key i = (cur[i] << 21) | cur[i + 1]: its first code point stays in cur[i], the second goes to oth[i]; sorted in place
            const GqKeysShared ksh{ GqShared{ &cur[0] }, GqShared{ &oth[0] } };
            for (int64_t i = 0; i + 1 < m; i++) oth[i] = cur[i + 1];
            gate_heapsort_v(ksh, nk);
            unsigned long long prev = ksh.get(0);
            ks[0] = prev;
            int64_t u = 1;
            for (int64_t i = 1; i < nk; i++) { const unsigned long long k = ksh.get(i); if (k != prev) { ks[u++] = k; prev = k; } }
            nk = u;
        } else {
            for (int64_t i = 0; i + 1 < m; i++) ks[i] = ((unsigned long long)cur[i] << 21) | cur[i + 1];
            gate_heapsort(ks, nk);
            int64_t u = 1; for (int64_t i = 1; i < nk; i++) if (ks[i] != ks[u - 1]) ks[u++] = ks[i];
            nk = u;
        }
    }
    *key_cnt_s = (int32_t)nk;
}

// one thread per question.  Scratch: two code-point buffers of cap_off[s+1]-cap_off[s] entries each (shared memory when
// that is <= GQ_CAP, else the question's region of bufA / bufB); keys of question s are written at keys[cap_off[s] ...],
// their count to key_cnt[s]; the normalised text (optional) to norm[cap_off[s] ...] / norm_len[s].
extern __shared__ uint16_t gq_smem[];
struct GqBatch {
    const uint8_t* utf8; const int64_t* off; int64_t n; const int64_t* cap_off; uint32_t* bufA; uint32_t* bufB;
    unsigned long long* keys; int32_t* key_cnt; uint32_t* norm; int32_t* norm_len;
};
// blocks [0, nblk) of a launch work through batch q, block `bid` of them
__device__ __forceinline__ void gq_run(const GqBatch& q, int64_t bid, int64_t nblk, const GateUnicode& U, const GateTablesDev& T) {
    const GqShared sa{ gq_smem + threadIdx.x }, sb{ gq_smem + GQ_CAP * GQ_THREADS + threadIdx.x };
    __shared__ int s_bmp;                                  // every synonym replacement writes BMP code points only
    if (threadIdx.x == 0) {
        int ok = 1;
        const int32_t nd = T.n_syn > 0 ? T.syn_dst_off[T.n_syn] : 0;
        for (int32_t j = 0; j < nd; j++) if (T.syn_dst[j] > 0xFFFFu) ok = 0;
        s_bmp = ok;
    }
    __syncthreads();
    const bool bmp_tables = s_bmp != 0;
    for (int64_t s = bid * blockDim.x + threadIdx.x; s < q.n; s += nblk * blockDim.x) {
        const uint8_t* t = q.utf8 + q.off[s];
        const int64_t nb = q.off[s + 1] - q.off[s];
        const int64_t c0 = q.cap_off[s];
        // code points (+1 for every 2-byte character: U+0130 lower-cases to two), times the synonym growth, bound the scratch use
        int64_t ncp = 2;
        // stepped exactly like the decoder of gate_question (a stray continuation byte starts a 2-byte sequence there), so raw bytes
        // that are not valid UTF-8 cannot produce more code points than counted here
        bool astral = false;                               // a 4-byte sequence: code points beyond 16 bits
        for (int64_t i = 0; i < nb && ncp <= GQ_CAP; ) { const uint32_t c = t[i]; const int len = c < 0x80u ? 1 : (c < 0xE0u ? 2 : (c < 0xF0u ? 3 : 4)); ncp += 1 + (len == 2); astral |= len == 4; i += len; }
        if (ncp * T.growth <= GQ_CAP && !astral && bmp_tables) gate_question(t, nb, sa, sb, U, T, q.keys + c0, q.key_cnt + s, q.norm ? q.norm + c0 : nullptr, q.norm ? q.norm_len + s : nullptr, true);
        else gate_question(t, nb, GqGlobal{ q.bufA + c0 }, GqGlobal{ q.bufB + c0 }, U, T, q.keys + c0, q.key_cnt + s, q.norm ? q.norm + c0 : nullptr, q.norm ? q.norm_len + s : nullptr, false);
    }
}
__global__ void __launch_bounds__(GQ_THREADS) k_question_keys(const uint8_t* __restrict__ utf8, const int64_t* __restrict__ off, int64_t n,
                                                        const int64_t* __restrict__ cap_off, uint32_t* __restrict__ bufA, uint32_t* __restrict__ bufB,
                                                        GateUnicode U, GateTablesDev T,
                                                        unsigned long long* __restrict__ keys, int32_t* __restrict__ key_cnt,
                                                        uint32_t* __restrict__ norm, int32_t* __restrict__ norm_len) {
    const GqBatch q{ utf8, off, n, cap_off, bufA, bufB, keys, key_cnt, norm, norm_len };
    gq_run(q, blockIdx.x, gridDim.x, U, T);
}
// history and candidates of one gate call in ONE launch (blocks [0, ga): batch a, the rest: batch b; separate scratch regions):
// at config-1 sizes each of the two launches is a latency chain of ~60 us, not work
__global__ void __launch_bounds__(GQ_THREADS) k_question_keys2(GqBatch a, int ga, GqBatch b, GateUnicode U, GateTablesDev T) {
    if ((int)blockIdx.x < ga) gq_run(a, blockIdx.x, ga, U, T);
    else gq_run(b, (int64_t)blockIdx.x - ga, (int64_t)gridDim.x - ga, U, T);
}

// (key, hid) pairs of the history questions, compacted at pair_off[h]
__global__ void k_gate_pairs(const unsigned long long* __restrict__ keys, const int64_t* __restrict__ cap_off, const int32_t* __restrict__ key_cnt,
                             const int64_t* __restrict__ pair_off, int64_t n_hist, unsigned long long* __restrict__ pk, uint32_t* __restrict__ pv) {
    for (int64_t h = blockIdx.x; h < n_hist; h += gridDim.x) {
        const int32_t c = key_cnt[h];
        for (int32_t i = threadIdx.x; i < c; i += blockDim.x) { pk[pair_off[h] + i] = keys[cap_off[h] + i]; pv[pair_off[h] + i] = (uint32_t)h; }
    }
}

#define GATE_THREADS 256
extern __shared__ uint32_t gate_acc[];
__global__ void __launch_bounds__(GATE_THREADS) k_gate_match(const unsigned long long* __restrict__ ckeys, const int64_t* __restrict__ ccap_off,
                                                              const int32_t* __restrict__ ccnt, int64_t n_cand,
                                                              const unsigned long long* __restrict__ pk, const uint32_t* __restrict__ pv, int64_t n_pairs,
                                                              const int32_t* __restrict__ hcnt, int64_t n_hist, int32_t tile_h, double thr,
                                                              int64_t* __restrict__ out, const int32_t* __restrict__ n_flagged, const int64_t* __restrict__ flagged) {
    __shared__ long long s_best;
    const int64_t n_items = n_flagged ? (int64_t)*n_flagged : n_cand;       // only what k_gate_match_prefix flagged, or everything
    if (n_items == 0) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int acc_words = (tile_h + 1) >> 1;
    for (int i = threadIdx.x; i < acc_words; i += GATE_THREADS) gate_acc[i] = 0u;
    // threshold <= 0: pairs with NO shared bigram also satisfy 0/union >= threshold (web/server.py:20490-20491), and they
    // never show up in the postings, so the answer is simply the last history entry with a non-empty token set.
    __shared__ long long s_last_nonempty;
    if (0.0 >= thr) {
        if (threadIdx.x == 0) s_last_nonempty = -1;
        __syncthreads();
        long long ln = -1;
        for (int64_t h = threadIdx.x; h < n_hist; h += GATE_THREADS) if (hcnt[h] > 0) ln = h;
        for (int s = 16; s > 0; s >>= 1) { const long long o = __shfl_xor_sync(0xffffffffu, ln, s); ln = o > ln ? o : ln; }
        if (lane == 0 && ln >= 0) atomicMax(&s_last_nonempty, ln);
    }
    __syncthreads();
    for (int64_t it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int64_t c = n_flagged ? flagged[it] : it;
        const int32_t L = ccnt[c];
        const unsigned long long* ck = ckeys + ccap_off[c];
        if (threadIdx.x == 0) s_best = -1;
        __syncthreads();
        if (L > 0 && 0.0 >= thr) {
            if (threadIdx.x == 0) s_best = s_last_nonempty;
        } else if (L > 0) {
            for (int64_t h0 = 0; h0 < n_hist; h0 += tile_h) {
                const int64_t h1 = min(n_hist, h0 + (int64_t)tile_h);
                for (int32_t j = warp; j < L; j += GATE_THREADS / 32) {
                    const unsigned long long key = ck[j];
                    int64_t a = 0, b = n_pairs;                                   // lower_bound of (key, h0)
                    while (a < b) { const int64_t m = (a + b) >> 1; const unsigned long long k = pk[m]; if (k < key || (k == key && (int64_t)pv[m] < h0)) a = m + 1; else b = m; }
                    for (int64_t p = a + lane; p < n_pairs; p += 32) {
                        if (pk[p] != key) break;
                        const int64_t h = pv[p];
                        if (h >= h1) break;
                        const uint32_t r = (uint32_t)(h - h0);
                        atomicAdd(&gate_acc[r >> 1], 1u << ((r & 1u) * 16));
                    }
                }
                __syncthreads();
                long long best = -1;
                for (int i = threadIdx.x; i < acc_words; i += GATE_THREADS) {
                    const uint32_t w = gate_acc[i];
                    if (w) {
                        gate_acc[i] = 0u;
#pragma unroll
                        for (int e = 0; e < 2; e++) {
                            const int32_t inter = (int32_t)((w >> (16 * e)) & 0xFFFFu);
                            const int64_t h = h0 + 2 * (int64_t)i + e;
                            if (inter > 0 && h < h1) {
                                int32_t un = L + hcnt[h] - inter; if (un < 1) un = 1;
                                if (__ddiv_rn((double)inter, (double)un) >= thr) best = h > best ? h : best;
                            }
                        }
                    }
                }
                for (int s = 16; s > 0; s >>= 1) { const long long o = __shfl_xor_sync(0xffffffffu, best, s); best = o > best ? o : best; }
                if (lane == 0 && best >= 0) atomicMax(&s_best, best);
                __syncthreads();
            }
        }
        if (threadIdx.x == 0) out[c] = s_best;
        __syncthreads();
    }
}


// ---- prefix-filtered match (threshold > 0): one WARP per candidate.
// J(L, R) >= t implies |L & R| >= t |L u R| >= t |L|, so R shares one of ANY |L| - ceil(t |L|) + 1 keys of L (pigeonhole):
// only the RAREST keys of the candidate (shortest posting lists) generate history ids, a size bound and "not above the
// best match so far" prune them, and the survivors are verified exactly (|L & R| by probing R's keys in a shared-memory
// set of L's keys, then the reference's fp64 test).  One key of slack keeps the bound safe against the rounding of the
// fp64 division.  Candidates with more than GP_MAXL keys are flagged (-2) for k_gate_match.
// Key -> posting range of the sorted (key, history id) pairs: an open-addressing table (empty = ~0) filled from the
// run heads / tails of the sorted pairs, so a candidate key costs one or two L2 probes instead of two 21-step binary
// searches over the pair array.
struct GateKeyTable { unsigned long long* key; uint32_t* lo; uint32_t* df; uint32_t mask; };
__device__ __forceinline__ uint32_t gkt_hash(unsigned long long key) { return (uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 32); }
__global__ void k_gate_table_heads(const unsigned long long* __restrict__ pk, int64_t n_pairs, GateKeyTable T) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = pk[i];
        if (i > 0 && pk[i - 1] == key) continue;
        uint32_t sl = gkt_hash(key) & T.mask;
        while (atomicCAS(&T.key[sl], ~0ULL, key) != ~0ULL) sl = (sl + 1) & T.mask;      // every head key is distinct
        T.lo[sl] = (uint32_t)i;
    }
}
__global__ void k_gate_table_tails(const unsigned long long* __restrict__ pk, int64_t n_pairs, GateKeyTable T) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_pairs; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = pk[i];
        if (i + 1 < n_pairs && pk[i + 1] == key) continue;
        uint32_t sl = gkt_hash(key) & T.mask;
        while (T.key[sl] != key) sl = (sl + 1) & T.mask;
        T.df[sl] = (uint32_t)(i + 1) - T.lo[sl];
    }
}

#define GP_MAXL 128
#ifndef GP_BATCH
#define GP_BATCH 8
#endif
#define GP_SET 256
struct GpWarp {
    unsigned long long key[GP_MAXL];
    unsigned long long set[GP_SET];
    uint32_t lo[GP_MAXL], df[GP_MAXL], cum[GP_MAXL + 1];
    uint16_t plist[GP_MAXL];
};
__global__ void __launch_bounds__(256) k_gate_match_prefix(const unsigned long long* __restrict__ ckeys, const int64_t* __restrict__ ccap_off,
                                                             const int32_t* __restrict__ ccnt, int64_t n_cand,
                                                             const unsigned long long* __restrict__ pk, const uint32_t* __restrict__ pv, int64_t n_pairs,
                                                             const unsigned long long* __restrict__ hkeys, const int64_t* __restrict__ hcap_off,
                                                             const int32_t* __restrict__ hcnt, double thr, int64_t* __restrict__ out,
                                                             int32_t* __restrict__ n_flagged, int64_t* __restrict__ flagged, GateKeyTable KT) {
    extern __shared__ __align__(16) unsigned char gp_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    GpWarp& W = reinterpret_cast<GpWarp*>(gp_raw)[warp];
    for (int64_t c = (int64_t)blockIdx.x * 8 + warp; c < n_cand; c += (int64_t)gridDim.x * 8) {
        const int32_t L = ccnt[c];
        if (L <= 0) { if (lane == 0) out[c] = -1; continue; }
        if (L > GP_MAXL) { if (lane == 0) { out[c] = -2; flagged[atomicAdd(n_flagged, 1)] = c; } continue; }
        const unsigned long long* ck = ckeys + ccap_off[c];
        __syncwarp();
        for (int i = lane; i < GP_SET; i += 32) W.set[i] = ~0ULL;
        __syncwarp();
        for (int j = lane; j < L; j += 32) {
            const unsigned long long key = ck[j];
            W.key[j] = key;
            uint32_t sl = (uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 56);
            while (atomicCAS(&W.set[sl], ~0ULL, key) != ~0ULL) sl = (sl + 1) & (GP_SET - 1);        // keys are distinct
            uint32_t ks = gkt_hash(key) & KT.mask, plo = 0u, pdf = 0u;          // posting range of the key (df = 0: not in the history)
            for (;;) {
                const unsigned long long tk = KT.key[ks];
                if (tk == key) { plo = KT.lo[ks]; pdf = KT.df[ks]; break; }
                if (tk == ~0ULL) break;
                ks = (ks + 1) & KT.mask;
            }
            W.lo[j] = plo; W.df[j] = pdf;
        }
        __syncwarp();
        int alpha = (int)ceil(thr * (double)L) - 1; if (alpha < 1) alpha = 1;
        int np_keys = L - alpha + 1; if (np_keys < 1) np_keys = 1; if (np_keys > L) np_keys = L;
        // the np_keys rarest keys: rank by (df, position)
        int n_pref = 0; uint32_t run = 0;
        for (int j0 = 0; j0 < L; j0 += 32) {
            const int j = j0 + lane;
            bool in = false; uint32_t d = 0;
            if (j < L) {
                d = W.df[j];
                int rank = 0;
                for (int i = 0; i < L; i++) { const uint32_t di = W.df[i]; rank += (di < d || (di == d && i < j)) ? 1 : 0; }
                in = rank < np_keys && d > 0;                      // a key without postings generates nothing
            }
            const uint32_t bal = __ballot_sync(0xffffffffu, in);
            uint32_t x = in ? d : 0u;                              // inclusive scan of the kept posting counts
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
            if (in) { const int slot = n_pref + __popc(bal & ((1u << lane) - 1u)); W.plist[slot] = (uint16_t)j; W.cum[slot + 1] = run + x; }
            run += __shfl_sync(0xffffffffu, x, 31);
            n_pref += __popc(bal);
        }
        if (lane == 0) W.cum[0] = 0;
        __syncwarp();
        const uint32_t total = run;
        long long best = -1;
        for (uint32_t base = 0; base < total; base += 32) {
            const uint32_t idx = base + lane;
            long long cand = -1;
            if (idx < total) {
                int a = 0, b = n_pref;                             // last m with cum[m] <= idx
                while (b - a > 1) { const int m = (a + b) >> 1; if (W.cum[m] <= idx) a = m; else b = m; }
                const int j = W.plist[a];
                const long long h = pv[(int64_t)W.lo[j] + W.df[j] - 1 - (idx - W.cum[a])];    // most recent history ids first
                if (h > best) {
                    const int32_t hc = hcnt[h];
                    const int32_t mn = hc < L ? hc : L;
                    int32_t un0 = L + hc - mn; if (un0 < 1) un0 = 1;
                    if (__ddiv_rn((double)mn, (double)un0) >= thr) {                          // size bound
                        const unsigned long long* hk = hkeys + hcap_off[h];
                        int32_t inter = 0;
                        // GP_BATCH keys per round: their loads are issued together (the one-key loop waited for every load in turn:
                        // long_scoreboard was 8 stall cycles per issue, profiles/README.md); a warp-cooperative verification of one
                        // survivor at a time was measured too and is slower (7.1 vs 5.3 ms: the survivors' load chains no longer overlap)
                        for (int32_t i = 0; i < hc; i += GP_BATCH) {
                            unsigned long long kq[GP_BATCH];
#pragma unroll
                            for (int u = 0; u < GP_BATCH; u++) kq[u] = i + u < hc ? hk[i + u] : 0ULL;
#pragma unroll
                            for (int u = 0; u < GP_BATCH; u++) {
                                if (i + u >= hc) break;
                                const unsigned long long key = kq[u];
                                uint32_t sl = (uint32_t)((key * 0x9E3779B97F4A7C15ULL) >> 56);
                                for (;;) { const unsigned long long v = W.set[sl]; if (v == key) { inter++; break; } if (v == ~0ULL) break; sl = (sl + 1) & (GP_SET - 1); }
                            }
                        }
                        int32_t un = L + hc - inter; if (un < 1) un = 1;
                        if (inter > 0 && __ddiv_rn((double)inter, (double)un) >= thr) cand = h;
                    }
                }
            }
            for (int o = 16; o > 0; o >>= 1) { const long long v = __shfl_xor_sync(0xffffffffu, cand, o); cand = v > cand ? v : cand; }
            best = cand > best ? cand : best;
        }
        if (lane == 0) out[c] = best;
    }
}

// exclusive scan of int32 counts into int64 offsets (single CTA)
static __global__ void __launch_bounds__(1024) k_gate_scan_counts(const int32_t* __restrict__ cnt, int64_t* __restrict__ off, int64_t n) {
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

// ------------------------------------------------------------------ host side
static int32_t gate_unicode(dvb_ctx* ctx, GateUnicode* U) {
    std::lock_guard<std::mutex> g(ctx->unicode_mu);
    if (!ctx->unicode_dev[0]) {
        std::vector<uint32_t> dst((size_t)DV_LOWER_N * 3);
        for (int i = 0; i < DV_LOWER_N; i++) for (int t = 0; t < 3; t++) dst[(size_t)i * 3 + t] = dv_lower_dst[i][t];
        const void* src[6] = { dv_lower_src, dst.data(), dv_cased_lo, dv_cased_hi, dv_cign_lo, dv_cign_hi };
        const size_t bytes[6] = { sizeof(dv_lower_src), dst.size() * 4, sizeof(dv_cased_lo), sizeof(dv_cased_hi), sizeof(dv_cign_lo), sizeof(dv_cign_hi) };
        void* dev[6] = { nullptr, nullptr, nullptr, nullptr, nullptr, nullptr };
        for (int i = 0; i < 6; i++) {
            if (cudaMalloc(&dev[i], bytes[i]) != cudaSuccess || cudaMemcpy(dev[i], src[i], bytes[i], cudaMemcpyHostToDevice) != cudaSuccess) {
                for (int j = 0; j <= i; j++) if (dev[j]) cudaFree(dev[j]);
                cudaGetLastError();
                return DVB_E_CUDA;
            }
        }
        for (int i = 5; i >= 0; i--) ctx->unicode_dev[i] = dev[i];
    }
    U->lower_src = (const uint32_t*)ctx->unicode_dev[0]; U->lower_dst = (const uint32_t*)ctx->unicode_dev[1];
    U->cased_lo = (const uint32_t*)ctx->unicode_dev[2]; U->cased_hi = (const uint32_t*)ctx->unicode_dev[3];
    U->cign_lo = (const uint32_t*)ctx->unicode_dev[4]; U->cign_hi = (const uint32_t*)ctx->unicode_dev[5];
    U->n_lower = DV_LOWER_N; U->n_cased = DV_CASED_RANGES_N; U->n_cign = DV_CIGN_RANGES_N;
    return DVB_OK;
}

struct GateHostTables {
    std::vector<int32_t> sso, sdo, sto;       // host copies of the offset arrays
    int64_t growth = 1;
};
// stages the tables and computes the worst-case growth factor of the synonym replaces
static int32_t gate_tables(dvb_ctx* ctx, dvb_io& io, const uint32_t* syn_src, const int32_t* syn_src_off, const uint32_t* syn_dst,
                           const int32_t* syn_dst_off, int32_t n_syn, const uint32_t* stop, const int32_t* stop_off, int32_t n_stop,
                           GateTablesDev* T, GateHostTables* H) {
    if (n_syn < 0 || n_stop < 0 || (n_syn > 0 && (!syn_src || !syn_src_off || !syn_dst || !syn_dst_off)) || (n_stop > 0 && (!stop || !stop_off))) {
        dvb_set_err(ctx, "gate tables: invalid argument%s%s"); return DVB_E_INVALID; }
    H->sso.assign((size_t)n_syn + 1, 0); H->sdo.assign((size_t)n_syn + 1, 0); H->sto.assign((size_t)n_stop + 1, 0);
    if (io.mem == DVB_MEM_HOST) {
        if (n_syn) { memcpy(H->sso.data(), syn_src_off, 4 * (size_t)(n_syn + 1)); memcpy(H->sdo.data(), syn_dst_off, 4 * (size_t)(n_syn + 1)); }
        if (n_stop) memcpy(H->sto.data(), stop_off, 4 * (size_t)(n_stop + 1));
    } else {
        if (n_syn) { DVB_CUDA(ctx, cudaMemcpyAsync(H->sso.data(), syn_src_off, 4 * (size_t)(n_syn + 1), cudaMemcpyDeviceToHost, io.st));
                     DVB_CUDA(ctx, cudaMemcpyAsync(H->sdo.data(), syn_dst_off, 4 * (size_t)(n_syn + 1), cudaMemcpyDeviceToHost, io.st)); }
        if (n_stop) DVB_CUDA(ctx, cudaMemcpyAsync(H->sto.data(), stop_off, 4 * (size_t)(n_stop + 1), cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    }
    H->growth = 1;
    for (int32_t y = 0; y < n_syn; y++) {
        const int32_t ns = H->sso[y + 1] - H->sso[y], nd = H->sdo[y + 1] - H->sdo[y];
        if (ns > 0 && nd > ns) { H->growth *= (nd + ns - 1) / ns; if (H->growth > 64) { dvb_set_err(ctx, "gate tables: synonym growth too large%s%s"); return DVB_E_INVALID; } }
    }
    T->n_syn = n_syn; T->n_stop = n_stop; T->growth = (int32_t)H->growth;
    T->syn_src = io.in(syn_src, (size_t)H->sso[n_syn]); T->syn_src_off = io.in(syn_src_off, (size_t)n_syn + 1);
    T->syn_dst = io.in(syn_dst, (size_t)H->sdo[n_syn]); T->syn_dst_off = io.in(syn_dst_off, (size_t)n_syn + 1);
    T->stop = io.in(stop, (size_t)H->sto[n_stop]); T->stop_off = io.in(stop_off, (size_t)n_stop + 1);
    return DVB_OK;
}

// capacity offsets (entries) per string: (bytes + 2) * growth
static int32_t gate_caps(dvb_ctx* ctx, dvb_io& io, const int64_t* off, int64_t n, int64_t growth, std::vector<int64_t>* hoff, std::vector<int64_t>* cap) {
    hoff->resize((size_t)n + 1);
    if (io.mem == DVB_MEM_HOST) memcpy(hoff->data(), off, 8 * (size_t)(n + 1));
    else { DVB_CUDA(ctx, cudaMemcpyAsync(hoff->data(), off, 8 * (size_t)(n + 1), cudaMemcpyDeviceToHost, io.st)); DVB_CUDA(ctx, cudaStreamSynchronize(io.st)); }
    cap->resize((size_t)n + 1);
    (*cap)[0] = 0;
    for (int64_t i = 0; i < n; i++) {
        const int64_t nb = (*hoff)[i + 1] - (*hoff)[i];
        if (nb < 0 || nb > 60000) { dvb_set_err(ctx, "question gate: a question is limited to 60000 bytes%s%s"); return DVB_E_INVALID; }
        (*cap)[i + 1] = (*cap)[i] + (nb + 2) * growth;
    }
    return DVB_OK;
}

extern "C" int32_t dvb_question_keys(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n,
                                      const uint32_t* syn_src, const int32_t* syn_src_off, const uint32_t* syn_dst, const int32_t* syn_dst_off, int32_t n_syn,
                                      const uint32_t* stop, const int32_t* stop_off, int32_t n_stop,
                                      int64_t* out_cap_off, uint64_t* out_keys, int32_t* out_key_cnt, uint32_t* out_norm, int32_t* out_norm_len, int64_t out_cap) {
    if (!ctx || n < 0 || !off || !out_cap_off || !out_keys || !out_key_cnt) { dvb_set_err(ctx, "dvb_question_keys: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    GateUnicode U; int32_t rc = gate_unicode(ctx, &U); if (rc) { dvb_set_err(ctx, "cannot upload unicode tables%s%s"); return rc; }
    GateTablesDev T; GateHostTables H;
    rc = gate_tables(ctx, io, syn_src, syn_src_off, syn_dst, syn_dst_off, n_syn, stop, stop_off, n_stop, &T, &H); if (rc) return rc;
    std::vector<int64_t> hoff, cap;
    rc = gate_caps(ctx, io, off, n, H.growth, &hoff, &cap); if (rc) return rc;
    if (cap[n] > out_cap) { dvb_set_err(ctx, "dvb_question_keys: out_cap too small%s%s"); return DVB_E_CAPACITY; }
    const int64_t total = hoff[n] - hoff[0];
    const uint8_t* d_utf8 = mem == DVB_MEM_HOST ? io.in(utf8 + hoff[0], (size_t)total + 16) - hoff[0] : utf8;
    const int64_t* d_off = io.in(off, (size_t)n + 1);
    int64_t* d_cap = io.tmp<int64_t>((size_t)n + 1);
    uint32_t* bufA = io.tmp<uint32_t>((size_t)cap[n] + 1); uint32_t* bufB = io.tmp<uint32_t>((size_t)cap[n] + 1);
    unsigned long long* d_keys = (unsigned long long*)io.out(out_keys, (size_t)cap[n]);
    int32_t* d_cnt = io.out(out_key_cnt, (size_t)n);
    uint32_t* d_norm = out_norm ? io.out(out_norm, (size_t)cap[n]) : nullptr;
    int32_t* d_nl = out_norm ? io.out(out_norm_len, (size_t)n) : nullptr;
    rc = io.check(ctx); if (rc) return rc;
    DVB_CUDA(ctx, cudaMemcpyAsync(d_cap, cap.data(), 8 * (size_t)(n + 1), cudaMemcpyHostToDevice, io.st));
    if (mem == DVB_MEM_HOST) memcpy(out_cap_off, cap.data(), 8 * (size_t)(n + 1));
    else DVB_CUDA(ctx, cudaMemcpyAsync(out_cap_off, cap.data(), 8 * (size_t)(n + 1), cudaMemcpyHostToDevice, io.st));
    DVB_CUDA(ctx, cudaFuncSetAttribute(k_question_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GQ_SMEM_BYTES));
    const int grid = (int)std::min<int64_t>((n + GQ_THREADS - 1) / GQ_THREADS, (int64_t)ctx->num_sm * 7);
    k_question_keys<<<grid, GQ_THREADS, GQ_SMEM_BYTES, io.st>>>(d_utf8, d_off, n, d_cap, bufA, bufB, U, T, d_keys, d_cnt, d_norm, d_nl);
    ctx->launches += 1;
    return io.finish(ctx);          // synchronises: `cap` stays alive until here
}

// Small histories (config 1: 60 questions, ~1 200 pairs): ONE CTA sorts the (key, history id) pairs in shared memory instead
// of the 24 launches of six radix passes, which cost more in launch latency than the whole rest of the call.  Bitonic network
// over the next power of two, ordered by (key, id): the same order the stable radix sort leaves (ids ascend inside a key).
#define GATE_SMALL_SORT 4096
__global__ void __launch_bounds__(1024) k_gate_sort_small(unsigned long long* __restrict__ pk, uint32_t* __restrict__ pv, int n) {
    extern __shared__ __align__(16) unsigned char gs_raw[];
    unsigned long long* key = reinterpret_cast<unsigned long long*>(gs_raw);
    uint32_t* val = reinterpret_cast<uint32_t*>(gs_raw + 8 * GATE_SMALL_SORT);
    int m = 1; while (m < n) m <<= 1;
    for (int i = threadIdx.x; i < m; i += blockDim.x) { key[i] = i < n ? pk[i] : ~0ULL; val[i] = i < n ? pv[i] : 0xFFFFFFFFu; }
    __syncthreads();
    for (int k = 2; k <= m; k <<= 1)
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < m; i += blockDim.x) {
                const int x = i ^ j;
                if (x > i) {
                    const unsigned long long ka = key[i], kb = key[x]; const uint32_t va = val[i], vb = val[x];
                    const bool gt = ka > kb || (ka == kb && va > vb);
                    if (gt == ((i & k) == 0)) { key[i] = kb; key[x] = ka; val[i] = vb; val[x] = va; }
                }
            }
            __syncthreads();
        }
    for (int i = threadIdx.x; i < n; i += blockDim.x) { pk[i] = key[i]; pv[i] = val[i]; }
}

// capacity offsets on the device: cap[i] = (bytes before string i + 2 i) * growth — a closed form of the running sum gate_caps
// builds on the host, so a 1 M-candidate call neither loops over the offsets on the host nor uploads 8 MB of capacities.
// err: a string longer than 60000 bytes or decreasing offsets.
__global__ void __launch_bounds__(256) k_gate_caps(const int64_t* __restrict__ off, int64_t n, int64_t growth, int64_t* __restrict__ cap, int32_t* __restrict__ err) {
    const int64_t o0 = off[0];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i <= n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t oi = off[i];
        cap[i] = (oi - o0 + 2 * i) * growth;
        if (i < n) { const int64_t nb = off[i + 1] - oi; if (nb < 0 || nb > 60000) atomicExch(err, 1); }
    }
}
// first and last offset of a CSR batch (host pointer or device pointer)
static int32_t gate_span(dvb_ctx* ctx, dvb_io& io, const int64_t* off, int64_t n, int64_t* first, int64_t* last) {
    if (io.mem == DVB_MEM_HOST) { *first = off[0]; *last = off[n]; return DVB_OK; }
    DVB_CUDA(ctx, cudaMemcpyAsync(first, off, 8, cudaMemcpyDeviceToHost, io.st));
    DVB_CUDA(ctx, cudaMemcpyAsync(last, off + n, 8, cudaMemcpyDeviceToHost, io.st));
    DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    return DVB_OK;
}

// DVB_GATE_TRACE=1: host-side checkpoints of a dvb_question_gate call on stderr (microseconds since entry)
struct GateTrace {
    bool on; std::chrono::steady_clock::time_point t0;
    GateTrace() : on([] { const char* e = getenv("DVB_GATE_TRACE"); return e && e[0] == '1'; }()), t0(std::chrono::steady_clock::now()) {}
    void at(const char* what) const {
        if (on) fprintf(stderr, "[gate] %-28s %8.0f us\n", what, std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count());
    }
};
extern "C" int32_t dvb_question_gate(dvb_ctx* ctx, int32_t mem, const uint8_t* cand_utf8, const int64_t* cand_off, int64_t n_cand,
                                      const uint8_t* hist_utf8, const int64_t* hist_off, int64_t n_hist,
                                      const uint32_t* syn_src, const int32_t* syn_src_off, const uint32_t* syn_dst, const int32_t* syn_dst_off, int32_t n_syn,
                                      const uint32_t* stop, const int32_t* stop_off, int32_t n_stop,
                                      double threshold, int64_t* out_match) {
    if (!ctx || n_cand < 0 || n_hist < 0 || !cand_off || !out_match || (n_hist > 0 && !hist_off)) { dvb_set_err(ctx, "dvb_question_gate: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_cand == 0) return DVB_OK;
    GateTrace tr;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    tr.at("scope");
    int64_t* d_out = io.out(out_match, (size_t)n_cand);
    if (n_hist == 0) {
        int32_t rc0 = io.check(ctx); if (rc0) return rc0;
        DVB_CUDA(ctx, cudaMemsetAsync(d_out, 0xFF, 8 * (size_t)n_cand, io.st));
        return io.finish(ctx);
    }
    GateUnicode U; int32_t rc = gate_unicode(ctx, &U); if (rc) { dvb_set_err(ctx, "cannot upload unicode tables%s%s"); return rc; }
    GateTablesDev T; GateHostTables H;
    rc = gate_tables(ctx, io, syn_src, syn_src_off, syn_dst, syn_dst_off, n_syn, stop, stop_off, n_stop, &T, &H); if (rc) return rc;
    int64_t c_first = 0, c_last = 0, h_first = 0, h_last = 0;
    rc = gate_span(ctx, io, cand_off, n_cand, &c_first, &c_last); if (rc) return rc;
    rc = gate_span(ctx, io, hist_off, n_hist, &h_first, &h_last); if (rc) return rc;
    const int64_t ctotal = c_last - c_first, htotal = h_last - h_first;
    if (ctotal < 0 || htotal < 0) { dvb_set_err(ctx, "question gate: offsets must not decrease%s%s"); return DVB_E_INVALID; }
    const int64_t ccap_n = (ctotal + 2 * n_cand) * H.growth, hcap_n = (htotal + 2 * n_hist) * H.growth;      // = cap[n] of k_gate_caps
    tr.at("tables + spans");
    // Host buffers with >= 16 MB of candidate text: the text crosses PCIe in GATE_SLICES slices on the lease's copy stream while
    // the key kernel works on the previous slice (the copy and the key kernel cost about the same, 3.4 ms each at config 4)
    const int n_sl = (mem == DVB_MEM_HOST && ctotal >= ((int64_t)16 << 20) && n_cand >= 8 * GATE_SLICES) ? GATE_SLICES : 1;
    uint8_t* d_cu_w = n_sl > 1 ? io.tmp<uint8_t>((size_t)ctotal + 16) : nullptr;
    const uint8_t* d_cu = n_sl > 1 ? d_cu_w - c_first : (mem == DVB_MEM_HOST ? io.in(cand_utf8 + c_first, (size_t)ctotal + 16) - c_first : cand_utf8);
    const uint8_t* d_hu = mem == DVB_MEM_HOST ? io.in(hist_utf8 + h_first, (size_t)htotal + 16) - h_first : hist_utf8;
    const int64_t* d_co = io.in(cand_off, (size_t)n_cand + 1); const int64_t* d_ho = io.in(hist_off, (size_t)n_hist + 1);
    int64_t* d_ccap = io.tmp<int64_t>((size_t)n_cand + 1); int64_t* d_hcap = io.tmp<int64_t>((size_t)n_hist + 1);
    const int64_t bufn = (n_sl == 1 ? ccap_n + hcap_n : std::max(ccap_n, hcap_n)) + 2;     // one launch for both batches: separate regions
    uint32_t* bufA = io.tmp<uint32_t>((size_t)bufn); uint32_t* bufB = io.tmp<uint32_t>((size_t)bufn);
    unsigned long long* d_ck = io.tmp<unsigned long long>((size_t)ccap_n + 1); int32_t* d_cc = io.tmp<int32_t>((size_t)n_cand);
    unsigned long long* d_hk = io.tmp<unsigned long long>((size_t)hcap_n + 1); int32_t* d_hc = io.tmp<int32_t>((size_t)n_hist);
    int64_t* d_poff = io.tmp<int64_t>((size_t)n_hist + 1);
    int32_t* d_err = io.tmp<int32_t>(1);
    rc = io.check(ctx); if (rc) return rc;
    tr.at("inputs staged (H2D queued)");
    DVB_CUDA(ctx, cudaMemsetAsync(d_err, 0, 4, io.st));
    k_gate_caps<<<(int)std::min<int64_t>((n_cand + 256) / 256, (int64_t)ctx->num_sm * 8), 256, 0, io.st>>>(d_co, n_cand, H.growth, d_ccap, d_err);
    k_gate_caps<<<(int)std::min<int64_t>((n_hist + 256) / 256, (int64_t)ctx->num_sm * 8), 256, 0, io.st>>>(d_ho, n_hist, H.growth, d_hcap, d_err);
    ctx->launches += 2;
    DVB_CUDA(ctx, cudaFuncSetAttribute(k_question_keys, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GQ_SMEM_BYTES));
    DVB_CUDA(ctx, cudaFuncSetAttribute(k_question_keys2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GQ_SMEM_BYTES));
    const int gh = (int)std::min<int64_t>((n_hist + GQ_THREADS - 1) / GQ_THREADS, (int64_t)ctx->num_sm * 7);
    if (n_sl == 1) {
        const int gc = (int)std::min<int64_t>((n_cand + GQ_THREADS - 1) / GQ_THREADS, (int64_t)ctx->num_sm * 7);
        const GqBatch qh{ d_hu, d_ho, n_hist, d_hcap, bufA, bufB, d_hk, d_hc, nullptr, nullptr };
        const GqBatch qc{ d_cu, d_co, n_cand, d_ccap, bufA + hcap_n + 1, bufB + hcap_n + 1, d_ck, d_cc, nullptr, nullptr };
        k_question_keys2<<<gh + gc, GQ_THREADS, GQ_SMEM_BYTES, io.st>>>(qh, gh, qc, U, T);
        ctx->launches -= 1;                                                  // the += 3 below counts two key launches
    } else {
        k_question_keys<<<gh, GQ_THREADS, GQ_SMEM_BYTES, io.st>>>(d_hu, d_ho, n_hist, d_hcap, bufA, bufB, U, T, d_hk, d_hc, nullptr, nullptr);
        dvb_lease* l = sc.l;
        if (!l->copy) DVB_CUDA(ctx, cudaStreamCreateWithFlags(&l->copy, cudaStreamNonBlocking));
        while ((int)l->pipe_ev.size() < n_sl + 1) { cudaEvent_t e; DVB_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); l->pipe_ev.push_back(e); }
        DVB_CUDA(ctx, cudaEventRecord(l->pipe_ev[0], io.st));                  // the copy stream starts behind the allocations' last use
        DVB_CUDA(ctx, cudaStreamWaitEvent(l->copy, l->pipe_ev[0], 0));
        const bool pageable = dvb_is_pageable(cand_utf8);
        cudaError_t pe = cudaSuccess;
        for (int k = 0; k < n_sl && pe == cudaSuccess; k++) {
            const int64_t s0 = n_cand * k / n_sl, s1 = n_cand * (k + 1) / n_sl;
            const int64_t b0 = cand_off[s0], b1 = cand_off[s1];
            if (b1 > b0) pe = (pageable && (size_t)(b1 - b0) >= DVB_H2D_MIN) ? dvb_h2d_staged(l, ctx->device, d_cu_w + (b0 - c_first), cand_utf8 + b0, (size_t)(b1 - b0), l->copy)
                                                                              : cudaMemcpyAsync(d_cu_w + (b0 - c_first), cand_utf8 + b0, (size_t)(b1 - b0), cudaMemcpyHostToDevice, l->copy);
            if (pe == cudaSuccess) pe = cudaEventRecord(l->pipe_ev[k + 1], l->copy);
            if (pe == cudaSuccess) pe = cudaStreamWaitEvent(io.st, l->pipe_ev[k + 1], 0);
            if (pe != cudaSuccess) break;
            const int gc = (int)std::min<int64_t>((s1 - s0 + GQ_THREADS - 1) / GQ_THREADS, (int64_t)ctx->num_sm * 7);
            k_question_keys<<<gc, GQ_THREADS, GQ_SMEM_BYTES, io.st>>>(d_cu, d_co + s0, s1 - s0, d_ccap + s0, bufA, bufB, U, T, d_ck, d_cc + s0, nullptr, nullptr);
            ctx->launches += 1;
        }
        if (pe != cudaSuccess) { cudaStreamSynchronize(l->copy); cudaStreamSynchronize(io.st); DVB_CUDA(ctx, pe); }
        ctx->launches -= 1;                                                  // the += 3 below counts one candidate launch
    }
    k_gate_scan_counts<<<1, 1024, 0, io.st>>>(d_hc, d_poff, n_hist);
    int64_t n_pairs = 0; int32_t bad_off = 0;
    DVB_CUDA(ctx, cudaMemcpyAsync(&n_pairs, d_poff + n_hist, 8, cudaMemcpyDeviceToHost, io.st));
    DVB_CUDA(ctx, cudaMemcpyAsync(&bad_off, d_err, 4, cudaMemcpyDeviceToHost, io.st));
    DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    tr.at("keys done (sync)");
    if (bad_off) { dvb_set_err(ctx, "question gate: a question is limited to 60000 bytes%s%s"); return DVB_E_INVALID; }
    ctx->launches += 3;
    if (n_pairs >= 0x7FFFFFFFLL) { dvb_set_err(ctx, "dvb_question_gate: history too large%s%s"); return DVB_E_CAPACITY; }
    unsigned long long* pk = io.tmp<unsigned long long>((size_t)n_pairs + 1); unsigned long long* pk2 = io.tmp<unsigned long long>((size_t)n_pairs + 1);
    uint32_t* pv = io.tmp<uint32_t>((size_t)n_pairs + 1); uint32_t* pv2 = io.tmp<uint32_t>((size_t)n_pairs + 1);
    uint32_t* hist = io.tmp<uint32_t>(dv_radix_hist_words(n_pairs) + 256);
    rc = io.check(ctx); if (rc) return rc;
    if (n_pairs > 0) {
        k_gate_pairs<<<(int)std::min<int64_t>(n_hist, (int64_t)ctx->num_sm * 8), 128, 0, io.st>>>(d_hk, d_hcap, d_hc, d_poff, n_hist, pk, pv);
        ctx->launches += 1;
        if (n_pairs <= GATE_SMALL_SORT) {
            DVB_CUDA(ctx, cudaFuncSetAttribute(k_gate_sort_small, cudaFuncAttributeMaxDynamicSharedMemorySize, 12 * GATE_SMALL_SORT));
            k_gate_sort_small<<<1, 1024, 12 * GATE_SMALL_SORT, io.st>>>(pk, pv, (int)n_pairs);
            ctx->launches += 1;
        } else {
            int64_t l = 0;
            const int which = dv_radix_sort_pairs(pk, pv, pk2, pv2, n_pairs, 0, 48, hist, io.st, &l);   // keys use 43 bits
            ctx->launches += l;
            if (which) { pk = pk2; pv = pv2; }
        }
    }
    const int32_t tile_h = (int32_t)std::min<int64_t>(n_hist, 96 * 1024);
    const size_t smem = (size_t)((tile_h + 1) / 2) * 4;
    DVB_CUDA(ctx, cudaFuncSetAttribute(k_gate_match, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1;
    DVB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_gate_match, GATE_THREADS, smem));
    const int grid = (int)std::min<int64_t>(n_cand, (int64_t)ctx->num_sm * std::max(1, occ));
    const bool timing = ctx->timing.load() != 0;
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev0, io.st));
    const char* force = getenv("DVB_GATE_KERNEL");
    const bool prefix = threshold > 0.0 && n_pairs > 0 && !(force && !strcmp(force, "dense"));
    int32_t* d_nflag = nullptr; int64_t* d_flagged = nullptr;
    GateKeyTable KT{ nullptr, nullptr, nullptr, 0u };
    if (prefix) {
        d_nflag = io.tmp<int32_t>(1); d_flagged = io.tmp<int64_t>((size_t)n_cand);
        if (n_pairs > (1LL << 30)) { dvb_set_err(ctx, "dvb_question_gate: history too large for the key table%s%s"); return DVB_E_CAPACITY; }
        uint32_t cap = 1024; while ((int64_t)cap < 2 * n_pairs) cap <<= 1;            // <= half full (distinct keys <= pairs); n_pairs <= 2^30 keeps cap in 32 bits
        KT.mask = cap - 1;
        KT.key = io.tmp<unsigned long long>((size_t)cap); KT.lo = io.tmp<uint32_t>((size_t)cap); KT.df = io.tmp<uint32_t>((size_t)cap);
        rc = io.check(ctx); if (rc) return rc;
        DVB_CUDA(ctx, cudaMemsetAsync(d_nflag, 0, 4, io.st));
        const size_t smem_p = 8 * sizeof(GpWarp);
        DVB_CUDA(ctx, cudaFuncSetAttribute(k_gate_match_prefix, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
        const int grid_p = (int)std::min<int64_t>((n_cand + 7) / 8, (int64_t)ctx->num_sm * 32);
        DVB_CUDA(ctx, cudaMemsetAsync(KT.key, 0xFF, 8 * (size_t)(KT.mask + 1), io.st));
        const int grid_t = (int)std::min<int64_t>((n_pairs + 255) / 256, (int64_t)ctx->num_sm * 16);
        k_gate_table_heads<<<grid_t, 256, 0, io.st>>>(pk, n_pairs, KT);
        k_gate_table_tails<<<grid_t, 256, 0, io.st>>>(pk, n_pairs, KT);
        k_gate_match_prefix<<<grid_p, 256, smem_p, io.st>>>(d_ck, d_ccap, d_cc, n_cand, pk, pv, n_pairs, d_hk, d_hcap, d_hc, threshold, d_out, d_nflag, d_flagged, KT);
        ctx->launches += 3;
    }
    // dense counters: everything when the prefix filter does not apply, else only the candidates it flagged (> GP_MAXL keys)
    k_gate_match<<<grid, GATE_THREADS, smem, io.st>>>(d_ck, d_ccap, d_cc, n_cand, pk, pv, n_pairs, d_hc, n_hist, tile_h, threshold, d_out, d_nflag, d_flagged);
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev1, io.st));
    ctx->launches += 1;
    tr.at("match queued");
    if (tr.on) { cudaStreamSynchronize(io.st); tr.at("match done (trace-only sync)"); }
    rc = io.finish(ctx);
    tr.at("finish (D2H + sync)");
    if (rc == DVB_OK && timing) { float ms = 0.f; cudaEventElapsedTime(&ms, sc.l->ev0, sc.l->ev1); ctx->last_ms = ms; }
    return rc;
}

// ================================================================== a10: solution signature Jaccard
// One thread per (left, right) pair of ALREADY clean_solution_text()'d strings (tiny volume: <= 8 pairs
// per report).  lower(), drop everything but [0-9a-z] and CJK, n-gram set (n <= 4; a text of <= n
// characters is its own single token), |L&R| and |L|R|.
__device__ int32_t sol_grams(const uint8_t* t, int64_t nb, int32_t n, unsigned long long* keys) {
    // 15-bit symbols: CJK 1..20992, [a-z0-9] 20993..21028; U+212A -> 'k', U+0130 -> 'i' (its U+0307 is dropped)
    int32_t m = 0;
    uint16_t* sym = reinterpret_cast<uint16_t*>(keys);            // symbols overlay the front of the key buffer
    for (int64_t i = 0; i < nb;) {
        uint32_t c = t[i], cp; int len;
        if (c < 0x80u) { cp = c; len = 1; } else if (c < 0xE0u) { cp = c & 0x1Fu; len = 2; } else if (c < 0xF0u) { cp = c & 0x0Fu; len = 3; } else { cp = c & 0x07u; len = 4; }
        for (int j = 1; j < len && i + j < nb; j++) cp = (cp << 6) | (t[i + j] & 0x3Fu);
        i += len;
        uint32_t s = 0;
        if (cp < 0x80u) { const uint32_t l = cp | 0x20u; if (l - 'a' < 26u) s = 20992u + (l - 'a' + 1u); else if (cp - '0' < 10u) s = 20992u + (cp - '0' + 27u); }
        else if (cp - 0x4E00u < 20992u) s = cp - 0x4E00u + 1u;
        else if (cp == 0x212Au) s = 20992u + 11u;
        else if (cp == 0x0130u) s = 20992u + 9u;
        if (s) sym[m++] = (uint16_t)s;                             // 2 B written per >= 1 B read: never overtakes the reader? (sym is a
    }                                                              // different buffer from t, so there is no hazard at all)
    if (m == 0) return 0;
    if (m <= n) {                                                  // a text of <= n characters is its own single token
        unsigned long long k = 0; for (int32_t j = 0; j < m; j++) k = (k << 15) | sym[j];
        keys[0] = m < n ? (((unsigned long long)m << 60) | k) : k;
        return 1;
    }
    const int32_t ng = m - n + 1;
    // keys[i] (8 B) overlays symbols 4i..4i+3.  Going from the back, n-gram i' < i needs symbols <= i+2 < 4i (i >= 1).
    for (int32_t i = ng - 1; i >= 0; i--) {
        unsigned long long k = 0; for (int32_t j = 0; j < n; j++) k = (k << 15) | sym[i + j];
        keys[i] = k;
    }
    gate_heapsort(keys, ng);
    int32_t u = 1; for (int32_t i = 1; i < ng; i++) if (keys[i] != keys[u - 1]) keys[u++] = keys[i];
    return u;
}

__global__ void k_solution_jaccard(const uint8_t* __restrict__ lu, const int64_t* __restrict__ lo, const uint8_t* __restrict__ ru, const int64_t* __restrict__ ro,
                                   int64_t n_pairs, int32_t ngram, const int64_t* __restrict__ cap_off, unsigned long long* __restrict__ bufL,
                                   unsigned long long* __restrict__ bufR, int32_t* __restrict__ out_inter, int32_t* __restrict__ out_union,
                                   int32_t* __restrict__ out_nl, int32_t* __restrict__ out_nr) {
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < n_pairs; p += (int64_t)gridDim.x * blockDim.x) {
        unsigned long long* kl = bufL + cap_off[p]; unsigned long long* kr = bufR + cap_off[p];
        const int32_t a = sol_grams(lu + lo[p], lo[p + 1] - lo[p], ngram, kl);
        const int32_t b = sol_grams(ru + ro[p], ro[p + 1] - ro[p], ngram, kr);
        int32_t i = 0, j = 0, c = 0;
        while (i < a && j < b) { if (kl[i] < kr[j]) i++; else if (kl[i] > kr[j]) j++; else { c++; i++; j++; } }
        out_inter[p] = c; out_union[p] = a + b - c; out_nl[p] = a; out_nr[p] = b;
    }
}

extern "C" int32_t dvb_solution_jaccard(dvb_ctx* ctx, int32_t mem, const uint8_t* left_utf8, const int64_t* left_off,
                                         const uint8_t* right_utf8, const int64_t* right_off, int64_t n_pairs, int32_t ngram,
                                         int32_t* out_inter, int32_t* out_union, int32_t* out_nleft, int32_t* out_nright) {
    if (!ctx || n_pairs < 0 || !left_off || !right_off || ngram < 1 || ngram > 4 || !out_inter || !out_union || !out_nleft || !out_nright) {
        dvb_set_err(ctx, "dvb_solution_jaccard: invalid argument (1 <= ngram <= 4)%s%s"); return DVB_E_INVALID; }
    if (n_pairs == 0) return DVB_OK;
    dvb_scope sc(ctx); dvb_io io(sc, mem);
    std::vector<int64_t> lo((size_t)n_pairs + 1), ro((size_t)n_pairs + 1), cap((size_t)n_pairs + 1);
    if (mem == DVB_MEM_HOST) { memcpy(lo.data(), left_off, 8 * (size_t)(n_pairs + 1)); memcpy(ro.data(), right_off, 8 * (size_t)(n_pairs + 1)); }
    else {
        DVB_CUDA(ctx, cudaMemcpyAsync(lo.data(), left_off, 8 * (size_t)(n_pairs + 1), cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaMemcpyAsync(ro.data(), right_off, 8 * (size_t)(n_pairs + 1), cudaMemcpyDeviceToHost, io.st));
        DVB_CUDA(ctx, cudaStreamSynchronize(io.st));
    }
    cap[0] = 0;
    for (int64_t p = 0; p < n_pairs; p++) cap[p + 1] = cap[p] + std::max(lo[p + 1] - lo[p], ro[p + 1] - ro[p]) + 2;
    const uint8_t* d_lu = mem == DVB_MEM_HOST ? io.in(left_utf8 + lo[0], (size_t)(lo[n_pairs] - lo[0]) + 16) - lo[0] : left_utf8;
    const uint8_t* d_ru = mem == DVB_MEM_HOST ? io.in(right_utf8 + ro[0], (size_t)(ro[n_pairs] - ro[0]) + 16) - ro[0] : right_utf8;
    const int64_t* d_lo = io.in(left_off, (size_t)n_pairs + 1); const int64_t* d_ro = io.in(right_off, (size_t)n_pairs + 1);
    int64_t* d_cap = io.tmp<int64_t>((size_t)n_pairs + 1);
    unsigned long long* bl = io.tmp<unsigned long long>((size_t)cap[n_pairs] + 1); unsigned long long* br = io.tmp<unsigned long long>((size_t)cap[n_pairs] + 1);
    int32_t* d_i = io.out(out_inter, (size_t)n_pairs); int32_t* d_u = io.out(out_union, (size_t)n_pairs);
    int32_t* d_nl = io.out(out_nleft, (size_t)n_pairs); int32_t* d_nr = io.out(out_nright, (size_t)n_pairs);
    int32_t rc = io.check(ctx); if (rc) return rc;
    DVB_CUDA(ctx, cudaMemcpyAsync(d_cap, cap.data(), 8 * (size_t)(n_pairs + 1), cudaMemcpyHostToDevice, io.st));
    k_solution_jaccard<<<(int)std::min<int64_t>((n_pairs + 63) / 64, 1024), 64, 0, io.st>>>(d_lu, d_lo, d_ru, d_ro, n_pairs, ngram, d_cap, bl, br, d_i, d_u, d_nl, d_nr);
    ctx->launches += 1;
    return io.finish(ctx);
}
