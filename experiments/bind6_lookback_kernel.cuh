// bind6.cuh — k_bind_claims: the fast variant of the fused claim->fact binding kernel (included by bind.cu).
// replaces: the scoring loop of infer_weak_evidence_refs_v3 (reference web/server.py:24813-24864)
//
// Same phases and results as k_bind_claims_general (bind.cu); what differs is where the claim tokens live:
//   * CJK bigrams (>= 85 % of all tokens) go to an EXACT bucketed table in shared memory: 4 x 30-bit ids per
//     16-byte bucket, one LDS.128 + 4 compares per fact bigram, no filter and no queue in front of it;
//   * pieces / ASCII tokens (64-bit keys, rare) keep a small open-addressing table in the L2-resident workspace
//     behind a 1 KB signature filter; a claim group with more than B6_TOK_LIMIT of them raises `overflow`
//     and the host re-runs the call on the general kernel;
//   * text is streamed by the lean tokenizer of text6.cuh;
//   * min(len(B), 18): the first 32 symbols of a fact usually hold >= 18 distinct bigrams (one MATCH.ANY).
#pragma once
#include "text6.cuh"

#define B6_CAPB 4096            // bigram-table slots (shared memory); claim groups that need more use the general kernel
#define B6_NBM (B6_CAPB / 4 - 1)
#define B6_BSHIFT 22            // 32 - log2(B6_CAPB / 4)
#define B6_CAPT 1024            // token-table slots (global workspace)
#define B6_TOK_LIMIT 704        // inserts stop here; up to 256 racing threads still find a free slot
#define B6_HW ((B6_CAPB + B6_CAPT) / 32)
#define B6_FT_BITS 15
#define B6_LIST 600             // hit-slot list entries (u16) that fit the idle tokenizer buffer of the warp
#define B6_SPILL 0x80000000u     // bit 31 of a bucket's last slot: an insert continued into the next bucket
#ifndef B6_CLAIM_RED
#define B6_CLAIM_RED 1          // claim marks as fire-and-forget REDs, len(A) through the warp's shared bitmap
#endif
#ifndef B6_MIN_CTAS
#define B6_MIN_CTAS 4
#endif

struct Bind6Shared {            // static shared memory: every access is [register + immediate]
    dv6_warp_buf wbuf[BIND_WARPS];
    uint32_t tab[B6_CAPB];                                   // bigram ids, 4 per 16-byte bucket, 0 = free
    uint32_t hit[BIND_WARPS][B6_HW];                         // per warp: table slots (bigram | token) hit by the current fact
    unsigned long long ktab[DV_KTAB];
    unsigned long long wset[BIND_WARPS][BIND_WSET];
    double rbest[BIND_WARPS][BIND_CLAIMS];
    uint32_t ft[1u << (B6_FT_BITS - 5)];
    uint32_t ov[BIND_WARPS][BIND_CLAIMS];
    int ridx[BIND_WARPS][BIND_CLAIMS];
    int wcnt[BIND_WARPS];
    int lenA[BIND_CLAIMS];
    int next_item, ntok, ovf;
    uint8_t lut[128];
};
static_assert(sizeof(dv6_warp_buf) >= B6_LIST * 2, "hit-slot list must fit the tokenizer buffer");
__shared__ Bind6Shared b6;

// token filter: one probe into 2^15 bits (false positives ~1 % at 350 tokens per claim group)
__device__ __forceinline__ bool b6_ft_pass(uint32_t sig) {
    const uint32_t h = sig >> (32 - B6_FT_BITS);
    return ((b6.ft[h >> 5] >> (h & 31)) & 1u) != 0u;
}

struct Emit6Claim {
    unsigned long long* tkeys; uint32_t* cmask; int c; int warp;
    __device__ __forceinline__ void restart() {}
    // slot -> claim bit in the L2-resident mask (fire-and-forget RED); len(A) counts the claim's DISTINCT slots through
    // the warp's shared-memory bitmap (one warp tokenises one claim), so no global atomic has to return a value
    __device__ __forceinline__ void mark(uint32_t slot) {
#if B6_CLAIM_RED
        atomicOr(&cmask[slot], 1u << c);
        const uint32_t bit = 1u << (slot & 31);
        const uint32_t old = atomicOr(&b6.hit[warp][slot >> 5], bit);
        if (!(old & bit)) atomicAdd(&b6.lenA[c], 1);
#else
        const uint32_t bit = 1u << c;
        const uint32_t old = atomicOr(&cmask[slot], bit);
        if (!(old & bit)) atomicAdd(&b6.lenA[c], 1);
#endif
    }
    __device__ __forceinline__ void bigram(bool v, uint32_t id) {
        if (!v) return;
        uint32_t b = (id * 0x9E3779B1u) >> B6_BSHIFT;
        for (uint32_t tries = 0; tries <= B6_NBM; tries++) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const uint32_t prev = atomicCAS(&b6.tab[b * 4u + j], 0u, id) & 0x7FFFFFFFu;
                if (prev == 0u || prev == id) { mark(b * 4u + j); return; }
            }
            atomicOr(&b6.tab[b * 4u + 3u], B6_SPILL);           // all four slots taken: the chain continues in the next bucket
            b = (b + 1u) & B6_NBM;
        }
        b6.ovf = 1;                                             // table sized from a wrong hint
    }
    __device__ __forceinline__ bool need_token(uint32_t) const { return true; }
    __device__ __forceinline__ void token(bool v, uint64_t key, uint32_t sig) {
        if (!v) return;
        if (*(volatile int*)&b6.ntok >= B6_TOK_LIMIT) { b6.ovf = 1; return; }
        const uint32_t h = sig >> (32 - B6_FT_BITS);
        atomicOr(&b6.ft[h >> 5], 1u << (h & 31));
        uint32_t slot = bind_hash(key) & (B6_CAPT - 1);
        for (;;) {
            const unsigned long long prev = atomicCAS(&tkeys[slot], 0ULL, (unsigned long long)key);
            if (prev == 0ULL) { atomicAdd(&b6.ntok, 1); break; }
            if (prev == key) break;
            slot = (slot + 1) & (B6_CAPT - 1);
        }
        mark(B6_CAPB + slot);
    }
};

struct Emit6Fact {
    const unsigned long long* tkeys; int warp;
    bool full, first; int lane;
    __device__ __forceinline__ void restart() { first = false; }                 // what was emitted is a correct prefix
    __device__ __forceinline__ void count_distinct(bool v, uint64_t key) {       // until 18 distinct tokens were seen
        if (v) {
            uint32_t s = (bind_hash(key) >> 20) & (BIND_WSET - 1);
            for (;;) {
                const unsigned long long prev = atomicCAS(&b6.wset[warp][s], 0ULL, (unsigned long long)key);
                if (prev == 0ULL) { atomicAdd(&b6.wcnt[warp], 1); break; }
                if (prev == key) break;
                s = (s + 1) & (BIND_WSET - 1);
            }
        }
        __syncwarp();
        full = (*(volatile int*)&b6.wcnt[warp] >= 18);
    }
    __device__ __forceinline__ void bigram(bool v, uint32_t id) {
        if (!full) {
            if (first) {                                        // first step of the fact: distinct bigrams among 32 symbols
                first = false;
                const uint32_t mm = __match_any_sync(0xffffffffu, v ? id : (0x80000000u | (uint32_t)lane));
                const uint32_t nd = __popc(__ballot_sync(0xffffffffu, v && (__ffs(mm) - 1 == lane)));
                if (nd >= 18u) { full = true; if (lane == 0) b6.wcnt[warp] = 18; }
            }
            if (!full) count_distinct(v, (uint64_t)id);
        }
        if (v) {
            uint32_t b = (id * 0x9E3779B1u) >> B6_BSHIFT;
            for (;;) {
                const uint4 q = *reinterpret_cast<const uint4*>(&b6.tab[b * 4u]);
                const uint32_t m = (q.x == id ? 1u : 0u) | (q.y == id ? 2u : 0u) | (q.z == id ? 4u : 0u) | ((q.w & 0x7FFFFFFFu) == id ? 8u : 0u);
                if (m) { const uint32_t hs = b * 4u + (uint32_t)(__ffs(m) - 1); atomicOr(&b6.hit[warp][hs >> 5], 1u << (hs & 31)); break; }
                if (!(q.w & B6_SPILL)) break;                   // no insert ever spilled out of this bucket: the chain ends here
                b = (b + 1u) & B6_NBM;
            }
        }
    }
    __device__ __forceinline__ bool need_token(uint32_t sig) const { return !full | b6_ft_pass(sig); }
    __device__ __forceinline__ void token(bool v, uint64_t key, uint32_t sig) {
        if (!full) count_distinct(v, key);
        if (v) {                                               // v implies the filter passed (or the count was still open)
            uint32_t slot = bind_hash(key) & (B6_CAPT - 1);
            for (;;) {
                const unsigned long long k = __ldcg(&tkeys[slot]);
                if (k == key) { const uint32_t hs = B6_CAPB + slot; atomicOr(&b6.hit[warp][hs >> 5], 1u << (hs & 31)); break; }
                if (k == 0ULL) break;
                slot = (slot + 1) & (B6_CAPT - 1);
            }
        }
    }
};

static inline size_t bind6_ws_u64_per_cta() { return (size_t)B6_CAPT + ((size_t)B6_CAPB + B6_CAPT) / 2; }

__global__ void __launch_bounds__(BIND_THREADS, B6_MIN_CTAS) k_bind_claims(BindParams p) {
    Bind6Shared& sh = b6;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned long long* tkeys = p.gtable + (size_t)blockIdx.x * ((size_t)B6_CAPT + ((size_t)B6_CAPB + B6_CAPT) / 2);
    uint32_t* cmask = reinterpret_cast<uint32_t*>(tkeys + B6_CAPT);
    dv_fill_ktab(sh.ktab);
    dv6_fill_lut(sh.lut);
    for (uint32_t i = lane; i < B6_HW; i += 32) sh.hit[warp][i] = 0u;
    __syncthreads();

    for (int64_t s = p.s_begin + blockIdx.x; s < p.s_end; s += gridDim.x) {
        const int64_t c0 = p.sess_claim_off[s], c1 = p.sess_claim_off[s + 1];
        const int64_t f0 = p.sess_fact_off[s], f1 = p.sess_fact_off[s + 1];
        for (int64_t cg = c0; cg < c1; cg += BIND_CLAIMS) {
            const int ncg = (int)min((int64_t)BIND_CLAIMS, c1 - cg);
            // ---- clear tables + filter
            for (uint32_t i = tid; i < (B6_CAPB >> 2); i += BIND_THREADS) reinterpret_cast<uint4*>(sh.tab)[i] = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t i = tid; i < B6_CAPT; i += BIND_THREADS) tkeys[i] = 0ULL;
            for (uint32_t i = tid; i < B6_CAPB + B6_CAPT; i += BIND_THREADS) cmask[i] = 0u;
            for (uint32_t i = tid; i < (1u << (B6_FT_BITS - 5)); i += BIND_THREADS) sh.ft[i] = 0u;
            if (tid < BIND_CLAIMS) sh.lenA[tid] = 0;
            if (tid == 0) { sh.next_item = 0; sh.ntok = 0; sh.ovf = 0; }
            __syncthreads();
            // ---- phase C
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&sh.next_item, 1);
                c = __shfl_sync(0xffffffffu, c, 0);
                if (c >= ncg) break;
                const int64_t b = p.claim_off[cg + c], e = p.claim_off[cg + c + 1];
                Emit6Claim emit{ tkeys, cmask, c, warp };
                dv6_tokenize_warp(p.claim_utf8 + b, (int32_t)(e - b), &sh.wbuf[warp], sh.ktab, sh.lut, emit);
#if B6_CLAIM_RED
                __syncwarp();
                for (uint32_t i = lane; i < B6_HW; i += 32) sh.hit[warp][i] = 0u;
                __syncwarp();
#endif
            }
            __syncthreads();
            if (tid == 0) { sh.next_item = 0; if (sh.ovf) *p.overflow = 1; }
            __syncthreads();
            // ---- phase F
            double best = -1.0; int best_f = 0x7fffffff;          // lane c (< ncg) tracks claim c
            const int la = lane < ncg ? sh.lenA[lane] : 0;
            const int da = max(3, min(la, 14));
            const int pref = lane < ncg ? p.claim_pref_dim[cg + lane] : -1;
            const int nf = (int)(f1 - f0);
            int f = 0;
            if (lane == 0) f = atomicAdd(&sh.next_item, 1);
            f = __shfl_sync(0xffffffffu, f, 0);
            int64_t b = 0, e = 0;
            if (f < nf) { b = p.fact_off[f0 + f]; e = p.fact_off[f0 + f + 1]; }
            while (f < nf) {
                sh.wset[warp][lane] = 0ULL; sh.wset[warp][lane + 32] = 0ULL;
                sh.ov[warp][lane] = 0u;
                if (lane == 0) sh.wcnt[warp] = 0;
                __syncwarp();
                Emit6Fact emit{ tkeys, warp, false, true, lane };
                dv6_tokenize_warp(p.fact_utf8 + b, (int32_t)(e - b), &sh.wbuf[warp], sh.ktab, sh.lut, emit);
                __syncwarp();
                const double fq = p.fact_quality[f0 + f];                 // issued here, consumed after the hit scan
                const int fdim = p.fact_dim[f0 + f];
                const uint32_t fhard = p.fact_hard[f0 + f];
                // hit bitmap -> per-claim overlap counts.  Pass 1 compacts the hit slots of the warp's bitmap into a list
                // (the tokenizer buffer is free now); pass 2 takes one slot per lane, so the L2 loads of the claim masks go
                // out 32 at a time and the per-claim increments run on full warps.  Slots beyond the list go the direct way.
                const int lb = min(sh.wcnt[warp], 18);
                {
                    uint16_t* list = reinterpret_cast<uint16_t*>(&sh.wbuf[warp]);
                    int* cnt = &sh.wcnt[warp];
                    __syncwarp();
                    if (lane == 0) *cnt = 0;
                    __syncwarp();
#pragma unroll
                    for (uint32_t wd = lane; wd < B6_HW; wd += 32) {
                        uint32_t bits = sh.hit[warp][wd];
                        if (bits) {
                            sh.hit[warp][wd] = 0u;
                            int pos = atomicAdd(cnt, __popc(bits));
                            const uint32_t base = wd * 32u - 1u;
                            do {
                                const uint32_t slot = base + __ffs(bits); bits &= bits - 1u;
                                if (pos < B6_LIST) list[pos] = (uint16_t)slot;
                                else { uint32_t cm = __ldcg(&cmask[slot]); while (cm) { atomicAdd(&sh.ov[warp][__ffs(cm) - 1], 1u); cm &= cm - 1u; } }
                                pos++;
                            } while (bits);
                        }
                    }
                    __syncwarp();
                    const int nl = min(*(volatile int*)cnt, B6_LIST);
                    for (int i0 = 0; i0 < nl; i0 += 64) {
                        uint32_t cm0 = 0u, cm1 = 0u;
                        if (i0 + lane < nl) cm0 = __ldcg(&cmask[list[i0 + lane]]);
                        if (i0 + 32 + lane < nl) cm1 = __ldcg(&cmask[list[i0 + 32 + lane]]);
                        while (cm0) { atomicAdd(&sh.ov[warp][__ffs(cm0) - 1], 1u); cm0 &= cm0 - 1u; }
                        while (cm1) { atomicAdd(&sh.ov[warp][__ffs(cm1) - 1], 1u); cm1 &= cm1 - 1u; }
                    }
                }
                __syncwarp();
                if (lane == 0 && p.len_fact && cg == c0) p.len_fact[f0 + f] = lb;
                const uint32_t ov = sh.ov[warp][lane];
                if (lane < ncg && ov > 0u) {
                    const int db = max(4, min(lb, 18));
                    const double coverage = __ddiv_rn((double)ov, (double)da);
                    const double precision = __ddiv_rn((double)ov, (double)db);
                    const double q = fq > 0.0 ? (fq < 1.0 ? fq : 1.0) : 0.0;
                    const double dimb = (pref >= 0 && fdim == pref) ? 0.12 : 0.0;
                    const double hardb = fhard ? 0.04 : 0.0;
                    const double sc = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(0.62, coverage), __dmul_rn(0.18, precision)),
                                                                     __dmul_rn(0.16, q)), dimb), hardb);
                    if (sc > best || (sc == best && f < best_f)) { best = sc; best_f = f; }
                }
                __syncwarp();
                if (lane == 0) f = atomicAdd(&sh.next_item, 1);
                f = __shfl_sync(0xffffffffu, f, 0);
                if (f < nf) { b = p.fact_off[f0 + f]; e = p.fact_off[f0 + f + 1]; }
            }
            sh.rbest[warp][lane] = best; sh.ridx[warp][lane] = best_f;
            __syncthreads();
            // ---- phase R
            if (tid < ncg) {
                double b = -1.0; int bf = 0x7fffffff;
                for (int w = 0; w < BIND_WARPS; w++) {
                    const double sc = sh.rbest[w][tid]; const int fi = sh.ridx[w][tid];
                    if (sc > b || (sc == b && fi < bf)) { b = sc; bf = fi; }
                }
                const int64_t c = cg + tid;
                const bool any = b >= 0.0;
                p.best_fact[c] = any ? bf : -1;
                p.best_score[c] = any ? b : 0.0;
                p.bound[c] = (any && !(__dadd_rn(b, 1e-9) < p.claim_threshold[c])) ? 1 : 0;
                if (p.len_claim) p.len_claim[c] = min(sh.lenA[tid], 14);
            }
            __syncthreads();
        }
    }
}
