// bind_general.cuh — k_bind_claims_general: the exact any-size claim->fact binding kernel (included by bind.cu).
// replaces: the scoring loop of infer_weak_evidence_refs_v3 (reference web/server.py:24813-24864)
// It is the re-run target of the sessions k_bind_claims (bind7.cuh) hands back; the phases are described in bind.cu.
#pragma once
#include "bind_common.cuh"
#include "text.cuh"

struct BindShared {       // fixed part; per-warp hit bitmaps follow in dynamic smem
    dv_warp_buf wbuf[BIND_WARPS];
    unsigned long long ktab[DV_KTAB];
    unsigned long long queue[BIND_WARPS][BIND_QCAP];
    unsigned long long wset[BIND_WARPS][BIND_WSET];
    uint32_t fb[1u << (BIND_FB_BITS - 5)];
    uint32_t ft[1u << (BIND_FT_BITS - 5)];
    uint32_t ov[BIND_WARPS][BIND_CLAIMS];
    double rbest[BIND_WARPS][BIND_CLAIMS];
    int ridx[BIND_WARPS][BIND_CLAIMS];
    int wcnt[BIND_WARPS];
    int lenA[BIND_CLAIMS];
    int next_item;
};

// ---- claims: insert every token, set filter bits
struct EmitClaim {
    unsigned long long* tkeys; uint32_t* tcmask; uint32_t mask; uint32_t* fb; uint32_t* ft; int* lenA; int c;
    __device__ __forceinline__ void restart() {}
    __device__ __forceinline__ void insert(uint64_t key) {
        uint32_t slot = bind_hash(key) & mask;
        for (;;) {
            const unsigned long long prev = atomicCAS(&tkeys[slot], 0ULL, (unsigned long long)key);
            if (prev == 0ULL || prev == key) break;
            slot = (slot + 1) & mask;
        }
        const uint32_t bit = 1u << c;
        const uint32_t old = atomicOr(&tcmask[slot], bit);
        if (!(old & bit)) atomicAdd(&lenA[c], 1);
    }
    __device__ __forceinline__ void bigram(bool v, uint32_t id) {
        if (!v) return;
        const uint32_t h = bind_fb_index(id);
        atomicOr(&fb[h >> 5], 1u << (h & 31));
        insert((uint64_t)id);
    }
    __device__ __forceinline__ bool need_token(uint32_t) const { return true; }
    __device__ __forceinline__ void token(bool v, uint64_t key, uint32_t sig) {
        if (!v) return;
        const uint32_t h = bind_ft_index(sig);
        atomicOr(&ft[h >> 5], 1u << (h & 31));
        insert(key);
    }
};

// ---- facts: filter, queue survivors, probe 32 at a time, mark hit slots
struct EmitFact {
    const unsigned long long* tkeys; uint32_t mask; const uint32_t* fb; const uint32_t* ft;
    unsigned long long* queue; uint32_t* hit; unsigned long long* wset; int* wcnt;
    int qn; bool full; int lane;
    __device__ __forceinline__ void restart() {}
    __device__ __forceinline__ void count_distinct(bool v, uint64_t key) {       // until 18 distinct tokens were seen
        if (v) {
            uint32_t s = (bind_hash(key) >> 20) & (BIND_WSET - 1);
            for (;;) {
                const unsigned long long prev = atomicCAS(&wset[s], 0ULL, (unsigned long long)key);
                if (prev == 0ULL) { atomicAdd(wcnt, 1); break; }
                if (prev == key) break;
                s = (s + 1) & (BIND_WSET - 1);
            }
        }
        __syncwarp();
        full = (*(volatile int*)wcnt >= 18);
    }
    __device__ __forceinline__ void probe(bool v, uint64_t key) {
        if (!v) return;
        uint32_t slot = bind_hash(key) & mask;
        for (;;) {
            const unsigned long long k = __ldcg(&tkeys[slot]);     // table is written by L2 atomics: bypass L1
            if (k == key) { atomicOr(&hit[slot >> 5], 1u << (slot & 31)); return; }
            if (k == 0ULL) return;
            slot = (slot + 1) & mask;
        }
    }
    __device__ __forceinline__ void push(bool v, uint64_t key) {
        const uint32_t m = __ballot_sync(0xffffffffu, v);
        if (!m) return;
        if (v) queue[qn + __popc(m & ((1u << lane) - 1u))] = key;
        qn += __popc(m);
        __syncwarp();
        if (qn >= 32) {
            probe(true, queue[lane]);
            __syncwarp();
            const unsigned long long rest = (lane + 32 < qn) ? queue[lane + 32] : 0ULL;
            __syncwarp();
            queue[lane] = rest;
            qn -= 32;
            __syncwarp();
        }
    }
    __device__ __forceinline__ void flush() {
        __syncwarp();
        probe(lane < qn, lane < qn ? queue[lane] : 0ULL);
        qn = 0;
        __syncwarp();
    }
    __device__ __forceinline__ void bigram(bool v, uint32_t id) {
        if (!full) count_distinct(v, (uint64_t)id);
        const uint32_t h = bind_fb_index(id);
        const bool pass = v && ((fb[h >> 5] >> (h & 31)) & 1u);
        push(pass, (uint64_t)id);
    }
    __device__ __forceinline__ bool ft_pass(uint32_t sig) const { const uint32_t h = bind_ft_index(sig); return (ft[h >> 5] >> (h & 31)) & 1u; }
    __device__ __forceinline__ bool need_token(uint32_t sig) const { return !full || ft_pass(sig); }
    __device__ __forceinline__ void token(bool v, uint64_t key, uint32_t sig) {
        if (!full) count_distinct(v, key);
        push(v && ft_pass(sig), key);
    }
};

#ifdef DV_EMU
static __align__(16) unsigned char bind_smem[192 * 1024];
#else
extern __shared__ __align__(16) unsigned char bind_smem[];
#endif

__global__ void __launch_bounds__(BIND_THREADS) k_bind_claims_general(BindParams p) {
    BindShared& sh = *reinterpret_cast<BindShared*>(bind_smem);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t hit_words = p.cap >> 5;
    unsigned long long* tkeys = p.gtable + (size_t)blockIdx.x * ((size_t)p.cap + (size_t)p.cap / 2);
    uint32_t* tcmask = reinterpret_cast<uint32_t*>(tkeys + p.cap);
    uint32_t* hit = p.ghit ? p.ghit + ((size_t)blockIdx.x * BIND_WARPS + warp) * hit_words
                           : reinterpret_cast<uint32_t*>(bind_smem + ((sizeof(BindShared) + 15) & ~size_t(15))) + (size_t)warp * hit_words;
    const uint32_t mask = p.cap - 1;
    dv_fill_ktab(sh.ktab);
    for (uint32_t i = lane; i < hit_words; i += 32) hit[i] = 0u;
    __syncthreads();

    for (int64_t s = p.s_begin + blockIdx.x; s < p.s_end; s += gridDim.x) {
        if (p.only_flagged && !p.only_flagged[s]) continue;       // re-run of the sessions the fast kernel handed back
        const int64_t c0 = p.sess_claim_off[s], c1 = p.sess_claim_off[s + 1];
        const int64_t f0 = p.sess_fact_off[s], f1 = p.sess_fact_off[s + 1];
        for (int64_t cg = c0; cg < c1; cg += BIND_CLAIMS) {
            const int ncg = (int)min((int64_t)BIND_CLAIMS, c1 - cg);
            // ---- clear table + filters
            for (uint32_t i = tid; i < p.cap; i += BIND_THREADS) { tkeys[i] = 0ULL; tcmask[i] = 0u; }
            for (uint32_t i = tid; i < (1u << (BIND_FB_BITS - 5)); i += BIND_THREADS) sh.fb[i] = 0u;
            for (uint32_t i = tid; i < (1u << (BIND_FT_BITS - 5)); i += BIND_THREADS) sh.ft[i] = 0u;
            if (tid < BIND_CLAIMS) sh.lenA[tid] = 0;
            if (tid == 0) sh.next_item = 0;
            __syncthreads();
            // ---- phase C
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&sh.next_item, 1);
                c = __shfl_sync(0xffffffffu, c, 0);
                if (c >= ncg) break;
                const int64_t b = p.claim_off[cg + c], e = p.claim_off[cg + c + 1];
                EmitClaim emit{ tkeys, tcmask, mask, sh.fb, sh.ft, sh.lenA, c };
                dv_tokenize_v3_warp(p.claim_utf8 + b, (int32_t)(e - b), &sh.wbuf[warp], sh.ktab, emit);
            }
            __syncthreads();
            if (tid == 0) sh.next_item = 0;
            __syncthreads();
            // ---- phase F
            double best = -1.0; int best_f = 0x7fffffff;          // lane c (< ncg) tracks claim c
            const int la = lane < ncg ? sh.lenA[lane] : 0;
            const int da = max(3, min(la, 14));
            const int pref = lane < ncg ? p.claim_pref_dim[cg + lane] : -1;
            for (;;) {
                int f = 0;
                if (lane == 0) f = atomicAdd(&sh.next_item, 1);
                f = __shfl_sync(0xffffffffu, f, 0);
                if ((int64_t)f >= f1 - f0) break;
                sh.wset[warp][lane] = 0ULL; sh.wset[warp][lane + 32] = 0ULL;
                sh.ov[warp][lane] = 0u;
                if (lane == 0) sh.wcnt[warp] = 0;
                __syncwarp();
                const int64_t b = p.fact_off[f0 + f], e = p.fact_off[f0 + f + 1];
                EmitFact emit{ tkeys, mask, sh.fb, sh.ft, sh.queue[warp], hit, sh.wset[warp], &sh.wcnt[warp], 0, false, lane };
                dv_tokenize_v3_warp(p.fact_utf8 + b, (int32_t)(e - b), &sh.wbuf[warp], sh.ktab, emit);
                emit.flush();
                // hit bitmap -> per-claim overlap counts; clear as we go
                for (uint32_t wd = lane; wd < hit_words; wd += 32) {
                    uint32_t bits = *(volatile uint32_t*)&hit[wd];
                    if (bits) {
                        hit[wd] = 0u;
                        while (bits) {
                            const uint32_t slot = wd * 32 + (__ffs(bits) - 1); bits &= bits - 1;
                            uint32_t cm = __ldcg(&tcmask[slot]);
                            while (cm) { const int c = __ffs(cm) - 1; cm &= cm - 1; atomicAdd(&sh.ov[warp][c], 1u); }
                        }
                    }
                }
                __syncwarp();
                const int lb = min(sh.wcnt[warp], 18);
                if (lane == 0 && p.len_fact && cg == c0) p.len_fact[f0 + f] = lb;
                const uint32_t ov = sh.ov[warp][lane];
                if (lane < ncg && ov > 0u) {
                    const int db = max(4, min(lb, 18));
                    const double coverage = __ddiv_rn((double)ov, (double)da);
                    const double precision = __ddiv_rn((double)ov, (double)db);
                    double q = p.fact_quality[f0 + f];
                    q = q > 0.0 ? (q < 1.0 ? q : 1.0) : 0.0;
                    const double dimb = (pref >= 0 && p.fact_dim[f0 + f] == pref) ? 0.12 : 0.0;
                    const double hardb = p.fact_hard[f0 + f] ? 0.04 : 0.0;
                    const double sc = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(0.62, coverage), __dmul_rn(0.18, precision)),
                                                                     __dmul_rn(0.16, q)), dimb), hardb);
                    if (sc > best || (sc == best && f < best_f)) { best = sc; best_f = f; }
                }
                __syncwarp();
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

