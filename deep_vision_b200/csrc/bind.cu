// bind.cu — dvb_bind_claims: fused tokenise + set-overlap + fp64 score + first-max argmax.
// replaces: the scoring loop of infer_weak_evidence_refs_v3 (reference web/server.py:24813-24864)
//
// One CTA per session (persistent grid, sessions strided).  Per group of <=32 claims:
//   phase C  warps tokenise the claims into a shared-memory open-addressing table
//            key -> bitmask of claims containing the token; len(A) counted on first insert.
//   phase F  per batch of <=64 facts: warps stream facts (coalesced 16-byte loads, text.cuh),
//            probe each token; a per-entry 64-bit `seen` mask makes the count a SET overlap;
//            a 64-slot per-warp set gives min(len(B), 18) exactly.
//   phase S  fp64 scores in the reference's operation order (no FMA contraction), strict '>'
//            scanning facts in index order (first maximum wins).
// No intermediate key arrays ever reach HBM: algorithmic traffic = claim + fact bytes in, 16 B/claim out.
#include "common.cuh"
#include "text.cuh"

#define BIND_WARPS 8
#define BIND_THREADS (BIND_WARPS * 32)
#define BIND_CLAIMS 32      // claims per group (bits of cmask)
#define BIND_FACTS 64       // facts per batch (bits of seen)
#define BIND_WSET 64        // per-warp distinct-token set slots (>= 17 + 32)

struct BindParams {
    const uint8_t* claim_utf8; const int64_t* claim_off; const uint8_t* fact_utf8; const int64_t* fact_off;
    const int64_t* sess_claim_off; const int64_t* sess_fact_off; int64_t n_sessions;
    const double* fact_quality; const int32_t* fact_dim; const uint8_t* fact_hard;
    const int32_t* claim_pref_dim; const double* claim_threshold;
    int32_t* best_fact; double* best_score; uint8_t* bound; int32_t* len_claim; int32_t* len_fact;
    uint32_t cap;                         // table capacity (power of two)
    unsigned long long* gtable;           // global-memory tables (per CTA) when the table does not fit smem, else null
};

__device__ __forceinline__ uint32_t bind_hash(uint64_t key) {
    uint32_t x = (uint32_t)key * 0x85EBCA6Bu ^ (uint32_t)(key >> 32) * 0xC2B2AE35u;
    x ^= x >> 16; x *= 0x9E3779B1u; x ^= x >> 15;
    return x;
}

struct BindShared {       // fixed part; tables follow in dynamic smem
    dv_warp_window win[BIND_WARPS];
    unsigned long long wset[BIND_WARPS][BIND_WSET];
    int wcnt[BIND_WARPS];
    uint32_t ov[BIND_FACTS * BIND_CLAIMS];
    int lenA[BIND_CLAIMS];
    int lenB[BIND_FACTS];
    int next_item;
};

struct EmitClaim {
    unsigned long long* tkeys; uint32_t* tcmask; uint32_t mask; int* lenA; int c;
    __device__ __forceinline__ void restart() {}
    __device__ __forceinline__ void operator()(bool v, uint64_t key) {
        if (!v) return;
        uint32_t slot = bind_hash(key) & mask;
        for (;;) {
            unsigned long long prev = atomicCAS(&tkeys[slot], 0ULL, (unsigned long long)key);
            if (prev == 0ULL || prev == key) break;
            slot = (slot + 1) & mask;
        }
        const uint32_t bit = 1u << c;
        const uint32_t old = atomicOr(&tcmask[slot], bit);
        if (!(old & bit)) atomicAdd(&lenA[c], 1);
    }
};

struct EmitFact {
    const unsigned long long* tkeys; const uint32_t* tcmask; unsigned long long* tseen; uint32_t mask;
    uint32_t* ov_row; int fl; unsigned long long* wset; int* wcnt; bool full;
    __device__ __forceinline__ void restart() {}
    __device__ __forceinline__ void operator()(bool v, uint64_t key) {
        const uint32_t h = bind_hash(key);
        if (!full) {                                             // distinct count up to 18 (warp-uniform branch)
            if (v) {
                uint32_t s = (h >> 20) & (BIND_WSET - 1);
                for (;;) {
                    unsigned long long prev = atomicCAS(&wset[s], 0ULL, (unsigned long long)key);
                    if (prev == 0ULL) { atomicAdd(wcnt, 1); break; }
                    if (prev == key) break;
                    s = (s + 1) & (BIND_WSET - 1);
                }
            }
            __syncwarp();
            full = (*(volatile int*)wcnt >= 18);
        }
        if (!v) return;
        uint32_t slot = h & mask;
        for (;;) {
            const unsigned long long k = tkeys[slot];
            if (k == key) {
                const unsigned long long bit = 1ULL << fl;
                const unsigned long long old = atomicOr(&tseen[slot], bit);
                if (!(old & bit)) {
                    uint32_t cm = tcmask[slot];
                    while (cm) { const int c = __ffs(cm) - 1; cm &= cm - 1; atomicAdd(&ov_row[c], 1u); }
                }
                return;
            }
            if (k == 0ULL) return;
            slot = (slot + 1) & mask;
        }
    }
};

extern __shared__ __align__(16) unsigned char bind_smem[];

__global__ void __launch_bounds__(BIND_THREADS) k_bind_claims(BindParams p) {
    BindShared& sh = *reinterpret_cast<BindShared*>(bind_smem);
    unsigned long long* tkeys; unsigned long long* tseen; uint32_t* tcmask;
    if (p.gtable) {
        unsigned long long* base = p.gtable + (size_t)blockIdx.x * ((size_t)p.cap * 2 + (size_t)p.cap / 2);
        tkeys = base; tseen = base + p.cap; tcmask = reinterpret_cast<uint32_t*>(base + 2 * (size_t)p.cap);
    } else {
        unsigned long long* base = reinterpret_cast<unsigned long long*>(bind_smem + ((sizeof(BindShared) + 15) & ~size_t(15)));
        tkeys = base; tseen = base + p.cap; tcmask = reinterpret_cast<uint32_t*>(base + 2 * (size_t)p.cap);
    }
    double* scores = reinterpret_cast<double*>(tseen);          // aliases `seen` during phase S (cap >= 2048 entries)
    const uint32_t mask = p.cap - 1;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    for (int64_t s = blockIdx.x; s < p.n_sessions; s += gridDim.x) {
        const int64_t c0 = p.sess_claim_off[s], c1 = p.sess_claim_off[s + 1];
        const int64_t f0 = p.sess_fact_off[s], f1 = p.sess_fact_off[s + 1];
        for (int64_t cg = c0; cg < c1; cg += BIND_CLAIMS) {
            const int ncg = (int)min((int64_t)BIND_CLAIMS, c1 - cg);
            // ---- clear table
            for (uint32_t i = tid; i < p.cap; i += BIND_THREADS) { tkeys[i] = 0ULL; tcmask[i] = 0u; }
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
                EmitClaim emit{ tkeys, tcmask, mask, sh.lenA, c };
                dv_tokenize_v3_warp(p.claim_utf8 + b, (int32_t)(e - b), &sh.win[warp], emit);
            }
            __syncthreads();
            double best = 0.0; int best_f = -1;                 // thread c (< ncg) owns claim c
            // ---- fact batches
            for (int64_t fb = f0; fb < f1; fb += BIND_FACTS) {
                const int nfb = (int)min((int64_t)BIND_FACTS, f1 - fb);
                for (uint32_t i = tid; i < p.cap; i += BIND_THREADS) tseen[i] = 0ULL;
                for (int i = tid; i < BIND_FACTS * BIND_CLAIMS; i += BIND_THREADS) sh.ov[i] = 0u;
                if (tid == 0) sh.next_item = 0;
                __syncthreads();
                // ---- phase F
                for (;;) {
                    int fl = 0;
                    if (lane == 0) fl = atomicAdd(&sh.next_item, 1);
                    fl = __shfl_sync(0xffffffffu, fl, 0);
                    if (fl >= nfb) break;
                    sh.wset[warp][lane] = 0ULL; sh.wset[warp][lane + 32] = 0ULL;
                    if (lane == 0) sh.wcnt[warp] = 0;
                    __syncwarp();
                    const int64_t b = p.fact_off[fb + fl], e = p.fact_off[fb + fl + 1];
                    EmitFact emit{ tkeys, tcmask, tseen, mask, &sh.ov[fl * BIND_CLAIMS], fl, sh.wset[warp], &sh.wcnt[warp], false };
                    dv_tokenize_v3_warp(p.fact_utf8 + b, (int32_t)(e - b), &sh.win[warp], emit);
                    __syncwarp();
                    if (lane == 0) {
                        const int n = min(sh.wcnt[warp], 18);
                        sh.lenB[fl] = n;
                        if (p.len_fact && cg == c0) p.len_fact[fb + fl] = n;
                    }
                }
                __syncthreads();
                // ---- phase S: scores for (fact, claim) pairs of this batch
                for (int i = tid; i < nfb * BIND_CLAIMS; i += BIND_THREADS) {
                    const int fl = i / BIND_CLAIMS, c = i % BIND_CLAIMS;
                    double sc = -1.0;
                    const uint32_t ov = sh.ov[i];
                    if (c < ncg && ov > 0u) {
                        const int la = sh.lenA[c], lb = sh.lenB[fl];
                        const int da = max(3, min(la, 14)), db = max(4, min(lb, 18));
                        const double coverage = __ddiv_rn((double)ov, (double)da);
                        const double precision = __ddiv_rn((double)ov, (double)db);
                        double q = p.fact_quality[fb + fl];
                        q = q > 0.0 ? (q < 1.0 ? q : 1.0) : 0.0;
                        const int pref = p.claim_pref_dim[cg + c];
                        const double dimb = (pref >= 0 && p.fact_dim[fb + fl] == pref) ? 0.12 : 0.0;
                        const double hardb = p.fact_hard[fb + fl] ? 0.04 : 0.0;
                        sc = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(0.62, coverage), __dmul_rn(0.18, precision)),
                                                           __dmul_rn(0.16, q)), dimb), hardb);
                    }
                    scores[i] = sc;
                }
                __syncthreads();
                if (tid < ncg) {
                    for (int fl = 0; fl < nfb; fl++) {
                        const double sc = scores[fl * BIND_CLAIMS + tid];
                        if (sc >= 0.0 && (best_f < 0 || sc > best)) { best = sc; best_f = (int)(fb - f0) + fl; }
                    }
                }
                __syncthreads();
            }
            if (tid < ncg) {
                const int64_t c = cg + tid;
                p.best_fact[c] = best_f;
                p.best_score[c] = best_f >= 0 ? best : 0.0;
                p.bound[c] = (best_f >= 0 && !(__dadd_rn(best, 1e-9) < p.claim_threshold[c])) ? 1 : 0;
                if (p.len_claim) p.len_claim[c] = min(sh.lenA[tid], 14);
            }
            __syncthreads();
        }
    }
}

static inline uint32_t bind_table_cap(int64_t max_group_bytes) {
    // distinct tokens of a text <= 0.45 * bytes + 1 (bigrams + pieces of a CJK run); load factor <= 2/3
    int64_t bound = (max_group_bytes * 9) / 20 + BIND_CLAIMS + 1;
    int64_t want = bound + bound / 2;
    uint32_t cap = 2048; while ((int64_t)cap < want) cap <<= 1;
    return cap;
}

extern "C" int32_t dvb_bind_claims(dvb_ctx* ctx, int32_t mem,
                                    const uint8_t* claim_utf8, const int64_t* claim_off,
                                    const uint8_t* fact_utf8, const int64_t* fact_off,
                                    const int64_t* sess_claim_off, const int64_t* sess_fact_off, int64_t n_sessions,
                                    const double* fact_quality, const int32_t* fact_dim, const uint8_t* fact_hard,
                                    const int32_t* claim_pref_dim, const double* claim_threshold,
                                    int32_t* out_best_fact, double* out_best_score, uint8_t* out_bound,
                                    int32_t* out_len_claim, int32_t* out_len_fact, int64_t max_claim_group_bytes) {
    if (!ctx || n_sessions < 0 || !claim_off || !fact_off || !sess_claim_off || !sess_fact_off ||
        !out_best_fact || !out_best_score || !out_bound) { dvb_set_err(ctx, "dvb_bind_claims: invalid argument%s%s"); return DVB_E_INVALID; }
    if (n_sessions == 0) return DVB_OK;
    dvb_scope sc(ctx);
    cudaStream_t st = sc.stream();
    // ---- host view of the session/claim offsets (sizes; table capacity)
    std::vector<int64_t> h_sco, h_sfo, h_co, h_fo;
    int64_t n_claims = 0, n_facts = 0, claim_bytes = 0, fact_bytes = 0, co0 = 0, fo0 = 0;
    if (mem == DVB_MEM_HOST) {
        n_claims = sess_claim_off[n_sessions]; n_facts = sess_fact_off[n_sessions];
        co0 = claim_off[0]; fo0 = fact_off[0];
        claim_bytes = claim_off[n_claims] - co0; fact_bytes = fact_off[n_facts] - fo0;
    } else {
        int64_t tmp[2];
        DVB_CUDA(ctx, cudaMemcpyAsync(&tmp[0], sess_claim_off + n_sessions, 8, cudaMemcpyDeviceToHost, st));
        DVB_CUDA(ctx, cudaMemcpyAsync(&tmp[1], sess_fact_off + n_sessions, 8, cudaMemcpyDeviceToHost, st));
        DVB_CUDA(ctx, cudaStreamSynchronize(st));
        n_claims = tmp[0]; n_facts = tmp[1];
    }
    if (n_claims < 0 || n_facts < 0) { dvb_set_err(ctx, "dvb_bind_claims: bad session offsets%s%s"); return DVB_E_INVALID; }
    if (max_claim_group_bytes <= 0) {
        const int64_t* sco = sess_claim_off; const int64_t* co = claim_off;
        if (mem == DVB_MEM_DEVICE) {
            h_sco.resize((size_t)n_sessions + 1); h_co.resize((size_t)n_claims + 1);
            DVB_CUDA(ctx, cudaMemcpyAsync(h_sco.data(), sess_claim_off, 8 * (size_t)(n_sessions + 1), cudaMemcpyDeviceToHost, st));
            DVB_CUDA(ctx, cudaMemcpyAsync(h_co.data(), claim_off, 8 * (size_t)(n_claims + 1), cudaMemcpyDeviceToHost, st));
            DVB_CUDA(ctx, cudaStreamSynchronize(st));
            sco = h_sco.data(); co = h_co.data();
        }
        int64_t mx = 0;
        for (int64_t s = 0; s < n_sessions; s++)
            for (int64_t c = sco[s]; c < sco[s + 1]; c += BIND_CLAIMS) {
                int64_t e = std::min<int64_t>(c + BIND_CLAIMS, sco[s + 1]);
                mx = std::max(mx, co[e] - co[c]);
            }
        max_claim_group_bytes = mx;
    }
    const uint32_t cap = bind_table_cap(max_claim_group_bytes);
    const size_t fixed = (sizeof(BindShared) + 15) & ~size_t(15);
    size_t smem = fixed + (size_t)cap * 20;
    bool use_global = smem > ctx->smem_optin;
    if (use_global) smem = fixed;
    DVB_CUDA(ctx, cudaFuncSetAttribute(k_bind_claims, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 1;
    DVB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_bind_claims, BIND_THREADS, smem));
    if (occ < 1) occ = 1;
    const int grid = (int)std::min<int64_t>(n_sessions, (int64_t)ctx->num_sm * occ);
    // ---- workspace
    size_t need = 8192;
    if (mem == DVB_MEM_HOST) {
        need += dvb_pad((size_t)claim_bytes) + dvb_pad((size_t)fact_bytes) + dvb_pad(8 * (size_t)(n_claims + 1)) + dvb_pad(8 * (size_t)(n_facts + 1))
              + 2 * dvb_pad(8 * (size_t)(n_sessions + 1)) + dvb_pad(8 * (size_t)n_facts) + dvb_pad(4 * (size_t)n_facts) + dvb_pad((size_t)n_facts)
              + dvb_pad(4 * (size_t)n_claims) * 3 + dvb_pad(8 * (size_t)n_claims) * 2 + dvb_pad((size_t)n_claims) + dvb_pad(4 * (size_t)n_facts);
    }
    if (use_global) need += dvb_pad((size_t)grid * (size_t)cap * 20);
    int32_t rc = sc.reserve(need); if (rc) return rc;
    BindParams p; memset(&p, 0, sizeof(p));
    p.n_sessions = n_sessions; p.cap = cap;
    cudaError_t cerr = cudaSuccess;
    int32_t *d_best = nullptr, *d_lc = nullptr, *d_lf = nullptr; double* d_score = nullptr; uint8_t* d_bound = nullptr;
    std::vector<int64_t> reb_c, reb_f;
    if (mem == DVB_MEM_HOST) {
        uint8_t* d_cu = sc.alloc<uint8_t>((size_t)claim_bytes + 64); uint8_t* d_fu = sc.alloc<uint8_t>((size_t)fact_bytes + 64);
        int64_t* d_co = sc.alloc<int64_t>((size_t)n_claims + 1); int64_t* d_fo = sc.alloc<int64_t>((size_t)n_facts + 1);
        int64_t* d_sco = sc.alloc<int64_t>((size_t)n_sessions + 1); int64_t* d_sfo = sc.alloc<int64_t>((size_t)n_sessions + 1);
        double* d_fq = sc.alloc<double>((size_t)n_facts); int32_t* d_fd = sc.alloc<int32_t>((size_t)n_facts); uint8_t* d_fh = sc.alloc<uint8_t>((size_t)n_facts);
        int32_t* d_cp = sc.alloc<int32_t>((size_t)n_claims); double* d_ct = sc.alloc<double>((size_t)n_claims);
        d_best = sc.alloc<int32_t>((size_t)n_claims); d_score = sc.alloc<double>((size_t)n_claims); d_bound = sc.alloc<uint8_t>((size_t)n_claims);
        d_lc = sc.alloc<int32_t>((size_t)n_claims); d_lf = sc.alloc<int32_t>((size_t)n_facts);
        if (!d_cu || !d_fu || !d_co || !d_fo || !d_sco || !d_sfo || !d_fq || !d_fd || !d_fh || !d_cp || !d_ct || !d_best || !d_score || !d_bound || !d_lc || !d_lf) {
            dvb_set_err(ctx, "workspace layout error%s%s"); return DVB_E_NOMEM; }
        const int64_t* co = claim_off; const int64_t* fo = fact_off;
        if (co0 != 0) { reb_c.resize((size_t)n_claims + 1); for (int64_t i = 0; i <= n_claims; i++) reb_c[i] = claim_off[i] - co0; co = reb_c.data(); }
        if (fo0 != 0) { reb_f.resize((size_t)n_facts + 1); for (int64_t i = 0; i <= n_facts; i++) reb_f[i] = fact_off[i] - fo0; fo = reb_f.data(); }
        p.claim_utf8 = dvb_in(mem, claim_utf8 + co0, d_cu, (size_t)claim_bytes, st, &cerr);
        p.fact_utf8 = dvb_in(mem, fact_utf8 + fo0, d_fu, (size_t)fact_bytes, st, &cerr);
        p.claim_off = dvb_in(mem, co, d_co, (size_t)n_claims + 1, st, &cerr);
        p.fact_off = dvb_in(mem, fo, d_fo, (size_t)n_facts + 1, st, &cerr);
        p.sess_claim_off = dvb_in(mem, sess_claim_off, d_sco, (size_t)n_sessions + 1, st, &cerr);
        p.sess_fact_off = dvb_in(mem, sess_fact_off, d_sfo, (size_t)n_sessions + 1, st, &cerr);
        p.fact_quality = dvb_in(mem, fact_quality, d_fq, (size_t)n_facts, st, &cerr);
        p.fact_dim = dvb_in(mem, fact_dim, d_fd, (size_t)n_facts, st, &cerr);
        p.fact_hard = dvb_in(mem, fact_hard, d_fh, (size_t)n_facts, st, &cerr);
        p.claim_pref_dim = dvb_in(mem, claim_pref_dim, d_cp, (size_t)n_claims, st, &cerr);
        p.claim_threshold = dvb_in(mem, claim_threshold, d_ct, (size_t)n_claims, st, &cerr);
        DVB_CUDA(ctx, cerr);
        p.best_fact = d_best; p.best_score = d_score; p.bound = d_bound; p.len_claim = out_len_claim ? d_lc : nullptr; p.len_fact = out_len_fact ? d_lf : nullptr;
    } else {
        p.claim_utf8 = claim_utf8; p.fact_utf8 = fact_utf8; p.claim_off = claim_off; p.fact_off = fact_off;
        p.sess_claim_off = sess_claim_off; p.sess_fact_off = sess_fact_off;
        p.fact_quality = fact_quality; p.fact_dim = fact_dim; p.fact_hard = fact_hard;
        p.claim_pref_dim = claim_pref_dim; p.claim_threshold = claim_threshold;
        p.best_fact = out_best_fact; p.best_score = out_best_score; p.bound = out_bound; p.len_claim = out_len_claim; p.len_fact = out_len_fact;
    }
    if (p.len_fact && n_facts) DVB_CUDA(ctx, cudaMemsetAsync(p.len_fact, 0, 4 * (size_t)n_facts, st));
    if (use_global) { p.gtable = sc.alloc<unsigned long long>(((size_t)grid * (size_t)cap * 20 + 7) / 8); if (!p.gtable) { dvb_set_err(ctx, "workspace layout error%s%s"); return DVB_E_NOMEM; } }
    const bool timing = ctx->timing.load() != 0;
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev0, st));
    k_bind_claims<<<grid, BIND_THREADS, smem, st>>>(p);
    if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev1, st));
    ctx->launches += 1;
    DVB_CUDA(ctx, cudaGetLastError());
    if (mem == DVB_MEM_HOST && n_claims) {
        DVB_CUDA(ctx, cudaMemcpyAsync(out_best_fact, d_best, 4 * (size_t)n_claims, cudaMemcpyDeviceToHost, st));
        DVB_CUDA(ctx, cudaMemcpyAsync(out_best_score, d_score, 8 * (size_t)n_claims, cudaMemcpyDeviceToHost, st));
        DVB_CUDA(ctx, cudaMemcpyAsync(out_bound, d_bound, (size_t)n_claims, cudaMemcpyDeviceToHost, st));
        if (out_len_claim) DVB_CUDA(ctx, cudaMemcpyAsync(out_len_claim, d_lc, 4 * (size_t)n_claims, cudaMemcpyDeviceToHost, st));
        if (out_len_fact && n_facts) DVB_CUDA(ctx, cudaMemcpyAsync(out_len_fact, d_lf, 4 * (size_t)n_facts, cudaMemcpyDeviceToHost, st));
    }
    DVB_CUDA(ctx, cudaStreamSynchronize(st));
    if (timing) { float ms = 0.f; DVB_CUDA(ctx, cudaEventElapsedTime(&ms, sc.l->ev0, sc.l->ev1)); ctx->last_ms = ms; }
    return DVB_OK;
}
