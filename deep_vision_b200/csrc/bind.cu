// bind.cu — dvb_bind_claims: fused tokenise + set-overlap + fp64 score + first-max argmax.
// replaces: the scoring loop of infer_weak_evidence_refs_v3 (reference web/server.py:24813-24864)
//
// One CTA per session (persistent grid, sessions strided).  Per group of <=32 claims:
//   phase C  warps tokenise the claims (text.cuh) and insert every token key into an exact
//            open-addressing table  key -> bitmask of claims  that lives in an L2-resident
//            per-CTA global workspace; two shared-memory bitmap filters (bigram ids, token keys)
//            record which keys exist.  len(A) is counted on first insert.
//   phase F  warps stream facts with coalesced 16-byte loads; a fact token is looked up in the
//            table only if it passes the filter (~1 % false positives), 32 survivors at a time;
//            hits set a bit in a per-warp bitmap over table slots, which makes the overlap a SET
//            overlap; a 64-slot per-warp set gives min(len(B), 18) exactly.
//            At the end of each fact, lane c scores claim c in fp64 in the reference's operation
//            order (no FMA contraction) and keeps its running best under the total order
//            (score desc, fact index asc)  ==  the reference's strict '>' scan in index order.
//   phase R  per-claim reduction of the warps' bests, outputs.
// No intermediate key arrays ever reach HBM: algorithmic traffic = claim + fact bytes in, 16 B/claim out.
#include <atomic>
#include "common.cuh"
#include "text.cuh"
#include "bind_common.cuh"
#include "bind_general.cuh"
#include "bind7.cuh"

static inline uint32_t bind_table_cap(int64_t max_group_bytes) {
    // distinct tokens of a text <= bytes/3 + 1 (<= one token per CJK char, <= one per 3 ASCII bytes); load factor <= 1/2
    int64_t bound = max_group_bytes / 3 + BIND_CLAIMS + 1;
    int64_t want = bound * 2;
    uint32_t cap = 1024; while ((int64_t)cap < want) cap <<= 1;
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
    dvb_trace tr("DVB_BIND_TRACE", "bind");
    dvb_scope sc(ctx);
    cudaStream_t st = sc.stream();
    tr.at("scope");
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
    const size_t hit_bytes = (size_t)BIND_WARPS * (cap >> 5) * 4;
    size_t smem = fixed + hit_bytes;
    const bool use_global = hit_bytes > 16384;          // huge claim groups: hit bitmaps move to global memory
    if (use_global) smem = fixed;
    // the general kernel runs only for handed-back sessions: its attribute and occupancy are queried when it is launched; the
    // workspace is sized for the most CTAs an SM can hold (8 x 256 threads).  A config-1 sized call is ~25 driver calls long,
    // so the ones that do not depend on the call are made once per device.
    int grid = (int)std::min<int64_t>(n_sessions, (int64_t)ctx->num_sm * 8);
    // ---- fast kernel (bind7.cuh): per-session capacity checks happen on the device (sess_fallback)
    const char* force = getenv("DVB_BIND_KERNEL");
    const bool fast = !(force && !strcmp(force, "general"));
    const size_t smem7 = bind7_smem_bytes(), smem7c = bind7_smem_bytes_claims();
    int grid7 = 1, grid7c = 1, grid7_max = 1;                      // fact kernel (3 CTAs per SM), claim kernel (4); resident fact CTAs of the device
    if (fast) {
        static std::atomic<int> occ7_of[64];                       // per device: 0 = not asked yet; fact occupancy | claim occupancy << 8
        const int dslot = ctx->device & 63;
        int occ7 = occ7_of[dslot].load();
        if (occ7 <= 0) {
            int of = 1, oc = 1;
            DVB_CUDA(ctx, cudaFuncSetAttribute(k_bind_claim_tables, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem7c));
            DVB_CUDA(ctx, cudaFuncSetAttribute(k_bind_facts, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem7));
            DVB_CUDA(ctx, cudaFuncSetAttribute(k_bind_facts_split, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem7));
            DVB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&of, k_bind_facts, BIND_THREADS, smem7));
            DVB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&oc, k_bind_claim_tables, BIND_THREADS, smem7c));
            occ7 = std::max(1, std::min(of, 255)) | (std::max(1, std::min(oc, 255)) << 8);
            occ7_of[dslot].store(occ7);
        }
        grid7c = (int)std::min<int64_t>(n_sessions + n_claims / BIND_CLAIMS, (int64_t)ctx->num_sm * (occ7 >> 8));   // CTAs take claim groups
        occ7 &= 255;
        grid7_max = ctx->num_sm * occ7;
        grid7 = (int)std::min<int64_t>(n_sessions + n_claims / BIND_CLAIMS, (int64_t)grid7_max);
    }
    // ---- workspace
    size_t need = 8192;
    if (mem == DVB_MEM_HOST) {
        need += dvb_pad((size_t)claim_bytes) + dvb_pad((size_t)fact_bytes) + dvb_pad(8 * (size_t)(n_claims + 1)) + dvb_pad(8 * (size_t)(n_facts + 1))
              + 2 * dvb_pad(8 * (size_t)(n_sessions + 1)) + dvb_pad(8 * (size_t)n_facts) + dvb_pad(4 * (size_t)n_facts) + dvb_pad((size_t)n_facts)
              + dvb_pad(4 * (size_t)n_claims) * 3 + dvb_pad(8 * (size_t)n_claims) * 2 + dvb_pad((size_t)n_claims) + dvb_pad(4 * (size_t)n_facts);
    }
    need += dvb_pad((size_t)grid * (size_t)cap * 12);
    if (use_global) need += dvb_pad((size_t)grid * hit_bytes);
    // fast kernels: one slab per claim group of a wave of sessions.  A session of n claims has ceil(n / 32) groups, so a
    // range of sessions holds at most sessions + claims / 32 of them; ranges are cut so that a wave's slabs stay below 1 GB.
    int64_t slab_budget = std::max<int64_t>(((int64_t)1 << 30) / B7_BLOB_BYTES, 64);
    if (const char* e = getenv("DVB_BIND_SLAB_BUDGET")) { const long long v = atoll(e); if (v >= 1) slab_budget = v; }       // tests: force waves on small inputs
    const int64_t groups_bound = n_sessions + n_claims / BIND_CLAIMS;
    const int64_t n_slabs = fast ? std::min(groups_bound, slab_budget) : 0;
    const int ws_ctas = std::max(std::max(grid7, grid7c), (int)std::min<int64_t>(grid7_max, (n_sessions + n_claims / BIND_CLAIMS) * BIND_MAX_PARTS));   // split fact kernel: groups x parts CTAs
    if (fast) need += dvb_pad((size_t)ws_ctas * bind7_ws_u64_per_cta() * 8) + dvb_pad((size_t)n_sessions) + dvb_pad(4 * (size_t)n_sessions) + dvb_pad(64)
                    + dvb_pad(4 * (size_t)n_slabs) * 2 + dvb_pad((size_t)grid7_max * BIND_CLAIMS * 8) + dvb_pad((size_t)grid7_max * BIND_CLAIMS * 4)
                    + dvb_pad((size_t)n_slabs * B7_BLOB_BYTES);
    tr.at("sizes, occupancy");
    int32_t rc = sc.reserve(need); if (rc) return rc;
    BindParams p; memset(&p, 0, sizeof(p));
    p.s_begin = 0; p.s_end = n_sessions; p.cap = cap;
    cudaError_t cerr = cudaSuccess;
    int32_t *d_best = nullptr, *d_lc = nullptr, *d_lf = nullptr; double* d_score = nullptr; uint8_t* d_bound = nullptr;
    std::vector<int64_t> reb_c, reb_f;
    const uint8_t *h_cu = nullptr, *h_fu = nullptr; uint8_t *d_cu_w = nullptr, *d_fu_w = nullptr;
    const int64_t *h_co_rel = nullptr, *h_fo_rel = nullptr;
    uint8_t* small_h = nullptr;                               // non-null: the call's inputs and outputs go through the pinned block
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
        p.claim_utf8 = d_cu; p.fact_utf8 = d_fu;          // text is copied slice by slice below
        h_cu = claim_utf8 + co0; h_fu = fact_utf8 + fo0; d_cu_w = d_cu; d_fu_w = d_fu; h_co_rel = co; h_fo_rel = fo;
        // A small call (one report: the reference's own call shape) is a dozen driver copies of a few hundred bytes each, ~6 us
        // apiece from pageable memory.  The inputs are consecutive in the workspace (d_cu .. d_ct) and so are the outputs
        // (d_best .. d_lf): they cross PCIe as ONE copy each through the lease's pinned block.
        const size_t in_span = (size_t)(reinterpret_cast<uint8_t*>(d_ct + n_claims) - d_cu);
        const size_t out_span = (size_t)(reinterpret_cast<uint8_t*>(d_lf + n_facts) - reinterpret_cast<uint8_t*>(d_best));
        if (reinterpret_cast<uint8_t*>(d_ct) > d_cu && reinterpret_cast<uint8_t*>(d_lf) > reinterpret_cast<uint8_t*>(d_best)
            && in_span <= DVB_SMALL_STAGE && out_span <= DVB_SMALL_STAGE && sc.in_arena(d_cu, in_span) && sc.in_arena(d_best, out_span)
            && !getenv("DVB_NO_SMALL_STAGE")) {
            dvb_lease* l = sc.l;
            if (!l->small_h && cudaHostAlloc((void**)&l->small_h, 2 * DVB_SMALL_STAGE, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); l->small_h = nullptr; }
            small_h = l->small_h;
        }
        if (small_h) {
            auto put = [&](const void* dev, const void* host, size_t bytes) { if (bytes) memcpy(small_h + (reinterpret_cast<const uint8_t*>(dev) - d_cu), host, bytes); };
            put(d_cu, h_cu, (size_t)claim_bytes); put(d_fu, h_fu, (size_t)fact_bytes);
            put(d_co, co, 8 * ((size_t)n_claims + 1)); put(d_fo, fo, 8 * ((size_t)n_facts + 1));
            put(d_sco, sess_claim_off, 8 * ((size_t)n_sessions + 1)); put(d_sfo, sess_fact_off, 8 * ((size_t)n_sessions + 1));
            put(d_fq, fact_quality, 8 * (size_t)n_facts); put(d_fd, fact_dim, 4 * (size_t)n_facts); put(d_fh, fact_hard, (size_t)n_facts);
            put(d_cp, claim_pref_dim, 4 * (size_t)n_claims); put(d_ct, claim_threshold, 8 * (size_t)n_claims);
            DVB_CUDA(ctx, cudaMemcpyAsync(d_cu, small_h, in_span, cudaMemcpyHostToDevice, st));
            p.claim_off = d_co; p.fact_off = d_fo; p.sess_claim_off = d_sco; p.sess_fact_off = d_sfo; p.fact_quality = d_fq; p.fact_dim = d_fd;
            p.fact_hard = d_fh; p.claim_pref_dim = d_cp; p.claim_threshold = d_ct;
        } else {
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
        }
        p.best_fact = d_best; p.best_score = d_score; p.bound = d_bound; p.len_claim = out_len_claim ? d_lc : nullptr; p.len_fact = out_len_fact ? d_lf : nullptr;
    } else {
        p.claim_utf8 = claim_utf8; p.fact_utf8 = fact_utf8; p.claim_off = claim_off; p.fact_off = fact_off;
        p.sess_claim_off = sess_claim_off; p.sess_fact_off = sess_fact_off;
        p.fact_quality = fact_quality; p.fact_dim = fact_dim; p.fact_hard = fact_hard;
        p.claim_pref_dim = claim_pref_dim; p.claim_threshold = claim_threshold;
        p.best_fact = out_best_fact; p.best_score = out_best_score; p.bound = out_bound; p.len_claim = out_len_claim; p.len_fact = out_len_fact;
    }
    tr.at("inputs staged (H2D queued)");
    if (p.len_fact && n_facts) DVB_CUDA(ctx, cudaMemsetAsync(p.len_fact, 0, 4 * (size_t)n_facts, st));
    p.gtable = sc.alloc<unsigned long long>(((size_t)grid * (size_t)cap * 12 + 7) / 8);
    if (!p.gtable) { dvb_set_err(ctx, "workspace layout error%s%s"); return DVB_E_NOMEM; }
    if (use_global) { p.ghit = sc.alloc<uint32_t>((size_t)grid * hit_bytes / 4); if (!p.ghit) { dvb_set_err(ctx, "workspace layout error%s%s"); return DVB_E_NOMEM; } }
    BindParams p7 = p;
    if (fast) {
        dvb_lease* l = sc.l;
        if (!l->flag_h) {                                     // overflow flag: pinned + mapped, the kernel writes it in place
            DVB_CUDA(ctx, cudaHostAlloc((void**)&l->flag_h, 64, cudaHostAllocMapped));
            DVB_CUDA(ctx, cudaHostGetDevicePointer((void**)&l->flag_d, l->flag_h, 0));
        }
        *(volatile int*)l->flag_h = 0;
        p7.gtable = sc.alloc<unsigned long long>((size_t)ws_ctas * bind7_ws_u64_per_cta());
        p7.sess_fallback = sc.alloc<uint8_t>((size_t)n_sessions);
        p7.sess_slab = sc.alloc<int32_t>((size_t)n_sessions);
        p7.slab_cursor = sc.alloc<int>(16);
        p7.slab_sess = sc.alloc<int32_t>((size_t)n_slabs);
        p7.part_cnt = sc.alloc<int>((size_t)n_slabs);
        p7.part_score = sc.alloc<double>((size_t)grid7_max * BIND_CLAIMS);          // groups x parts <= the fact kernel's resident CTAs when it splits
        p7.part_idx = sc.alloc<int>((size_t)grid7_max * BIND_CLAIMS);
        p7.slabs = sc.alloc<uint8_t>((size_t)n_slabs * B7_BLOB_BYTES);
        if (!p7.gtable || !p7.sess_fallback || !p7.sess_slab || !p7.slab_cursor || !p7.slab_sess || !p7.part_cnt || !p7.part_score || !p7.part_idx || !p7.slabs) { dvb_set_err(ctx, "workspace layout error%s%s"); return DVB_E_NOMEM; }
        DVB_CUDA(ctx, cudaMemsetAsync(p7.sess_fallback, 0, (size_t)n_sessions, st));
        p7.overflow = l->flag_d;
    }
    // sessions [s0, s1): the fast pair (claim tables, then facts; in waves whose claim groups fit the slabs) unless the call
    // was routed to the general kernel
    std::vector<int64_t> wave_sco;                            // session -> first claim, on the host, only when waves have to be cut
    auto launch = [&](bool use_fast, int64_t s0, int64_t s1) -> cudaError_t {
        BindParams q = use_fast ? p7 : p;
        if (!use_fast) {
            q.s_begin = s0; q.s_end = s1;
            cudaError_t ge = cudaFuncSetAttribute(k_bind_claims_general, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            int occ = 1;
            if (ge == cudaSuccess) ge = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_bind_claims_general, BIND_THREADS, smem);
            if (ge != cudaSuccess) return ge;
            grid = (int)std::min<int64_t>((int64_t)grid, (int64_t)ctx->num_sm * std::max(1, occ));      // never above the workspace's CTA count
            k_bind_claims_general<<<(int)std::min<int64_t>(s1 - s0, (int64_t)grid), BIND_THREADS, smem, st>>>(q);
            ctx->launches += 1;
            return cudaGetLastError();
        }
        const int64_t* sco_h = mem == DVB_MEM_HOST ? sess_claim_off : nullptr;
        if (groups_bound > n_slabs && !sco_h) {               // device offsets and more groups than slabs: fetch the offsets once
            if (wave_sco.empty()) {
                wave_sco.resize((size_t)n_sessions + 1);
                cudaError_t e = cudaMemcpyAsync(wave_sco.data(), sess_claim_off, 8 * (size_t)(n_sessions + 1), cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess) e = cudaStreamSynchronize(st);
                if (e != cudaSuccess) return e;
            }
            sco_h = wave_sco.data();
        }
        for (int64_t a = s0; a < s1;) {
            int64_t b = s1;
            int64_t units = std::min<int64_t>(groups_bound, n_slabs);   // claim groups of [a, b): exact with host offsets, else the bound
            if (groups_bound > n_slabs) {                     // longest prefix of sessions whose groups fit
                int64_t g = 0; b = a;
                while (b < s1) {
                    const int64_t gs = (sco_h[b + 1] - sco_h[b] + BIND_CLAIMS - 1) / BIND_CLAIMS;
                    if (g + gs > n_slabs) break;
                    g += gs; b++;
                }
                if (b == a) return cudaErrorInvalidValue;     // one session with more groups than a wave holds (> 1 GB of tables)
                units = g;
            } else if (sco_h) {
                units = 0;
                for (int64_t x = a; x < b; x++) units += (sco_h[x + 1] - sco_h[x] + BIND_CLAIMS - 1) / BIND_CLAIMS;
            }
            q.s_begin = a; q.s_end = b;
            if (units > 0) {
                k_bind_groups<<<1, B7_GROUPS_THREADS, 0, st>>>(q);
                k_bind_claim_tables<<<(int)std::min<int64_t>(units, (int64_t)grid7c), BIND_THREADS, smem7c, st>>>(q);
                // fewer groups than half the resident CTAs (one report per call): the facts of every group are cut into `parts` ranges
                const int parts = (int)std::min<int64_t>(BIND_MAX_PARTS, (int64_t)grid7_max / units);
                if (parts >= 2 && !getenv("DVB_BIND_NO_SPLIT")) {
                    q.parts = parts;
                    k_bind_facts_split<<<(int)(units * parts), BIND_THREADS, smem7, st>>>(q);
                } else k_bind_facts<<<(int)std::min<int64_t>(units, (int64_t)grid7), BIND_THREADS, smem7, st>>>(q);
                ctx->launches += 3;
                const cudaError_t e = cudaGetLastError();
                if (e != cudaSuccess) return e;
            }
            a = b;
        }
        return cudaSuccess;
    };
    const bool timing = ctx->timing.load() != 0;
    if (mem == DVB_MEM_DEVICE) {
        if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev0, st));
        DVB_CUDA(ctx, launch(fast, 0, n_sessions));
        if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev1, st));
    } else {
        // Host buffers: cut the sessions into slices of ~1/24 of the fact text (8..64 MB); slice i+1 crosses PCIe on
        // the copy stream while slice i is scored, so a call costs ~max(H2D, kernel) instead of their sum.
        const int64_t target = std::min<int64_t>(std::max<int64_t>(fact_bytes / 24, (int64_t)8 << 20), (int64_t)64 << 20);
        std::vector<int64_t> cuts(1, 0);
        for (int64_t sidx = 1; sidx < n_sessions; sidx++)
            if (h_fo_rel[sess_fact_off[sidx]] - h_fo_rel[sess_fact_off[cuts.back()]] >= target) cuts.push_back(sidx);
        cuts.push_back(n_sessions);
        const size_t n_slices = cuts.size() - 1;
        dvb_lease* l = sc.l;
        cudaStream_t cp = st;                                              // one slice: everything on the call's stream
        if (n_slices > 1) {
            if (!l->copy) DVB_CUDA(ctx, cudaStreamCreateWithFlags(&l->copy, cudaStreamNonBlocking));
            while (l->pipe_ev.size() < n_slices + 1) {
                cudaEvent_t e; DVB_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); l->pipe_ev.push_back(e);
            }
            cp = l->copy;
        }
        const bool facts_pageable = dvb_is_pageable(h_fu);
        auto run = [&]() -> cudaError_t {
            cudaError_t e = cudaSuccess;
            #define BIND_TRY(x) do { e = (x); if (e != cudaSuccess) return e; } while (0)
            if (n_slices > 1) { BIND_TRY(cudaEventRecord(l->pipe_ev[0], st)); BIND_TRY(cudaStreamWaitEvent(cp, l->pipe_ev[0], 0)); }
            if (timing) BIND_TRY(cudaEventRecord(l->ev0, st));
            for (size_t i = 0; i < n_slices; i++) {
                const int64_t s0 = cuts[i], s1 = cuts[i + 1];
                const int64_t cb0 = h_co_rel[sess_claim_off[s0]], cb1 = h_co_rel[sess_claim_off[s1]];
                const int64_t fb0 = h_fo_rel[sess_fact_off[s0]], fb1 = h_fo_rel[sess_fact_off[s1]];
                if (small_h) { BIND_TRY(launch(fast, s0, s1)); continue; }        // text went with the one copy above
                if (cb1 > cb0) BIND_TRY(cudaMemcpyAsync(d_cu_w + cb0, h_cu + cb0, (size_t)(cb1 - cb0), cudaMemcpyHostToDevice, cp));
                // fact text of a slice from PAGEABLE memory (numpy behind the ctypes shim): 4 host threads stage it through pinned buffers
                if (fb1 > fb0 && (size_t)(fb1 - fb0) >= DVB_H2D_MIN && facts_pageable) BIND_TRY(dvb_h2d_staged(l, ctx->device, d_fu_w + fb0, h_fu + fb0, (size_t)(fb1 - fb0), cp));
                else if (fb1 > fb0) BIND_TRY(cudaMemcpyAsync(d_fu_w + fb0, h_fu + fb0, (size_t)(fb1 - fb0), cudaMemcpyHostToDevice, cp));
                if (n_slices > 1) { BIND_TRY(cudaEventRecord(l->pipe_ev[i + 1], cp)); BIND_TRY(cudaStreamWaitEvent(st, l->pipe_ev[i + 1], 0)); }
                BIND_TRY(launch(fast, s0, s1));
            }
            if (timing) BIND_TRY(cudaEventRecord(l->ev1, st));
            #undef BIND_TRY
            return e;
        };
        const cudaError_t pe = run();
        if (pe != cudaSuccess) { if (n_slices > 1) cudaStreamSynchronize(cp); cudaStreamSynchronize(st); DVB_CUDA(ctx, pe); }
    }
    auto queue_host_out = [&]() -> cudaError_t {
        cudaError_t e = cudaSuccess;
        if (mem != DVB_MEM_HOST || !n_claims) return e;
        if (small_h) return cudaMemcpyAsync(small_h + DVB_SMALL_STAGE, d_best, (size_t)(reinterpret_cast<uint8_t*>(d_lf + n_facts) - reinterpret_cast<uint8_t*>(d_best)),
                                            cudaMemcpyDeviceToHost, st);
        if ((e = cudaMemcpyAsync(out_best_fact, d_best, 4 * (size_t)n_claims, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        if ((e = cudaMemcpyAsync(out_best_score, d_score, 8 * (size_t)n_claims, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        if ((e = cudaMemcpyAsync(out_bound, d_bound, (size_t)n_claims, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        if (out_len_claim && (e = cudaMemcpyAsync(out_len_claim, d_lc, 4 * (size_t)n_claims, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        if (out_len_fact && n_facts && (e = cudaMemcpyAsync(out_len_fact, d_lf, 4 * (size_t)n_facts, cudaMemcpyDeviceToHost, st)) != cudaSuccess) return e;
        return e;
    };
    tr.at("kernels queued");
    DVB_CUDA(ctx, queue_host_out());
    tr.at("D2H queued");
    DVB_CUDA(ctx, cudaStreamSynchronize(st));
    tr.at("stream synchronised");
    if (fast && *(volatile int*)sc.l->flag_h) {               // sessions the cell walk handed back: exact re-run of exactly those
        p.only_flagged = p7.sess_fallback;
        DVB_CUDA(ctx, launch(false, 0, n_sessions));
        if (timing) DVB_CUDA(ctx, cudaEventRecord(sc.l->ev1, st));
        DVB_CUDA(ctx, queue_host_out());
        DVB_CUDA(ctx, cudaStreamSynchronize(st));
    }
    if (small_h && n_claims) {                                // after the stream sync: pinned block -> the caller's arrays
        const uint8_t* o = small_h + DVB_SMALL_STAGE; const uint8_t* base = reinterpret_cast<const uint8_t*>(d_best);
        auto get = [&](void* host, const void* dev, size_t bytes) { if (host && bytes) memcpy(host, o + (reinterpret_cast<const uint8_t*>(dev) - base), bytes); };
        get(out_best_fact, d_best, 4 * (size_t)n_claims); get(out_best_score, d_score, 8 * (size_t)n_claims); get(out_bound, d_bound, (size_t)n_claims);
        get(out_len_claim, d_lc, 4 * (size_t)n_claims); get(out_len_fact, d_lf, 4 * (size_t)n_facts);
    }
    if (timing) { float ms = 0.f; DVB_CUDA(ctx, cudaEventElapsedTime(&ms, sc.l->ev0, sc.l->ev1)); ctx->last_ms = ms; }
    return DVB_OK;
}
