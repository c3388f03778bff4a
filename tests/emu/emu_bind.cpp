// emu_bind.cpp — runs the PRODUCT source of k_bind_claim_tables + k_bind_facts (bind7.cuh) and k_bind_claims_general (bind_general.cuh) on the
// CPU SIMT emulator, with the launch flow of dvb_bind_claims (fast kernel, then the general kernel over the sessions it
// handed back).  TEST INFRASTRUCTURE: built by tests/emu/emu.py, called by tests/test_emu_bind.py.
#include "simt_emu.h"
#include "../../deep_vision_b200/csrc/bind_general.cuh"
#include "../../deep_vision_b200/csrc/bind7.cuh"

static uint32_t emu_table_cap(int64_t max_group_bytes) {
    int64_t want = (max_group_bytes / 3 + BIND_CLAIMS + 1) * 2;
    uint32_t cap = 1024; while ((int64_t)cap < want) cap <<= 1;
    return cap;
}

// mode: 0 = fast kernel + general re-run of the handed-back sessions (product flow), 1 = fast kernel only, 2 = general only
extern "C" int emu_bind_claims(const uint8_t* claim_utf8, const int64_t* claim_off, const uint8_t* fact_utf8, const int64_t* fact_off,
                               const int64_t* sess_claim_off, const int64_t* sess_fact_off, int64_t n_sessions,
                               const double* fact_quality, const int32_t* fact_dim, const uint8_t* fact_hard,
                               const int32_t* claim_pref_dim, const double* claim_threshold,
                               int32_t* best_fact, double* best_score, uint8_t* bound, int32_t* len_claim, int32_t* len_fact,
                               uint8_t* sess_fallback, int grid, int mode) {
    BindParams p; memset(&p, 0, sizeof(p));
    p.claim_utf8 = claim_utf8; p.claim_off = claim_off; p.fact_utf8 = fact_utf8; p.fact_off = fact_off;
    p.sess_claim_off = sess_claim_off; p.sess_fact_off = sess_fact_off; p.s_begin = 0; p.s_end = n_sessions;
    p.fact_quality = fact_quality; p.fact_dim = fact_dim; p.fact_hard = fact_hard;
    p.claim_pref_dim = claim_pref_dim; p.claim_threshold = claim_threshold;
    p.best_fact = best_fact; p.best_score = best_score; p.bound = bound; p.len_claim = len_claim; p.len_fact = len_fact;
    int overflow = 0;
    p.overflow = &overflow;
    p.sess_fallback = sess_fallback;
    if (n_sessions <= 0) return 0;
    memset(sess_fallback, 0, (size_t)n_sessions);
    if (grid < 1) grid = 1;
    if ((int64_t)grid > n_sessions) grid = (int)n_sessions;
    int64_t mx = 0;
    for (int64_t s = 0; s < n_sessions; s++)
        for (int64_t c = sess_claim_off[s]; c < sess_claim_off[s + 1]; c += BIND_CLAIMS) {
            const int64_t e = std::min<int64_t>(c + BIND_CLAIMS, sess_claim_off[s + 1]);
            mx = std::max(mx, claim_off[e] - claim_off[c]);
        }
    if (mode != 2) {
        std::vector<unsigned long long> ws((size_t)grid * bind7_ws_u64_per_cta());
        p.gtable = ws.data();
        int64_t n_groups = 0;
        for (int64_t s = 0; s < n_sessions; s++) n_groups += (sess_claim_off[s + 1] - sess_claim_off[s] + BIND_CLAIMS - 1) / BIND_CLAIMS;
        std::vector<unsigned long long> slabs(((size_t)n_groups * B7_BLOB_BYTES + 7) / 8 + 1);
        std::vector<int32_t> sess_slab((size_t)n_sessions), slab_sess((size_t)n_groups + 1);
        int cursor = 0;
        p.slabs = reinterpret_cast<uint8_t*>(slabs.data()); p.slab_cursor = &cursor; p.sess_slab = sess_slab.data(); p.slab_sess = slab_sess.data();
        // EMU_BIND_PARTS=n: the split fact kernel with n CTAs per claim group (dvb_bind_claims takes it for launches of few groups)
        const char* pe = getenv("EMU_BIND_PARTS");
        const int parts = pe ? std::max(1, std::min(atoi(pe), BIND_MAX_PARTS)) : 1;
        std::vector<int> part_cnt((size_t)n_groups + 1), part_idx((size_t)(n_groups + 1) * parts * BIND_CLAIMS);
        std::vector<double> part_score((size_t)(n_groups + 1) * parts * BIND_CLAIMS);
        p.part_cnt = part_cnt.data(); p.part_score = part_score.data(); p.part_idx = part_idx.data(); p.parts = parts;
        emu::launch(1u, B7_GROUPS_THREADS, [&] { k_bind_groups(p); });
        emu::launch((unsigned)grid, BIND_THREADS, [&] { k_bind_claim_tables(p); });
        if (getenv("EMU_BIND_VERBOSE")) fprintf(stderr, "emu_bind: %d part(s) per claim group\n", parts);
        if (parts >= 2) emu::launch((unsigned)grid, BIND_THREADS, [&] { k_bind_facts_split(p); });
        else emu::launch((unsigned)grid, BIND_THREADS, [&] { k_bind_facts(p); });
    }
    if (mode == 2 || (mode == 0 && overflow)) {
        const uint32_t cap = emu_table_cap(mx);
        const size_t fixed = (sizeof(BindShared) + 15) & ~size_t(15);
        const size_t hit_bytes = (size_t)BIND_WARPS * (cap >> 5) * 4;
        std::vector<unsigned long long> ws(((size_t)grid * cap * 12 + 7) / 8);
        std::vector<uint32_t> gh;
        p.cap = cap; p.gtable = ws.data(); p.ghit = nullptr;
        if (fixed + hit_bytes > sizeof(bind_smem)) { gh.resize((size_t)grid * hit_bytes / 4); p.ghit = gh.data(); }
        p.only_flagged = mode == 2 ? nullptr : sess_fallback;
        emu::launch((unsigned)grid, BIND_THREADS, [&] { k_bind_claims_general(p); });
    }
    return overflow;
}
