/* TEST INFRASTRUCTURE ONLY (CPU oracle). See oracle/dv_oracle.c.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (deep_vision_b200) never does. */
#ifndef DV_ORACLE_H
#define DV_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- strings ---- */
int64_t dvo_utf8_decode(const uint8_t* s, int64_t n, uint32_t* out);          /* -> #code points */
int64_t dvo_lower(const uint32_t* in, int64_t n, uint32_t* out, int64_t cap); /* CPython str.lower() */
int     dvo_is_space(uint32_t cp);
int     dvo_is_digit(uint32_t cp);

/* ---- a1: _tokenize_similarity_text_v3 (server.py:24722-24732) ----
 * Emits the DISTINCT tokens of one text.  tok_start/tok_len index the lowered
 * code-point buffer `low` (caller-provided, cap >= 3*n_cp).  Returns #tokens or <0. */
int64_t dvo_tokenize_v3(const uint8_t* utf8, int64_t nbytes,
                        uint32_t* low, int64_t low_cap, int64_t* n_low,
                        int32_t* tok_start, int32_t* tok_len, int64_t tok_cap);
/* 64-bit token key of a lowered token (restates include/dvb200.h "token key" spec). */
uint64_t dvo_token_key(const uint32_t* tok, int32_t len);
/* sorted distinct keys of one text; returns count or <0 */
int64_t dvo_tokenize_v3_keys(const uint8_t* utf8, int64_t nbytes, uint64_t* keys, int64_t cap);

/* ---- a2 core: infer_weak_evidence_refs_v3 scoring loop (server.py:24813-24864) ----
 * Sessions are CSR over claims and over facts; texts are CSR over UTF-8 bytes.
 * out_overlap (optional, may be NULL) is laid out per session as claims x facts
 * in the order of pair_off (int64[n_sessions+1], prefix sums of nc*nf). */
int dvo_bind_claims(const uint8_t* claim_utf8, const int64_t* claim_off,
                    const uint8_t* fact_utf8, const int64_t* fact_off,
                    const int64_t* sess_claim_off, const int64_t* sess_fact_off, int64_t n_sessions,
                    const double* fact_quality, const int32_t* fact_dim, const uint8_t* fact_hard,
                    const int32_t* claim_pref_dim, const double* claim_threshold,
                    int32_t* out_best_fact, double* out_best_score, uint8_t* out_bound,
                    int32_t* out_len_claim, int32_t* out_len_fact,
                    int32_t* out_overlap, const int64_t* pair_off,
                    int n_threads);

/* ---- a4: _build_reference_material_chunks (server.py:31932-31970) ---- */
int64_t dvo_chunk_build(const uint8_t* utf8, int64_t nbytes, int32_t chunk_size, int32_t overlap_cfg,
                        int32_t* out_start, int32_t* out_end, int32_t* out_tstart, int32_t* out_tend,
                        int64_t cap);

/* ---- a6: _tokenize_reference_material_query (server.py:32037-32051) ----
 * terms are written as lowered code points into term_cp (CSR by term_off); returns #terms (<=80) */
int64_t dvo_query_terms(const uint8_t* utf8, int64_t nbytes, uint32_t* term_cp, int64_t cp_cap,
                        int32_t* term_off, int64_t term_cap);

/* ---- a7: scoring loop of select_reference_material_context (server.py:32073-32087) ----
 * score[c] = sum_t lower(chunk_c).count(term_t)  (non-overlapping substring count) */
int dvo_chunk_scores(const uint8_t* chunk_utf8, const int64_t* chunk_off, int64_t n_chunks,
                     const uint32_t* term_cp, const int32_t* term_off, int64_t n_terms,
                     int32_t* out_score, int n_threads);
/* top-k by (-score, index) over score>0; returns #selected */
int64_t dvo_topk(const int32_t* score, const int64_t* index, int64_t n, int32_t k,
                 int64_t* out_index, int32_t* out_score);

/* ---- a8/a9: question normaliser + bigram Jaccard gate (server.py:20462-20501) ----
 * syn_src/syn_dst and stop are CSR tables of code points in the CALLER's iteration order. */
typedef struct {
    const uint32_t* syn_src; const int32_t* syn_src_off;
    const uint32_t* syn_dst; const int32_t* syn_dst_off; int32_t n_syn;
    const uint32_t* stop;    const int32_t* stop_off;    int32_t n_stop;
} dvo_gate_tables;
int64_t dvo_normalize_question(const uint8_t* utf8, int64_t nbytes, const dvo_gate_tables* t,
                               uint32_t* out, int64_t cap);
/* sorted distinct bigram keys of the normalised question (key = cp0<<21|cp1, or 1<<42|cp for len 1) */
int64_t dvo_question_keys(const uint8_t* utf8, int64_t nbytes, const dvo_gate_tables* t,
                          uint64_t* keys, int64_t cap);
/* for each candidate: largest history index h with J(hist_h, cand) >= thr, else -1 */
int dvo_question_gate(const uint8_t* cand_utf8, const int64_t* cand_off, int64_t n_cand,
                      const uint8_t* hist_utf8, const int64_t* hist_off, int64_t n_hist,
                      const dvo_gate_tables* t, double threshold, int64_t* out_match, int n_threads);

/* ---- a10: _solution_signature_ngrams / _solution_jaccard_similarity (server.py:36798-36813)
 * inputs are ALREADY clean_solution_text()'d; returns unrounded J, and set sizes */
double dvo_solution_jaccard(const uint8_t* l, int64_t nl, const uint8_t* r, int64_t nr, int32_t ngram,
                            int32_t* inter, int32_t* uni);

#ifdef __cplusplus
}
#endif
#endif
