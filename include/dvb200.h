/* dvb200.h — C-ABI of libdvb200.so: B200 (sm_100a) evidence-indexing kernels for DeepVision.
 *
 * The reference (hehaivictor/deep-vision) has no FFI: its hot functions are module-level
 * Python `def`s in web/server.py that callers resolve by name at call time.
 * Each entry point below is what a ctypes binding for one of those functions would call;
 * the `replaces:` line cites the reference interface.  INTEGRATION.md shows the binding.
 *
 * Conventions
 *  - plain pointers and sizes; no torch / C++ types.  Every function returns int32
 *    (DVB_OK or a negative DVB_E_*); dvb_last_error(ctx) gives the message.
 *  - text batches are CSR: `utf8` bytes + int64 `off[n+1]` byte offsets.  Each text must be
 *    valid UTF-8 (Python's "surrogatepass" 3-byte forms are accepted).  The byte buffer must
 *    be readable up to the next 16-byte boundary past its end (device buffers from cudaMalloc
 *    or torch always are; host buffers are copied by the library, which pads).
 *  - `mem` says where ALL pointer arguments of a call live: DVB_MEM_HOST (library copies
 *    in/out on its stream; this is the end-to-end path) or DVB_MEM_DEVICE (resident data,
 *    nothing is copied; outputs are written on the device).
 *  - the caller allocates every output array.  The library never retains caller pointers.
 *  - re-entrant: each call leases a stream + workspace from the ctx; no default stream, no
 *    device-wide sync.  One ctx per (process, device).
 *  - there is NO CPU fallback: without a CUDA device dvb_create fails.
 *
 * Token key (64-bit), produced by dvb_tokenize_v3 and used inside dvb_bind_claims:
 *   symbols are the lowered code points of class ASCII [a-z0-9_] or CJK [U+4E00..U+9FFF];
 *   code(sym) = cp-0x4E00+1 (CJK, 1..20992) or a-z 1..26, 0-9 27..36, _ 37 (ASCII)
 *   tag 0 (bits 63:62): CJK bigram           : (code0 << 15) | code1              (injective)
 *   tag 1: ASCII token (len >= 2)            : (1<<62) | (M >> 2)
 *   tag 3: CJK piece (3..12 chars)           : (3<<62) | (M >> 2)
 *   M = sum_d code_d * K(d) + len * K_LEN mod 2^64, K(d) = splitmix64(d+1)|1, K_LEN = splitmix64(0x4C454E)|1
 *   (multilinear hashing: any two distinct tokens collide with probability 2^-62; the oracle
 *   compares token STRINGS, so a collision would surface as a parity failure, not silently)
 */
#ifndef DVB200_H
#define DVB200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define DVB_OK          0
#define DVB_E_INVALID  (-1)
#define DVB_E_CAPACITY (-2)
#define DVB_E_CUDA     (-3)
#define DVB_E_NOMEM    (-4)

#define DVB_MEM_HOST   0
#define DVB_MEM_DEVICE 1

typedef struct dvb_ctx dvb_ctx;

/* ---- handles ---- */
int32_t     dvb_create(int32_t device, dvb_ctx** out);
void        dvb_destroy(dvb_ctx* ctx);
const char* dvb_last_error(dvb_ctx* ctx);           /* per thread: valid until the next failing call of this thread */
int32_t     dvb_version(void);
/* Unicode version of the case / class tables compiled into the library (CPython's unicodedata.unidata_version at
 * generation time): str.lower() of web/server.py:24723, 32039, 20463 must agree with the host interpreter's. */
const char* dvb_unicode_version(void);
/* number of library kernels launched on this ctx since creation (bench.py's gpu_launches) */
int64_t     dvb_launch_count(dvb_ctx* ctx);
/* device time (ms) of the dominant kernel of the most recent call on this ctx, measured with CUDA
 * events on the launching stream (0 if timing was not enabled with dvb_set_timing). */
void        dvb_set_timing(dvb_ctx* ctx, int32_t enabled);
float       dvb_last_kernel_ms(dvb_ctx* ctx);
/* run the subsequent calls THIS THREAD makes on ctx on the caller's CUDA stream (a cudaStream_t; NULL restores the
 * internal per-call streams).  Lets the caller bracket calls with its own CUDA events; other threads are not affected. */
void        dvb_set_stream(dvb_ctx* ctx, void* stream);

/* ---- a1  replaces: _tokenize_similarity_text_v3(text) -> set[str]   web/server.py:24722-24732
 * For each text: the sorted distinct token keys.  out_keys is CSR by out_off[n+1] (int64).
 * key_cap = capacity of out_keys; DVB_E_CAPACITY if too small (sum over texts of bytes/2+1 always fits). */
int32_t dvb_tokenize_v3(dvb_ctx* ctx, int32_t mem,
                        const uint8_t* utf8, const int64_t* off, int64_t n_texts,
                        uint64_t* out_keys, int64_t key_cap, int64_t* out_off);

/* ---- a2  replaces: the per-fact scoring loop of infer_weak_evidence_refs_v3   web/server.py:24813-24864
 * Batch of sessions.  Session s owns claims [sess_claim_off[s], sess_claim_off[s+1]) and facts
 * [sess_fact_off[s], sess_fact_off[s+1]) (facts in evidence_pack order, already filtered to q_id ~ Q\d+).
 *   claim text  = " ".join(item[k] for k in text_fields[field])          (24798-24803)
 *   fact text   = question + " " + answer + " " + dimension_name + " " + dimension   (24821-24826)
 *   fact_quality: quality_score already clamped to [0,1]; fact_dim: id of fact["dimension"];
 *   fact_hard: bool(hard_triggered); claim_pref_dim: id of the preferred dimension or -1;
 *   claim_threshold: the clamped threshold of 24848-24859.
 * Outputs per claim: best_fact = session-local index of the first fact with the maximal score
 * (-1 if no fact overlaps), best_score (fp64, same operation order as the reference),
 * bound = !(best_score + 1e-9 < threshold).  Optional (may be NULL): len_claim / len_fact =
 * distinct-token counts CAPPED at 14 / 18 (all the reference's formula observes).
 * max_claim_group_bytes: max over sessions of the UTF-8 bytes of any 32 consecutive claims of a
 * session (sizes the on-chip token table); <= 0 lets the library compute it (costs a copy of the
 * offset arrays to the host when mem == DVB_MEM_DEVICE). */
int32_t dvb_bind_claims(dvb_ctx* ctx, int32_t mem,
                        const uint8_t* claim_utf8, const int64_t* claim_off,
                        const uint8_t* fact_utf8, const int64_t* fact_off,
                        const int64_t* sess_claim_off, const int64_t* sess_fact_off, int64_t n_sessions,
                        const double* fact_quality, const int32_t* fact_dim, const uint8_t* fact_hard,
                        const int32_t* claim_pref_dim, const double* claim_threshold,
                        int32_t* out_best_fact, double* out_best_score, uint8_t* out_bound,
                        int32_t* out_len_claim, int32_t* out_len_fact,
                        int64_t max_claim_group_bytes);

/* ---- a7  replaces: scoring loop + top-k of select_reference_material_context   web/server.py:32073-32087
 * Symbols: the lowered text as uint16 per character: 0 = separator (anything that is not a token
 * character), 1..20992 = CJK U+4E00+code-1, 0x8000|code = ASCII [a-z0-9_] (code a-z 1..26, 0-9 27..36, _ 37).
 * U+0130 yields 'i' + separator, U+212A yields 'k' (CPython str.lower()).  Query terms (web/server.py:32039)
 * contain only token characters, so lowered.count(term) == non-overlapping substring count on symbols. */

/* ---- a6  replaces: _tokenize_reference_material_query(query_text) -> list[str]   web/server.py:32037-32051
 * For a batch of query texts: the ordered, de-duplicated term list of each (terms of >= 2 characters of one class;
 * pure-CJK terms of >= 6 characters add their 4-grams, then their 3-grams; the first max_terms <= 80 terms) as the
 * symbol CSR dvb_chunk_index_topk / dvb_chunk_scores take: query q owns terms [out_query_term_off[q],
 * out_query_term_off[q+1]), term t = out_term_sym[out_term_off[t] .. out_term_off[t+1]).  Capacities: out_term_off
 * nq * max_terms + 1 entries, out_query_term_off nq + 1, out_term_sym sym_cap (DVB_E_CAPACITY names the size needed;
 * 8 * (bytes of the batch) + 16 always suffices for texts of <= 80 terms worth of n-grams ... the call reports it).
 * out_totals2 = { terms, symbols } of the batch is ALWAYS a host pointer. */
int32_t dvb_query_terms(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t nq, int32_t max_terms,
                        uint16_t* out_term_sym, int64_t sym_cap, int32_t* out_term_off, int32_t* out_query_term_off, int64_t* out_totals2);

/* text t -> out_sym[off[t]-off[0] .. +out_cnt[t])   (out_sym needs off[n]-off[0] entries) */
int32_t dvb_symbolize(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_texts,
                      uint16_t* out_sym, int32_t* out_cnt);

/* out_score[q*n_chunks + c] = sum over the terms of query q of count(term in lower(chunk c)).
 * Terms are symbol strings: term t = term_sym[term_off[t] .. term_off[t+1]); query q owns terms
 * [query_term_off[q], query_term_off[q+1]).  Exact scan; intended for API-sized inputs. */
int32_t dvb_chunk_scores(dvb_ctx* ctx, int32_t mem, const uint8_t* chunk_utf8, const int64_t* chunk_off, int64_t n_chunks,
                         const uint16_t* term_sym, const int32_t* term_off, int32_t n_terms,
                         const int32_t* query_term_off, int32_t nq, int32_t* out_score);

/* Per query: the k (<= 32) best candidates with score > 0 under (score desc, index asc) ==
 * sorted(positive, key=(-score, index))[:k].  score is [nq][n]; idx is [nq][n] explicit indices or NULL
 * (index = index_base + position).  Also merges candidate lists (per shard / per GPU after the
 * all-gather): pass the gathered lists as score/idx.  out_idx/out_score are [nq][k] (idx -1 / score 0 =
 * empty), out_n[q] = number of valid entries. */
int32_t dvb_topk(dvb_ctx* ctx, int32_t mem, const int32_t* score, const int64_t* idx, int64_t n, int32_t nq,
                 int64_t index_base, int32_t k, int64_t* out_idx, int32_t* out_score, int32_t* out_n);

/* ---- dvb_topk_static: NO counterpart in the reference — added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as dvb_topk(DVB_MEM_DEVICE) — same kernels, same total order — but the call allocates nothing: no
 * cudaMalloc / cudaFree, no workspace lease, no host<->device copy, no synchronisation.  The per-split partial lists
 * live in one file-scope device array of DVB_TOPK_STATIC_SCRATCH_BYTES that every GPU loading the library holds until
 * exit; its device address is resolved once per device at dvb_create (cudaGetSymbolAddress with that device current).
 *  - ALL pointers are device pointers on ctx's device; the two kernels are queued on `stream` (a cudaStream_t owned
 *    by the caller; NULL = the legacy default stream) and the call returns without waiting: the outputs are valid once
 *    `stream` has passed this point.
 *  - limit: nq * k * 12 bytes <= DVB_TOPK_STATIC_SCRATCH_BYTES (nq * k <= 349 525 entries), else DVB_E_CAPACITY with the
 *    sizes in dvb_last_error; within the limit the split count shrinks to what the array holds (results do not change).
 *  - every call on a device shares the array and finds the previous call's contents in it.  Nothing has to start at
 *    zero (every list slot a call reads was written by that call first), but two calls that OVERLAP on the GPU — different
 *    streams, or graph branches without a dependency — race on it: order them. */
#define DVB_TOPK_STATIC_SCRATCH_BYTES (4 << 20)
int32_t dvb_topk_static(dvb_ctx* ctx, void* stream, const int32_t* d_score, const int64_t* d_idx, int64_t n, int32_t nq,
                        int64_t index_base, int32_t k, int64_t* d_out_idx, int32_t* d_out_score, int32_t* d_out_n);

/* ---- a4  replaces: _build_reference_material_chunks(content)   web/server.py:31932-31970
 * Per document d the chunks are written at [cap_off[d], cap_off[d] + out_count[d]) of the output arrays
 * (cap_off[n_docs+1]: caller-chosen capacities; a document of B bytes never yields more than
 * B / max(1, min(size, max(200,size/2)+1) - min(overlap,size/3)) + 2 chunks).  start/end/tstart/tend are CODE-POINT
 * indices: chunk["start"], chunk["end"], and chunk["text"] == content[tstart:tend] (== content[start:end].strip());
 * tbyte0/tbyte1 are the byte offsets of tstart/tend inside the document.  Chunks whose stripped text is
 * empty are skipped (web/server.py:31952), so `index` is the position in the output. */
int32_t dvb_chunk_build(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_docs,
                        int32_t chunk_size, int32_t overlap, const int64_t* cap_off,
                        int32_t* out_count, int32_t* out_start, int32_t* out_end, int32_t* out_tstart, int32_t* out_tend,
                        int64_t* out_tbyte0, int64_t* out_tbyte1);

/* ---- a8/a9  replaces: normalize_interview_question_text / question_token_set /
 *                        is_similar_interview_question / find_similar_interview_question_log   web/server.py:20462-20501
 * Synonym and stop-word tables are code-point CSR arrays IN THE CALLER'S ITERATION ORDER (the reference
 * iterates a dict and a set: web/server.py:20464, 20467 — the set order depends on PYTHONHASHSEED, so the
 * caller passes list(QUESTION_SIMILARITY_STOPWORDS) of its own process).  Empty synonym sources are skipped.
 * Bigram key: (cp0 << 21) | cp1 over ALL code points of the normalised text; a 1-char text is (1<<42)|cp. */

/* Sorted distinct bigram keys of each question at out_keys[out_cap_off[i] .. + out_key_cnt[i]); optionally the
 * normalised text (code points) at out_norm[out_cap_off[i] .. + out_norm_len[i]).  out_cap >= sum (bytes_i+2)*G,
 * G = worst-case growth of the synonym replaces (1 for the reference's tables). */
int32_t dvb_question_keys(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n,
                          const uint32_t* syn_src, const int32_t* syn_src_off, const uint32_t* syn_dst, const int32_t* syn_dst_off, int32_t n_syn,
                          const uint32_t* stop, const int32_t* stop_off, int32_t n_stop,
                          int64_t* out_cap_off, uint64_t* out_keys, int32_t* out_key_cnt, uint32_t* out_norm, int32_t* out_norm_len, int64_t out_cap);

/* out_match[c] = the LARGEST history index h with  |L&R| / max(1, |L|R|) >= threshold  (fp64, as the reference),
 * L = token set of history[h], R = token set of candidate c; -1 if none (empty sets never match). */
int32_t dvb_question_gate(dvb_ctx* ctx, int32_t mem, const uint8_t* cand_utf8, const int64_t* cand_off, int64_t n_cand,
                          const uint8_t* hist_utf8, const int64_t* hist_off, int64_t n_hist,
                          const uint32_t* syn_src, const int32_t* syn_src_off, const uint32_t* syn_dst, const int32_t* syn_dst_off, int32_t n_syn,
                          const uint32_t* stop, const int32_t* stop_off, int32_t n_stop,
                          double threshold, int64_t* out_match);

/* ---- a10  replaces: _solution_signature_ngrams / _solution_jaccard_similarity   web/server.py:36798-36813
 * Inputs are pairs of ALREADY clean_solution_text()'d strings.  Per pair: |L&R|, |L|R|, |L|, |R| of the
 * n-gram sets (1 <= ngram <= 4) over lower(text) restricted to [0-9a-z] and CJK; the caller rounds
 * inter / max(1, union) to 3 places (0.0 if either set is empty). */
int32_t dvb_solution_jaccard(dvb_ctx* ctx, int32_t mem, const uint8_t* left_utf8, const int64_t* left_off,
                             const uint8_t* right_utf8, const int64_t* right_off, int64_t n_pairs, int32_t ngram,
                             int32_t* out_inter, int32_t* out_union, int32_t* out_nleft, int32_t* out_nright);

/* ---- SURVEY 8(f)4  replaces the scoring, ordering and the three greedy passes of: select_slimmed_facts_for_prompt
 *               (web/server.py:22886-22993), for n_reports evidence packs per launch (one warp per pack).
 * Facts of report r = [fact_off[r], fact_off[r+1]) (at most 1024).  Per fact, computed by the caller the way the reference does:
 *   quality   max(0.0, min(1.0, _safe_float(quality_score)))                                  (server.py:22923)
 *   flags     bit 0 hard_triggered, 1 q_id in the contradiction refs, 2 q_id in the unknown refs, 3 is_follow_up
 *   sig_id    dense id in [0, n) of _fact_signature_for_prompt (equal ids <=> equal signatures)  (server.py:22876-22883)
 *   dim_id    dense id in [0, n) of the lower-cased dimension ("__unknown__" included)
 *   sort_key  _fact_sort_key_for_prompt                                                        (server.py:22867-22874)
 * limit[r] = max(1, int(facts_limit or 1)); dim_quota / min_quality / keep_hard / dedup = REPORT_V3_EVIDENCE_DIM_QUOTA,
 * _MIN_QUALITY, _KEEP_HARD_TRIGGERED, _DEDUP_ENABLED (server.py:992-998).
 * out_sel[out_off[r] ..) receives the selected fact indices (0-based inside the report) in the reference's final order,
 * out_cnt[r] their number; out_off[r+1] - out_off[r] >= min(limit[r], n) (3 n when dedup == 0: a fact may then be taken by
 * more than one pass, as in the reference). */
int32_t dvb_slim_facts(dvb_ctx* ctx, int32_t mem, int64_t n_reports, const int64_t* fact_off, const double* quality, const uint8_t* flags,
                       const int32_t* sig_id, const int32_t* dim_id, const int64_t* sort_key, const int32_t* limit, const int64_t* out_off,
                       int32_t dim_quota, double min_quality, int32_t keep_hard, int32_t dedup, int32_t* out_sel, int32_t* out_cnt);

/* ---- a11-a13  replaces the string scans of: evaluate_answer_depth (web/server.py:21221-21401), _score_rich_option_answer
 *               (20318-20367), evaluate_answer_quality (20370-20426), _infer_legacy_log_capture_contract (21419-21466)
 * Per text: [b0, b1) = byte bounds of Python's text.strip(); out_len = len(text.strip()) in code points;
 * out_flags bit 0 = any(c.isdigit() for c in text.strip()), bit 1 = stripped text non-empty;
 * out_mask bit g = some pattern of group g (0..63) occurs in the stripped text as a substring, or — for patterns with
 * pat_whole != 0 — equals the stripped text.  Patterns (UTF-8, 1..255 bytes each, CSR by int32 pat_off[n_pat+1]) are the
 * reference's literal keyword lists; the pattern arrays are ALWAYS host pointers (they are tiny), whatever `mem` says.
 * out_b0 / out_b1 may both be NULL. */
int32_t dvb_text_features(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_texts,
                          const uint8_t* pat_utf8, const int32_t* pat_off, const uint8_t* pat_group, const uint8_t* pat_whole, int32_t n_pat,
                          uint64_t* out_mask, int32_t* out_len, uint8_t* out_flags, int32_t* out_b0, int32_t* out_b1);

/* Needle j is searched in haystack needle_hay[j] (`option in answer`, web/server.py:20326-20327, 21321).  out_bits[j]:
 * bit 0 = needle (raw) in haystack.strip()  ("" is contained in everything, as in Python);
 * bit 1 = needle.strip() non-empty and in haystack.strip();  bit 2 = needle.strip() == haystack.strip() != "". */
int32_t dvb_contains(dvb_ctx* ctx, int32_t mem, const uint8_t* hay_utf8, const int64_t* hay_off, int64_t n_hay,
                     const uint8_t* needle_utf8, const int64_t* needle_off, const int32_t* needle_hay, int64_t n_needles,
                     uint8_t* out_bits);

/* ---- a11 / a12 decision stage  replaces, per interview-log turn: _infer_legacy_log_capture_contract (web/server.py:
 *      21419-21466), the contract resolution of backfill_session_interview_log_evidence_annotations (21484-21520),
 *      _score_rich_option_answer (20318-20367), evaluate_answer_depth (21221-21401), evaluate_answer_quality (20370-20426) and
 *      classify_answer_evidence_class_v3 (22552-22565), over the feature bits of dvb_text_features / dvb_contains.
 * Inputs per turn: masks / stripped length / flags of the question and the answer; the options of turn t are entries
 * [opt_off[t], opt_off[t+1]) of the option arrays (mask, stripped length, opt_flags bit 0 = has digit, bit 1 = the option is a
 * str; opt_bits = dvb_contains bits against the turn's answer).  The masks use the group numbers DVB_AN_G_* below.
 * turn_flags: bit 0 bool(options), 1 multi_select, 2 bool(is_follow_up), 3 bool(log hard_triggered), 4 bool(log requires_rationale),
 *   5 dimension in ("tech_constraints", "project_constraints"), 6 the same test on str(dimension).strip().lower(),
 *   8-9 answer_mode of the log (0 = none / invalid, 1 pick_only, 2 pick_with_reason), 10-11 evidence_intent of the log
 *   (0 invalid, 1 low, 2 medium, 3 high), 13 is_follow_up or follow_up_round > 0.
 * formal_prior = number of earlier logs of the session with the same dimension and is_follow_up false (21432);
 * sensitivity = DIMENSION_FOLLOW_UP_SENSITIVITY.get(dimension, 0.5); hard_signal_mask = signals of HARD_FOLLOW_UP_SIGNALS.
 * Signal ids (the order evaluate_answer_depth appends them): 0 too_short, 1 vague_expression, 2 generic_answer, 3 option_only,
 *   4 rich_option_answer, 5 no_quantification, 6 single_selection, 7 contradiction_detected.
 * out_bits: 0-7 signals, 8 needs_follow_up, 9 suggest_ai_eval, 10 hard_triggered, 11 has_numbers, 12 requires_rationale,
 *   13-14 answer_mode, 15-16 evidence_intent, 17-19 evidence class (0 explicit, 1 rich_option, 2 weak_inferred, 3 pending_follow_up),
 *   20 a non-str option reached `opt in answer` (the reference raises TypeError there), 21 quality >= 0.75 before rounding,
 *   22 stripped answer >= 80 characters.  Scores are fp64 in the reference's operation order, rounded like Python round(x, 2). */
#define DVB_AN_G_VAGUE 0
#define DVB_AN_G_INFORMATIVE 1
#define DVB_AN_G_SEMICOLON 2
#define DVB_AN_G_CONTRA_L 3   /* six pairs: left words 3..8, right words 9..14 */
#define DVB_AN_G_CONTRA_R 9
#define DVB_AN_G_SCENARIO 15
#define DVB_AN_G_GENERIC 16
#define DVB_AN_G_OPT_PUNCT 17
#define DVB_AN_G_CAUSE 18
#define DVB_AN_G_CONTEXT 19
#define DVB_AN_G_ROLE 20
#define DVB_AN_G_PROCESS 21
#define DVB_AN_G_UNITS 22
#define DVB_AN_G_Q_RICH 24
#define DVB_AN_G_Q_EVIDENCE 25
int32_t dvb_annotate_turns(dvb_ctx* ctx, int32_t mem, int64_t n_turns,
                           const uint64_t* q_mask, const uint64_t* a_mask, const int32_t* a_len, const uint8_t* a_flags,
                           const int32_t* opt_off, const uint64_t* opt_mask, const int32_t* opt_len, const uint8_t* opt_flags, const uint8_t* opt_bits,
                           const uint32_t* turn_flags, const int32_t* follow_up_round, const int32_t* formal_prior, const double* sensitivity,
                           uint32_t hard_signal_mask, int32_t overwrite_contract,
                           uint32_t* out_bits, double* out_follow_up_score, double* out_quality_score);

/* ---- a7 at corpus scale (BASELINE configs 3/5)  replaces: the scoring loop + top-k of select_reference_material_context
 *      (web/server.py:32073-32087) for many queries over a resident chunk corpus.
 * dvb_chunk_index_build symbolises the chunks and builds a positional inverted index over adjacent same-class symbol
 * pairs (radix sort + segmented reduce); the index lives in device memory owned by the returned handle (free it with
 * dvb_chunk_index_free; the caller's text is not retained).  dvb_chunk_index_topk scores term lists (symbol strings as
 * for dvb_chunk_scores; every term must be >= 2 symbols of one class, which is what web/server.py:32039-32049 emits)
 * exactly — non-overlapping substring counts, summed per query — and returns per query the k (<= 32) best chunks with
 * score > 0 under (score desc, index asc), index = index_base + chunk position (so shards of a corpus can be merged
 * with dvb_topk).  stats: out6 = { n_chunks, text bytes, postings, distinct pair keys, resident device bytes, device }. */
typedef struct dvb_index dvb_index;
int32_t dvb_chunk_index_build(dvb_ctx* ctx, int32_t mem, const uint8_t* utf8, const int64_t* off, int64_t n_chunks, dvb_index** out);
void    dvb_chunk_index_free(dvb_index* index);
int32_t dvb_chunk_index_stats(const dvb_index* index, int64_t* out6);
int32_t dvb_chunk_index_topk(dvb_ctx* ctx, const dvb_index* index, int32_t mem, const uint16_t* term_sym, const int32_t* term_off, int32_t n_terms,
                             const int32_t* query_term_off, int32_t nq, int64_t index_base, int32_t k,
                             int64_t* out_idx, int32_t* out_score, int32_t* out_n);

/* ---- §8(f)1  persistent sidecar of a chunk index, written next to chunks.json (web/server.py:31986-31999) so that a
 * later process does not re-parse and re-lower every chunk per prompt build (web/server.py:32028, 32075).
 * save: writes `path` atomically (path + ".tmp", then rename).  tag32 = 32 caller bytes identifying the content the index
 * was built from (may be NULL = all zero).  load: refuses a file whose tag differs from tag32 (NULL = accept any), whose
 * device-side checksums or structural validation fail, or whose format version differs: DVB_E_INVALID, never a stale or
 * partial index.  The loaded index belongs to ctx's device and is freed with dvb_chunk_index_free. */
int32_t dvb_chunk_index_save(dvb_ctx* ctx, const dvb_index* index, const char* path, const uint8_t* tag32);
int32_t dvb_chunk_index_load(dvb_ctx* ctx, const char* path, const uint8_t* tag32, dvb_index** out);

#ifdef __cplusplus
}
#endif
#endif
