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
const char* dvb_last_error(dvb_ctx* ctx);           /* valid until the next call on ctx from this thread */
int32_t     dvb_version(void);
/* number of library kernels launched on this ctx since creation (bench.py's gpu_launches) */
int64_t     dvb_launch_count(dvb_ctx* ctx);
/* device time (ms) of the dominant kernel of the most recent call on this ctx, measured with CUDA
 * events on the launching stream (0 if timing was not enabled with dvb_set_timing). */
void        dvb_set_timing(dvb_ctx* ctx, int32_t enabled);
float       dvb_last_kernel_ms(dvb_ctx* ctx);
/* run all subsequent calls on the caller's CUDA stream (a cudaStream_t; NULL restores the internal
 * per-call streams).  Lets the caller bracket calls with its own CUDA events. */
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

#ifdef __cplusplus
}
#endif
#endif
