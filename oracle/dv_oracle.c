/* dv_oracle.c — CPU ORACLE for the DeepVision evidence-indexing hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a plain-C restatement of the reference's
 * algorithms (reference = /root/reference/web/server.py, pure Python); each
 * function cites the reference lines it follows.  It is pinned against the real
 * reference by tests/test_reference_diff.py (AST-executed reference functions,
 * run in the build container) and by the committed fixtures in tests/golden/
 * (generated from the reference by tools/make_golden.py).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may load this library.  The product never does.
 *
 * String semantics are CPython's: code-point indexing, full-Unicode lower()
 * (including U+0130 -> "i̇" and the final-sigma rule), Unicode
 * whitespace for strip()/\s, str.isdigit().  Tables: include/dv_unicode_tables.h.
 */
#define _GNU_SOURCE
#include "dv_oracle.h"
#include "../include/dv_unicode_tables.h"
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ strings */

int64_t dvo_utf8_decode(const uint8_t* s, int64_t n, uint32_t* out) {
    int64_t i = 0, k = 0;
    while (i < n) {
        uint32_t b = s[i], cp; int len;
        if (b < 0x80) { cp = b; len = 1; }
        else if ((b & 0xE0) == 0xC0) { cp = b & 0x1F; len = 2; }
        else if ((b & 0xF0) == 0xE0) { cp = b & 0x0F; len = 3; }
        else if ((b & 0xF8) == 0xF0) { cp = b & 0x07; len = 4; }
        else { cp = 0xFFFD; len = 1; }              /* stray continuation byte */
        int j;
        for (j = 1; j < len; j++) {
            if (i + j >= n || (s[i + j] & 0xC0) != 0x80) break;
            cp = (cp << 6) | (s[i + j] & 0x3F);
        }
        if (j != len) { cp = 0xFFFD; }              /* truncated sequence */
        out[k++] = cp;
        i += j;
    }
    return k;
}

static int in_ranges(uint32_t cp, const uint32_t* lo, const uint32_t* hi, int n) {
    int a = 0, b = n - 1;
    while (a <= b) {
        int m = (a + b) >> 1;
        if (cp < lo[m]) b = m - 1; else if (cp > hi[m]) a = m + 1; else return 1;
    }
    return 0;
}
int dvo_is_space(uint32_t cp) {
    for (int i = 0; i < DV_SPACE_N; i++) if (dv_space_cp[i] == cp) return 1;
    return 0;
}
int dvo_is_digit(uint32_t cp) { return in_ranges(cp, dv_digit_lo, dv_digit_hi, DV_DIGIT_RANGES_N); }
static int is_cased(uint32_t cp) { return in_ranges(cp, dv_cased_lo, dv_cased_hi, DV_CASED_RANGES_N); }
static int is_cign(uint32_t cp)  { return in_ranges(cp, dv_cign_lo, dv_cign_hi, DV_CIGN_RANGES_N); }

static int lower_lookup(uint32_t cp) {
    int a = 0, b = DV_LOWER_N - 1;
    while (a <= b) {
        int m = (a + b) >> 1;
        if (cp < dv_lower_src[m]) b = m - 1; else if (cp > dv_lower_src[m]) a = m + 1; else return m;
    }
    return -1;
}

/* CPython str.lower(): per-char full mapping; U+03A3 by the Final_Sigma context. */
int64_t dvo_lower(const uint32_t* in, int64_t n, uint32_t* out, int64_t cap) {
    int64_t k = 0;
    for (int64_t i = 0; i < n; i++) {
        uint32_t cp = in[i];
        if (k + 3 > cap) return -1;
        if (cp < 0x80) { out[k++] = (cp >= 'A' && cp <= 'Z') ? cp + 32 : cp; continue; }
        if (cp == 0x3A3) {
            int64_t j; uint32_t c = 0; int fin;
            for (j = i - 1; j >= 0; j--) { c = in[j]; if (!is_cign(c)) break; }
            fin = (j >= 0 && is_cased(c));
            if (fin) {
                for (j = i + 1; j < n; j++) { c = in[j]; if (!is_cign(c)) break; }
                fin = (j == n || !is_cased(c));
            }
            out[k++] = fin ? 0x3C2 : 0x3C3;
            continue;
        }
        int m = lower_lookup(cp);
        if (m < 0) { out[k++] = cp; continue; }
        for (int t = 0; t < 3 && dv_lower_dst[m][t]; t++) out[k++] = dv_lower_dst[m][t];
    }
    return k;
}

static inline int is_ascii_tok(uint32_t c) {
    return (c >= 'a' && c <= 'z') || (c >= '0' && c <= '9') || c == '_';
}
static inline int is_cjk(uint32_t c) { return c >= 0x4E00 && c <= 0x9FFF; }

/* ---------------------------------------------------- a1 raw token scanning */

/* Emit (start,len) of every token occurrence (duplicates included) of the lowered text.
 * server.py:24727  re.findall(r"[a-z0-9_]{2,}")            -> maximal ASCII runs, len >= 2
 * server.py:24728  re.findall(r"[一-鿿]{2,12}")    -> greedy non-overlapping pieces
 * server.py:24729-24731  piece itself + its len-1 bigrams */
typedef void (*tok_cb)(void* ctx, int32_t start, int32_t len);

static void scan_tokens_v3(const uint32_t* low, int64_t n, tok_cb cb, void* ctx) {
    int64_t i = 0;
    while (i < n) {
        if (is_ascii_tok(low[i])) {
            int64_t j = i; while (j < n && is_ascii_tok(low[j])) j++;
            if (j - i >= 2) cb(ctx, (int32_t)i, (int32_t)(j - i));
            i = j;
        } else if (is_cjk(low[i])) {
            int64_t j = i; while (j < n && is_cjk(low[j])) j++;
            for (int64_t p = i; p < j; p += 12) {
                int64_t len = j - p < 12 ? j - p : 12;
                if (len < 2) break;
                cb(ctx, (int32_t)p, (int32_t)len);
                for (int64_t b = p; b + 1 < p + len; b++) cb(ctx, (int32_t)b, 2);
            }
            i = j;
        } else i++;
    }
}

/* -------------------------------------------------------------- token keys */

static uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
static inline uint32_t ascii_code(uint32_t c) {        /* a-z 1..26, 0-9 27..36, _ 37 */
    if (c >= 'a' && c <= 'z') return c - 'a' + 1;
    if (c >= '0' && c <= '9') return c - '0' + 27;
    return 37;
}
/* Token key spec (include/dvb200.h), restated:
 *  code(sym) = cp-0x4E00+1 for CJK, 1..37 for ASCII [a-z0-9_]
 *  tag 0: CJK bigram                : (code0 << 15) | code1                      (injective)
 *  tag 1: ASCII token, len >= 2     : (1<<62) | (M >> 2)
 *  tag 3: CJK piece, len 3..12      : (3<<62) | (M >> 2)
 *  M = sum_d code_d * K(d) + len * K_LEN  (mod 2^64),  K(d) = splitmix64(d+1)|1  (multilinear hash) */
uint64_t dvo_token_key(const uint32_t* tok, int32_t len) {
    int cjk = is_cjk(tok[0]);
    if (cjk && len == 2) return ((uint64_t)(tok[0] - 0x4E00 + 1) << 15) | (uint64_t)(tok[1] - 0x4E00 + 1);
    uint64_t m = 0;
    for (int32_t d = 0; d < len; d++) {
        uint64_t code = cjk ? (uint64_t)(tok[d] - 0x4E00 + 1) : (uint64_t)ascii_code(tok[d]);
        m += code * (splitmix64((uint64_t)d + 1) | 1ULL);
    }
    m += (uint64_t)len * (splitmix64(0x4C454EULL) | 1ULL);
    return ((cjk ? 3ULL : 1ULL) << 62) | (m >> 2);
}

/* ------------------------------------------- content-exact token interning */

typedef struct { uint64_t h; const uint32_t* p; int32_t len; int32_t id; } intern_ent;
typedef struct {
    intern_ent* e; int64_t cap; int32_t n;     /* cap is a power of two */
} intern_tab;

static uint64_t content_hash(const uint32_t* p, int32_t len) {
    uint64_t h = 1469598103934665603ULL;
    for (int32_t i = 0; i < len; i++) { h ^= p[i]; h *= 1099511628211ULL; }
    h ^= h >> 29; return h | 1;                   /* never 0: 0 marks an empty slot */
}
static void intern_init(intern_tab* t, int64_t want) {
    int64_t cap = 64; while (cap < want * 2) cap <<= 1;
    t->e = (intern_ent*)calloc((size_t)cap, sizeof(intern_ent)); t->cap = cap; t->n = 0;
}
static void intern_clear(intern_tab* t) { memset(t->e, 0, (size_t)t->cap * sizeof(intern_ent)); t->n = 0; }
static void intern_free(intern_tab* t) { free(t->e); t->e = NULL; }
static void intern_grow(intern_tab* t) {
    intern_ent* old = t->e; int64_t ocap = t->cap;
    t->cap <<= 1; t->e = (intern_ent*)calloc((size_t)t->cap, sizeof(intern_ent));
    for (int64_t i = 0; i < ocap; i++) if (old[i].h) {
        int64_t s = (int64_t)(old[i].h & (uint64_t)(t->cap - 1));
        while (t->e[s].h) s = (s + 1) & (t->cap - 1);
        t->e[s] = old[i];
    }
    free(old);
}
/* returns id; *is_new set when inserted.  Equality is by CONTENT (len + code points). */
static int32_t intern_get(intern_tab* t, const uint32_t* p, int32_t len, int insert, int* is_new) {
    if (insert && (int64_t)t->n * 2 >= t->cap) intern_grow(t);
    uint64_t h = content_hash(p, len);
    int64_t s = (int64_t)(h & (uint64_t)(t->cap - 1));
    while (t->e[s].h) {
        if (t->e[s].h == h && t->e[s].len == len && memcmp(t->e[s].p, p, (size_t)len * 4) == 0) {
            if (is_new) *is_new = 0;
            return t->e[s].id;
        }
        s = (s + 1) & (t->cap - 1);
    }
    if (!insert) { if (is_new) *is_new = 0; return -1; }
    t->e[s].h = h; t->e[s].p = p; t->e[s].len = len; t->e[s].id = t->n;
    if (is_new) *is_new = 1;
    return t->n++;
}

/* ------------------------------------------------------------- a1 (public) */

typedef struct { intern_tab* tab; const uint32_t* low; int32_t* ts; int32_t* tl; int64_t cap; int64_t n; int ovf; } tokv3_ctx;
static void tokv3_cb(void* c, int32_t start, int32_t len) {
    tokv3_ctx* x = (tokv3_ctx*)c; int is_new = 0;
    intern_get(x->tab, x->low + start, len, 1, &is_new);
    if (is_new) {
        if (x->n >= x->cap) { x->ovf = 1; return; }
        x->ts[x->n] = start; x->tl[x->n] = len; x->n++;
    }
}

int64_t dvo_tokenize_v3(const uint8_t* utf8, int64_t nbytes, uint32_t* low, int64_t low_cap, int64_t* n_low,
                        int32_t* tok_start, int32_t* tok_len, int64_t tok_cap) {
    uint32_t* cps = (uint32_t*)malloc((size_t)(nbytes + 1) * 4);
    int64_t n = dvo_utf8_decode(utf8, nbytes, cps);
    int64_t nl = dvo_lower(cps, n, low, low_cap);
    free(cps);
    if (nl < 0) return -1;
    if (n_low) *n_low = nl;
    intern_tab tab; intern_init(&tab, 2 * nl + 16);
    tokv3_ctx ctx = { &tab, low, tok_start, tok_len, tok_cap, 0, 0 };
    scan_tokens_v3(low, nl, tokv3_cb, &ctx);
    intern_free(&tab);
    return ctx.ovf ? -2 : ctx.n;
}

static int cmp_u64(const void* a, const void* b) {
    uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b; return x < y ? -1 : x > y;
}
int64_t dvo_tokenize_v3_keys(const uint8_t* utf8, int64_t nbytes, uint64_t* keys, int64_t cap) {
    int64_t low_cap = 3 * (nbytes + 1);
    uint32_t* low = (uint32_t*)malloc((size_t)low_cap * 4);
    int64_t tcap = 2 * (nbytes + 1);
    int32_t* ts = (int32_t*)malloc((size_t)tcap * 4); int32_t* tl = (int32_t*)malloc((size_t)tcap * 4);
    int64_t nl = 0, nt = dvo_tokenize_v3(utf8, nbytes, low, low_cap, &nl, ts, tl, tcap);
    int64_t ret = nt;
    if (nt >= 0) {
        if (nt > cap) ret = -2;
        else {
            for (int64_t i = 0; i < nt; i++) keys[i] = dvo_token_key(low + ts[i], tl[i]);
            qsort(keys, (size_t)nt, 8, cmp_u64);
        }
    }
    free(low); free(ts); free(tl);
    return ret;
}

/* -------------------------------------------------------------- a2 scoring */

typedef struct {
    intern_tab tab;
    uint32_t* cps; int64_t cps_cap;            /* decode scratch */
    uint32_t* arena; int64_t arena_cap, arena_n; /* lowered text of the whole session */
    int32_t* stamp; uint64_t* cmask; int64_t ids_cap; int32_t mask_words;
    int32_t cur; int32_t distinct; int phase;  /* callback state */
    int32_t* ov_row;                            /* overlap[claim] for the current fact */
} bind_ws;

static void bind_reserve_ids(bind_ws* w, int64_t n) {
    if (n <= w->ids_cap) return;
    int64_t nc = w->ids_cap ? w->ids_cap : 1024; while (nc < n) nc <<= 1;
    w->stamp = (int32_t*)realloc(w->stamp, (size_t)nc * 4);
    w->cmask = (uint64_t*)realloc(w->cmask, (size_t)nc * 8 * (size_t)w->mask_words);
    w->ids_cap = nc;
}
static void bind_cb(void* c, int32_t start, int32_t len) {
    bind_ws* w = (bind_ws*)c; int is_new = 0;
    int32_t id = intern_get(&w->tab, w->arena + start, len, 1, &is_new);
    if (is_new) {
        bind_reserve_ids(w, (int64_t)id + 1);
        w->stamp[id] = -1;
        memset(w->cmask + (size_t)id * w->mask_words, 0, 8 * (size_t)w->mask_words);
    }
    if (w->stamp[id] == w->cur) return;          /* duplicate inside this text */
    w->stamp[id] = w->cur; w->distinct++;
    uint64_t* m = w->cmask + (size_t)id * w->mask_words;
    if (w->phase == 0) { m[(w->cur) >> 6] |= 1ULL << (w->cur & 63); }
    else {
        for (int32_t wd = 0; wd < w->mask_words; wd++) {
            uint64_t v = m[wd];
            while (v) { int b = __builtin_ctzll(v); v &= v - 1; w->ov_row[wd * 64 + b]++; }
        }
    }
}
/* lower one text into the session arena, scan it; returns 0 or <0 */
static int bind_text(bind_ws* w, const uint8_t* s, int64_t nb) {
    if (nb + 1 > w->cps_cap) { w->cps_cap = 2 * (nb + 1); w->cps = (uint32_t*)realloc(w->cps, (size_t)w->cps_cap * 4); }
    int64_t n = dvo_utf8_decode(s, nb, w->cps);
    /* arena must not move while interned pointers are live: it is sized per session up front */
    int64_t nl = dvo_lower(w->cps, n, w->arena + w->arena_n, w->arena_cap - w->arena_n);
    if (nl < 0) return -1;
    int64_t base = w->arena_n; w->arena_n += nl;
    const uint32_t* low = w->arena + base;
    /* bias the arena pointer so callback offsets (relative to `low`) resolve */
    uint32_t* saved = w->arena; w->arena = (uint32_t*)low;
    scan_tokens_v3(low, nl, bind_cb, w);
    w->arena = saved;
    return 0;
}

typedef struct {
    const uint8_t* claim_utf8; const int64_t* claim_off; const uint8_t* fact_utf8; const int64_t* fact_off;
    const int64_t* sc_off; const int64_t* sf_off; int64_t n_sessions;
    const double* fq; const int32_t* fdim; const uint8_t* fhard; const int32_t* cpref; const double* cthr;
    int32_t* best_fact; double* best_score; uint8_t* bound; int32_t* len_claim; int32_t* len_fact;
    int32_t* overlap; const int64_t* pair_off;
    int64_t next; pthread_mutex_t mu; int err;
} bind_job;

static void bind_session(bind_job* j, bind_ws* w, int64_t s) {
    int64_t c0 = j->sc_off[s], c1 = j->sc_off[s + 1], f0 = j->sf_off[s], f1 = j->sf_off[s + 1];
    int64_t nc = c1 - c0, nf = f1 - f0;
    int64_t bytes = (j->claim_off[c1] - j->claim_off[c0]) + (j->fact_off[f1] - j->fact_off[f0]);
    int64_t need = 3 * bytes + 64;
    if (need > w->arena_cap) { w->arena_cap = need; free(w->arena); w->arena = (uint32_t*)malloc((size_t)need * 4); }
    w->arena_n = 0;
    int32_t words = (int32_t)((nc + 63) / 64); if (words < 1) words = 1;
    if (words != w->mask_words) { w->mask_words = words; free(w->cmask); free(w->stamp); w->cmask = NULL; w->stamp = NULL; w->ids_cap = 0; }
    intern_clear(&w->tab);
    int32_t* ov_row = (int32_t*)calloc((size_t)(words * 64), 4);
    w->ov_row = ov_row;
    int32_t* lenA = (int32_t*)malloc((size_t)(nc + 1) * 4);
    /* claims: server.py:24803-24804 */
    w->phase = 0;
    for (int64_t c = 0; c < nc; c++) {
        w->cur = (int32_t)c; w->distinct = 0;
        if (bind_text(w, j->claim_utf8 + j->claim_off[c0 + c], j->claim_off[c0 + c + 1] - j->claim_off[c0 + c])) j->err = 1;
        lenA[c] = w->distinct;
        if (j->len_claim) j->len_claim[c0 + c] = w->distinct;
        j->best_fact[c0 + c] = -1; j->best_score[c0 + c] = 0.0; j->bound[c0 + c] = 0;
    }
    /* facts in index order: server.py:24814-24843 (strict '>' keeps the first maximum) */
    w->phase = 1;
    for (int64_t f = 0; f < nf; f++) {
        w->cur = (int32_t)(nc + f); w->distinct = 0;
        memset(ov_row, 0, (size_t)(words * 64) * 4);
        if (bind_text(w, j->fact_utf8 + j->fact_off[f0 + f], j->fact_off[f0 + f + 1] - j->fact_off[f0 + f])) j->err = 1;
        int32_t lenB = w->distinct;
        if (j->len_fact) j->len_fact[f0 + f] = lenB;
        double q = j->fq[f0 + f]; if (!(q > 0.0)) q = 0.0; if (q > 1.0) q = 1.0;   /* max(0.0, min(1.0, q)) */
        for (int64_t c = 0; c < nc; c++) {
            int32_t ov = ov_row[c];
            if (j->overlap) j->overlap[j->pair_off[s] + c * nf + f] = ov;
            if (ov <= 0 || lenA[c] == 0 || lenB == 0) continue;
            int32_t da = lenA[c] < 14 ? lenA[c] : 14; if (da < 3) da = 3;      /* max(3, min(len(A),14)) */
            int32_t db = lenB < 18 ? lenB : 18; if (db < 4) db = 4;            /* max(4, min(len(B),18)) */
            volatile double coverage = (double)ov / (double)da;
            volatile double precision = (double)ov / (double)db;
            double dimb = (j->cpref[c0 + c] >= 0 && j->fdim[f0 + f] == j->cpref[c0 + c]) ? 0.12 : 0.0;
            double hardb = j->fhard[f0 + f] ? 0.04 : 0.0;
            volatile double t0 = 0.62 * coverage;      /* volatile: one rounding per Python operation */
            volatile double t1 = 0.18 * precision;
            volatile double t2 = 0.16 * q;
            volatile double sc = t0 + t1; sc = sc + t2; sc = sc + dimb; sc = sc + hardb;
            if (j->best_fact[c0 + c] < 0 || sc > j->best_score[c0 + c]) {
                j->best_fact[c0 + c] = (int32_t)f; j->best_score[c0 + c] = sc;
            }
        }
    }
    for (int64_t c = 0; c < nc; c++) {            /* server.py:24861 */
        if (j->best_fact[c0 + c] >= 0)
            j->bound[c0 + c] = !(j->best_score[c0 + c] + 1e-9 < j->cthr[c0 + c]);
    }
    free(ov_row); free(lenA);
}

static void* bind_worker(void* arg) {
    bind_job* j = (bind_job*)arg;
    bind_ws w; memset(&w, 0, sizeof(w)); intern_init(&w.tab, 1 << 15); w.mask_words = 0;
    for (;;) {
        pthread_mutex_lock(&j->mu); int64_t s = j->next; j->next += 8; pthread_mutex_unlock(&j->mu);
        if (s >= j->n_sessions) break;
        for (int64_t t = s; t < s + 8 && t < j->n_sessions; t++) bind_session(j, &w, t);
    }
    intern_free(&w.tab); free(w.cps); free(w.arena); free(w.stamp); free(w.cmask);
    return NULL;
}

int dvo_bind_claims(const uint8_t* claim_utf8, const int64_t* claim_off, const uint8_t* fact_utf8, const int64_t* fact_off,
                    const int64_t* sess_claim_off, const int64_t* sess_fact_off, int64_t n_sessions,
                    const double* fact_quality, const int32_t* fact_dim, const uint8_t* fact_hard,
                    const int32_t* claim_pref_dim, const double* claim_threshold,
                    int32_t* out_best_fact, double* out_best_score, uint8_t* out_bound,
                    int32_t* out_len_claim, int32_t* out_len_fact, int32_t* out_overlap, const int64_t* pair_off,
                    int n_threads) {
    bind_job j = { claim_utf8, claim_off, fact_utf8, fact_off, sess_claim_off, sess_fact_off, n_sessions,
                   fact_quality, fact_dim, fact_hard, claim_pref_dim, claim_threshold,
                   out_best_fact, out_best_score, out_bound, out_len_claim, out_len_fact, out_overlap, pair_off,
                   0, PTHREAD_MUTEX_INITIALIZER, 0 };
    if (n_threads < 1) n_threads = 1;
    if (n_threads == 1) { bind_worker(&j); return j.err ? -1 : 0; }
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)n_threads);
    for (int i = 0; i < n_threads; i++) pthread_create(&th[i], NULL, bind_worker, &j);
    for (int i = 0; i < n_threads; i++) pthread_join(th[i], NULL);
    free(th);
    return j.err ? -1 : 0;
}

/* ------------------------------------------------------------ a4 chunking */

static int64_t rfind1(const uint32_t* t, uint32_t ch, int64_t lo, int64_t hi) {      /* str.rfind(ch, lo, hi) */
    for (int64_t i = hi - 1; i >= lo; i--) if (t[i] == ch) return i;
    return -1;
}
static int64_t rfind2(const uint32_t* t, uint32_t ch, int64_t lo, int64_t hi) {      /* str.rfind(ch+ch, lo, hi) */
    for (int64_t i = hi - 2; i >= lo; i--) if (t[i] == ch && t[i + 1] == ch) return i;
    return -1;
}

/* server.py:31932-31970.  Offsets are code-point indices.  out_tstart/out_tend give the
 * stripped text range (chunk["text"] == text[tstart:tend]); empty chunks are skipped (31952). */
int64_t dvo_chunk_build(const uint8_t* utf8, int64_t nbytes, int32_t chunk_size, int32_t overlap_cfg,
                        int32_t* out_start, int32_t* out_end, int32_t* out_tstart, int32_t* out_tend, int64_t cap) {
    uint32_t* t = (uint32_t*)malloc((size_t)(nbytes + 1) * 4);
    int64_t n = dvo_utf8_decode(utf8, nbytes, t), cnt = 0;
    int64_t third = chunk_size / 3; if (third < 0) third = 0;
    int64_t overlap = overlap_cfg < third ? overlap_cfg : third;
    int64_t half = chunk_size / 2; if (half < 200) half = 200;                   /* max(200, size // 2) */
    int all_ws = 1;
    for (int64_t i = 0; i < n; i++) if (!dvo_is_space(t[i])) { all_ws = 0; break; }
    if (all_ws) { free(t); return 0; }
    int64_t start = 0;
    while (start < n) {
        int64_t hard_end = start + chunk_size < n ? start + chunk_size : n;
        int64_t end = hard_end;
        if (hard_end < n) {
            int64_t floor_ = start + half;
            int64_t ps = rfind2(t, '\n', floor_, hard_end), ls = rfind1(t, '\n', floor_, hard_end);
            int64_t split_at = ps > ls ? ps : ls;
            if (split_at > start) end = split_at + 1 < n ? split_at + 1 : n;
        }
        int64_t a = start, b = end;
        while (a < b && dvo_is_space(t[a])) a++;
        while (b > a && dvo_is_space(t[b - 1])) b--;
        if (b > a) {
            if (cnt >= cap) { free(t); return -2; }
            out_start[cnt] = (int32_t)start; out_end[cnt] = (int32_t)end;
            out_tstart[cnt] = (int32_t)a; out_tend[cnt] = (int32_t)b; cnt++;
        }
        if (end >= n) break;
        int64_t next = end - overlap > start + 1 ? end - overlap : start + 1;
        if (next <= start) next = end;
        start = next;
    }
    free(t);
    return cnt;
}

/* --------------------------------------------------------- a6 query terms */

static int term_seen(const uint32_t* cp, const int32_t* off, int64_t n, const uint32_t* t, int32_t len) {
    for (int64_t i = 0; i < n; i++)
        if (off[i + 1] - off[i] == len && memcmp(cp + off[i], t, (size_t)len * 4) == 0) return 1;
    return 0;
}
/* server.py:32037-32051.  Terms beyond the first 80 are never observable ([:80]) and are not stored. */
int64_t dvo_query_terms(const uint8_t* utf8, int64_t nbytes, uint32_t* term_cp, int64_t cp_cap,
                        int32_t* term_off, int64_t term_cap) {
    uint32_t* cps = (uint32_t*)malloc((size_t)(nbytes + 1) * 4);
    uint32_t* low = (uint32_t*)malloc((size_t)(nbytes + 1) * 12);
    int64_t n = dvo_lower(cps, dvo_utf8_decode(utf8, nbytes, cps), low, 3 * (nbytes + 1));
    int64_t nt = 0; int32_t used = 0; int64_t lim = term_cap < 80 ? term_cap : 80;
    term_off[0] = 0;
#define ADD_TERM(ptr, len_) do { \
        if (nt < lim && !term_seen(term_cp, term_off, nt, (ptr), (len_))) { \
            if (used + (len_) > cp_cap) { free(cps); free(low); return -2; } \
            memcpy(term_cp + used, (ptr), (size_t)(len_) * 4); used += (len_); term_off[++nt] = used; } } while (0)
    int64_t i = 0;
    while (i < n && nt < lim) {
        int cls = is_ascii_tok(low[i]) ? 1 : (is_cjk(low[i]) ? 2 : 0);
        if (!cls) { i++; continue; }
        int64_t j = i;
        while (j < n && (cls == 1 ? is_ascii_tok(low[j]) : is_cjk(low[j]))) j++;
        int32_t len = (int32_t)(j - i);
        if (len >= 2) {
            ADD_TERM(low + i, len);
            if (cls == 2 && len >= 6)
                for (int32_t size = 4; size >= 3; size--)
                    for (int32_t k = 0; k + size <= len; k++) ADD_TERM(low + i + k, size);
        }
        i = j;
    }
#undef ADD_TERM
    free(cps); free(low);
    return nt;
}

/* ------------------------------------------------------ a7 chunk scoring */

static int32_t count_nonoverlap(const uint32_t* t, int64_t n, const uint32_t* p, int32_t m) {   /* str.count */
    int32_t c = 0; int64_t i = 0;
    if (m <= 0 || m > n) return 0;
    while (i + m <= n) {
        if (t[i] == p[0] && memcmp(t + i, p, (size_t)m * 4) == 0) { c++; i += m; } else i++;
    }
    return c;
}
typedef struct {
    const uint8_t* utf8; const int64_t* off; int64_t n; const uint32_t* tcp; const int32_t* toff; int64_t nt;
    int32_t* score; int64_t next; pthread_mutex_t mu;
} score_job;
static void* score_worker(void* arg) {
    score_job* j = (score_job*)arg; uint32_t *cps = NULL, *low = NULL; int64_t cap = 0;
    for (;;) {
        pthread_mutex_lock(&j->mu); int64_t c0 = j->next; j->next += 256; pthread_mutex_unlock(&j->mu);
        if (c0 >= j->n) break;
        for (int64_t c = c0; c < c0 + 256 && c < j->n; c++) {
            int64_t nb = j->off[c + 1] - j->off[c];
            if (nb + 1 > cap) { cap = 2 * (nb + 1); cps = (uint32_t*)realloc(cps, (size_t)cap * 4); low = (uint32_t*)realloc(low, (size_t)cap * 12); }
            int64_t n = dvo_lower(cps, dvo_utf8_decode(j->utf8 + j->off[c], nb, cps), low, 3 * cap);   /* 32075 */
            int32_t s = 0;
            for (int64_t t = 0; t < j->nt; t++)                                                       /* 32076 */
                s += count_nonoverlap(low, n, j->tcp + j->toff[t], j->toff[t + 1] - j->toff[t]);
            j->score[c] = s;
        }
    }
    free(cps); free(low);
    return NULL;
}
int dvo_chunk_scores(const uint8_t* chunk_utf8, const int64_t* chunk_off, int64_t n_chunks,
                     const uint32_t* term_cp, const int32_t* term_off, int64_t n_terms, int32_t* out_score, int n_threads) {
    score_job j = { chunk_utf8, chunk_off, n_chunks, term_cp, term_off, n_terms, out_score, 0, PTHREAD_MUTEX_INITIALIZER };
    if (n_threads <= 1) { score_worker(&j); return 0; }
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)n_threads);
    for (int i = 0; i < n_threads; i++) pthread_create(&th[i], NULL, score_worker, &j);
    for (int i = 0; i < n_threads; i++) pthread_join(th[i], NULL);
    free(th);
    return 0;
}

/* sorted(positive, key=(-score, index))[:k]  (server.py:32081-32086) */
int64_t dvo_topk(const int32_t* score, const int64_t* index, int64_t n, int32_t k, int64_t* out_index, int32_t* out_score) {
    int64_t m = 0;
    for (int64_t i = 0; i < n; i++) {
        if (score[i] <= 0) continue;
        int64_t idx = index ? index[i] : i, p = m;
        while (p > 0 && (out_score[p - 1] < score[i] || (out_score[p - 1] == score[i] && out_index[p - 1] > idx))) p--;
        if (p >= k) continue;
        int64_t last = m < k ? m : k - 1;
        for (int64_t q = last; q > p; q--) { out_score[q] = out_score[q - 1]; out_index[q] = out_index[q - 1]; }
        out_score[p] = score[i]; out_index[p] = idx;
        if (m < k) m++;
    }
    return m;
}

/* ------------------------------------------------- a8 question normaliser */

static int is_gate_punct(uint32_t c) {   /* server.py:20466 character class (besides \s) */
    static const uint32_t P[] = { 0xFF0C, 0x3002, 0xFF1F, 0xFF01, 0x3001, 0xFF1B, 0xFF1A, ',', '.', '?', '!', ';', ':',
                                  0xFF08, 0xFF09, '(', ')', 0x3010, 0x3011, '[', ']', '"', '\'', 0x201C, 0x201D, 0x2018, 0x2019 };
    for (size_t i = 0; i < sizeof(P) / sizeof(P[0]); i++) if (P[i] == c) return 1;
    return 0;
}
/* str.replace(src, dst): left-to-right, non-overlapping.  Returns new length or -1 on overflow. */
static int64_t replace_all(const uint32_t* in, int64_t n, const uint32_t* src, int32_t ns,
                           const uint32_t* dst, int32_t nd, uint32_t* out, int64_t cap) {
    int64_t i = 0, k = 0;
    if (ns <= 0) { if (n > cap) return -1; memcpy(out, in, (size_t)n * 4); return n; }
    while (i < n) {
        if (i + ns <= n && in[i] == src[0] && memcmp(in + i, src, (size_t)ns * 4) == 0) {
            if (k + nd > cap) return -1;
            if (nd) memcpy(out + k, dst, (size_t)nd * 4);
            k += nd; i += ns;
        } else { if (k + 1 > cap) return -1; out[k++] = in[i++]; }
    }
    return k;
}
/* server.py:20462-20469 */
int64_t dvo_normalize_question(const uint8_t* utf8, int64_t nbytes, const dvo_gate_tables* t, uint32_t* out, int64_t cap) {
    int64_t grow = 1;
    for (int32_t s = 0; s < t->n_syn; s++) {
        int32_t ns = t->syn_src_off[s + 1] - t->syn_src_off[s], nd = t->syn_dst_off[s + 1] - t->syn_dst_off[s];
        if (ns > 0 && nd > ns) grow *= (nd + ns - 1) / ns;
    }
    int64_t wcap = (3 * (nbytes + 1)) * grow + 8;
    uint32_t* a = (uint32_t*)malloc((size_t)wcap * 4); uint32_t* b = (uint32_t*)malloc((size_t)wcap * 4);
    int64_t n = dvo_utf8_decode(utf8, nbytes, a);
    int64_t lo = 0, hi = n;                                   /* .strip() */
    while (lo < hi && dvo_is_space(a[lo])) lo++;
    while (hi > lo && dvo_is_space(a[hi - 1])) hi--;
    n = dvo_lower(a + lo, hi - lo, b, wcap);                  /* .lower() */
    for (int32_t s = 0; s < t->n_syn && n >= 0; s++) {
        n = replace_all(b, n, t->syn_src + t->syn_src_off[s], t->syn_src_off[s + 1] - t->syn_src_off[s],
                        t->syn_dst + t->syn_dst_off[s], t->syn_dst_off[s + 1] - t->syn_dst_off[s], a, wcap);
        uint32_t* tmp = a; a = b; b = tmp;
    }
    if (n >= 0) { int64_t k = 0; for (int64_t i = 0; i < n; i++) if (!dvo_is_space(b[i]) && !is_gate_punct(b[i])) b[k++] = b[i]; n = k; }
    for (int32_t s = 0; s < t->n_stop && n >= 0; s++) {
        n = replace_all(b, n, t->stop + t->stop_off[s], t->stop_off[s + 1] - t->stop_off[s], NULL, 0, a, wcap);
        uint32_t* tmp = a; a = b; b = tmp;
    }
    if (n > cap) n = -2;
    if (n > 0) memcpy(out, b, (size_t)n * 4);
    free(a); free(b);
    return n;
}

/* server.py:20472-20482: distinct bigrams over ALL code points; a 1-char text is its own token */
int64_t dvo_question_keys(const uint8_t* utf8, int64_t nbytes, const dvo_gate_tables* t, uint64_t* keys, int64_t cap) {
    int64_t ncap = 3 * (nbytes + 1) * 4 + 8;
    uint32_t* norm = (uint32_t*)malloc((size_t)ncap * 4);
    int64_t n = dvo_normalize_question(utf8, nbytes, t, norm, ncap), k = 0;
    if (n < 0) { free(norm); return n; }
    if (n == 1) { if (cap < 1) { free(norm); return -2; } keys[0] = (1ULL << 42) | norm[0]; free(norm); return 1; }
    if (n - 1 > cap) { free(norm); return -2; }
    for (int64_t i = 0; i + 1 < n; i++) keys[k++] = ((uint64_t)norm[i] << 21) | norm[i + 1];
    free(norm);
    if (k > 1) {
        qsort(keys, (size_t)k, 8, cmp_u64);
        int64_t u = 1; for (int64_t i = 1; i < k; i++) if (keys[i] != keys[u - 1]) keys[u++] = keys[i];
        k = u;
    }
    return k;
}

static int32_t sorted_intersect(const uint64_t* a, int32_t na, const uint64_t* b, int32_t nb) {
    int32_t i = 0, j = 0, c = 0;
    while (i < na && j < nb) { if (a[i] < b[j]) i++; else if (a[i] > b[j]) j++; else { c++; i++; j++; } }
    return c;
}
typedef struct {
    const uint8_t* cu; const int64_t* co; int64_t nc; const uint64_t* hk; const int64_t* hko; int64_t nh;
    const dvo_gate_tables* t; double thr; int64_t* out; int64_t next; pthread_mutex_t mu;
} gate_job;
static void* gate_worker(void* arg) {
    gate_job* j = (gate_job*)arg; uint64_t* ck = NULL; int64_t ckcap = 0;
    for (;;) {
        pthread_mutex_lock(&j->mu); int64_t c0 = j->next; j->next += 16; pthread_mutex_unlock(&j->mu);
        if (c0 >= j->nc) break;
        for (int64_t c = c0; c < c0 + 16 && c < j->nc; c++) {
            int64_t nb = j->co[c + 1] - j->co[c];
            int64_t need = 3 * (nb + 1) * 4 + 8;
            if (need > ckcap) { ckcap = need; ck = (uint64_t*)realloc(ck, (size_t)ckcap * 8); }
            int64_t nk = dvo_question_keys(j->cu + j->co[c], nb, j->t, ck, ckcap);
            int64_t hit = -1;
            /* server.py:20495-20497: a blank candidate never matches */
            int blank = 1;
            {
                uint32_t* tmp = (uint32_t*)malloc((size_t)(nb + 1) * 4);
                int64_t n = dvo_utf8_decode(j->cu + j->co[c], nb, tmp);
                for (int64_t i = 0; i < n; i++) if (!dvo_is_space(tmp[i])) { blank = 0; break; }
                free(tmp);
            }
            if (!blank && nk > 0) {
                for (int64_t h = j->nh - 1; h >= 0; h--) {                       /* reversed(logs): 20498 */
                    int32_t nhk = (int32_t)(j->hko[h + 1] - j->hko[h]);
                    if (nhk == 0) continue;                                       /* 20488 */
                    int32_t in = sorted_intersect(j->hk + j->hko[h], nhk, ck, (int32_t)nk);
                    int32_t un = nhk + (int32_t)nk - in; if (un < 1) un = 1;
                    if ((double)in / (double)un >= j->thr) { hit = h; break; }    /* 20490-20491 */
                }
            }
            j->out[c] = hit;
        }
    }
    free(ck);
    return NULL;
}
int dvo_question_gate(const uint8_t* cand_utf8, const int64_t* cand_off, int64_t n_cand,
                      const uint8_t* hist_utf8, const int64_t* hist_off, int64_t n_hist,
                      const dvo_gate_tables* t, double threshold, int64_t* out_match, int n_threads) {
    int64_t total = 0;
    for (int64_t h = 0; h < n_hist; h++) total += 3 * (hist_off[h + 1] - hist_off[h] + 1) * 4 + 8;
    uint64_t* hk = (uint64_t*)malloc((size_t)(total + 1) * 8);
    int64_t* hko = (int64_t*)malloc((size_t)(n_hist + 1) * 8);
    int64_t k = 0; hko[0] = 0;
    for (int64_t h = 0; h < n_hist; h++) {
        int64_t nk = dvo_question_keys(hist_utf8 + hist_off[h], hist_off[h + 1] - hist_off[h], t, hk + k, total - k);
        if (nk < 0) { free(hk); free(hko); return -1; }
        k += nk; hko[h + 1] = k;
    }
    gate_job j = { cand_utf8, cand_off, n_cand, hk, hko, n_hist, t, threshold, out_match, 0, PTHREAD_MUTEX_INITIALIZER };
    if (n_threads <= 1) gate_worker(&j);
    else {
        pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)n_threads);
        for (int i = 0; i < n_threads; i++) pthread_create(&th[i], NULL, gate_worker, &j);
        for (int i = 0; i < n_threads; i++) pthread_join(th[i], NULL);
        free(th);
    }
    free(hk); free(hko);
    return 0;
}

/* -------------------------------------------------- a10 solution Jaccard */

typedef struct { const uint32_t* p; int32_t len; } gram;
static int cmp_gram(const void* a, const void* b) {
    const gram* x = (const gram*)a; const gram* y = (const gram*)b;
    if (x->len != y->len) return x->len < y->len ? -1 : 1;
    return memcmp(x->p, y->p, (size_t)x->len * 4);
}
/* server.py:36798-36805 on an already clean_solution_text()'d string */
static int32_t solution_grams(const uint8_t* s, int64_t nb, int32_t n, uint32_t** buf, gram** out) {
    uint32_t* cps = (uint32_t*)malloc((size_t)(nb + 1) * 4);
    uint32_t* low = (uint32_t*)malloc((size_t)(nb + 1) * 12);
    int64_t m = dvo_lower(cps, dvo_utf8_decode(s, nb, cps), low, 3 * (nb + 1)), k = 0;
    for (int64_t i = 0; i < m; i++) {                      /* drop \s, then keep [0-9a-z一-鿿] */
        uint32_t c = low[i];
        if ((c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || is_cjk(c)) low[k++] = c;
    }
    free(cps); *buf = low;
    gram* g = (gram*)malloc(sizeof(gram) * (size_t)(k + 1)); int32_t ng = 0;
    if (k == 0) { *out = g; return 0; }
    if (k <= n) { g[0].p = low; g[0].len = (int32_t)k; ng = 1; }
    else for (int64_t i = 0; i + n <= k; i++) { g[ng].p = low + i; g[ng].len = n; ng++; }
    qsort(g, (size_t)ng, sizeof(gram), cmp_gram);
    int32_t u = ng ? 1 : 0; for (int32_t i = 1; i < ng; i++) if (cmp_gram(&g[i], &g[u - 1]) != 0) g[u++] = g[i];
    *out = g; return u;
}
double dvo_solution_jaccard(const uint8_t* l, int64_t nl, const uint8_t* r, int64_t nr, int32_t ngram, int32_t* inter, int32_t* uni) {
    uint32_t *bl, *br; gram *gl, *gr;
    int32_t a = solution_grams(l, nl, ngram, &bl, &gl), b = solution_grams(r, nr, ngram, &br, &gr);
    int32_t i = 0, j = 0, c = 0;
    while (i < a && j < b) { int d = cmp_gram(&gl[i], &gr[j]); if (d < 0) i++; else if (d > 0) j++; else { c++; i++; j++; } }
    int32_t u = a + b - c;
    if (inter) *inter = c;
    if (uni) *uni = u;
    free(bl); free(br); free(gl); free(gr);
    if (a == 0 || b == 0) return 0.0;
    return (double)c / (double)(u < 1 ? 1 : u);
}
