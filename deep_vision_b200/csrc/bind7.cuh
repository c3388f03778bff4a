// bind7.cuh — k_bind_claims: the fused claim->fact binding kernel, cell tokenizer (included by bind.cu).
// replaces: the scoring loop of infer_weak_evidence_refs_v3 (reference web/server.py:24813-24864) with the token
//           semantics of _tokenize_similarity_text_v3 (web/server.py:24722-24732).
//
// Shape of the work (what differs from the look-back tokenizer it replaces): a lane owns a 12-byte CELL of the text
// (a warp step = 384 bytes) and walks the <= 4 three-byte lead bytes (>= 0xE0) of its cell one after the other in
// registers — no character compaction across lanes, no symbol buffer, no history slide.  Per lead ("slot"):
//   decode -> CJK symbol; adjacency with the previous lead (3 bytes apart, both CJK); position in the run mod 12
//   (the greedy {2,12} pieces of the reference); bigram (previous, this) valid iff adjacent and position % 12 != 0.
// The position a lane starts with comes from the nearest lane below it that is not "transparent" (4 adjacent CJK
// characters continuing the previous lane's run): one ballot, one clz, one shuffle per step.
//   facts : valid bigram -> ONE 16-byte bucket load of the exact claim table in shared memory -> on a hit a byte flag
//           H[t] = 1 (t = dense index of the claim token; plain store, no atomic).  Pieces (3..12 characters) and ASCII
//           tokens are rare: a piece end is queued only when its last bigram hit an entry that ends a claim piece of the
//           same length; cells holding ASCII bytes are queued; both queues are worked off 32 items at a time.
//   claims: the same walk inserts instead of probing; the first claim to insert a token owns it (t inside the claim's
//           own range), other claims list it as shared.  len(A) = own + shared.
//   per fact: lane c adds the flags of claim c's own range (16 flags per LDS.128 + 4 dp4a) and of its shared list,
//           scores in fp64 in the reference's operation order and keeps its running first-max.
// Texts the cell walk cannot treat exactly (U+212A, a 0xC4 lead byte (U+0130), malformed UTF-8, tables too small)
// raise sess_fallback[s]: the host re-runs exactly those sessions on k_bind_claims_general.
#pragma once
#include "bind_common.cuh"
#include "text.cuh"

#define B7_CAPB 4096                 // bigram-table slots (4 ids per 16-byte bucket)
#define B7_NBM (B7_CAPB / 4 - 1)
#define B7_BSHIFT 22                 // 32 - log2(B7_CAPB / 4)
#define B7_CAPT 1024                 // piece / ASCII token slots (64-bit keys, L2-resident workspace)
#define B7_TOK_LIMIT 704
#define B7_TCAP 3072                 // dense token indices per claim group = flag bytes per warp
#define B7_FT_BITS 14                // signature filter in front of the ASCII-token probes
#define B7_SPILL 0x80000000u
#define B7_CELL 12
#define B7_STEP (32 * B7_CELL)
#define B7_FULL 0xffffffffu

struct __align__(16) B7Scratch {     // per warp
    union {
        struct { uint32_t aq[64]; uint32_t pq[160]; };                                  // cell queue, piece-end queue (< 32 + 4 x 32 per step)
        struct { unsigned long long queue[32]; uint32_t qsig[32]; uint32_t qkind; uint32_t pad_;
                 unsigned long long wset[BIND_WSET]; };                                 // exact serial count (text.cuh)
    };
};

struct __align__(16) B7Shared {
    uint32_t tab[B7_CAPB];                         // 30-bit bigram ids, 0 = free; bit 31 of a bucket's last slot = spilled
    uint8_t H[BIND_WARPS][B7_TCAP];                // facts: hit flags by token index; claims: "already listed" flags
    uint16_t tix[B7_CAPB + B7_CAPT];               // slot -> (t + 1) | piece-length flag << 12
    uint16_t slist[B7_TCAP];                       // claim c's shared tokens at [rbase[c], rbase[c] + shc[c])
    uint32_t ft[1u << (B7_FT_BITS - 5)];
    B7Scratch scr[BIND_WARPS];
    unsigned long long ktab[DV_KTAB];
    double rbest[BIND_WARPS][BIND_CLAIMS];
    int ridx[BIND_WARPS][BIND_CLAIMS];
    uint16_t rbase[BIND_CLAIMS + 1];
    uint16_t ownc[BIND_CLAIMS], shc[BIND_CLAIMS];
    uint8_t lut[128];
    int next_item, ntok, ovf, exotic, wcnt[BIND_WARPS];
};

#ifdef DV_EMU
static B7Shared b7_emu_shared;
#define B7_SH() (b7_emu_shared)
#else
#define B7_SH() (*reinterpret_cast<B7Shared*>(bind_smem))
#endif

struct B7Tok {                       // per-warp tokenizer state (registers)
    B7Shared* sh; unsigned long long* tkeys; uint8_t* H; B7Scratch* scr;
    const uint8_t* base; int32_t a, total;          // text = bytes [a, total) from the 4-byte aligned base
    int lane, aqn, pqn;
    uint32_t rb, re; int own, shd;                  // claims: own token range [rb, re), counts
    bool cert;                                      // facts: >= 18 distinct tokens certified
    bool exotic;
};

// ---- words of the text with everything outside [a, total) read as spaces
__device__ __forceinline__ uint32_t b7_ldw(const B7Tok& t, int32_t o) {
    if (o < 0 || o >= t.total) return 0x20202020u;
    uint32_t w = *reinterpret_cast<const uint32_t*>(t.base + o);
    const int lo = t.a - o, hi = t.total - o;
    if (lo > 0 || hi < 4) {
        uint32_t km = hi >= 4 ? 0xFFFFFFFFu : ((1u << (8 * hi)) - 1u);
        if (lo > 0) km &= ~((1u << (8 * lo)) - 1u);
        w = (w & km) | (0x20202020u & ~km);
    }
    return w;
}

// 4-bit mask of the bytes >= 0xE0 (leads of 3- and 4-byte sequences) of a word
__device__ __forceinline__ uint32_t b7_lead4(uint32_t w) {
    const uint32_t t = w & (w << 1) & (w << 2) & 0x80808080u;
    return (((t >> 7) * 0x00204081u) >> 21) & 0xFu;
}
// bytes q..q+3 of the 16-byte register window
__device__ __forceinline__ uint32_t b7_fetch(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3, uint32_t q) {
    const uint32_t r = q >> 2;
    const uint32_t lo = r == 0u ? w0 : (r == 1u ? w1 : w2);
    const uint32_t hi = r == 0u ? w1 : (r == 1u ? w2 : w3);
    return __funnelshift_r(lo, hi, (q & 3u) * 8u);
}
// 4-bit mask of the bytes of a word that are [A-Za-z0-9_]
__device__ __forceinline__ uint32_t b7_alnum4(uint32_t w) {
    const uint32_t v = w & 0x7F7F7F7Fu, na = ~w & 0x80808080u;          // 7-bit values; bit 7 set where the byte is ASCII
    const uint32_t f = v | 0x20202020u;                                   // letters folded to lower case
    const uint32_t let = (f + 0x1F1F1F1Fu) & ~(f + 0x05050505u);          // 0x61 <= f <= 0x7A
    const uint32_t dig = (v + 0x50505050u) & ~(v + 0x46464646u);          // 0x30 <= v <= 0x39
    const uint32_t x5f = v ^ 0x5F5F5F5Fu;
    const uint32_t und = ~(x5f + 0x7F7F7F7Fu);                            // v == 0x5F
    const uint32_t m = (let | dig | und) & na;
    return (((m >> 7) * 0x00204081u) >> 21) & 0xFu;
}
__device__ __forceinline__ bool b7_has_byte(uint32_t w, uint32_t b) {
    const uint32_t x = w ^ (b * 0x01010101u);
    return ((x - 0x01010101u) & ~x & 0x80808080u) != 0u;
}
__device__ __forceinline__ bool b7_ft_pass(const B7Shared& sh, uint32_t sig) {
    const uint32_t h = sig >> (32 - B7_FT_BITS);
    return ((sh.ft[h >> 5] >> (h & 31)) & 1u) != 0u;
}

// ---------------------------------------------------------------- claim side: insert, token index, shared lists
// Convergent.  `act` lanes hold DISTINCT slots of tokens of the warp's claim; `created` = this lane made the entry.
// A created token gets the next index of the claim's own range; a token another claim (or an earlier step of this
// claim) created is listed as shared unless it is the claim's own or already listed (flag in the warp's H array).
__device__ __forceinline__ void b7_claim_slot(B7Tok& t, bool act, uint32_t slot, bool created) {
    B7Shared& sh = *t.sh;
    const uint32_t lt = (1u << t.lane) - 1u;
    const uint32_t cm = __ballot_sync(B7_FULL, act && created);
    if (act && created) {
        uint32_t tt = t.rb + (uint32_t)t.own + (uint32_t)__popc(cm & lt);
        if (tt >= t.re) { sh.ovf = 1; tt = t.rb; }                 // cannot happen (range bound = bytes / 3 + 2); keeps waiters from spinning
        *(volatile uint16_t*)&sh.tix[slot] = (uint16_t)(tt + 1u);
    }
    t.own += __popc(cm);
    bool isnew = false; uint32_t tt = 0;
    if (act && !created) {
        uint32_t v;
        for (;;) { v = *(volatile uint16_t*)&sh.tix[slot]; if ((v & 0xFFFu) != 0u) break; DV_SPIN_PAUSE(); }   // the creator (another warp) is a few instructions away
        tt = (v & 0xFFFu) - 1u;
        if ((tt < t.rb || tt >= t.re) && !t.H[tt]) { t.H[tt] = 1; isnew = true; }
    }
    const uint32_t sm = __ballot_sync(B7_FULL, isnew);
    if (isnew) {
        const uint32_t j = (uint32_t)t.shd + (uint32_t)__popc(sm & lt);
        if (t.rb + j < t.re) sh.slist[t.rb + j] = (uint16_t)tt; else sh.ovf = 1;
    }
    t.shd += __popc(sm);
}

// Convergent: valid lanes insert their bigram id into the bucketed table.
__device__ __forceinline__ void b7_bigram_insert(B7Tok& t, bool v, uint32_t id) {
    B7Shared& sh = *t.sh;
    const uint32_t mm = __match_any_sync(B7_FULL, v ? id : (0x80000000u | (uint32_t)t.lane));
    const bool leader = v && (__ffs(mm) - 1 == t.lane);
    uint32_t slot = 0; bool created = false, ok = false;
    if (leader) {
        uint32_t b = (id * 0x9E3779B1u) >> B7_BSHIFT;
        for (uint32_t tries = 0; tries <= B7_NBM && !ok; tries++) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                if (ok) break;
                const uint32_t prev = atomicCAS(&sh.tab[b * 4u + j], 0u, id) & 0x7FFFFFFFu;
                if (prev == 0u || prev == id) { slot = b * 4u + j; created = (prev == 0u); ok = true; }
            }
            if (!ok) { atomicOr(&sh.tab[b * 4u + 3u], B7_SPILL); b = (b + 1u) & B7_NBM; }
        }
        if (!ok) sh.ovf = 1;
    }
    b7_claim_slot(t, ok, slot, created);
}

// slot of a bigram id that is in the table (claims: the last bigram of a piece was inserted before the piece is keyed)
__device__ __forceinline__ int b7_bigram_find(const B7Shared& sh, uint32_t id) {
    uint32_t b = (id * 0x9E3779B1u) >> B7_BSHIFT;
    for (uint32_t tries = 0; tries <= B7_NBM; tries++) {
        const uint4 e = *reinterpret_cast<const uint4*>(&sh.tab[b * 4u]);
        if (e.x == id) return (int)(b * 4u);
        if (e.y == id) return (int)(b * 4u + 1u);
        if (e.z == id) return (int)(b * 4u + 2u);
        if ((e.w & 0x7FFFFFFFu) == id) return (int)(b * 4u + 3u);
        if (!(e.w & B7_SPILL)) return -1;
        b = (b + 1u) & B7_NBM;
    }
    return -1;
}

// Convergent: piece / ASCII tokens of the claim (64-bit keys) go to the open-addressing table in the workspace.
__device__ __forceinline__ void b7_token_insert(B7Tok& t, bool v, uint64_t key, uint32_t sig) {
    B7Shared& sh = *t.sh;
    const unsigned long long mk = v ? (unsigned long long)key : (unsigned long long)t.lane;   // real keys have bit 62 set
    const uint32_t mm = __match_any_sync(B7_FULL, mk);
    const bool leader = v && (__ffs(mm) - 1 == t.lane);
    uint32_t slot = 0; bool created = false, ok = false;
    if (leader) {
        if (*(volatile int*)&sh.ntok >= B7_TOK_LIMIT) sh.ovf = 1;
        else {
            const uint32_t h = sig >> (32 - B7_FT_BITS);
            atomicOr(&sh.ft[h >> 5], 1u << (h & 31));
            slot = bind_hash(key) & (B7_CAPT - 1);
            for (;;) {
                const unsigned long long prev = atomicCAS(&t.tkeys[slot], 0ULL, (unsigned long long)key);
                if (prev == 0ULL) { atomicAdd(&sh.ntok, 1); created = true; break; }
                if (prev == key) break;
                slot = (slot + 1) & (B7_CAPT - 1);
            }
            ok = true;
        }
    }
    b7_claim_slot(t, ok, B7_CAPB + slot, created);
}

// piece-length flag of the entry of a piece's last bigram: 0 none, f = len - 2 (1..10), 15 = several lengths
__device__ __forceinline__ void b7_set_lenflag(B7Shared& sh, int slot, uint32_t f) {
    uint32_t* wp = reinterpret_cast<uint32_t*>(&sh.tix[slot & ~1]);
    const uint32_t shf = (uint32_t)(slot & 1) * 16u + 12u;
    for (;;) {
        const uint32_t old = *(volatile uint32_t*)wp;
        const uint32_t cur = (old >> shf) & 15u;
        const uint32_t nf = cur == 0u ? f : (cur == f ? f : 15u);
        if (nf == cur) return;
        const uint32_t nw = (old & ~(15u << shf)) | (nf << shf);
        if (atomicCAS(wp, old, nw) == old) return;
    }
}

// ---------------------------------------------------------------- fact side: probe of a piece / ASCII token key
__device__ __forceinline__ void b7_token_probe(B7Tok& t, bool v, uint64_t key) {
    if (!v) return;
    uint32_t slot = bind_hash(key) & (B7_CAPT - 1);
    for (;;) {
        const unsigned long long k = __ldcg(&t.tkeys[slot]);
        if (k == key) { t.H[(t.sh->tix[B7_CAPB + slot] & 0xFFFu) - 1u] = 1; return; }
        if (k == 0ULL) return;
        slot = (slot + 1) & (B7_CAPT - 1);
    }
}

// ---------------------------------------------------------------- deferred work: ASCII cells and piece ends, 32 at a time
__device__ __forceinline__ uint32_t b7_code3(const B7Tok& t, int32_t p) {      // symbol (code + 1) of the CJK character at byte p, else 0
    if (p < t.a || p + 3 > t.total) return 0u;
    const uint32_t b0 = t.base[p], b1 = t.base[p + 1], b2 = t.base[p + 2];
    if ((b0 & 0xF0u) != 0xE0u || (b1 & 0xC0u) != 0x80u || (b2 & 0xC0u) != 0x80u) return 0u;
    const uint32_t code = (((b0 & 0x0Fu) << 12) | ((b1 & 0x3Fu) << 6) | (b2 & 0x3Fu)) - 0x4E00u;
    return code < 20992u ? code + 1u : 0u;
}
// word at offset o (4-aligned, o < total) with the bytes at and beyond `total` read as spaces
__device__ __forceinline__ uint32_t b7_ldw_tail(const B7Tok& t, int32_t o) {
    uint32_t w = *reinterpret_cast<const uint32_t*>(t.base + o);
    const int hi = t.total - o;
    if (hi < 4) { const uint32_t km = (1u << (8 * hi)) - 1u; w = (w & km) | (0x20202020u & ~km); }
    return w;
}

// ASCII tokens are keyed by two 32-bit multiplicative hashes over the case-folded bytes (internal to this kernel: claims
// and facts of a launch use the same function; the general kernel has its own tables and keys)
#define B7_AH1 0x811C9DC5u
#define B7_AH2 0x2F0B4A37u
__device__ __forceinline__ uint64_t b7_ascii_key(uint32_t h1, uint32_t h2, uint32_t len) {
    h2 = (h2 ^ len) * 0xC2B2AE35u;
    return (1ULL << 62) | ((((uint64_t)h1 << 32) | h2) >> 2);
}

// Lanes [0, cnt) own one queued cell each: every ASCII token ([a-z0-9_]{2,} after lower-casing) that STARTS in the cell
// is keyed by walking its bytes; a 0xC4 byte (lead of U+0130) hands the session back.
template <bool CLAIM>
__device__ void b7_ascii_round(B7Tok& t, int head, int cnt) {
    B7Shared& sh = *t.sh;
    const bool act = t.lane < cnt;
    const int32_t o = act ? (int32_t)t.scr->aq[head + t.lane] : 0;
    uint32_t starts = 0;
    if (act) {
        const uint32_t wp = b7_ldw(t, o - 4), w0 = b7_ldw(t, o), w1 = b7_ldw(t, o + 4), w2 = b7_ldw(t, o + 8);
        if (b7_has_byte(w0, 0xC4u) || b7_has_byte(w1, 0xC4u) || b7_has_byte(w2, 0xC4u)) t.exotic = true;
        const uint32_t al = b7_alnum4(w0) | (b7_alnum4(w1) << 4) | (b7_alnum4(w2) << 8);
        const uint32_t pa = b7_alnum4(wp) >> 3;
        starts = al & ~((al << 1) | pa) & 0xFFFu;
    }
    while (__any_sync(B7_FULL, starts != 0u)) {
        const bool has = starts != 0u;
        const int j = __ffs(starts) - 1;
        starts &= starts - 1u;
        uint32_t h1 = B7_AH1, h2 = B7_AH2, len = 0;
        if (has) {
            const uint8_t* bp = t.base + o + j;
            const int32_t room = t.total - (o + j);
            for (;;) {
                if ((int32_t)len >= room) break;
                const uint32_t b = bp[len];
                const uint32_t code = b < 0x80u ? (uint32_t)sh.lut[b] : 0u;
                if (!code) break;
                h1 = (h1 ^ code) * 0x01000193u;
                h2 = (h2 + code) * 0x9E3779B1u;
                len++;
            }
        }
        const bool tk = has && len >= 2u;
        const uint64_t key = b7_ascii_key(h1, h2, len);
        const uint32_t sig = (h1 ^ (h2 >> 7)) * 0x85EBCA6Bu;
        if (CLAIM) b7_token_insert(t, tk, key, sig);
        else b7_token_probe(t, tk && b7_ft_pass(sh, sig), key);
    }
}

// Lanes [0, cnt) own one queued piece end each: entry = byte offset of the last character << 4 | (length - 1).
// The end is verified (the next character is not an adjacent CJK character unless the piece is 12 long), the piece is
// keyed by walking back over its characters (all CJK by construction of the position counter).
template <bool CLAIM>
__device__ void b7_piece_round(B7Tok& t, int head, int cnt) {
    B7Shared& sh = *t.sh;
    const bool act = t.lane < cnt;
    const uint32_t e = act ? t.scr->pq[head + t.lane] : 0u;
    const int32_t p = (int32_t)(e >> 4);
    const uint32_t L = (e & 15u) + 1u;
    bool ok = act;
    if (ok && L < 12u && b7_code3(t, p + 3) != 0u) ok = false;
    uint64_t mm = 0; uint32_t sp = 0, sl = 0;
    if (ok) {
        const uint8_t* bp = t.base + p - 3 * (int32_t)(L - 1u);
        for (uint32_t d = 0; d < L; d++) {
            const uint32_t b0 = bp[0], b1 = bp[1], b2 = bp[2];
            const uint32_t s = ((((b0 & 0x0Fu) << 12) | ((b1 & 0x3Fu) << 6) | (b2 & 0x3Fu)) - 0x4E00u) + 1u;
            mm += (uint64_t)s * sh.ktab[d];
            sp = sl; sl = s; bp += 3;
        }
    }
    const uint64_t key = dv_token_key(mm, L, true);
    const uint32_t sig = dv_token_sig(sp, sl, L);
    if (CLAIM) {
        b7_token_insert(t, ok, key, sig);
        if (ok) { const int slot = b7_bigram_find(sh, (sp << 15) | sl); if (slot >= 0) b7_set_lenflag(sh, slot, L - 2u); else sh.ovf = 1; }
    } else b7_token_probe(t, ok, key);
}

template <bool CLAIM>
__device__ __forceinline__ void b7_flush_queues(B7Tok& t, bool all) {
    int head = 0;
    __syncwarp();
    while (t.aqn - head >= 32 || (all && t.aqn > head)) {
        const int n = t.aqn - head < 32 ? t.aqn - head : 32;
        b7_ascii_round<CLAIM>(t, head, n);
        head += n;
    }
    if (head) {
        __syncwarp();
        const uint32_t rest = (t.lane < t.aqn - head) ? t.scr->aq[head + t.lane] : 0u;
        __syncwarp();
        t.scr->aq[t.lane] = rest;
        t.aqn -= head;
    }
    head = 0;
    while (t.pqn - head >= 32 || (all && t.pqn > head)) {
        const int n = t.pqn - head < 32 ? t.pqn - head : 32;
        b7_piece_round<CLAIM>(t, head, n);
        head += n;
    }
    if (head) {
        __syncwarp();
        const uint32_t rest = (t.lane < t.pqn - head) ? t.scr->pq[head + t.lane] : 0u;
        __syncwarp();
        t.scr->pq[t.lane] = rest;
        t.pqn -= head;
    }
    __syncwarp();
}

// ---------------------------------------------------------------- the cell walk
// Leads of 3-byte characters are >= 3 bytes apart, so each of the four 3-byte GROUPS of a cell (bytes 3i..3i+2) holds at
// most one: slot i = group i, lead offset inside the group g_i in {0, 1, 2}; no search loop, constant-shift windows.
__device__ __forceinline__ void b7_load_cell(const B7Tok& t, int32_t o, uint32_t& w0, uint32_t& w1, uint32_t& w2, uint32_t& w3) {
    if (o + 16 <= t.total) {
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(t.base + o);
        w0 = wp[0]; w1 = wp[1]; w2 = wp[2]; w3 = wp[3];
    } else {
        w0 = o < t.total ? b7_ldw_tail(t, o) : 0x20202020u;
        w1 = o + 4 < t.total ? b7_ldw_tail(t, o + 4) : 0x20202020u;
        w2 = o + 8 < t.total ? b7_ldw_tail(t, o + 8) : 0x20202020u;
        w3 = o + 12 < t.total ? b7_ldw_tail(t, o + 12) : 0x20202020u;
    }
}
__device__ __forceinline__ uint32_t b7_gather4(uint32_t m) { return (((m >> 7) * 0x00204081u) >> 21) & 0xFu; }   // bits 7,15,23,31 -> 0..3

template <bool CLAIM>
__device__ void b7_text(B7Tok& t, const uint8_t* tx, int32_t n) {
    B7Shared& sh = *t.sh;
    const int lane = t.lane;
    const uint32_t lt = (1u << lane) - 1u;
    t.a = (int32_t)((uintptr_t)tx & 3u);
    t.base = tx - t.a;
    t.total = t.a + n;
    t.aqn = 0; t.pqn = 0;
    if (n <= 0) return;
    const int nsteps = (t.total + B7_STEP - 1) / B7_STEP;
    uint32_t carry_pk = 0;                         // previous step's last group: symbol << 2 | g, 0 = no CJK character there
    int carryP = 0;                                // position in the run (mod 12) of the previous step's last character
    uint32_t n0, n1, n2, n3;
    b7_load_cell(t, lane * B7_CELL, n0, n1, n2, n3);
    if (t.a && lane == 0) { const uint32_t km = (1u << (8 * t.a)) - 1u; n0 = (n0 & ~km) | (0x20202020u & km); }
    const uint4* tab4 = reinterpret_cast<const uint4*>(sh.tab);
    const uint2* tix2 = reinterpret_cast<const uint2*>(sh.tix);
    for (int s = 0; s < nsteps; s++) {
        const int32_t o = s * B7_STEP + lane * B7_CELL;
        const uint32_t w0 = n0, w1 = n1, w2 = n2, w3 = n3;
        if (s + 1 < nsteps) b7_load_cell(t, o + B7_STEP, n0, n1, n2, n3);
        // ---- cell masks: leads >= 0xE0 (12 bits), any ASCII byte or 2-byte lead
        const uint32_t u0 = w0 & (w0 << 1), u1 = w1 & (w1 << 1), u2 = w2 & (w2 << 1);
        const uint32_t mm = b7_gather4(u0 & (w0 << 2) & 0x80808080u) | (b7_gather4(u1 & (w1 << 2) & 0x80808080u) << 4) |
                            (b7_gather4(u2 & (w2 << 2) & 0x80808080u) << 8);
        const uint32_t rare = (~(w0 & w1 & w2) | (u0 & ~(w0 << 2)) | (u1 & ~(w1 << 2)) | (u2 & ~(w2 << 2))) & 0x80808080u;
        bool bad = (((mm & (mm >> 1)) & 0x6DBu) | ((mm & (mm >> 2)) & 0x249u)) != 0u;      // two leads in one group: malformed
        // ---- slots: windows, decode
        const uint32_t v1l = __funnelshift_r(w0, w1, 24), v1h = __funnelshift_r(w1, w2, 24);
        const uint32_t v2l = __funnelshift_r(w1, w2, 16), v2h = __funnelshift_r(w2, w3, 16);
        const uint32_t v3l = __funnelshift_r(w2, w3, 8), v3h = w3 >> 8;
        uint32_t g[4], x[4], sy[4];
        g[0] = (mm >> 1) & 3u; g[1] = (mm >> 4) & 3u; g[2] = (mm >> 7) & 3u; g[3] = (mm >> 10) & 3u;     // bits {1,2,4} >> 1
        x[0] = __funnelshift_r(w0, w1, g[0] * 8u); x[1] = __funnelshift_r(v1l, v1h, g[1] * 8u);
        x[2] = __funnelshift_r(v2l, v2h, g[2] * 8u); x[3] = __funnelshift_r(v3l, v3h, g[3] * 8u);
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t xi = x[i];
            const uint32_t code = (((xi & 0x0Fu) << 12) | ((xi >> 2) & 0xFC0u) | ((xi >> 16) & 0x3Fu)) - 0x4E00u;
            sy[i] = ((xi & 0x00C0C0F0u) == 0x008080E0u && code < 20992u) ? code + 1u : 0u;
            bad |= (xi & 0x00FFFFFFu) == 0x00AA84E2u;                                    // U+212A lower-cases to 'k'
        }
        if (bad) t.exotic = true;
        // ---- adjacency, position of the lane's first character in its run
        const uint32_t pk = sy[3] ? ((sy[3] << 2) | g[3]) : 0u;
        uint32_t pv = __shfl_up_sync(B7_FULL, pk, 1);
        if (lane == 0) pv = carry_pk;
        carry_pk = __shfl_sync(B7_FULL, pk, 31);
        bool ad[4];
        ad[0] = sy[0] != 0u && pv != 0u && (pv & 3u) == g[0];
#pragma unroll
        for (int i = 1; i < 4; i++) ad[i] = sy[i] != 0u && sy[i - 1] != 0u && g[i] == g[i - 1];
        const bool transparent = ad[0] && ad[1] && ad[2] && ad[3];
        const int ploc = ad[3] ? (ad[2] ? (ad[1] ? (ad[0] ? 4 : 3) : 2) : 1) : 0;
        const uint32_t T = __ballot_sync(B7_FULL, transparent);
        const uint32_t nt = ~T & lt;
        const int jn = nt ? 31 - __clz((int)nt) : 0;
        const int pj = __shfl_sync(B7_FULL, ploc, jn);
        int p = nt ? pj + 4 * (lane - 1 - jn) : carryP + 4 * lane;
        p = (int)((uint32_t)p % 12u);
        // ---- slots: positions, bigrams
        uint32_t candm = 0, candp = 0, lfp = 0, pp = 0;
        uint32_t prev = pv >> 2;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            p = ad[i] ? p + 1 : 0;
            if (p == 12) p = 0;
            const bool v = ad[i] && p != 0;
            const uint32_t id = (prev << 15) | sy[i];
            pp |= (uint32_t)p << (4 * i);
            if (CLAIM) b7_bigram_insert(t, v, id);
            else {
                if (!t.cert) {                    // min(len(B), 18): >= 18 distinct valid bigrams in one slot round certify it
                    const uint32_t ma = __match_any_sync(B7_FULL, v ? id : (0x80000000u | (uint32_t)lane));
                    const uint32_t nd = (uint32_t)__popc(__ballot_sync(B7_FULL, v && (__ffs(ma) - 1 == lane)));
                    if (nd >= 18u) t.cert = true;
                }
                uint32_t b = (id * 0x9E3779B1u) >> B7_BSHIFT;
                const uint4 e = tab4[b];
                const uint2 tq = tix2[b];
                const bool m1 = e.y == id, m2 = e.z == id, m3 = (e.w & 0x7FFFFFFFu) == id;
                const uint32_t wsel = (m2 || m3) ? tq.y : tq.x;
                uint32_t t16 = (m1 || m3) ? (wsel >> 16) : (wsel & 0xFFFFu);
                bool hit = v && (e.x == id || m1 || m2 || m3);
                if (v && !hit && (e.w & B7_SPILL)) {                                     // rare: the chain continues in the next bucket(s)
                    for (;;) {
                        b = (b + 1u) & B7_NBM;
                        const uint4 f = tab4[b];
                        const int j = f.x == id ? 0 : (f.y == id ? 1 : (f.z == id ? 2 : ((f.w & 0x7FFFFFFFu) == id ? 3 : -1)));
                        if (j >= 0) { t16 = sh.tix[b * 4u + (uint32_t)j]; hit = true; break; }
                        if (!(f.w & B7_SPILL)) break;
                    }
                }
                if (hit) { t.H[(t16 & 0xFFFu) - 1u] = 1; lfp |= (t16 >> 12) << (4 * i); }
            }
            prev = sy[i];
        }
        carryP = __shfl_sync(B7_FULL, p, 31);
        // ---- piece ends: claims queue every end of a piece of >= 3, facts only ends whose last bigram hit an entry that
        //      ends a claim piece of the same length
        if (CLAIM || lfp != 0u) {
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const uint32_t pi = (pp >> (4 * i)) & 15u;
                const uint32_t lf = (lfp >> (4 * i)) & 15u;
                const bool ends = pi == 11u || (i < 3 ? !ad[i < 3 ? i + 1 : 3] : ((x[3] >> 24) - 0xE5u) > 4u);
                const bool want = CLAIM ? true : (lf == 15u || lf == pi - 1u);          // flag = length - 2 = position - 1
                if (ad[i] && pi >= 2u && ends && want) { candm |= 1u << i; candp |= pi << (4 * i); }
            }
        }
        while (__any_sync(B7_FULL, candm != 0u)) {
            const bool c = candm != 0u;
            const int i = __ffs(candm) - 1;
            candm &= candm - 1u;
            const uint32_t bal = __ballot_sync(B7_FULL, c);
            if (c) {
                const uint32_t gi = i == 0 ? g[0] : (i == 1 ? g[1] : (i == 2 ? g[2] : g[3]));
                t.scr->pq[t.pqn + __popc(bal & lt)] = ((uint32_t)(o + 3 * i + (int32_t)gi) << 4) | ((candp >> (4 * i)) & 15u);
            }
            t.pqn += __popc(bal);
        }
        // ---- cells holding ASCII bytes or 2-byte leads are looked at 32 at a time
        {
            const bool need = rare != 0u && o < t.total;
            const uint32_t bal = __ballot_sync(B7_FULL, need);
            if (bal) {
                if (need) t.scr->aq[t.aqn + __popc(bal & lt)] = (uint32_t)o;
                t.aqn += __popc(bal);
            }
        }
        // the one call site of the deferred rounds: whenever a queue holds a full round, and at the end of the text
        if (t.aqn >= 32 || t.pqn >= 32 || s + 1 == nsteps) b7_flush_queues<CLAIM>(t, s + 1 == nsteps);
    }
}

// exact min(len(B), 18) of a fact the walk could not certify (short or repetitive facts): the serial tokenizer of
// text.cuh counts distinct keys in a 64-slot set until 18 were seen
struct B7EmitCount {
    B7Scratch* scr; int* wcnt; bool full;
    __device__ __forceinline__ void restart() {}
    __device__ __forceinline__ void count(bool v, uint64_t key) {
        if (v && !full) {
            uint32_t s = (bind_hash(key) >> 20) & (BIND_WSET - 1);
            for (;;) {
                const unsigned long long prev = atomicCAS(&scr->wset[s], 0ULL, (unsigned long long)key);
                if (prev == 0ULL) { atomicAdd(wcnt, 1); break; }
                if (prev == key) break;
                s = (s + 1) & (BIND_WSET - 1);
            }
        }
        __syncwarp();
        full = (*(volatile int*)wcnt >= 18);
    }
    __device__ __forceinline__ void bigram(bool v, uint32_t id) { count(v, (uint64_t)id); }
    __device__ __forceinline__ bool need_token(uint32_t) const { return true; }
    __device__ __forceinline__ void token(bool v, uint64_t key, uint32_t) { count(v, key); }
};

static inline size_t bind7_ws_u64_per_cta() { return (size_t)B7_CAPT; }
static inline size_t bind7_smem_bytes() { return sizeof(B7Shared); }

__global__ void __launch_bounds__(BIND_THREADS, 3) k_bind_claims(BindParams p) {
    B7Shared& sh = B7_SH();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned long long* tkeys = p.gtable + (size_t)blockIdx.x * (size_t)B7_CAPT;
    dv_fill_ktab(sh.ktab);
    for (int b = tid; b < 128; b += BIND_THREADS) sh.lut[b] = (uint8_t)(dv_ascii_sym((uint32_t)b) & 0xFFu);
    for (uint32_t i = tid; i < BIND_WARPS * B7_TCAP / 16; i += BIND_THREADS) reinterpret_cast<uint4*>(&sh.H[0][0])[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    B7Tok t;
    t.sh = &sh; t.tkeys = tkeys; t.H = sh.H[warp]; t.scr = &sh.scr[warp]; t.lane = lane;
    t.base = nullptr; t.a = 0; t.total = 0; t.aqn = 0; t.pqn = 0; t.rb = 0; t.re = 0; t.own = 0; t.shd = 0; t.cert = false; t.exotic = false;

    for (int64_t s = p.s_begin + blockIdx.x; s < p.s_end; s += gridDim.x) {
        const int64_t c0 = p.sess_claim_off[s], c1 = p.sess_claim_off[s + 1];
        const int64_t f0 = p.sess_fact_off[s], f1 = p.sess_fact_off[s + 1];
        bool handback = false;
        for (int64_t cg = c0; cg < c1 && !handback; cg += BIND_CLAIMS) {
            const int ncg = (int)min((int64_t)BIND_CLAIMS, c1 - cg);
            // ---- clear tables, token-index ranges of the claims
            for (uint32_t i = tid; i < (B7_CAPB >> 2); i += BIND_THREADS) reinterpret_cast<uint4*>(sh.tab)[i] = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t i = tid; i < (B7_CAPB + B7_CAPT) / 8; i += BIND_THREADS) reinterpret_cast<uint4*>(sh.tix)[i] = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t i = tid; i < B7_CAPT; i += BIND_THREADS) tkeys[i] = 0ULL;
            for (uint32_t i = tid; i < (1u << (B7_FT_BITS - 5)); i += BIND_THREADS) sh.ft[i] = 0u;
            if (warp == 0) {
                uint32_t bound = 0;
                if (lane < ncg) {
                    const int64_t nb = p.claim_off[cg + lane + 1] - p.claim_off[cg + lane];
                    bound = (uint32_t)min((int64_t)65535, ((nb / 3 + 2) + 15) & ~(int64_t)15);       // distinct tokens <= bytes / 3 + 1
                }
                uint32_t incl = bound;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(B7_FULL, incl, d); if (lane >= d) incl += y; }
                const uint32_t tot = __shfl_sync(B7_FULL, incl, 31);
                if (tot <= B7_TCAP) { sh.rbase[lane + 1] = (uint16_t)incl; if (lane == 0) sh.rbase[0] = 0; }
                if (lane == 0) { sh.next_item = 0; sh.ntok = 0; sh.ovf = tot > B7_TCAP ? 1 : 0; sh.exotic = 0; }
            }
            __syncthreads();
            const bool too_big = sh.ovf != 0;                         // uniform: written before the barrier, next writer is behind one
            __syncthreads();
            if (too_big) { handback = true; break; }
            // ---- phase C: one warp per claim
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&sh.next_item, 1);
                c = __shfl_sync(B7_FULL, c, 0);
                if (c >= ncg) break;
                const int64_t b = p.claim_off[cg + c], e = p.claim_off[cg + c + 1];
                t.rb = sh.rbase[c]; t.re = sh.rbase[c + 1]; t.own = 0; t.shd = 0;
                b7_text<true>(t, p.claim_utf8 + b, (int32_t)(e - b));
                __syncwarp();
                for (int j = lane; j < t.shd && t.rb + (uint32_t)j < t.re; j += 32) t.H[sh.slist[t.rb + j]] = 0;      // the "already listed" flags
                if (lane == 0) { sh.ownc[c] = (uint16_t)t.own; sh.shc[c] = (uint16_t)t.shd; }
                __syncwarp();
            }
            if (__any_sync(B7_FULL, t.exotic)) sh.exotic = 1;
            __syncthreads();
            const bool give_up = sh.ovf || sh.exotic;
            if (tid == 0) sh.next_item = 0;
            __syncthreads();
            if (give_up) { handback = true; break; }
            // ---- phase F: one warp per fact
            double best = -1.0; int best_f = 0x7fffffff;              // lane c (< ncg) tracks claim c
            const uint32_t rb = lane < ncg ? sh.rbase[lane] : 0u;
            const int no = lane < ncg ? sh.ownc[lane] : 0, ns = lane < ncg ? sh.shc[lane] : 0;
            const int la = no + ns;
            const int da = max(3, min(la, 14));
            const int pref = lane < ncg ? p.claim_pref_dim[cg + lane] : -1;
            const uint32_t ttot = sh.rbase[ncg];
            const int nf = (int)(f1 - f0);
            for (;;) {
                int f = 0;
                if (lane == 0) f = atomicAdd(&sh.next_item, 1);
                f = __shfl_sync(B7_FULL, f, 0);
                if (f >= nf) break;
                const int64_t b = p.fact_off[f0 + f], e = p.fact_off[f0 + f + 1];
                t.cert = false;
                b7_text<false>(t, p.fact_utf8 + b, (int32_t)(e - b));
                __syncwarp();
                const double fq = p.fact_quality[f0 + f];
                const int fdim = p.fact_dim[f0 + f];
                const uint32_t fhard = p.fact_hard[f0 + f];
                int lb = 18;
                if (!t.cert) {
                    B7Scratch* scr = t.scr;
                    scr->wset[lane] = 0ULL; scr->wset[lane + 32] = 0ULL;
                    if (lane == 0) sh.wcnt[warp] = 0;
                    __syncwarp();
                    B7EmitCount emit{ scr, &sh.wcnt[warp], false };
                    dv_tokenize_serial(p.fact_utf8 + b, (int32_t)(e - b), scr, sh.ktab, emit);
                    __syncwarp();
                    { const int wc = *(volatile int*)&sh.wcnt[warp]; lb = wc < 18 ? wc : 18; }
                    __syncwarp();
                }
                // ---- overlap of claim `lane` with this fact: its own index range + its shared list
                uint32_t ov = 0;
                if (lane < ncg) {
                    const uint8_t* hp = t.H + rb;
                    for (int j = 0; j < no; j += 16) {
                        const uint4 h4 = *reinterpret_cast<const uint4*>(hp + j);
                        ov = __dp4a(h4.x, 0x01010101u, ov); ov = __dp4a(h4.y, 0x01010101u, ov);
                        ov = __dp4a(h4.z, 0x01010101u, ov); ov = __dp4a(h4.w, 0x01010101u, ov);
                    }
                    const uint16_t* sl = sh.slist + rb;
                    for (int j = 0; j < ns; j++) ov += t.H[sl[j]];
                }
                __syncwarp();
                for (uint32_t i = lane * 16u; i < ttot; i += 512u) *reinterpret_cast<uint4*>(t.H + i) = make_uint4(0u, 0u, 0u, 0u);
                if (lane == 0 && p.len_fact && cg == c0) p.len_fact[f0 + f] = lb;
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
            }
            if (__any_sync(B7_FULL, t.exotic)) sh.exotic = 1;
            sh.rbest[warp][lane] = best; sh.ridx[warp][lane] = best_f;
            __syncthreads();
            if (sh.exotic) handback = true;                          // uniform; no writer until the next barrier
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
                if (p.len_claim) p.len_claim[c] = min((int)sh.ownc[tid] + (int)sh.shc[tid], 14);
            }
            __syncthreads();
        }
        t.exotic = false;
        if (handback) {
            if (tid == 0) { p.sess_fallback[s] = 1; *p.overflow = 1; }
            // the H arrays may hold stale flags of an abandoned claim phase
            __syncthreads();
            for (uint32_t i = tid; i < BIND_WARPS * B7_TCAP / 16; i += BIND_THREADS) reinterpret_cast<uint4*>(&sh.H[0][0])[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
        }
    }
}
