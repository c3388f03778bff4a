// bind7.cuh — k_bind_claim_tables + k_bind_facts: the claim->fact binding kernels, cell tokenizer (included by bind.cu).
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
//   facts : valid bigram -> ONE 16-byte bucket load of the exact claim table in shared memory, four XORs and three
//           minima (entry layout below) -> on a hit a byte flag H[t] = 1 (t = dense index of the claim token; plain
//           store, no atomic).  Pieces (3..12 characters) and ASCII tokens are rare: a piece end is queued only when its
//           last bigram hit an entry that ends a claim piece; cells holding ASCII bytes are queued; both queues are
//           worked off 32 items at a time.
//   claims: the same walk inserts instead of probing; the first claim to insert a token owns it (t inside the claim's
//           own range), other claims list it as shared.  len(A) = own + shared.
//   per fact: lane c adds the flags of claim c's own range (16 flags per LDS.128 + 4 dp4a) and of its shared list,
//           scores in fp64 in the reference's operation order and keeps its running first-max.
// Texts the cell walk cannot treat exactly (U+212A, a 0xC4 lead byte (U+0130), malformed UTF-8, tables too small)
// raise sess_fallback[s]: the host re-runs exactly those sessions on k_bind_claims_general.
#pragma once
#include <stddef.h>
#include "bind_common.cuh"
#include "text.cuh"

// Cost ladder (profiles/r02_bind_ladder.json): builds with -DB7_RUNG=n keep the persistent-CTA skeleton and the claim phase
// and stop the FACT side after stage n: 0 no walk, 1 cell loads + checksum, 2 + lead masks and decode, 3 + adjacency, bigram
// id, table probe, hit flag, 4 + run positions mod 12, valid bigrams, piece ends, certification, 5 + ASCII cells and the
// deferred piece / ASCII rounds (the whole walk), 6 + per-fact epilogue = the product kernel (the default).
#ifndef B7_RUNG
#define B7_RUNG 6
#endif
#ifndef B7_NI_INSERT
#define B7_NI_INSERT 1
#endif
#ifndef B7_B_APFX
#define B7_B_APFX 1
#endif
#ifndef B7_C_PPFX
#define B7_C_PPFX 1
#endif
#define B7_NB 2048                   // bigram-table buckets of four 32-bit entries (32 KB)
#define B7_HASH 0x9E3779B1u          // id -> h = id * B7_HASH: a bijection on the 30-bit ids (mod 2^30)
#define B7_STASH 32                  // exact side list for the keys of buckets that are full
#define B7_CAPT 1024                 // piece / ASCII token slots (64-bit keys, L2-resident workspace)
#define B7_TOK_LIMIT 704
#define B7_TCAP 3072                 // dense token indices per claim group = flag bytes per warp
#define B7_HSTRIDE 4096              // per warp: B7_TCAP flag bytes, then the warp's scratch
#define B7_FT_BITS 12                // signature filter in front of the ASCII-token probes
#define B7_APFX_WORDS 48             // 38 * 38 = 1444 bits, padded to a multiple of 16 bytes
#define B7_PPFX_BITS 12
#define B7_CELL 12
#define B7_STEP (32 * B7_CELL)
#define B7_FULL 0xffffffffu
#define B7_SLOT_STASH 0x8000u        // claim-side slot numbers: [0, 4 B7_NB) table entry, | STASH side list, | TOKEN token table
#define B7_SLOT_TOKEN 0x10000u

// Bigram table entry (32 bits):  chk << 13 | piece << 12 | (t + 1)
//   h = id * B7_HASH;  bucket = bits 19..29 of h;  chk = bits 0..18 of h  — (bucket, chk) <-> id is one to one
//   t = dense index of the claim token (t + 1 in 1..B7_TCAP; 0 = entry claimed, index not published yet)
//   piece = the bigram ends a claim piece of 3..12 characters (facts look at piece ends only behind such a hit)
// A probe XORs the four entries of its bucket with k = chk << 13: the entry of the probed id — and nothing else, a key
// with chk == 0 is never stored (the session is handed back) — comes out below 8192 and IS piece << 12 | (t + 1).
struct __align__(16) B7Scratch {     // per warp
    union {
        struct { uint32_t aq[64]; uint32_t pq[160]; };                                  // cell queue, piece-end queue (< 32 + 4 x 32 per step)
        struct { unsigned long long queue[32]; uint32_t qsig[32]; uint32_t qkind; uint32_t pad_;
                 unsigned long long wset[BIND_WSET]; };                                 // exact serial count (text.cuh)
        struct { double rbest[BIND_CLAIMS]; int ridx[BIND_CLAIMS]; };                   // end of the fact phase: the warp's best per claim
    };
};

// What the claim kernel hands to the fact kernel besides the packed tables: copied as it stands (uint4s) between shared
// memory and the group's slab.
struct __align__(16) B7Head {
    uint32_t n_tab, n_tok, n_sl, ttot;             // packed section sizes: table entries, token keys, shared-list entries; token indices used
    uint16_t rbase[BIND_CLAIMS + 8];               // token-index range of claim c = [rbase[c], rbase[c + 1])
    uint16_t ownc[BIND_CLAIMS], shc[BIND_CLAIMS];  // own / shared token counts of the claims
    uint32_t ovb[B7_NB / 32];                      // bucket b has keys in the side list
    uint32_t stash_id[B7_STASH], stash_e[B7_STASH];
    uint32_t ft[1u << (B7_FT_BITS - 5)];
    uint32_t apfx[B7_APFX_WORDS];                  // first two characters of the claims' ASCII tokens (38 x 38 bits)
    uint32_t ppfx[1u << (B7_PPFX_BITS - 5)];       // (first two characters, length) of the claims' pieces, hashed
};
static_assert(sizeof(B7Head) % 16 == 0, "copied as uint4s");

struct __align__(16) B7Shared {
    uint32_t tab[B7_NB * 4];
    uint16_t tixT[B7_CAPT];                        // token-table slot -> t + 1
    uint16_t slist[B7_TCAP];                       // claim c's shared tokens at [rbase[c], rbase[c] + shc[c])
    B7Head hd;
    unsigned long long ktab[DV_KTAB];
    uint8_t lut[128];
    int next_item, ntok, ovf, exotic, wcnt[BIND_WARPS];
    uint32_t tsum[32];                             // pack: token keys per 32-slot word
    uint32_t nent, last;                           // claim kernel: bigram-table entries created so far in this group | split fact kernel: this CTA reduces
    uint16_t* oslot;                               // claim kernel: their slots, appended to the group's slab as they are created
    // LAST member: the fact kernel uses all of it (per warp [0, B7_TCAP) hit flags by token index, then its B7Scratch); the claim
    // kernel allocates only BIND_WARPS * B7_HSTRIDE_C bytes of it (per warp its B7Scratch, then B7_TCAP "already listed" BITS),
    // which takes its CTA from 70.7 to 53.8 KB: four CTAs per SM instead of three (it needs 64 registers, no spills)
    uint8_t H[BIND_WARPS][B7_HSTRIDE];
};
#define B7_HSTRIDE_C (1024 + B7_TCAP / 8)
static_assert(offsetof(B7Shared, H) % 16 == 0 && B7_HSTRIDE_C % 16 == 0, "uint4 clears");
// Slab of one claim group in the workspace (bytes): head | token bitmap | packed shared lists | packed token indices |
// packed token keys | table entries | their slots.  Sections are sized for the kernel's own limits (B7_TCAP token indices,
// B7_TOK_LIMIT token keys), so a group that passes the claim kernel always fits.
#define B7_BLOB_TOKMAP 2048
#define B7_BLOB_SLIST (B7_BLOB_TOKMAP + B7_CAPT / 8)
#define B7_BLOB_TOKTIX (B7_BLOB_SLIST + 2 * B7_TCAP)
#define B7_BLOB_TOKKEY (B7_BLOB_TOKTIX + 2 * B7_TOK_LIMIT)
#define B7_BLOB_TABENT (B7_BLOB_TOKKEY + 8 * B7_TOK_LIMIT)
#define B7_BLOB_TABSLOT (B7_BLOB_TABENT + 4 * B7_TCAP)
#define B7_BLOB_BYTES (B7_BLOB_TABSLOT + 2 * B7_TCAP)
static_assert(sizeof(B7Head) <= B7_BLOB_TOKMAP && B7_BLOB_TOKKEY % 8 == 0 && B7_BLOB_TABENT % 16 == 0 && B7_BLOB_BYTES % 16 == 0, "slab layout");
static_assert(sizeof(B7Scratch) <= B7_HSTRIDE - B7_TCAP && sizeof(B7Scratch) <= 1024, "the warp scratch lives behind the warp's flag bytes");
static_assert(sizeof(B7Shared) <= 76800, "three CTAs per SM");

#ifdef DV_EMU
static B7Shared b7_emu_shared;
#define B7_SH() (b7_emu_shared)
typedef uintptr_t b7_saddr;                         // the emulator has no shared window: plain pointers
static inline b7_saddr b7_sa(const void* p) { return (uintptr_t)p; }
static inline uint4 b7_lds128(b7_saddr a) { return *reinterpret_cast<const uint4*>(a); }
static inline void b7_sts8(b7_saddr a, uint32_t v) { *reinterpret_cast<volatile uint8_t*>(a) = (uint8_t)v; }
static inline void b7_sts8_if(bool c, b7_saddr a, uint32_t v) { if (c) *reinterpret_cast<volatile uint8_t*>(a) = (uint8_t)v; }
#else
#define B7_SH() (*reinterpret_cast<B7Shared*>(bind_smem))
// 32-bit shared-window addresses kept in registers: the compiler otherwise rebuilds the window base (S2R + MOV + LEA) at
// every access it reaches through a pointer
typedef uint32_t b7_saddr;
__device__ __forceinline__ b7_saddr b7_sa(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint4 b7_lds128(b7_saddr a) {
    uint4 v;
    asm("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));    // the table is read-only while facts are walked: free to schedule
    return v;
}
__device__ __forceinline__ void b7_sts8(b7_saddr a, uint32_t v) { asm volatile("st.shared.u8 [%0], %1;" :: "r"(a), "r"(v) : "memory"); }
// one predicated store: an `if` around the hit-flag store of the slot loop became a divergent region per slot (BSSY / BRA /
// BSYNC / WARPSYNC: 13 instructions per slot, 52 of the 180 control instructions of a step)
__device__ __forceinline__ void b7_sts8_if(bool c, b7_saddr a, uint32_t v) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p st.shared.u8 [%0], %1;\n\t}" :: "r"(a), "r"(v), "r"((uint32_t)c) : "memory");
}
#endif

struct B7Tok {                       // per-warp tokenizer state (registers)
    B7Shared* sh; unsigned long long* tkeys; uint8_t* H; uint32_t* hbits; B7Scratch* scr;   // H: fact kernel (flag bytes); hbits: claim kernel (flag bits)
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
    return ((sh.hd.ft[h >> 5] >> (h & 31)) & 1u) != 0u;
}

// ---------------------------------------------------------------- claim side: insert, token index, shared lists
__device__ __forceinline__ uint32_t* b7_entry(B7Shared& sh, uint32_t slot) {       // bigram slots only
    return (slot & B7_SLOT_STASH) ? &sh.hd.stash_e[slot & (B7_SLOT_STASH - 1u)] : &sh.tab[slot];
}
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
        if (slot & B7_SLOT_TOKEN) *(volatile uint16_t*)&sh.tixT[slot & 0xFFFFu] = (uint16_t)(tt + 1u);
        else atomicOr(b7_entry(sh, slot), tt + 1u);                // the piece bit of the entry may be raised concurrently
    }
    t.own += __popc(cm);
    bool isnew = false; uint32_t tt = 0;
    if (act && !created) {
        uint32_t v;
        for (;;) {                                                 // the creator (another warp) is a few instructions away
            v = (slot & B7_SLOT_TOKEN) ? (uint32_t)*(volatile uint16_t*)&sh.tixT[slot & 0xFFFFu] : *(volatile uint32_t*)b7_entry(sh, slot);
            if ((v & 0xFFFu) != 0u) break;
            DV_SPIN_PAUSE();
        }
        tt = (v & 0xFFFu) - 1u;
        if (tt < t.rb || tt >= t.re) {                             // another claim's token: listed once (test-and-set of its bit)
            const uint32_t bit = 1u << (tt & 31u);
            isnew = (atomicOr(&t.hbits[tt >> 5], bit) & bit) == 0u;
        }
    }
    const uint32_t sm = __ballot_sync(B7_FULL, isnew);
    if (isnew) {
        const uint32_t j = (uint32_t)t.shd + (uint32_t)__popc(sm & lt);
        if (t.rb + j < t.re) sh.slist[t.rb + j] = (uint16_t)tt; else sh.ovf = 1;
    }
    t.shd += __popc(sm);
}

// slot of bigram id (claim side): finds it or claims a free entry of its bucket, else of the side list.  -1: no room.
__device__ __forceinline__ int b7_bigram_slot(B7Shared& sh, uint32_t id, bool insert, bool& created) {
    const uint32_t h = id * B7_HASH, k = h << 13, b = (h >> 19) & (B7_NB - 1u);
    created = false;
    if (k == 0u) return -1;                                        // chk == 0 would look like a free entry
#pragma unroll
    for (int j = 0; j < 4; j++) {
        uint32_t prev = *(volatile uint32_t*)&sh.tab[b * 4u + j];
        if (prev == 0u && insert) prev = atomicCAS(&sh.tab[b * 4u + j], 0u, k);
        if (prev == 0u) { created = insert; return insert ? (int)(b * 4u + j) : -1; }
        if (((prev ^ k) >> 13) == 0u) return (int)(b * 4u + j);
    }
    if (insert) atomicOr(&sh.hd.ovb[b >> 5], 1u << (b & 31u));
    else if (!((sh.hd.ovb[b >> 5] >> (b & 31u)) & 1u)) return -1;
    for (int j = 0; j < B7_STASH; j++) {
        uint32_t prev = *(volatile uint32_t*)&sh.hd.stash_id[j];
        if (prev == 0u && insert) prev = atomicCAS(&sh.hd.stash_id[j], 0u, id);
        if (prev == 0u) { created = insert; return insert ? (int)(B7_SLOT_STASH | (uint32_t)j) : -1; }
        if (prev == id) return (int)(B7_SLOT_STASH | (uint32_t)j);
    }
    return -1;
}

// Convergent: valid lanes insert their bigram id into the bucketed table.
__device__ __forceinline__ void b7_bigram_insert(B7Tok& t, bool v, uint32_t id) {
    B7Shared& sh = *t.sh;
    const uint32_t mm = __match_any_sync(B7_FULL, v ? id : (0x80000000u | (uint32_t)t.lane));
    const bool leader = v && (__ffs(mm) - 1 == t.lane);
    int slot = -1; bool created = false;
    if (leader) {
        slot = b7_bigram_slot(sh, id, true, created);
        if (slot < 0) sh.ovf = 1;
        else if (created && !((uint32_t)slot & B7_SLOT_STASH)) {      // b7_pack walks this list instead of scanning the 8192 slots
            const uint32_t n = atomicAdd(&sh.nent, 1u);
            if (n < B7_TCAP) sh.oslot[n] = (uint16_t)slot; else sh.ovf = 1;      // cannot happen: every created entry takes a token index
        }
    }
    b7_claim_slot(t, slot >= 0, (uint32_t)slot, created);
}

// One out-of-line copy for the four slots of the claim walk (the inlined copies were 2 500 of the kernel's 8 600 instructions;
// the fact walk shares the instruction cache with them).  vid = valid << 31 | id, own_shd = own | shared << 16.
#ifdef DV_EMU
#define B7_NOINLINE
#else
#define B7_NOINLINE __noinline__
#endif
__device__ B7_NOINLINE uint32_t b7_bigram_insert_ni(B7Shared* sh, uint32_t* hbits, int lane, uint32_t rb, uint32_t re, uint32_t own_shd, uint32_t vid) {
    B7Tok t;
    t.sh = sh; t.hbits = hbits; t.lane = lane; t.rb = rb; t.re = re; t.own = (int)(own_shd & 0xFFFFu); t.shd = (int)(own_shd >> 16);
    b7_bigram_insert(t, (vid >> 31) != 0u, vid & 0x7FFFFFFFu);
    return (uint32_t)t.own | ((uint32_t)t.shd << 16);
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
            atomicOr(&sh.hd.ft[h >> 5], 1u << (h & 31));
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
    b7_claim_slot(t, ok, B7_SLOT_TOKEN | slot, created);
}

// ---------------------------------------------------------------- fact side: a miss in a full bucket
// piece << 12 | (t + 1) of id if it lives in the side list, else 0.  ~1 % of the probes get here; the loop stays rolled.
__device__ __forceinline__ uint32_t b7_side_lookup(const B7Shared* sh, uint32_t h, uint32_t id) {
    const uint32_t bk = (h >> 19) & (B7_NB - 1u);
    if (!((sh->hd.ovb[bk >> 5] >> (bk & 31u)) & 1u)) return 0u;
#pragma unroll 1
    for (int j = 0; j < B7_STASH; j++) {
        const uint32_t sid = sh->hd.stash_id[j];
        if (sid == 0u) break;
        if (sid == id) return sh->hd.stash_e[j] & 0x1FFFu;
    }
    return 0u;
}

// ---------------------------------------------------------------- fact side: probe of a piece / ASCII token key
__device__ __forceinline__ void b7_token_probe(B7Tok& t, bool v, uint64_t key) {
    if (!v) return;
    uint32_t slot = bind_hash(key) & (B7_CAPT - 1);
    for (;;) {
        const unsigned long long k = __ldcg(&t.tkeys[slot]);
        if (k == key) { t.H[(uint32_t)t.sh->tixT[slot] - 1u] = 1; return; }
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
// is keyed by walking its bytes; a 0xC4 byte (lead of U+0130) hands the session back.  Facts walk a token only when its
// first two characters start an ASCII token of the claim group (apfx): most fact tokens end after two table reads.
template <bool CLAIM>
__device__ void b7_ascii_round(B7Tok& t, int head, int cnt) {
    B7Shared& sh = *t.sh;
    const bool act = t.lane < cnt;
    const int32_t o = act ? (int32_t)t.scr->aq[head + t.lane] : 0;
    uint32_t starts = 0, w0 = 0, w1 = 0, w2 = 0, w3 = 0;
    if (act) {
        const uint32_t wp = b7_ldw(t, o - 4);
        w0 = b7_ldw(t, o); w1 = b7_ldw(t, o + 4); w2 = b7_ldw(t, o + 8); w3 = b7_ldw(t, o + 12);
        if (b7_has_byte(w0, 0xC4u) || b7_has_byte(w1, 0xC4u) || b7_has_byte(w2, 0xC4u)) t.exotic = true;
        const uint32_t al = b7_alnum4(w0) | (b7_alnum4(w1) << 4) | (b7_alnum4(w2) << 8) | (b7_alnum4(w3) << 12);
        const uint32_t pa = b7_alnum4(wp) >> 3;
        starts = al & ~((al << 1) | pa) & (al >> 1) & 0xFFFu;          // starts of tokens of >= 2 characters
    }
    while (__any_sync(B7_FULL, starts != 0u)) {
        bool has = starts != 0u;
        const int j = __ffs(starts) - 1;
        starts &= starts - 1u;
        if (has) {
            const uint32_t x = b7_fetch(w0, w1, w2, w3, (uint32_t)j);
            const uint32_t pi = (uint32_t)sh.lut[x & 0x7Fu] * 38u + (uint32_t)sh.lut[(x >> 8) & 0x7Fu];
            if (CLAIM) atomicOr(&sh.hd.apfx[pi >> 5], 1u << (pi & 31u));
            else if (B7_B_APFX) has = ((sh.hd.apfx[pi >> 5] >> (pi & 31u)) & 1u) != 0u;
        }
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
        const uint64_t key = b7_ascii_key(h1, h2, len);
        const uint32_t sig = (h1 ^ (h2 >> 7)) * 0x85EBCA6Bu;
        if (CLAIM) b7_token_insert(t, has, key, sig);
        else b7_token_probe(t, has && b7_ft_pass(sh, sig), key);
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
    if (ok) {                                                       // (first two characters, length) of the piece: facts walk it only if a claim piece has them
        const int32_t q = p - 3 * (int32_t)(L - 1u);                // bytes q .. q + 8 lie inside the text (L >= 3)
        const uint32_t* wq = reinterpret_cast<const uint32_t*>(t.base + (q & ~3));
        const uint32_t sft = ((uint32_t)q & 3u) * 8u;
        const uint32_t q0 = wq[0], q1 = wq[1], q2 = wq[2];
        const uint32_t x0 = __funnelshift_r(q0, q1, sft), x1 = __funnelshift_r(q1, q2, sft);
        const uint32_t r0 = x0 & 0xFFFFFFu, r1 = ((x0 >> 24) | (x1 << 8)) & 0xFFFFFFu;
        const uint32_t fi = ((r0 * 0x9E3779B1u) ^ (r1 * 0x85EBCA6Bu) ^ (L * 0xC2B2AE35u)) >> (32 - B7_PPFX_BITS);
        if (CLAIM) atomicOr(&sh.hd.ppfx[fi >> 5], 1u << (fi & 31u));
        else if (B7_C_PPFX) ok = ((sh.hd.ppfx[fi >> 5] >> (fi & 31u)) & 1u) != 0u;
    }
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
        if (ok) { bool cr; const int slot = b7_bigram_slot(sh, (sp << 15) | sl, false, cr); if (slot >= 0) atomicOr(b7_entry(sh, (uint32_t)slot), 1u << 12); else sh.ovf = 1; }
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
    const b7_saddr tabS = b7_sa(sh.tab), hitS = b7_sa(t.H) - 1u;                  // entry t + 1 -> flag byte t
    uint32_t ladder = 0;
    for (int s = 0; s < nsteps; s++) {
        const int32_t o = s * B7_STEP + lane * B7_CELL;
        const uint32_t w0 = n0, w1 = n1, w2 = n2, w3 = n3;
        if (s + 1 < nsteps) b7_load_cell(t, o + B7_STEP, n0, n1, n2, n3);
        if (!CLAIM && B7_RUNG == 1) { ladder ^= w0 ^ w1 ^ w2 ^ w3; continue; }
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
        if (!CLAIM && B7_RUNG == 2) { ladder ^= sy[0] ^ sy[1] ^ sy[2] ^ sy[3] ^ rare; continue; }
        // ---- adjacency, position of the lane's first character in its run
        const uint32_t pk = sy[3] ? ((sy[3] << 2) | g[3]) : 0u;
        uint32_t pv = __shfl_up_sync(B7_FULL, pk, 1);
        if (lane == 0) pv = carry_pk;
        carry_pk = __shfl_sync(B7_FULL, pk, 31);
        bool ad[4];
        ad[0] = sy[0] != 0u && pv != 0u && (pv & 3u) == g[0];
#pragma unroll
        for (int i = 1; i < 4; i++) ad[i] = sy[i] != 0u && sy[i - 1] != 0u && g[i] == g[i - 1];
        int p = 0;
        if (CLAIM || B7_RUNG >= 4) {
        const bool transparent = ad[0] && ad[1] && ad[2] && ad[3];
        const int ploc = ad[3] ? (ad[2] ? (ad[1] ? (ad[0] ? 4 : 3) : 2) : 1) : 0;
        const uint32_t T = __ballot_sync(B7_FULL, transparent);
        const uint32_t nt = ~T & lt;
        const int jn = nt ? 31 - __clz((int)nt) : 0;
        const int pj = __shfl_sync(B7_FULL, ploc, jn);
        p = nt ? pj + 4 * (lane - 1 - jn) : carryP + 4 * lane;
        p = (int)((uint32_t)p % 12u);
        }
        // ---- slots: positions, bigrams; piece ends: claims queue every end of a piece of >= 3, facts only ends whose last
        //      bigram hit an entry that ends a claim piece
        uint32_t candm = 0, candp = 0, vm = 0;
        uint32_t prev = pv >> 2;
        const int p_in = p;
        uint32_t sidem = 0;                          // slots whose bigram missed in a FULL bucket (~1 % of the probes)
        bool adn[4];                                 // the next character continues the run (last slot: possibly)
        adn[0] = ad[1]; adn[1] = ad[2]; adn[2] = ad[3]; adn[3] = ((x[3] >> 24) - 0xE5u) <= 4u;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            p = ad[i] ? p + 1 : 0;
            if (p == 12) p = 0;
            const bool v = ad[i] && (p != 0 || (!CLAIM && B7_RUNG == 3));
            const uint32_t id = (prev << 15) | sy[i];
            bool pend = v && p >= 2 && (p == 11 || !adn[i]);
            if (CLAIM) {
#if B7_NI_INSERT
                const uint32_t os = b7_bigram_insert_ni(t.sh, t.hbits, lane, t.rb, t.re, (uint32_t)t.own | ((uint32_t)t.shd << 16), (v ? 0x80000000u : 0u) | id);
                t.own = (int)(os & 0xFFFFu); t.shd = (int)(os >> 16);
#else
                b7_bigram_insert(t, v, id);
#endif
            } else {
                const uint32_t h = id * B7_HASH, k = h << 13;
                const uint4 e = b7_lds128(tabS + ((h >> 15) & ((B7_NB - 1u) << 4)));
                uint32_t m = min(min(e.x ^ k, e.y ^ k), min(e.z ^ k, e.w ^ k));
                const bool hit = v && (m - 1u) < 8191u;
                if (v && !hit && e.w != 0u) sidem |= 1u << i;                            // rare: full bucket, the key may be in the side list
                b7_sts8_if(hit, hitS + (m & 0xFFFu), 1u);
                pend = pend && hit && m >= 4096u;
                vm |= (v ? 1u : 0u) << i;
            }
            if (pend) { candm |= 1u << i; candp |= (uint32_t)p << (4 * i); }
            prev = sy[i];
        }
        if (!CLAIM && sidem) {                       // most steps have one such lane: only its flagged slots look at the side list
            int pp = p_in; uint32_t pr = pv >> 2;
#pragma unroll
            for (int i = 0; i < 4; i++) {
                pp = ad[i] ? pp + 1 : 0;
                if (pp == 12) pp = 0;
                if ((sidem >> i) & 1u) {
                    const uint32_t id = (pr << 15) | sy[i];
                    const uint32_t sm_ = b7_side_lookup(t.sh, id * B7_HASH, id);
                    if (sm_) {
                        b7_sts8(hitS + (sm_ & 0xFFFu), 1u);
                        if (pp >= 2 && (pp == 11 || !adn[i]) && sm_ >= 4096u) { candm |= 1u << i; candp |= (uint32_t)pp << (4 * i); }
                    }
                }
                pr = sy[i];
            }
        }
        if (!CLAIM && B7_RUNG == 3) continue;
        carryP = __shfl_sync(B7_FULL, p, 31);
        if (!CLAIM && !t.cert) {                 // min(len(B), 18): >= 18 distinct valid bigrams in one slot round certify it
#pragma unroll 1
            for (int i = 0; i < 4; i++) {
                const uint32_t pr = i == 0 ? (pv >> 2) : (i == 1 ? sy[0] : (i == 2 ? sy[1] : sy[2]));
                const uint32_t sc = i == 0 ? sy[0] : (i == 1 ? sy[1] : (i == 2 ? sy[2] : sy[3]));
                const bool v = ((vm >> i) & 1u) != 0u;
                const uint32_t ma = __match_any_sync(B7_FULL, v ? ((pr << 15) | sc) : (0x80000000u | (uint32_t)lane));
                const uint32_t nd = (uint32_t)__popc(__ballot_sync(B7_FULL, v && (__ffs(ma) - 1 == lane)));
                if (nd >= 18u) { t.cert = true; break; }
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
        if (!CLAIM && B7_RUNG == 4) { t.pqn = 0; continue; }
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
    if (!CLAIM && B7_RUNG < 3 && ladder == 0x9E3779B1u) t.exotic = true;          // keeps the ladder builds' loads alive
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

// out of line: short or repetitive facts only, and 500 instructions that would otherwise sit inside the fact loop
__device__ B7_NOINLINE int b7_exact_count(B7Shared& sh, B7Scratch* scr, int* wcnt, const uint8_t* tx, int32_t n, int lane) {
    scr->wset[lane] = 0ULL; scr->wset[lane + 32] = 0ULL;
    if (lane == 0) *wcnt = 0;
    __syncwarp();
    B7EmitCount emit{ scr, wcnt, false };
    dv_tokenize_serial(tx, n, scr, sh.ktab, emit);
    __syncwarp();
    const int wc = *(volatile int*)wcnt;
    __syncwarp();
    return wc < 18 ? wc : 18;
}

static inline size_t bind7_ws_u64_per_cta() { return (size_t)B7_CAPT; }
static inline size_t bind7_smem_bytes() { return sizeof(B7Shared); }
static inline size_t bind7_smem_bytes_claims() { return offsetof(B7Shared, H) + (size_t)BIND_WARPS * B7_HSTRIDE_C; }

// ---------------------------------------------------------------- pack / unpack of a group's tables (all threads of the CTA)
// Lane k of warp w owns word 32 w + k of a 256-word bitmap over the 8192 table slots; the slots of a word are read and
// written by the 32 lanes side by side (no bank conflicts), the packed entries sit in slot order.
__device__ __forceinline__ uint32_t b7_excl_scan32(uint32_t v, int lane, uint32_t& total) {
    uint32_t x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t y = __shfl_up_sync(B7_FULL, x, d); if (lane >= d) x += y; }
    total = __shfl_sync(B7_FULL, x, 31);
    return x - v;
}
__device__ void b7_pack(B7Shared& sh, const unsigned long long* tkeys, uint8_t* slab, int ncg) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t* tokmap = reinterpret_cast<uint32_t*>(slab + B7_BLOB_TOKMAP);
    uint16_t* o_slot = reinterpret_cast<uint16_t*>(slab + B7_BLOB_TABSLOT);
    uint16_t* o_sl = reinterpret_cast<uint16_t*>(slab + B7_BLOB_SLIST);
    uint16_t* o_tix = reinterpret_cast<uint16_t*>(slab + B7_BLOB_TOKTIX);
    unsigned long long* o_key = reinterpret_cast<unsigned long long*>(slab + B7_BLOB_TOKKEY);
    uint32_t* o_ent = reinterpret_cast<uint32_t*>(slab + B7_BLOB_TABENT);
    // ---- table: the slots of the entries are in the slab already (appended by their creators, in no particular order; the
    // barrier in front of this call made them visible); their entries go next to them
    const uint32_t ntab = min(sh.nent, (uint32_t)B7_TCAP);
    for (uint32_t i = tid; i < ntab; i += BIND_THREADS) o_ent[i] = sh.tab[o_slot[i]];
    // ---- token keys: warp w owns words 4 w .. 4 w + 3 of the 32-word bitmap over the 1024 key slots
    uint32_t tw[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        tw[k] = __ballot_sync(B7_FULL, tkeys[(uint32_t)(4 * warp + k) * 32u + (uint32_t)lane] != 0ULL);
        if (lane == 0) sh.tsum[4 * warp + k] = (uint32_t)__popc(tw[k]);
    }
    __syncthreads();
    uint32_t tbase = 0, ntok = 0;
    for (int j = 0; j < 32; j++) { if (j < 4 * warp) tbase += sh.tsum[j]; ntok += sh.tsum[j]; }
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t slot = (uint32_t)(4 * warp + k) * 32u + (uint32_t)lane;
        if (lane == 0) tokmap[4 * warp + k] = tw[k];
        if ((tw[k] >> lane) & 1u) {
            const uint32_t o = tbase + (uint32_t)__popc(tw[k] & lt);
            o_key[o] = tkeys[slot]; o_tix[o] = sh.tixT[slot];
        }
        tbase += (uint32_t)__popc(tw[k]);
    }
    // ---- shared lists: claim c's entries follow those of the claims before it
    if (tid <= ncg) {                                              // thread c: claim c's list; thread ncg: the total
        uint32_t before = 0;
        for (int c = 0; c < tid; c++) before += sh.hd.shc[c];
        if (tid < ncg) { const uint32_t n = sh.hd.shc[tid], rb = sh.hd.rbase[tid]; for (uint32_t j = 0; j < n; j++) o_sl[before + j] = sh.slist[rb + j]; }
        else { sh.hd.n_tab = ntab; sh.hd.n_tok = ntok; sh.hd.n_sl = before; sh.hd.ttot = sh.hd.rbase[ncg]; }
    }
    __syncthreads();
    for (uint32_t i = tid; i < sizeof(B7Head) / 16; i += BIND_THREADS) reinterpret_cast<uint4*>(slab)[i] = reinterpret_cast<const uint4*>(&sh.hd)[i];
}
__device__ void b7_unpack(B7Shared& sh, unsigned long long* tkeys, const uint8_t* slab, int ncg) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const uint32_t* tokmap = reinterpret_cast<const uint32_t*>(slab + B7_BLOB_TOKMAP);
    const uint16_t* i_sl = reinterpret_cast<const uint16_t*>(slab + B7_BLOB_SLIST);
    const uint16_t* i_tix = reinterpret_cast<const uint16_t*>(slab + B7_BLOB_TOKTIX);
    const unsigned long long* i_key = reinterpret_cast<const unsigned long long*>(slab + B7_BLOB_TOKKEY);
    const uint32_t* i_ent = reinterpret_cast<const uint32_t*>(slab + B7_BLOB_TABENT);
    const uint16_t* i_slot = reinterpret_cast<const uint16_t*>(slab + B7_BLOB_TABSLOT);
    for (uint32_t i = tid; i < sizeof(B7Head) / 16; i += BIND_THREADS) reinterpret_cast<uint4*>(&sh.hd)[i] = reinterpret_cast<const uint4*>(slab)[i];
    for (uint32_t i = tid; i < B7_NB; i += BIND_THREADS) reinterpret_cast<uint4*>(sh.tab)[i] = make_uint4(0u, 0u, 0u, 0u);
    // token keys: every slot of the CTA's key table is rewritten (key or 0)
    const uint32_t tword = tokmap[lane];
    uint32_t ttot_;
    const uint32_t tex = b7_excl_scan32((uint32_t)__popc(tword), lane, ttot_);
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const uint32_t wd = __shfl_sync(B7_FULL, tword, 4 * warp + k), base = __shfl_sync(B7_FULL, tex, 4 * warp + k);
        const uint32_t slot = (uint32_t)(4 * warp + k) * 32u + (uint32_t)lane;
        const bool on = ((wd >> lane) & 1u) != 0u;
        const uint32_t o = base + (uint32_t)__popc(wd & lt);
        tkeys[slot] = on ? i_key[o] : 0ULL;
        sh.tixT[slot] = on ? i_tix[o] : (uint16_t)0;
    }
    __syncthreads();
    const uint32_t ntab = sh.hd.n_tab;
    for (uint32_t i = tid; i < ntab; i += BIND_THREADS) sh.tab[i_slot[i]] = i_ent[i];
    if (tid < ncg) {
        uint32_t before = 0;
        for (int c = 0; c < tid; c++) before += sh.hd.shc[c];
        const uint32_t n = sh.hd.shc[tid], rb = sh.hd.rbase[tid];
        for (uint32_t j = 0; j < n; j++) sh.slist[rb + j] = i_sl[before + j];
    }
    if (tid == 0) { sh.next_item = 0; sh.exotic = 0; }               // every warp left the previous fact loop and read the flag before the barrier above
    __syncthreads();
}

template <bool CLAIMS>
__device__ __forceinline__ void b7_cta_init(B7Shared& sh, B7Tok& t, unsigned long long* tkeys) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    dv_fill_ktab(sh.ktab);
    for (int b = tid; b < 128; b += BIND_THREADS) sh.lut[b] = (uint8_t)(dv_ascii_sym((uint32_t)b) & 0xFFu);
    uint8_t* hbase = &sh.H[0][0];
    const uint32_t hbytes = BIND_WARPS * (CLAIMS ? B7_HSTRIDE_C : B7_HSTRIDE);
    for (uint32_t i = tid; i < hbytes / 16; i += BIND_THREADS) reinterpret_cast<uint4*>(hbase)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    t.sh = &sh; t.tkeys = tkeys; t.lane = lane;
    if (CLAIMS) { t.H = nullptr; t.scr = reinterpret_cast<B7Scratch*>(hbase + warp * B7_HSTRIDE_C); t.hbits = reinterpret_cast<uint32_t*>(hbase + warp * B7_HSTRIDE_C + 1024); }
    else { t.H = sh.H[warp]; t.hbits = nullptr; t.scr = reinterpret_cast<B7Scratch*>(sh.H[warp] + B7_TCAP); }
    t.base = nullptr; t.a = 0; t.total = 0; t.aqn = 0; t.pqn = 0; t.rb = 0; t.re = 0; t.own = 0; t.shd = 0; t.cert = false; t.exotic = false;
}

// ---------------------------------------------------------------- kernel 0: numbering of the claim groups
// One CTA.  The unit of work of the two kernels below is a claim GROUP (<= 32 claims of one session), not a session: 2000
// sessions over 592 resident CTAs are 3.4 rounds of which the slowest CTA runs 4; their 8000 groups are 13.5 of which it runs 14.
// sess_slab[s] = groups in front of session s (in [s_begin, s_end)), slab_sess[u] = session of group u, *slab_cursor = groups.
#define B7_GROUPS_THREADS 1024
#define B7_GROUPS_PER_THREAD 4
__global__ void __launch_bounds__(B7_GROUPS_THREADS) k_bind_groups(BindParams p) {
    __shared__ uint32_t wtot[B7_GROUPS_THREADS / 32];
    __shared__ uint32_t carry;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int64_t s0 = p.s_begin; s0 < p.s_end; s0 += B7_GROUPS_THREADS * B7_GROUPS_PER_THREAD) {
        const int64_t sa = s0 + (int64_t)tid * B7_GROUPS_PER_THREAD;          // this thread: sessions sa .. sa + 3
        uint32_t ng[B7_GROUPS_PER_THREAD], mine = 0;
#pragma unroll
        for (int k = 0; k < B7_GROUPS_PER_THREAD; k++) {
            ng[k] = 0;
            if (sa + k < p.s_end) {
                const int64_t nc = p.sess_claim_off[sa + k + 1] - p.sess_claim_off[sa + k];
                ng[k] = nc > 0 ? (uint32_t)((nc + BIND_CLAIMS - 1) / BIND_CLAIMS) : 0u;
            }
            mine += ng[k];
        }
        uint32_t tot;
        const uint32_t excl = b7_excl_scan32(mine, lane, tot);
        if (lane == 0) wtot[warp] = tot;
        __syncthreads();
        uint32_t base = carry + excl;
        for (int w = 0; w < warp; w++) base += wtot[w];
#pragma unroll
        for (int k = 0; k < B7_GROUPS_PER_THREAD; k++) {
            if (sa + k < p.s_end) {
                p.sess_slab[sa + k] = (int32_t)base;
                for (uint32_t g = 0; g < ng[k]; g++) { p.slab_sess[base + g] = (int32_t)(sa + k); p.part_cnt[base + g] = 0; }
            }
            base += ng[k];
        }
        __syncthreads();
        if (tid == B7_GROUPS_THREADS - 1) carry = base;
        __syncthreads();
    }
    if (tid == 0) *p.slab_cursor = (int)carry;
}

// ---------------------------------------------------------------- kernel 1 of 2: the claims of every group -> tables -> slab
// Two kernels instead of one fused kernel: the SM's instruction fetch is the limiter of this code (profiles/README.md,
// experiments/icache/): 24 warps streaming more than ~12 KB of code run at 2 instructions per clock per SM instead of 4, and
// one CTA's claim phase (insert paths, 60 KB of code) used to run next to the fact walk of its two neighbours.
__global__ void __launch_bounds__(BIND_THREADS, 4) k_bind_claim_tables(BindParams p) {
    B7Shared& sh = B7_SH();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned long long* tkeys = p.gtable + (size_t)blockIdx.x * (size_t)B7_CAPT;
    B7Tok t;
    b7_cta_init<true>(sh, t, tkeys);
    const int n_units = *p.slab_cursor;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x) {
        const int64_t s = p.slab_sess[u];
        const int64_t c0 = p.sess_claim_off[s], c1 = p.sess_claim_off[s + 1];
        bool handback = false;
        do {
            const int64_t cg = c0 + (int64_t)(u - p.sess_slab[s]) * BIND_CLAIMS;
            const int ncg = (int)min((int64_t)BIND_CLAIMS, c1 - cg);
            // ---- clear tables, token-index ranges of the claims
            for (uint32_t i = tid; i < B7_NB; i += BIND_THREADS) reinterpret_cast<uint4*>(sh.tab)[i] = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t i = tid; i < B7_CAPT / 8; i += BIND_THREADS) reinterpret_cast<uint4*>(sh.tixT)[i] = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t i = tid; i < sizeof(B7Head) / 16; i += BIND_THREADS) reinterpret_cast<uint4*>(&sh.hd)[i] = make_uint4(0u, 0u, 0u, 0u);
            for (uint32_t i = tid; i < B7_CAPT; i += BIND_THREADS) tkeys[i] = 0ULL;
            __syncthreads();
            if (warp == 0) {
                uint32_t bound = 0;
                if (lane < ncg) {
                    const int64_t nb = p.claim_off[cg + lane + 1] - p.claim_off[cg + lane];
                    bound = (uint32_t)min((int64_t)65535, ((nb / 3 + 2) + 15) & ~(int64_t)15);       // distinct tokens <= bytes / 3 + 1
                }
                uint32_t tot;
                const uint32_t excl = b7_excl_scan32(bound, lane, tot);
                if (tot <= B7_TCAP) { sh.hd.rbase[lane + 1] = (uint16_t)(excl + bound); if (lane == 0) sh.hd.rbase[0] = 0; }
                if (lane == 0) {
                    sh.next_item = 0; sh.ntok = 0; sh.ovf = tot > B7_TCAP ? 1 : 0; sh.exotic = 0;
                    sh.nent = 0; sh.oslot = reinterpret_cast<uint16_t*>(p.slabs + (size_t)u * B7_BLOB_BYTES + B7_BLOB_TABSLOT);
                }
            }
            __syncthreads();
            const bool too_big = sh.ovf != 0;                         // uniform: written before the barrier, next writer is behind one
            __syncthreads();
            if (too_big) { handback = true; break; }
            // ---- one warp per claim
            for (;;) {
                int c = 0;
                if (lane == 0) c = atomicAdd(&sh.next_item, 1);
                c = __shfl_sync(B7_FULL, c, 0);
                if (c >= ncg) break;
                const int64_t b = p.claim_off[cg + c], e = p.claim_off[cg + c + 1];
                t.rb = sh.hd.rbase[c]; t.re = sh.hd.rbase[c + 1]; t.own = 0; t.shd = 0;
                b7_text<true>(t, p.claim_utf8 + b, (int32_t)(e - b));
                __syncwarp();
                for (int j = lane; j < t.shd && t.rb + (uint32_t)j < t.re; j += 32) {                                     // the "already listed" bits
                    const uint32_t tt = sh.slist[t.rb + j];
                    atomicAnd(&t.hbits[tt >> 5], ~(1u << (tt & 31u)));
                }
                if (lane == 0) { sh.hd.ownc[c] = (uint16_t)t.own; sh.hd.shc[c] = (uint16_t)t.shd; }
                __syncwarp();
            }
            if (__any_sync(B7_FULL, t.exotic)) sh.exotic = 1;
            __syncthreads();
            if (sh.ovf || sh.exotic) { handback = true; break; }      // uniform; the next writers are behind the barrier of the pack
            b7_pack(sh, tkeys, p.slabs + (size_t)u * B7_BLOB_BYTES, ncg);
            __syncthreads();
        } while (0);
        t.exotic = false;
        if (handback) {
            if (tid == 0) { p.sess_fallback[s] = 1; *p.overflow = 1; p.slab_sess[u] = (int32_t)(-1 - s); }     // every thread read slab_sess[u] before the barriers above
            // the "already listed" bits may be stale after an abandoned claim phase
            __syncthreads();
            for (uint32_t i = tid; i < BIND_WARPS * B7_HSTRIDE_C / 16; i += BIND_THREADS) reinterpret_cast<uint4*>(&sh.H[0][0])[i] = make_uint4(0u, 0u, 0u, 0u);
            __syncthreads();
        }
    }
}

// ---------------------------------------------------------------- kernel 2 of 2: the facts of every session against its groups' tables
// SPLIT: a launch with fewer groups than SMs (one report per call) cuts the facts of every group into p.parts ranges, one CTA
// each; a part leaves its best (score, fact) per claim in scratch and the LAST part of a group to finish (ticket counter,
// no waiting) reduces them under the same order and writes the outputs.  The plain kernel is the same code with one part.
template <bool SPLIT>
__device__ __forceinline__ void b7_facts_body(const BindParams& p) {
    B7Shared& sh = B7_SH();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned long long* tkeys = p.gtable + (size_t)blockIdx.x * (size_t)B7_CAPT;
    B7Tok t;
    b7_cta_init<false>(sh, t, tkeys);
    const int n_units = *p.slab_cursor;
    const int parts = SPLIT ? p.parts : 1;
    for (int w = blockIdx.x; w < n_units * parts; w += gridDim.x) {
        const int u = SPLIT ? w / parts : w, part = SPLIT ? w % parts : 0;
        // a group the claim kernel handed back has no tables (its session number is stored negated); nothing writes slab_sess
        // in this kernel, so the branch is uniform.  The other groups of such a session are scored here and again by the general kernel
        if (p.slab_sess[u] < 0) continue;
        const int64_t s = p.slab_sess[u];
        const int64_t c0 = p.sess_claim_off[s], c1 = p.sess_claim_off[s + 1];
        const int64_t f0 = p.sess_fact_off[s], f1 = p.sess_fact_off[s + 1];
        bool handback = false;
        {
            const int64_t cg = c0 + (int64_t)(u - p.sess_slab[s]) * BIND_CLAIMS;
            const int ncg = (int)min((int64_t)BIND_CLAIMS, c1 - cg);
            // no barrier here: behind the barrier that ends the previous fact phase only warp 0 still works (the reduction),
            // on registers and the warps' scratch, which nothing writes before the two barriers of the unpack
            b7_unpack(sh, tkeys, p.slabs + (size_t)u * B7_BLOB_BYTES, ncg);
            // ---- one warp per fact
            double best = -1.0; int best_f = 0x7fffffff;              // lane c (< ncg) tracks claim c
            const uint32_t rb = lane < ncg ? sh.hd.rbase[lane] : 0u;
            const int no = lane < ncg ? sh.hd.ownc[lane] : 0, ns = lane < ncg ? sh.hd.shc[lane] : 0;
            const int la = no + ns;
            const int da = max(3, min(la, 14));
            const int pref = lane < ncg ? p.claim_pref_dim[cg + lane] : -1;
            const uint32_t ttot = sh.hd.rbase[ncg];
            const int nf = (int)(f1 - f0);
            const int fa = SPLIT ? (int)((int64_t)nf * part / parts) : 0, fb = SPLIT ? (int)((int64_t)nf * (part + 1) / parts) : nf;   // this CTA's facts
            for (;;) {
                int f = 0;
                if (lane == 0) f = fa + atomicAdd(&sh.next_item, 1);
                f = __shfl_sync(B7_FULL, f, 0);
                if (f >= fb) break;
                const int64_t b = p.fact_off[f0 + f], e = p.fact_off[f0 + f + 1];
                t.cert = false;
                if (B7_RUNG >= 1) b7_text<false>(t, p.fact_utf8 + b, (int32_t)(e - b));
                __syncwarp();
                if (B7_RUNG < 6) continue;
                const double fq = p.fact_quality[f0 + f];
                const int fdim = p.fact_dim[f0 + f];
                const uint32_t fhard = p.fact_hard[f0 + f];
                int lb = 18;
                if (!t.cert) lb = b7_exact_count(sh, t.scr, &sh.wcnt[warp], p.fact_utf8 + b, (int32_t)(e - b), lane);
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
            __syncwarp();
            t.scr->rbest[lane] = best; t.scr->ridx[lane] = best_f;
            __syncthreads();
            if (sh.exotic) handback = true;                          // uniform; no writer until the next barrier
            // ---- per-claim reduction over the warps
            double b = -1.0; int bf = 0x7fffffff;
            if (tid < ncg) {
                for (int wi = 0; wi < BIND_WARPS; wi++) {
                    const B7Scratch* ws = reinterpret_cast<const B7Scratch*>(sh.H[wi] + B7_TCAP);
                    const double sc = ws->rbest[tid]; const int fi = ws->ridx[tid];
                    if (sc > b || (sc == b && fi < bf)) { b = sc; bf = fi; }
                }
            }
            bool writer = true;                                       // this CTA writes the group's outputs
            if (SPLIT) {
                const size_t slot0 = ((size_t)u * (size_t)parts) * BIND_CLAIMS;
                if (tid < ncg) { p.part_score[slot0 + (size_t)part * BIND_CLAIMS + tid] = b; p.part_idx[slot0 + (size_t)part * BIND_CLAIMS + tid] = bf; }
                __threadfence();
                __syncthreads();
                if (tid == 0) sh.last = (atomicAdd(&p.part_cnt[u], 1) == parts - 1) ? 1u : 0u;
                __syncthreads();
                writer = sh.last != 0u;                               // uniform; the next writer of sh.last is behind the barriers of the unpack
                if (writer && tid < ncg) {
                    __threadfence();
                    b = -1.0; bf = 0x7fffffff;
                    for (int q = 0; q < parts; q++) {
                        const double sc = *(volatile const double*)&p.part_score[slot0 + (size_t)q * BIND_CLAIMS + tid];
                        const int fi = *(volatile const int*)&p.part_idx[slot0 + (size_t)q * BIND_CLAIMS + tid];
                        if (sc > b || (sc == b && fi < bf)) { b = sc; bf = fi; }
                    }
                }
            }
            if (writer && tid < ncg) {
                const int64_t c = cg + tid;
                const bool any = b >= 0.0;
                p.best_fact[c] = any ? bf : -1;
                p.best_score[c] = any ? b : 0.0;
                p.bound[c] = (any && !(__dadd_rn(b, 1e-9) < p.claim_threshold[c])) ? 1 : 0;
                if (p.len_claim) p.len_claim[c] = min(la, 14);        // tid == lane in warp 0
            }
        }
        t.exotic = false;
        if (handback && tid == 0) { p.sess_fallback[s] = 1; *p.overflow = 1; }
    }
}
__global__ void __launch_bounds__(BIND_THREADS, 3) k_bind_facts(BindParams p) { b7_facts_body<false>(p); }
__global__ void __launch_bounds__(BIND_THREADS, 3) k_bind_facts_split(BindParams p) { b7_facts_body<true>(p); }
