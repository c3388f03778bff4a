// text.cuh — warp-level UTF-8/CJK decode + similarity-token key generation (kernel stage K1).
//
// Follows _tokenize_similarity_text_v3 (reference web/server.py:24722-24732):
//   lower(); ASCII tokens [a-z0-9_]{2,}; CJK runs split greedily into pieces of <=12 chars,
//   each piece (len>=2) and all bigrams inside a piece are tokens.
//
// One warp streams one text.
//   pass A  coalesced 16-byte loads (512 B per iteration).  Each lane finds the bytes at which a
//           symbol is produced with SWAR bit tricks on its four words (ASCII bytes, 2-/4-byte
//           leads = separators, third byte of a 3-byte char), then decodes one symbol per set bit
//           from a conflict-free shared-memory copy of its bytes; one warp prefix sum compacts
//           the SYMBOLS (u16: 0 = separator, 1..20992 = CJK code, 0x8000|code = ASCII) densely.
//   pass B  32 symbols per step, one per lane: run starts by ballot/clz, position inside the
//           12-char piece, exact bigram ids; piece / ASCII-run tokens are detected at their last
//           symbol and their 64-bit multilinear key  M = sum_d code_d * K(d)  is computed (a short
//           loop over the symbol history) only for tokens the consumer asks for — every token
//           of a claim, but only filter survivors of a fact.
// Tokens are emitted warp-convergently: emit.bigram(valid, id30), emit.token(valid, key64, sig).
//
// Restart on the exact serial path dv_tokenize_serial() when the text contains
//   * a byte that may lower-case INTO the ASCII class from outside it (U+0130 -> 'i'+U+0307,
//     U+212A -> 'k'; detected as lead 0xC4 / bytes E2 84), or
//   * a needed ASCII token longer than the 64-symbol history.
#pragma once
#include <stdint.h>

#define DV_SYM_ASCII 0x8000u
#define DV_TILE   512                    // bytes staged per iteration (32 lanes x 16 B)
#define DV_HIST   64                     // processed symbols kept for key loops (max fast-path token length)
#define DV_SYMCAP (DV_HIST + DV_TILE + 8) // history + held-back + new symbols + terminator, padded

struct __align__(16) dv_warp_buf {
    uint32_t raw[32 * 5];                // per lane: [previous word, 4 data words]; stride 5 words = conflict-free
    uint16_t sym[DV_SYMCAP];
    unsigned long long queue[32];        // serial-path key queue
    uint32_t qsig[32];                   // serial-path token signatures
    uint32_t qkind;                      // bit i: queue[i] is a bigram id
};

__host__ __device__ __forceinline__ uint64_t dv_splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
#define DV_KTAB 64
#define DV_KLEN_SEED 0x4C454EULL
// ktab[d] = K(d) for d < DV_KTAB (shared memory, filled by dv_fill_ktab); larger d computed on the fly
__device__ __forceinline__ void dv_fill_ktab(unsigned long long* ktab) {
    for (int d = threadIdx.x; d < DV_KTAB; d += blockDim.x) ktab[d] = dv_splitmix64((uint64_t)d + 1) | 1ULL;
}
__device__ __forceinline__ uint64_t dv_kpos(uint32_t d, const unsigned long long* ktab) {
    return d < DV_KTAB ? ktab[d] : (dv_splitmix64((uint64_t)d + 1) | 1ULL);
}
__device__ __forceinline__ uint64_t dv_token_key(uint64_t m, uint32_t len, bool cjk) {
    const uint64_t klen = dv_splitmix64(DV_KLEN_SEED) | 1ULL;          // folded to a constant by the compiler
    return ((cjk ? 3ULL : 1ULL) << 62) | ((m + (uint64_t)len * klen) >> 2);
}

__device__ __forceinline__ uint32_t dv_ascii_sym(uint32_t b) {          // b < 0x80
    uint32_t l = b | 0x20u;
    if (l - 'a' < 26u) return DV_SYM_ASCII | (l - 'a' + 1u);
    if (b - '0' < 10u) return DV_SYM_ASCII | (b - '0' + 27u);
    return b == '_' ? (DV_SYM_ASCII | 37u) : 0u;
}
__device__ __forceinline__ bool dv_is_cjk_sym(uint32_t s) { return s - 1u < 20992u; }
__device__ __forceinline__ bool dv_is_ascii_sym(uint32_t s) { return s >= DV_SYM_ASCII; }

__device__ __forceinline__ uint32_t dv_token_sig(uint32_t prev_sym, uint32_t last_sym, uint32_t len) {
    return (prev_sym * 0x9E3779B1u) ^ (last_sym * 0x85EBCA6Bu) ^ (len * 0xC2B2AE35u);   // filter signature (top bits used)
}

// ---- pass A building blocks (shared by the tokenizer and the symboliser in chunk.cu)
// Load this lane's 16 bytes of the tile, stash [prev word, 4 words] in R, return the 16-bit mask of
// bytes at which a symbol is produced: ASCII bytes, 2-/4-byte leads (separator at the lead) and the
// third byte of 3-byte chars.  Bytes outside [0, n) never produce symbols.
__device__ __forceinline__ uint32_t dv_tile_masks(const uint8_t* text, int32_t n, int32_t tile, int lane,
                                                  uint32_t& carry_word, uint32_t* R) {
    const int32_t pos = tile + lane * 16;
    uint4 w = make_uint4(0, 0, 0, 0);
    uint32_t inmask = 0;
    if (pos < n && pos + 16 > 0) {
        w = *reinterpret_cast<const uint4*>(text + pos);
        const int lo = pos < 0 ? -pos : 0, hi = (n - pos) < 16 ? (n - pos) : 16;
        inmask = (0xFFFFu >> (16 - hi)) & (0xFFFFu << lo);
    }
    uint32_t prev = __shfl_up_sync(0xffffffffu, w.w, 1);
    if (lane == 0) prev = carry_word;
    carry_word = __shfl_sync(0xffffffffu, w.w, 31);
    R[0] = prev; R[1] = w.x; R[2] = w.y; R[3] = w.z; R[4] = w.w;
    uint32_t m16 = 0;
    const uint32_t ws[5] = { prev, w.x, w.y, w.z, w.w };
    uint32_t l3_prev;
    {
        const uint32_t x = ws[0], Hh = x & 0x80808080u, b6 = (x << 1) & 0x80808080u, b5 = (x << 2) & 0x80808080u, b4 = (x << 3) & 0x80808080u;
        l3_prev = Hh & b6 & b5 & ~b4;
    }
#pragma unroll
    for (int k = 1; k < 5; k++) {
        const uint32_t x = ws[k], Hh = x & 0x80808080u, b6 = (x << 1) & 0x80808080u, b5 = (x << 2) & 0x80808080u, b4 = (x << 3) & 0x80808080u;
        const uint32_t lead = Hh & b6, cont = Hh & ~b6;
        const uint32_t l3 = lead & b5 & ~b4;
        const uint32_t l24 = lead & ~(b5 & ~b4);                                  // 2-byte or 4-byte lead: separator at the lead
        const uint32_t l3s = (l3 << 16) | (l3_prev >> 16);                       // lead3 two bytes earlier
        const uint32_t ends = (~x & 0x80808080u) | l24 | (cont & l3s);
        m16 |= ((((ends >> 7) * 0x00204081u) >> 21) & 0xFu) << ((k - 1) * 4);    // gather bits 0,8,16,24 -> 4 bits
        l3_prev = l3;
    }
    return m16 & inmask;
}
// Symbol produced at byte t (a set bit of dv_tile_masks) of this lane.  `c4` is set for a 0xC4 lead
// (U+0100..U+013F: U+0130 lower-cases to 'i' + U+0307, which needs the serial path).
__device__ __forceinline__ uint32_t dv_decode_sym(const uint32_t* R, int t, bool& c4) {
    const int ao = t + 1;                                               // offset of byte t-3 in the lane's 20-byte block
    const uint32_t x = __funnelshift_r(R[ao >> 2], R[(ao >> 2) + 1], (ao & 3) * 8);   // bytes t-3..t, byte t on top
    const uint32_t b = x >> 24, p1 = (x >> 16) & 0xFFu, p2 = (x >> 8) & 0xFFu;
    const uint32_t code = (((p2 & 0x0Fu) << 12) | ((p1 & 0x3Fu) << 6) | (b & 0x3Fu)) - 0x4E00u;
    c4 |= (b == 0xC4u);
    if (b < 0x80u) return dv_ascii_sym(b);
    if (b >= 0xC0u) return 0u;                                          // 2-/4-byte lead
    if (code < 20992u) return code + 1u;                                // CJK
    return ((x >> 8) == 0x00AA84E2u) ? (DV_SYM_ASCII | 11u) : 0u;       // U+212A KELVIN SIGN lower-cases to 'k'
}

// ---------------------------------------------------------------- exact serial path
// Lane 0 walks the text; keys are queued and emitted by the whole warp.  Handles every
// code point exactly (U+0130 -> 'i' + separator, U+212A -> 'k').
template <class Emit, class WB>
__device__ void dv_tokenize_serial(const uint8_t* text, int32_t n, WB* wb, const unsigned long long* ktab, Emit& emit) {
    const int lane = threadIdx.x & 31;
    int32_t i = 0;
    int qn = 0; uint32_t qkind = 0;
    int cls = 0, alen = 0, plen = 0; uint64_t asum = 0, psum = 0; uint32_t prev_code = 0;      // lane-0 state
    uint32_t s_prev = 0, s_last = 0;                                                            // last two symbols (for signatures)
    bool done = false;
    while (!done) {
        if (lane == 0) {
            qn = 0; qkind = 0;
            while (qn < 29) {
                uint32_t code = 0; int scls = 0; bool brk_after = false;
                if (i >= n) done = true;
                else {
                    uint32_t b = text[i], cp; int len;
                    if (b < 0x80u) { cp = b; len = 1; }
                    else if (b < 0xE0u) { cp = b & 0x1Fu; len = 2; }
                    else if (b < 0xF0u) { cp = b & 0x0Fu; len = 3; }
                    else { cp = b & 0x07u; len = 4; }
                    for (int k = 1; k < len && i + k < n; k++) cp = (cp << 6) | (text[i + k] & 0x3Fu);
                    i += len;
                    if (cp < 0x80u) { uint32_t s = dv_ascii_sym(cp); if (s) { scls = 1; code = s & 0x7FFFu; } }
                    else if (cp - 0x4E00u < 20992u) { scls = 2; code = cp - 0x4E00u + 1u; }
                    else if (cp == 0x212Au) { scls = 1; code = 11; }                       // KELVIN SIGN -> 'k'
                    else if (cp == 0x0130u) { scls = 1; code = 9; brk_after = true; }      // -> 'i' + U+0307
                }
                if (cls == 1 && scls != 1) {                                               // ASCII run closes
                    if (alen >= 2) { wb->qsig[qn] = dv_token_sig(s_prev, s_last, alen); wb->queue[qn++] = dv_token_key(asum, alen, false); }
                    alen = 0; asum = 0;
                }
                if (cls == 2 && scls != 2) {                                               // CJK run closes
                    if (plen >= 3) { wb->qsig[qn] = dv_token_sig(s_prev, s_last, plen); wb->queue[qn++] = dv_token_key(psum, plen, true); }
                    plen = 0; psum = 0;
                }
                cls = scls;
                if (done) break;
                if (scls == 1) {
                    asum += (uint64_t)code * dv_kpos(alen, ktab); alen++;
                    s_prev = s_last; s_last = DV_SYM_ASCII | code;
                    if (brk_after) {
                        if (alen >= 2) { wb->qsig[qn] = dv_token_sig(s_prev, s_last, alen); wb->queue[qn++] = dv_token_key(asum, alen, false); }
                        alen = 0; asum = 0; cls = 0;
                    }
                } else if (scls == 2) {
                    if (plen == 12) { wb->qsig[qn] = dv_token_sig(s_prev, s_last, 12); wb->queue[qn++] = dv_token_key(psum, 12, true); plen = 0; psum = 0; }
                    if (plen >= 1) { qkind |= 1u << qn; wb->queue[qn++] = ((uint64_t)prev_code << 15) | code; }
                    psum += (uint64_t)code * dv_kpos(plen, ktab); plen++;
                    prev_code = code; s_prev = s_last; s_last = code;
                }
            }
        }
        qn = __shfl_sync(0xffffffffu, qn, 0);
        qkind = __shfl_sync(0xffffffffu, qkind, 0);
        done = __shfl_sync(0xffffffffu, (int)done, 0) != 0;
        __syncwarp();
        const uint64_t key = lane < qn ? wb->queue[lane] : 0ULL;
        const bool is_bg = (qkind >> lane) & 1u;
        emit.bigram(lane < qn && is_bg, (uint32_t)key);
        emit.token(lane < qn && !is_bg, key, lane < qn ? wb->qsig[lane] : 0u);
        __syncwarp();
    }
}

// ---------------------------------------------------------------- parallel fast path
// Returns true (warp-uniform) if the text must restart on the serial path.  Exotic bytes are
// detected in pass A of the tile holding them, i.e. before any token they could influence is
// emitted (a token's last symbol and its one-symbol look-ahead lie at most 4 bytes apart and a
// tile's last symbol is held back), so everything emitted up to that point is a correct prefix;
// consumers are idempotent or reset via emit.restart().
template <class Emit>
__device__ bool dv_tokenize_fast(const uint8_t* text, int32_t n, dv_warp_buf* wb, const unsigned long long* ktab, Emit& emit) {
    const int lane = threadIdx.x & 31;
    const int32_t a = (int32_t)((uintptr_t)text & 15u);
    uint16_t* B = wb->sym;                         // B[0..H) processed history, B[H..] pending symbols
    uint32_t* R = wb->raw + lane * 5;
    B[lane] = 0; B[lane + 32] = 0;
    int pending = 0;
    uint32_t carry_word = 0;
    int32_t gbase = 0;                             // global symbol index of B[H]
    int32_t cstart = 0, astart = 0;                // global index of the most recent CJK / ASCII run start
    __syncwarp();
    for (int32_t tile = -a; tile < n; tile += DV_TILE) {
        // ---- pass A: load, find symbol positions (SWAR), decode, compact
        uint32_t m16 = dv_tile_masks(text, n, tile, lane, carry_word, R);
        const int cnt = __popc(m16);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += y; }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        int widx = DV_HIST + pending + incl - cnt;
        bool ex = false;
        while (m16) {
            const int t = __ffs(m16) - 1; m16 &= m16 - 1;
            B[widx++] = (uint16_t)dv_decode_sym(R, t, ex);
        }
        if (__any_sync(0xffffffffu, ex)) return true;
        const bool last = (tile + DV_TILE >= n);
        const int m = pending + total;
        int m_proc = m;
        if (last) { if (lane == 0) B[DV_HIST + m] = 0; } else m_proc = m > 0 ? m - 1 : 0;
        __syncwarp();
        // ---- pass B: 32 symbols per step
        bool too_long = false;
        for (int s0 = 0; s0 < m_proc; s0 += 32) {
            const int i = DV_HIST + s0 + lane;
            const bool act = s0 + lane < m_proc;
            const uint32_t sv = act ? B[i] : 0u;
            const uint32_t sm1 = act ? B[i - 1] : 0u;
            const uint32_t sp1 = act ? B[i + 1] : 0u;
            const bool isC = dv_is_cjk_sym(sv), isA = dv_is_ascii_sym(sv);
            const bool prevC = isC && dv_is_cjk_sym(sm1), nextC = isC && dv_is_cjk_sym(sp1);
            const bool prevA = isA && dv_is_ascii_sym(sm1), nextA = isA && dv_is_ascii_sym(sp1);
            const int32_t step_g = gbase + s0, gi = step_g + lane;
            const uint32_t le = 0xffffffffu >> (31 - lane);
            const uint32_t cs_mask = __ballot_sync(0xffffffffu, isC && !prevC);
            const uint32_t as_mask = __ballot_sync(0xffffffffu, isA && !prevA);
            const int32_t my_cs = (cs_mask & le) ? step_g + (31 - __clz(cs_mask & le)) : cstart;
            const int32_t my_as = (as_mask & le) ? step_g + (31 - __clz(as_mask & le)) : astart;
            if (cs_mask) cstart = step_g + (31 - __clz(cs_mask));
            if (as_mask) astart = step_g + (31 - __clz(as_mask));
            const uint32_t d = isC ? (uint32_t)(gi - my_cs) : (uint32_t)(gi - my_as);
            const uint32_t dist = isC ? d % 12u : (isA ? d : 0u);     // distance to the token's first symbol
            const bool bg = prevC && dist != 0u;
            emit.bigram(bg, (sm1 << 15) | sv);
            const bool tk = isC ? ((dist == 11u || !nextC) && dist >= 2u) : (isA && !nextA && dist >= 1u);
            const uint32_t len = dist + 1u;
            const uint32_t sig = dv_token_sig(sm1, sv, len);
            const bool need = tk && emit.need_token(sig);
            uint64_t key = 0;
            if (__any_sync(0xffffffffu, need)) {
                if (need) {
                    if (len > DV_HIST) too_long = true;
                    else {
                        uint64_t mm = 0;
                        const uint16_t* tp = B + (i - (int)dist);
                        for (uint32_t q = 0; q < len; q++) mm += (uint64_t)(tp[q] & 0x7FFFu) * ktab[q];
                        key = dv_token_key(mm, len, isC);
                    }
                }
                if (__any_sync(0xffffffffu, too_long)) return true;
                emit.token(need, key, sig);
            }
        }
        // ---- slide: keep the last DV_HIST processed symbols and the held-back one
        __syncwarp();
        {
            const int keep = DV_HIST + ((!last && m > 0) ? 1 : 0);
            uint16_t t0 = 0, t1 = 0, t2 = 0;
            if (lane < keep) t0 = B[m_proc + lane];
            if (lane + 32 < keep) t1 = B[m_proc + lane + 32];
            if (lane + 64 < keep) t2 = B[m_proc + lane + 64];
            __syncwarp();
            if (lane < keep) B[lane] = t0;
            if (lane + 32 < keep) B[lane + 32] = t1;
            if (lane + 64 < keep) B[lane + 64] = t2;
            pending = keep - DV_HIST;
        }
        gbase += m_proc;
        __syncwarp();
    }
    return false;
}

template <class Emit>
__device__ __forceinline__ void dv_tokenize_v3_warp(const uint8_t* text, int32_t n, dv_warp_buf* wb, const unsigned long long* ktab, Emit& emit) {
    if (n <= 0) return;
    if (dv_tokenize_fast(text, n, wb, ktab, emit)) {
        emit.restart();
        dv_tokenize_serial(text, n, wb, ktab, emit);
    }
}
