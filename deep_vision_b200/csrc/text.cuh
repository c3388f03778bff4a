// text.cuh — warp-level UTF-8/CJK decode + similarity-token key generation (kernel stage K1).
//
// Follows _tokenize_similarity_text_v3 (reference web/server.py:24722-24732):
//   lower(); ASCII tokens [a-z0-9_]{2,}; CJK runs split greedily into pieces of <=12 chars,
//   each piece (len>=2) and all bigrams inside a piece are tokens.
//
// One warp streams one text.
//   pass A  coalesced 16-byte loads (512 B per iteration); each lane classifies its 16 bytes in
//           registers (look-behind through the neighbour lane's last word) and the warp compacts
//           the resulting SYMBOLS (u16: 0 = separator, 1..20992 = CJK code, 0x8000|code = ASCII)
//           into a dense shared-memory buffer with one prefix sum.
//   pass B  32 symbols per step, one per lane: run starts by ballot/clz, position inside the
//           12-char piece, exact bigram ids, and the multilinear token hash
//               M = sum_d code_d * K(d)      (K(d) = splitmix64(d+1)|1)
//           of every piece / ASCII run by ONE segmented 64-bit warp scan whose carry crosses
//           steps and tiles — no per-lane loops, no look-behind beyond one symbol.
// Tokens are emitted at their last symbol: emit.bigram(valid, id30) / emit.token(valid, key64),
// both called warp-convergently.
//
// Texts containing a byte that may lower-case INTO the ASCII class from outside it
// (U+0130 -> 'i'+U+0307, U+212A -> 'k'; conservatively: any 0xC4 lead or E2 84 xx) restart on
// the exact serial path dv_tokenize_serial().
#pragma once
#include <stdint.h>

#define DV_SYM_ASCII 0x8000u
#define DV_TILE   512                    // bytes staged per iteration (32 lanes x 16 B)
#define DV_SYMCAP (DV_TILE + 40)         // history + pending + new symbols + terminator, padded

struct __align__(16) dv_warp_buf {
    uint16_t sym[DV_SYMCAP];
    unsigned long long queue[32];        // serial-path key queue
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

#define DV_NOSYM 0xFFFFu
// Symbols ENDING at each of the 16 bytes of one lane (DV_NOSYM where no char ends), packed 2 per word.
// prev = the word preceding the 16 bytes.  Returns the number of symbols.
__device__ __forceinline__ int dv_symbolize16(uint4 w, uint32_t prev, uint32_t packed[8], bool& exotic) {
    const uint32_t words[5] = { prev, w.x, w.y, w.z, w.w };
    int cnt = 0;
#pragma unroll
    for (int t = 0; t < 16; t++) {
        const int q = 4 + t;
        const uint32_t b  = (words[q >> 2] >> ((q & 3) * 8)) & 0xFFu;
        const uint32_t p1 = (words[(q - 1) >> 2] >> (((q - 1) & 3) * 8)) & 0xFFu;
        const uint32_t p2 = (words[(q - 2) >> 2] >> (((q - 2) & 3) * 8)) & 0xFFu;
        uint32_t s;
        if (b < 0x80u) s = dv_ascii_sym(b);
        else if (b < 0xC0u) {                                           // continuation byte
            if (p1 >= 0xC0u) s = (p1 < 0xE0u) ? 0u : DV_NOSYM;          // end of a 2-byte char / middle of a longer one
            else if (p1 >= 0x80u) {
                if (p2 >= 0xE0u) {
                    if (p2 < 0xF0u) {                                   // end of a 3-byte char
                        const uint32_t code = (((p2 & 0x0Fu) << 12) | ((p1 & 0x3Fu) << 6) | (b & 0x3Fu)) - 0x4E00u;
                        s = code < 20992u ? code + 1u : 0u;
                    } else s = DV_NOSYM;                                // third byte of a 4-byte char
                } else s = 0u;                                          // end of a 4-byte char (or invalid)
            } else s = 0u;                                              // invalid: continuation after ASCII
            if (p1 == 0xE2u && b == 0x84u) exotic = true;
        } else { s = DV_NOSYM; if (b == 0xC4u) exotic = true; }         // lead byte
        cnt += (s != DV_NOSYM);
        if (t & 1) packed[t >> 1] |= s << 16; else packed[t >> 1] = s;
    }
    return cnt;
}

// ---------------------------------------------------------------- exact serial path
// Lane 0 walks the text; keys are queued and emitted by the whole warp.  Handles every
// code point exactly (U+0130 -> 'i' + separator, U+212A -> 'k').
template <class Emit>
__device__ void dv_tokenize_serial(const uint8_t* text, int32_t n, dv_warp_buf* wb, const unsigned long long* ktab, Emit& emit) {
    const int lane = threadIdx.x & 31;
    int32_t i = 0;
    int qn = 0; uint32_t qkind = 0;
    int cls = 0, alen = 0, plen = 0; uint64_t asum = 0, psum = 0; uint32_t prev_code = 0;      // lane-0 state
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
                    if (alen >= 2) wb->queue[qn++] = dv_token_key(asum, alen, false);
                    alen = 0; asum = 0;
                }
                if (cls == 2 && scls != 2) {                                               // CJK run closes
                    if (plen >= 3) wb->queue[qn++] = dv_token_key(psum, plen, true);
                    plen = 0; psum = 0;
                }
                cls = scls;
                if (done) break;
                if (scls == 1) {
                    asum += (uint64_t)code * dv_kpos(alen, ktab); alen++;
                    if (brk_after) {
                        if (alen >= 2) wb->queue[qn++] = dv_token_key(asum, alen, false);
                        alen = 0; asum = 0; cls = 0;
                    }
                } else if (scls == 2) {
                    if (plen == 12) { wb->queue[qn++] = dv_token_key(psum, 12, true); plen = 0; psum = 0; }
                    if (plen >= 1) { qkind |= 1u << qn; wb->queue[qn++] = ((uint64_t)prev_code << 15) | code; }
                    psum += (uint64_t)code * dv_kpos(plen, ktab); plen++;
                    prev_code = code;
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
        emit.token(lane < qn && !is_bg, key);
        __syncwarp();
    }
}

// ---------------------------------------------------------------- parallel fast path
// Returns true (warp-uniform) if an exotic byte was met: the routine then stops at once and the
// caller restarts the text on the serial path.  Detection happens in pass A of the tile holding
// the byte, i.e. before any token it could influence is emitted (a token's last symbol and its
// one-symbol look-ahead lie at most 4 bytes apart and a tile's last symbol is held back), so
// everything emitted up to that point is a correct prefix; consumers are idempotent or reset
// via emit.restart().
template <class Emit>
__device__ bool dv_tokenize_fast(const uint8_t* text, int32_t n, dv_warp_buf* wb, const unsigned long long* ktab, Emit& emit) {
    const int lane = threadIdx.x & 31;
    const int32_t a = (int32_t)((uintptr_t)text & 15u);
    uint16_t* B = wb->sym;                         // B[0] = last processed symbol, B[1..] = pending symbols
    if (lane == 0) B[0] = 0;
    int pending = 0;
    uint32_t carry_word = 0;
    int32_t gbase = 0;                             // global symbol index of B[1]
    int32_t cstart = 0, astart = 0;                // global index of the most recent CJK / ASCII run start
    uint64_t scarry = 0;                           // scan value at the last processed symbol
    __syncwarp();
    for (int32_t tile = -a; tile < n; tile += DV_TILE) {
        // ---- pass A: load, classify, compact
        const int32_t pos = tile + lane * 16;
        uint4 w = make_uint4(0, 0, 0, 0);
        if (pos < n && pos + 16 > 0) {
            w = *reinterpret_cast<const uint4*>(text + pos);
            if (pos < 0 || pos + 16 > n) {                             // mask bytes outside the text
                uint32_t ws[4] = { w.x, w.y, w.z, w.w };
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    uint32_t m = 0;
#pragma unroll
                    for (int bb = 0; bb < 4; bb++) { const int32_t q = pos + k * 4 + bb; if (q >= 0 && q < n) m |= 0xFFu << (bb * 8); }
                    ws[k] &= m;
                }
                w = make_uint4(ws[0], ws[1], ws[2], ws[3]);
            }
        }
        uint32_t prev = __shfl_up_sync(0xffffffffu, w.w, 1);
        if (lane == 0) prev = carry_word;
        carry_word = __shfl_sync(0xffffffffu, w.w, 31);
        bool ex = false;
        uint32_t packed[8];
        int cnt = dv_symbolize16(w, prev, packed, ex);
        if (__any_sync(0xffffffffu, ex)) return true;
        // bytes outside the text were zeroed -> they produced separator symbols; drop them
        if (pos < 0 || pos + 16 > n) {
            cnt = 0;
#pragma unroll
            for (int t = 0; t < 16; t++) {
                const int32_t q = pos + t;
                const uint32_t s = (packed[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
                if (q < 0 || q >= n) { if (t & 1) packed[t >> 1] |= 0xFFFF0000u; else packed[t >> 1] |= 0xFFFFu; }
                else cnt += (s != DV_NOSYM);
            }
        }
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += y; }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        int widx = 1 + pending + incl - cnt;
#pragma unroll
        for (int t = 0; t < 16; t++) {
            const uint32_t s = (packed[t >> 1] >> ((t & 1) * 16)) & 0xFFFFu;
            if (s != DV_NOSYM) B[widx++] = (uint16_t)s;
        }
        const bool last = (tile + DV_TILE >= n);
        const int m = pending + total;
        int m_proc = m;
        if (last) { if (lane == 0) B[1 + m] = 0; } else m_proc = m > 0 ? m - 1 : 0;
        __syncwarp();
        // ---- pass B: 32 symbols per step
        for (int s0 = 0; s0 < m_proc; s0 += 32) {
            const int i = 1 + s0 + lane;
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
            // multilinear contribution + segmented inclusive scan (segments = pieces / ASCII runs)
            uint64_t v = 0;
            if (isC || isA) v = (uint64_t)(sv & 0x7FFFu) * dv_kpos(dist, ktab);
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint64_t y = __shfl_up_sync(0xffffffffu, v, off);
                if (lane >= off && dist >= (uint32_t)off) v += y;
            }
            if (dist > (uint32_t)lane) v += scarry;                    // token began before this step
            const int nact = min(32, m_proc - s0);
            scarry = __shfl_sync(0xffffffffu, v, nact - 1);
            // tokens ending here
            const bool bg = prevC && dist != 0u;
            const uint32_t bg_id = (sm1 << 15) | sv;
            const bool tk = isC ? ((dist == 11u || !nextC) && dist >= 2u) : (isA && !nextA && dist >= 1u);
            const uint64_t key = dv_token_key(v, dist + 1u, isC);
            emit.bigram(bg, bg_id);
            emit.token(tk, key);
        }
        // ---- slide: keep the last processed symbol and the held-back one
        __syncwarp();
        if (!last && m > 0) {
            if (lane == 0) {
                const uint16_t h0 = m_proc >= 1 ? B[m_proc] : B[0];
                const uint16_t h1 = B[m];
                B[0] = h0; B[1] = h1;
            }
            pending = 1;
        } else pending = last ? 0 : pending;
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
