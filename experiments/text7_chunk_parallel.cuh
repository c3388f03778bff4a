// text7.cuh — chunk-parallel warp tokenizer for long texts (facts): every lane is a small scalar automaton.
//
// Same token semantics as text.cuh / text6.cuh (_tokenize_similarity_text_v3, reference web/server.py:24722-24732).
// A tile (<= 2 KB, balanced over the text) is staged in shared memory; lane l owns the characters that START in the
// l-th of 32 equal chunks of the tile.  It first recovers the tokenizer state at its chunk start by scanning BACKWARDS
// over the run that reaches into the chunk (previous two symbols, position inside the run / 12-character piece),
// then walks its characters forwards: per character one bigram (previous, current) and the token that ends at the
// previous character, exactly the look-back formulation of text6.cuh.  No warp-wide compaction, no position lists,
// no per-step ballots: 32 characters advance per loop iteration.
// Restart (return true) when a run reaches further back than the staged history, on exotic case mappings
// (U+0130, U+212A) and on ASCII tokens longer than 64: the caller resets its state and re-runs the exact tokenizer.
#pragma once
#include "text6.cuh"

#define DV7_TILE 2048
#define DV7_HIST 128
#define DV7_BYTES (DV7_HIST + DV7_TILE + 16)       // history | tile | 16 B look-ahead

struct __align__(16) dv7_warp_buf { uint32_t w[DV7_BYTES / 4]; };

// 4 buffer bytes starting at byte index i
__device__ __forceinline__ uint32_t dv7_win(const uint32_t* W, int i) {
    return __funnelshift_r(W[i >> 2], W[(i >> 2) + 1], (uint32_t)(i & 3) * 8u);
}
// symbol of the character whose last byte is buffer byte i-1 (U+212A is handled by the restart, here it is a separator)
__device__ __forceinline__ uint32_t dv7_back_sym(const uint32_t* W, const uint8_t* lut, int i) {
    const uint32_t x = dv7_win(W, i - 4);
    const uint32_t b = x >> 24;
    const uint32_t ac = lut[b & 0x7Fu];
    const uint32_t code = ((((x >> 8) & 0x0Fu) << 12) | (((x >> 16) & 0x3Fu) << 6) | (b & 0x3Fu)) - 0x4E00u;
    const bool is3 = (x & 0x00C0F000u) == 0x0080E000u;
    const uint32_t s = (is3 & (code < 20992u)) ? code + 1u : 0u;
    return (b < 0x80u) ? (ac ? (DV_SYM_ASCII | ac) : 0u) : s;
}
__device__ __forceinline__ uint32_t dv7_cls(uint32_t s) { return dv_is_cjk_sym(s) ? 2u : (dv_is_ascii_sym(s) ? 1u : 0u); }

template <class Emit>
__device__ bool dv7_tokenize_fast(const uint8_t* text, int32_t n, dv7_warp_buf* wb, const unsigned long long* ktab,
                                  const uint8_t* lut, Emit& emit) {
    const int lane = threadIdx.x & 31;
    uint32_t* W = wb->w;
    const int32_t a = (int32_t)((uintptr_t)text & 15u);
    const uint8_t* g0 = text - a;                          // 16-byte aligned; stream coordinate s = byte offset from g0
    const int32_t end_s = a + n;
    const int32_t end16 = (end_s + 15) & ~15;
    bool bad = false;                                      // exotic / history exhausted / token too long
    // balanced tiles: the stream positions [0, end_s] (end_s = terminator) are cut into equal tiles of <= DV7_TILE bytes,
    // each tile into 32 equal chunks, so every lane walks about the same number of characters
    const int32_t ntiles = (end_s + DV7_TILE) / DV7_TILE;
    const int32_t tb = (((end_s + ntiles) / ntiles) + 15) & ~15;
    const int32_t ch = (tb + 31) >> 5;
    for (int32_t t0 = 0; t0 <= end_s; t0 += tb) {
        // ---- stage the tile (history = last DV7_HIST bytes of the previous tile)
        __syncwarp();
        if (t0 > 0) {
            uint4 h = make_uint4(0u, 0u, 0u, 0u);
            if (lane < DV7_HIST / 16) h = reinterpret_cast<const uint4*>(W)[(tb >> 4) + lane];
            __syncwarp();
            if (lane < DV7_HIST / 16) reinterpret_cast<uint4*>(W)[lane] = h;
        }
        for (int32_t blk = lane; blk <= (tb >> 4); blk += 32) {            // tile + one 16-byte block of look-ahead
            const int32_t s = t0 + blk * 16;
            if (s < end16) reinterpret_cast<uint4*>(W)[DV7_HIST / 16 + blk] = *reinterpret_cast<const uint4*>(g0 + s);
        }
        __syncwarp();
        const int32_t base = DV7_HIST - t0;                // buffer byte index = s + base
        const int32_t cs = t0 + lane * ch, ce = min(cs + ch, t0 + tb);
        const int32_t lo = max(cs, a), hi = min(ce, end_s);
        bool term_pending = (cs <= end_s) && (end_s < ce);
        // ---- first character that starts in the chunk
        int32_t p = lo;
        if (lo < hi && lo > a) {
            const uint32_t x = dv7_win(W, lo + base);
            const uint32_t cont = x & ~(x << 1) & 0x80808080u;             // continuation bytes
            const uint32_t lead = ~cont & 0x80808080u;
            p = lead ? min(lo + ((__ffs(lead) - 1) >> 3), hi) : hi;          // first non-continuation byte (<= 3 skipped)
        } else if (lo >= hi) p = hi;
        // ---- tokenizer state just before p: previous two symbols, position of the previous symbol in its run
        uint32_t sm1 = 0, sm2 = 0, rp = 0;
        if ((p < hi || term_pending) && p > a) {
            sm1 = dv7_back_sym(W, lut, p + base);
            const uint32_t c1 = dv7_cls(sm1);
            if (c1) {
                const int32_t step = c1 == 2u ? 3 : 1;
                int32_t q = p - step;
                uint32_t cnt = 0;
                while (q > a) {
                    if (q < t0 - DV7_HIST + 4 || cnt > 64u) { bad = true; break; }
                    const uint32_t s = dv7_back_sym(W, lut, q + base);
                    if (dv7_cls(s) != c1) break;
                    if (cnt == 0u) sm2 = s;
                    cnt++; q -= step;
                }
                rp = c1 == 2u ? cnt % 12u : cnt;
            }
        }
        // ---- forward walk: one character per lane per iteration
        for (;;) {
            const bool act = p < hi;
            const bool term = !act && term_pending;
            if (!__any_sync(0xffffffffu, act || term)) break;
            uint32_t sv = 0, clen = 0;
            if (act) {
                const uint32_t x = dv7_win(W, p + base);
                const uint32_t b0 = x & 0xFFu;
                const uint32_t ac = lut[b0 & 0x7Fu];
                const uint32_t code = (((b0 & 0x0Fu) << 12) | (((x >> 8) & 0x3Fu) << 6) | ((x >> 16) & 0x3Fu)) - 0x4E00u;
                clen = b0 < 0x80u ? 1u : (b0 < 0xE0u ? 2u : (b0 < 0xF0u ? 3u : 4u));
                sv = b0 < 0x80u ? (ac ? (DV_SYM_ASCII | ac) : 0u) : ((clen == 3u && code < 20992u) ? code + 1u : 0u);
                bad |= (b0 == 0xC4u) | ((x & 0x00FFFFFFu) == 0x00AA84E2u);  // U+0130 / U+212A lower-case into the ASCII class
            }
            if (term) term_pending = false;
            const bool v = act | term;
            const bool curC = dv_is_cjk_sym(sv), curA = dv_is_ascii_sym(sv);
            const bool pC = dv_is_cjk_sym(sm1), pA = dv_is_ascii_sym(sm1);
            emit.bigram(v & pC & curC & (rp != 11u), (sm1 << 15) | sv);
            const bool tk = v & (pC ? (((rp == 11u) | !curC) & (rp >= 2u)) : (pA & !curA & (rp >= 1u)));
            const uint32_t len = rp + 1u;
            const uint32_t sig = dv_token_sig(sm2, sm1, len);
            const bool need = tk & emit.need_token(sig);
            if (__any_sync(0xffffffffu, need)) {
                uint64_t key = 0;
                if (need) {
                    if (len > 64u) bad = true;
                    else {
                        uint64_t mm = 0;
                        const int step = pC ? 3 : 1;
                        int q = p + base;
                        for (uint32_t d = len; d-- > 0u; q -= step) mm += (uint64_t)(dv7_back_sym(W, lut, q) & 0x7FFFu) * ktab[d];
                        key = dv_token_key(mm, len, pC);
                    }
                }
                if (__any_sync(0xffffffffu, bad)) return true;
                emit.token(need, key, sig);
            }
            if (v) {
                const bool same = (curC & pC) | (curA & pA);
                rp = same ? ((curC & (rp == 11u)) ? 0u : rp + 1u) : 0u;
                sm2 = sm1; sm1 = sv; p += (int32_t)clen;
            }
        }
        if (__any_sync(0xffffffffu, bad)) return true;
    }
    return false;
}
