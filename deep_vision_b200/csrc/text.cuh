// text.cuh — warp-level UTF-8/CJK decode + similarity-token key generation (kernel stage K1).
//
// Follows _tokenize_similarity_text_v3 (reference web/server.py:24722-24732):
//   lower(); ASCII tokens [a-z0-9_]{2,}; CJK runs split greedily into pieces of <=12 chars,
//   each piece (len>=2) and all bigrams inside a piece are tokens.
//
// One warp streams one text.  Bytes are fetched with coalesced 16-byte loads, staged in a
// per-warp shared-memory window and classified IN BYTE SPACE (no compaction): a "cell" per
// byte says which symbol ENDS at that byte.  Within an ASCII run symbols are 1 byte apart and
// within a CJK run exactly 3 bytes apart, so neighbours are cell[j±1] / cell[j±3].
// Tokens are emitted at their LAST symbol, so only 3 bytes of look-ahead are needed.
//
// Texts containing a byte that may lower-case INTO the ASCII class from outside it
// (U+0130 -> 'i'+U+0307, U+212A -> 'k'; conservatively: any 0xC4 lead or E2 84 xx) take the
// exact serial path tokenize_serial() instead.
#pragma once
#include <stdint.h>

#define DV_CELL_NONE  0u     // lead or middle byte
#define DV_CELL_OTHER 64u    // a separator symbol ends here
#define DV_CELL_CJK   128u   // a CJK symbol (U+4E00..U+9FFF) ends here; ASCII codes are 1..37

#define DV_TILE   512        // bytes staged per iteration (32 lanes x 16 B)
#define DV_BACK   64         // look-behind kept in the window (>= 16 lag + 36 for a 12-char piece)
#define DV_WIN    (DV_BACK + DV_TILE)

struct __align__(16) dv_warp_window {
    uint8_t raw[DV_WIN];
    uint8_t cell[DV_WIN];
    unsigned long long queue[32];   // serial-path key queue
};

__device__ __forceinline__ uint32_t dv_ascii_cell(uint32_t b) {
    uint32_t l = b | 0x20u;
    if (l - 'a' < 26u) return l - 'a' + 1u;
    if (b - '0' < 10u) return b - '0' + 27u;
    return b == '_' ? 37u : DV_CELL_OTHER;
}
__device__ __forceinline__ bool dv_is_ascii_cell(uint32_t c) { return c - 1u < 37u; }

__device__ __forceinline__ uint64_t dv_fnv_step(uint64_t h, uint32_t sym) { return (h ^ (uint64_t)sym) * 0x100000001b3ULL; }
__device__ __forceinline__ uint64_t dv_hash62_finish(uint64_t h, uint32_t len) {
    h = dv_fnv_step(h, len);
    h ^= h >> 33; h *= 0xff51afd7ed558ccdULL; h ^= h >> 33; h *= 0xc4ceb9fe1a85ec53ULL; h ^= h >> 33;
    return h & 0x3FFFFFFFFFFFFFFFULL;
}
#define DV_FNV_SEED 0xcbf29ce484222325ULL
__device__ __forceinline__ uint32_t dv_ascii_code_to_cp(uint32_t code) {
    return code <= 26u ? code + ('a' - 1u) : (code <= 36u ? code + ('0' - 27u) : (uint32_t)'_');
}

// classify the 16 bytes of one lane.  prev = the word preceding them (bytes -4..-1).
__device__ __forceinline__ uint4 dv_classify16(uint4 w, uint32_t prev, bool& exotic) {
    uint32_t words[5] = { prev, w.x, w.y, w.z, w.w };
    uint32_t out[4];
#pragma unroll
    for (int wi = 0; wi < 4; wi++) {
        uint32_t o = 0;
#pragma unroll
        for (int bi = 0; bi < 4; bi++) {
            const int t = 4 + wi * 4 + bi;                       // byte index into words[] stream
            uint32_t b  = (words[t >> 2] >> ((t & 3) * 8)) & 0xFFu;
            uint32_t p1 = (words[(t - 1) >> 2] >> (((t - 1) & 3) * 8)) & 0xFFu;
            uint32_t p2 = (words[(t - 2) >> 2] >> (((t - 2) & 3) * 8)) & 0xFFu;
            uint32_t p3 = (words[(t - 3) >> 2] >> (((t - 3) & 3) * 8)) & 0xFFu;
            uint32_t cell;
            if (b < 0x80u) cell = dv_ascii_cell(b);
            else if ((b & 0xC0u) == 0x80u) {
                if (p1 >= 0xC0u) cell = (p1 < 0xE0u) ? DV_CELL_OTHER : DV_CELL_NONE;
                else if (p1 >= 0x80u) {
                    if (p2 >= 0xE0u) {
                        if (p2 < 0xF0u) cell = ((p2 - 0xE5u < 5u) || (p2 == 0xE4u && p1 >= 0xB8u)) ? DV_CELL_CJK : DV_CELL_OTHER;
                        else cell = DV_CELL_NONE;
                    } else cell = DV_CELL_OTHER;                 // end of a 4-byte char (or invalid)
                } else cell = DV_CELL_OTHER;
                if (p1 == 0xE2u && b == 0x84u) exotic = true;
                (void)p3;
            } else { cell = DV_CELL_NONE; if (b == 0xC4u) exotic = true; }
            o |= cell << (bi * 8);
        }
        out[wi] = o;
    }
    return make_uint4(out[0], out[1], out[2], out[3]);
}

__device__ __forceinline__ uint32_t dv_cjk_code(const uint8_t* raw, int bi) {   // symbol ending at window index bi
    uint32_t cp = ((raw[bi - 2] & 0x0Fu) << 12) | ((raw[bi - 1] & 0x3Fu) << 6) | (raw[bi] & 0x3Fu);
    return cp;
}

// ---------------------------------------------------------------- exact serial path
// Lane 0 walks the text; keys are queued and emitted by the whole warp.  Handles every
// code point exactly (U+0130 -> 'i' + separator, U+212A -> 'k').
template <class Emit>
__device__ void dv_tokenize_serial(const uint8_t* text, int32_t n, dv_warp_window* win, Emit& emit) {
    const int lane = threadIdx.x & 31;
    int32_t i = 0;
    int qn = 0;
    // state (lane 0 only)
    int cls = 0, alen = 0, plen = 0; uint64_t ahash = DV_FNV_SEED, apack = 0; uint32_t piece[12];
    bool done = false;
    while (!done) {
        if (lane == 0) {
            while (qn < 30) {
                uint32_t sym = 0; int scls = 0; bool brk_after = false;
                if (i >= n) { done = true; }
                else {
                    uint32_t b = text[i], cp; int len;
                    if (b < 0x80u) { cp = b; len = 1; }
                    else if (b < 0xE0u) { cp = b & 0x1Fu; len = 2; }
                    else if (b < 0xF0u) { cp = b & 0x0Fu; len = 3; }
                    else { cp = b & 0x07u; len = 4; }
                    for (int k = 1; k < len && i + k < n; k++) cp = (cp << 6) | (text[i + k] & 0x3Fu);
                    i += len;
                    if (cp < 0x80u) { uint32_t c = dv_ascii_cell(cp); if (c != DV_CELL_OTHER) { scls = 1; sym = c; } }
                    else if (cp >= 0x4E00u && cp <= 0x9FFFu) { scls = 2; sym = cp; }
                    else if (cp == 0x212Au) { scls = 1; sym = 11; }                 // KELVIN SIGN -> 'k'
                    else if (cp == 0x0130u) { scls = 1; sym = 9; brk_after = true; } // -> 'i' + U+0307
                }
                // close runs that do not continue with this symbol
                if (cls == 1 && scls != 1) {
                    if (alen >= 2) win->queue[qn++] = alen <= 10 ? ((1ULL << 62) | (apack << (6 * (10 - alen))))
                                                                 : ((2ULL << 62) | dv_hash62_finish(ahash, alen));
                    alen = 0; ahash = DV_FNV_SEED; apack = 0;
                }
                if (cls == 2 && scls != 2) {
                    if (plen >= 3) {
                        if (plen <= 4) { uint64_t k = 0; for (int t = 0; t < 4; t++) k = (k << 15) | (t < plen ? (uint64_t)(piece[t] - 0x4E00u + 1u) : 0ULL); win->queue[qn++] = k; }
                        else { uint64_t h = DV_FNV_SEED; for (int t = 0; t < plen; t++) h = dv_fnv_step(h, piece[t]); win->queue[qn++] = (3ULL << 62) | dv_hash62_finish(h, plen); }
                    }
                    plen = 0;
                }
                cls = scls;
                if (done) break;
                if (scls == 1) {
                    if (alen < 10) apack = (apack << 6) | sym;
                    ahash = dv_fnv_step(ahash, dv_ascii_code_to_cp(sym)); alen++;
                    if (brk_after) {        // the separator U+0307 follows immediately
                        if (alen >= 2) win->queue[qn++] = alen <= 10 ? ((1ULL << 62) | (apack << (6 * (10 - alen))))
                                                                     : ((2ULL << 62) | dv_hash62_finish(ahash, alen));
                        alen = 0; ahash = DV_FNV_SEED; apack = 0; cls = 0;
                    }
                } else if (scls == 2) {
                    if (plen == 12) {       // piece full: emit it and start the next piece
                        uint64_t h = DV_FNV_SEED; for (int t = 0; t < 12; t++) h = dv_fnv_step(h, piece[t]);
                        win->queue[qn++] = (3ULL << 62) | dv_hash62_finish(h, 12);
                        plen = 0;
                    }
                    if (plen >= 1) win->queue[qn++] = ((uint64_t)(piece[plen - 1] - 0x4E00u + 1u) << 45) | ((uint64_t)(sym - 0x4E00u + 1u) << 30);
                    piece[plen++] = sym;
                }
            }
        }
        qn = __shfl_sync(0xffffffffu, qn, 0);
        done = __shfl_sync(0xffffffffu, (int)done, 0) != 0;
        __syncwarp();
        uint64_t key = lane < qn ? win->queue[lane] : 0ULL;
        emit(lane < qn, key);
        __syncwarp();
        qn = 0;
    }
}

// ---------------------------------------------------------------- parallel fast path
// emit(valid, key) is called warp-convergently (all 32 lanes, twice per 32-byte step).
// Returns true (warp-uniform) if an exotic byte was met: the routine then stops at once and the
// caller restarts the text on the serial path.  Detection happens in pass A of the tile holding
// the byte, i.e. before any token it could influence is emitted, so everything emitted up to
// that point is a correct prefix; consumers are idempotent or are reset via emit.restart().
template <class Emit>
__device__ bool dv_tokenize_fast(const uint8_t* text, int32_t n, dv_warp_window* win, Emit& emit) {
    const int lane = threadIdx.x & 31;
    const int32_t a = (int32_t)((uintptr_t)text & 15u);
    // window index wi  <->  text-relative byte position  p = origin + wi
    for (int t = lane; t < DV_BACK / 4; t += 32) {
        reinterpret_cast<uint32_t*>(win->raw)[t] = 0u;
        reinterpret_cast<uint32_t*>(win->cell)[t] = 0x40404040u;   // DV_CELL_OTHER
    }
    uint32_t carry_word = 0;
    int32_t cstart = -0x40000000, astart = -0x40000000;            // most recent run starts (warp-uniform)
    __syncwarp();
    for (int32_t tile = -a; tile < n + 16; tile += DV_TILE) {
        const int32_t origin = tile - DV_BACK;
        // ---- stage + classify (pass A)
        {
            int32_t pos = tile + lane * 16;
            uint4 w = make_uint4(0, 0, 0, 0);
            if (pos < n && pos + 16 > 0) {
                w = *reinterpret_cast<const uint4*>(text + pos);
                if (pos < 0 || pos + 16 > n) {                        // mask bytes outside the text
                    uint32_t ws[4] = { w.x, w.y, w.z, w.w };
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        uint32_t m = 0;
#pragma unroll
                        for (int bb = 0; bb < 4; bb++) { int32_t q = pos + k * 4 + bb; if (q >= 0 && q < n) m |= 0xFFu << (bb * 8); }
                        ws[k] &= m;
                    }
                    w = make_uint4(ws[0], ws[1], ws[2], ws[3]);
                }
            }
            uint32_t prev = __shfl_up_sync(0xffffffffu, w.w, 1);
            if (lane == 0) prev = carry_word;
            carry_word = __shfl_sync(0xffffffffu, w.w, 31);
            bool ex = false;
            uint4 c = dv_classify16(w, prev, ex);
            *reinterpret_cast<uint4*>(win->raw + DV_BACK + lane * 16) = w;
            *reinterpret_cast<uint4*>(win->cell + DV_BACK + lane * 16) = c;
            if (__any_sync(0xffffffffu, ex)) return true;
        }
        __syncwarp();
        // ---- emit (pass B) for positions [tile-16, tile+DV_TILE-16) within the text
#pragma unroll 1
        for (int s = 0; s < DV_TILE / 32; s++) {
            const int32_t step_base = tile - 16 + s * 32;
            if (step_base >= n) break;
            if (step_base + 32 <= 0) continue;
            const int32_t j = step_base + lane;
            const int wi = j - origin;                                // >= DV_BACK-16
            const bool inside = (j >= 0) && (j < n);
            const uint32_t c0 = inside ? win->cell[wi] : DV_CELL_NONE;
            const bool isC = (c0 == DV_CELL_CJK);
            const bool isA = dv_is_ascii_cell(c0);
            const bool prevC = isC && (win->cell[wi - 3] == DV_CELL_CJK);
            const bool nextC = isC && (win->cell[wi + 3] == DV_CELL_CJK);
            const bool prevA = isA && dv_is_ascii_cell(win->cell[wi - 1]);
            const bool nextA = isA && dv_is_ascii_cell(win->cell[wi + 1]);
            const uint32_t le = 0xffffffffu >> (31 - lane);
            const uint32_t cs_mask = __ballot_sync(0xffffffffu, isC && !prevC);
            const uint32_t as_mask = __ballot_sync(0xffffffffu, isA && !prevA);
            int32_t my_cs = (cs_mask & le) ? step_base + (31 - __clz(cs_mask & le)) : cstart;
            int32_t my_as = (as_mask & le) ? step_base + (31 - __clz(as_mask & le)) : astart;
            if (cs_mask) cstart = step_base + (31 - __clz(cs_mask));
            if (as_mask) astart = step_base + (31 - __clz(as_mask));

            bool v1 = false, v2 = false; uint64_t k1 = 0, k2 = 0;
            if (isC) {
                const int d = (j - my_cs) / 3;
                const int dm = d % 12;
                const uint32_t code = dv_cjk_code(win->raw, wi) - 0x4E00u + 1u;
                if (prevC && dm != 0) {
                    const uint32_t pcode = dv_cjk_code(win->raw, wi - 3) - 0x4E00u + 1u;
                    k1 = ((uint64_t)pcode << 45) | ((uint64_t)code << 30); v1 = true;
                }
                if ((dm == 11 || !nextC) && dm >= 2) {
                    const int len = dm + 1;
                    if (len <= 4) {
                        uint64_t k = 0;
                        for (int t = 0; t < 4; t++) k = (k << 15) | (t < len ? (uint64_t)(dv_cjk_code(win->raw, wi - 3 * (len - 1 - t)) - 0x4E00u + 1u) : 0ULL);
                        k2 = k;
                    } else {
                        uint64_t h = DV_FNV_SEED;
                        for (int t = 0; t < len; t++) h = dv_fnv_step(h, dv_cjk_code(win->raw, wi - 3 * (len - 1 - t)));
                        k2 = (3ULL << 62) | dv_hash62_finish(h, len);
                    }
                    v2 = true;
                }
            } else if (isA && !nextA) {
                const int len = j - my_as + 1;
                if (len >= 2) {
                    if (len <= 10) {
                        uint64_t k = 0;
                        for (int t = 0; t < len; t++) k = (k << 6) | (uint64_t)win->cell[wi - (len - 1 - t)];
                        k2 = (1ULL << 62) | (k << (6 * (10 - len)));
                    } else {                                          // rare: read the run from global memory
                        uint64_t h = DV_FNV_SEED;
                        for (int32_t q = my_as; q <= j; q++) h = dv_fnv_step(h, dv_ascii_code_to_cp(dv_ascii_cell(text[q])));
                        k2 = (2ULL << 62) | dv_hash62_finish(h, len);
                    }
                    v2 = true;
                }
            }
            emit(v1, k1);
            emit(v2, k2);
        }
        __syncwarp();
        // ---- slide: keep the last DV_BACK bytes of the window as look-behind
        if (lane < DV_BACK / 16) {
            uint4 r = *reinterpret_cast<uint4*>(win->raw + DV_TILE + lane * 16);
            uint4 c = *reinterpret_cast<uint4*>(win->cell + DV_TILE + lane * 16);
            *reinterpret_cast<uint4*>(win->raw + lane * 16) = r;
            *reinterpret_cast<uint4*>(win->cell + lane * 16) = c;
        }
        __syncwarp();
    }
    return false;
}

template <class Emit>
__device__ __forceinline__ void dv_tokenize_v3_warp(const uint8_t* text, int32_t n, dv_warp_window* win, Emit& emit) {
    if (n <= 0) return;
    if (dv_tokenize_fast(text, n, win, emit)) {
        emit.restart();
        dv_tokenize_serial(text, n, win, emit);
    }
}
