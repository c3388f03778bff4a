// text6.cuh — lean warp tokenizer for the bind kernel (same token semantics as text.cuh, fewer instructions per byte).
//
// Follows _tokenize_similarity_text_v3 (reference web/server.py:24722-24732); symbols, bigram ids, token keys and
// signatures are the ones text.cuh defines.  One warp streams one text in 256-byte tiles:
//   stage 1   8 bytes per lane; a byte ends a character iff the NEXT byte is not a UTF-8 continuation byte (two SWAR
//             words + the first byte of the next lane); the byte offsets of the character ends are compacted to pos[].
//   stage 2a  one character end per lane: the 4 bytes ending there are fetched from the shared-memory copy of the
//             tile and decoded to a symbol (full lane utilisation; no per-lane decode loops).
//   stage 2b  32 symbols per step, only look-BACK: the lane of symbol i emits the bigram (i-1, i) and the token that
//             ENDS at i-1 (piece / ASCII run); a terminator symbol after the text closes the last token.  Symbols left
//             over at a tile end (< 32) are carried into the next tile so every step runs on 32 lanes.
// Restart conditions (exotic case mappings, ASCII token longer than the history) as in text.cuh.
#pragma once
#include "text.cuh"

#define DV6_TILE 256
#define DV6_HIST 64
#define DV6_SYMCAP (DV6_HIST + 32 + DV6_TILE + 8)

struct __align__(16) dv6_warp_buf {
    uint32_t raw[2 + DV6_TILE / 4 + 2];   // [1] = last word of the previous tile, [2..66) = tile bytes
    union {
        struct { uint16_t sym[DV6_SYMCAP]; uint8_t pos[DV6_TILE]; };
        struct { unsigned long long queue[32]; uint32_t qsig[32]; uint32_t qkind; };     // serial-path scratch
    };
};

// lut[b] for b < 0x80: ASCII class code (1..37) or 0
__device__ __forceinline__ void dv6_fill_lut(uint8_t* lut) {
    for (int b = threadIdx.x; b < 128; b += blockDim.x) lut[b] = (uint8_t)(dv_ascii_sym((uint32_t)b) & 0xFFu);
}

// Returns true (warp-uniform) if the text must restart on the serial path.
template <class Emit>
__device__ bool dv6_tokenize_fast(const uint8_t* text, int32_t n, dv6_warp_buf* wb, const unsigned long long* ktab,
                                  const uint8_t* lut, Emit& emit) {
    const int lane = threadIdx.x & 31;
    const int32_t a = (int32_t)((uintptr_t)text & 7u);
    uint16_t* B = wb->sym;
    uint32_t* R = wb->raw;
    uint8_t* P = wb->pos;
    B[lane] = 0; B[lane + 32] = 0;
    int pending = 0;
    uint32_t carry_word = 0;
    int32_t gbase = 0;                             // global symbol index of B[DV6_HIST]
    int32_t rstart = 0;                            // global index of the most recent run start
    const uint32_t lt = (1u << lane) - 1u;
    // software pipeline: the loads of tile t+1 are issued before tile t is processed
    uint2 wn = make_uint2(0u, 0u); uint32_t en = 0u;
    {
        const int32_t p0 = -a + lane * 8;
        if (p0 < n && p0 + 8 > 0) wn = *reinterpret_cast<const uint2*>(text + p0);
        if (lane == 31 && -a + DV6_TILE < n) en = (uint32_t)text[-a + DV6_TILE];
    }
    for (int32_t tile = -a; tile < n; tile += DV6_TILE) {
        // ---- stage 1
        const int32_t p0 = tile + lane * 8;
        const uint2 w = wn;
        uint32_t nbw = __shfl_down_sync(0xffffffffu, w.x, 1);
        if (lane == 31) nbw = en;
        {
            const int32_t q0 = p0 + DV6_TILE;
            wn = make_uint2(0u, 0u); en = 0u;
            if (q0 < n) wn = *reinterpret_cast<const uint2*>(text + q0);
            if (lane == 31 && tile + 2 * DV6_TILE < n) en = (uint32_t)text[tile + 2 * DV6_TILE];
        }
        const uint32_t cw = __shfl_sync(0xffffffffu, w.y, 31);
        __syncwarp();                                                   // previous tile's decodes are done
        if (lane == 0) R[1] = carry_word;
        carry_word = cw;
        *reinterpret_cast<uint2*>(R + 2 + lane * 2) = w;
        const uint32_t cx = w.x & ~(w.x << 1) & 0x80808080u;            // continuation bytes (10xxxxxx)
        const uint32_t cy = w.y & ~(w.y << 1) & 0x80808080u;
        const uint32_t cn = nbw & ~(nbw << 1) & 0x80u;
        const uint32_t ex_ = ~__funnelshift_r(cx, cy, 8) & 0x80808080u; // byte t ends a character: byte t+1 is no continuation
        const uint32_t ey_ = ~__funnelshift_r(cy, cn, 8) & 0x80808080u;
        uint32_t m8 = ((((ex_ >> 7) * 0x00204081u) >> 21) & 0xFu) | (((((ey_ >> 7) * 0x00204081u) >> 21) & 0xFu) << 4);
        {
            const int lo = p0 < 0 ? -p0 : 0;
            int hi = n - p0;
            if (hi >= 1 && hi <= 8) m8 |= 1u << (hi - 1);               // the text's last byte always ends a character
            hi = hi > 8 ? 8 : hi;
            const uint32_t inm = hi > lo ? ((0xFFu >> (8 - hi)) & (0xFFu << lo)) : 0u;
            m8 &= inm;
        }
        const int cnt = __popc(m8);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += y; }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        {
            int widx = incl - cnt;
            const uint32_t tb = (uint32_t)lane * 8u - 1u;
            while (m8) { P[widx++] = (uint8_t)(tb + __ffs(m8)); m8 &= m8 - 1; }
        }
        __syncwarp();
        // ---- stage 2a: decode one character end per lane
        bool exotic = false;
        uint16_t* Bn = B + DV6_HIST + pending;
        for (int k0 = 0; k0 < total; k0 += 32) {
            const int k = k0 + lane;
            if (k < total) {
                const uint32_t o = (uint32_t)P[k] + 5u;                 // tile byte t sits at raw byte 8 + t; window = t-3..t
                const uint32_t x = __funnelshift_r(R[o >> 2], R[(o >> 2) + 1], (o & 3u) * 8u);
                const uint32_t b = x >> 24;
                const uint32_t ac = lut[b & 0x7Fu];                     // ASCII class code (valid when b < 0x80)
                const uint32_t code = ((((x >> 8) & 0x0Fu) << 12) | (((x >> 16) & 0x3Fu) << 6) | (b & 0x3Fu)) - 0x4E00u;
                const bool is3 = (x & 0x00C0F000u) == 0x0080E000u;      // E-lead two bytes back, continuation one byte back
                uint32_t s = (is3 & (code < 20992u)) ? code + 1u : (((x >> 8) == 0x00AA84E2u) ? (DV_SYM_ASCII | 11u) : 0u);   // U+212A -> 'k'
                s = (b < 0x80u) ? (ac ? (DV_SYM_ASCII | ac) : 0u) : s;
                exotic |= (x & 0x80FF0000u) == 0x80C40000u;             // C4 lead (U+0130 lower-cases to 'i' + U+0307)
                Bn[k] = (uint16_t)s;
            }
        }
        if (__any_sync(0xffffffffu, exotic)) return true;
        const bool last = (tile + DV6_TILE >= n);
        int avail = pending + total;
        if (last) { if (lane == 0) B[DV6_HIST + avail] = 0; avail++; }   // terminator closes the last token
        const int nproc = last ? avail : (avail & ~31);
        __syncwarp();
        // ---- stage 2b
        bool too_long = false;
        for (int s0 = 0; s0 < nproc; s0 += 32) {
            const int i = DV6_HIST + s0 + lane;
            const bool act = s0 + lane < nproc;
            const uint32_t sv = act ? B[i] : 0u;
            const uint32_t sm1 = act ? B[i - 1] : 0u;
            const uint32_t sm2 = B[i - 2];
            const bool curC = dv_is_cjk_sym(sv), curA = dv_is_ascii_sym(sv);
            const bool pC = dv_is_cjk_sym(sm1), pA = dv_is_ascii_sym(sm1);
            const int32_t step_g = gbase + s0;
            const uint32_t smask = __ballot_sync(0xffffffffu, (curC && !pC) || (curA && !pA));
            const uint32_t before = smask & lt;
            const int32_t ps = before ? step_g + (31 - __clz(before)) : rstart;      // start of the run holding symbol i-1
            if (smask) rstart = step_g + (31 - __clz(smask));
            const uint32_t rp = (uint32_t)(step_g + lane - 1 - ps);                  // position of symbol i-1 inside its run
            const uint32_t dp = rp % 12u;
            emit.bigram(pC && curC && dp != 11u, (sm1 << 15) | sv);
            const bool tk = pC ? ((dp == 11u || !curC) && dp >= 2u) : (pA && !curA && rp >= 1u);
            const uint32_t len = (pC ? dp : rp) + 1u;
            const uint32_t sig = dv_token_sig(sm2, sm1, len);
            const bool need = tk & emit.need_token(sig);
            if (__any_sync(0xffffffffu, need)) {
                uint64_t key = 0;
                if (need) {
                    if (len > DV6_HIST) too_long = true;
                    else {
                        uint64_t mm = 0;
                        const uint16_t* tp = B + (i - (int)len);
                        for (uint32_t q = 0; q < len; q++) mm += (uint64_t)(tp[q] & 0x7FFFu) * ktab[q];
                        key = dv_token_key(mm, len, pC);
                    }
                }
                if (__any_sync(0xffffffffu, too_long)) return true;
                emit.token(need, key, sig);
            }
        }
        // ---- slide: keep the last DV6_HIST processed symbols and the unprocessed remainder
        __syncwarp();
        {
            const int keep = DV6_HIST + (avail - nproc);
            uint16_t t0 = 0, t1 = 0, t2 = 0;
            if (lane < keep) t0 = B[nproc + lane];
            if (lane + 32 < keep) t1 = B[nproc + lane + 32];
            if (lane + 64 < keep) t2 = B[nproc + lane + 64];
            __syncwarp();
            if (lane < keep) B[lane] = t0;
            if (lane + 32 < keep) B[lane + 32] = t1;
            if (lane + 64 < keep) B[lane + 64] = t2;
            pending = avail - nproc;
        }
        gbase += nproc;
    }
    return false;
}

template <class Emit>
__device__ __forceinline__ void dv6_tokenize_warp(const uint8_t* text, int32_t n, dv6_warp_buf* wb, const unsigned long long* ktab,
                                                  const uint8_t* lut, Emit& emit) {
    if (n <= 0) return;
    if (dv6_tokenize_fast(text, n, wb, ktab, lut, emit)) {
        emit.restart();
        __syncwarp();
        dv_tokenize_serial(text, n, wb, ktab, emit);
    }
}
