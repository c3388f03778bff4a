// text8.cuh — byte-mask scanner for long texts (the facts of the bind kernel): no per-character compaction.
//
// Same token semantics as text.cuh / text6.cuh (_tokenize_similarity_text_v3, reference web/server.py:24722-24732).
// A text is cut into balanced tiles of 32 lane chunks (8..32 bytes each, 8-byte aligned, read straight from global
// memory).  Every lane classifies its chunk with SWAR arithmetic on the raw bytes into two bit masks:
//   C  byte is the lead of a CJK character U+4E00..U+9FFF (E4 B8 80 .. E9 BF BF: lead E4..E9, and not E4 + second byte < B8)
//   A  byte is in [0-9A-Za-z_]
// Because every CJK character is 3 bytes long, everything the tokenizer needs is shift arithmetic on these masks
// (look-BACK only, so a tile never needs the classification of the next tile):
//   bigram (previous, this) ends at p      B  = C & (C << 3)
//   run start                              S  = C & ~(C << 3)
//   position right after a CJK run         PE = (C << 3) & ~C
//   position right after an ASCII run >= 2 AE = (A << 1) & (A << 2) & ~A
// A run that began `rel` bytes before the chunk is cut into 12-character pieces at its byte offsets 36 m: at that
// position the bigram is dropped (it would span two pieces) and a 12-character piece ends.  Only a run carried into
// the chunk can reach such an offset inside it (36 > 32).  Run starts / last non-alphanumeric positions of earlier
// lanes travel by ballot + one shuffle.
// Work left per lane: one table probe per set bit of B (the two characters are decoded from a shared-memory copy of the
// tile), and per set bit of PE | AE a signature + filter test, the 64-bit key only for filter survivors.
//
// The scanner only handles what the bind kernel's fact phase needs (probe, never insert) and hands the text back
// (returns true) when it cannot be exact or cheap: U+0130 / U+212A (bytes C4 xx / E2 84 xx), an ASCII token longer
// than 64, or fewer than 18 distinct bigrams among the first bigram of every lane (min(len(B), 18) is then counted by
// the exact path).  Hits are bits in a set, so whatever was marked before a hand-back stays correct.
//
// MEASURED AND NOT ADOPTED (round 1, session 10).  Wired into bind6.cuh as a first pass over every fact >= 192 bytes
// (Probe8Fact = the probe halves of Emit6Fact), bit-exact on the whole GPU parity suite including
// tests/test_gpu_parity.py::test_a2_every_pair_of_long_facts, it executed 0.60 G warp instructions per 2 000 sessions
// (current kernel 0.66 G) but ran 1.04 ms instead of 0.84 ms (4.58 vs 4.29 ms at 10 000 sessions): the per-lane
// loops leave 15 of 32 threads active on average (token loop: 6-7, key loops: 1.3), the larger code adds
// instruction-cache stalls (no_instruction 1.9 cycles per issue), and the probe + epilogue + claim phase (35 % of the
// instructions) are common to both designs.  Per 768-byte tile: classify 300, bigram loop 9 trips x 77, token loop 625.
#pragma once
#include "text6.cuh"

#define DV8_HIST 64                                   // bytes of the previous tile kept in front of the tile
#define DV8_WORDS (DV8_HIST / 4 + 256 + 4)            // history | tile (<= 1024 B) | look-ahead
#define DV8_MIN_BYTES 192
#define DV8_NONE (-(1 << 28))

// 4 bytes of the staged tile starting at byte address ba (little endian: byte ba in bits 0..7)
__device__ __forceinline__ uint32_t dv8_win(const uint32_t* W, int32_t ba) {
    return __funnelshift_r(W[ba >> 2], W[(ba >> 2) + 1], (uint32_t)(ba & 3) * 8u);
}
// symbol (code - 0x4E00 + 1) of the 3-byte character in bits 0..23 of x
__device__ __forceinline__ uint32_t dv8_cjk_sym(uint32_t x) {
    return (((x & 0x0Fu) << 12) | ((x >> 2) & 0xFC0u) | ((x >> 16) & 0x3Fu)) - 0x4DFFu;
}
__device__ __forceinline__ uint32_t dv8_gather(uint32_t flags) {            // bits 7,15,23,31 -> bits 0..3
    return (((flags >> 7) & 0x01010101u) * 0x00204081u >> 21) & 0xFu;
}

// Probe: bool ft_pass(sig); void bigram(id); void token(key)   (all callable from divergent code)
template <class Probe>
__device__ bool dv8_scan(const uint8_t* text, int32_t n, uint32_t* W, const unsigned long long* ktab, const uint8_t* lut,
                         Probe& probe) {
    const int lane = threadIdx.x & 31;
    const uint32_t lt = (1u << lane) - 1u;
    const int32_t a = (int32_t)((uintptr_t)text & 7u);
    const uint8_t* g0 = text - a;                         // 8-byte aligned; stream coordinate = byte offset from g0
    const int32_t total = a + n;                          // stream position of the terminator
    const int32_t ntiles = (total + 1024) >> 10;          // positions 0..total inclusive
    const int32_t nu = (total + ntiles * 256) / (ntiles * 256);     // 8-byte units per lane, 1..4
    const int32_t CB = nu * 8, TB = CB * 32;
    const uint32_t CM = CB == 32 ? 0xFFFFFFFFu : ((1u << CB) - 1u);
    const uint8_t* Wb = reinterpret_cast<const uint8_t*>(W);
    uint32_t carryC = 0u, carryA = 0u;
    int32_t rs_carry = DV8_NONE, na_carry = a - 1;
    bool bad = false;
    if (lane < DV8_HIST / 4) W[lane] = 0u;
    for (int32_t t0 = 0; t0 <= total; t0 += TB) {
        const int32_t s0 = t0 + lane * CB;
        // ---- load the chunk (+ 8 bytes of look-ahead = the next lane's first unit)
        uint32_t w[9];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            uint2 v = make_uint2(0u, 0u);
            if (u < nu && s0 + 8 * u < total) v = *reinterpret_cast<const uint2*>(g0 + s0 + 8 * u);
            w[2 * u] = v.x; w[2 * u + 1] = v.y;
        }
        uint32_t nx0 = __shfl_down_sync(0xffffffffu, w[0], 1), nx1 = __shfl_down_sync(0xffffffffu, w[1], 1);
        if (lane == 31) {
            uint2 v = make_uint2(0u, 0u);
            if (t0 + TB < total) v = *reinterpret_cast<const uint2*>(g0 + t0 + TB);
            nx0 = v.x; nx1 = v.y;
        }
        // ---- stage: history = last 64 bytes of the previous tile
        __syncwarp();
        if (t0 > 0) {
            uint32_t h = 0u;
            if (lane < DV8_HIST / 4) h = W[DV8_HIST / 4 + (TB >> 2) - DV8_HIST / 4 + lane];
            __syncwarp();
            if (lane < DV8_HIST / 4) W[lane] = h;
        }
        {
            uint32_t* dst = W + DV8_HIST / 4 + lane * (CB >> 2);
#pragma unroll
            for (int u = 0; u < 4; u++) if (u < nu) *reinterpret_cast<uint2*>(dst + 2 * u) = make_uint2(w[2 * u], w[2 * u + 1]);
            if (lane == 31) { dst[CB >> 2] = nx0; dst[(CB >> 2) + 1] = nx1; }
        }
        // ---- classify
        uint32_t C = 0u, A = 0u, exo = 0u;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (j < 2 * nu) {
                const uint32_t x = w[j];
                const uint32_t nxt = (j + 1 == 2 * nu) ? nx0 : w[j + 1];
                const uint32_t nb = __funnelshift_r(x, nxt, 8);                 // the byte after each byte
                const uint32_t l7 = x & 0x7F7F7F7Fu, n7 = nb & 0x7F7F7F7Fu;
                const uint32_t inr = (l7 + 0x1C1C1C1Cu) & ~(l7 + 0x16161616u);   // low 7 bits in 0x64..0x69
                const uint32_t e4 = ~((l7 ^ 0x64646464u) + 0x7F7F7F7Fu);         // == 0x64
                const uint32_t nlt = ~(n7 + 0x48484848u);                        // next low 7 bits < 0x38
                const uint32_t c = x & inr & ~(e4 & nlt);
                const uint32_t f = l7 | 0x20202020u;
                const uint32_t letter = (f + 0x1F1F1F1Fu) & ~(f + 0x05050505u);   // 0x61..0x7A
                const uint32_t digit = (l7 + 0x50505050u) & ~(l7 + 0x46464646u); // 0x30..0x39
                const uint32_t us = ~((l7 ^ 0x5F5F5F5Fu) + 0x7F7F7F7Fu);
                const uint32_t al = ~x & (letter | digit | us);
                const uint32_t c4 = ~((l7 ^ 0x44444444u) + 0x7F7F7F7Fu);          // C4 lead (U+0130)
                const uint32_t e2 = ~((l7 ^ 0x62626262u) + 0x7F7F7F7Fu) & ~((n7 ^ 0x04040404u) + 0x7F7F7F7Fu);   // E2 84 (U+212A)
                exo |= x & (c4 | e2);
                C |= dv8_gather(c) << (4 * j);
                A |= dv8_gather(al) << (4 * j);
            }
        }
        {
            const int32_t lo = a > s0 ? a - s0 : 0, hi = total - s0;             // valid positions [lo, hi) of the chunk
            uint32_t V = 0u;
            if (hi > lo && lo < CB) V = (hi >= 32 ? 0xFFFFFFFFu : ((1u << hi) - 1u)) & (0xFFFFFFFFu << lo) & CM;
            C &= V; A &= V;
            exo &= 0x80808080u;
        }
        if (__any_sync(0xffffffffu, exo != 0u && s0 < total)) return true;
        // ---- neighbours
        uint32_t Cp = __shfl_up_sync(0xffffffffu, C, 1), Ap = __shfl_up_sync(0xffffffffu, A, 1);
        if (lane == 0) { Cp = carryC; Ap = carryA; }
        carryC = __shfl_sync(0xffffffffu, C, 31); carryA = __shfl_sync(0xffffffffu, A, 31);
        const uint32_t Csh = ((C << 3) | (Cp >> (CB - 3))) & CM;
        const uint32_t Ash1 = ((A << 1) | (Ap >> (CB - 1))) & CM, Ash2 = ((A << 2) | (Ap >> (CB - 2))) & CM;
        uint32_t B = C & Csh;
        const uint32_t S = C & ~Csh;
        uint32_t PE = Csh & ~C;
        const uint32_t AE = Ash1 & Ash2 & ~A;
        const uint32_t NA = ~A & CM;
        int32_t rs_in, na_in;
        {
            const int32_t lastS = S ? s0 + 31 - __clz(S) : DV8_NONE;
            const uint32_t hasS = __ballot_sync(0xffffffffu, S != 0u);
            const uint32_t bef = hasS & lt;
            const int32_t got = __shfl_sync(0xffffffffu, lastS, bef ? 31 - __clz(bef) : 0);
            rs_in = bef ? got : rs_carry;
            const int32_t top = __shfl_sync(0xffffffffu, lastS, hasS ? 31 - __clz(hasS) : 0);
            if (hasS) rs_carry = top;
            const int32_t lastN = NA ? s0 + 31 - __clz(NA) : DV8_NONE;
            const uint32_t hasN = __ballot_sync(0xffffffffu, NA != 0u);
            const uint32_t befn = hasN & lt;
            const int32_t gotn = __shfl_sync(0xffffffffu, lastN, befn ? 31 - __clz(befn) : 0);
            na_in = befn ? gotn : na_carry;
            const int32_t topn = __shfl_sync(0xffffffffu, lastN, hasN ? 31 - __clz(hasN) : 0);
            if (hasN) na_carry = topn;
        }
        // ---- 12-character piece boundary of the run carried into the chunk
        {
            const uint32_t rel = (uint32_t)(s0 - rs_in);
            const uint32_t r36 = rel % 36u;
            const uint32_t e = r36 ? 36u - r36 : 0u;
            if (e < (uint32_t)CB && rel != 0u && !(S & ((2u << e) - 1u))) {
                B &= ~(1u << e);
                PE |= Csh & (1u << e);
            }
        }
        __syncwarp();                                                            // the staged tile is visible
        const int32_t cb0 = DV8_HIST + lane * CB;                                // shared-memory byte address of chunk position 0
        // ---- min(len(B), 18): distinct ids among the first bigram of every lane (first tile only)
        if (t0 == 0) {
            uint32_t id = 0x80000000u | (uint32_t)lane;
            if (B) {
                const int32_t ba = cb0 + (__ffs(B) - 1) - 3;
                const uint32_t lo = dv8_win(W, ba), hi = dv8_win(W, ba + 3);
                id = (dv8_cjk_sym(lo) << 15) | dv8_cjk_sym(hi);
            }
            const uint32_t mm = __match_any_sync(0xffffffffu, id);
            const uint32_t nd = __popc(__ballot_sync(0xffffffffu, B != 0u && (__ffs(mm) - 1 == lane)));
            if (nd < 18u) return true;
        }
        // ---- bigrams
        {
            uint32_t bb = B;
            while (bb) {
                const int32_t p = __ffs(bb) - 1; bb &= bb - 1u;
                const int32_t ba = cb0 + p - 3;
                const uint32_t* q = W + (ba >> 2);
                const uint32_t sh = (uint32_t)(ba & 3) * 8u;
                const uint32_t w0 = q[0], w1 = q[1], w2 = q[2];
                const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);   // bytes ba..ba+3, ba+4..ba+7
                const uint32_t c1 = dv8_cjk_sym(lo);
                const uint32_t c2 = ((((lo >> 24) & 0x0Fu) << 12) | ((hi & 0x3Fu) << 6) | ((hi >> 8) & 0x3Fu)) - 0x4DFFu;
                probe.bigram((c1 << 15) | c2);
            }
        }
        // ---- tokens: pieces of 3..12 characters and ASCII runs
        {
            uint32_t te = PE | AE;
            while (te) {
                const int32_t p = __ffs(te) - 1; te &= te - 1u;
                const bool isC = (PE >> p) & 1u;
                const int32_t endb = cb0 + p;
                const uint32_t below = (1u << p) - 1u;
                uint32_t len, s_prev, s_last; int32_t startb;
                if (isC) {
                    const uint32_t sb = S & below;
                    const int32_t rs = sb ? s0 + 31 - __clz(sb) : rs_in;
                    const uint32_t k = (uint32_t)(s0 + p - rs) / 3u;
                    const uint32_t r = k % 12u;
                    len = r ? r : 12u;
                    if (len < 3u) continue;
                    const uint32_t lo = dv8_win(W, endb - 6), hi = dv8_win(W, endb - 3);
                    s_prev = dv8_cjk_sym(lo); s_last = dv8_cjk_sym(hi);
                    startb = endb - 3 * (int32_t)len;
                } else {
                    const uint32_t nb_ = NA & below;
                    const int32_t nap = nb_ ? s0 + 31 - __clz(nb_) : na_in;
                    len = (uint32_t)(s0 + p - nap - 1);
                    if (len > 64u) { bad = true; continue; }
                    s_prev = DV_SYM_ASCII | lut[Wb[endb - 2] & 0x7Fu]; s_last = DV_SYM_ASCII | lut[Wb[endb - 1] & 0x7Fu];
                    startb = endb - (int32_t)len;
                }
                if (!probe.ft_pass(dv_token_sig(s_prev, s_last, len))) continue;
                uint64_t mm = 0;
                if (isC) for (uint32_t q = 0; q < len; q++) mm += (uint64_t)dv8_cjk_sym(dv8_win(W, startb + 3 * (int32_t)q)) * ktab[q];
                else for (uint32_t q = 0; q < len; q++) mm += (uint64_t)lut[Wb[startb + (int32_t)q] & 0x7Fu] * ktab[q];
                probe.token(dv_token_key(mm, len, isC));
            }
        }
        if (__any_sync(0xffffffffu, bad)) return true;
    }
    return false;
}
