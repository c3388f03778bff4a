// simt_emu_intrinsics.h — CUDA intrinsics used by the repo's kernels, for the CPU SIMT emulator (test infrastructure).
#pragma once

#define EMU_FULL(m) do { if ((m) != 0xffffffffu) { fprintf(stderr, "simt_emu: partial-mask intrinsic at %s:%d\n", __FILE__, __LINE__); abort(); } } while (0)

static inline void __syncthreads() { emu::block_barrier(); }
static inline void __syncwarp(unsigned m = 0xffffffffu) { EMU_FULL(m); emu::warp_barrier(); }

// all lanes publish v, meet, and read what they need
#define EMU_XCHG_BEGIN(v) emu::Warp* w_ = emu::tw; const int b_ = emu::warp_arrive_prepare(); \
    w_->xchg[b_][emu::tlane] = (unsigned long long)(v); emu::warp_barrier();

static inline unsigned __ballot_sync(unsigned m, int pred) {
    EMU_FULL(m);
    EMU_XCHG_BEGIN(pred ? 1 : 0)
    unsigned r = 0;
    for (int l = 0; l < 32; l++) if (!w_->done[l] && w_->xchg[b_][l]) r |= 1u << l;
    return r;
}
static inline int __any_sync(unsigned m, int pred) { return __ballot_sync(m, pred) != 0u; }
static inline int __all_sync(unsigned m, int pred) { return __ballot_sync(m, !pred) == 0u; }
static inline unsigned __activemask() { return 0xffffffffu; }

template <typename T> static inline T emu_shfl_get(unsigned m, T v, int src, bool ok) {
    EMU_FULL(m);
    static_assert(sizeof(T) <= 8, "shuffle of <= 64-bit values");
    unsigned long long bits = 0; memcpy(&bits, &v, sizeof(T));
    EMU_XCHG_BEGIN(bits)
    if (!ok) return v;
    unsigned long long r = w_->xchg[b_][src & 31];
    T out; memcpy(&out, &r, sizeof(T));
    return out;
}
template <typename T> static inline T __shfl_sync(unsigned m, T v, int src) { return emu_shfl_get(m, v, src & 31, true); }
template <typename T> static inline T __shfl_up_sync(unsigned m, T v, unsigned d) { const int s = emu::tlane - (int)d; return emu_shfl_get(m, v, s, s >= 0); }
template <typename T> static inline T __shfl_down_sync(unsigned m, T v, unsigned d) { const int s = emu::tlane + (int)d; return emu_shfl_get(m, v, s, s < 32); }
template <typename T> static inline T __shfl_xor_sync(unsigned m, T v, int x) { return emu_shfl_get(m, v, emu::tlane ^ x, true); }

template <typename T> static inline unsigned __match_any_sync(unsigned m, T v) {
    EMU_FULL(m);
    unsigned long long bits = 0; memcpy(&bits, &v, sizeof(T));
    EMU_XCHG_BEGIN(bits)
    unsigned r = 0;
    for (int l = 0; l < 32; l++) if (!w_->done[l] && w_->xchg[b_][l] == bits) r |= 1u << l;
    return r;
}
static inline unsigned __reduce_or_sync(unsigned m, unsigned v) {
    EMU_FULL(m);
    EMU_XCHG_BEGIN(v)
    unsigned r = 0;
    for (int l = 0; l < 32; l++) if (!w_->done[l]) r |= (unsigned)w_->xchg[b_][l];
    return r;
}
static inline unsigned __reduce_add_sync(unsigned m, unsigned v) {
    EMU_FULL(m);
    EMU_XCHG_BEGIN(v)
    unsigned r = 0;
    for (int l = 0; l < 32; l++) if (!w_->done[l]) r += (unsigned)w_->xchg[b_][l];
    return r;
}

// ---- bit tricks
static inline int __popc(unsigned x) { return __builtin_popcount(x); }
static inline int __popcll(unsigned long long x) { return __builtin_popcountll(x); }
static inline int __clz(int x) { return x ? __builtin_clz((unsigned)x) : 32; }
static inline int __ffs(int x) { return __builtin_ffs(x); }
static inline unsigned __brev(unsigned x) { unsigned r = 0; for (int i = 0; i < 32; i++) if (x & (1u << i)) r |= 1u << (31 - i); return r; }
static inline unsigned __funnelshift_r(unsigned lo, unsigned hi, unsigned sh) { sh &= 31u; return sh ? (lo >> sh) | (hi << (32 - sh)) : lo; }
static inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned sh) { sh &= 31u; return sh ? (hi << sh) | (lo >> (32 - sh)) : hi; }
static inline unsigned __byte_perm(unsigned a, unsigned b, unsigned s) {
    const unsigned long long v = ((unsigned long long)b << 32) | a;
    unsigned r = 0;
    for (int i = 0; i < 4; i++) r |= (unsigned)((v >> (8 * ((s >> (4 * i)) & 7u))) & 0xFFu) << (8 * i);
    return r;
}
static inline unsigned __dp4a(unsigned a, unsigned b, unsigned c) {
    for (int i = 0; i < 4; i++) c += ((a >> (8 * i)) & 0xFFu) * ((b >> (8 * i)) & 0xFFu);
    return c;
}
static inline unsigned __umulhi(unsigned a, unsigned b) { return (unsigned)(((unsigned long long)a * b) >> 32); }

// ---- fp64 in the reference's operation order (compile the emulator with -ffp-contract=off)
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
static inline double __ddiv_rn(double a, double b) { return a / b; }

// ---- memory
template <typename T> static inline T __ldcg(const T* p) { return __atomic_load_n(p, __ATOMIC_RELAXED); }
template <typename T> static inline T __ldg(const T* p) { return *p; }
static inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
static inline void __threadfence_block() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }

template <typename T> static inline T atomicCAS(T* p, T cmp, T val) {
    __atomic_compare_exchange_n(p, &cmp, val, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST);
    return cmp;
}
template <typename T> static inline T atomicOr(T* p, T v) { return __atomic_fetch_or(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicAnd(T* p, T v) { return __atomic_fetch_and(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicAdd(T* p, T v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicExch(T* p, T v) { return __atomic_exchange_n(p, v, __ATOMIC_SEQ_CST); }
template <typename T> static inline T atomicMax(T* p, T v) {
    T old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
    while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {}
    return old;
}
