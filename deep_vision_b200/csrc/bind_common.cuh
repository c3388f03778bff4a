// bind_common.cuh — launch parameters and constants shared by the bind kernels (bind_general.cuh, bind7.cuh) and the CPU
// SIMT emulator of the tests (tests/emu).  replaces: nothing in the reference; plumbing of dvb_bind_claims.
#pragma once
#include <stdint.h>

#define BIND_WARPS 8
#define BIND_THREADS (BIND_WARPS * 32)
#define BIND_CLAIMS 32          // claims per group (bits of cmask)
#define BIND_WSET 64            // per-warp distinct-token set slots (>= 17 + 32)
#define BIND_FB_BITS 16         // bigram filter: 2^16 bits = 8 KB
#define BIND_FT_BITS 13         // token filter: 2^13 bits = 1 KB
#define BIND_QCAP 64
#define BIND_MAX_PARTS 8        // split fact kernel: at most this many CTAs per claim group

// a lane that polls memory another lane (possibly of its own warp) is about to write: the GPU schedules diverged lanes
// independently; the CPU emulator of the tests runs lanes as cooperative fibers and has to be told to switch
#ifdef DV_EMU
#define DV_SPIN_PAUSE() emu::yield()
#else
#define DV_SPIN_PAUSE() ((void)0)
#endif

struct BindParams {
    const uint8_t* claim_utf8; const int64_t* claim_off; const uint8_t* fact_utf8; const int64_t* fact_off;
    const int64_t* sess_claim_off; const int64_t* sess_fact_off; int64_t s_begin, s_end;   // session range of this launch
    const double* fact_quality; const int32_t* fact_dim; const uint8_t* fact_hard;
    const int32_t* claim_pref_dim; const double* claim_threshold;
    int32_t* best_fact; double* best_score; uint8_t* bound; int32_t* len_claim; int32_t* len_fact;
    uint32_t cap;                         // table capacity (power of two)
    unsigned long long* gtable;           // per CTA: cap keys (u64) then cap cmasks (u32)
    uint32_t* ghit;                       // per-warp hit bitmaps in global memory when cap/32 words do not fit smem, else null
    int* overflow;                        // fast kernel: raised (host-mapped word) when at least one session was handed back
    uint8_t* sess_fallback;               // fast kernel: sess_fallback[s] = 1 -> session s must be scored by the general kernel
    const uint8_t* only_flagged;          // general kernel: when non-null, score only sessions with only_flagged[s] != 0
    uint8_t* slabs;                       // fast kernels: one slab of B7_BLOB_BYTES per claim group (claim kernel -> fact kernel)
    int* slab_cursor;                     // fast kernels: claim groups (= slabs in use) of this launch
    int32_t* sess_slab;                   // fast kernels: first slab of session s
    int32_t* slab_sess;                   // fast kernels: session of slab u (-1 - s once the claim kernel handed the group back)
    int parts;                            // split fact kernel: CTAs per claim group (each walks 1 / parts of the session's facts)
    int* part_cnt;                        // fast kernels: per group, parts that have finished (zeroed by k_bind_groups)
    double* part_score; int* part_idx;    // split fact kernel: [group * parts + part][32] best score / fact per claim
};

__device__ __forceinline__ uint32_t bind_hash(uint64_t key) {
    uint32_t x = (uint32_t)key * 0x85EBCA6Bu ^ (uint32_t)(key >> 32) * 0xC2B2AE35u;
    x ^= x >> 16; x *= 0x9E3779B1u; x ^= x >> 15;
    return x;
}
__device__ __forceinline__ uint32_t bind_fb_index(uint32_t bigram_id) { return (bigram_id * 0x9E3779B1u) >> (32 - BIND_FB_BITS); }
__device__ __forceinline__ uint32_t bind_ft_index(uint32_t sig) { return sig >> (32 - BIND_FT_BITS); }

