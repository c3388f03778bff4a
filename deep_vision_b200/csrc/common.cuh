// common.cuh — ctx, stream/workspace leases and error plumbing for libdvb200.so (sm_100a).
#pragma once
#include <chrono>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include <string>
#include <vector>
#include <atomic>
#include <thread>
#include <algorithm>
#include "../../include/dvb200.h"

#define DVB_NUM_SM_FALLBACK 148
#define DVB_LEASES 4

struct dvb_arena {
    uint8_t* base = nullptr;
    size_t cap = 0, used = 0;
};

struct dvb_lease {
    cudaStream_t stream = nullptr;
    dvb_arena arena;
    std::vector<void*> overflow;         // extra blocks of the previous call (arena was too small)
    size_t high_water = 0;               // bytes the previous calls needed in total
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaStream_t copy = nullptr;         // H2D stream of pipelined host-buffer calls (created on first use)
    std::vector<cudaEvent_t> pipe_ev;    // one per in-flight input slice
    int* flag_h = nullptr;               // pinned + mapped word a kernel can raise (read by the host after the stream sync)
    int* flag_d = nullptr;
    uint8_t* small_h = nullptr;          // pinned 2 x DVB_SMALL_STAGE bytes: all inputs / all outputs of a small host-buffer call in ONE copy each
    uint8_t* bounce = nullptr;           // pinned staging of large PAGEABLE inputs: DVB_H2D_THREADS x DVB_H2D_CHUNK bytes (first use)
    cudaEvent_t bounce_ev[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool busy = false;
};

struct dvb_ctx {
    int device = 0;
    int num_sm = DVB_NUM_SM_FALLBACK;
    size_t smem_optin = 0;
    std::mutex mu;
    dvb_lease leases[DVB_LEASES];
    std::atomic<int64_t> launches{0};
    std::atomic<int> timing{0};
    float last_ms = 0.f;
    std::mutex unicode_mu;
    void* unicode_dev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};   // CPython case tables (gate.cu)
    // last error text of any thread (each thread also keeps its own: dvb_tls_err)
    char err[512] = {0};
};

// Error text: one buffer per THREAD (the header promises "valid until the next call on ctx from this thread": gunicorn gthread
// workers share a ctx) plus a copy in the ctx for readers on another thread.  dvb_tls_err lives in api.cu.
char* dvb_tls_err(dvb_ctx* owner);                           // this thread's 512-byte buffer, from now on speaking for `owner`
static inline void dvb_set_err(dvb_ctx* ctx, const char* fmt, const char* a = "", const char* b = "") {
    if (!ctx) return;
    char* buf = dvb_tls_err(ctx);
    snprintf(buf, 512, fmt, a, b);
    std::lock_guard<std::mutex> g(ctx->mu);
    memcpy(ctx->err, buf, sizeof(ctx->err));
}

// dvb_set_stream is per (thread, ctx): a bench thread that brackets calls with its own events does not redirect the calls of
// a worker thread on the same ctx.  Defined in api.cu.
void* dvb_tls_stream(const dvb_ctx* ctx);

#define DVB_CUDA(ctx, expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { \
    dvb_set_err(ctx, "CUDA error %s at %s", cudaGetErrorString(e__), #expr); return DVB_E_CUDA; } } while (0)

// RAII lease of a stream + workspace arena
struct dvb_scope {
    dvb_ctx* ctx; dvb_lease* l = nullptr;
    explicit dvb_scope(dvb_ctx* c) : ctx(c) {
        for (;;) {
            {
                std::lock_guard<std::mutex> g(ctx->mu);
                for (int i = 0; i < DVB_LEASES; i++) if (!ctx->leases[i].busy) { l = &ctx->leases[i]; l->busy = true; break; }
            }
            if (l) break;
            std::this_thread::yield();
        }
        cudaSetDevice(ctx->device);
        // the previous call on this lease synchronised its stream before returning
        for (void* b : l->overflow) cudaFree(b);
        l->overflow.clear();
        l->arena.used = 0;
        call_bytes = 0;
        if (l->high_water > l->arena.cap) reserve(l->high_water);
    }
    size_t call_bytes = 0;
    ~dvb_scope() {
        // a call that returned early (error path) may still have kernels in flight on the lease's stream: the next call on this
        // lease resets the arena, so the stream is drained here (a no-op after a normal finish)
        cudaStreamSynchronize(l->stream);
        if (l->copy) cudaStreamSynchronize(l->copy);
        // the workspace follows the recent calls instead of the largest call ever: the high-water mark decays by 1/8 per call and
        // an arena more than 4x (and 256 MB) above it is given back
        l->high_water = std::max(call_bytes, l->high_water - l->high_water / 8);
        if (l->arena.base && l->arena.cap > (size_t(256) << 20) && l->arena.cap / 4 > l->high_water) {
            cudaFree(l->arena.base); l->arena.base = nullptr; l->arena.cap = 0;
        }
        std::lock_guard<std::mutex> g(ctx->mu); l->busy = false;
    }
    cudaStream_t stream() const { void* u = dvb_tls_stream(ctx); return u ? (cudaStream_t)u : l->stream; }
    // reserve total workspace up front (grows the arena if needed; invalidates earlier allocs)
    int32_t reserve(size_t bytes) {
        bytes = (bytes + 255) & ~size_t(255);
        if (bytes <= l->arena.cap) return DVB_OK;
        cudaStreamSynchronize(stream());
        if (l->arena.base) cudaFree(l->arena.base);
        l->arena.base = nullptr; l->arena.cap = 0;
        size_t want = bytes + (bytes >> 2);
        if (cudaMalloc(&l->arena.base, want) != cudaSuccess) {
            cudaGetLastError();
            cudaMemPool_t pool;                                  // give cached pool blocks (index builds) back and retry
            if (cudaDeviceGetDefaultMemPool(&pool, ctx->device) == cudaSuccess) cudaMemPoolTrimTo(pool, 0);
            if (cudaMalloc(&l->arena.base, bytes) != cudaSuccess) { cudaGetLastError(); dvb_set_err(ctx, "out of device memory%s%s"); return DVB_E_NOMEM; }
            want = bytes;
        }
        l->arena.cap = want;
        return DVB_OK;
    }
    // [p, p + bytes) lies in the arena block (not in a one-off overflow block): ranges between two such pointers are one allocation
    bool in_arena(const void* p, size_t bytes) const {
        const uint8_t* a = reinterpret_cast<const uint8_t*>(p); const uint8_t* base = reinterpret_cast<const uint8_t*>(l->arena.base);
        return base && a >= base && a + bytes <= base + l->arena.cap;
    }
    template <typename T> T* alloc(size_t n) {
        size_t bytes = (n * sizeof(T) + 255 + 64) & ~size_t(255);
        call_bytes += bytes;
        if (call_bytes > l->high_water) l->high_water = call_bytes;
        if (l->arena.used + bytes <= l->arena.cap) {
            T* p = reinterpret_cast<T*>(l->arena.base + l->arena.used);
            l->arena.used += bytes;
            return p;
        }
        void* b = nullptr;                                       // arena too small this time: one-off block,
        if (cudaMalloc(&b, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }   // the arena grows on the next call
        l->overflow.push_back(b);
        return reinterpret_cast<T*>(b);
    }
};

static inline size_t dvb_pad(size_t bytes) { return (bytes + 255 + 64) & ~size_t(255); }

// Large inputs in PAGEABLE host memory (what numpy hands the ctypes shim): cudaMemcpyAsync stages them through the
// driver's own bounce buffer with one copying thread (12 GB/s on the B200 boxes against 40 GB/s from pinned memory).
// Here DVB_H2D_THREADS host threads copy alternate 2 MB chunks into pinned buffers of the lease and queue the DMA of each
// chunk on the call's stream, so the host memcpy of one chunk overlaps the DMA of the others.  Returns after every chunk
// has left its bounce buffer (the stream still orders later kernels behind the copies).
#define DVB_SMALL_STAGE (size_t(256) << 10)
#define DVB_H2D_THREADS 8
#define DVB_H2D_CHUNK (size_t(2) << 20)
#define DVB_H2D_MIN (size_t(8) << 20)
static inline bool dvb_is_pageable(const void* p) {
    static const bool off = [] { const char* e = getenv("DVB_H2D_STAGED"); return e && e[0] == '0'; }();     // A/B switch
    if (off) return false;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return at.type == cudaMemoryTypeUnregistered;
}
static inline cudaError_t dvb_h2d_staged(dvb_lease* l, int device, void* dst, const void* src, size_t bytes, cudaStream_t st) {
    if (!l->bounce) {
        if (cudaHostAlloc((void**)&l->bounce, DVB_H2D_THREADS * DVB_H2D_CHUNK, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); l->bounce = nullptr; }
        for (int i = 0; l->bounce && i < DVB_H2D_THREADS; i++)
            if (cudaEventCreateWithFlags(&l->bounce_ev[i], cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); cudaFreeHost(l->bounce); l->bounce = nullptr; }
    }
    if (!l->bounce) return cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st);
    const size_t n_chunks = (bytes + DVB_H2D_CHUNK - 1) / DVB_H2D_CHUNK;
    cudaError_t errs[DVB_H2D_THREADS];
    auto work = [&](int t) {
        cudaError_t e = cudaSetDevice(device);
        uint8_t* buf = l->bounce + (size_t)t * DVB_H2D_CHUNK;
        for (size_t c = (size_t)t; c < n_chunks && e == cudaSuccess; c += DVB_H2D_THREADS) {
            const size_t o = c * DVB_H2D_CHUNK, nb = std::min(DVB_H2D_CHUNK, bytes - o);
            memcpy(buf, (const uint8_t*)src + o, nb);
            e = cudaMemcpyAsync((uint8_t*)dst + o, buf, nb, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) e = cudaEventRecord(l->bounce_ev[t], st);
            if (e == cudaSuccess) e = cudaEventSynchronize(l->bounce_ev[t]);       // the buffer is free again
        }
        errs[t] = e;
    };
    std::thread th[DVB_H2D_THREADS - 1];
    for (int t = 1; t < DVB_H2D_THREADS; t++) th[t - 1] = std::thread(work, t);
    work(0);
    for (int t = 1; t < DVB_H2D_THREADS; t++) th[t - 1].join();
    for (int t = 0; t < DVB_H2D_THREADS; t++) if (errs[t] != cudaSuccess) return errs[t];
    return cudaSuccess;
}

// Stage a caller array: HOST -> async copy into `dst` (device); DEVICE -> use in place.
template <typename T>
static inline const T* dvb_in(int mem, const T* src, T* dst, size_t n, cudaStream_t s, cudaError_t* err) {
    if (mem == DVB_MEM_DEVICE) return src;
    if (n) { cudaError_t e = cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyHostToDevice, s); if (e != cudaSuccess) *err = e; }
    return dst;
}

// Host/device staging of call arguments.  HOST: inputs are copied to fresh workspace, outputs are
// produced in workspace and copied back by finish().  DEVICE: caller pointers are used in place.
struct dvb_io {
    dvb_scope& sc; int mem; cudaStream_t st; cudaError_t err = cudaSuccess; bool oom = false;
    struct Out { void* host; void* dev; size_t bytes; };
    std::vector<Out> outs;
    // Small host arrays (a one-report call is a handful of arrays of a few hundred bytes) are gathered in the lease's pinned block
    // at the offsets their device copies have in the workspace and cross PCIe as ONE copy at check(); the outputs come back the
    // same way in finish().  A driver copy from pageable memory costs ~6 us whatever its size.  `coalesce` = false: every array
    // is copied where in() is called (entry points that launch between their in() calls).
    bool coalesce; uint8_t* stage = nullptr; uint8_t* first = nullptr; size_t span = 0;
    dvb_io(dvb_scope& s, int m, bool coalesce_small = true) : sc(s), mem(m), st(s.stream()), coalesce(coalesce_small && m == DVB_MEM_HOST) {}
    uint8_t* stage_block() {
        dvb_lease* l = sc.l;
        if (!l->small_h && cudaHostAlloc((void**)&l->small_h, 2 * DVB_SMALL_STAGE, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); l->small_h = nullptr; }
        return l->small_h;
    }
    template <typename T> const T* in(const T* p, size_t n) {
        if (mem == DVB_MEM_DEVICE || !p) return p;
        T* d = sc.alloc<T>(n ? n : 1);
        if (!d) { oom = true; return nullptr; }
        if (n) {
            const size_t bytes = n * sizeof(T);
            if (coalesce && bytes <= DVB_SMALL_STAGE && sc.in_arena(d, bytes) && (stage || (stage = stage_block()))) {
                if (!first) first = reinterpret_cast<uint8_t*>(d);
                const ptrdiff_t o = reinterpret_cast<uint8_t*>(d) - first;
                if (o >= 0 && (size_t)o + bytes <= DVB_SMALL_STAGE) {
                    memcpy(stage + o, p, bytes);
                    span = std::max(span, (size_t)o + bytes);
                    return d;
                }
            }
            cudaError_t e = (bytes >= DVB_H2D_MIN && dvb_is_pageable(p)) ? dvb_h2d_staged(sc.l, sc.ctx->device, d, p, bytes, st)
                                                                         : cudaMemcpyAsync(d, p, bytes, cudaMemcpyHostToDevice, st);
            if (e != cudaSuccess) err = e;
        }
        return d;
    }
    template <typename T> T* out(T* p, size_t n) {
        if (mem == DVB_MEM_DEVICE || !p) return p;
        T* d = sc.alloc<T>(n ? n : 1);
        if (!d) { oom = true; return nullptr; }
        outs.push_back({ p, d, n * sizeof(T) });
        return d;
    }
    template <typename T> T* tmp(size_t n) { T* d = sc.alloc<T>(n ? n : 1); if (!d) oom = true; return d; }
    // the gathered inputs go out; from here on in() copies directly (the pinned block is in flight until the stream is synchronised)
    void flush() {
        if (span) { const cudaError_t e = cudaMemcpyAsync(first, stage, span, cudaMemcpyHostToDevice, st); if (e != cudaSuccess) err = e; }
        span = 0; coalesce = false;
    }
    int32_t check(dvb_ctx* ctx) {
        flush();
        if (oom) { dvb_set_err(ctx, "out of device memory%s%s"); return DVB_E_NOMEM; }
        if (err != cudaSuccess) { dvb_set_err(ctx, "CUDA error %s while staging arguments%s", cudaGetErrorString(err)); return DVB_E_CUDA; }
        return DVB_OK;
    }
    int32_t finish(dvb_ctx* ctx) {
        cudaError_t e = cudaGetLastError();
        // outputs that lie within DVB_SMALL_STAGE bytes of each other in the workspace: one copy into the pinned block
        uint8_t *lo = nullptr, *hi = nullptr; bool arena_only = true;
        for (auto& o : outs) if (o.bytes) {
            uint8_t* a = reinterpret_cast<uint8_t*>(o.dev);
            arena_only = arena_only && sc.in_arena(a, o.bytes);
            if (!lo || a < lo) lo = a;
            if (!hi || a + o.bytes > hi) hi = a + o.bytes;
        }
        if (!stage && mem == DVB_MEM_HOST && outs.size() > 1) stage = stage_block();
        const bool one = stage && lo && arena_only && outs.size() > 1 && (size_t)(hi - lo) <= DVB_SMALL_STAGE;
        if (one) { if (e == cudaSuccess) e = cudaMemcpyAsync(stage + DVB_SMALL_STAGE, lo, (size_t)(hi - lo), cudaMemcpyDeviceToHost, st); }
        else for (auto& o : outs) if (o.bytes && e == cudaSuccess) e = cudaMemcpyAsync(o.host, o.dev, o.bytes, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { dvb_set_err(ctx, "CUDA error %s at finish%s", cudaGetErrorString(e)); return DVB_E_CUDA; }
        if (one) for (auto& o : outs) if (o.bytes) memcpy(o.host, stage + DVB_SMALL_STAGE + (reinterpret_cast<uint8_t*>(o.dev) - lo), o.bytes);
        return DVB_OK;
    }
};

// Host-side trace of one call (where the microseconds of a small call go): DVB_<NAME>_TRACE=1 prints the time since the call began
// at every at().  Costs one getenv per call when off.
struct dvb_trace {
    bool on; const char* tag; std::chrono::steady_clock::time_point t0;
    dvb_trace(const char* env, const char* tag_) : on([env] { const char* e = getenv(env); return e && e[0] == '1'; }()), tag(tag_), t0(std::chrono::steady_clock::now()) {}
    void at(const char* what) const {
        if (on) fprintf(stderr, "[%s] %-32s %8.0f us\n", tag, what, std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count());
    }
};

// chunk.cu: address of the static scratch of dvb_topk_static on `device` (which must be current)
cudaError_t dv_topk_static_setup(int device, unsigned char** out);
// launchers defined in chunk.cu, used by index.cu
void dv_launch_symbolize(dvb_ctx* ctx, cudaStream_t st, const uint8_t* d_utf8, const int64_t* d_off, int64_t n_texts, uint16_t* d_sym, int32_t* d_cnt);
int64_t dv_topk_splits(dvb_ctx* ctx, int64_t n, int32_t nq);
void dv_topk_device(dvb_ctx* ctx, cudaStream_t st, const int32_t* d_score, const int64_t* d_idx, int64_t n, int32_t nq, int64_t index_base, int32_t k,
                    int64_t n_split, int32_t* d_ps, int64_t* d_pi, int64_t* d_oidx, int32_t* d_oscore, int32_t* d_on, bool merge_prev = false);
