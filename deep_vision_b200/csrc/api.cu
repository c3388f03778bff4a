// api.cu — ctx lifecycle for libdvb200.so.  No CPU fallback: dvb_create fails without a CUDA device.
#include "common.cuh"

#include "../../include/dv_unicode_tables.h"

extern "C" int32_t dvb_version(void) { return 101; }
extern "C" const char* dvb_unicode_version(void) { return DV_UNICODE_VERSION; }

extern "C" int32_t dvb_create(int32_t device, dvb_ctx** out) {
    if (!out) return DVB_E_INVALID;
    *out = nullptr;
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) { cudaGetLastError(); return DVB_E_CUDA; }
    if (cudaSetDevice(device) != cudaSuccess) return DVB_E_CUDA;
    dvb_ctx* c = new dvb_ctx();
    c->device = device;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) {
        c->num_sm = prop.multiProcessorCount;
        c->smem_optin = prop.sharedMemPerBlockOptin;
    }
    {                                                    // index builds allocate stream-ordered from the default pool: keep freed blocks
        cudaMemPool_t pool; unsigned long long keep = ~0ULL;
        if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        cudaGetLastError();
    }
    for (int i = 0; i < DVB_LEASES; i++) {
        if (cudaStreamCreateWithFlags(&c->leases[i].stream, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreate(&c->leases[i].ev0) != cudaSuccess || cudaEventCreate(&c->leases[i].ev1) != cudaSuccess) {
            dvb_destroy(c); return DVB_E_CUDA;
        }
    }
    if (dv_topk_static_setup(device, nullptr) != cudaSuccess) { cudaGetLastError(); dvb_destroy(c); return DVB_E_CUDA; }   // set-up of dvb_topk_static
    *out = c;
    return DVB_OK;
}

extern "C" void dvb_destroy(dvb_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    for (int i = 0; i < DVB_LEASES; i++) {
        if (c->leases[i].stream) { cudaStreamSynchronize(c->leases[i].stream); cudaStreamDestroy(c->leases[i].stream); }
        if (c->leases[i].ev0) cudaEventDestroy(c->leases[i].ev0);
        if (c->leases[i].ev1) cudaEventDestroy(c->leases[i].ev1);
        if (c->leases[i].copy) { cudaStreamSynchronize(c->leases[i].copy); cudaStreamDestroy(c->leases[i].copy); }
        for (cudaEvent_t e : c->leases[i].pipe_ev) cudaEventDestroy(e);
        if (c->leases[i].flag_h) cudaFreeHost(c->leases[i].flag_h);
        if (c->leases[i].small_h) cudaFreeHost(c->leases[i].small_h);
        if (c->leases[i].bounce) { cudaFreeHost(c->leases[i].bounce); for (cudaEvent_t e : c->leases[i].bounce_ev) if (e) cudaEventDestroy(e); }
        if (c->leases[i].arena.base) cudaFree(c->leases[i].arena.base);
        for (void* b : c->leases[i].overflow) cudaFree(b);
    }
    for (int i = 0; i < 6; i++) if (c->unicode_dev[i]) cudaFree(c->unicode_dev[i]);
    delete c;
}

// ---- per-thread state (common.cuh)
struct DvbTls { char err[512]; dvb_ctx* err_ctx; const dvb_ctx* stream_ctx; void* stream; };
static thread_local DvbTls g_tls = { { 0 }, nullptr, nullptr, nullptr };
char* dvb_tls_err(dvb_ctx* owner) { g_tls.err_ctx = owner; return g_tls.err; }
void* dvb_tls_stream(const dvb_ctx* ctx) { return g_tls.stream_ctx == ctx ? g_tls.stream : nullptr; }

// this thread's last message for c if it has one, else the last message any thread left on c
extern "C" const char* dvb_last_error(dvb_ctx* c) {
    if (!c) return "null ctx";
    if (g_tls.err_ctx == c && g_tls.err[0]) return g_tls.err;
    return c->err;
}
extern "C" int64_t dvb_launch_count(dvb_ctx* c) { return c ? c->launches.load() : 0; }
extern "C" void dvb_set_timing(dvb_ctx* c, int32_t enabled) { if (c) c->timing.store(enabled ? 1 : 0); }
extern "C" float dvb_last_kernel_ms(dvb_ctx* c) { return c ? c->last_ms : 0.f; }
extern "C" void dvb_set_stream(dvb_ctx* c, void* stream) { if (c) { g_tls.stream_ctx = stream ? c : nullptr; g_tls.stream = stream; } }
