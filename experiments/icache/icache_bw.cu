// Instruction-fetch microbenchmark (sm_100a): a straight-line loop body of N independent ALU/FMA-pipe instructions,
// 24 warps per SM (3 CTAs x 256 threads) like k_bind_claims, warps de-phased.  Prints issued warp instructions per clock per SM.
// usage: icache_bw   (prints one line per body size)
#include <cstdio>
#include <cuda_runtime.h>
#define OP2 asm volatile("add.u32 %0, %0, %4;\n\tmad.lo.u32 %1, %1, %5, %4;\n\tadd.u32 %2, %2, %5;\n\tmad.lo.u32 %3, %3, %4, %5;" : "+r"(a), "+r"(b), "+r"(c), "+r"(d) : "r"(x), "r"(y));
#define R4(X) X X X X
#define R16(X) R4(R4(X))
#define R64(X) R4(R16(X))
template <int N64>   // body = N64 * 64 * ... instructions
__global__ void __launch_bounds__(256, 3) k(unsigned* out, int iters, unsigned x, unsigned y, long long* cyc) {
    unsigned a = threadIdx.x, b = 1, c = 2, d = 3;
    // de-phase the warps
    const int w = threadIdx.x >> 5;
    for (int i = 0; i < w * 37; i++) asm volatile("add.u32 %0, %0, %1;" : "+r"(a) : "r"(x));
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < N64; r++) { R16(OP2) }     // 16 * 4 = 64 instructions
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = a + b + c + d;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int N64> void run(unsigned* out, long long* cyc, int sms) {
    const int grid = sms * 3;
    const int iters = 200000 / N64;
    k<N64><<<grid, 256>>>(out, 10, 1, 2, cyc);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<N64><<<grid, 256>>>(out, iters, 1, 2, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double inst = (double)iters * N64 * 64 * 8 * 3;           // warp instructions per SM
    long long h[8]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("body %5d instr (%6.1f KB): %.3f ms, IPC/SM by clock64 %.2f (by events at %d MHz: %.2f)\n", N64 * 64, N64 * 64 * 16 / 1024.0, ms,
           inst / (double)h[0], khz / 1000, inst / (ms * 1e-3 * khz * 1e3));
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    unsigned* out; long long* cyc; cudaMalloc(&out, sms * 3 * 256 * 4); cudaMalloc(&cyc, sms * 3 * 8);
    run<1>(out, cyc, sms); run<2>(out, cyc, sms); run<4>(out, cyc, sms); run<5>(out, cyc, sms); run<6>(out, cyc, sms); run<8>(out, cyc, sms);
    run<12>(out, cyc, sms); run<16>(out, cyc, sms); run<24>(out, cyc, sms); run<32>(out, cyc, sms); run<64>(out, cyc, sms); run<128>(out, cyc, sms);
    return 0;
}
