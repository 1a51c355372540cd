// popc_bench.cu — measures the POPC32 issue rate of this GPU (the denominator of the matcher's roofline,
// "must be confirmed on sm_100 with a POPC microbenchmark").
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o popc_bench tools/popc_bench.cu && ./popc_bench
// Each thread runs 8 independent chains of x = popc(x ^ k) + x style updates; prints POPC32/s and per SM per clock.
#include <cstdio>
#include <cuda_runtime.h>

template <bool WITH_XOR_ADD>
__global__ void k_popc(unsigned* out, unsigned seed, int iters) {
    unsigned a[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u + blockIdx.x;
    unsigned acc = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (WITH_XOR_ADD) { acc += __popc(a[i] ^ acc); a[i] += 0x01010101u; }  // the matcher's XOR + POPC + ADD
            else a[i] = __popc(a[i]) + a[i];                                       // POPC + one add per chain
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += a[i];
    if (acc == 0x12345678u) out[0] = acc;
}

int main() {
    int dev = 0, sms = 0, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    unsigned* d; cudaMalloc(&d, 4);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 4096, threads = 256, blocks = sms * 8;
    for (int mode = 0; mode < 2; ++mode) {
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep) {
            cudaEventRecord(e0);
            if (mode == 0) k_popc<false><<<blocks, threads>>>(d, 12345u, iters); else k_popc<true><<<blocks, threads>>>(d, 12345u, iters);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        const double popc = (double)blocks * threads * iters * 8;
        const double per_s = popc / (best * 1e-3);
        printf("{\"mode\": \"%s\", \"popc32_per_s\": %.4g, \"popc32_per_clk_per_sm\": %.2f, \"sms\": %d, \"clock_mhz\": %.0f, \"ms\": %.3f}\n",
               mode == 0 ? "popc+add" : "xor+popc+add+add", per_s, per_s / sms / (khz * 1e3), sms, khz / 1e3, best);
    }
    return 0;
}
