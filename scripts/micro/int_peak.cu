// Integer-pipe peaks of this GPU, measured: dependent-free streams of IMAD (FMA pipe: 32-bit multiply-add),
// IMAD.WIDE (32 x 32 -> 64), IADD3 and LOP3 (ALU pipe), 8 independent chains per thread, 148 x 4 CTAs of 1024 threads.
// Prints G thread-instructions / s per kind and the per-SM per-clock rate.  build: nvcc -O3 -arch=sm_100a int_peak.cu
#include <cstdio>
#include <cuda_runtime.h>
#define CH 8
template <int KIND>
__global__ void spin(unsigned* out, int iters, unsigned a, unsigned b) {
    unsigned x[CH]; unsigned long long w[CH];
#pragma unroll
    for (int c = 0; c < CH; c++) { x[c] = threadIdx.x + c; w[c] = x[c]; }
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int c = 0; c < CH; c++) {
            if (KIND == 0) x[c] = x[c] * a + b;                                   // IMAD
            if (KIND == 1) w[c] = (unsigned long long)(unsigned)w[c] * a + w[c];  // IMAD.WIDE
            if (KIND == 2) x[c] = x[c] + x[(c + 1) % CH] + a;                     // IADD3
            if (KIND == 3) x[c] = (x[c] & a) ^ (x[c] | b);                        // LOP3
        }
    }
    unsigned acc = 0;
#pragma unroll
    for (int c = 0; c < CH; c++) acc += x[c] + (unsigned)w[c];
    if (acc == 0x12345678u) out[0] = acc;
}
template <int KIND> double run(unsigned* out, int sms, const char* name, double per_iter) {
    const int iters = 4096, blocks = sms * 2, threads = 1024;
    spin<KIND><<<blocks, threads>>>(out, 64, 3u, 5u); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0); spin<KIND><<<blocks, threads>>>(out, iters, 3u, 5u); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double ops = (double)blocks * threads * iters * CH * per_iter;
    printf("{\"kind\": \"%s\", \"G_thread_ops_per_s\": %.1f, \"ms\": %.3f}\n", name, ops / ms / 1e6, ms);
    return ops / ms / 1e6;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    unsigned* out; cudaMalloc(&out, 64);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d}\n", p.name, p.multiProcessorCount, clk);
    run<0>(out, p.multiProcessorCount, "IMAD (32-bit multiply-add, FMA pipe)", 1);
    run<1>(out, p.multiProcessorCount, "IMAD.WIDE (32x32->64 multiply-add)", 1);
    run<2>(out, p.multiProcessorCount, "IADD3 (ALU pipe)", 1);
    run<3>(out, p.multiProcessorCount, "LOP3 (ALU pipe)", 1);
    return 0;
}
