// dependent-load latency on this GPU: pointer chase over arrays of different footprints (cycles and ns)
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
__global__ void chase(const unsigned* a, int steps, unsigned start, unsigned long long* out) {
    unsigned i = start;
    unsigned long long t0, t1; long long c0 = clock64();
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    for (int s = 0; s < steps; s++) i = a[i];
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
    long long c1 = clock64();
    out[0] = t1 - t0; out[1] = (unsigned long long)(c1 - c0); out[2] = i;
}
__global__ void sync_cost(unsigned long long* out, int reps) {
    long long c0 = clock64();
    for (int r = 0; r < reps; r++) __syncthreads();
    long long c1 = clock64();
    if (threadIdx.x == 0) out[0] = (unsigned long long)(c1 - c0);
}
int main() {
    unsigned long long* out; cudaMallocManaged(&out, 64);
    size_t sizes[] = {1u << 14, 1u << 20, 1u << 23, 1u << 26, 1u << 29};   // elements (4 B): 64 KB .. 2 GB
    for (size_t n : sizes) {
        std::vector<unsigned> h(n);
        // random cyclic permutation with stride >= 64 elements apart (Sattolo)
        std::vector<unsigned> p(n); for (size_t i = 0; i < n; i++) p[i] = (unsigned)i;
        for (size_t i = n - 1; i > 0; i--) { size_t j = (size_t)(((unsigned long long)rand() << 16 ^ rand()) % i); std::swap(p[i], p[j]); }
        for (size_t i = 0; i < n; i++) h[p[i]] = p[(i + 1) % n];
        unsigned* d; cudaMalloc(&d, n * 4); cudaMemcpy(d, h.data(), n * 4, cudaMemcpyHostToDevice);
        for (int rep = 0; rep < 3; rep++) { chase<<<1, 1>>>(d, 2000, p[0], out); cudaDeviceSynchronize(); }
        printf("footprint %8.1f MB: %.1f ns / load, %.1f cycles / load\n", n * 4 / 1048576.0, out[0] / 2000.0, out[1] / 2000.0);
        cudaFree(d);
    }
    sync_cost<<<1, 512>>>(out, 1000); cudaDeviceSynchronize();
    printf("__syncthreads (512 threads): %.1f cycles\n", out[0] / 1000.0);
    return 0;
}
