// Micro-benchmarks for the B200 FP64 / predicate / shared-memory paths the canary kernels lean on.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o tools/ubench tools/ubench.cu
#include <cstdio>
#include <cuda_runtime.h>

#define N_ITER 4096

template <int MODE>
__global__ void lat_kernel(double* out, double a, double b, long long* cyc) {
    double x = a + threadIdx.x, y = b;
    double x2 = a * 3 + threadIdx.x, x3 = a * 5, x4 = a * 7;
    long long t0 = clock64();
#pragma unroll 16
    for (int i = 0; i < N_ITER; ++i) {
        if (MODE == 0) x = x + y;                                  // DADD chain
        if (MODE == 1) x = x * y;                                  // DMUL chain
        if (MODE == 2) x = (x < y) ? x + 1.0 : x;                  // DSETP -> DADD/FSEL chain
        if (MODE == 3) { x = x + y; x2 = x2 + y; x3 = x3 + y; x4 = x4 + y; }   // 4 independent DADD chains
        if (MODE == 4) x = x / y;                                  // DDIV chain
        if (MODE == 5) x = fmax(x, y) + 1e-9;                      // max + add
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = x + x2 + x3 + x4;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int MODE>
void run(const char* name, int blocks, int threads, int ops_per_iter) {
    double* out; long long* cyc;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    cudaMalloc(&cyc, 8);
    lat_kernel<MODE><<<blocks, threads>>>(out, 1.0, 1.0000001, cyc);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    lat_kernel<MODE><<<blocks, threads>>>(out, 1.0, 1.0000001, cyc);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    double warp_ops = (double)blocks * threads / 32 * N_ITER * ops_per_iter;
    printf("%-28s blocks=%4d thr=%4d  cycles/iter(warp0)=%7.2f  warp-ops/s=%.3e  (%.2f warp-ops/clk/SM @1.9GHz)\n", name, blocks,
           threads, (double)c / N_ITER, warp_ops / (ms * 1e-3), warp_ops / (ms * 1e-3) / 148 / 1.9e9);
    cudaFree(out); cudaFree(cyc);
}

int main() {
    run<0>("DADD chain, 1 warp", 1, 32, 1);
    run<1>("DMUL chain, 1 warp", 1, 32, 1);
    run<2>("DSETP+sel chain, 1 warp", 1, 32, 1);
    run<3>("4 indep DADD, 1 warp", 1, 32, 4);
    run<4>("DDIV chain, 1 warp", 1, 32, 1);
    run<5>("fmax+DADD chain, 1 warp", 1, 32, 1);
    run<0>("DADD chain, full chip", 148 * 4, 512, 1);
    run<3>("4 indep DADD, full chip", 148 * 4, 512, 4);
    run<4>("DDIV chain, full chip", 148 * 4, 512, 1);
    run<2>("DSETP+sel, full chip", 148 * 4, 512, 1);
    return 0;
}
