// Microbenchmark: FP64 tensor (mma.sync m8n8k4 f64, "DMMA") vs vector DFMA throughput and latency on sm_100a.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_vs_dfma dmma_vs_dfma.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int CHAINS>
__global__ void k_dmma(double *out, int iters, double a, double b) {
    double c[CHAINS][2];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) c[i][0] = c[i][1] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) dmma(c[i][0], c[i][1], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CHAINS>
__global__ void k_dfma(double *out, int iters, double a, double b) {
    double c[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) c[i] = threadIdx.x * 1e-3 + i;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < CHAINS; ++i) c[i] = fma(c[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <typename F>
float timeit(F f) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    f(); cudaDeviceSynchronize();
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}

int main() {
    double *out; cudaMalloc(&out, 148 * 32 * 1024 * sizeof(double));
    const int iters = 20000;
    for (int warps : {1, 2, 4, 8, 16, 32}) {
        int blocks = 148 * 2, threads = warps * 32 / 2; if (threads < 32) { threads = 32; blocks = 148; }
        float m1 = timeit([&] { k_dmma<1><<<blocks, threads>>>(out, iters, 1.0000001, 0.999999); });
        float m4 = timeit([&] { k_dmma<4><<<blocks, threads>>>(out, iters, 1.0000001, 0.999999); });
        float f1 = timeit([&] { k_dfma<1><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); });
        float f8 = timeit([&] { k_dfma<8><<<blocks, threads>>>(out, iters, 1.0000001, 1e-9); });
        double nw = (double)blocks * threads / 32;
        printf("warps/SM %2d: DMMA x1 %.2f TFLOP/s (%.1f clk/inst/warp @1.9GHz) x4 %.2f TFLOP/s | DFMA x1 %.2f x8 %.2f TFLOP/s (x1 %.1f clk)\n",
               (int)(nw / 148), nw * iters * 512.0 / m1 / 1e9, m1 * 1e-3 * 1.9e9 / iters, nw * iters * 4 * 512.0 / m4 / 1e9,
               nw * iters * 64.0 / f1 / 1e9, nw * iters * 8 * 64.0 / f8 / 1e9, f1 * 1e-3 * 1.9e9 / iters);
    }
    return 0;
}
