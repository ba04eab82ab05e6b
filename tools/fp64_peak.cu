// Micro-benchmarks that calibrate the FP64 side of the roofline on the box:
// DFMA issue rate, and the cost of the special functions the sweep kernel uses.
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void k(double *out, int iters, double seed) {
    double a[8];
    for (int i = 0; i < 8; ++i) a[i] = seed + threadIdx.x * 1e-3 + i * 0.1;
    const double b = 0.999999, c = 1e-7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0) a[i] = fma(a[i], b, c);
            if (MODE == 1) a[i] = exp(-a[i] * a[i]) + 0.3;
            if (MODE == 2) a[i] = erfcx(a[i]) + 0.2;
            if (MODE == 3) a[i] = 1.0 / (a[i] + 1.5);
            if (MODE == 4) a[i] = erfc(a[i]) + 0.2;
            if (MODE == 5) a[i] = log(a[i] + 1.5);
        }
    }
    double s = 0;
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, int iters) {
    const int grid = 148 * 8, block = 256;
    double *out;
    cudaMalloc(&out, sizeof(double) * grid * block);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<grid, block>>>(out, iters, 0.5);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k<MODE><<<grid, block>>>(out, iters, 0.5);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    const double calls = (double)grid * block * 8.0 * iters;
    printf("{\"op\": \"%s\", \"ms\": %.4f, \"Gcalls_per_s\": %.2f}\n", name, best, calls / best * 1e-6);
    cudaFree(out);
}

int main() {
    run<0>("dfma", 4096);
    run<1>("exp(-x*x)+c", 512);
    run<2>("erfcx+c", 256);
    run<3>("1/(x+c)", 512);
    run<4>("erfc+c", 256);
    run<5>("log(x+c)", 512);
    return 0;
}
