// Can a non-FP64 instruction issue in the second cycle a DFMA occupies its pipe?  DFMA chains mixed with
// independent integer (LOP3/IADD) or FP32 work, 24 warps per SM.  (development aid)
#include <cstdio>
#include <cuda_runtime.h>
template <int NI>
__global__ void k(double *out, int *iout, int iters, double rb, double rc, int seed) {
    double a[4];
    for (int i = 0; i < 4; ++i) a[i] = 0.5 + threadIdx.x * 1e-3 + i * 0.1;
    int z[4];
    for (int i = 0; i < 4; ++i) z[i] = seed + threadIdx.x * (i + 1);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                a[i] = fma(a[i], rb, rc);
#pragma unroll
                for (int j = 0; j < NI; ++j) z[(i + j) & 3] = (z[(i + j) & 3] ^ (z[(i + j + 1) & 3] + 0x9e3779b9)) + r;
            }
    }
    double s = 0; int zi = 0;
    for (int i = 0; i < 4; ++i) { s += a[i]; zi += z[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    iout[blockIdx.x * blockDim.x + threadIdx.x] = zi;
}
template <int NI>
void run() {
    const int block = 128, grid = 148 * 6, iters = 2048;
    double *out; int *iout;
    cudaMalloc(&out, sizeof(double) * grid * block); cudaMalloc(&iout, sizeof(int) * grid * block);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<NI><<<grid, block>>>(out, iout, iters, 0.999999, 1e-7, 3); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0); k<NI><<<grid, block>>>(out, iout, iters, 0.999999, 1e-7, 3); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    printf("%d integer statements per DFMA: %.2f T DFMA/s  (%.3f ms)\n", NI, (double)grid * block * 32.0 * iters / best * 1e-9, best);
    cudaFree(out); cudaFree(iout);
}
int main() { run<0>(); run<1>(); run<2>(); run<3>(); return 0; }
