// DFMA throughput as a function of resident warps per SM and independent chains per thread:
// tells how much parallelism the sweep needs to fill the FP64 pipe (development aid).
#include <cstdio>
#include <cuda_runtime.h>
template <int ILP>
__global__ void k(double *out, int iters) {
    double a[ILP];
    for (int i = 0; i < ILP; ++i) a[i] = 0.5 + threadIdx.x * 1e-3 + i * 0.1;
    const double b = 0.999999, c = 1e-7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 16; ++r)
#pragma unroll
            for (int i = 0; i < ILP; ++i) a[i] = fma(a[i], b, c);
    }
    double s = 0;
    for (int i = 0; i < ILP; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int ILP>
void run(int warps_per_sm) {
    const int block = 128, per_sm = warps_per_sm / 4, grid = 148 * per_sm, iters = 2048;
    double *out; cudaMalloc(&out, sizeof(double) * grid * block);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<ILP><<<grid, block>>>(out, iters); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0); k<ILP><<<grid, block>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    const double fma = (double)grid * block * ILP * 16.0 * iters;
    printf("warps/SM %2d ILP %d : %.2f T DFMA/s  (%.1f%% of 18.1)\n", warps_per_sm, ILP, fma / best * 1e-9, fma / best * 1e-9 / 18.1 * 100);
    cudaFree(out);
}
int main() {
    for (int w : {4, 8, 16, 24, 32, 64}) { run<1>(w); run<2>(w); run<4>(w); }
    return 0;
}
