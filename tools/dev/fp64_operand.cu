// Does a DFMA run slower when one operand comes from the constant bank / a uniform register /
// an immediate instead of a register?  (development aid)
#include <cstdio>
#include <cuda_runtime.h>
__constant__ double kc[8] = {1e-7, 2e-7, 3e-7, 4e-7, 5e-7, 6e-7, 7e-7, 8e-7};
template <int MODE>
__global__ void k(double *out, int iters, double rb, double rc) {
    double a[4];
    for (int i = 0; i < 4; ++i) a[i] = 0.5 + threadIdx.x * 1e-3 + i * 0.1;
    double rr[8];
    for (int i = 0; i < 8; ++i) rr[i] = rc * (i + 1) + threadIdx.x * 1e-12;      // per-thread registers
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (MODE == 0) a[i] = fma(a[i], rb, rr[r]);          // all registers
                if (MODE == 1) a[i] = fma(a[i], rb, kc[r]);          // addend from the constant bank
                if (MODE == 2) a[i] = fma(a[i], 0.999999, 0.5);      // immediates
                if (MODE == 3) a[i] = fma(a[i], rb, 1.0 / (3 + r));  // full 64-bit literals
                if (MODE == 4) a[i] = a[i] * rb;                     // DMUL
                if (MODE == 5) a[i] = a[i] + rr[r];                  // DADD
            }
    }
    double s = 0;
    for (int i = 0; i < 4; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char *name) {
    const int block = 128, grid = 148 * 6, iters = 2048;          // 24 warps per SM, ILP 4
    double *out; cudaMalloc(&out, sizeof(double) * grid * block);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<grid, block>>>(out, iters, 0.999999, 1e-7); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
        cudaEventRecord(e0); k<MODE><<<grid, block>>>(out, iters, 0.999999, 1e-7); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    printf("%-28s %.2f T op/s\n", name, (double)grid * block * 32.0 * iters / best * 1e-9);
    cudaFree(out);
}
int main() {
    run<0>("dfma reg,reg,reg"); run<1>("dfma reg,reg,const-bank"); run<2>("dfma reg,imm,imm");
    run<3>("dfma reg,reg,64-bit literal"); run<4>("dmul"); run<5>("dadd");
    return 0;
}
