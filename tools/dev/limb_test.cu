// Development check of the limb accumulator (fx128.cuh) against exact integer arithmetic on the host.
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include "../../melissa_b200/csrc/fx128.cuh"
using namespace melissa;
__global__ void k(const double *v, int n, unsigned long long *out) {
    __shared__ unsigned int acc[4 * 64];
    for (int i = threadIdx.x; i < 256; i += blockDim.x) acc[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) limb_add_shared(acc + (i & 63), 64, v[i]);
    __syncthreads();
    if (threadIdx.x < 64) {
        Fx128 f = limb_to_fx(acc[threadIdx.x], acc[64 + threadIdx.x], acc[128 + threadIdx.x], acc[192 + threadIdx.x]);
        out[2 * threadIdx.x] = f.lo;
        out[2 * threadIdx.x + 1] = (unsigned long long)f.hi;
    }
}
int main() {
    const int n = 64 * 3000;
    std::vector<double> v(n);
    srand(1);
    for (int i = 0; i < n; ++i) {
        double u = (rand() / (double)RAND_MAX - 0.5);
        v[i] = (i % 7 == 0) ? u * 1e-6 : (i % 5 == 0 ? u * 1e5 : u * 300.0);
    }
    double *dv; unsigned long long *dout;
    cudaMalloc(&dv, n * 8); cudaMalloc(&dout, 128 * 8);
    cudaMemcpy(dv, v.data(), n * 8, cudaMemcpyHostToDevice);
    k<<<1, 256>>>(dv, n, dout);
    unsigned long long out[128];
    cudaMemcpy(out, dout, sizeof out, cudaMemcpyDeviceToHost);
    double worst = 0;
    for (int j = 0; j < 64; ++j) {
        long double ref = 0;
        __int128 X = 0;
        for (int i = j; i < n; i += 64) { ref += v[i]; X += (__int128)floorl((long double)v[i] * 0x1p44L); }
        double got = fma((double)(out[2 * j] >> 11), 0x1p-53, (double)(long long)out[2 * j + 1]);
        __int128 G = ((__int128)(long long)out[2 * j + 1] << 44) + (__int128)(out[2 * j] >> 20);
        if (G != X) printf("slot %d: integer mismatch got-ref = %lld\n", j, (long long)(G - X));
        worst = fmax(worst, fabs(got - (double)ref));
    }
    printf("worst abs error vs long double sum: %.3e (3000 adds per slot, expected < 3000 * 2^-44 = %.3e)\n", worst, 3000 * 0x1p-44);
    return 0;
}
