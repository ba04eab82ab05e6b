// Micro-benchmark: what the sweep's instruction classes cost a scheduler, alone and next to DFMA (development aid).
// Every statement is one volatile inline-PTX instruction, so the SASS has exactly the mix named in the output.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/issue_mix.cu -o tools/bin/issue_mix
#include <cstdio>
#include <cuda_runtime.h>

// NF DFMA + NL LOP3 + NM IMAD + NS LDS per slot, 8 slots per trip
template <int NF, int NL, int NM, int NS>
__global__ void k(double *out, int iters, double seed) {
    __shared__ double sm[1024];
    for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = i;
    __syncthreads();
    double f[8], acc = 0;
    unsigned z[8];
    for (int i = 0; i < 8; ++i) { f[i] = seed + threadIdx.x * 1e-3 + i; z[i] = threadIdx.x * 7 + i; }
    const double a = 0.999999 + threadIdx.x * 1e-9, b = 1e-7;
    const unsigned saddr = (unsigned)__cvta_generic_to_shared(sm) + (threadIdx.x & 31) * 8;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
#pragma unroll
            for (int j = 0; j < NF; ++j) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(f[(i + j) & 7]) : "d"(a), "d"(b));
#pragma unroll
            for (int j = 0; j < NL; ++j) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(z[(i + j) & 7]) : "r"(z[(i + j + 1) & 7]), "r"(z[(i + j + 3) & 7]));
#pragma unroll
            for (int j = 0; j < NM; ++j) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(z[(i + j) & 7]) : "r"(z[(i + j + 1) & 7]), "r"(z[(i + j + 3) & 7]));
#pragma unroll
            for (int j = 0; j < NS; ++j) { double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(saddr + ((i + j) & 7) * 256)); acc += 0 * v; }
        }
    }
    double s = acc;
    for (int i = 0; i < 8; ++i) s += f[i] + (double)z[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int NF, int NL, int NM, int NS>
void run(int warps_per_sm) {
    const int block = 32 * warps_per_sm, grid = 148, iters = 2048;
    double *out;
    cudaMalloc(&out, sizeof(double) * grid * block);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<NF, NL, NM, NS><<<grid, block>>>(out, iters, 0.5);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k<NF, NL, NM, NS><<<grid, block>>>(out, iters, 0.5);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    printf("{\"dfma\": %d, \"lop3\": %d, \"imad\": %d, \"lds\": %d, \"warps_per_sm\": %d, \"ms\": %.4f, \"cycles_per_slot_per_smsp_at_1.93GHz\": %.2f}\n",
           NF, NL, NM, NS, warps_per_sm, best, best * 1e-3 * 1.93e9 / ((double)warps_per_sm / 4.0 * 8.0 * iters));
    cudaFree(out);
}

int main() {
    run<1, 0, 0, 0>(20);
    run<0, 1, 0, 0>(20); run<0, 2, 0, 0>(20); run<0, 4, 0, 0>(20);
    run<0, 0, 1, 0>(20); run<0, 0, 2, 0>(20);
    run<0, 0, 0, 1>(20); run<0, 0, 0, 2>(20);
    run<1, 1, 0, 0>(20); run<1, 2, 0, 0>(20); run<1, 3, 0, 0>(20); run<1, 4, 0, 0>(20);
    run<1, 0, 1, 0>(20); run<1, 0, 2, 0>(20);
    run<1, 0, 0, 1>(20); run<1, 0, 0, 2>(20);
    run<1, 1, 0, 1>(20); run<2, 1, 0, 1>(20); run<2, 1, 1, 0>(20); run<2, 2, 0, 0>(20);
    run<4, 1, 1, 1>(20); run<4, 2, 1, 1>(20);
    run<1, 1, 0, 0>(8); run<1, 2, 0, 0>(8);
    return 0;
}
