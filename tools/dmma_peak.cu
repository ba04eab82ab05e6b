// Micro-benchmark: what one mma.sync.m8n8k4.f64 (DMMA) costs next to DFMA on this GPU, alone and mixed with integer work.
// Decides whether the sweep's per-pair products ([K x R] . [R x 32]) would gain from the tensor-core path.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/dmma_peak.cu -o /tmp/dmma_peak
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// MODE 5 / 6: DFMA + NI LOP3 per DFMA / LOP3 only.  MODE 0: 8 independent DMMA chains; 1: 8 independent DFMA chains; 2: DMMA + NI integer ops per DMMA; 3: DFMA + NI integer
// ops per DFMA; 4: integer ops only (NI per slot)
template <int MODE, int NI>
__global__ void k(double *out, int iters, double seed) {
    double c[8][2], f[8];
    unsigned z[8];
    for (int i = 0; i < 8; ++i) { c[i][0] = seed + i; c[i][1] = seed - i; f[i] = seed + threadIdx.x * 1e-3 + i; z[i] = threadIdx.x * 7 + i; }
    const double a = 0.999999 + threadIdx.x * 1e-9, b = 1e-7;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (MODE == 0 || MODE == 2) dmma(c[i][0], c[i][1], a, b);
            if (MODE == 1 || MODE == 3 || MODE == 5) f[i] = fma(f[i], a, b);
            if (MODE == 5 || MODE == 6) {                                  // full-rate integer work: one LOP3 per statement
#pragma unroll
                for (int j = 0; j < NI; ++j) z[(i + j) & 7] = (z[(i + j) & 7] & z[(i + j + 1) & 7]) ^ z[(i + j + 3) & 7];
            } else if (MODE >= 2) {
#pragma unroll
                for (int j = 0; j < NI; ++j) z[i] = z[i] * z[i] + 1u;                 // not an affine map: cannot be folded
            }
        }
    }
    double s = 0;
    for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1] + f[i] + (double)z[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE, int NI>
void run(const char *name, int iters, int warps_per_sm) {
    const int block = 32 * warps_per_sm, grid = 148;
    double *out;
    cudaMalloc(&out, sizeof(double) * grid * block);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE, NI><<<grid, block>>>(out, iters, 0.5);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k<MODE, NI><<<grid, block>>>(out, iters, 0.5);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    // warp instructions of the FP64 kind per second, and FMAs per second (a DMMA is 8 x 8 x 4 = 256 FMAs, a DFMA 32)
    const double winstr = (double)grid * warps_per_sm * 8.0 * iters;
    const double fmas = winstr * ((MODE == 0 || MODE == 2) ? 256.0 : ((MODE == 4 || MODE == 6) ? 0.0 : 32.0));
    printf("{\"op\": \"%s\", \"warps_per_sm\": %d, \"int_per_fp64\": %d, \"ms\": %.4f, \"G_fp64_winstr_per_s\": %.2f, \"T_fma_per_s\": %.3f, "
           "\"cycles_per_slot_per_smsp_at_1.93GHz\": %.2f}\n",
           name, warps_per_sm, NI, best, winstr / best * 1e-6, fmas / best * 1e-9,
           best * 1e-3 * 1.93e9 / ((double)warps_per_sm / 4.0 * 8.0 * iters));
    cudaFree(out);
}

int main() {
    run<0, 0>("dmma", 2048, 20);
    run<0, 0>("dmma", 2048, 8);
    run<0, 0>("dmma", 2048, 4);
    run<1, 0>("dfma", 4096, 20);
    run<1, 0>("dfma", 4096, 4);
    run<2, 2>("dmma+int", 2048, 20);
    run<2, 4>("dmma+int", 2048, 20);
    run<2, 8>("dmma+int", 2048, 20);
    run<2, 16>("dmma+int", 2048, 20);
    run<3, 1>("dfma+int", 4096, 20);
    run<3, 2>("dfma+int", 4096, 20);
    run<3, 4>("dfma+int", 4096, 20);
    run<4, 2>("int", 4096, 20);
    run<6, 1>("lop3", 4096, 20);
    run<6, 4>("lop3", 4096, 20);
    run<5, 1>("dfma+lop3", 4096, 20);
    run<5, 2>("dfma+lop3", 4096, 20);
    run<5, 3>("dfma+lop3", 4096, 20);
    run<5, 4>("dfma+lop3", 4096, 20);
    return 0;
}
