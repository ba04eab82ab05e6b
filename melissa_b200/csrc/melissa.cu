// melissa_b200: CUDA (sm_100a) implementation of Melissa's mean-field VB loop
// behind the C ABI declared in include/melissa_b200.h.
//
// Reference algorithm: andreaskapou/Melissa R/melissa_vb.R:232-529
// (melissa_inner).  Nothing here is translated R: the loop is re-cut into a
// per-CpG sweep, a per-cell soft-max, a region-major contraction and a batched
// D x D solve, see DESIGN.md.
//
// Two data layouts live in this translation unit:
//   * region-sliced ELL (sliced.cuh): the default whenever the RBF centres are equally
//     spaced -- persistent sweep with table-driven basis / truncated-normal mean,
//     FP64 tensor-core contraction;
//   * flat cell-major CSR (the kernels below without a _v2 / _mma suffix): the first
//     correct path, kept for hand-placed centres and as an independent cross-check in
//     the parity tests.
//
// Per-pair record ("stats"), R = D + T doubles, T = D(D+1)/2:
//     [ g = H^T E[z]  (D) | packed upper triangle of H^T H  (T) ]
// Per-(region, cluster) record ("U"), same length:
//     [ m = E[w]      (D) | -1/2 * (m m^T + S) packed, off-diagonals doubled (T) ]
// so that   U[m,k] . stats[p]  =  m^T H^T E[z] - 1/2 tr(H^T H (m m^T + S))
// is the pair's E-step term (R/melissa_vb.R:325-326), and
//     acc[m,k] = sum_n r_nk * stats[p]
// is everything the M-step needs from the data (:368-373).
#include "melissa_b200.h"
#include "special.cuh"
#include "fx128.cuh"

#include <cuda.h>                  // types of the driver's virtual-memory API (multi.cuh); no libcuda at link time
#include <cuda_runtime.h>
#include <cub/cub.cuh>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <mutex>
#include <thread>
#include <chrono>
#include <string>
#include <vector>

namespace {

thread_local std::string g_err;
thread_local melissa_progress_fn g_progress = nullptr;      // melissa_set_progress: called at every bound of a run started
thread_local void *g_progress_user = nullptr;               // from this thread

int fail(int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(MELISSA_ERR_CUDA, "%s:%d: %s failed: %s", __FILE__, __LINE__, #call, \
                        cudaGetErrorString(e_));                                          \
    } while (0)

constexpr int MAX_D = 16;       // basis$M <= 15 (the reference's scripts go up to M = 11) (the reference's experiments use M = 3, 4, 7)
constexpr int MAX_MUS = 15;

struct BasisDev {
    int D;
    double gamma;
    double mus[MAX_MUS];
};

// Small per-fit scalar block in device memory, all length-K arrays.
struct Scalars {
    double *beta, *delta0, *delta, *elogpi, *Eww, *pi, *logdet, *lb_out;
    // cross-block sums, 128-bit fixed point so that they do not depend on block order (fx128.cuh):
    // [E_ww K | log det 1 | colSums(r) K | sum r log r 1] pairs, then the "blocks done" counter of resp_kernel
    unsigned long long *fx;
    __host__ __device__ unsigned long long *fx_Eww(int k) const { return fx + 2 * k; }
    __host__ __device__ unsigned long long *fx_logdet(int K) const { return fx + 2 * K; }
    __host__ __device__ unsigned long long *fx_colsum(int K, int k) const { return fx + 2 * K + 2 + 2 * k; }
    __host__ __device__ unsigned long long *fx_rlogr(int K) const { return fx + 4 * K + 2; }
    __host__ __device__ unsigned long long *fx_done(int K) const { return fx + 4 * K + 4; }
    static constexpr int fx_words(int K) { return 4 * K + 6; }
};

}  // namespace

namespace {
struct Sliced;
}

struct melissa_handle {
    int device = 0;
    int N = 0, M = 0, D = 0, T = 0, R = 0;
    int64_t P = 0, nnz = 0;
    BasisDev basis{};
    // packed observations (device)
    int64_t *pair_ptr = nullptr;
    int32_t *pair_cell = nullptr, *pair_region = nullptr;
    double *x = nullptr;
    uint8_t *y = nullptr;
    int64_t *reg_ptr = nullptr;
    int32_t *reg_pair = nullptr, *reg_cell = nullptr;
    double *stats = nullptr;       // [P][R]
    // VB state (device)
    int K = 0;
    int64_t N_total = 0;
    double alpha_0 = 0, beta_0 = 0, alpha_k = 0;
    double *U = nullptr;           // [M][K][R]
    double *Spk = nullptr;         // [M][K][T]
    double *r = nullptr, *log_r = nullptr;                       // [N][K]
    unsigned long long *log_rho_fx = nullptr;                    // [N][K][2] E-step sums, 128-bit fixed point (fx128.cuh)
    unsigned long long *lb_fx = nullptr;                         // [4] lb_pz_qz partial (fixed point), wrong-side count
    unsigned int *poison = nullptr;                              // a non-finite / out-of-range term was met
    double *red = nullptr;         // [M*K*R | K colsum | 1 sum r log r]
    int64_t red_len = 0;
    double *lbbuf = nullptr;       // [lb_pz_qz partial | #CpG with s*mu < -5]
    double *scal_mem = nullptr;
    Scalars sc{};
    cudaStream_t stream = nullptr;
    int64_t launches = 0;
    bool begun = false;
    int sm_count = 148;
    // optional per-kernel timing (CUDA events on h->stream), see melissa_profile_*
    bool prof_on = false;
    std::vector<cudaEvent_t> prof_ev;      // start/stop pairs
    std::vector<int> prof_id;
    Sliced *sl = nullptr;                  // v2 layout (sliced.cuh); null = v1 path
    int sweep_cfg[2] = {0, 0};             // sweep variants whose function attributes have been set
    int sweep_warps = 0;                   // warps per CTA of the sweep for this fit's K (0 = does not fit)
    bool sweep_direct = false;             // the sweep's E-step sums go straight to global memory (very large K)
};

namespace {

using namespace melissa;

__host__ __device__ inline int tri_index(int D, int i, int j) {   // i <= j, row-major upper triangle
    return i * D - (i * (i - 1)) / 2 + (j - i);
}

template <int D>
__device__ __forceinline__ void eval_basis(double x, const BasisDev &b, double (&h)[D]) {
    h[0] = 1.0;
#pragma unroll
    for (int j = 1; j < D; ++j) {
        const double d = x - b.mus[j - 1];
        h[j] = exp(-b.gamma * d * d);
    }
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum, result valid in thread 0.  blockDim.x multiple of 32, <= 1024.
__device__ inline double block_sum(double v, double *smem32) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) smem32[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (lane < (int)(blockDim.x >> 5)) ? smem32[lane] : 0.0;
        v = warp_sum(v);
    }
    return v;
}

// ---------------------------------------------------------------------------
// H^T H per pair, once at pack time (R/melissa_vb.R:254-255).
// ---------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(128) hh_kernel(const int64_t *__restrict__ pair_ptr,
                                                 const double *__restrict__ x, double *__restrict__ stats,
                                                 int64_t P, BasisDev basis) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
         p += (int64_t)gridDim.x * blockDim.x) {
        double hh[T];
#pragma unroll
        for (int t = 0; t < T; ++t) hh[t] = 0.0;
        const int64_t b = pair_ptr[p], e = pair_ptr[p + 1];
        for (int64_t i = b; i < e; ++i) {
            double h[D];
            eval_basis<D>(x[i], basis, h);
            int t = 0;
#pragma unroll
            for (int a = 0; a < D; ++a)
#pragma unroll
                for (int c = a; c < D; ++c) hh[t++] += h[a] * h[c];
        }
        double *rec = stats + p * R;
#pragma unroll
        for (int d = 0; d < D; ++d) rec[d] = 0.0;
#pragma unroll
        for (int t = 0; t < T; ++t) rec[D + t] = hh[t];
    }
}

// ---------------------------------------------------------------------------
// The per-CpG sweep (R/melissa_vb.R:418-435 refresh, fused with the NEXT
// iteration's E-step terms :322-327 and, when WITH_LB, lb_pz_qz :469-473).
// v1: one thread per pair.
// ---------------------------------------------------------------------------
template <int D, bool WITH_LB, bool UNIFORM_R>
__global__ void __launch_bounds__(128)
sweep_kernel(const int64_t *__restrict__ pair_ptr, const int32_t *__restrict__ pair_cell,
             const int32_t *__restrict__ pair_region, const double *__restrict__ x,
             const uint8_t *__restrict__ y, double *__restrict__ stats, const double *__restrict__ U,
             const double *__restrict__ r, unsigned long long *__restrict__ log_rho_fx,
             unsigned long long *__restrict__ lb_fx, unsigned int *__restrict__ poison, int64_t P, int K,
             BasisDev basis) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    __shared__ double red_smem[32];
    double lb_local = 0.0, wrong_side = 0.0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    const int64_t P_round = (P + 31) & ~int64_t(31);
    const double inv_K = 1.0 / (double)K;
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P_round; p += stride) {
        const bool valid = p < P;
        int n = -1, m = 0;
        double g[D];
#pragma unroll
        for (int d = 0; d < D; ++d) g[d] = 0.0;
        const double *Um = U;
        if (valid) {
            n = pair_cell[p];
            m = pair_region[p];
            Um = U + (int64_t)m * K * R;
            double wbar[D];
#pragma unroll
            for (int d = 0; d < D; ++d) wbar[d] = 0.0;
            for (int k = 0; k < K; ++k) {
                const double rk = UNIFORM_R ? 1.0 : r[(int64_t)n * K + k];
#pragma unroll
                for (int d = 0; d < D; ++d) wbar[d] += rk * Um[k * R + d];
            }
            if (UNIFORM_R) {
#pragma unroll
                for (int d = 0; d < D; ++d) wbar[d] *= inv_K;      // rowMeans, R/melissa_vb.R:291
            }
            const int64_t b = pair_ptr[p], e = pair_ptr[p + 1];
            for (int64_t i = b; i < e; ++i) {
                double h[D];
                eval_basis<D>(x[i], basis, h);
                double mu = wbar[0];
#pragma unroll
                for (int d = 1; d < D; ++d) mu += h[d] * wbar[d];
                const bool yi = y[i] != 0;
                const double t = yi ? mu : -mu;
                const double G = trunc_mean(t);
                const double Ez = yi ? G : -G;
                g[0] += Ez;
#pragma unroll
                for (int d = 1; d < D; ++d) g[d] += h[d] * Ez;
                if (WITH_LB) {
                    lb_local += 0.5 * mu * mu + loglik_term(mu, yi);
                    if (t < -5.0) wrong_side += 1.0;
                }
            }
            double *rec = stats + p * R;
#pragma unroll
            for (int d = 0; d < D; ++d) rec[d] = g[d];
        }
        // E-step terms of the next iteration, reduced per cell.
        const unsigned full = 0xffffffffu;
        const int n0 = __shfl_sync(full, n, 0);
        const bool uniform = __all_sync(full, n == n0 || !valid);
        double hh[T];
        if (valid) {
            const double *rec = stats + p * R + D;
#pragma unroll
            for (int t = 0; t < T; ++t) hh[t] = rec[t];
        } else {
#pragma unroll
            for (int t = 0; t < T; ++t) hh[t] = 0.0;
        }
        for (int k = 0; k < K; ++k) {
            double lin = 0.0, trc = 0.0;
            if (valid) {
                const double *u = Um + k * R;
#pragma unroll
                for (int d = 0; d < D; ++d) lin += u[d] * g[d];
#pragma unroll
                for (int t = 0; t < T; ++t) trc += u[D + t] * hh[t];
                if (WITH_LB) lb_local += (UNIFORM_R ? inv_K : r[(int64_t)n * K + k]) * trc;
            }
            double v = lin + trc;
            if (uniform) {
                v = warp_sum(v);
                if ((threadIdx.x & 31) == 0 && n0 >= 0) {
                    if (!(fabs(v) < FX_LIMIT)) atomicOr(poison, 1u);
                    fx_add(log_rho_fx + 2 * ((int64_t)n0 * K + k), fx_from_double(v));
                }
            } else if (valid) {
                if (!(fabs(v) < FX_LIMIT)) atomicOr(poison, 1u);
                fx_add(log_rho_fx + 2 * ((int64_t)n * K + k), fx_from_double(v));
            }
        }
    }
    if (WITH_LB) {
        // per-thread sums are fixed by the launch shape; the cross-warp part is order-free
        const double s0 = warp_sum(lb_local);
        if ((threadIdx.x & 31) == 0) {
            if (!(fabs(s0) < FX_LIMIT)) atomicOr(poison, 1u);
            fx_add(lb_fx, fx_from_double(s0));
        }
        const double s = block_sum(wrong_side, red_smem);
        if (threadIdx.x == 0 && s != 0.0) atomicAdd(lb_fx + 2, (unsigned long long)s);
    }
}

// ---------------------------------------------------------------------------
// Responsibilities (R/melissa_vb.R:330-331) + colSums(r) (:336) + sum r log r (:482).
// One thread per cell.  Clears log_rho for the next sweep and the E_ww / logdet
// accumulators for the solve that follows.
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(128)
resp_kernel(unsigned long long *__restrict__ log_rho_fx, const unsigned int *__restrict__ poison,
            const double *__restrict__ elogpi, double *__restrict__ r,
            double *__restrict__ log_r, double *__restrict__ colsum, double *__restrict__ rlogr,
            Scalars sc, int N, int K) {
    __shared__ double red_smem[32];
    __shared__ bool is_last;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (blockIdx.x == 0)
        for (int i = threadIdx.x; i < 2 * K + 2; i += blockDim.x) sc.fx[i] = 0ull;     // E_ww and log det sums of the solve
    double lse = 0.0, rl = 0.0;
    // a term the fixed-point sums could not hold (non-finite or >= 2^50) poisons the fit the way the
    // reference's own arithmetic would: everything downstream becomes NaN
    const double bias = *poison ? nan("") : 0.0;
    auto rho = [&](int k) {
        const unsigned long long *p = log_rho_fx + 2 * ((int64_t)n * K + k);
        return fx_to_double(p[0], p[1]) + elogpi[k] + bias;
    };
    if (n < N) {
        double mx = -INFINITY;
        for (int k = 0; k < K; ++k) mx = fmax(mx, rho(k));
        double s = 0.0;
        for (int k = 0; k < K; ++k) s += exp(rho(k) - mx);
        lse = log(s) + mx;                                     // R/utils.R:16-20
    }
    for (int k = 0; k < K; ++k) {
        double rk = 0.0;
        if (n < N) {
            const double lr = rho(k) - lse;
            rk = exp(lr);
            log_r[(int64_t)n * K + k] = lr;
            r[(int64_t)n * K + k] = rk;
            rl += rk * lr;
        }
        const double s = block_sum(rk, red_smem);
        if (threadIdx.x == 0) fx_add(sc.fx_colsum(K, k), fx_from_double(s));
    }
    if (n < N)
        for (int k = 0; k < 2 * K; ++k) log_rho_fx[2 * (int64_t)n * K + k] = 0ull;       // ready for the next sweep
    const double s = block_sum(rl, red_smem);
    if (threadIdx.x == 0) {
        fx_add(sc.fx_rlogr(K), fx_from_double(s));
        __threadfence();
        is_last = atomicAdd(sc.fx_done(K), 1ull) == (unsigned long long)gridDim.x - 1;
    }
    __syncthreads();
    if (is_last) {                       // the block that finishes last turns the sums into the doubles the all-reduce carries
        __threadfence();
        for (int k = threadIdx.x; k <= K; k += blockDim.x) {
            unsigned long long *p = k < K ? sc.fx_colsum(K, k) : sc.fx_rlogr(K);
            const double v = fx_to_double(__ldcg(p), __ldcg(p + 1));
            if (k < K) colsum[k] = v; else *rlogr = v;
            p[0] = p[1] = 0ull;
        }
        if (threadIdx.x == 0) *sc.fx_done(K) = 0ull;
    }
}

// ---------------------------------------------------------------------------
// Region-major contraction  acc[m,k,:] = sum_{pairs of region m} r[n,k] * stats[p,:]
// (the [K x N] x [N x M(D^2+D)] product of R/melissa_vb.R:368-373).
// v1: one warp per slice of a region's pair list; lane j owns record component(s) j, j+32, ...
// ---------------------------------------------------------------------------
template <int J, int KC>
__global__ void __launch_bounds__(128)
contract_kernel(const int64_t *__restrict__ reg_ptr, const int32_t *__restrict__ reg_pair,
                const int32_t *__restrict__ reg_cell, const double *__restrict__ stats,
                const double *__restrict__ r, double *__restrict__ acc_out, int K, int R) {
    const int m = blockIdx.x;
    const int lane = threadIdx.x & 31;
    const int slice = blockIdx.y * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int nslice = gridDim.y * (blockDim.x >> 5);
    const int64_t q0 = reg_ptr[m], L = reg_ptr[m + 1] - q0;
    const int64_t qb = q0 + (L * slice) / nslice, qe = q0 + (L * (slice + 1)) / nslice;
    if (qb >= qe) return;
    for (int k0 = 0; k0 < K; k0 += KC) {
        double acc[J][KC];
#pragma unroll
        for (int j = 0; j < J; ++j)
#pragma unroll
            for (int c = 0; c < KC; ++c) acc[j][c] = 0.0;
        for (int64_t base = qb; base < qe; base += 32) {
            const int64_t q = base + lane;
            const int32_t my_p = q < qe ? reg_pair[q] : 0;
            const int32_t my_n = q < qe ? reg_cell[q] : 0;
            const int cnt = (int)min((int64_t)32, qe - base);
#pragma unroll 4
            for (int it = 0; it < cnt; ++it) {
                const int64_t p = __shfl_sync(0xffffffffu, my_p, it);
                const int64_t n = __shfl_sync(0xffffffffu, my_n, it);
                double rk[KC];
#pragma unroll
                for (int c = 0; c < KC; ++c) rk[c] = (k0 + c < K) ? r[n * K + k0 + c] : 0.0;
#pragma unroll
                for (int j = 0; j < J; ++j) {
                    const int idx = lane + 32 * j;
                    const double s = idx < R ? stats[p * R + idx] : 0.0;
#pragma unroll
                    for (int c = 0; c < KC; ++c) acc[j][c] += rk[c] * s;
                }
            }
        }
#pragma unroll
        for (int j = 0; j < J; ++j) {
            const int idx = lane + 32 * j;
            if (idx < R) {
#pragma unroll
                for (int c = 0; c < KC; ++c)
                    if (k0 + c < K) atomicAdd(&acc_out[((int64_t)m * K + k0 + c) * R + idx], acc[j][c]);
            }
        }
    }
}

// ---------------------------------------------------------------------------
// q(w_mk): S = ((alpha/beta_k + 1e-7) I + sum r HH)^-1, m = S sum r H^T E[z]
// (R/melissa_vb.R:368-375), by Cholesky.  One thread per (region, cluster).
// ---------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(128)
solve_kernel(const double *__restrict__ acc, double *__restrict__ U, double *__restrict__ Spk,
             const double *__restrict__ beta, Scalars sc, int M, int K, double alpha_k) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    __shared__ double red_smem[32];
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    double eww = 0.0, logdetS = 0.0;
    if (m < M) {
        const double *rec = acc + ((int64_t)m * K + k) * R;
        const double ridge = alpha_k / beta[k] + 1e-7;
        double A[D][D];      // lower triangle used
        double b[D];
#pragma unroll
        for (int i = 0; i < D; ++i) b[i] = rec[i];
        {
            int t = 0;
#pragma unroll
            for (int i = 0; i < D; ++i)
#pragma unroll
                for (int j = i; j < D; ++j) {
                    A[j][i] = rec[D + t] + (i == j ? ridge : 0.0);
                    ++t;
                }
        }
        // Cholesky A = L L^T, in place (lower).
#pragma unroll
        for (int j = 0; j < D; ++j) {
            double d = A[j][j];
#pragma unroll
            for (int c = 0; c < j; ++c) d -= A[j][c] * A[j][c];
            const double ljj = sqrt(d);
            A[j][j] = ljj;
            logdetS -= 2.0 * log(ljj);
            const double inv = 1.0 / ljj;
#pragma unroll
            for (int i = j + 1; i < D; ++i) {
                double s = A[i][j];
#pragma unroll
                for (int c = 0; c < j; ++c) s -= A[i][c] * A[j][c];
                A[i][j] = s * inv;
            }
        }
        // Li = L^-1 (lower), in place.
#pragma unroll
        for (int j = 0; j < D; ++j) {
            A[j][j] = 1.0 / A[j][j];
#pragma unroll
            for (int i = j + 1; i < D; ++i) {
                double s = 0.0;
#pragma unroll
                for (int c = j; c < i; ++c) s -= A[i][c] * A[c][j];
                A[i][j] = s / A[i][i];
            }
        }
        // S = Li^T Li ; mvec = S b.
        double S[T];
        {
            int t = 0;
#pragma unroll
            for (int i = 0; i < D; ++i)
#pragma unroll
                for (int j = i; j < D; ++j) {
                    double s = 0.0;
#pragma unroll
                    for (int c = j; c < D; ++c) s += A[c][i] * A[c][j];
                    S[t++] = s;
                }
        }
        double mv[D];
#pragma unroll
        for (int i = 0; i < D; ++i) {
            double s = 0.0;
#pragma unroll
            for (int j = 0; j < D; ++j) s += S[i <= j ? tri_index(D, i, j) : tri_index(D, j, i)] * b[j];
            mv[i] = s;
        }
        double *u = U + ((int64_t)m * K + k) * R;
        double *sp = Spk + ((int64_t)m * K + k) * T;
        {
            int t = 0;
#pragma unroll
            for (int i = 0; i < D; ++i) {
                u[i] = mv[i];
                eww += mv[i] * mv[i] + S[tri_index(D, i, i)];
#pragma unroll
                for (int j = i; j < D; ++j) {
                    const double c = mv[i] * mv[j] + S[t];
                    u[D + t] = (i == j ? -0.5 : -1.0) * c;
                    sp[t] = S[t];
                    ++t;
                }
            }
        }
    }
    double s = block_sum(eww, red_smem);
    if (threadIdx.x == 0) fx_add(sc.fx_Eww(k), fx_from_double(s));
    s = block_sum(logdetS, red_smem);
    if (threadIdx.x == 0) fx_add(sc.fx_logdet(K), fx_from_double(s));
}

// Initial q(w): m = w0, S = solve(diag(50, D)) (R/melissa_vb.R:266-269, :279-280).
template <int D>
__global__ void init_params_kernel(const double *__restrict__ w0, double *__restrict__ U,
                                   double *__restrict__ Spk, int M, int K) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    const int m = blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (m >= M) return;
    double mv[D];
#pragma unroll
    for (int d = 0; d < D; ++d) mv[d] = w0[m + (int64_t)M * (d + D * k)];
    double *u = U + ((int64_t)m * K + k) * R;
    double *sp = Spk + ((int64_t)m * K + k) * T;
    int t = 0;
#pragma unroll
    for (int i = 0; i < D; ++i) {
        u[i] = mv[i];
#pragma unroll
        for (int j = i; j < D; ++j) {
            const double s = (i == j) ? 1.0 / 50.0 : 0.0;
            u[D + t] = (i == j ? -0.5 : -1.0) * (mv[i] * mv[j] + s);
            sp[t] = s;
            ++t;
        }
    }
}

// beta / delta / pi / E[log pi]  (R/melissa_vb.R:336-338, :386-392, :443).  One thread.
__global__ void scalars_kernel(Scalars sc, const double *__restrict__ colsum, int K, double alpha_k,
                               double beta_0, double N_total, int init) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double sum_delta = 0.0;
    for (int k = 0; k < K; ++k) {
        if (init) {
            sc.beta[k] = beta_0;
            sc.Eww[k] = 0.0;
            sc.delta[k] = sc.delta0[k];
            sc.pi[k] = 0.0;
        } else {
            sc.Eww[k] = fx_to_double(sc.fx_Eww(k)[0], sc.fx_Eww(k)[1]);
            double b = beta_0 + 0.5 * sc.Eww[k];
            if (b > 10.0 * alpha_k) b = 10.0 * alpha_k;
            sc.beta[k] = b;
            sc.delta[k] = sc.delta0[k] + colsum[k];
            sc.pi[k] = (sc.delta0[k] + colsum[k]) / (K * sc.delta0[k] + N_total);
        }
        sum_delta += sc.delta[k];
    }
    const double dsum = digamma_pos(sum_delta);
    for (int k = 0; k < K; ++k) sc.elogpi[k] = digamma_pos(sc.delta[k]) - dsum;
    if (!init) *sc.logdet = fx_to_double(sc.fx_logdet(K)[0], sc.fx_logdet(K)[1]);
}

// Lower bound assembly (R/melissa_vb.R:475-492).  One thread.
__global__ void lb_kernel(Scalars sc, const double *__restrict__ colsum, const double *__restrict__ rlogr,
                          const double *__restrict__ lbbuf, int K, int M, int D, double alpha_0,
                          double alpha_k, double beta_0) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const double LOG2PI = 1.83787706640934548356;
    const double MD = (double)M * D;
    const double dga = digamma_pos(alpha_k);
    double lb_p_w = 0, lb_p_c = 0, lb_p_pi = 0, lb_p_tau = 0, lb_q_pi = 0, lb_q_tau = 0;
    double sum_d0 = 0, sum_dk = 0, sum_lg0 = 0, sum_lgk = 0;
    for (int k = 0; k < K; ++k) {
        const double dg = dga - log(sc.beta[k]);
        const double E_alpha = alpha_k / sc.beta[k];
        lb_p_w += -0.5 * MD * LOG2PI + 0.5 * MD * dg - 0.5 * E_alpha * sc.Eww[k];
        lb_p_c += colsum[k] * sc.elogpi[k];
        lb_p_pi += (sc.delta0[k] - 1.0) * sc.elogpi[k];
        lb_q_pi += (sc.delta[k] - 1.0) * sc.elogpi[k];
        sum_d0 += sc.delta0[k];
        sum_dk += sc.delta[k];
        sum_lg0 += lgamma(sc.delta0[k]);
        sum_lgk += lgamma(sc.delta[k]);
        lb_p_tau += alpha_0 * log(beta_0) + (alpha_0 - 1.0) * dg - beta_0 * E_alpha - lgamma(alpha_0);
        lb_q_tau += -lgamma(alpha_k) + (alpha_k - 1.0) * dga + log(sc.beta[k]) - alpha_k;
    }
    lb_p_pi += lgamma(sum_d0) - sum_lg0;
    lb_q_pi += lgamma(sum_dk) - sum_lgk;
    const double lb_q_w = -0.5 * (*sc.logdet) - 0.5 * D * (1.0 + LOG2PI) * (double)M * K;
    const double lb_q_c = *rlogr;
    *sc.lb_out = lbbuf[0] + lb_p_c + lb_p_pi + lb_p_w + lb_p_tau - lb_q_c - lb_q_pi - lb_q_w - lb_q_tau;
}

// Imputation (R/utils.R:246-265): pred = round(sum_k r_nk Phi(h . W[m,,k]), 3).
// W and r_nk are in R's column-major layouts (dim c(M, D, K) and c(N, K)).
template <int D>
__global__ void impute_kernel(const int64_t *__restrict__ pair_ptr, const int32_t *__restrict__ pair_cell,
                              const int32_t *__restrict__ pair_region, const double *__restrict__ x,
                              const double *__restrict__ W, const double *__restrict__ r_nk,
                              double *__restrict__ pred, int64_t P, int N, int M, int K, BasisDev basis) {
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < P;
         p += (int64_t)gridDim.x * blockDim.x) {
        const int n = pair_cell[p], m = pair_region[p];
        for (int64_t i = pair_ptr[p]; i < pair_ptr[p + 1]; ++i) {
            double h[D];
            eval_basis<D>(x[i], basis, h);
            double s = 0.0;
            for (int k = 0; k < K; ++k) {
                const double *w = W + m + (int64_t)M * D * k;
                double z = w[0];
#pragma unroll
                for (int d = 1; d < D; ++d) z += h[d] * w[(int64_t)M * d];
                s += r_nk[n + (int64_t)N * k] * norm_cdf(z);
            }
            pred[i] = rint(s * 1000.0) / 1000.0;
        }
    }
}

#include "sliced.cuh"
#include "sweep3.cuh"
#include "contract_tma.cuh"

// Per-lane values back into the caller's pair order: out[pair][d] = src[slice * 32 + lane][d], or, with
// from_stats, the g = H^T E[z] part of the lane's record in the tile-ordered stats array.
__global__ void lanes_to_pairs_kernel(const double *__restrict__ src, const int32_t *__restrict__ lane_pair,
                                      double *__restrict__ out, int64_t n_lanes, int D, int R, int from_stats) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_lanes) return;
    const int32_t p = lane_pair[t];
    if (p < 0) return;
    const int64_t s = t >> 5;
    const int lane = (int)(t & 31);
    for (int d = 0; d < D; ++d)
        out[(int64_t)p * D + d] = from_stats ? src[s * ((int64_t)R * 32) + ((int64_t)(lane >> 2) * R + d) * 4 + (lane & 3)]
                                             : src[t * D + d];
}

struct StageTimer {       // MELISSA_PACK_TIMING=1 prints where melissa_pack spends its time
    bool on = std::getenv("MELISSA_PACK_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    bool sync = on && std::strcmp(std::getenv("MELISSA_PACK_TIMING"), "2") != 0;     // "2": host clock only, no device sync
    void lap(const char *what) {
        if (!on) return;
        if (sync) cudaDeviceSynchronize();
        const auto t1 = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[melissa_pack] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

struct HostLap {          // MELISSA_PACK_TIMING set: host-clock laps of the multi-GPU entry points (no device sync)
    bool on = std::getenv("MELISSA_PACK_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    void lap(const char *what) {
        if (!on) return;
        const auto t1 = std::chrono::steady_clock::now();
        std::fprintf(stderr, "[melissa_multi] %-28s %8.1f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

// library-wide options (melissa_set_option)
int g_opt_cell_block = 0;     // 0 = 256 cells
int g_opt_sliced = -1;        // -1 auto, 0 never, 1 require
int g_opt_sweep_chunk = 0;    // slices per chunk of the sweep, 0 = from the problem size
int g_opt_sweep_taper = 0;    // 1 = the three-step chunk list instead of the guided one (comparison)
int g_opt_sweep_group = 0;    // consecutive slices a warp claims at a time, 0 = from the chunk size
int g_opt_sweep_warps = 0;    // warps per CTA of the sweep, 0 = as many as fit
int g_opt_pack_host = 0;      // 1 = build the slice index with host threads (round 1's builder) instead of on the device

void prof_begin(melissa_handle *h, int id) {
    h->launches++;
    if (!h->prof_on) return;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    h->prof_ev.push_back(a);
    h->prof_ev.push_back(b);
    h->prof_id.push_back(id);
    cudaEventRecord(a, h->stream);
}
void prof_end(melissa_handle *h) {
    if (!h->prof_on) return;
    cudaEventRecord(h->prof_ev.back(), h->stream);
}

// ---------------------------------------------------------------------------
// launch helpers
// ---------------------------------------------------------------------------
int grid_for(int64_t work, int block, int sm_count, int per_sm) {
    const int64_t need = (work + block - 1) / block;
    return (int)std::max<int64_t>(1, std::min<int64_t>(need, (int64_t)sm_count * per_sm));
}

#define DISPATCH_D(D_, ...)                                                       \
    switch (D_) {                                                                 \
        case 1: { constexpr int DD = 1; __VA_ARGS__; } break;                     \
        case 2: { constexpr int DD = 2; __VA_ARGS__; } break;                     \
        case 3: { constexpr int DD = 3; __VA_ARGS__; } break;                     \
        case 4: { constexpr int DD = 4; __VA_ARGS__; } break;                     \
        case 5: { constexpr int DD = 5; __VA_ARGS__; } break;                     \
        case 6: { constexpr int DD = 6; __VA_ARGS__; } break;                     \
        case 7: { constexpr int DD = 7; __VA_ARGS__; } break;                     \
        case 8: { constexpr int DD = 8; __VA_ARGS__; } break;                     \
        case 9: { constexpr int DD = 9; __VA_ARGS__; } break;                     \
        case 10: { constexpr int DD = 10; __VA_ARGS__; } break;                   \
        case 11: { constexpr int DD = 11; __VA_ARGS__; } break;                   \
        case 12: { constexpr int DD = 12; __VA_ARGS__; } break;                   \
        case 13: { constexpr int DD = 13; __VA_ARGS__; } break;                   \
        case 14: { constexpr int DD = 14; __VA_ARGS__; } break;                   \
        case 15: { constexpr int DD = 15; __VA_ARGS__; } break;                   \
        case 16: { constexpr int DD = 16; __VA_ARGS__; } break;                   \
        default: return fail(MELISSA_ERR_ARG, "unsupported D = %d", D_);          \
    }

int launch_sweep(melissa_handle *h, bool with_lb, bool uniform_r) {
    const int block = 128;
    prof_begin(h, MELISSA_KERNEL_SWEEP);
    const int grid = grid_for(h->P, block, h->sm_count, 16);
#define SWEEP(LB_, UNI_)                                                                        \
    DISPATCH_D(h->D, sweep_kernel<DD, LB_, UNI_><<<grid, block, 0, h->stream>>>(                 \
                         h->pair_ptr, h->pair_cell, h->pair_region, h->x, h->y, h->stats, h->U, \
                         h->r, h->log_rho_fx, h->lb_fx, h->poison, h->P, h->K, h->basis))
    if (with_lb && uniform_r) { SWEEP(true, true) }
    else if (with_lb) { SWEEP(true, false) }
    else if (uniform_r) { SWEEP(false, true) }
    else { SWEEP(false, false) }
#undef SWEEP
    prof_end(h);
    CU(cudaGetLastError());
    return MELISSA_OK;
}

int launch_contract(melissa_handle *h) {
    const int J = (h->R + 31) / 32;
    int slices = (int)std::max<int64_t>(1, std::min<int64_t>(64, ((int64_t)h->sm_count * 32) / std::max(1, h->M)));
    // each slice should still have a few dozen pairs
    const int64_t avg = h->P / std::max(1, h->M);
    while (slices > 1 && avg / slices < 32) slices >>= 1;
    const dim3 grid(h->M, (slices + 3) / 4);
    const int block = 128;
    double *acc = h->red;
    prof_begin(h, MELISSA_KERNEL_CONTRACT);
#define CONTRACT(J_, KC_)                                                                     \
    contract_kernel<J_, KC_><<<grid, block, 0, h->stream>>>(h->reg_ptr, h->reg_pair, h->reg_cell, \
                                                          h->stats, h->r, acc, h->K, h->R)
    if (J == 1) {
        if (h->K <= 4) CONTRACT(1, 4);
        else if (h->K <= 8) CONTRACT(1, 8);
        else CONTRACT(1, 16);
    } else if (J == 2) {
        if (h->K <= 4) CONTRACT(2, 4);
        else CONTRACT(2, 8);
    } else if (J == 3) {
        CONTRACT(3, 4);
    } else if (J == 4) {
        CONTRACT(4, 4);
    } else if (J == 5) {
        CONTRACT(5, 4);
    } else {
        return fail(MELISSA_ERR_ARG, "record length %d not supported", h->R);
    }
#undef CONTRACT
    prof_end(h);
    CU(cudaGetLastError());
    return MELISSA_OK;
}

template <typename Tp>
int dev_alloc(Tp **p, int64_t n) {
    *p = nullptr;
    CU(cudaMalloc((void **)p, std::max<int64_t>(1, n) * sizeof(Tp)));
    return MELISSA_OK;
}


// ---------------------------------------------------------------------------
// v2 (sliced) host side: layout construction and launchers
// ---------------------------------------------------------------------------
struct SliceBuild {   // slices of a contiguous range of cell blocks, built by one host thread
    std::vector<int32_t> slice_region, slice_len, slice_cb, lane_cell, lane_pair, lane_meta;
    std::vector<int64_t> lane_src;
};

// blk_cells[cb * Tn + j] = global cell at position j of block cb (-1 = none).
void build_slices_range(int cb_lo, int cb_hi, int Tn, int M, const int32_t *blk_cells, const int64_t *cell_ptr,
                        const int64_t *pair_ptr, const int32_t *pair_region, SliceBuild &out) {
    std::vector<int32_t> cnt((size_t)M + 1), local_of;
    std::vector<int64_t> members, sorted;
    for (int cb = cb_lo; cb < cb_hi; ++cb) {
        // pairs of the block's cells, bucketed by region (counting sort), remembering the cell's position
        std::fill(cnt.begin(), cnt.end(), 0);
        int64_t total = 0;
        for (int j = 0; j < Tn; ++j) {
            const int c = blk_cells[(size_t)cb * Tn + j];
            if (c < 0) continue;
            for (int64_t p = cell_ptr[c]; p < cell_ptr[c + 1]; ++p) cnt[(size_t)pair_region[p] + 1]++;
            total += cell_ptr[c + 1] - cell_ptr[c];
        }
        for (int m = 0; m < M; ++m) cnt[(size_t)m + 1] += cnt[m];
        members.resize((size_t)total);
        local_of.resize((size_t)total);
        {
            std::vector<int32_t> cur(cnt.begin(), cnt.end() - 1);
            for (int j = 0; j < Tn; ++j) {
                const int c = blk_cells[(size_t)cb * Tn + j];
                if (c < 0) continue;
                for (int64_t p = cell_ptr[c]; p < cell_ptr[c + 1]; ++p) {
                    const int32_t q = cur[pair_region[p]]++;
                    members[(size_t)q] = p;
                    local_of[(size_t)q] = j;
                }
            }
        }
        std::vector<int32_t> idx;
        for (int m = 0; m < M; ++m) {
            const int b = cnt[m], e = cnt[(size_t)m + 1];
            if (b == e) continue;
            idx.resize((size_t)(e - b));
            for (int q = 0; q < e - b; ++q) idx[(size_t)q] = b + q;
            // longest first; ties keep block order
            std::stable_sort(idx.begin(), idx.end(), [&](int32_t u, int32_t v) {
                const int64_t pu = members[(size_t)u], pv = members[(size_t)v];
                return (pair_ptr[pu + 1] - pair_ptr[pu]) > (pair_ptr[pv + 1] - pair_ptr[pv]);
            });
            for (int q = 0; q < e - b; q += 32) {
                const int fill = std::min(32, e - b - q);
                const int64_t pl = members[(size_t)idx[(size_t)q]];
                out.slice_region.push_back(m);
                out.slice_cb.push_back(cb);
                out.slice_len.push_back((int32_t)(pair_ptr[pl + 1] - pair_ptr[pl]));
                for (int l = 0; l < 32; ++l) {
                    if (l < fill) {
                        const int32_t k = idx[(size_t)(q + l)];
                        const int64_t p = members[(size_t)k];
                        out.lane_cell.push_back(blk_cells[(size_t)cb * Tn + local_of[(size_t)k]]);
                        out.lane_pair.push_back((int32_t)p);
                        out.lane_meta.push_back((int32_t)(pair_ptr[p + 1] - pair_ptr[p]) | (local_of[(size_t)k] << 16));
                        out.lane_src.push_back(pair_ptr[p]);
                    } else {
                        out.lane_cell.push_back(-1);
                        out.lane_pair.push_back(-1);
                        out.lane_meta.push_back(0);
                        out.lane_src.push_back(0);
                    }
                }
            }
        }
    }
}

// log2 of the RBF table's points per unit of x (sliced.cuh, rbf_row3), or -1 if gamma is too large for a
// table of at most 2^13 + 3 entries.  The series there are cut where |w| = gamma s u_max stays below
// w_max, which keeps their truncation error under 3e-16.
int rbf_table_log2(const BasisDev &b) {
    if (b.D <= 1) return 8;
    const double u_max = std::max(std::fabs(-1.0 - b.mus[0]), std::fabs(1.0 - b.mus[0]));
    const double w_max = b.D <= 9 ? 0.02 : 0.055;
    for (int k = 8; k <= 12; ++k)
        if (b.gamma * std::ldexp(1.0, -k) * u_max <= w_max) return k;
    return -1;
}

bool rbf_equally_spaced(const BasisDev &b) {
    const int nb = b.D - 1;
    if (nb <= 1) return true;
    const double delta = b.mus[1] - b.mus[0];
    if (!(delta > 0.0)) return false;
    for (int j = 2; j < nb; ++j)
        if (std::fabs((b.mus[j] - b.mus[0]) - j * delta) > 1e-12) return false;
    return true;
}

void free_sliced(melissa_handle *h) {
    Sliced *sl = h->sl;
    if (!sl) return;
    cudaFree(sl->x_sell); cudaFree(sl->ybits); cudaFree(sl->slice_xbase); cudaFree(sl->slice_ybase);
    cudaFree(sl->slice_len); cudaFree(sl->slice_region); cudaFree(sl->slice_cb); cudaFree(sl->lane_cell);
    cudaFree(sl->lane_meta); cudaFree(sl->lane_pair); cudaFree(sl->blk_cells); cudaFree(sl->reg_slice_ptr);
    cudaFree(sl->reg_slice); cudaFree(sl->stats); cudaFree(sl->mills_tab); cudaFree(sl->lphi_tab); cudaFree(sl->rbf_tab);
    cudaFree(sl->chunk_ptr); cudaFree(sl->chunk_counter);
    delete sl;
    h->sl = nullptr;
}

template <typename Tp>
int upload(Tp **dst, const std::vector<Tp> &v) {
    int rc = dev_alloc(dst, (int64_t)v.size());
    if (rc) return rc;
    if (!v.empty()) CU(cudaMemcpy(*dst, v.data(), v.size() * sizeof(Tp), cudaMemcpyHostToDevice));
    return MELISSA_OK;
}

// Builds the sliced layout from the host index lists and the device-resident x / y.
#include "pack_gpu.cuh"
#include "multi.cuh"

// The slice index built by host threads from HOST index lists (round 1's builder; the default is the device
// builder in pack_gpu.cuh, this one stays as its cross-check and for the "pack_host" option).
int build_slice_index_host(melissa_handle *h, Sliced *sl, const int64_t *pair_ptr, const int32_t *pair_cell,
                           const int32_t *pair_region, cudaStream_t st, int64_t **d_src_out) {
    const int N = h->N, M = h->M;
    const int64_t P = h->P;
    StageTimer tm;
    const int Tn = sl->Tn;
    std::vector<int64_t> cell_ptr((size_t)N + 1, 0);
    for (int n = 0; n <= N; ++n)       // pair_cell is sorted: N + 1 binary searches instead of a pass over P
        cell_ptr[(size_t)n] = std::lower_bound(pair_cell, pair_cell + P, (int32_t)n) - pair_cell;
    // ---- blocks hold cells of similar coverage (sorted by CpG count), so that the pairs a slice groups
    //      have similar lengths and little of the lock-step loop is padding
    std::vector<int32_t> blk_cells((size_t)sl->NB * Tn, -1);
    {
        std::vector<int32_t> order((size_t)N);
        for (int n = 0; n < N; ++n) order[(size_t)n] = n;
        std::stable_sort(order.begin(), order.end(), [&](int32_t u, int32_t v) {
            return (pair_ptr[cell_ptr[(size_t)u + 1]] - pair_ptr[cell_ptr[u]]) >
                   (pair_ptr[cell_ptr[(size_t)v + 1]] - pair_ptr[cell_ptr[v]]);
        });
        for (int n = 0; n < N; ++n) blk_cells[(size_t)n] = order[(size_t)n];
    }
    tm.lap("cell order");
    // ---- slices, built by a few host threads over ranges of cell blocks
    int nthr = (int)std::min<int64_t>(std::max(1u, std::thread::hardware_concurrency()), 16);
    nthr = std::max(1, std::min(nthr, sl->NB));
    if (P < 200000) nthr = 1;
    std::vector<SliceBuild> parts((size_t)nthr);
    {
        std::vector<std::thread> th;
        for (int t = 0; t < nthr; ++t) {
            const int lo = (int)((int64_t)sl->NB * t / nthr), hi = (int)((int64_t)sl->NB * (t + 1) / nthr);
            th.emplace_back([&, t, lo, hi]() {
                build_slices_range(lo, hi, Tn, M, blk_cells.data(), cell_ptr.data(), pair_ptr, pair_region,
                                   parts[(size_t)t]);
            });
        }
        for (auto &t : th) t.join();
    }
    tm.lap("build slices (threads)");
    // the per-thread pieces are never concatenated on the host: they go to the device piece by piece
    std::vector<int64_t> soff((size_t)nthr + 1, 0);
    for (int t = 0; t < nthr; ++t) soff[(size_t)t + 1] = soff[(size_t)t] + (int64_t)parts[(size_t)t].slice_region.size();
    const int64_t S = soff[(size_t)nthr];
    sl->S = S;
    if (S >= (int64_t)std::numeric_limits<int32_t>::max()) return fail(MELISSA_ERR_ARG, "too many slices");
    // x / y offsets of every slice: local prefix sums per piece, then the piece's base
    std::vector<std::vector<int64_t>> xbase((size_t)nthr), ybase((size_t)nthr);
    std::vector<int64_t> xtot((size_t)nthr + 1, 0), ytot((size_t)nthr + 1, 0);
    {
        std::vector<std::thread> th;
        for (int t = 0; t < nthr; ++t)
            th.emplace_back([&, t]() {
                const auto &len = parts[(size_t)t].slice_len;
                xbase[(size_t)t].resize(len.size());
                ybase[(size_t)t].resize(len.size());
                int64_t xo = 0, yo = 0;
                for (size_t q = 0; q < len.size(); ++q) {
                    xbase[(size_t)t][q] = xo;
                    ybase[(size_t)t][q] = yo;
                    xo += (int64_t)len[q] * 32;
                    yo += (int64_t)((len[q] + 31) / 32) * 32;
                }
                xtot[(size_t)t + 1] = xo;
                ytot[(size_t)t + 1] = yo;
            });
        for (auto &t : th) t.join();
        for (int t = 0; t < nthr; ++t) { xtot[(size_t)t + 1] += xtot[(size_t)t]; ytot[(size_t)t + 1] += ytot[(size_t)t]; }
        th.clear();
        for (int t = 0; t < nthr; ++t)
            th.emplace_back([&, t]() {
                for (auto &v : xbase[(size_t)t]) v += xtot[(size_t)t];
                for (auto &v : ybase[(size_t)t]) v += ytot[(size_t)t];
            });
        for (auto &t : th) t.join();
    }
    sl->x_len = xtot[(size_t)nthr];
    sl->y_len = ytot[(size_t)nthr];
    tm.lap("offsets");
    // ---- region-major list of slices
    std::vector<int64_t> reg_slice_ptr((size_t)M + 1, 0);
    std::vector<int32_t> reg_slice((size_t)S);
    for (int t = 0; t < nthr; ++t)
        for (int32_t m : parts[(size_t)t].slice_region) reg_slice_ptr[(size_t)m + 1]++;
    for (int m = 0; m < M; ++m) reg_slice_ptr[(size_t)m + 1] += reg_slice_ptr[m];
    {
        std::vector<int64_t> cur(reg_slice_ptr.begin(), reg_slice_ptr.end() - 1);
        for (int t = 0; t < nthr; ++t) {
            const auto &reg = parts[(size_t)t].slice_region;
            for (size_t q = 0; q < reg.size(); ++q) reg_slice[(size_t)cur[reg[q]]++] = (int32_t)(soff[(size_t)t] + (int64_t)q);
        }
    }
    // ---- chunks of the sweep (make_chunks, pack_gpu.cuh)
    std::vector<int64_t> blk_first((size_t)sl->NB + 1, S);
    {
        int64_t s0 = 0;
        int last = -1;
        for (int t = 0; t < nthr; ++t) {
            const auto &cbv = parts[(size_t)t].slice_cb;
            for (size_t q = 0; q < cbv.size(); ++q)
                if (cbv[q] != last) {
                    for (int c = last + 1; c <= cbv[q]; ++c) blk_first[(size_t)c] = s0 + (int64_t)q;
                    last = cbv[q];
                }
            s0 += (int64_t)cbv.size();
        }
    }
    const std::vector<int32_t> chunk_ptr = make_chunks(blk_first, S, h->sm_count, g_opt_sweep_chunk, g_opt_sweep_taper != 0);
    sl->NC = (int)chunk_ptr.size() - 1;
    tm.lap("region index");
    // ---- upload, piece by piece
    int rc;
    int64_t *d_src = nullptr;
    if ((rc = dev_alloc(&sl->slice_xbase, S)) || (rc = dev_alloc(&sl->slice_ybase, S)) ||
        (rc = dev_alloc(&sl->slice_len, S)) || (rc = dev_alloc(&sl->slice_region, S)) ||
        (rc = dev_alloc(&sl->slice_cb, S)) || (rc = dev_alloc(&sl->lane_cell, S * 32)) ||
        (rc = dev_alloc(&sl->lane_meta, S * 32)) || (rc = dev_alloc(&sl->lane_pair, S * 32)) ||
        (rc = dev_alloc(&d_src, S * 32)) || (rc = upload(&sl->chunk_ptr, chunk_ptr)) ||
        (rc = upload(&sl->blk_cells, blk_cells)) || (rc = upload(&sl->reg_slice_ptr, reg_slice_ptr)) ||
        (rc = upload(&sl->reg_slice, reg_slice))) {
        cudaFree(d_src);
        return rc;
    }
    for (int t = 0; t < nthr; ++t) {
        const SliceBuild &pt = parts[(size_t)t];
        const int64_t o = soff[(size_t)t];
        const size_t n = pt.slice_region.size();
        if (!n) continue;
#define PUT(dst, vec, mul, Tp)                                                                              \
        if (cudaMemcpyAsync(dst + o * mul, vec.data(), n * mul * sizeof(Tp), cudaMemcpyHostToDevice, st) != \
            cudaSuccess) {                                                                                    \
            cudaFree(d_src);                                                                                  \
            return fail(MELISSA_ERR_CUDA, "upload of the slice index failed: %s",                             \
                        cudaGetErrorString(cudaGetLastError()));                                              \
        }
        PUT(sl->slice_xbase, xbase[(size_t)t], 1, int64_t)
        PUT(sl->slice_ybase, ybase[(size_t)t], 1, int64_t)
        PUT(sl->slice_len, pt.slice_len, 1, int32_t)
        PUT(sl->slice_region, pt.slice_region, 1, int32_t)
        PUT(sl->slice_cb, pt.slice_cb, 1, int32_t)
        PUT(sl->lane_cell, pt.lane_cell, 32, int32_t)
        PUT(sl->lane_meta, pt.lane_meta, 32, int32_t)
        PUT(sl->lane_pair, pt.lane_pair, 32, int32_t)
        PUT(d_src, pt.lane_src, 32, int64_t)
#undef PUT
    }
    if (cudaStreamSynchronize(st) != cudaSuccess) {            // the host vectors go out of scope
        cudaFree(d_src);
        return fail(MELISSA_ERR_CUDA, "upload of the slice index failed: %s", cudaGetErrorString(cudaGetLastError()));
    }
    *d_src_out = d_src;
    tm.lap("upload index arrays");
    return MELISSA_OK;
}

// Tables, the re-laid observations and the per-pair H^T H, once the slice index exists (either builder).
int finish_sliced(melissa_handle *h, Sliced *sl, int64_t *d_src, cudaStream_t st, cudaEvent_t xy_arrived) {
    const int64_t S = sl->S;
    StageTimer tm;
    int rc;
    // ---- RBF table (host exp) and the Mills table (device erfcx)
    const BasisDev &b = h->basis;
    const int nb = b.D - 1;
    RbfTab &rt = sl->rt;
    std::vector<double2> rbf_tab;
    if (nb >= 1) {
        const double delta = nb >= 2 ? (b.mus[1] - b.mus[0]) : 1.0;
        const int k = rbf_table_log2(b);               // grid spacing s = 2^-k (pack_impl has checked k >= 0)
        rt.s = std::ldexp(1.0, -k);
        rt.inv_s = std::ldexp(1.0, k);
        rt.neg_s = -rt.s;
        rt.mu1 = b.mus[0];
        rt.neg_gamma = -b.gamma;
        rt.cA = -2.0 * b.gamma;                        // -2 gamma u0 = cA x0 + cB,  u0 = x0 - mu_1
        rt.cB = 2.0 * b.gamma * rt.mu1;
        rt.g2d = 2.0 * b.gamma * delta;
        rt.q = std::exp(-b.gamma * delta * delta);
        rt.q2 = rt.q * rt.q;
        const int j_hi = (1 << k) + 1;                 // x in [-1, 1] -> j in [-2^k, 2^k], one guard entry each side
        rt.idx0 = j_hi;
        rt.n_entries = 2 * j_hi + 1;
        rt.magic = (6755399441055744.0 + (double)j_hi) * rt.s;     // (1.5 * 2^52 + idx0) s, exact
        rbf_tab.resize((size_t)rt.n_entries);
        for (int j = -j_hi; j <= j_hi; ++j) {
            const double u0 = (double)j * rt.s - rt.mu1;
            rbf_tab[(size_t)(j + j_hi)] = make_double2(std::exp(-b.gamma * u0 * u0), rt.q * std::exp(rt.g2d * u0));
        }
    } else {
        rt = RbfTab{};
        rt.n_entries = 1;
        rbf_tab.assign(1, make_double2(1.0, 1.0));
    }
    if ((rc = upload(&sl->rbf_tab, rbf_tab))) { cudaFree(d_src); return rc; }
    // slack: the sweep's x ring reads up to NST groups past the end of a slice
    if ((rc = dev_alloc(&sl->x_sell, sl->x_len + 2 * SWEEP3_NST * SWEEP3_XU * 32)) || (rc = dev_alloc(&sl->ybits, sl->y_len)) ||
        (rc = dev_alloc(&sl->stats, S * h->R * 32)) || (rc = dev_alloc(&sl->mills_tab, MILLS_N)) || (rc = dev_alloc(&sl->lphi_tab, MILLS_N)) ||
        (rc = dev_alloc(&sl->chunk_counter, 1))) {
        cudaFree(d_src);
        return rc;
    }
    cudaMemsetAsync(sl->x_sell + sl->x_len, 0, (size_t)2 * SWEEP3_NST * SWEEP3_XU * 32 * sizeof(double), st);
    CU(cudaStreamWaitEvent(st, xy_arrived, 0));                 // the flat x / y must be on the device from here on
    const int grid = (int)((S * 32 + 255) / 256);
    sell_fill_kernel<<<(int)std::min<int64_t>((S + 3) / 4, (int64_t)h->sm_count * 32), 128, 0, st>>>(
        h->x, h->y, d_src, sl->lane_meta, sl->slice_xbase, sl->slice_ybase, sl->slice_len, sl->x_sell, sl->ybits, S);
    {
        auto run = [&]() -> int {
            DISPATCH_D(h->D, hh_sliced_kernel<DD><<<grid, 256, 0, st>>>(sl->x_sell, sl->slice_xbase, sl->lane_meta, sl->stats,
                                                                S, h->basis, sl->rt, sl->rbf_tab));
            return MELISSA_OK;
        };
        if ((rc = run())) { cudaFree(d_src); return rc; }
    }
    mills_table_kernel<<<(MILLS_N + 255) / 256, 256, 0, st>>>(sl->mills_tab);
    lphi_table_kernel<<<(MILLS_N + 255) / 256, 256, 0, st>>>(sl->lphi_tab);
    h->launches += 4;
    const cudaError_t e = cudaStreamSynchronize(st);
    cudaFree(d_src);
    if (e != cudaSuccess) return fail(MELISSA_ERR_CUDA, "sliced layout build failed: %s", cudaGetErrorString(e));
    tm.lap("fill + HtH kernels");
    sl->on = true;
    return MELISSA_OK;
}

// Shared memory of one sweep CTA with nw warps: the two tables, the block's fixed-point sums,
// one U[m] staging area per warp, lb sums and control words.
size_t sweep3_smem(const melissa_handle *h, int nw, bool direct = false) {
    const Sliced *sl = h->sl;
    const size_t UR = ((size_t)h->K * h->R + 1) & ~(size_t)1;
    return (size_t)sl->rt.n_entries * sizeof(double2) + ((size_t)MILLS_N + 1) * sizeof(double) +
           (direct ? 0 : 4 * (size_t)sl->Tn * h->K * sizeof(unsigned int)) + (size_t)nw * UR * sizeof(double) +
           (size_t)nw * SWEEP3_NST * SWEEP3_XU * 32 * sizeof(double) + 4 * 8 + 16;
}

// Most warps (of SWEEP3_THREADS / 32) whose staging areas still fit next to the rest; 0 = not even four.
// When the block sums (Tn x K x 16 bytes) leave no room for at least eight warps the sweep runs without them
// ("direct": its E-step terms go to global memory with 64-bit atomics -- slower, but any K fits).
int sweep3_warps(const melissa_handle *h, bool *direct) {
    int nw0 = SWEEP3_THREADS / 32;
    if (g_opt_sweep_warps > 0) nw0 = std::min(nw0, g_opt_sweep_warps);
    *direct = false;
    for (int nw = nw0; nw >= std::min(8, nw0); nw -= 2)
        if (sweep3_smem(h, nw) <= 227u * 1024) return nw;
    *direct = true;
    for (int nw = nw0; nw >= 2; nw -= 2)
        if (sweep3_smem(h, nw, true) <= 227u * 1024) return nw;
    return 0;
}

__global__ void fill_kernel(double *p, int64_t n, double v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// ---- development check ("debug_estep" option): the E-step sums of the last sweep recomputed pair by pair from
//      the stored records and compared, accumulator by accumulator, with what the sweep left in log_rho_fx
template <int D>
__global__ void estep_check_kernel(const double *__restrict__ stats, const int32_t *__restrict__ slice_region,
                                   const int32_t *__restrict__ lane_cell, const double *__restrict__ U,
                                   unsigned long long *__restrict__ dbg, int64_t n_lanes, int K) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_lanes) return;
    const int cell = lane_cell[t];
    if (cell < 0) return;
    const int64_t s = t >> 5;
    const int lane = (int)(t & 31), m = slice_region[s];
    const double *rec = stats + s * (R * 32) + (lane >> 2) * (R * 4) + (lane & 3);
    for (int k = 0; k < K; ++k) {
        const double *u = U + ((int64_t)m * K + k) * R;
        double t0 = 0.0, t1 = 0.0, t2 = 0.0, lin = 0.0;       // the sweep's order of operations
        for (int q = 0; q < T; q += 3) {
            t0 = fma(u[D + q], rec[(D + q) * 4], t0);
            if (q + 1 < T) t1 = fma(u[D + q + 1], rec[(D + q + 1) * 4], t1);
            if (q + 2 < T) t2 = fma(u[D + q + 2], rec[(D + q + 2) * 4], t2);
        }
        for (int d = 0; d < D; ++d) lin = fma(u[d], rec[d * 4], lin);
        const double v = lin + ((t0 + t1) + t2);
        // same 2^-44 image as the sweep's limbs
        const double f = floor(v * 0x1p44);
        const long long hi = (long long)floor(f * 0x1p-44);
        Fx128 fx;
        fx.hi = hi;
        fx.lo = (unsigned long long)(f - (double)hi * 0x1p44) << 20;
        fx_add(dbg + 2 * ((int64_t)cell * K + k), fx);
    }
}
// the same check for g = H^T E[z]: recomputed per lane with exp / erfcx from x, r and U
template <int D>
__global__ void g_check_kernel(const double *__restrict__ x_sell, const uint32_t *__restrict__ ybits,
                               const int64_t *__restrict__ slice_xbase, const int64_t *__restrict__ slice_ybase,
                               const int32_t *__restrict__ slice_len, const int32_t *__restrict__ lane_meta,
                               const double *__restrict__ stats, const int32_t *__restrict__ slice_region,
                               const int32_t *__restrict__ lane_cell, const double *__restrict__ U, const double *__restrict__ r,
                               unsigned long long *__restrict__ out, int64_t n_lanes, int K, BasisDev basis) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_lanes) return;
    const int cell = lane_cell[t];
    if (cell < 0) return;
    const int64_t s = t >> 5;
    const int lane = (int)(t & 31), m = slice_region[s];
    const double *rec = stats + s * (R * 32) + (lane >> 2) * (R * 4) + (lane & 3);
    WBar<D> wb;
    for (int d = 0; d < D; ++d) wb.w[d] = 0.0;
    for (int k = 0; k < K; ++k)
        for (int d = 0; d < D; ++d) wb.w[d] = fma(r[(int64_t)cell * K + k], U[((int64_t)m * K + k) * R + d], wb.w[d]);
    const SliceSums<D> o = sweep_slice_slow<D, false>(x_sell + slice_xbase[s] + lane, ybits + slice_ybase[s] + lane, slice_len[s],
                                                      lane_meta[t] & 0xffff, wb, &basis);
    for (int d = 0; d < D; ++d) {
        const double got = rec[d * 4];
        if (!(fabs(got - o.g[d]) <= 1e-9 * (1.0 + fabs(o.g[d])))) {
            const unsigned long long c = atomicAdd(out, 1ull);
            if (c < 4) {
                out[1 + 5 * c] = (unsigned long long)t;
                out[2 + 5 * c] = (unsigned long long)d;
                out[3 + 5 * c] = (unsigned long long)__double_as_longlong(got);
                out[4 + 5 * c] = (unsigned long long)__double_as_longlong(o.g[d]);
                out[5 + 5 * c] = (unsigned long long)slice_len[s];
            }
        }
    }
}
__global__ void estep_compare_kernel(const unsigned long long *a, const unsigned long long *b, int64_t n, unsigned long long *out) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (a[2 * i] != b[2 * i] || a[2 * i + 1] != b[2 * i + 1]) {
        const unsigned long long c = atomicAdd(out, 1ull);
        if (c < 4) {
            out[1 + 5 * c] = (unsigned long long)i;
            out[2 + 5 * c] = a[2 * i]; out[3 + 5 * c] = a[2 * i + 1];
            out[4 + 5 * c] = b[2 * i]; out[5 + 5 * c] = b[2 * i + 1];
        }
    }
}
int g_opt_debug_estep = 0;
int debug_estep_check(melissa_handle *h) {
    Sliced *sl = h->sl;
    const int64_t NK = (int64_t)h->N * h->K;
    unsigned long long *dbg = nullptr, *out = nullptr;
    int rc;
    if ((rc = dev_alloc(&dbg, 2 * NK)) || (rc = dev_alloc(&out, 32))) return rc;
    cudaMemsetAsync(dbg, 0, 2 * NK * 8, h->stream);
    cudaMemsetAsync(out, 0, 32 * 8, h->stream);
    auto run = [&]() -> int {
        DISPATCH_D(h->D, estep_check_kernel<DD><<<(unsigned)((sl->S * 32 + 255) / 256), 256, 0, h->stream>>>(
                             sl->stats, sl->slice_region, sl->lane_cell, h->U, dbg, sl->S * 32, h->K));
        return MELISSA_OK;
    };
    if ((rc = run())) return rc;
    estep_compare_kernel<<<(unsigned)((NK + 255) / 256), 256, 0, h->stream>>>(h->log_rho_fx, dbg, NK, out);
    unsigned long long host[32];
    cudaMemcpyAsync(host, out, sizeof host, cudaMemcpyDeviceToHost, h->stream);
    cudaStreamSynchronize(h->stream);
    static int sweep_no = 0;
    ++sweep_no;
    if (host[0]) {
        std::fprintf(stderr, "[debug_estep] sweep %d: %llu of %lld accumulators differ\n", sweep_no, host[0], (long long)NK);
        for (unsigned long long c = 0; c < host[0] && c < 4; ++c)
            std::fprintf(stderr, "   cell %lld k %lld  sweep hi=%lld lo=%016llx   check hi=%lld lo=%016llx   diff = %.6e\n",
                         (long long)(host[1 + 5 * c] / h->K), (long long)(host[1 + 5 * c] % h->K), (long long)host[3 + 5 * c],
                         host[2 + 5 * c], (long long)host[5 + 5 * c], host[4 + 5 * c],
                         ((double)(long long)host[3 + 5 * c] - (double)(long long)host[5 + 5 * c]) +
                             ((double)host[2 + 5 * c] - (double)host[4 + 5 * c]) * 0x1p-64);
    }
    // g (note: r is still the r the sweep used: resp_kernel runs later)
    cudaMemsetAsync(out, 0, 32 * 8, h->stream);
    auto run2 = [&]() -> int {
        DISPATCH_D(h->D, g_check_kernel<DD><<<(unsigned)((sl->S * 32 + 255) / 256), 256, 0, h->stream>>>(
                             sl->x_sell, sl->ybits, sl->slice_xbase, sl->slice_ybase, sl->slice_len, sl->lane_meta, sl->stats,
                             sl->slice_region, sl->lane_cell, h->U, h->r, out, sl->S * 32, h->K, h->basis));
        return MELISSA_OK;
    };
    if ((rc = run2())) return rc;
    cudaMemcpyAsync(host, out, sizeof host, cudaMemcpyDeviceToHost, h->stream);
    cudaStreamSynchronize(h->stream);
    if (host[0]) {
        std::fprintf(stderr, "[debug_estep] sweep %d: %llu values of g differ\n", sweep_no, host[0]);
        for (unsigned long long c = 0; c < host[0] && c < 4; ++c) {
            double got, want;
            std::memcpy(&got, &host[3 + 5 * c], 8);
            std::memcpy(&want, &host[4 + 5 * c], 8);
            std::fprintf(stderr, "   slice %lld lane %lld d %lld (slice length %lld): sweep %.12g  check %.12g\n",
                         (long long)(host[1 + 5 * c] >> 5), (long long)(host[1 + 5 * c] & 31), (long long)host[2 + 5 * c],
                         (long long)host[5 + 5 * c], got, want);
        }
    }
    cudaFree(dbg);
    cudaFree(out);
    return MELISSA_OK;
}

int launch_sweep3(melissa_handle *h, bool with_lb) {
    Sliced *sl = h->sl;
    const int nw = h->sweep_warps;
    if (nw == 0)
        return fail(MELISSA_ERR_ARG, "K = %d with cell blocks of %d does not fit the sweep's shared memory; "
                    "pack with a smaller \"cell_block\" option", h->K, sl->Tn);
    const size_t smem = sweep3_smem(h, nw, h->sweep_direct);
    Sweep3Args a;
    a.direct = h->sweep_direct ? 1 : 0;
    a.x_sell = sl->x_sell; a.ybits = sl->ybits; a.slice_xbase = sl->slice_xbase; a.slice_ybase = sl->slice_ybase;
    a.slice_len = sl->slice_len; a.slice_region = sl->slice_region; a.slice_cb = sl->slice_cb;
    a.lane_cell = sl->lane_cell; a.lane_meta = sl->lane_meta; a.blk_cells = sl->blk_cells;
    a.chunk_ptr = sl->chunk_ptr; a.stats = sl->stats; a.U = h->U; a.r = h->r;
    a.log_rho_fx = h->log_rho_fx; a.lb_fx = h->lb_fx; a.poison = h->poison;
    a.mills_tab = sl->mills_tab; a.lphi_tab = sl->lphi_tab; a.rbf_tab = sl->rbf_tab; a.rt = sl->rt; a.basis = h->basis;
    a.K = h->K; a.Tn = sl->Tn; a.NC = sl->NC; a.G = g_opt_sweep_group > 0 ? g_opt_sweep_group : 2;    // (chunks shorter than 640 slices are claimed slice by slice)
    a.chunk_counter = sl->chunk_counter;
    const int block = 32 * nw;
    const int grid = std::min(h->sm_count, sl->NC);
    CU(cudaMemsetAsync(sl->chunk_counter, 0, sizeof(unsigned int), h->stream));
    prof_begin(h, MELISSA_KERNEL_SWEEP);
#define SWEEP3(LB_)                                                                                         \
    DISPATCH_D(h->D, {                                                                                      \
        auto kfn = sweep3_kernel<DD, LB_>;                                                                  \
        if (!h->sweep_cfg[LB_ ? 1 : 0]) {      /* the opt-in limit is per function: ask for the most any fit can need */ \
            CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));         \
            h->sweep_cfg[LB_ ? 1 : 0] = 1;                                                                  \
        }                                                                                                   \
        kfn<<<grid, block, smem, h->stream>>>(a);                                                           \
    })
    if (with_lb) { SWEEP3(true) } else { SWEEP3(false) }
#undef SWEEP3
    prof_end(h);
    CU(cudaGetLastError());
    if (g_opt_debug_estep) return debug_estep_check(h);
    return MELISSA_OK;
}

int g_opt_contract_mma = 1;
int g_opt_contract_tma = 1;
int g_opt_contract_warps = 0;     // 0 = chosen from the slices per region

template <int MT, int NT>
int launch_contract_mma_t(melissa_handle *h, int block) {
    Sliced *sl = h->sl;
    const size_t smem = (size_t)(block / 32) * MT * NT * 64 * sizeof(double);
    auto kfn = contract_mma_kernel<MT, NT>;
    if (smem > 48 * 1024) CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kfn<<<h->M, block, smem, h->stream>>>(sl->reg_slice_ptr, sl->reg_slice, sl->lane_cell, sl->stats, h->r, h->red,
                                         h->K, h->R);
    return MELISSA_OK;
}

template <int MT, int NT>
int launch_contract_tma_t(melissa_handle *h, int block) {
    Sliced *sl = h->sl;
    const int nw = block / 32;
    const size_t ring = (size_t)nw * CTMA_NST * ((size_t)h->R * 256 + 128) + (size_t)nw * CTMA_NST * 8;
    const size_t smem = std::max(ring, (size_t)nw * MT * NT * 64 * sizeof(double));
    auto kfn = contract_tma_kernel<MT, NT>;
    if (smem > 48 * 1024) CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kfn<<<h->M, block, smem, h->stream>>>(sl->reg_slice_ptr, sl->reg_slice, sl->lane_cell, sl->stats, h->r, h->red,
                                         h->K, h->R);
    return MELISSA_OK;
}

int launch_contract_v2(melissa_handle *h) {
    Sliced *sl = h->sl;
    const int MT = (h->K + 7) / 8, NT = (h->R + 7) / 8;
    prof_begin(h, MELISSA_KERNEL_CONTRACT);
    bool done = false;
    if (g_opt_contract_mma && MT <= 4 && NT <= 6) {
        // warps per region: enough to keep >= ~32 warps per SM in flight, not more than the region has slices
        const int64_t avg_slices = std::max<int64_t>(1, sl->S / std::max(1, h->M));
        // (a region with few slices -- a small shard -- is spread over fewer warps so that more CTAs share an SM and
        // their start-up and final cross-warp sums overlap: 0.370 -> 0.344 ms on an eighth of C4)
        int nw = avg_slices < 128 ? 4 : 8;
        while (nw > 1 && avg_slices / nw < 4) nw >>= 1;
        if (g_opt_contract_warps > 0) nw = std::min(8, g_opt_contract_warps);
        const int block = 32 * nw;
        int rc = MELISSA_OK;
        // operand staged by the bulk-copy engine (contract_tma.cuh) unless switched off ("contract_tma" option)
#define CMMA(MT_, NT_) if (MT == MT_ && NT == NT_) { rc = g_opt_contract_tma ? launch_contract_tma_t<MT_, NT_>(h, block) \
                                                                             : launch_contract_mma_t<MT_, NT_>(h, block); done = true; }
        // every K <= 32 (MT <= 4) with every D <= 8 (R <= 44, NT <= 6); beyond that the CUDA-core kernel below
        CMMA(1, 1) CMMA(1, 2) CMMA(1, 3) CMMA(1, 4) CMMA(1, 5) CMMA(1, 6)
        CMMA(2, 1) CMMA(2, 2) CMMA(2, 3) CMMA(2, 4) CMMA(2, 5) CMMA(2, 6)
        CMMA(3, 1) CMMA(3, 2) CMMA(3, 3) CMMA(3, 4) CMMA(3, 5) CMMA(3, 6)
        CMMA(4, 1) CMMA(4, 2) CMMA(4, 3) CMMA(4, 4) CMMA(4, 5) CMMA(4, 6)
#undef CMMA
        if (rc) return rc;
    }
    if (!done) {
        constexpr int JC = 5, KC = 4;
        const int combos = ((h->R + JC - 1) / JC) * ((h->K + KC - 1) / KC);
        const int block = 32 * std::min(8, combos);
        contract_v2_kernel<JC, KC><<<h->M, block, 0, h->stream>>>(sl->reg_slice_ptr, sl->reg_slice, sl->lane_cell,
                                                                 sl->stats, h->r, h->red, h->K, h->R);
    }
    prof_end(h);
    CU(cudaGetLastError());
    return MELISSA_OK;
}

void free_fit(melissa_handle *h) {
    cudaFree(h->U); cudaFree(h->Spk); cudaFree(h->r); cudaFree(h->log_r); cudaFree(h->log_rho_fx);
    cudaFree(h->red); cudaFree(h->lbbuf); cudaFree(h->scal_mem); cudaFree(h->lb_fx);
    h->U = h->Spk = h->r = h->log_r = h->red = h->lbbuf = h->scal_mem = nullptr;
    h->log_rho_fx = h->lb_fx = nullptr;
    h->poison = nullptr;
    h->begun = false;
}

__global__ void count_bad_y_kernel(const uint8_t *__restrict__ y, int64_t n, unsigned long long *bad) {
    unsigned long long c = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        c += y[i] > 1;
    if (c) atomicAdd(bad, c);
}

}  // namespace

// ===========================================================================
// C ABI
// ===========================================================================
extern "C" {

const char *melissa_last_error(void) { return g_err.c_str(); }

int melissa_set_progress(melissa_progress_fn fn, void *user) {
    g_progress = fn;
    g_progress_user = user;
    return MELISSA_OK;
}
int melissa_abi_version(void) { return 1; }

static const char *pair_list_error(int code) {
    switch (code) {
        case 1: return "pair %lld: cell/region index out of range";
        case 2: return "pairs must be cell-major (pair %lld)";
        case 4: return "regions must be strictly increasing inside a cell (pair %lld)";
        case 5: return "pair %lld has no CpG";
        default: return "pair %lld has more than 65535 CpGs";
    }
}

static int pack_impl(int32_t n_cells, int32_t n_regions, int64_t n_pairs, const int64_t *pair_ptr_in,
                     const int32_t *pair_cell_in, const int32_t *pair_region_in, const double *x,
                     const uint8_t *y, int32_t basis_kind, int32_t n_basis, const double *mus,
                     double gamma, int32_t device, int32_t allow_empty_regions, bool on_device,
                     melissa_handle **out, int64_t ptr_base = 0, int32_t cell_base = 0) {
    // ptr_base / cell_base: the lists are a window of a larger job's lists (melissa_pack_multi): offsets start at
    // ptr_base instead of 0 and cells at cell_base; x and y already point at the window's first CpG
    if (!out) return fail(MELISSA_ERR_ARG, "out is NULL");
    StageTimer tm_pack;
    *out = nullptr;
    if (n_cells <= 0 || n_regions <= 0) return fail(MELISSA_ERR_ARG, "N > 0 and M > 0 are required");
    if (n_pairs <= 0 || !pair_ptr_in || !pair_cell_in || !pair_region_in || !x || !y)
        return fail(MELISSA_ERR_ARG, "empty or NULL observation arrays");
    if (n_pairs >= (int64_t)std::numeric_limits<int32_t>::max())
        return fail(MELISSA_ERR_ARG, "more than 2^31-1 pairs on one GPU; shard the cells");
    if (basis_kind != MELISSA_BASIS_RBF)
        return fail(MELISSA_ERR_ARG, "only the RBF basis object is supported (polynomial / Fourier are not pinned)");
    if (n_basis < 0 || n_basis + 1 > MAX_D || (n_basis > 0 && !mus))
        return fail(MELISSA_ERR_ARG, "basis$M must be in 0..%d", MAX_D - 1);
    if (!(gamma > 0.0) && n_basis > 0) return fail(MELISSA_ERR_ARG, "basis$gamma must be positive");
    const int64_t P = n_pairs;
    BasisDev basis_chk{};
    basis_chk.D = n_basis + 1;
    basis_chk.gamma = gamma;
    for (int j = 0; j < n_basis && j < MAX_MUS; ++j) basis_chk.mus[j] = mus[j];
    const bool use_sliced = g_opt_sliced != 0 && rbf_equally_spaced(basis_chk) && rbf_table_log2(basis_chk) >= 0;
    if (g_opt_sliced == 1 && !use_sliced)
        return fail(MELISSA_ERR_ARG, "the sliced path needs equally spaced RBF centres (and a gamma its table can resolve)");
    // Where the layout is built: on the device (default for the sliced layout), or by host threads from host copies
    // of the index lists (the flat layout's region-major index, and the "pack_host" cross-check).
    const bool host_build = !use_sliced || g_opt_pack_host != 0;
    const int64_t *pair_ptr = pair_ptr_in;
    const int32_t *pair_cell = pair_cell_in, *pair_region = pair_region_in;
    std::vector<int64_t> host_ptr;
    std::vector<int32_t> host_cell, host_region;
    if (on_device && host_build) {
        CU(cudaSetDevice(device));
        host_ptr.resize((size_t)P + 1);
        host_cell.resize((size_t)P);
        host_region.resize((size_t)P);
        CU(cudaMemcpy(host_ptr.data(), pair_ptr_in, host_ptr.size() * sizeof(int64_t), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(host_cell.data(), pair_cell_in, host_cell.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(host_region.data(), pair_region_in, host_region.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
        pair_ptr = host_ptr.data();
        pair_cell = host_cell.data();
        pair_region = host_region.data();
    }
    std::vector<int64_t> rb_ptr;
    std::vector<int32_t> rb_cell;
    if ((ptr_base != 0 || cell_base != 0) && (host_build || on_device)) {
        // the host builders want plain lists: make them (the device path below rebases after the upload instead)
        if (on_device) return fail(MELISSA_ERR_ARG, "a window of device-resident lists is not supported");
        rb_ptr.resize((size_t)P + 1);
        rb_cell.resize((size_t)P);
        for (int64_t p = 0; p <= P; ++p) rb_ptr[(size_t)p] = pair_ptr_in[p] - ptr_base;
        for (int64_t p = 0; p < P; ++p) rb_cell[(size_t)p] = pair_cell_in[p] - cell_base;
        pair_ptr = pair_ptr_in = rb_ptr.data();
        pair_cell = pair_cell_in = rb_cell.data();
        ptr_base = 0;
        cell_base = 0;
    }
    const bool have_host_lists = !on_device || host_build;
    std::vector<int64_t> reg_cnt((size_t)n_regions + 1, 0);
    int64_t nnz = 0;
    if (have_host_lists) {
    // ---- validate the pair list on the host (before the first CUDA call when the lists are host memory)
    if (pair_ptr[0] != ptr_base) return fail(MELISSA_ERR_ARG, "pair_ptr[0] must be 0");
    {
        // every rule relates pair p to pair p-1 only, so chunks are checked by separate host threads;
        // the failure with the lowest pair index is the one reported
        int nthr = (int)std::min<int64_t>(std::max(1u, std::thread::hardware_concurrency()), 16);
        if (P < 400000) nthr = 1;
        struct Bad { int64_t p = -1; int code = 0; };
        std::vector<Bad> bad((size_t)nthr);
        std::vector<std::vector<int64_t>> cnt((size_t)nthr);
        auto work = [&](int t) {
            const int64_t lo = P * t / nthr, hi = P * (t + 1) / nthr;
            std::vector<int64_t> &c = cnt[(size_t)t];
            c.assign((size_t)n_regions, 0);
            for (int64_t p = lo; p < hi; ++p) {
                const int32_t n = pair_cell[p] - cell_base, m = pair_region[p];
                int code = 0;
                if (n < 0 || n >= n_cells || m < 0 || m >= n_regions) code = 1;
                else {
                    const int32_t pn = p ? pair_cell[p - 1] - cell_base : -1, pm = p ? pair_region[p - 1] : -1;
                    if (n < pn) code = 2;
                    else if (n > pn + 1) code = 3;
                    else if (n == pn && m <= pm) code = 4;
                    else if (pair_ptr[p + 1] <= pair_ptr[p]) code = 5;
                    else if (pair_ptr[p + 1] - pair_ptr[p] > 65535) code = 6;
                }
                if (code) { bad[(size_t)t].p = p; bad[(size_t)t].code = code; return; }
                c[(size_t)m]++;
            }
        };
        if (nthr == 1) work(0);
        else {
            std::vector<std::thread> th;
            for (int t = 0; t < nthr; ++t) th.emplace_back(work, t);
            for (auto &t : th) t.join();
        }
        for (int t = 0; t < nthr; ++t) {
            const Bad &b = bad[(size_t)t];
            if (b.p < 0) continue;
            const long long p = (long long)b.p;
            switch (b.code) {
                case 1: return fail(MELISSA_ERR_ARG, "pair %lld: cell/region index out of range", p);
                case 2: return fail(MELISSA_ERR_ARG, "pairs must be cell-major (pair %lld)", p);
                case 3: return fail(MELISSA_ERR_ARG, "cell %d: No coverage across all regions. This cell should be discarded!",
                                    (b.p ? pair_cell[b.p - 1] : cell_base - 1) + 1);
                case 4: return fail(MELISSA_ERR_ARG, "regions must be strictly increasing inside a cell (pair %lld)", p);
                case 5: return fail(MELISSA_ERR_ARG, "pair %lld has no CpG", p);
                default: return fail(MELISSA_ERR_ARG, "pair %lld has more than 65535 CpGs", p);
            }
        }
        if (pair_cell[P - 1] - cell_base != n_cells - 1)
            return fail(MELISSA_ERR_ARG, "cell %d: No coverage across all regions. This cell should be discarded!",
                        pair_cell[P - 1] + 1);
        for (int t = 0; t < nthr; ++t)
            for (int32_t m = 0; m < n_regions; ++m) reg_cnt[(size_t)m + 1] += cnt[(size_t)t][(size_t)m];
    }
    nnz = pair_ptr[P] - ptr_base;
    if (!allow_empty_regions)
        for (int32_t m = 0; m < n_regions; ++m)
            if (reg_cnt[(size_t)m + 1] == 0)
                return fail(MELISSA_ERR_ARG, "region %d is not covered by any cell", m);
    for (int32_t m = 0; m < n_regions; ++m) reg_cnt[(size_t)m + 1] += reg_cnt[m];
    }
    std::vector<int32_t> reg_pair, reg_cell;
    if (!use_sliced) {
        reg_pair.resize((size_t)P);
        reg_cell.resize((size_t)P);
        std::vector<int64_t> cursor(reg_cnt.begin(), reg_cnt.end() - 1);
        for (int64_t p = 0; p < P; ++p) {
            const int64_t q = cursor[pair_region[p]]++;
            reg_pair[(size_t)q] = (int32_t)p;
            reg_cell[(size_t)q] = pair_cell[p];
        }
    }
    tm_pack.lap("validate (host)");
    // ---- device side.  Two streams: `sc` carries the uploads (index lists first, then the 9 bytes per CpG of x / y),
    //      `sm` builds the layout from the index lists while x / y are still on their way.
    CU(cudaSetDevice(device));
    melissa_handle *h = new melissa_handle();
    h->device = device;
    h->N = n_cells; h->M = n_regions; h->P = P; h->nnz = nnz;
    h->D = n_basis + 1; h->T = h->D * (h->D + 1) / 2; h->R = h->D + h->T;
    h->basis.D = h->D; h->basis.gamma = gamma;
    for (int j = 0; j < MAX_MUS; ++j) h->basis.mus[j] = j < n_basis ? mus[j] : 0.0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) h->sm_count = prop.multiProcessorCount;
    int rc = MELISSA_OK;
    unsigned long long *bad_y = nullptr, *first_bad = nullptr;
    int *reg_seen = nullptr;
    int64_t *d_ptr = nullptr, *d_src = nullptr;
    int32_t *d_cell = nullptr, *d_reg = nullptr;
    cudaStream_t sc = nullptr, sm = nullptr;
    cudaEvent_t e_idx = nullptr, e_xy = nullptr;
    auto release = [&]() {
        cudaFree(bad_y); cudaFree(first_bad); cudaFree(reg_seen); cudaFree(d_src);
        if (!on_device) { cudaFree(d_ptr); cudaFree(d_cell); cudaFree(d_reg); }
        if (sc) { cudaStreamSynchronize(sc); cudaStreamDestroy(sc); }
        if (sm) { cudaStreamSynchronize(sm); cudaStreamDestroy(sm); }
        if (e_idx) cudaEventDestroy(e_idx);
        if (e_xy) cudaEventDestroy(e_xy);
    };
    auto bail = [&](int code) { release(); melissa_free(h); return code; };
#define PK(call)                                                                                          \
    if ((call) != cudaSuccess)                                                                             \
        return bail(fail(MELISSA_ERR_CUDA, "pack: %s failed: %s", #call, cudaGetErrorString(cudaGetLastError())))
    PK(cudaStreamCreateWithFlags(&sc, cudaStreamNonBlocking));
    PK(cudaStreamCreateWithFlags(&sm, cudaStreamNonBlocking));
    PK(cudaEventCreateWithFlags(&e_idx, cudaEventDisableTiming));
    PK(cudaEventCreateWithFlags(&e_xy, cudaEventDisableTiming));
    const bool device_build = use_sliced && !host_build;
    if (device_build) {
        if (on_device) {
            d_ptr = const_cast<int64_t *>(pair_ptr_in); d_cell = const_cast<int32_t *>(pair_cell_in); d_reg = const_cast<int32_t *>(pair_region_in);
        } else {
            if ((rc = dev_alloc(&d_ptr, P + 1)) || (rc = dev_alloc(&d_cell, P)) || (rc = dev_alloc(&d_reg, P))) return bail(rc);
            PK(cudaMemcpyAsync(d_ptr, pair_ptr_in, (size_t)(P + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, sc));
            PK(cudaMemcpyAsync(d_cell, pair_cell_in, (size_t)P * sizeof(int32_t), cudaMemcpyHostToDevice, sc));
            PK(cudaMemcpyAsync(d_reg, pair_region_in, (size_t)P * sizeof(int32_t), cudaMemcpyHostToDevice, sc));
            if (ptr_base != 0 || cell_base != 0)
                pk_rebase_kernel<<<grid_for(P + 1, 256, h->sm_count, 8), 256, 0, sc>>>(d_ptr, d_cell, P, ptr_base, cell_base);
        }
    }
    PK(cudaEventRecord(e_idx, sc));
    if (!have_host_lists) {
        // lists that live on the device are validated there (same rules, same messages)
        if ((rc = dev_alloc(&first_bad, 1)) || (rc = dev_alloc(&reg_seen, n_regions))) return bail(rc);
        PK(cudaMemsetAsync(first_bad, 0xff, sizeof(unsigned long long), sm));
        PK(cudaMemsetAsync(reg_seen, 0, (size_t)n_regions * sizeof(int), sm));
        pk_validate_kernel<<<grid_for(P, 256, h->sm_count, 8), 256, 0, sm>>>(d_ptr, d_cell, d_reg, P, n_cells, n_regions, first_bad, reg_seen);
        unsigned long long fb = 0;
        int64_t head_tail[2] = {0, 0};
        int32_t last_cell = 0;
        std::vector<int> seen((size_t)n_regions);
        PK(cudaMemcpyAsync(&fb, first_bad, sizeof fb, cudaMemcpyDeviceToHost, sm));
        PK(cudaMemcpyAsync(&head_tail[0], d_ptr, 8, cudaMemcpyDeviceToHost, sm));
        PK(cudaMemcpyAsync(&head_tail[1], d_ptr + P, 8, cudaMemcpyDeviceToHost, sm));
        PK(cudaMemcpyAsync(&last_cell, d_cell + (P - 1), 4, cudaMemcpyDeviceToHost, sm));
        PK(cudaMemcpyAsync(seen.data(), reg_seen, (size_t)n_regions * sizeof(int), cudaMemcpyDeviceToHost, sm));
        PK(cudaStreamSynchronize(sm));
        if (head_tail[0] != 0) return bail(fail(MELISSA_ERR_ARG, "pair_ptr[0] must be 0"));
        if (fb != ~0ull) {
            const long long bp = (long long)(fb >> 3);
            const int code = (int)(fb & 7);
            if (code == 3) {
                int32_t prev = -1;
                if (bp > 0) PK(cudaMemcpy(&prev, d_cell + (bp - 1), 4, cudaMemcpyDeviceToHost));
                return bail(fail(MELISSA_ERR_ARG, "cell %d: No coverage across all regions. This cell should be discarded!", prev + 1));
            }
            return bail(fail(MELISSA_ERR_ARG, pair_list_error(code), bp));
        }
        if (last_cell != n_cells - 1)
            return bail(fail(MELISSA_ERR_ARG, "cell %d: No coverage across all regions. This cell should be discarded!", last_cell + 1));
        if (!allow_empty_regions)
            for (int32_t m = 0; m < n_regions; ++m)
                if (!seen[(size_t)m]) return bail(fail(MELISSA_ERR_ARG, "region %d is not covered by any cell", m));
        nnz = head_tail[1];
        h->nnz = nnz;
        tm_pack.lap("validate (device)");
    }
    if ((rc = dev_alloc(&h->x, nnz))) return bail(rc);
    if ((rc = dev_alloc(&h->y, nnz))) return bail(rc);
    PK(cudaMemcpyAsync(h->x, x, (size_t)nnz * sizeof(double), cudaMemcpyDefault, sc));
    PK(cudaMemcpyAsync(h->y, y, (size_t)nnz * sizeof(uint8_t), cudaMemcpyDefault, sc));
    // y is checked on the device once it has arrived; the verdict is read at the end
    if ((rc = dev_alloc(&bad_y, 1))) return bail(rc);
    PK(cudaMemsetAsync(bad_y, 0, sizeof(unsigned long long), sc));
    count_bad_y_kernel<<<grid_for(nnz, 256, h->sm_count, 8), 256, 0, sc>>>(h->y, nnz, bad_y);
    PK(cudaEventRecord(e_xy, sc));
    tm_pack.lap("uploads queued");
    if (use_sliced) {
        // re-lay the observations into region slices, then drop the flat copies
        Sliced *sl = new Sliced();
        h->sl = sl;
        // Cells per block = the pool in which the pairs of one region are sorted by length before they are cut
        // into slices: the larger, the less padding.  256 keeps the block's E-step sums (Tn x K x 16 bytes of
        // shared memory in the sweep) affordable up to K ~ 24; beyond that the sweep runs with fewer warps.
        sl->Tn = g_opt_cell_block > 0 ? g_opt_cell_block : 256;
        sl->NB = (n_cells + sl->Tn - 1) / sl->Tn;
        if (device_build) {
            PK(cudaStreamWaitEvent(sm, e_idx, 0));
            if ((rc = build_slice_index_device(h, sl, d_ptr, d_cell, d_reg, sm, &d_src, g_opt_sweep_chunk))) return bail(rc);
        } else {
            if ((rc = build_slice_index_host(h, sl, pair_ptr, pair_cell, pair_region, sm, &d_src))) return bail(rc);
        }
        tm_pack.lap("slice index");
        rc = finish_sliced(h, sl, d_src, sm, e_xy);
        d_src = nullptr;                       // freed by finish_sliced
        if (rc) return bail(rc);
        tm_pack.lap("tables, fill, H^T H");
        cudaFree(h->x); cudaFree(h->y);
        h->x = nullptr; h->y = nullptr;
    } else {
        if ((rc = dev_alloc(&h->pair_ptr, P + 1))) return bail(rc);
        if ((rc = dev_alloc(&h->pair_cell, P))) return bail(rc);
        if ((rc = dev_alloc(&h->pair_region, P))) return bail(rc);
        if ((rc = dev_alloc(&h->reg_ptr, (int64_t)n_regions + 1))) return bail(rc);
        if ((rc = dev_alloc(&h->reg_pair, P))) return bail(rc);
        if ((rc = dev_alloc(&h->reg_cell, P))) return bail(rc);
        if ((rc = dev_alloc(&h->stats, P * h->R))) return bail(rc);
        PK(cudaMemcpyAsync(h->pair_ptr, pair_ptr_in, (size_t)(P + 1) * sizeof(int64_t), cudaMemcpyDefault, sc));
        PK(cudaMemcpyAsync(h->pair_cell, pair_cell_in, (size_t)P * sizeof(int32_t), cudaMemcpyDefault, sc));
        PK(cudaMemcpyAsync(h->pair_region, pair_region_in, (size_t)P * sizeof(int32_t), cudaMemcpyDefault, sc));
        PK(cudaMemcpyAsync(h->reg_ptr, reg_cnt.data(), (size_t)(n_regions + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, sc));
        PK(cudaMemcpyAsync(h->reg_pair, reg_pair.data(), (size_t)P * sizeof(int32_t), cudaMemcpyHostToDevice, sc));
        PK(cudaMemcpyAsync(h->reg_cell, reg_cell.data(), (size_t)P * sizeof(int32_t), cudaMemcpyHostToDevice, sc));
        const int grid = grid_for(P, 128, h->sm_count, 16);
        auto run = [&]() -> int {
            DISPATCH_D(h->D, hh_kernel<DD><<<grid, 128, 0, sc>>>(h->pair_ptr, h->x, h->stats, P, h->basis));
            return MELISSA_OK;
        };
        if ((rc = run())) return bail(rc);
        h->launches++;
    }
    {
        unsigned long long bad_h = 0;
        PK(cudaMemcpyAsync(&bad_h, bad_y, sizeof bad_h, cudaMemcpyDeviceToHost, sc));
        PK(cudaStreamSynchronize(sc));
        if (bad_h) return bail(fail(MELISSA_ERR_ARG, "%llu methylation states are not 0/1", bad_h));
    }
    PK(cudaStreamSynchronize(sm));
#undef PK
    release();
    *out = h;
    return MELISSA_OK;
}

int melissa_pack(int32_t n_cells, int32_t n_regions, int64_t n_pairs, const int64_t *pair_ptr,
                 const int32_t *pair_cell, const int32_t *pair_region, const double *x,
                 const uint8_t *y, int32_t basis_kind, int32_t n_basis, const double *mus,
                 double gamma, int32_t device, int32_t allow_empty_regions, melissa_handle **out) {
    return pack_impl(n_cells, n_regions, n_pairs, pair_ptr, pair_cell, pair_region, x, y, basis_kind, n_basis, mus,
                     gamma, device, allow_empty_regions, false, out);
}

int melissa_pack_device(int32_t n_cells, int32_t n_regions, int64_t n_pairs, const int64_t *pair_ptr,
                        const int32_t *pair_cell, const int32_t *pair_region, const double *x,
                        const uint8_t *y, int32_t basis_kind, int32_t n_basis, const double *mus,
                        double gamma, int32_t device, int32_t allow_empty_regions, melissa_handle **out) {
    return pack_impl(n_cells, n_regions, n_pairs, pair_ptr, pair_cell, pair_region, x, y, basis_kind, n_basis, mus,
                     gamma, device, allow_empty_regions, true, out);
}

static int profiles_mle_impl(melissa_handle *h, double lambda, int32_t n_iter, double *w_out, bool out_on_device) {
    if (!h) return fail(MELISSA_ERR_ARG, "handle is NULL");
    if (!w_out) return fail(MELISSA_ERR_ARG, "w_out is NULL");
    if (!(lambda >= 0.0) || n_iter < 1) return fail(MELISSA_ERR_ARG, "lambda >= 0 and n_iter >= 1 are required");
    if (!h->sl) return fail(MELISSA_ERR_STATE, "melissa_profiles_mle needs the sliced layout (equally spaced RBF centres)");
    CU(cudaSetDevice(h->device));
    Sliced *sl = h->sl;
    const int D = h->D;
    double *d_out = nullptr;
    int rc = dev_alloc(&d_out, sl->S * 32 * D);
    if (rc) return rc;
    const size_t smem = (size_t)sl->rt.n_entries * sizeof(double2) + ((size_t)MILLS_N + 1) * sizeof(double);
    const int grid = (int)std::min<int64_t>((sl->S + 7) / 8, (int64_t)h->sm_count * 2);
    auto run = [&]() -> int {
        DISPATCH_D(D, {
            auto kfn = mle_kernel<DD>;
            CU(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            kfn<<<grid, 256, smem, h->stream>>>(sl->x_sell, sl->ybits, sl->slice_xbase, sl->slice_ybase, sl->slice_len,
                                                sl->lane_meta, sl->mills_tab, sl->rbf_tab, sl->rt, h->basis, d_out, sl->S,
                                                lambda, n_iter);
        });
        return MELISSA_OK;
    };
    if ((rc = run())) { cudaFree(d_out); return rc; }
    double *d_pairs = out_on_device ? w_out : nullptr;
    if (!out_on_device && (rc = dev_alloc(&d_pairs, h->P * D))) { cudaFree(d_out); return rc; }
    lanes_to_pairs_kernel<<<(unsigned)((sl->S * 32 + 255) / 256), 256, 0, h->stream>>>(d_out, sl->lane_pair, d_pairs,
                                                                                         sl->S * 32, D, h->R, 0);
    h->launches += 2;
    cudaError_t e = cudaSuccess;
    if (!out_on_device) e = cudaMemcpyAsync(w_out, d_pairs, (size_t)h->P * D * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d_out);
    if (!out_on_device) cudaFree(d_pairs);
    if (e != cudaSuccess) return fail(MELISSA_ERR_CUDA, "melissa_profiles_mle failed: %s", cudaGetErrorString(e));
    return MELISSA_OK;
}

int melissa_profiles_mle(melissa_handle *h, double lambda, int32_t n_iter, double *w_out) {
    return profiles_mle_impl(h, lambda, n_iter, w_out, false);
}

int melissa_profiles_mle_device(melissa_handle *h, double lambda, int32_t n_iter, double *w_out_device) {
    return profiles_mle_impl(h, lambda, n_iter, w_out_device, true);
}

int melissa_set_option(const char *key, int64_t value) {
    if (!key) return fail(MELISSA_ERR_ARG, "key is NULL");
    if (!std::strcmp(key, "cell_block")) {
        if (value < 0 || value > 4096) return fail(MELISSA_ERR_ARG, "cell_block must be in 0..4096");
        g_opt_cell_block = (int)value;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "clusters_hint")) return MELISSA_OK;      // round-1 knob: the layout no longer depends on K
    if (!std::strcmp(key, "pack_host")) {
        g_opt_pack_host = value != 0;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "debug_estep")) {
        g_opt_debug_estep = value != 0;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "sweep_chunk")) {
        if (value < 0 || value > (1 << 20)) return fail(MELISSA_ERR_ARG, "sweep_chunk must be in 0..2^20");
        g_opt_sweep_chunk = (int)value;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "sweep_taper")) {
        g_opt_sweep_taper = value != 0;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "sweep_group")) {
        if (value < 0 || value > 64) return fail(MELISSA_ERR_ARG, "sweep_group must be in 0..64 (0 = automatic)");
        g_opt_sweep_group = (int)value;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "sweep_warps")) {
        if (value < 0 || value > SWEEP3_THREADS / 32) return fail(MELISSA_ERR_ARG, "sweep_warps must be in 0..%d", SWEEP3_THREADS / 32);
        g_opt_sweep_warps = (int)value;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "contract_mma")) {
        g_opt_contract_mma = value != 0;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "contract_tma")) {
        g_opt_contract_tma = value != 0;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "contract_warps")) {
        if (value < 0 || value > 8) return fail(MELISSA_ERR_ARG, "contract_warps must be in 0..8 (0 = automatic)");
        g_opt_contract_warps = (int)value;
        return MELISSA_OK;
    }
    if (!std::strcmp(key, "sliced")) {
        if (value < -1 || value > 1) return fail(MELISSA_ERR_ARG, "sliced must be -1 (auto), 0 or 1");
        g_opt_sliced = (int)value;
        return MELISSA_OK;
    }
    return fail(MELISSA_ERR_ARG, "unknown option '%s'", key);
}

int melissa_profile_enable(melissa_handle *h, int32_t on) {
    if (!h) return fail(MELISSA_ERR_ARG, "handle is NULL");
    cudaSetDevice(h->device);
    for (cudaEvent_t e : h->prof_ev) cudaEventDestroy(e);
    h->prof_ev.clear();
    h->prof_id.clear();
    h->prof_on = on != 0;
    return MELISSA_OK;
}

int melissa_profile_read(melissa_handle *h, int32_t kernel_id, double *total_ms, int64_t *count) {
    if (!h) return fail(MELISSA_ERR_ARG, "handle is NULL");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    double tot = 0.0;
    int64_t cnt = 0;
    for (size_t i = 0; i < h->prof_id.size(); ++i) {
        if (h->prof_id[i] != kernel_id) continue;
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, h->prof_ev[2 * i], h->prof_ev[2 * i + 1]));
        tot += ms;
        ++cnt;
    }
    if (total_ms) *total_ms = tot;
    if (count) *count = cnt;
    return MELISSA_OK;
}

int melissa_free(melissa_handle *h) {
    if (!h) return MELISSA_OK;
    cudaSetDevice(h->device);
    for (cudaEvent_t e : h->prof_ev) cudaEventDestroy(e);
    free_fit(h);
    free_sliced(h);
    cudaFree(h->pair_ptr); cudaFree(h->pair_cell); cudaFree(h->pair_region); cudaFree(h->x);
    cudaFree(h->y); cudaFree(h->reg_ptr); cudaFree(h->reg_pair); cudaFree(h->reg_cell);
    cudaFree(h->stats);
    delete h;
    return MELISSA_OK;
}

int melissa_vb_begin(melissa_handle *h, int32_t K, const double *w0, const double *delta_0,
                     double alpha_0, double beta_0, int64_t n_cells_total, void *stream) {
    if (!h) return fail(MELISSA_ERR_ARG, "handle is NULL");
    if (K <= 0 || K > 256) return fail(MELISSA_ERR_ARG, "K must be in 1..256");
    if (!w0 || !delta_0) return fail(MELISSA_ERR_ARG, "w0 / delta_0 is NULL");
    if (n_cells_total < h->N) return fail(MELISSA_ERR_ARG, "n_cells_total < cells in this shard");
    for (int k = 0; k < K; ++k)
        if (!(delta_0[k] > 0.0)) return fail(MELISSA_ERR_ARG, "delta_0 must be positive");
    CU(cudaSetDevice(h->device));
    free_fit(h);
    h->K = K;
    h->stream = (cudaStream_t)stream;
    h->N_total = n_cells_total;
    h->alpha_0 = alpha_0;
    h->alpha_k = alpha_0 + (double)h->M * h->D / 2.0;                       // R/melissa_vb.R:264
    h->beta_0 = std::isnan(beta_0) ? std::sqrt(h->alpha_k) : beta_0;        // :271-273
    if (!(h->beta_0 > 0.0) || !(alpha_0 > 0.0)) return fail(MELISSA_ERR_ARG, "alpha_0 and beta_0 must be positive");
    const int64_t MK = (int64_t)h->M * K;
    int rc;
    if ((rc = dev_alloc(&h->U, MK * h->R))) return rc;
    if ((rc = dev_alloc(&h->Spk, MK * h->T))) return rc;
    if ((rc = dev_alloc(&h->r, (int64_t)h->N * K))) return rc;
    if ((rc = dev_alloc(&h->log_r, (int64_t)h->N * K))) return rc;
    if ((rc = dev_alloc(&h->log_rho_fx, (int64_t)h->N * K * 2))) return rc;
    if ((rc = dev_alloc(&h->lb_fx, 6))) return rc;
    h->poison = reinterpret_cast<unsigned int *>(h->lb_fx + 4);
    h->red_len = MK * h->R + K + 1;
    if ((rc = dev_alloc(&h->red, h->red_len))) return rc;
    if ((rc = dev_alloc(&h->lbbuf, 2))) return rc;
    const int64_t n_scal = (int64_t)6 * K + 2 + Scalars::fx_words(K);
    if ((rc = dev_alloc(&h->scal_mem, n_scal))) return rc;
    double *s = h->scal_mem;
    h->sc.beta = s; h->sc.delta0 = s + K; h->sc.delta = s + 2 * K; h->sc.elogpi = s + 3 * K;
    h->sc.Eww = s + 4 * K; h->sc.pi = s + 5 * K;
    h->sc.logdet = s + 6 * K; h->sc.lb_out = s + 6 * K + 1;
    h->sc.fx = reinterpret_cast<unsigned long long *>(s + 6 * K + 2);
    cudaStream_t st = h->stream;
    CU(cudaMemsetAsync(h->scal_mem, 0, (size_t)n_scal * sizeof(double), st));
    CU(cudaMemsetAsync(h->log_rho_fx, 0, (size_t)h->N * K * 2 * sizeof(unsigned long long), st));
    CU(cudaMemsetAsync(h->lb_fx, 0, 6 * sizeof(unsigned long long), st));
    CU(cudaMemsetAsync(h->r, 0, (size_t)h->N * K * sizeof(double), st));
    CU(cudaMemsetAsync(h->log_r, 0, (size_t)h->N * K * sizeof(double), st));
    CU(cudaMemsetAsync(h->red, 0, (size_t)h->red_len * sizeof(double), st));
    CU(cudaMemsetAsync(h->lbbuf, 0, 2 * sizeof(double), st));
    CU(cudaMemcpyAsync(h->sc.delta0, delta_0, (size_t)K * sizeof(double), cudaMemcpyHostToDevice, st));
    // stage w0 through the reduction buffer's tail-free head (M*D*K <= M*K*R doubles)
    CU(cudaMemcpyAsync(h->red, w0, (size_t)MK * h->D * sizeof(double), cudaMemcpyHostToDevice, st));
    {
        const dim3 grid((h->M + 127) / 128, K);
        auto run = [&]() -> int {
            DISPATCH_D(h->D, init_params_kernel<DD><<<grid, 128, 0, st>>>(h->red, h->U, h->Spk, h->M, K));
            return MELISSA_OK;
        };
        if ((rc = run())) return rc;
        h->launches++;
    }
    CU(cudaMemsetAsync(h->red, 0, (size_t)h->red_len * sizeof(double), st));
    scalars_kernel<<<1, 32, 0, st>>>(h->sc, h->red + MK * h->R, K, h->alpha_k, h->beta_0, (double)h->N_total, 1);
    h->launches++;
    CU(cudaGetLastError());
    h->begun = true;
    // mu from rowMeans(w), E[z], E-step terms for i = 2  (R/melissa_vb.R:283-302)
    if (h->sl) {
        h->sweep_warps = sweep3_warps(h, &h->sweep_direct);
        // rowMeans over clusters = responsibilities 1/K for every cell in this first sweep
        fill_kernel<<<grid_for((int64_t)h->N * K, 256, h->sm_count, 8), 256, 0, st>>>(h->r, (int64_t)h->N * K, 1.0 / (double)K);
        h->launches++;
        return launch_sweep3(h, false);
    }
    return launch_sweep(h, false, true);
}

int melissa_vb_phase_a(melissa_handle *h) {
    if (!h || !h->begun) return fail(MELISSA_ERR_STATE, "melissa_vb_begin has not been called");
    CU(cudaSetDevice(h->device));
    const int64_t MK = (int64_t)h->M * h->K;
    CU(cudaMemsetAsync(h->red, 0, (size_t)h->red_len * sizeof(double), h->stream));
    double *colsum = h->red + MK * h->R, *rlogr = colsum + h->K;
    prof_begin(h, MELISSA_KERNEL_RESP);
    resp_kernel<<<(h->N + 127) / 128, 128, 0, h->stream>>>(h->log_rho_fx, h->poison, h->sc.elogpi, h->r, h->log_r,
                                                              colsum, rlogr, h->sc, h->N, h->K);
    prof_end(h);
    CU(cudaGetLastError());
    return h->sl ? launch_contract_v2(h) : launch_contract(h);
}

int melissa_vb_phase_b(melissa_handle *h, int32_t with_lb) {
    if (!h || !h->begun) return fail(MELISSA_ERR_STATE, "melissa_vb_begin has not been called");
    CU(cudaSetDevice(h->device));
    const int64_t MK = (int64_t)h->M * h->K;
    double *colsum = h->red + MK * h->R;
    {
        const dim3 grid((h->M + 127) / 128, h->K);
        auto run = [&]() -> int {
            DISPATCH_D(h->D, solve_kernel<DD><<<grid, 128, 0, h->stream>>>(
                                 h->red, h->U, h->Spk, h->sc.beta, h->sc, h->M, h->K, h->alpha_k));
            return MELISSA_OK;
        };
        int rc;
        prof_begin(h, MELISSA_KERNEL_SOLVE);
        if ((rc = run())) return rc;
        prof_end(h);
    }
    prof_begin(h, MELISSA_KERNEL_SCALARS);
    scalars_kernel<<<1, 32, 0, h->stream>>>(h->sc, colsum, h->K, h->alpha_k, h->beta_0, (double)h->N_total, 0);
    prof_end(h);
    CU(cudaGetLastError());
    int rc = h->sl ? launch_sweep3(h, with_lb != 0) : launch_sweep(h, with_lb != 0, false);
    if (rc) return rc;
    if (with_lb) {                      // this shard's lb_pz_qz as a double, ready for the all-reduce
        lb_finish_kernel<<<1, 32, 0, h->stream>>>(h->lb_fx, h->lbbuf);
        h->launches++;
        CU(cudaGetLastError());
    }
    return MELISSA_OK;
}

int melissa_vb_lower_bound(melissa_handle *h, double *lb_out) {
    if (!h || !h->begun) return fail(MELISSA_ERR_STATE, "melissa_vb_begin has not been called");
    CU(cudaSetDevice(h->device));
    const int64_t MK = (int64_t)h->M * h->K;
    double *colsum = h->red + MK * h->R, *rlogr = colsum + h->K;
    prof_begin(h, MELISSA_KERNEL_LB);
    lb_kernel<<<1, 32, 0, h->stream>>>(h->sc, colsum, rlogr, h->lbbuf, h->K, h->M, h->D, h->alpha_0, h->alpha_k,
                                       h->beta_0);
    prof_end(h);
    CU(cudaGetLastError());
    double v = 0.0;
    CU(cudaMemcpyAsync(&v, h->sc.lb_out, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    CU(cudaStreamSynchronize(h->stream));
    if (lb_out) *lb_out = v;
    return MELISSA_OK;
}

int melissa_vb_reduce_buffer(melissa_handle *h, void **dev_ptr, int64_t *n_doubles) {
    if (!h || !h->begun) return fail(MELISSA_ERR_STATE, "melissa_vb_begin has not been called");
    if (dev_ptr) *dev_ptr = h->red;
    if (n_doubles) *n_doubles = h->red_len;
    return MELISSA_OK;
}

int melissa_vb_lb_buffer(melissa_handle *h, void **dev_ptr, int64_t *n_doubles) {
    if (!h || !h->begun) return fail(MELISSA_ERR_STATE, "melissa_vb_begin has not been called");
    if (dev_ptr) *dev_ptr = h->lbbuf;
    if (n_doubles) *n_doubles = 2;
    return MELISSA_OK;
}

int melissa_sync(melissa_handle *h) {
    if (!h) return fail(MELISSA_ERR_ARG, "handle is NULL");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    return MELISSA_OK;
}

int64_t melissa_launch_count(const melissa_handle *h) { return h ? h->launches : 0; }

int melissa_layout_info(const melissa_handle *h, int64_t *out, int32_t n) {
    if (!h || !out) return fail(MELISSA_ERR_ARG, "handle / out is NULL");
    const int64_t v[8] = {h->sl ? 1 : 0, h->sl ? h->sl->Tn : 0, h->sl ? h->sl->S : 0, h->sl ? h->sl->x_len : h->nnz,
                          h->nnz, h->P, h->sl ? h->sl->NC : 0, h->sl ? h->sweep_warps : 0};
    for (int i = 0; i < n && i < 8; ++i) out[i] = v[i];
    return MELISSA_OK;
}

int melissa_vb_fetch(melissa_handle *h, double *W, double *W_Sigma, double *r_nk, double *delta,
                     double *alpha, double *beta, double *pi_k) {
    if (!h || !h->begun) return fail(MELISSA_ERR_STATE, "melissa_vb_begin has not been called");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    const int K = h->K, M = h->M, D = h->D, T = h->T, R = h->R, N = h->N;
    const int64_t MK = (int64_t)M * K;
    if (W) {
        std::vector<double> U((size_t)MK * R);
        CU(cudaMemcpy(U.data(), h->U, U.size() * sizeof(double), cudaMemcpyDeviceToHost));
        for (int m = 0; m < M; ++m)
            for (int k = 0; k < K; ++k)
                for (int d = 0; d < D; ++d)
                    W[m + (int64_t)M * (d + (int64_t)D * k)] = U[((size_t)m * K + k) * R + d];
    }
    if (W_Sigma) {
        std::vector<double> S((size_t)MK * T);
        CU(cudaMemcpy(S.data(), h->Spk, S.size() * sizeof(double), cudaMemcpyDeviceToHost));
        for (int k = 0; k < K; ++k)
            for (int m = 0; m < M; ++m) {
                double *blk = W_Sigma + ((int64_t)k * M + m) * D * D;
                const double *sp = S.data() + ((size_t)m * K + k) * T;
                for (int i = 0; i < D; ++i)
                    for (int j = i; j < D; ++j) {
                        const double v = sp[tri_index(D, i, j)];
                        blk[i + D * j] = v;
                        blk[j + D * i] = v;
                    }
            }
    }
    if (r_nk) {
        std::vector<double> r((size_t)N * K);
        CU(cudaMemcpy(r.data(), h->r, r.size() * sizeof(double), cudaMemcpyDeviceToHost));
        for (int n = 0; n < N; ++n)
            for (int k = 0; k < K; ++k) r_nk[n + (int64_t)N * k] = r[(size_t)n * K + k];
    }
    if (delta) CU(cudaMemcpy(delta, h->sc.delta, K * sizeof(double), cudaMemcpyDeviceToHost));
    if (beta) CU(cudaMemcpy(beta, h->sc.beta, K * sizeof(double), cudaMemcpyDeviceToHost));
    if (pi_k) CU(cudaMemcpy(pi_k, h->sc.pi, K * sizeof(double), cudaMemcpyDeviceToHost));
    if (alpha)
        for (int k = 0; k < K; ++k) alpha[k] = h->alpha_k;
    return MELISSA_OK;
}

int melissa_vb_fetch_HtEz(melissa_handle *h, double *out) {
    if (!h || !h->begun) return fail(MELISSA_ERR_STATE, "melissa_vb_begin has not been called");
    if (!out) return fail(MELISSA_ERR_ARG, "out is NULL");
    CU(cudaSetDevice(h->device));
    CU(cudaStreamSynchronize(h->stream));
    if (h->sl) {
        const Sliced *sl = h->sl;
        double *d_pairs = nullptr;
        int rc = dev_alloc(&d_pairs, h->P * h->D);
        if (rc) return rc;
        lanes_to_pairs_kernel<<<(unsigned)((sl->S * 32 + 255) / 256), 256, 0, h->stream>>>(sl->stats, sl->lane_pair, d_pairs,
                                                                                             sl->S * 32, h->D, h->R, 1);
        h->launches++;
        cudaError_t e = cudaMemcpyAsync(out, d_pairs, (size_t)h->P * h->D * sizeof(double), cudaMemcpyDeviceToHost, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        cudaFree(d_pairs);
        if (e != cudaSuccess) return fail(MELISSA_ERR_CUDA, "melissa_vb_fetch_HtEz failed: %s", cudaGetErrorString(e));
        return MELISSA_OK;
    }
    CU(cudaMemcpy2D(out, (size_t)h->D * sizeof(double), h->stats, (size_t)h->R * sizeof(double),
                    (size_t)h->D * sizeof(double), (size_t)h->P, cudaMemcpyDeviceToHost));
    return MELISSA_OK;
}

int melissa_vb_run(melissa_handle *h, int32_t K, const double *w0, const double *delta_0,
                   double alpha_0, double beta_0, int32_t vb_max_iter, double epsilon_conv,
                   int32_t lb_every, double *W, double *W_Sigma, double *r_nk, double *delta,
                   double *alpha, double *beta, double *pi_k, double *lb, int32_t lb_cap,
                   int32_t *n_lb, int32_t *n_iter, int32_t *flags) {
    if (!h) return fail(MELISSA_ERR_ARG, "handle is NULL");
    if (vb_max_iter < 2) return fail(MELISSA_ERR_ARG, "vb_max_iter must be >= 2");
    if (lb_every <= 0) return fail(MELISSA_ERR_ARG, "lb_every must be positive");
    int rc;
    if ((rc = melissa_vb_begin(h, K, w0, delta_0, alpha_0, beta_0, h->N, nullptr))) return rc;
    double lb_prev = -std::numeric_limits<double>::infinity();              // LB <- c(-Inf), :242
    int nlb = 0, fl = 0, it = 1;
    if (lb && lb_cap > 0) lb[0] = lb_prev;
    nlb = 1;
    for (int i = 2; i <= vb_max_iter; ++i) {                                // :309
        it = i;
        if ((rc = melissa_vb_phase_a(h))) return rc;
        const bool with_lb = (i % lb_every == 0) || i == vb_max_iter;       // :458
        if ((rc = melissa_vb_phase_b(h, with_lb))) return rc;
        if (with_lb) {
            double v;
            if ((rc = melissa_vb_lower_bound(h, &v))) return rc;
            if (lb && nlb < lb_cap) lb[nlb] = v;
            ++nlb;
            if (v < lb_prev) fl |= MELISSA_FLAG_LB_DECREASED;               // :495-497
            const bool conv = std::fabs(v - lb_prev) < epsilon_conv;        // :499-501
            const double diff = v - lb_prev;
            lb_prev = v;
            if (conv) break;
            if (g_progress && g_progress(i, v, diff, g_progress_user)) {    // the host asks to stop (user interrupt)
                fl |= MELISSA_FLAG_INTERRUPTED;
                break;
            }
        }
        if (i == vb_max_iter) fl |= MELISSA_FLAG_NOT_CONVERGED;             // :511-513
    }
    if ((rc = melissa_vb_fetch(h, W, W_Sigma, r_nk, delta, alpha, beta, pi_k))) return rc;
    if (n_lb) *n_lb = nlb;
    if (n_iter) *n_iter = it;
    if (flags) *flags = fl;
    return MELISSA_OK;
}

int melissa_impute(int32_t n_cells, int32_t n_regions, int32_t K, int32_t basis_kind, int32_t n_basis,
                   const double *mus, double gamma, const double *W, const double *r_nk, int64_t n_pairs,
                   const int64_t *pair_ptr, const int32_t *pair_cell, const int32_t *pair_region,
                   const double *x, double *pred, int32_t device) {
    if (n_cells <= 0 || n_regions <= 0 || K <= 0) return fail(MELISSA_ERR_ARG, "N, M and K must be positive");
    if (basis_kind != MELISSA_BASIS_RBF) return fail(MELISSA_ERR_ARG, "only the RBF basis object is supported");
    if (n_basis < 0 || n_basis + 1 > MAX_D || (n_basis > 0 && !mus))
        return fail(MELISSA_ERR_ARG, "basis$M must be in 0..%d", MAX_D - 1);
    if (n_pairs <= 0 || !pair_ptr || !pair_cell || !pair_region || !x || !pred || !W || !r_nk)
        return fail(MELISSA_ERR_ARG, "empty or NULL arrays");
    for (int64_t p = 0; p < n_pairs; ++p) {
        if (pair_cell[p] < 0 || pair_cell[p] >= n_cells || pair_region[p] < 0 || pair_region[p] >= n_regions)
            return fail(MELISSA_ERR_ARG, "test pair %lld: cell/region index out of range", (long long)p);
        if (pair_ptr[p + 1] < pair_ptr[p]) return fail(MELISSA_ERR_ARG, "test pair_ptr must be non-decreasing");
    }
    if (pair_ptr[0] != 0) return fail(MELISSA_ERR_ARG, "test pair_ptr[0] must be 0");
    CU(cudaSetDevice(device));
    BasisDev basis{};
    const int D = n_basis + 1;
    basis.D = D; basis.gamma = gamma;
    for (int j = 0; j < MAX_MUS; ++j) basis.mus[j] = j < n_basis ? mus[j] : 0.0;
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device);
    const int64_t nnz = pair_ptr[n_pairs];
    const int64_t nW = (int64_t)n_regions * D * K, nr = (int64_t)n_cells * K;
    int64_t *d_ptr = nullptr; int32_t *d_cell = nullptr, *d_reg = nullptr;
    double *d_x = nullptr, *d_pred = nullptr, *d_W = nullptr, *d_r = nullptr;
    int rc = MELISSA_OK;
    auto cleanup = [&]() {
        cudaFree(d_ptr); cudaFree(d_cell); cudaFree(d_reg); cudaFree(d_x); cudaFree(d_pred); cudaFree(d_W); cudaFree(d_r);
    };
    if ((rc = dev_alloc(&d_ptr, n_pairs + 1)) || (rc = dev_alloc(&d_cell, n_pairs)) ||
        (rc = dev_alloc(&d_reg, n_pairs)) || (rc = dev_alloc(&d_x, nnz)) || (rc = dev_alloc(&d_pred, nnz)) ||
        (rc = dev_alloc(&d_W, nW)) || (rc = dev_alloc(&d_r, nr))) {
        cleanup();
        return rc;
    }
    cudaMemcpyAsync(d_ptr, pair_ptr, (size_t)(n_pairs + 1) * sizeof(int64_t), cudaMemcpyHostToDevice, nullptr);
    cudaMemcpyAsync(d_cell, pair_cell, (size_t)n_pairs * sizeof(int32_t), cudaMemcpyHostToDevice, nullptr);
    cudaMemcpyAsync(d_reg, pair_region, (size_t)n_pairs * sizeof(int32_t), cudaMemcpyHostToDevice, nullptr);
    cudaMemcpyAsync(d_x, x, (size_t)nnz * sizeof(double), cudaMemcpyHostToDevice, nullptr);
    cudaMemcpyAsync(d_W, W, (size_t)nW * sizeof(double), cudaMemcpyHostToDevice, nullptr);
    cudaMemcpyAsync(d_r, r_nk, (size_t)nr * sizeof(double), cudaMemcpyHostToDevice, nullptr);
    const int grid = grid_for(n_pairs, 128, sm_count, 16);
    auto run = [&]() -> int {
        DISPATCH_D(D, impute_kernel<DD><<<grid, 128>>>(d_ptr, d_cell, d_reg, d_x, d_W, d_r, d_pred, n_pairs,
                                                      n_cells, n_regions, K, basis));
        return MELISSA_OK;
    };
    if ((rc = run())) { cleanup(); return rc; }
    cudaMemcpyAsync(pred, d_pred, (size_t)nnz * sizeof(double), cudaMemcpyDeviceToHost, nullptr);
    const cudaError_t e = cudaDeviceSynchronize();
    cleanup();
    if (e != cudaSuccess) return fail(MELISSA_ERR_CUDA, "impute failed: %s", cudaGetErrorString(e));
    return MELISSA_OK;
}


// ---------------------------------------------------------------------------
// Imputation without any device allocation.
//
// melissa_impute() above stages the caller's host arrays in device buffers it
// cudaMalloc's and frees on every call.  A host that must not (or cannot)
// allocate device memory -- a process whose GPU memory has been handed in full
// to another runtime's pool, so that a further cudaMalloc fails; a service that
// imputes one small batch per request and has to keep the allocator (and the
// implicit synchronisation of cudaFree) out of its request path -- gets the same
// result from the entry point below: the staging area is this fixed array, which
// lives in the library image (every GPU that loads the library holds these bytes
// until the process exits).  The copies and the launch go through the default
// stream and block the calling thread.
//
// `g_noalloc_scratch` is only a host-side handle here; the kernel needs the
// device address, obtained once per device with cudaGetSymbolAddress while that
// device is current.
// ---------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_noalloc_scratch[MELISSA_NOALLOC_SCRATCH_BYTES];

static std::mutex g_noalloc_mu;
static unsigned char *g_noalloc_ptr[64] = {nullptr};

static int noalloc_device_ptr(int device, unsigned char **out) {
    if (device < 0 || device >= 64) return fail(MELISSA_ERR_ARG, "device index %d out of range", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (!g_noalloc_ptr[device]) {
        CU(cudaSetDevice(device));
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return fail(MELISSA_ERR_CUDA, "cudaGetSymbolAddress(scratch) failed on device %d: %s", device,
                        cudaGetErrorString(e));
        g_noalloc_ptr[device] = static_cast<unsigned char *>(p);
    }
    *out = g_noalloc_ptr[device];
    return MELISSA_OK;
}

int melissa_impute_noalloc(int32_t n_cells, int32_t n_regions, int32_t K, int32_t basis_kind, int32_t n_basis,
                           const double *mus, double gamma, const double *W, const double *r_nk, int64_t n_pairs,
                           const int64_t *pair_ptr, const int32_t *pair_cell, const int32_t *pair_region,
                           const double *x, double *pred, int32_t device) {
    if (n_cells <= 0 || n_regions <= 0 || K <= 0) return fail(MELISSA_ERR_ARG, "N, M and K must be positive");
    if (basis_kind != MELISSA_BASIS_RBF) return fail(MELISSA_ERR_ARG, "only the RBF basis object is supported");
    if (n_basis < 0 || n_basis + 1 > MAX_D || (n_basis > 0 && !mus))
        return fail(MELISSA_ERR_ARG, "basis$M must be in 0..%d", MAX_D - 1);
    if (n_pairs <= 0 || !pair_ptr || !pair_cell || !pair_region || !x || !pred || !W || !r_nk)
        return fail(MELISSA_ERR_ARG, "empty or NULL arrays");
    for (int64_t p = 0; p < n_pairs; ++p) {
        if (pair_cell[p] < 0 || pair_cell[p] >= n_cells || pair_region[p] < 0 || pair_region[p] >= n_regions)
            return fail(MELISSA_ERR_ARG, "test pair %lld: cell/region index out of range", (long long)p);
        if (pair_ptr[p + 1] < pair_ptr[p]) return fail(MELISSA_ERR_ARG, "test pair_ptr must be non-decreasing");
    }
    if (pair_ptr[0] != 0) return fail(MELISSA_ERR_ARG, "test pair_ptr[0] must be 0");
    const int D = n_basis + 1;
    const int64_t nnz = pair_ptr[n_pairs];
    // carve the fixed array: every piece starts on a 256-byte boundary
    auto up = [](int64_t b) { return (b + 255) & ~int64_t(255); };
    const int64_t sz_ptr = up((n_pairs + 1) * (int64_t)sizeof(int64_t)), sz_idx = up(n_pairs * (int64_t)sizeof(int32_t));
    const int64_t sz_x = up(nnz * (int64_t)sizeof(double));
    const int64_t sz_W = up((int64_t)n_regions * D * K * (int64_t)sizeof(double));
    const int64_t sz_r = up((int64_t)n_cells * K * (int64_t)sizeof(double));
    const int64_t need = sz_ptr + 2 * sz_idx + 2 * sz_x + sz_W + sz_r;
    if (need > (int64_t)MELISSA_NOALLOC_SCRATCH_BYTES)
        return fail(MELISSA_ERR_ARG,
                    "melissa_impute_noalloc: this call needs %lld bytes of device staging but the fixed scratch "
                    "array holds %lld; split the batch or use melissa_impute",
                    (long long)need, (long long)MELISSA_NOALLOC_SCRATCH_BYTES);
    // the caller's current device is put back on every way out
    struct DeviceGuard {
        int prev = -1;
        DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    unsigned char *base = nullptr;
    int rc = noalloc_device_ptr(device, &base);
    if (rc) return rc;
    CU(cudaSetDevice(device));
    unsigned char *cur = base;
    int64_t *d_ptr = reinterpret_cast<int64_t *>(cur); cur += sz_ptr;
    int32_t *d_cell = reinterpret_cast<int32_t *>(cur); cur += sz_idx;
    int32_t *d_reg = reinterpret_cast<int32_t *>(cur); cur += sz_idx;
    double *d_x = reinterpret_cast<double *>(cur); cur += sz_x;
    double *d_pred = reinterpret_cast<double *>(cur); cur += sz_x;
    double *d_W = reinterpret_cast<double *>(cur); cur += sz_W;
    double *d_r = reinterpret_cast<double *>(cur);
    // The array keeps whatever the previous call left in it.  Nothing here relies on
    // zeros: each input piece is overwritten in full before the launch, and because
    // pair_ptr starts at 0 and is non-decreasing (checked above) the pairs tile
    // [0, nnz), so the kernel writes every pred[i] before it is copied out.
    CU(cudaMemcpy(d_ptr, pair_ptr, (size_t)(n_pairs + 1) * sizeof(int64_t), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_cell, pair_cell, (size_t)n_pairs * sizeof(int32_t), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_reg, pair_region, (size_t)n_pairs * sizeof(int32_t), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_x, x, (size_t)nnz * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_W, W, (size_t)n_regions * D * K * sizeof(double), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(d_r, r_nk, (size_t)n_cells * K * sizeof(double), cudaMemcpyHostToDevice));
    BasisDev basis{};
    basis.D = D; basis.gamma = gamma;
    for (int j = 0; j < MAX_MUS; ++j) basis.mus[j] = j < n_basis ? mus[j] : 0.0;
    int sm_count = 148;
    cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, device);
    const int grid = grid_for(n_pairs, 128, sm_count, 16);
    auto run = [&]() -> int {
        DISPATCH_D(D, impute_kernel<DD><<<grid, 128>>>(d_ptr, d_cell, d_reg, d_x, d_W, d_r, d_pred, n_pairs,
                                                      n_cells, n_regions, K, basis));
        return MELISSA_OK;
    };
    if ((rc = run())) return rc;
    CU(cudaGetLastError());
    CU(cudaMemcpy(pred, d_pred, (size_t)nnz * sizeof(double), cudaMemcpyDeviceToHost));
    return MELISSA_OK;
}

// ===========================================================================
// Several GPUs behind one call (multi.cuh)
// ===========================================================================
// Fallback when the exchange buffers cannot be shared through the virtual-memory API (multi.cuh): whole-device peer
// access, switched on for as long as some run needs it and off again afterwards (while it is on, every device
// allocation of the process is also mapped into the peers).  Links that were already on when the library first
// looked (another library's doing) are left alone.
namespace {
std::mutex g_peer_mu;
int g_peer_ref[64][64];          // runs using the link dg -> dp
bool g_peer_ours[64][64];        // the library switched it on (so it switches it off)
}

struct melissa_multi {
    std::vector<std::pair<int, int>> links;    // (dg, dp) counted in g_peer_ref
    int G = 0;
    std::vector<melissa_handle *> h;
    std::vector<int64_t> cell_lo;          // first global cell of every shard (G + 1 entries)
    bool own_handles = false;
    int64_t N_total = 0;
    double last_loop_ms = 0.0;             // iteration loop of the last run: CUDA events per GPU, max over GPUs
};

// The multi-GPU entry points visit every GPU from the calling thread; the caller's current device is put back.
struct CallerDevice {
    int prev = -1;
    CallerDevice() { if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); } }
    ~CallerDevice() { if (prev >= 0) cudaSetDevice(prev); }
};

static void peer_links_release(melissa_multi *mm) {
    std::lock_guard<std::mutex> lock(g_peer_mu);
    for (const auto &l : mm->links)
        if (--g_peer_ref[l.first][l.second] == 0 && g_peer_ours[l.first][l.second]) {
            cudaSetDevice(l.first);
            cudaDeviceDisablePeerAccess(l.second);
            cudaGetLastError();
            g_peer_ours[l.first][l.second] = false;
        }
    mm->links.clear();
}

static int peer_links_acquire(melissa_multi *mm) {
    for (int g = 0; g < mm->G; ++g)
        for (int p = 0; p < mm->G; ++p) {
            const int dg = mm->h[(size_t)g]->device, dp = mm->h[(size_t)p]->device;
            if (dg == dp) continue;
            cudaError_t e = cudaSetDevice(dg);
            if (e == cudaSuccess) {
                std::lock_guard<std::mutex> lock(g_peer_mu);
                if (g_peer_ref[dg][dp] == 0) {
                    e = cudaDeviceEnablePeerAccess(dp, 0);
                    cudaGetLastError();
                    if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) {
                        g_peer_ours[dg][dp] = e == cudaSuccess;
                        e = cudaSuccess;
                    }
                }
                if (e == cudaSuccess) {
                    g_peer_ref[dg][dp]++;
                    mm->links.emplace_back(dg, dp);
                }
            }
            if (e != cudaSuccess) {
                peer_links_release(mm);
                return fail(MELISSA_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d -> %d) failed: %s", dg, dp, cudaGetErrorString(e));
            }
        }
    return MELISSA_OK;
}

int melissa_multi_create(int32_t n_gpus, melissa_handle **handles, melissa_multi **out) {
    if (!out || !handles) return fail(MELISSA_ERR_ARG, "handles / out is NULL");
    if (n_gpus < 1 || n_gpus > MULTI_MAX_GPUS) return fail(MELISSA_ERR_ARG, "n_gpus must be in 1..%d", MULTI_MAX_GPUS);
    CallerDevice caller_device;
    melissa_multi *mm = new melissa_multi();
    mm->G = n_gpus;
    mm->cell_lo.push_back(0);
    for (int g = 0; g < n_gpus; ++g) {
        melissa_handle *h = handles[g];
        if (!h) { delete mm; return fail(MELISSA_ERR_ARG, "handle %d is NULL", g); }
        if (g && (h->M != handles[0]->M || h->D != handles[0]->D)) {
            delete mm;
            return fail(MELISSA_ERR_ARG, "shard %d has another number of regions / basis functions than shard 0", g);
        }
        mm->h.push_back(h);
        mm->cell_lo.push_back(mm->cell_lo.back() + h->N);
    }
    mm->N_total = mm->cell_lo.back();
    // every GPU reads every other GPU's partial sums directly: the links must exist (they are opened per run)
    for (int g = 0; g < n_gpus; ++g) {
        CU(cudaSetDevice(mm->h[g]->device));
        for (int p = 0; p < n_gpus; ++p) {
            const int dg = mm->h[g]->device, dp = mm->h[p]->device;
            if (dg == dp) continue;
            int can = 0;
            CU(cudaDeviceCanAccessPeer(&can, dg, dp));
            if (!can) { melissa_multi_free(mm); return fail(MELISSA_ERR_CUDA, "GPU %d cannot read GPU %d's memory (no peer access)", dg, dp); }
            if (dg >= 64 || dp >= 64) { melissa_multi_free(mm); return fail(MELISSA_ERR_ARG, "device ordinal beyond 63"); }
        }
    }
    *out = mm;
    return MELISSA_OK;
}

int melissa_multi_free(melissa_multi *mm) {
    if (!mm) return MELISSA_OK;
    CallerDevice caller_device;
    HostLap tm;
    if (mm->own_handles)
        for (melissa_handle *h : mm->h) melissa_free(h);
    peer_links_release(mm);
    delete mm;
    tm.lap("free");
    return MELISSA_OK;
}

int32_t melissa_multi_gpus(const melissa_multi *mm) { return mm ? mm->G : 0; }
double melissa_multi_last_loop_ms(const melissa_multi *mm) { return mm ? mm->last_loop_ms : 0.0; }
int64_t melissa_multi_launch_count(const melissa_multi *mm) {
    int64_t n = 0;
    if (mm) for (const melissa_handle *h : mm->h) n += h->launches;
    return n;
}

int melissa_pack_multi(int32_t n_cells, int32_t n_regions, int64_t n_pairs, const int64_t *pair_ptr,
                       const int32_t *pair_cell, const int32_t *pair_region, const double *x, const uint8_t *y,
                       int32_t basis_kind, int32_t n_basis, const double *mus, double gamma, int32_t n_gpus,
                       const int32_t *devices, melissa_multi **out) {
    if (!out) return fail(MELISSA_ERR_ARG, "out is NULL");
    *out = nullptr;
    if (n_gpus < 1 || n_gpus > MULTI_MAX_GPUS) return fail(MELISSA_ERR_ARG, "n_gpus must be in 1..%d", MULTI_MAX_GPUS);
    if (n_cells < n_gpus) return fail(MELISSA_ERR_ARG, "more GPUs than cells");
    if (n_pairs <= 0 || !pair_ptr || !pair_cell || !pair_region || !x || !y) return fail(MELISSA_ERR_ARG, "empty or NULL observation arrays");
    if (n_gpus == 1) {
        melissa_handle *h = nullptr;
        int rc = melissa_pack(n_cells, n_regions, n_pairs, pair_ptr, pair_cell, pair_region, x, y, basis_kind, n_basis, mus, gamma,
                              devices ? devices[0] : 0, 0, &h);
        if (rc) return rc;
        if ((rc = melissa_multi_create(1, &h, out))) { melissa_free(h); return rc; }
        (*out)->own_handles = true;
        return MELISSA_OK;
    }
    // contiguous cell ranges balanced by CpG count (coverage varies several-fold across cells)
    HostLap tm;
    const int64_t P = n_pairs;
    if (pair_ptr[0] != 0) return fail(MELISSA_ERR_ARG, "pair_ptr[0] must be 0");
    std::vector<int64_t> cell_first((size_t)n_cells + 1);
    for (int n = 0; n <= n_cells; ++n) cell_first[(size_t)n] = std::lower_bound(pair_cell, pair_cell + P, (int32_t)n) - pair_cell;
    const int64_t nnz = pair_ptr[P];
    std::vector<int> bound((size_t)n_gpus + 1, 0);
    bound[(size_t)n_gpus] = n_cells;
    for (int g = 1; g < n_gpus; ++g) {
        const int64_t target = nnz / n_gpus * g;
        int lo = bound[(size_t)g - 1] + 1, hi = n_cells - (n_gpus - g);
        int b = lo;
        while (lo <= hi) {                                      // first cell whose CpG offset reaches the target
            const int mid = (lo + hi) / 2;
            if (pair_ptr[cell_first[(size_t)mid]] < target) lo = mid + 1; else { b = mid; hi = mid - 1; }
        }
        bound[(size_t)g] = std::max(bound[(size_t)g - 1] + 1, std::min(b, n_cells - (n_gpus - g)));
    }
    tm.lap("pack_multi: shard bounds");
    std::vector<melissa_handle *> hs((size_t)n_gpus, nullptr);
    std::vector<int> rcs((size_t)n_gpus, MELISSA_OK);
    std::vector<std::string> errs((size_t)n_gpus);
    std::vector<std::thread> th;
    for (int g = 0; g < n_gpus; ++g)
        th.emplace_back([&, g]() {
            const int c0 = bound[(size_t)g], c1 = bound[(size_t)g + 1];
            const int64_t p0 = cell_first[(size_t)c0], p1 = cell_first[(size_t)c1], i0 = pair_ptr[p0];
            // the shard is a window of the caller's lists: no host copy, the offsets are taken off on the device
            rcs[(size_t)g] = pack_impl(c1 - c0, n_regions, p1 - p0, pair_ptr + p0, pair_cell + p0, pair_region + p0, x + i0, y + i0,
                                       basis_kind, n_basis, mus, gamma, devices ? devices[g] : g, 1, false, &hs[(size_t)g], i0, c0);
            if (rcs[(size_t)g]) errs[(size_t)g] = g_err;
        });
    for (auto &t : th) t.join();
    tm.lap("pack_multi: shards packed");
    {
        // a fault in the caller's lists is reported before a CUDA failure of another shard; pair numbers in the
        // message count from the shard's first pair
        int bad = -1;
        for (int g = 0; g < n_gpus && bad < 0; ++g) if (rcs[(size_t)g] == MELISSA_ERR_ARG) bad = g;
        for (int g = 0; g < n_gpus && bad < 0; ++g) if (rcs[(size_t)g]) bad = g;
        if (bad >= 0) {
            for (melissa_handle *h : hs) melissa_free(h);
            const int c0 = bound[(size_t)bad];
            return fail(rcs[(size_t)bad], "shard %d (cells from %d, pairs from %lld): %s", bad, c0,
                        (long long)cell_first[(size_t)c0], errs[(size_t)bad].c_str());
        }
    }
    // a region no cell of ANY shard covers is an error, as in the single-GPU pack (each shard alone may miss regions)
    int rc = melissa_multi_create(n_gpus, hs.data(), out);
    if (rc) { for (melissa_handle *h : hs) melissa_free(h); return rc; }
    (*out)->own_handles = true;
    return MELISSA_OK;
}

int melissa_vb_run_multi(melissa_multi *mm, int32_t K, const double *w0, const double *delta_0, double alpha_0,
                         double beta_0, int32_t vb_max_iter, double epsilon_conv, int32_t lb_every, double *W,
                         double *W_Sigma, double *r_nk, double *delta, double *alpha, double *beta, double *pi_k,
                         double *lb, int32_t lb_cap, int32_t *n_lb, int32_t *n_iter, int32_t *flags) {
    if (!mm) return fail(MELISSA_ERR_ARG, "multi handle is NULL");
    if (vb_max_iter < 2) return fail(MELISSA_ERR_ARG, "vb_max_iter must be >= 2");
    if (lb_every <= 0) return fail(MELISSA_ERR_ARG, "lb_every must be positive");
    const int G = mm->G;
    if (G == 1 && mm->h[0]->N == mm->N_total)
        return melissa_vb_run(mm->h[0], K, w0, delta_0, alpha_0, beta_0, vb_max_iter, epsilon_conv, lb_every, W, W_Sigma, r_nk,
                              delta, alpha, beta, pi_k, lb, lb_cap, n_lb, n_iter, flags);
    struct PerGpu {
        cudaStream_t st = nullptr;
        double *part[2] = {nullptr, nullptr}, *lbpart = nullptr, *sum = nullptr, *lbsum = nullptr, *orig_red = nullptr, *orig_lb = nullptr;
        cudaEvent_t ev_part[2] = {nullptr, nullptr}, ev_lb = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
        float loop_ms = 0.f;
        int rc = MELISSA_OK;
        std::string err;
    };
    std::vector<PerGpu> pg((size_t)G);
    CallerDevice caller_device;
    HostLap tm;
    HostBarrier bar;
    bar.n = G;
    // shared between the host threads (written by thread 0 between two barriers)
    std::vector<double> lbs;
    lbs.push_back(-std::numeric_limits<double>::infinity());        // LB <- c(-Inf), :242
    int fl = 0, it_last = 1;
    std::atomic<int> stop{0};
    const melissa_progress_fn progress = g_progress;
    void *const progress_user = g_progress_user;
    // the memory the peers read: both parities of the partial sums and the bound's two partial terms, one block per GPU
    static const DrvApi drv;
    const int64_t n_red_all = (int64_t)mm->h[0]->M * K * mm->h[0]->R + K + 1, part_stride = (n_red_all + 1) & ~(int64_t)1;
    std::vector<SharedBuf> xbuf((size_t)G);
    bool shared_ok = true;
    {
        int devs[MULTI_MAX_GPUS];
        for (int g = 0; g < G; ++g) devs[g] = mm->h[(size_t)g]->device;
        for (int g = 0; g < G && shared_ok; ++g)
            shared_ok = cudaSetDevice(devs[g]) == cudaSuccess && cudaFree(nullptr) == cudaSuccess &&
                        xbuf[(size_t)g].alloc(drv, devs[g], (size_t)(2 * part_stride + 8) * sizeof(double), devs, G);
        if (!shared_ok) {
            for (SharedBuf &b : xbuf) b.free(drv);
            cudaGetLastError();
            const int rc_links = peer_links_acquire(mm);           // whole-device peer access instead
            if (rc_links) return rc_links;
        }
    }
    tm.lap("run_multi: exchange buffers");
    auto worker = [&](int g) {
        melissa_handle *h = mm->h[(size_t)g];
        PerGpu &me = pg[(size_t)g];
        auto bad = [&](int rc) { me.rc = rc; me.err = g_err; bar.fail_all(); };
#define MG(call)                                                                                         \
        do {                                                                                             \
            if ((call) != cudaSuccess) {                                                                 \
                bad(fail(MELISSA_ERR_CUDA, "GPU %d: %s failed: %s", h->device, #call, cudaGetErrorString(cudaGetLastError()))); \
                return;                                                                                  \
            }                                                                                            \
        } while (0)
        MG(cudaSetDevice(h->device));
        MG(cudaStreamCreateWithFlags(&me.st, cudaStreamNonBlocking));
        int rc = melissa_vb_begin(h, K, w0, delta_0, alpha_0, beta_0, mm->N_total, me.st);
        if (rc) { bad(rc); return; }
        const int64_t n_red = h->red_len;
        me.orig_red = h->red;
        me.orig_lb = h->lbbuf;
        if (n_red != n_red_all) { bad(fail(MELISSA_ERR_STATE, "GPU %d: unexpected length of the partial sums", h->device)); return; }
        if (shared_ok) {
            me.part[0] = reinterpret_cast<double *>(xbuf[(size_t)g].ptr);
            me.part[1] = me.part[0] + part_stride;
            me.lbpart = me.part[0] + 2 * part_stride;
            MG(cudaMemsetAsync(me.lbpart, 0, 2 * sizeof(double), me.st));
        } else {
            me.part[0] = h->red;
            me.lbpart = h->lbbuf;
            if ((rc = dev_alloc(&me.part[1], n_red))) { bad(rc); return; }
        }
        h->lbbuf = me.lbpart;
        if ((rc = dev_alloc(&me.sum, n_red)) || (rc = dev_alloc(&me.lbsum, 2))) { bad(rc); return; }
        for (int q = 0; q < 2; ++q) MG(cudaEventCreateWithFlags(&me.ev_part[q], cudaEventDisableTiming));
        MG(cudaEventCreateWithFlags(&me.ev_lb, cudaEventDisableTiming));
        MG(cudaEventCreate(&me.ev_t0));
        MG(cudaEventCreate(&me.ev_t1));
        MG(cudaStreamSynchronize(me.st));                           // set-up and the first sweep are not part of the loop
        if (!bar.wait()) return;                                    // everybody's buffers and events exist
        if (g == 0) tm.lap("run_multi: begin + 1st sweep");
        MG(cudaEventRecord(me.ev_t0, me.st));
        for (int i = 2; i <= vb_max_iter; ++i) {                    // :309
            const int par = i & 1;
            h->red = me.part[par];
            if ((rc = melissa_vb_phase_a(h))) { bad(rc); return; }  // this shard's partial sums
            MG(cudaEventRecord(me.ev_part[par], me.st));
            if (!bar.wait()) return;                                // every GPU's event of this iteration has been recorded
            PartPtrs pp;
            pp.G = G;
            for (int p = 0; p < G; ++p) {
                if (p != g) MG(cudaStreamWaitEvent(me.st, pg[(size_t)p].ev_part[par], 0));
                pp.p[p] = pg[(size_t)p].part[par];
            }
            reduce_parts_kernel<<<h->sm_count * 2, 256, 0, me.st>>>(pp, n_red, me.sum);
            h->launches++;
            h->red = me.sum;
            const bool with_lb = (i % lb_every == 0) || i == vb_max_iter;       // :458
            if ((rc = melissa_vb_phase_b(h, with_lb))) { bad(rc); return; }
            if (with_lb) {
                MG(cudaEventRecord(me.ev_lb, me.st));
                if (!bar.wait()) return;
                PartPtrs lp;
                lp.G = G;
                for (int p = 0; p < G; ++p) {
                    if (p != g) MG(cudaStreamWaitEvent(me.st, pg[(size_t)p].ev_lb, 0));
                    lp.p[p] = pg[(size_t)p].lbpart;
                }
                reduce_parts_kernel<<<1, 32, 0, me.st>>>(lp, 2, me.lbsum);
                h->launches++;
                h->lbbuf = me.lbsum;
                double v = 0.0;
                rc = melissa_vb_lower_bound(h, &v);                 // (synchronises this GPU's stream)
                h->lbbuf = me.lbpart;
                if (rc) { bad(rc); return; }
                if (!bar.wait()) return;                            // nobody overwrites lbbuf before everybody has read it
                if (g == 0) {
                    const double prev = lbs.back();
                    lbs.push_back(v);
                    if (v < prev) fl |= MELISSA_FLAG_LB_DECREASED;  // :495-497
                    if (std::fabs(v - prev) < epsilon_conv) stop.store(1);      // :499-501
                    else if (progress && progress(i, v, v - prev, progress_user)) {       // (worker 0 IS the calling thread)
                        fl |= MELISSA_FLAG_INTERRUPTED;
                        stop.store(1);
                    }
                    it_last = i;
                }
                if (!bar.wait()) return;
                if (stop.load()) break;
            }
            if (g == 0) {
                it_last = i;
                if (i == vb_max_iter) fl |= MELISSA_FLAG_NOT_CONVERGED;         // :511-513
            }
        }
        h->red = me.sum;                                            // what melissa_vb_fetch and a later caller see
        MG(cudaEventRecord(me.ev_t1, me.st));
        MG(cudaStreamSynchronize(me.st));
        MG(cudaEventElapsedTime(&me.loop_ms, me.ev_t0, me.ev_t1));
#undef MG
    };
    std::vector<std::thread> th;
    for (int g = 1; g < G; ++g) th.emplace_back(worker, g);
    worker(0);                                                      // GPU 0 is driven by the caller's own thread
    for (auto &t : th) t.join();
    tm.lap("run_multi: loop");
    int rc = MELISSA_OK;
    for (int g = 0; g < G && !rc; ++g)
        if (pg[(size_t)g].rc) rc = fail(pg[(size_t)g].rc, "%s", pg[(size_t)g].err.c_str());
    // ---- results: q(w) and the scalars are identical on every GPU, responsibilities come shard by shard
    if (!rc) {
        melissa_handle *h0 = mm->h[0];
        rc = melissa_vb_fetch(h0, W, W_Sigma, nullptr, delta, alpha, beta, pi_k);
        if (!rc && r_nk) {
            for (int g = 0; g < G && !rc; ++g) {
                melissa_handle *h = mm->h[(size_t)g];
                std::vector<double> r((size_t)h->N * K);
                if ((rc = melissa_vb_fetch(h, nullptr, nullptr, r.data(), nullptr, nullptr, nullptr, nullptr))) break;
                for (int k = 0; k < K; ++k)
                    std::memcpy(r_nk + (size_t)k * mm->N_total + (size_t)mm->cell_lo[(size_t)g], r.data() + (size_t)k * h->N,
                                (size_t)h->N * sizeof(double));
            }
        }
    }
    tm.lap("run_multi: fetch");
    // ---- put the handles back the way melissa_vb_begin left them, release the per-run resources
    for (int g = 0; g < G; ++g) {
        melissa_handle *h = mm->h[(size_t)g];
        PerGpu &me = pg[(size_t)g];
        cudaSetDevice(h->device);
        if (me.st) cudaStreamSynchronize(me.st);
        if (me.orig_red) h->red = me.orig_red;
        if (me.orig_lb) h->lbbuf = me.orig_lb;
        if (!shared_ok) cudaFree(me.part[1]);
        cudaFree(me.sum); cudaFree(me.lbsum);
        for (int q = 0; q < 2; ++q) if (me.ev_part[q]) cudaEventDestroy(me.ev_part[q]);
        if (me.ev_lb) cudaEventDestroy(me.ev_lb);
        if (me.ev_t0) cudaEventDestroy(me.ev_t0);
        if (me.ev_t1) cudaEventDestroy(me.ev_t1);
        mm->last_loop_ms = g == 0 ? (double)me.loop_ms : std::max(mm->last_loop_ms, (double)me.loop_ms);
        h->stream = nullptr;                                        // the run's stream goes away with the run
        h->begun = false;
        if (me.st) cudaStreamDestroy(me.st);
    }
    for (SharedBuf &b : xbuf) b.free(drv);
    if (!shared_ok) peer_links_release(mm);
    if (rc) return rc;
    const int nlb = (int)lbs.size();
    if (lb) for (int q = 0; q < nlb && q < lb_cap; ++q) lb[q] = lbs[(size_t)q];
    if (n_lb) *n_lb = nlb;
    if (n_iter) *n_iter = it_last;
    if (flags) *flags = fl;
    return MELISSA_OK;
}

}  // extern "C"
