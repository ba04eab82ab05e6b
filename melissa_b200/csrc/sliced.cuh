// v2 data path: "region-sliced ELL".
//
// Cells are cut into blocks of Tn; inside a block, the cells that cover region m are
// sorted by their CpG count and grouped 32 at a time into a SLICE.  A slice is one
// warp's unit of work: lane l owns pair (cell_l, m); CpG i of every lane sits in one
// 256-byte row x_sell[xbase + 32 i + l], so each load is coalesced and all lanes walk
// their CpGs in lock-step with (almost) equal trip counts.  Because the region is
// the same for the whole warp, the q(w) record U[m] is a broadcast operand.  The per-pair
// records of a slice are stored in groups of four lanes,
//     stats[slice][lane / 4][j][lane % 4]          (j = record component, R of them),
// which is exactly the B-operand tile order of the tensor-core contraction (4 pairs x 8
// components = 32 consecutive doubles), while the sweep still touches whole 32-byte
// sectors when lane l reads or writes component j of its own pair.
//
// Included from melissa.cu inside its anonymous namespace.
#pragma once

struct RbfTab {
    double inv_s, neg_s, neg_gamma, cA, cB, g2d, q2;
    double s, mu1, q;
    double magic;            // 1.5 * 2^52 s + idx0 s: x + magic has the table index in its low word and (x + magic) - magic = x0
    int n_entries, idx0;
};

constexpr int MILLS_SP = 512;                 // table points per unit of t
constexpr int MILLS_HALF = 8 * MILLS_SP;      // t in [-8, 8]
constexpr int MILLS_N = 2 * MILLS_HALF + 1;
constexpr double MAGIC = 6755399441055744.0;  // 1.5 * 2^52: (v + MAGIC) - MAGIC rounds v to an integer
// the same with the scale and the index offset folded in: t + MILLS_MAGIC has round(t * MILLS_SP) + MILLS_HALF in its low word
constexpr double MILLS_MAGIC = MAGIC / MILLS_SP + (double)MILLS_HALF / MILLS_SP;
#ifndef SWEEP3_THREADS
#define SWEEP3_THREADS 640                    // one persistent CTA per SM; 20 warps x 96 registers measured fastest (768 x 80: +3 %, 512 x 128: +-0)
#endif

struct Sliced {
    bool on = false;
    int Tn = 256, NB = 0;
    int64_t S = 0, x_len = 0, y_len = 0;
    double *x_sell = nullptr;
    uint32_t *ybits = nullptr;
    int64_t *slice_xbase = nullptr, *slice_ybase = nullptr;
    int32_t *slice_len = nullptr, *slice_region = nullptr;
    int32_t *slice_cb = nullptr;          // [S] cell block of the slice
    int32_t *lane_cell = nullptr;         // [S*32] global cell index, -1 = padding lane
    int32_t *lane_meta = nullptr;         // [S*32] CpG count | (position of the cell in its block << 16)
    int32_t *lane_pair = nullptr;         // [S*32] pair (in the caller's order) of the lane, -1 = padding lane
    int32_t *blk_cells = nullptr;         // [NB*Tn] global cell index of each block position, -1 = none
    int64_t *reg_slice_ptr = nullptr;     // [M + 1]
    int32_t *reg_slice = nullptr;         // [S]
    double *stats = nullptr;              // [S][8][R][4]
    double *mills_tab = nullptr;          // [MILLS_N]
    double2 *lphi_tab = nullptr;          // [MILLS_N] (L, L') of L(t) = log(Phi(t) + 1e-10) on the same grid
    double2 *rbf_tab = nullptr;           // [rt.n_entries] (E(u0), q B(u0))
    RbfTab rt{};
    int32_t *chunk_ptr = nullptr;         // [NC + 1] first slice of each chunk of the sweep
    int NC = 0;
    unsigned int *chunk_counter = nullptr;         // chunk dispenser of the persistent sweep
};

// R(t0) = Phi(t0) / phi(t0) on the grid t0 = (i - MILLS_HALF) / MILLS_SP.
__global__ void mills_table_kernel(double *tab) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= MILLS_N) return;
    const double t0 = (double)(i - MILLS_HALF) / (double)MILLS_SP;
    const double v = fabs(t0) * 0.70710678118654752440;
    double R = 1.25331413731550025121 * erfcx(v);
    if (t0 > 0.0) R = 2.50662827463100050242 * exp(0.5 * t0 * t0) - R;
    tab[i] = R;
}

// L(t0) = log(Phi(t0) + 1e-10) and q(t0) = L'(t0) = phi / (Phi + 1e-10) on the same grid.  q obeys q' = -q (t + q)
// whatever the constant next to Phi is, so the sweep gets every further Taylor coefficient of L from these two.
__global__ void lphi_table_kernel(double2 *tab) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= MILLS_N) return;
    const double t0 = (double)(i - MILLS_HALF) / (double)MILLS_SP;
    const double Phi = 0.5 * erfc(-t0 * 0.70710678118654752440);
    const double phi = 0.39894228040143267794 * exp(-0.5 * t0 * t0);
    tab[i] = make_double2(log(Phi + 1e-10), phi / (Phi + 1e-10));
}

// Series coefficients that are not exactly representable in a 32-bit immediate live in
// constant memory: a DFMA can take one operand straight from the constant bank, whereas a
// literal costs two register moves every time the (register-starved) loop needs it again.
__constant__ double kInvIntC[7] = {0.0, 1.0, 0.5, 1.0 / 3, 0.25, 0.2, 1.0 / 6};
#define kInvInt(i) kInvIntC[i]
__constant__ double kInvFact3[10] = {1.0, 1.0, 0.5, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040, 1.0 / 40320,
                                     1.0 / 362880};

// log(Phi(t) + 1e-10) at t = t0 + dt, |dt| <= 1 / (2 MILLS_SP), from the table entry (L0, a0) of t0:
//   L = L0 + a0 dt + a1 dt^2 / 2 + a2 dt^3 / 3 + a3 dt^4 / 4,   (n + 1) a_{n+1} = -(t0 a_n + a_{n-1}) - sum_i a_i a_{n-i}
// (the Taylor coefficients of q = L' about t0).  The first term left out is a4 dt^5 / 5 with dt^5 / 5 <= 1.8e-16.
__device__ __forceinline__ double log_phi_eps(double t0, double dt, double2 e) {
    const double a0 = e.y;
    const double s2 = fma(2.0, a0, t0);
    const double a1 = -a0 * (t0 + a0);
    const double a2 = -0.5 * fma(a1, s2, a0);
    const double a3 = -kInvInt(3) * fma(a2, s2, fma(a1, a1, a1));
    double p = fma(dt, 0.25 * a3, kInvInt(3) * a2);
    p = fma(dt, p, 0.5 * a1);
    p = fma(dt, p, a0);
    return fma(dt, p, e.x);
}


template <int N>
__device__ __forceinline__ double exp_series(double w) {         // sum_{i <= N} w^i / i!
    double c = fma(w, kInvFact3[N], kInvFact3[N - 1]);
#pragma unroll
    for (int i = N - 2; i >= 2; --i) c = fma(w, c, kInvFact3[i]);
    c = fma(w, c, 1.0);
    return fma(w, c, 1.0);
}

// RBF row from the position x in [-1, 1], equally spaced centres mu_j = mu_1 + (j - 1) Delta:
//   h_1 = exp(-gamma u^2),  h_{j+1} = h_j * b * q^(2j-1),  u = x - mu_1,  b = exp(2 gamma Delta u),
//   q = exp(-gamma Delta^2).
// The table sits on the grid x0 = j s, s a power of two (so x - x0 is exact), and holds
// E = exp(-gamma u0^2) and B = q exp(2 gamma Delta u0); the remainder d = x - x0 enters through
// exp(-gamma d (2 u0 + d)) and exp(2 gamma Delta d) as short series.  s is chosen by the packer
// so that the truncation error stays below 3e-16 (rbf_table_log2 in melissa.cu).
// Positions outside the table are reported through `bad` (the caller then uses exp()).
template <int D>
__device__ __forceinline__ void rbf_row3(double x, const RbfTab &rt, const double2 *__restrict__ tab,
                                         double (&h)[D], unsigned &bad) {
    constexpr int NW = D <= 9 ? 6 : 8;
    h[0] = 1.0;
    if (D == 1) return;
    const double v = x + rt.magic;                                         // low word = table index
    const unsigned j = (unsigned)__double2loint(v);
    const unsigned jc = min(j, (unsigned)(rt.n_entries - 1));
    bad |= j ^ jc;
    const double x0 = v - rt.magic;
    const double d = x - x0;
    const double2 eb = tab[jc];
    const double w = d * fma(rt.neg_gamma, d, fma(x0, rt.cA, rt.cB));      // -gamma d (2 u0 + d), u0 = x0 - mu_1
    h[1] = eb.x * exp_series<NW>(w);
    if (D > 2) {
        double bj = eb.y * exp_series<NW - 1>(rt.g2d * d);
#pragma unroll
        for (int q = 2; q < D; ++q) {
            h[q] = h[q - 1] * bj;
            if (q + 1 < D) bj *= rt.q2;
        }
    }
}

// E[z | z > 0], z ~ N(t, 1) = t + 1 / R(t), R = Phi / phi.  R' = 1 + t R, so the Taylor
// coefficients about a grid point follow from the table value alone:
//   c_0 = R0, c_1 = 1 + t0 R0, c_{n+1} = (c_{n-1} + t0 c_n) / (n + 1).
// Spacing 1/512, order 4: indistinguishable from order 5 at 1/128 and from the erfcx closed form
// against mpmath (tests/test_table_functions.py).  The reciprocal is MUFU.RCP64H (2^-23) plus one
// cubically convergent step (2^-69).  |t| beyond the table is reported through `bad`.
__device__ __forceinline__ double trunc_mean3(double t, const double *__restrict__ tab, unsigned &bad, double &R_out) {
    const double v = t + MILLS_MAGIC;                                      // low word = table index
    const unsigned i = (unsigned)__double2loint(v);
    const unsigned ic = min(i, 2u * MILLS_HALF);
    bad |= i ^ ic;
    const double t0 = v - MILLS_MAGIC;
    const double d = t - t0;
    const double R0 = tab[ic];
    const double c1 = fma(t0, R0, 1.0);
    const double c2 = 0.5 * fma(t0, c1, R0);
    const double c3 = fma(t0, c2, c1) * kInvInt(3);
    const double c4 = 0.25 * fma(t0, c3, c2);
    double R = fma(d, c4, c3);
    R = fma(d, R, c2);
    R = fma(d, R, c1);
    R = fma(d, R, R0);
    R_out = R;
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(R));
    const double e = fma(-R, y, 1.0);
    y = fma(y, fma(e, e, e), y);
    return t + y;
}

// Re-lay x / y into slices.  One WARP per slice, 32 rows at a time through a shared-memory tile: the
// flat arrays are read in runs (the CpGs of one pair are contiguous), the slice is written row by row
// (256 contiguous bytes), so both sides of the transposition are coalesced.
__global__ void __launch_bounds__(128)
sell_fill_kernel(const double *__restrict__ x, const uint8_t *__restrict__ y, const int64_t *__restrict__ lane_src,
                 const int32_t *__restrict__ lane_meta, const int64_t *__restrict__ slice_xbase,
                 const int64_t *__restrict__ slice_ybase, const int32_t *__restrict__ slice_len,
                 double *__restrict__ x_sell, uint32_t *__restrict__ ybits, int64_t S) {
    __shared__ double tile[4][32][33];                     // [warp][pair][row], padded against bank conflicts
    __shared__ uint32_t ytile[4][32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int64_t s = (int64_t)blockIdx.x * 4 + warp; s < S; s += (int64_t)gridDim.x * 4) {
        const int L = slice_len[s];
        const int my_len = lane_meta[s * 32 + lane] & 0xffff;
        const int64_t my_src = lane_src[s * 32 + lane], xb = slice_xbase[s], yb = slice_ybase[s];
        for (int i0 = 0; i0 < L; i0 += 32) {
            // rows i0 .. i0 + 31 of every pair of the slice: lane = row while reading, lane = pair while writing
            // (eight pairs' loads are issued before the first of them is used: the runs are short and scattered,
            // so it is the number of requests in flight that decides the speed)
            const int i = i0 + lane;
            for (int l0 = 0; l0 < 32; l0 += 8) {
                double xv[8];
                uint8_t yv[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int len = __shfl_sync(0xffffffffu, my_len, l0 + u);
                    const int64_t src = __shfl_sync(0xffffffffu, my_src, l0 + u);
                    xv[u] = 0.0;
                    yv[u] = 0;
                    if (i < len) {
                        xv[u] = __ldg(x + src + i);
                        yv[u] = __ldg(y + src + i);
                    }
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    tile[warp][l0 + u][lane] = xv[u];
                    const uint32_t word = __ballot_sync(0xffffffffu, yv[u] != 0);
                    if (lane == 0) ytile[warp][l0 + u] = word;
                }
            }
            __syncwarp();
            const int rows = min(32, L - i0);
            for (int r = 0; r < rows; ++r) x_sell[xb + (int64_t)(i0 + r) * 32 + lane] = tile[warp][lane][r];
            ybits[yb + (int64_t)(i0 >> 5) * 32 + lane] = ytile[warp][lane];
            __syncwarp();
        }
    }
}

// H^T H per pair into stats[slice][D + t][lane]; g rows zeroed.  One thread per (slice, lane).
template <int D>
__global__ void hh_sliced_kernel(const double *__restrict__ x_sell, const int64_t *__restrict__ slice_xbase,
                                 const int32_t *__restrict__ lane_meta, double *__restrict__ stats, int64_t S,
                                 BasisDev basis, RbfTab rt, const double2 *__restrict__ rbf_tab) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S * 32) return;
    const int64_t s = t >> 5;
    const int lane = (int)(t & 31);
    const int len = lane_meta[t] & 0xffff;
    const int64_t xb = slice_xbase[s];
    double hh[T];
#pragma unroll
    for (int q = 0; q < T; ++q) hh[q] = 0.0;
    for (int i = 0; i < len; ++i) {
        // the same table-driven row the sweep uses (the table sits in L1 here); exp() for a position off the table
        double h[D];
        const double x = x_sell[xb + (int64_t)i * 32 + lane];
        unsigned bad = 0;
        rbf_row3<D>(x, rt, rbf_tab, h, bad);
        if (bad) eval_basis<D>(x, basis, h);
        int q = 0;
#pragma unroll
        for (int a = 0; a < D; ++a)
#pragma unroll
            for (int c = a; c < D; ++c) hh[q++] += h[a] * h[c];
    }
    double *rec = stats + s * R * 32 + (lane >> 2) * (R * 4) + (lane & 3);      // see stats layout above
#pragma unroll
    for (int d = 0; d < D; ++d) rec[d * 4] = 0.0;
#pragma unroll
    for (int q = 0; q < T; ++q) rec[(D + q) * 4] = hh[q];
}

// Rare path of the sweep (some x outside the table, or |s mu| >= 8, in some lane of the slice):
// the whole slice is redone with plain exp / erfcx evaluation, after and outside the row loop.
// (Inlined: an out-of-line call anywhere in the kernel makes ptxas keep the hot loop's state in
// ABI registers and spill around it -- measured 1.1 -> 1.6 ms.)
template <int D>
struct SliceSums {
    double g[D];
    double lb, wrong;
};
template <int D>
struct WBar {
    double w[D];
};
#ifndef SLICE_SLOW_NOINLINE
#define SLICE_SLOW_NOINLINE 0
#endif
#if SLICE_SLOW_NOINLINE
#define SLICE_SLOW_ATTR __noinline__
#else
#define SLICE_SLOW_ATTR __forceinline__
#endif
template <int D, bool WITH_LB>
__device__ SLICE_SLOW_ATTR SliceSums<D> sweep_slice_slow(const double *xrow, const uint32_t *yrow, int L, int len,
                                                      WBar<D> wb, const BasisDev *basis) {
    SliceSums<D> o;
#pragma unroll
    for (int d = 0; d < D; ++d) o.g[d] = 0.0;
    o.lb = 0.0;
    o.wrong = 0.0;
    for (int i = 0; i < len && i < L; ++i) {
        const bool yi = (yrow[(int64_t)(i >> 5) * 32] >> (i & 31)) & 1u;
        double h[D];
        eval_basis<D>(xrow[(int64_t)i * 32], *basis, h);
        double mu = wb.w[0];
#pragma unroll
        for (int d = 1; d < D; ++d) mu = fma(h[d], wb.w[d], mu);
        const double t = yi ? mu : -mu;
        const double G = trunc_mean(t);
        const double Ez = yi ? G : -G;
#pragma unroll
        for (int d = 0; d < D; ++d) o.g[d] = fma(h[d], Ez, o.g[d]);
        if (WITH_LB) {
            o.lb += 0.5 * mu * mu + loglik_term(mu, yi);
            if (t < -5.0) o.wrong += 1.0;
        }
    }
    return o;
}

// Contraction on slices: acc[m,k,j] = sum over the region's slices and lanes of
// r[cell,k] * stats[slice][j][lane]  (R/melissa_vb.R:368-373 sums).  One CTA per region;
// warp w takes the (component chunk, cluster chunk) combinations w, w + nw, ...; lanes
// are pairs, JC x KC accumulators per lane, one butterfly at the end.  The slice ids of
// the next 32 slices are fetched with one coalesced load and handed round by shuffle,
// and two slices are in flight per loop trip, so the dependent index -> cell -> r loads
// of one slice overlap the row loads of the other.
template <int JC, int KC>
__global__ void __launch_bounds__(256)
contract_v2_kernel(const int64_t *__restrict__ reg_slice_ptr, const int32_t *__restrict__ reg_slice,
                   const int32_t *__restrict__ lane_cell, const double *__restrict__ stats,
                   const double *__restrict__ r, double *__restrict__ acc_out, int K, int R) {
    const int m = blockIdx.x;
    const int nw = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t q0 = reg_slice_ptr[m], q1 = reg_slice_ptr[m + 1];
    const int njc = (R + JC - 1) / JC, nkc = (K + KC - 1) / KC;
    for (int combo = warp; combo < njc * nkc; combo += nw) {
        const int j0 = (combo % njc) * JC, k0 = (combo / njc) * KC;
        double acc[JC][KC];
#pragma unroll
        for (int j = 0; j < JC; ++j)
#pragma unroll
            for (int c = 0; c < KC; ++c) acc[j][c] = 0.0;
        for (int64_t base = q0; base < q1; base += 32) {
            const int cnt = (int)min((int64_t)32, q1 - base);
            const int my_s = lane < cnt ? reg_slice[base + lane] : 0;
            for (int it = 0; it < cnt; it += 2) {
                const int64_t sa = __shfl_sync(0xffffffffu, my_s, it);
                const bool has_b = it + 1 < cnt;
                const int64_t sb = __shfl_sync(0xffffffffu, my_s, has_b ? it + 1 : it);
                const int cell_a = lane_cell[sa * 32 + lane];
                const int cell_b = has_b ? lane_cell[sb * 32 + lane] : -1;
                const double *rec_a = stats + sa * ((int64_t)R * 32) + (lane >> 2) * (R * 4) + (lane & 3);
                const double *rec_b = stats + sb * ((int64_t)R * 32) + (lane >> 2) * (R * 4) + (lane & 3);
                double va[JC], vb[JC], ra[KC], rb[KC];
#pragma unroll
                for (int j = 0; j < JC; ++j) {
                    va[j] = (j0 + j < R) ? rec_a[(j0 + j) * 4] : 0.0;
                    vb[j] = (has_b && j0 + j < R) ? rec_b[(j0 + j) * 4] : 0.0;
                }
#pragma unroll
                for (int c = 0; c < KC; ++c) {
                    ra[c] = (cell_a >= 0 && k0 + c < K) ? r[(int64_t)cell_a * K + k0 + c] : 0.0;
                    rb[c] = (cell_b >= 0 && k0 + c < K) ? r[(int64_t)cell_b * K + k0 + c] : 0.0;
                }
#pragma unroll
                for (int j = 0; j < JC; ++j)
#pragma unroll
                    for (int c = 0; c < KC; ++c) acc[j][c] = fma(ra[c], va[j], fma(rb[c], vb[j], acc[j][c]));
            }
        }
#pragma unroll
        for (int j = 0; j < JC; ++j)
#pragma unroll
            for (int c = 0; c < KC; ++c) {
                const double v = warp_sum(acc[j][c]);
                if (lane == 0 && j0 + j < R && k0 + c < K) acc_out[((int64_t)m * K + k0 + c) * R + j0 + j] = v;
            }
    }
}

// ---------------------------------------------------------------------------
// Contraction on FP64 tensor cores.  For one region the r-weighted statistics are a
// small GEMM with a long inner dimension,
//     acc[K x R] = r^T [K x pairs] * stats [pairs x R],
// so each warp walks its share of the region's slices and feeds them to
// mma.sync.m8n8k4.f64 four pairs at a time: A = r^T tile (8 clusters x 4 pairs), B =
// stats tile (4 pairs x 8 record components), C accumulates in 2 registers per tile.
// Every stats row is read exactly once, the reduction over pairs happens inside the MMA,
// and the register footprint (MT*NT*2 doubles) does not grow with K*R the way the
// per-lane accumulators of contract_v2_kernel do.
// ---------------------------------------------------------------------------
__device__ __forceinline__ void dmma_m8n8k4(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

#ifndef CMMA_UNROLL
#define CMMA_UNROLL 8
#endif
#ifndef CMMA_MINB
#define CMMA_MINB 4
#endif
#define CMMA_STR2(x) #x
#define CMMA_STR(x) CMMA_STR2(x)
#define CMMA_PRAGMA_UNROLL(n) _Pragma(CMMA_STR(unroll n))
template <int MT, int NT>
__global__ void __launch_bounds__(256, CMMA_MINB)
contract_mma_kernel(const int64_t *__restrict__ reg_slice_ptr, const int32_t *__restrict__ reg_slice,
                    const int32_t *__restrict__ lane_cell, const double *__restrict__ stats,
                    const double *__restrict__ r, double *__restrict__ acc_out, int K, int R) {
    extern __shared__ double s_red[];                 // [nw][MT*NT*64]
    const int m = blockIdx.x;
    const int nw = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int quad = lane >> 2, tid4 = lane & 3;
    const int64_t q0 = reg_slice_ptr[m], q1 = reg_slice_ptr[m + 1];
    double c[MT][NT][2];
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) c[mt][nt][0] = c[mt][nt][1] = 0.0;
    // warp w takes slices q0 + w, q0 + w + nw, ...; their ids are fetched 32 at a time (one per lane)
    for (int64_t base = q0 + warp; base < q1; base += (int64_t)nw * 32) {
        const int64_t mine_q = base + (int64_t)lane * nw;
        const int my_s = mine_q < q1 ? reg_slice[mine_q] : 0;
        const int cnt = (int)min((int64_t)32, (q1 - base + nw - 1) / nw);
        for (int it = 0; it < cnt; ++it) {
            const int64_t s = __shfl_sync(0xffffffffu, my_s, it);
            const int my_cell = lane_cell[s * 32 + lane];
            const double *rec = stats + s * ((int64_t)R * 32);
CMMA_PRAGMA_UNROLL(CMMA_UNROLL)
            for (int ks = 0; ks < 8; ++ks) {
                const int pair = 4 * ks + tid4;
                const int cell = __shfl_sync(0xffffffffu, my_cell, pair);
                double a[MT], b[NT];
#pragma unroll
                for (int mt = 0; mt < MT; ++mt) {
                    const int cl = mt * 8 + quad;
                    a[mt] = (cell >= 0 && cl < K) ? r[(int64_t)cell * K + cl] : 0.0;
                }
#pragma unroll
                for (int nt = 0; nt < NT; ++nt) {
                    const int j = nt * 8 + quad;
                    b[nt] = j < R ? rec[(ks * R + j) * 4 + tid4] : 0.0;   // 32 consecutive doubles per tile
                }
#pragma unroll
                for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                    for (int nt = 0; nt < NT; ++nt) dmma_m8n8k4(c[mt][nt][0], c[mt][nt][1], a[mt], b[nt]);
            }
        }
    }
    // C fragment: lane holds rows quad (cluster), columns 2*tid4 + {0,1} (component) of each tile
    double *mine = s_red + (size_t)warp * (MT * NT * 64);
#pragma unroll
    for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
            mine[(mt * NT + nt) * 64 + quad * 8 + 2 * tid4] = c[mt][nt][0];
            mine[(mt * NT + nt) * 64 + quad * 8 + 2 * tid4 + 1] = c[mt][nt][1];
        }
    __syncthreads();
    for (int i = threadIdx.x; i < MT * NT * 64; i += blockDim.x) {
        double v = 0.0;
        for (int w = 0; w < nw; ++w) v += s_red[(size_t)w * (MT * NT * 64) + i];
        const int tile = i >> 6, row = (i >> 3) & 7, col = i & 7;
        const int cl = (tile / NT) * 8 + row, j = (tile % NT) * 8 + col;
        if (cl < K && j < R) acc_out[((int64_t)m * K + cl) * R + j] = v;
    }
}

// ---------------------------------------------------------------------------
// Per-pair penalised probit MLE (the role of BPRMeth::infer_profiles_mle(model = "bernoulli",
// lambda, opt_itnmax) in the default initialisation, reference R/melissa_vb.R:136-145): Newton
// iterations on  sum_i log Phi(s_i h_i.w) - lambda |w|^2  for every covered (cell, region) pair.
// One warp per slice, one lane per pair; gradient and Hessian are accumulated over the pair's
// CpGs with the same table-driven basis / Mills-ratio code as the sweep, then a D x D Cholesky
// solve per lane.  out[(slice * 32 + lane) * D + d]; padding lanes write zeros.
// ---------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(256)
mle_kernel(const double *__restrict__ x_sell, const uint32_t *__restrict__ ybits,
           const int64_t *__restrict__ slice_xbase, const int64_t *__restrict__ slice_ybase,
           const int32_t *__restrict__ slice_len, const int32_t *__restrict__ lane_meta,
           const double *__restrict__ mills_tab, const double2 *__restrict__ rbf_tab, RbfTab rt, BasisDev basis,
           double *__restrict__ out, int64_t S, double lambda, int n_iter) {
    constexpr int T = D * (D + 1) / 2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2 *s_rbf = reinterpret_cast<double2 *>(smem_raw);
    double *s_mills = reinterpret_cast<double *>(s_rbf + rt.n_entries);
    for (int i = threadIdx.x; i < rt.n_entries; i += blockDim.x) s_rbf[i] = rbf_tab[i];
    for (int i = threadIdx.x; i < MILLS_N; i += blockDim.x) s_mills[i] = mills_tab[i];
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int64_t warp0 = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
    for (int64_t s = warp0; s < S; s += nwarps) {
        const int L = slice_len[s];
        const int len = lane_meta[s * 32 + lane] & 0xffff;
        const double *xrow = x_sell + slice_xbase[s] + lane;
        const uint32_t *yrow = ybits + slice_ybase[s] + lane;
        double w[D];
#pragma unroll
        for (int d = 0; d < D; ++d) w[d] = 0.0;
        for (int it = 0; it < n_iter; ++it) {
            double g[D], A[T];
#pragma unroll
            for (int d = 0; d < D; ++d) g[d] = 0.0;
#pragma unroll
            for (int q = 0; q < T; ++q) A[q] = 0.0;
            uint32_t yw = 0;
            for (int i = 0; i < L; ++i) {
                if ((i & 31) == 0) yw = yrow[(int64_t)(i >> 5) * 32];
                const bool yi = (yw >> (i & 31)) & 1u;
                double h[D];
                unsigned bad = 0;
                const double xi = xrow[(int64_t)i * 32];
                rbf_row3<D>(xi, rt, s_rbf, h, bad);
                if (bad) eval_basis<D>(xi, basis, h);                // position outside the table
                double eta = w[0];
#pragma unroll
                for (int d = 1; d < D; ++d) eta = fma(h[d], w[d], eta);
                const double t = yi ? eta : -eta;
                double Rt;
                bad = 0;
                double G = trunc_mean3(t, s_mills, bad, Rt);     // t + phi/Phi
                if (bad) G = trunc_mean(t);                      // |t| beyond the table
                double lam = G - t;                              // phi(t) / Phi(t)
                double hess = lam * G;                           // -(d^2/d eta^2) log Phi(s eta)
                if (i >= len) { lam = 0.0; hess = 0.0; }
                const double gr = yi ? lam : -lam;
                int q = 0;
#pragma unroll
                for (int a2 = 0; a2 < D; ++a2) {
                    g[a2] = fma(h[a2], gr, g[a2]);
                    const double ha = h[a2] * hess;
#pragma unroll
                    for (int c2 = a2; c2 < D; ++c2) {
                        A[q] = fma(ha, h[c2], A[q]);
                        ++q;
                    }
                }
            }
            // Newton step: (A + 2 lambda I) step = g - 2 lambda w
            double Lm[D][D];
            {
                int q = 0;
#pragma unroll
                for (int a2 = 0; a2 < D; ++a2) {
                    g[a2] -= 2.0 * lambda * w[a2];
#pragma unroll
                    for (int c2 = a2; c2 < D; ++c2) {
                        Lm[c2][a2] = A[q] + (a2 == c2 ? 2.0 * lambda : 0.0);
                        ++q;
                    }
                }
            }
#pragma unroll
            for (int j = 0; j < D; ++j) {
                double dd = Lm[j][j];
#pragma unroll
                for (int c2 = 0; c2 < j; ++c2) dd -= Lm[j][c2] * Lm[j][c2];
                const double ljj = sqrt(dd), inv = 1.0 / ljj;
                Lm[j][j] = ljj;
#pragma unroll
                for (int i2 = j + 1; i2 < D; ++i2) {
                    double sacc = Lm[i2][j];
#pragma unroll
                    for (int c2 = 0; c2 < j; ++c2) sacc -= Lm[i2][c2] * Lm[j][c2];
                    Lm[i2][j] = sacc * inv;
                }
            }
            double z[D];
#pragma unroll
            for (int i2 = 0; i2 < D; ++i2) {                     // forward
                double sacc = g[i2];
#pragma unroll
                for (int c2 = 0; c2 < i2; ++c2) sacc -= Lm[i2][c2] * z[c2];
                z[i2] = sacc / Lm[i2][i2];
            }
#pragma unroll
            for (int i2 = D - 1; i2 >= 0; --i2) {                // backward
                double sacc = z[i2];
#pragma unroll
                for (int c2 = i2 + 1; c2 < D; ++c2) sacc -= Lm[c2][i2] * z[c2];
                z[i2] = sacc / Lm[i2][i2];
            }
#pragma unroll
            for (int d = 0; d < D; ++d) w[d] += z[d];
        }
#pragma unroll
        for (int d = 0; d < D; ++d) out[(s * 32 + lane) * D + d] = len > 0 ? w[d] : 0.0;
    }
}
