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
    double mu1, s, inv_s, neg_gamma, g2d, q, q2;
    int n_entries, idx0;
};

constexpr int MILLS_SP = 128;                 // table points per unit of t
constexpr int MILLS_HALF = 8 * MILLS_SP;      // t in [-8, 8]
constexpr int MILLS_N = 2 * MILLS_HALF + 1;
#ifndef SWEEP_XU
#define SWEEP_XU 2
#endif
#ifndef SWEEP_MINB
#define SWEEP_MINB 3
#endif
#ifndef SWEEP_STAGE_U
#define SWEEP_STAGE_U 1     // 1: copy U[m] to shared memory per slice; 0: read it through L1 (uniform loads)
#endif
constexpr int RBF_SUB = 128;                  // table points per centre spacing
constexpr double MAGIC = 6755399441055744.0;  // 1.5 * 2^52: (v + MAGIC) - MAGIC rounds v to an integer

struct Sliced {
    bool on = false;
    int Tn = 128, NB = 0;
    int64_t S = 0, x_len = 0, y_len = 0;
    double *x_sell = nullptr;
    uint32_t *ybits = nullptr;
    int64_t *slice_xbase = nullptr, *slice_ybase = nullptr;
    int32_t *slice_len = nullptr, *slice_region = nullptr;
    int32_t *slice_cb = nullptr;          // [S] cell block of the slice
    int32_t *lane_cell = nullptr;         // [S*32] global cell index, -1 = padding lane
    int32_t *lane_meta = nullptr;         // [S*32] CpG count | (position of the cell in its block << 16)
    int32_t *blk_cells = nullptr;         // [NB*Tn] global cell index of each block position, -1 = none
    int64_t *reg_slice_ptr = nullptr;     // [M + 1]
    int32_t *reg_slice = nullptr;         // [S]
    double *stats = nullptr;              // [S][8][R][4]
    double *mills_tab = nullptr;          // [MILLS_N]
    double2 *rbf_tab = nullptr;           // [rt.n_entries] (E(u0), q B(u0))
    RbfTab rt{};
    unsigned long long *slice_counter = nullptr;   // slice dispenser of the persistent sweep
    std::vector<std::vector<int32_t>> lane_pair;   // host, per build piece: original pair of (slice, lane), -1 = padding
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

// E[z | z > 0], z ~ N(t, 1), from the table: R = Phi/phi satisfies R' = 1 + t R, so all
// Taylor coefficients about a grid point follow from R(t0) alone:
//   c_0 = R0, c_1 = 1 + t0 R0, c_{n+1} = (c_{n-1} + t0 c_n) / (n + 1).
// Order 5 at spacing 1/128 is at the rounding floor: against mpmath it is indistinguishable from order 6
// and from the erfcx closed form (order 4 would lose a digit: 5e-14); tests/test_table_functions.py.
__device__ __forceinline__ double trunc_mean_tab(double t, const double *__restrict__ tab) {
    if (!(fabs(t) < 7.99)) return trunc_mean(t);
    const double v = fma(t, (double)MILLS_SP, MAGIC);
    const int i = __double2loint(v);
    const double t0 = (v - MAGIC) * (1.0 / MILLS_SP);
    const double d = t - t0;
    const double R0 = tab[i + MILLS_HALF];
    const double c1 = fma(t0, R0, 1.0);
    const double c2 = 0.5 * fma(t0, c1, R0);
    const double c3 = (1.0 / 3.0) * fma(t0, c2, c1);
    const double c4 = 0.25 * fma(t0, c3, c2);
    const double c5 = 0.2 * fma(t0, c4, c3);
    double R = fma(d, c5, c4);
    R = fma(d, R, c3);
    R = fma(d, R, c2);
    R = fma(d, R, c1);
    R = fma(d, R, R0);
    return t + __drcp_rn(R);
}

// RBF row from the position, equally spaced centres mu_j = mu_1 + (j-1) Delta:
//   h_1 = exp(-gamma u^2),  h_{j+1} = h_j * b * q^(2j-1),  u = x - mu_1,   (the table stores q b)
//   b = exp(2 gamma Delta u), q = exp(-gamma Delta^2).
// exp(-gamma u^2) and b come from a table at u0 = k s (s = Delta / RBF_SUB) times short
// series in the remainder d = u - u0 (|arguments| < 0.05).
template <int D>
__device__ __forceinline__ void eval_basis_tab(double x, const RbfTab &rt, const double2 *__restrict__ tab,
                                               const BasisDev &basis, double (&h)[D]) {
    h[0] = 1.0;
    if (D == 1) return;
    const double u = x - rt.mu1;
    const double v = fma(u, rt.inv_s, MAGIC);
    const int idx = __double2loint(v) + rt.idx0;
    const double u0 = (v - MAGIC) * rt.s;
    const double d = u - u0;
    if ((unsigned)idx >= (unsigned)rt.n_entries) {           // x outside [-1, 1]: plain evaluation
        eval_basis<D>(x, basis, h);
        return;
    }
    const double2 eb = tab[idx];
    const double w = (rt.neg_gamma * d) * fma(2.0, u0, d);   // -gamma (2 u0 d + d^2)
    double c = fma(w, 1.0 / 5040, 1.0 / 720);
    c = fma(w, c, 1.0 / 120);
    c = fma(w, c, 1.0 / 24);
    c = fma(w, c, 1.0 / 6);
    c = fma(w, c, 0.5);
    c = fma(w, c, 1.0);
    c = fma(w, c, 1.0);
    h[1] = eb.x * c;
    if (D > 2) {
        const double z = rt.g2d * d;
        double e = fma(z, 1.0 / 720, 1.0 / 120);
        e = fma(z, e, 1.0 / 24);
        e = fma(z, e, 1.0 / 6);
        e = fma(z, e, 0.5);
        e = fma(z, e, 1.0);
        e = fma(z, e, 1.0);
        double bj = eb.y * e;                          // the table's second entry carries the factor q
#pragma unroll
        for (int j = 2; j < D; ++j) {
            h[j] = h[j - 1] * bj;
            bj *= rt.q2;
        }
    }
}

// Series coefficients that are not exactly representable in a 32-bit immediate live in
// constant memory: a DFMA can take one operand straight from the constant bank, whereas a
// literal costs two register moves every time the (register-starved) loop needs it again.
#ifndef COEF_CONST
#define COEF_CONST 1
#endif
#if COEF_CONST
__constant__ double kInvFactC[8] = {1.0, 1.0, 0.5, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040};
__constant__ double kInvIntC[7] = {0.0, 1.0, 0.5, 1.0 / 3, 0.25, 0.2, 1.0 / 6};
#define kInvFact(i) kInvFactC[i]
#define kInvInt(i) kInvIntC[i]
#else
__host__ __device__ constexpr double inv_fact_lit(int i) { return i <= 1 ? 1.0 : inv_fact_lit(i - 1) / i; }
#define kInvFact(i) inv_fact_lit(i)
#define kInvInt(i) (1.0 / (i))
#endif

// Branch-free twins of the two table functions for the sweep's inner loop: the table
// index is clamped instead of tested, and the (never expected) out-of-range case is
// reported through the return value so the caller can redo that CpG on the slow path.
// Straight-line code lets the scheduler interleave the CpGs of one prefetch group.
template <int D>
__device__ __forceinline__ bool eval_basis_fast(double x, const RbfTab &rt, const double2 *__restrict__ tab,
                                                double (&h)[D]) {
    h[0] = 1.0;
    if (D == 1) return false;
    const double u = x - rt.mu1;
    const double v = fma(u, rt.inv_s, MAGIC);
    int idx = __double2loint(v) + rt.idx0;
    const bool bad = (unsigned)idx >= (unsigned)rt.n_entries;
    idx = min(max(idx, 0), rt.n_entries - 1);
    const double u0 = (v - MAGIC) * rt.s;
    const double d = u - u0;
    const double2 eb = tab[idx];
    const double w = (rt.neg_gamma * d) * fma(2.0, u0, d);
    double c = fma(w, kInvFact(7), kInvFact(6));
    c = fma(w, c, kInvFact(5));
    c = fma(w, c, kInvFact(4));
    c = fma(w, c, kInvFact(3));
    c = fma(w, c, 0.5);
    c = fma(w, c, 1.0);
    c = fma(w, c, 1.0);
    h[1] = eb.x * c;
    if (D > 2) {
        const double z = rt.g2d * d;
        double e = fma(z, kInvFact(6), kInvFact(5));
        e = fma(z, e, kInvFact(4));
        e = fma(z, e, kInvFact(3));
        e = fma(z, e, 0.5);
        e = fma(z, e, 1.0);
        e = fma(z, e, 1.0);
        double bj = eb.y * e;                          // the table's second entry carries the factor q
#pragma unroll
        for (int j = 2; j < D; ++j) {
            h[j] = h[j - 1] * bj;
            bj *= rt.q2;
        }
    }
    return bad;
}

__device__ __forceinline__ double trunc_mean_fast(double t, const double *__restrict__ tab, bool &bad, double &R_out) {
    bad = !(fabs(t) < 7.99);
    const double v = fma(t, (double)MILLS_SP, MAGIC);
    const int i = min(max(__double2loint(v), -MILLS_HALF), MILLS_HALF);
    const double t0 = (v - MAGIC) * (1.0 / MILLS_SP);
    const double d = t - t0;
    const double R0 = tab[i + MILLS_HALF];
    const double c1 = fma(t0, R0, 1.0);
    const double c2 = 0.5 * fma(t0, c1, R0);
    const double c3 = fma(t0, c2, c1) * kInvInt(3);
    const double c4 = 0.25 * fma(t0, c3, c2);
    const double c5 = fma(t0, c4, c3) * kInvInt(5);
    double R = fma(d, c5, c4);
    R = fma(d, R, c3);
    R = fma(d, R, c2);
    R = fma(d, R, c1);
    R = fma(d, R, R0);
    R_out = R;
    return t + __drcp_rn(R);
}

// Re-lay x / y into slices.  One thread per (slice, lane).
__global__ void sell_fill_kernel(const double *__restrict__ x, const uint8_t *__restrict__ y,
                                 const int64_t *__restrict__ lane_src, const int32_t *__restrict__ lane_meta,
                                 const int64_t *__restrict__ slice_xbase, const int64_t *__restrict__ slice_ybase,
                                 const int32_t *__restrict__ slice_len, double *__restrict__ x_sell,
                                 uint32_t *__restrict__ ybits, int64_t S) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= S * 32) return;
    const int64_t s = t >> 5;
    const int lane = (int)(t & 31);
    const int L = slice_len[s], len = lane_meta[t] & 0xffff;
    const int64_t src = lane_src[t], xb = slice_xbase[s], yb = slice_ybase[s];
    uint32_t word = 0;
    for (int i = 0; i < L; ++i) {
        double xv = 0.0;
        if (i < len) {
            xv = x[src + i];
            word |= (uint32_t)(y[src + i] != 0) << (i & 31);
        }
        x_sell[xb + (int64_t)i * 32 + lane] = xv;
        if ((i & 31) == 31 || i == L - 1) {
            ybits[yb + (int64_t)(i >> 5) * 32 + lane] = word;
            word = 0;
        }
    }
}

// H^T H per pair into stats[slice][D + t][lane]; g rows zeroed.  One thread per (slice, lane).
template <int D>
__global__ void hh_sliced_kernel(const double *__restrict__ x_sell, const int64_t *__restrict__ slice_xbase,
                                 const int32_t *__restrict__ lane_meta, double *__restrict__ stats, int64_t S,
                                 BasisDev basis) {
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
        double h[D];
        eval_basis<D>(x_sell[xb + (int64_t)i * 32 + lane], basis, h);
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

struct SweepV2Args {
    const double *x_sell;
    const uint32_t *ybits;
    const int64_t *slice_xbase, *slice_ybase;
    const int32_t *slice_len, *slice_region, *slice_cb, *lane_cell, *lane_meta, *blk_cells;
    double *stats;
    const double *U, *r;
    double *log_rho, *lbbuf;
    const double *mills_tab;
    const double2 *rbf_tab;
    RbfTab rt;
    BasisDev basis;
    int K, Tn, N;
    unsigned long long *slice_counter;
    int64_t S, chunk;
};

// The per-CpG sweep on slices (R/melissa_vb.R:418-435 refresh + the next iteration's
// E-step terms :322-327 + lb_pz_qz :469-473 when WITH_LB).
// Persistent CTAs load the two function tables once; after that every WARP is on its
// own: it claims slices one at a time from a global counter (slices are short, ~7 us,
// so there is no tail and no CTA barrier), keeps the E-step sums of its current cell
// block in a private shared-memory array indexed by the cell's position in the block,
// and adds that array to log_rho when it moves on to another block.
template <int D, bool WITH_LB, bool UNIFORM_R>
__global__ void __launch_bounds__(256, SWEEP_MINB) sweep_v2_kernel(SweepV2Args a) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    constexpr int XU = SWEEP_XU;                // CpG rows fetched per group (prefetch distance, ILP)
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K = a.K, Tn = a.Tn;
    const int nw = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // shared: rbf table | mills table | per-warp U staging | per-warp accumulators
    double2 *s_rbf = reinterpret_cast<double2 *>(smem_raw);
    double *s_mills = reinterpret_cast<double *>(s_rbf + a.rt.n_entries);
    double *s_u = s_mills + MILLS_N + 1;
    double *s_acc = s_u + (SWEEP_STAGE_U ? nw * K * R : 0);
    for (int i = threadIdx.x; i < a.rt.n_entries; i += blockDim.x) s_rbf[i] = a.rbf_tab[i];
    for (int i = threadIdx.x; i < MILLS_N; i += blockDim.x) s_mills[i] = a.mills_tab[i];
    for (int i = threadIdx.x; i < nw * Tn * K; i += blockDim.x) s_acc[i] = 0.0;
    __syncthreads();
    const double inv_K = 1.0 / (double)K;
    double *my_u = s_u + warp * K * R;
    double *my_acc = s_acc + warp * Tn * K;
    double lb_local = 0.0, wrong_side = 0.0;
    int cur_cb = -1, cur_m = -1;
    // Slices are handed out in runs of `chunk` consecutive ones: a run stays inside one or two cell
    // blocks, so the private sums are flushed once per run instead of once per slice (with ~3,500
    // warps in flight, single slices would land every warp in a new block each time).
    const int64_t chunk = a.chunk;
    auto claim = [&]() -> int64_t {
        unsigned long long c = 0;
        if (lane == 0) c = atomicAdd(a.slice_counter, (unsigned long long)chunk);
        return (int64_t)__shfl_sync(0xffffffffu, c, 0);
    };
    auto flush = [&]() {                        // private sums of block cur_cb -> log_rho, then clear
        if (cur_cb < 0) return;
        __syncwarp();
        for (int j = lane; j < Tn; j += 32) {
            const int cell = a.blk_cells[(int64_t)cur_cb * Tn + j];
            for (int k = 0; k < K; ++k) {
                const double v = my_acc[j * K + k];
                if (cell >= 0 && v != 0.0) atomicAdd(&a.log_rho[(int64_t)cell * K + k], v);
                my_acc[j * K + k] = 0.0;
            }
        }
        __syncwarp();
    };
    int64_t run = claim();
    while (run < a.S) {
        const int64_t run_following = claim();  // claimed early: its round trip hides behind this run
        const int64_t run_end = min(run + chunk, a.S);
        // slice metadata is fetched one slice ahead, so that only the first slice of a run waits for it
        int n_cb = a.slice_cb[run], n_m = a.slice_region[run], n_L = a.slice_len[run];
        int64_t n_xb = a.slice_xbase[run], n_yb = a.slice_ybase[run];
        int n_cell = a.lane_cell[run * 32 + lane], n_meta = a.lane_meta[run * 32 + lane];
        for (int64_t s = run; s < run_end; ++s) {
        const int cb = n_cb, m = n_m, L = n_L, cell = n_cell, meta = n_meta;
        const int64_t xb = n_xb, yb = n_yb;
        if (s + 1 < run_end) {
            n_cb = a.slice_cb[s + 1]; n_m = a.slice_region[s + 1]; n_L = a.slice_len[s + 1];
            n_xb = a.slice_xbase[s + 1]; n_yb = a.slice_ybase[s + 1];
            n_cell = a.lane_cell[(s + 1) * 32 + lane]; n_meta = a.lane_meta[(s + 1) * 32 + lane];
        }
        if (cb != cur_cb) {
            flush();
            cur_cb = cb;
        }
        const int len = meta & 0xffff, nl = meta >> 16;
        const bool valid = cell >= 0;
        const double *xrow = a.x_sell + xb + lane;
        double xa[XU];
#pragma unroll
        for (int u = 0; u < XU; ++u) xa[u] = xrow[u * 32];
        uint32_t yw = a.ybits[yb + lane];
        constexpr int RK = 8;                   // responsibilities of the lane's cell, fetched up front when K <= 8
        double rkv[RK];
#pragma unroll
        for (int k = 0; k < RK; ++k)
            rkv[k] = UNIFORM_R ? inv_K : ((valid && k < K && K <= RK) ? a.r[(int64_t)cell * K + k] : 0.0);
        double *rec = a.stats + s * (R * 32) + (lane >> 2) * (R * 4) + (lane & 3);
        double hh[T];
#pragma unroll
        for (int q = 0; q < T; ++q) hh[q] = rec[(D + q) * 4];
        const double *um = SWEEP_STAGE_U ? my_u : a.U + (int64_t)m * K * R;
        if (SWEEP_STAGE_U && m != cur_m) {      // consecutive slices of a run mostly share the region
            __syncwarp();
            for (int i = lane; i < K * R; i += 32) my_u[i] = a.U[(int64_t)m * K * R + i];
            __syncwarp();
            cur_m = m;
        }
        // -1/2 tr(H^T H (m m^T + S)) needs no E[z]: do it while the first CpG rows are in flight
        double wbar[D];
#pragma unroll
        for (int d = 0; d < D; ++d) wbar[d] = 0.0;
        auto pre_k = [&](int k, double rk) {
            const double *u = um + k * R;
            double trc = 0.0;
#pragma unroll
            for (int q = 0; q < T; ++q) trc = fma(u[D + q], hh[q], trc);
#pragma unroll
            for (int d = 0; d < D; ++d) wbar[d] = fma(rk, u[d], wbar[d]);
            if (valid) {
                my_acc[nl * K + k] += trc;                 // lanes of one slice are distinct cells
                if (WITH_LB) lb_local = fma(rk, trc, lb_local);
            }
        };
        if (K <= RK) {
#pragma unroll
            for (int k = 0; k < RK; ++k)
                if (k < K) pre_k(k, rkv[k]);
        } else {
            for (int k = 0; k < K; ++k) pre_k(k, UNIFORM_R ? inv_K : (valid ? a.r[(int64_t)cell * K + k] : 0.0));
        }
        double g[D];
#pragma unroll
        for (int d = 0; d < D; ++d) g[d] = 0.0;
        // Rows past a lane's own length are masked to E[z] = 0 and rows past L (read from whatever
        // follows in x_sell, which is over-allocated by one group) are past every lane's length,
        // so each group runs all XU rows with no trip-count test and no predicated load.
        double lb_s = 0.0, wrong_s = 0.0;
        bool bad_any = false;
        const double *xp = xrow;
        for (int i0 = 0; i0 < L; i0 += XU) {
            double xnext[XU];
#pragma unroll
            for (int u = 0; u < XU; ++u) xnext[u] = xp[(XU + u) * 32];
            xp += XU * 32;
            uint32_t ywn = yw;
            if (((i0 + XU) & 31) == 0 && i0 + XU < L) ywn = a.ybits[yb + (int64_t)((i0 + XU) >> 5) * 32 + lane];
#pragma unroll
            for (int u = 0; u < XU; ++u) {
                const int i = i0 + u;
                const uint32_t ybit = (yw >> (i & 31)) & 1u;
                const uint32_t sflip = (ybit ^ 1u) << 31;             // s = +1 for y = 1, -1 for y = 0
                const uint32_t live = i < len ? 0xffffffffu : 0u;
                double h[D];
                const bool bad = eval_basis_fast<D>(xa[u], a.rt, s_rbf, h);
                double mu = wbar[0];
#pragma unroll
                for (int d = 1; d < D; ++d) mu = fma(h[d], wbar[d], mu);
                const double t = __hiloint2double(__double2hiint(mu) ^ sflip, __double2loint(mu));
                bool bad_t;
                double Rt;
                const double G = trunc_mean_fast(t, s_mills, bad_t, Rt);
                bad_any |= (bad || bad_t) && live;
                const double Ez = __hiloint2double((__double2hiint(G) ^ sflip) & live, __double2loint(G) & live);
                g[0] += Ez;
#pragma unroll
                for (int d = 1; d < D; ++d) g[d] = fma(h[d], Ez, g[d]);
                if (WITH_LB) {
                    if (live) {
                        // log(Phi(s mu) + 1e-10): both branches of R/melissa_vb.R:471-472 reduce to it, and
                        // Phi = R * phi with the R the E[z] evaluation just produced (one exp + one log)
                        lb_s += 0.5 * mu * mu + log(fma(Rt, 0.39894228040143267794 * exp(-0.5 * t * t), 1e-10));
                        if (t < -5.0) wrong_s += 1.0;
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < XU; ++u) xa[u] = xnext[u];
            yw = ywn;
        }
        if (__any_sync(0xffffffffu, bad_any)) {                    // never taken on sane data
            WBar<D> wb;
#pragma unroll
            for (int d = 0; d < D; ++d) wb.w[d] = wbar[d];
            const SliceSums<D> o = sweep_slice_slow<D, WITH_LB>(xrow, a.ybits + yb + lane, L, len, wb, &a.basis);
#pragma unroll
            for (int d = 0; d < D; ++d) g[d] = o.g[d];
            lb_s = o.lb;
            wrong_s = o.wrong;
        }
        if (WITH_LB) {
            lb_local += lb_s;
            wrong_side += wrong_s;
        }
#pragma unroll
        for (int d = 0; d < D; ++d) rec[d * 4] = g[d];
        if (valid) {
            for (int k = 0; k < K; ++k) {
                const double *u = um + k * R;
                double lin = 0.0;
#pragma unroll
                for (int d = 0; d < D; ++d) lin = fma(u[d], g[d], lin);
                my_acc[nl * K + k] += lin;
            }
        }
        }
        run = run_following;
    }
    flush();
    if (WITH_LB) {
        lb_local = warp_sum(lb_local);
        wrong_side = warp_sum(wrong_side);
        if (lane == 0) {
            atomicAdd(&a.lbbuf[0], lb_local);
            if (wrong_side != 0.0) atomicAdd(&a.lbbuf[1], wrong_side);
        }
    }
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
                eval_basis_tab<D>(xrow[(int64_t)i * 32], rt, s_rbf, basis, h);
                double eta = w[0];
#pragma unroll
                for (int d = 1; d < D; ++d) eta = fma(h[d], w[d], eta);
                const double t = yi ? eta : -eta;
                const double G = trunc_mean_tab(t, s_mills);     // t + phi/Phi
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
