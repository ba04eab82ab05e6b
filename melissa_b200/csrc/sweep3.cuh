// The per-CpG sweep, third generation (R/melissa_vb.R:418-435 refresh + the next iteration's
// E-step terms :322-327 + lb_pz_qz :469-473 when WITH_LB).
//
// What changed against sweep_v2 (round 1) and why:
//   * ONE persistent CTA per SM.  The two function tables are therefore paid once per SM and can
//     be finer (truncated-normal table at spacing 1/512 with a 4th-order series, RBF table at a
//     power-of-two spacing with 6th/5th-order series): 52 FP64 operations per CpG instead of 74.
//   * The E-step sums of a cell block live in ONE CTA-wide shared array, not one per warp, so
//     the cell block (= the pool the slices of a region are sorted in) no longer shrinks with K:
//     256 cells at every K, which takes the row padding at K = 10 from 1.39x to 1.07x.
//   * The CTA claims CHUNKS (runs of slices inside one cell block) from a global counter, its
//     warps claim slices of the chunk from a shared counter; the block's sums are flushed to
//     global memory once per chunk.  The chunk list is guided self-scheduling (make_chunks,
//     pack_gpu.cuh): long chunks while much is left, short ones at the end; chunks shorter than
//     640 slices are claimed slice by slice, so that little is lost at the barrier that ends a chunk.
//   * WITH_LB: log(Phi(s mu) + 1e-10) comes from a second table (L, L') on the Mills grid and the
//     recurrence of L' (log_phi_eps, sliced.cuh) instead of one exp and one log per CpG.
//   * Those sums are accumulated in 128-bit fixed point (integer atomics, fx128.cuh), so the
//     result does not depend on the order in which warps and CTAs arrive: log rho, and with it
//     the hard labels (first arg-max, R/melissa_vb.R:212-213), are reproducible bit for bit.
//
// Included from melissa.cu inside its anonymous namespace, after sliced.cuh.
#pragma once

struct Sweep3Args {
    const double *x_sell;
    const uint32_t *ybits;
    const int64_t *slice_xbase, *slice_ybase;
    const int32_t *slice_len, *slice_region, *slice_cb, *lane_cell, *lane_meta, *blk_cells;
    const int32_t *chunk_ptr;                 // [NC + 1] first slice of every chunk; a chunk stays inside one cell block
    double *stats;
    const double *U, *r;
    unsigned long long *log_rho_fx;           // [N][K][2]
    unsigned long long *lb_fx;                // [4]: lb_pz_qz (lo, hi), #CpG with s mu < -5, unused
    unsigned int *poison;                     // set when a term is non-finite or beyond the fixed-point range
    const double *mills_tab;
    const double2 *lphi_tab;
    const double2 *rbf_tab;
    RbfTab rt;
    BasisDev basis;
    int K, Tn, NC, G;
    int direct;                               // 1: no shared block sums (K too large for them): E-step terms go straight to log_rho_fx
    unsigned int *chunk_counter;
};

#ifndef SWEEP3_XU
#define SWEEP3_XU 2           // CpG rows per group: one 16-byte cp.async per lane moves a group (2 rows x 256 bytes)
#endif
#ifndef SWEEP3_NST
#define SWEEP3_NST 4          // groups of the x ring per warp: NST - 1 groups (6 rows) are in flight ahead of the arithmetic
#endif
static_assert(SWEEP3_XU % 2 == 0, "the x ring moves two rows per cp.async");
static_assert((SWEEP3_NST & (SWEEP3_NST - 1)) == 0, "ring size must be a power of two");

// One group of x (rows 2g, 2g + 1 of a slice: 512 contiguous bytes) global -> shared, no registers
// involved: lanes 0-15 move the first row, lanes 16-31 the second, 16 bytes each, bypassing L1.
__device__ __forceinline__ void x_group_async(uint32_t dst_smem, const double *src) {
#pragma unroll
    for (int c = 0; c < SWEEP3_XU / 2; ++c)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_smem + c * 512), "l"(src + c * 64) : "memory");
}
__device__ __forceinline__ void x_group_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void x_group_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int D, bool WITH_LB>
__global__ void __launch_bounds__(SWEEP3_THREADS, 1) sweep3_kernel(Sweep3Args a) {
    constexpr int T = D * (D + 1) / 2, R = D + T;
    constexpr int XU = SWEEP3_XU, NST = SWEEP3_NST;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int K = a.K, Tn = a.Tn;
    const int nw = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int UR = (K * R + 1) & ~1;
    // shared: rbf table | mills table | block sums (fixed point) | per-warp U[m] | lb sums | control words
    double2 *s_rbf = reinterpret_cast<double2 *>(smem_raw);
    double *s_mills = reinterpret_cast<double *>(s_rbf + a.rt.n_entries);
    unsigned int *s_acc = reinterpret_cast<unsigned int *>(s_mills + MILLS_N + 1);      // [4 limbs][K][Tn]
    double *s_u = reinterpret_cast<double *>(s_acc + (a.direct ? 0 : 4 * (size_t)Tn * K));
    double *s_x = s_u + (size_t)nw * UR;                                                         // per-warp ring of x groups
    unsigned long long *s_lb = reinterpret_cast<unsigned long long *>(s_x + (size_t)nw * NST * XU * 32);   // [2] + wrong-side count
    int *s_ctl = reinterpret_cast<int *>(s_lb + 4);                                             // cursor, next chunk
    for (int i = threadIdx.x; i < a.rt.n_entries; i += blockDim.x) s_rbf[i] = a.rbf_tab[i];
    for (int i = threadIdx.x; i < MILLS_N; i += blockDim.x) s_mills[i] = a.mills_tab[i];
    if (!a.direct)
        for (int i = threadIdx.x; i < 4 * Tn * K; i += blockDim.x) s_acc[i] = 0u;
    if (threadIdx.x < 4) s_lb[threadIdx.x] = 0ull;
    if (threadIdx.x == 0) s_ctl[1] = (int)atomicAdd(a.chunk_counter, 1u);
    __syncthreads();
    double *my_u = s_u + (size_t)warp * UR;
    // x ring of this warp.  Groups are issued and consumed strictly in order: `iss` / `con` count them,
    // slot = counter mod NST.  Between slices NST - 1 groups stay in flight (`pre` of them belong to the
    // slice that comes next), so that a slice does not start with an empty pipeline.
    // (NST - 1 groups are always between the issue and the consume position, so one slot offset serves both.)
    constexpr uint32_t SLOT_BYTES = XU * 256, RING_MASK = (NST - 1) * SLOT_BYTES;
    const uint32_t my_x_smem = (uint32_t)__cvta_generic_to_shared(s_x + (size_t)warp * NST * XU * 32);
    const uint32_t my_x_fill = my_x_smem + (uint32_t)lane * 16;      // this lane's 16 bytes of a group
    const uint32_t my_x_read = my_x_smem + (uint32_t)lane * 8;       // this lane's CpG of a row
    const uint32_t rbf_addr = (uint32_t)__cvta_generic_to_shared(s_rbf), mills_addr = (uint32_t)__cvta_generic_to_shared(s_mills);
    uint32_t slot = 0;                                               // byte offset of the next slot to fill
    int pre = 0;
    int cur_m = -1;
    int chunk = s_ctl[1];
    while (chunk < a.NC) {
        const int c_beg = a.chunk_ptr[chunk], c_end = a.chunk_ptr[chunk + 1];
        const int G = (c_end - c_beg) < 640 ? 1 : a.G;     // short chunks: single slices, less waiting at the chunk's end
        const int cb = a.slice_cb[c_beg];
        __syncthreads();                                   // everybody has read s_ctl[1]; the last flush is done
        if (threadIdx.x == 0) {
            s_ctl[0] = c_beg;
            s_ctl[1] = (int)atomicAdd(a.chunk_counter, 1u);        // the chunk after this one, claimed early
        }
        __syncthreads();
        auto claim = [&]() -> int {
            int c = 0;
            if (lane == 0) c = atomicAdd(&s_ctl[0], G);
            return __shfl_sync(0xffffffffu, c, 0);
        };
        int run = claim();
        // slice metadata is fetched one slice ahead
        int n_m = 0, n_L = 0, n_cell = -1, n_meta = 0;
        int64_t n_xb = 0, n_yb = 0;
        auto fetch_meta = [&](int s) {
            n_m = a.slice_region[s]; n_L = a.slice_len[s];
            n_xb = a.slice_xbase[s]; n_yb = a.slice_ybase[s];
            n_cell = a.lane_cell[(int64_t)s * 32 + lane]; n_meta = a.lane_meta[(int64_t)s * 32 + lane];
        };
        if (run < c_end) fetch_meta(run);
        while (run < c_end) {
            const int run_following = claim();
            const int run_end = min(run + G, c_end);
            for (int s = run; s < run_end; ++s) {
                const int m = n_m, L = n_L, cell = n_cell, meta = n_meta;
                const int64_t xb = n_xb, yb = n_yb;
                bool has_next;
                {
                    const int s_next = s + 1 < run_end ? s + 1 : run_following;
                    has_next = s_next < c_end;
                    if (has_next) fetch_meta(s_next);
                }
                const int len = meta & 0xffff, nl = meta >> 16;
                const bool valid = cell >= 0;
                const double *xrow = a.x_sell + xb + lane;
                const double *xgrp = a.x_sell + xb + 2 * lane;              // + 32 XU g: this lane's 16 bytes of each row pair of group g
                const int ng = (L + XU - 1) / XU;
                const double *fsrc = xgrp + 32 * XU * (NST - 1);            // next group to fetch
                if (pre != NST - 1) {                                       // cold start: nothing usable in flight
                    __syncwarp();                                           // (every lane is done reading the ring)
                    x_group_wait<0>();
#pragma unroll
                    for (int gq = 0; gq < NST - 1; ++gq) {                 // (groups past the slice read the slack / the next slice: harmless)
                        x_group_async(my_x_fill + slot, xgrp + 32 * XU * gq);
                        x_group_commit();
                        slot = (slot + SLOT_BYTES) & RING_MASK;
                    }
                }
                uint32_t yw = a.ybits[yb + lane];
                if (m != cur_m) {                           // consecutive slices of a run mostly share the region
                    __syncwarp();
                    const double *src = a.U + (int64_t)m * K * R;
                    for (int i = lane; i < K * R; i += 32) my_u[i] = src[i];
                    __syncwarp();
                    cur_m = m;
                }
                const double *rp = a.r + (int64_t)(valid ? cell : 0) * K;
                double *rec = a.stats + (int64_t)s * (R * 32) + (lane >> 2) * (R * 4) + (lane & 3);
                {   // the H^T H part of the record is wanted after the row loop: start it on its way to L2 now
                    // (the 4 lanes of a record group share its 32-byte rows, so each lane asks for every 4th one)
                    const double *hrow = rec - (lane & 3) + (D + (lane & 3)) * 4;
#pragma unroll
                    for (int q = 0; q < T; q += 4)
                        if (q + (lane & 3) < T) asm volatile("prefetch.global.L2 [%0];" ::"l"(hrow + q * 4));
                }
                double wbar[D];
#pragma unroll
                for (int d = 0; d < D; ++d) wbar[d] = 0.0;
                for (int k0 = 0; k0 < K; k0 += 8) {         // the cell's responsibilities, eight loads in flight at a time
                    double rk[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) rk[j] = (valid && k0 + j < K) ? rp[k0 + j] : 0.0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        if (k0 + j < K) {
                            const double *u = my_u + (k0 + j) * R;
#pragma unroll
                            for (int d = 0; d < D; ++d) wbar[d] = fma(rk[j], u[d], wbar[d]);
                        }
                    }
                }
                double g[D];
#pragma unroll
                for (int d = 0; d < D; ++d) g[d] = 0.0;
                // Rows past a lane's own length contribute nothing (predicated), rows past L (read from
                // whatever follows in x_sell, which is over-allocated by two groups) are past every lane's
                // length, so each group runs all XU rows with no trip-count test and no predicated load.
                double lb_s = 0.0;
                int wrong_s = 0;
                unsigned bad = 0;
                // the slice after this one is prefetched into the ring while this one's last groups are consumed
                // (n_* already hold its metadata); short slices skip that and start cold
#ifdef SWEEP3_NOCHAIN
                const bool chain = false;
#else
                const bool chain = has_next && ng >= NST - 1;
#endif
                const int i_switch = chain ? XU * (ng - (NST - 1)) : 0x7fffffff;   // from this row on, fetch the next slice
                const double *nsrc = chain ? a.x_sell + n_xb + 2 * lane : fsrc;
                // methylation states come 32 rows per word: the row loop runs word by word
                for (int wbase = 0; wbase < L; wbase += 32) {
                    const int iend = min(wbase + 32, L);
                    uint32_t ywn = 0;
                    if (iend < L) ywn = a.ybits[yb + (int64_t)((wbase >> 5) + 1) * 32 + lane];
                    // bit 31 of ycur = "flip the sign" for the current row (y = 0), shifted up row by row
                    uint32_t ycur = __brev(~yw);
                    for (int i0 = wbase; i0 < iend; i0 += XU) {
                        if (i0 == i_switch) fsrc = nsrc;
                        // The slot filled now is the one read in the previous trip.  Lanes of a warp need not run
                        // in lock-step (a lane that is behind would find the NEXT group's bytes in its slot), so
                        // nobody refills before everybody has read:
                        __syncwarp();
                        x_group_async(my_x_fill + slot, fsrc);
                        x_group_commit();
                        fsrc += 32 * XU;
                        slot = (slot + SLOT_BYTES) & RING_MASK;             // = the slot NST - 1 groups back: the one to consume
                        x_group_wait<NST - 1>();
                        __syncwarp();
                        double xa[XU];
#ifdef SWEEP3_NORING
#pragma unroll
                        for (int u = 0; u < XU; ++u) xa[u] = xrow[(int64_t)(i0 + u) * 32];
#else
#pragma unroll
                        for (int u = 0; u < XU; ++u)
                            asm volatile("ld.shared.f64 %0, [%1];" : "=d"(xa[u]) : "r"(my_x_read + slot + u * 256) : "memory");
#endif
#pragma unroll
                        for (int u = 0; u < XU; ++u) {
                            const int i = i0 + u;
                            const uint32_t yu = ycur << u;
                            // ---- RBF row (see rbf_row3 in sliced.cuh for the maths; here with the grid offset folded
                            //      into the rounding constant and shared-memory addresses instead of pointers)
                            double h[D];
                            unsigned bad_row = 0;
                            h[0] = 1.0;
                            if (D > 1) {
                                constexpr int NW = D <= 9 ? 6 : 8;
                                const double v = xa[u] + a.rt.magic;                       // low word = table index
                                const unsigned j = (unsigned)__double2loint(v);
                                const unsigned jc = min(j, (unsigned)(a.rt.n_entries - 1));
                                bad_row = j ^ jc;
                                const double x0 = v - a.rt.magic;
                                const double dx = xa[u] - x0;
                                double e0, b0;
                                asm("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(e0), "=d"(b0) : "r"(rbf_addr + jc * 16));
                                const double w = dx * fma(a.rt.neg_gamma, dx, fma(x0, a.rt.cA, a.rt.cB));
                                h[1] = e0 * exp_series<NW>(w);
                                if (D > 2) {
                                    double bj = b0 * exp_series<NW - 1>(a.rt.g2d * dx);
#pragma unroll
                                    for (int q = 2; q < D; ++q) {
                                        h[q] = h[q - 1] * bj;
                                        if (q + 1 < D) bj *= a.rt.q2;
                                    }
                                }
                            }
                            double mu = wbar[0];
#pragma unroll
                            for (int d = 1; d < D; ++d) mu = fma(h[d], wbar[d], mu);
                            // t = s mu, s = +1 for y = 1 and -1 for y = 0: one three-input logic op on the high word
                            const double t = __hiloint2double(__double2hiint(mu) ^ (int)(yu & 0x80000000u), __double2loint(mu));
                            // ---- E[z | z > 0] = t + 1 / R(t) (trunc_mean3 in sliced.cuh)
                            const double vt = t + MILLS_MAGIC;                             // low word = table index
                            const unsigned it = (unsigned)__double2loint(vt);
                            const unsigned ic = min(it, 2u * MILLS_HALF);
                            bad_row |= it ^ ic;
                            const double t0 = vt - MILLS_MAGIC;
                            const double dt = t - t0;
                            double R0;
                            asm("ld.shared.f64 %0, [%1];" : "=d"(R0) : "r"(mills_addr + ic * 8));
                            // the bound's table entry is asked for here, a Mills series ahead of its use: ncu had 24 %
                            // of the bound sweep's samples waiting for it when it was loaded where it is needed
                            double2 lphi = make_double2(0.0, 0.0);
                            if (WITH_LB) lphi = __ldg(a.lphi_tab + ic);
                            const double c1 = fma(t0, R0, 1.0);
                            const double c2 = 0.5 * fma(t0, c1, R0);
                            const double c3 = fma(t0, c2, c1) * kInvInt(3);
                            const double c4 = 0.25 * fma(t0, c3, c2);
                            double Rt = fma(dt, c4, c3);
                            Rt = fma(dt, Rt, c2);
                            Rt = fma(dt, Rt, c1);
                            Rt = fma(dt, Rt, R0);
                            double y;
                            asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(Rt));
                            const double e = fma(-Rt, y, 1.0);
                            y = fma(y, fma(e, e, e), y);
                            const double Gt = t + y;
                            const double Ez = __hiloint2double(__double2hiint(Gt) ^ (int)(yu & 0x80000000u), __double2loint(Gt));
                            if (i < len) {
                                bad |= bad_row;
                                g[0] += Ez;
#pragma unroll
                                for (int d = 1; d < D; ++d) g[d] = fma(h[d], Ez, g[d]);
                                if (WITH_LB) {
                                    // log(Phi(s mu) + 1e-10): both branches of R/melissa_vb.R:471-472 reduce to it; taken
                                    // from its own table on the grid point t0 the E[z] evaluation just used (the table
                                    // is read through L1 / L2: this branch runs on bound iterations only)
                                    lb_s += fma(0.5 * mu, mu, log_phi_eps(t0, dt, lphi));
                                    wrong_s += t < -5.0;
                                }
                            }
                        }
                        ycur <<= XU;
                    }
                    yw = ywn;
                }
                pre = chain ? NST - 1 : 0;
                if (__any_sync(0xffffffffu, bad != 0)) {                   // never taken on sane data
                    WBar<D> wb;
#pragma unroll
                    for (int d = 0; d < D; ++d) wb.w[d] = wbar[d];
                    const SliceSums<D> o = sweep_slice_slow<D, WITH_LB>(xrow, a.ybits + yb + lane, L, len, wb, &a.basis);
#pragma unroll
                    for (int d = 0; d < D; ++d) g[d] = o.g[d];
                    lb_s = o.lb;
                    wrong_s = (int)o.wrong;
                }
#pragma unroll
                for (int d = 0; d < D; ++d) rec[d * 4] = g[d];
                // E-step terms of the next iteration (R/melissa_vb.R:325-326): U[m,k] . [g | H^T H]
                double hh[T];
#pragma unroll
                for (int q = 0; q < T; ++q) hh[q] = rec[(D + q) * 4];
                if (valid) {
                    unsigned int *acc = s_acc + nl;
                    bool over = false;
                    auto term = [&](int k, double &trc) {                   // four independent chains per cluster
                        const double *u = my_u + k * R;
                        double t0 = 0.0, t1 = 0.0, t2 = 0.0, lin = 0.0;
#pragma unroll
                        for (int q = 0; q < T; q += 3) {
                            t0 = fma(u[D + q], hh[q], t0);
                            if (q + 1 < T) t1 = fma(u[D + q + 1], hh[q + 1], t1);
                            if (q + 2 < T) t2 = fma(u[D + q + 2], hh[q + 2], t2);
                        }
#pragma unroll
                        for (int d = 0; d < D; ++d) lin = fma(u[d], g[d], lin);
                        trc = (t0 + t1) + t2;
                        return lin + trc;
                    };
                    for (int k = 0; k < K; k += 2) {                        // two clusters at a time
                        double trc0, trc1 = 0.0;
                        const double v0 = term(k, trc0);
                        const bool two = k + 1 < K;
                        const double v1 = two ? term(k + 1, trc1) : 0.0;
                        over |= !(fabs(v0) < LIMB_LIMIT) || !(fabs(v1) < LIMB_LIMIT);
                        if (!a.direct) {
                            limb_add_shared(acc + k * Tn, K * Tn, v0);
                            if (two) limb_add_shared(acc + (k + 1) * Tn, K * Tn, v1);
                        } else {
                            unsigned long long *gacc = a.log_rho_fx + 2 * ((int64_t)cell * K + k);
                            fx_add(gacc, fx_from_double(v0));
                            if (two) fx_add(gacc + 2, fx_from_double(v1));
                        }
                        if (WITH_LB) {
                            lb_s = fma(rp[k], trc0, lb_s);
                            if (two) lb_s = fma(rp[k + 1], trc1, lb_s);
                        }
                    }
                    if (over) atomicOr(a.poison, 1u);
                }
                if (WITH_LB) {
                    // per-slice sums go through the fixed butterfly and then into the fixed-point
                    // accumulator, so the bound does not depend on which warp took which slice either
                    if (!valid) lb_s = 0.0;
                    lb_s = warp_sum(lb_s);
                    wrong_s = __reduce_add_sync(0xffffffffu, wrong_s);
                    if (lane == 0) {
                        if (!(fabs(lb_s) < FX_LIMIT)) atomicOr(a.poison, 1u);
                        fx_add(s_lb, fx_from_double(lb_s));
                        if (wrong_s) atomicAdd(s_lb + 2, (unsigned long long)wrong_s);
                    }
                }
            }
            run = run_following;
        }
        __syncthreads();
        const int chunk_next = s_ctl[1];
        // the block's sums -> global, then clear
        const int KT = a.direct ? 0 : K * Tn;
        for (int idx = threadIdx.x; idx < KT; idx += blockDim.x) {
            const int k = idx / Tn, j = idx - k * Tn;
            const unsigned int l0 = s_acc[idx], l1 = s_acc[KT + idx], l2 = s_acc[2 * KT + idx], l3 = s_acc[3 * KT + idx];
            if (l0 | l1 | l2 | l3) {
                const int cell = a.blk_cells[(int64_t)cb * Tn + j];
                fx_add(a.log_rho_fx + 2 * ((int64_t)cell * K + k), limb_to_fx(l0, l1, l2, l3));
                s_acc[idx] = 0u;
                s_acc[KT + idx] = 0u;
                s_acc[2 * KT + idx] = 0u;
                s_acc[3 * KT + idx] = 0u;
            }
        }
        chunk = chunk_next;
    }
    if (WITH_LB) {
        __syncthreads();
        if (threadIdx.x == 0) {
            Fx128 f;
            f.lo = s_lb[0];
            f.hi = (long long)s_lb[1];
            if (f.lo | (unsigned long long)f.hi) fx_add(a.lb_fx, f);
            if (s_lb[2]) atomicAdd(a.lb_fx + 2, s_lb[2]);
        }
    }
}

// lb_fx -> lbbuf (doubles, the all-reduce payload of bound iterations), then clear.
__global__ void lb_finish_kernel(unsigned long long *lb_fx, double *lbbuf) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    lbbuf[0] = fx_to_double(lb_fx[0], lb_fx[1]);
    lbbuf[1] = (double)lb_fx[2];
    lb_fx[0] = lb_fx[1] = lb_fx[2] = lb_fx[3] = 0ull;
}
