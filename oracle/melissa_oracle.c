/*
 * CPU restatement in plain C (FP64, optional OpenMP) of the reference's VB loop.
 * TEST INFRASTRUCTURE ONLY: linked by tests/ and by bench.py's cpu_baseline /
 * `--impl reference` legs, never by the product (melissa_b200/).
 *
 * PARITY UNPINNED against R itself (R is not installed here and the reference's
 * own tests assert nothing on this path); it is pinned against
 * oracle/melissa_oracle.py, which carries the anchoring described in its header.
 *
 * Follows andreaskapou/Melissa R/melissa_vb.R:232-529 (melissa_inner) phase by
 * phase, keeping what the reference keeps in memory: the design matrices H
 * (R/utils.R:64-75), H^T H per pair (:254-255), mu / E[z] per CpG (:283-302).
 * Parallel regions mirror the reference's mclapply loops: over cells for the
 * E-step (:315) and the refresh (:398), over regions for the M-step (:343).
 *
 * Build:  make -C oracle      ->  oracle/liboracle.so
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define MAXD 16

static double digamma_pos(double x) {
    double acc = 0.0;
    while (x < 10.0) { acc -= 1.0 / x; x += 1.0; }
    double i = 1.0 / x, i2 = i * i;
    double s = i2 * (1.0 / 12 - i2 * (1.0 / 120 - i2 * (1.0 / 252 - i2 * (1.0 / 240 - i2 * (1.0 / 132
               - i2 * (691.0 / 32760 - i2 * (1.0 / 12)))))));
    return acc + log(x) - 0.5 * i - s;
}

static double pnorm(double z) { return 0.5 * erfc(-z * M_SQRT1_2); }
static double dnorm(double z) { return exp(-0.5 * z * z) * 0.39894228040143267794; }

/* R/melissa_vb.R:222-226, the reference's own expressions */
static double update_Ez(double mu, int y) {
    if (y) return mu + dnorm(-mu) / (1.0 - pnorm(-mu));
    return mu - dnorm(-mu) / pnorm(-mu);
}

/* solve(A) and det(A) by LU with partial pivoting (base R uses LAPACK dgesv / dgetrf) */
static double inv_det(int D, const double *A, double *inv) {
    double a[MAXD][2 * MAXD];
    double det = 1.0;
    for (int i = 0; i < D; ++i)
        for (int j = 0; j < D; ++j) { a[i][j] = A[i * D + j]; a[i][D + j] = (i == j); }
    for (int c = 0; c < D; ++c) {
        int piv = c;
        for (int r = c + 1; r < D; ++r) if (fabs(a[r][c]) > fabs(a[piv][c])) piv = r;
        if (piv != c) { for (int j = 0; j < 2 * D; ++j) { double t = a[c][j]; a[c][j] = a[piv][j]; a[piv][j] = t; } det = -det; }
        det *= a[c][c];
        double ip = 1.0 / a[c][c];
        for (int j = 0; j < 2 * D; ++j) a[c][j] *= ip;
        for (int r = 0; r < D; ++r) if (r != c) {
            double f = a[r][c];
            if (f != 0.0) for (int j = 0; j < 2 * D; ++j) a[r][j] -= f * a[c][j];
        }
    }
    for (int i = 0; i < D; ++i) for (int j = 0; j < D; ++j) inv[i * D + j] = a[i][D + j];
    return det;
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/*
 * One VB run.  Layouts as in include/melissa_b200.h: w0 / W dim c(M, D, K)
 * column-major, r_nk dim c(N, K) column-major, W_Sigma block (k, m) D x D.
 * Returns the number of lower bounds written (lb[0] = -Inf).
 */
int oracle_vb_run(int N, int M, int64_t P, const int64_t *pair_ptr, const int32_t *pair_cell,
                  const int32_t *pair_region, const double *x, const uint8_t *y, int n_basis,
                  const double *mus, double gamma, int K, const double *w0, const double *delta_0,
                  double alpha_0, double beta_0, int vb_max_iter, double epsilon_conv, int lb_every,
                  int n_threads, double *W, double *W_Sigma, double *r_out, double *delta_out,
                  double *beta_out, double *pi_out, double *lb, int lb_cap, int *n_iter_out, int *flags_out,
                  double *loop_seconds) {
    const int D = n_basis + 1;
    const int64_t nnz = pair_ptr[P];
    if (D > MAXD) return -1;
#ifdef _OPENMP
    if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
    /* index lists: region_ind (:119) is the cell-major pair order, cell_ind (:121) the region-major one */
    int64_t *cell_ptr = calloc((size_t)N + 1, sizeof(int64_t));
    int64_t *reg_ptr = calloc((size_t)M + 1, sizeof(int64_t));
    int64_t *reg_pair = malloc((size_t)P * sizeof(int64_t));
    for (int64_t p = 0; p < P; ++p) { cell_ptr[pair_cell[p] + 1]++; reg_ptr[pair_region[p] + 1]++; }
    for (int n = 0; n < N; ++n) cell_ptr[n + 1] += cell_ptr[n];
    for (int m = 0; m < M; ++m) reg_ptr[m + 1] += reg_ptr[m];
    {
        int64_t *cur = malloc((size_t)M * sizeof(int64_t));
        memcpy(cur, reg_ptr, (size_t)M * sizeof(int64_t));
        for (int64_t p = 0; p < P; ++p) reg_pair[cur[pair_region[p]]++] = p;
        free(cur);
    }
    double *H = malloc((size_t)nnz * D * sizeof(double));          /* R/utils.R:64-75 */
    double *HH = calloc((size_t)P * D * D, sizeof(double));        /* :254-255 */
    double *mu = malloc((size_t)nnz * sizeof(double));
    double *Ez = malloc((size_t)nnz * sizeof(double));
    double *Ezz = malloc((size_t)nnz * sizeof(double));
    double *HtEz = malloc((size_t)P * D * sizeof(double));
    double *m_k = malloc((size_t)M * D * K * sizeof(double));      /* [m][k][d] */
    double *S_k = malloc((size_t)M * K * D * D * sizeof(double));  /* [m][k][D*D] */
    double *mkSk = malloc((size_t)M * K * D * D * sizeof(double));
    double *r = calloc((size_t)N * K, sizeof(double));
    double *log_r = calloc((size_t)N * K, sizeof(double));
    double *Eww_m = malloc((size_t)M * sizeof(double));
    double *logdet_m = malloc((size_t)M * sizeof(double));
    double beta_k[64], delta_k[64], e_log_pi[64], E_ww[64], pi_k[64], colsum[64];
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < P; ++p) {
        double *hh = HH + p * D * D;
        for (int64_t i = pair_ptr[p]; i < pair_ptr[p + 1]; ++i) {
            double *h = H + i * D;
            h[0] = 1.0;
            for (int j = 1; j < D; ++j) { double d = x[i] - mus[j - 1]; h[j] = exp(-gamma * d * d); }
            for (int a = 0; a < D; ++a) for (int b = 0; b < D; ++b) hh[a * D + b] += h[a] * h[b];
        }
    }
    const double alpha_k = alpha_0 + (double)M * D / 2.0;          /* :264 */
    if (isnan(beta_0)) beta_0 = sqrt(alpha_k);                     /* :271-273 */
    double sum_delta = 0.0;
    for (int k = 0; k < K; ++k) { beta_k[k] = beta_0; delta_k[k] = delta_0[k]; E_ww[k] = 0.0; pi_k[k] = 0.0; sum_delta += delta_0[k]; }
    for (int k = 0; k < K; ++k) e_log_pi[k] = digamma_pos(delta_k[k]) - digamma_pos(sum_delta);   /* :278 */
    for (int m = 0; m < M; ++m)
        for (int k = 0; k < K; ++k) {
            double *mv = m_k + ((size_t)m * K + k) * D, *S = S_k + ((size_t)m * K + k) * D * D, *C = mkSk + ((size_t)m * K + k) * D * D;
            for (int d = 0; d < D; ++d) mv[d] = w0[m + (size_t)M * (d + (size_t)D * k)];
            for (int a = 0; a < D; ++a) for (int b = 0; b < D; ++b) {
                S[a * D + b] = (a == b) ? 1.0 / 50.0 : 0.0;                                         /* :268-269 */
                C[a * D + b] = mv[a] * mv[b] + S[a * D + b];                                        /* :279-280 */
            }
        }
    /* mu from rowMeans(m_k[m,,]) (:283-293), E_z, E_zz (:295-302) */
#pragma omp parallel for schedule(static)
    for (int64_t p = 0; p < P; ++p) {
        const int m = pair_region[p];
        double wbar[MAXD];
        for (int d = 0; d < D; ++d) {
            double s = 0.0;
            for (int k = 0; k < K; ++k) s += m_k[((size_t)m * K + k) * D + d];
            wbar[d] = s / K;
        }
        for (int64_t i = pair_ptr[p]; i < pair_ptr[p + 1]; ++i) {
            double s = 0.0;
            for (int d = 0; d < D; ++d) s += H[i * D + d] * wbar[d];
            mu[i] = s; Ez[i] = update_Ez(s, y[i]); Ezz[i] = 1.0 + s * Ez[i];
        }
    }
    int n_lb = 1, flags = 0, it = 1;
    struct timespec ts0, ts1;
    clock_gettime(CLOCK_MONOTONIC, &ts0);                          /* the loop alone is timed (set-up excluded) */
    if (lb_cap > 0) lb[0] = -INFINITY;                             /* :242 */
    double lb_prev = -INFINITY;
    for (int i = 2; i <= vb_max_iter; ++i) {                       /* :309 */
        it = i;
        /* ---- E-step :322-331, over cells */
#pragma omp parallel for schedule(dynamic, 8)
        for (int n = 0; n < N; ++n) {
            double log_rho[64];
            for (int k = 0; k < K; ++k) log_rho[k] = 0.0;
            for (int64_t p = cell_ptr[n]; p < cell_ptr[n + 1]; ++p) {
                const int m = pair_region[p];
                double g[MAXD], ezz2 = 0.0;
                for (int d = 0; d < D; ++d) g[d] = 0.0;
                for (int64_t c = pair_ptr[p]; c < pair_ptr[p + 1]; ++c) {
                    for (int d = 0; d < D; ++d) g[d] += H[c * D + d] * Ez[c];
                    ezz2 += Ezz[c] * Ezz[c];
                }
                for (int d = 0; d < D; ++d) HtEz[p * D + d] = g[d];
                for (int k = 0; k < K; ++k) {
                    const double *mv = m_k + ((size_t)m * K + k) * D, *C = mkSk + ((size_t)m * K + k) * D * D, *hh = HH + p * D * D;
                    double lin = 0.0, tr = 0.0;
                    for (int d = 0; d < D; ++d) lin += mv[d] * g[d];
                    for (int a = 0; a < D; ++a) for (int b = 0; b < D; ++b) tr += hh[a * D + b] * C[b * D + a];
                    log_rho[k] += -0.5 * ezz2 + lin - 0.5 * tr;
                }
            }
            double mx = -INFINITY, s = 0.0;
            for (int k = 0; k < K; ++k) { log_rho[k] += e_log_pi[k]; if (log_rho[k] > mx) mx = log_rho[k]; }
            for (int k = 0; k < K; ++k) s += exp(log_rho[k] - mx);
            const double lse = log(s) + mx;                        /* R/utils.R:16-20 */
            for (int k = 0; k < K; ++k) { log_r[(size_t)n * K + k] = log_rho[k] - lse; r[(size_t)n * K + k] = exp(log_rho[k] - lse); }
        }
        /* ---- M-step :336-392 */
        for (int k = 0; k < K; ++k) {
            double s = 0.0;
            for (int n = 0; n < N; ++n) s += r[(size_t)n * K + k];
            colsum[k] = s;
            delta_k[k] = delta_0[k] + s;
            pi_k[k] = (delta_0[k] + s) / (K * delta_0[k] + N);
        }
        for (int k = 0; k < K; ++k) {
            const double ridge = alpha_k / beta_k[k] + 1e-7;
#pragma omp parallel for schedule(dynamic, 4)
            for (int m = 0; m < M; ++m) {
                double A[MAXD * MAXD], b[MAXD], Sm[MAXD * MAXD];
                for (int a = 0; a < D * D; ++a) A[a] = 0.0;
                for (int d = 0; d < D; ++d) { b[d] = 0.0; A[d * D + d] = ridge; }
                for (int64_t q = reg_ptr[m]; q < reg_ptr[m + 1]; ++q) {
                    const int64_t p = reg_pair[q];
                    const double rk = r[(size_t)pair_cell[p] * K + k];
                    for (int a = 0; a < D * D; ++a) A[a] += HH[p * D * D + a] * rk;
                    for (int d = 0; d < D; ++d) b[d] += HtEz[p * D + d] * rk;
                }
                const double det = inv_det(D, A, Sm);
                double *mv = m_k + ((size_t)m * K + k) * D, *S = S_k + ((size_t)m * K + k) * D * D;
                double eww = 0.0;
                for (int a = 0; a < D; ++a) {
                    double s = 0.0;
                    for (int c = 0; c < D; ++c) s += Sm[a * D + c] * b[c];
                    mv[a] = s;
                }
                for (int a = 0; a < D; ++a) { eww += mv[a] * mv[a] + Sm[a * D + a]; }
                memcpy(S, Sm, sizeof(double) * D * D);
                Eww_m[m] = eww;
                if (k == 0) logdet_m[m] = 0.0;
                logdet_m[m] += log(1.0 / det);                     /* det(S) = 1 / det(A) */
            }
            double s = 0.0;
            for (int m = 0; m < M; ++m) s += Eww_m[m];
            E_ww[k] = s;
            beta_k[k] = beta_0 + 0.5 * s;                          /* :386 */
            if (beta_k[k] > 10 * alpha_k) beta_k[k] = 10 * alpha_k; /* :390-392 */
        }
        /* ---- refresh :418-435, mk_Sk :450-451, e_log_pi :443 */
#pragma omp parallel for schedule(static)
        for (int64_t mk = 0; mk < (int64_t)M * K; ++mk) {
            const double *mv = m_k + mk * D, *S = S_k + mk * D * D;
            double *C = mkSk + mk * D * D;
            for (int a = 0; a < D; ++a) for (int b = 0; b < D; ++b) C[a * D + b] = mv[a] * mv[b] + S[a * D + b];
        }
        sum_delta = 0.0;
        for (int k = 0; k < K; ++k) sum_delta += delta_k[k];
        for (int k = 0; k < K; ++k) e_log_pi[k] = digamma_pos(delta_k[k]) - digamma_pos(sum_delta);
        const int with_lb = (i % lb_every == 0) || (i == vb_max_iter);   /* :458 */
        double lb_pz_qz = 0.0;
#pragma omp parallel for schedule(dynamic, 8) reduction(+ : lb_pz_qz)
        for (int n = 0; n < N; ++n) {
            for (int64_t p = cell_ptr[n]; p < cell_ptr[n + 1]; ++p) {
                const int m = pair_region[p];
                double wbar[MAXD];
                for (int d = 0; d < D; ++d) {
                    double s = 0.0;
                    for (int k = 0; k < K; ++k) s += r[(size_t)n * K + k] * m_k[((size_t)m * K + k) * D + d];
                    wbar[d] = s;
                }
                double acc = 0.0;
                for (int64_t c = pair_ptr[p]; c < pair_ptr[p + 1]; ++c) {
                    double s = 0.0;
                    for (int d = 0; d < D; ++d) s += H[c * D + d] * wbar[d];
                    mu[c] = s; Ez[c] = update_Ez(s, y[c]); Ezz[c] = 1.0 + s * Ez[c];
                    if (with_lb) {
                        const double pm = pnorm(-s);
                        acc += 0.5 * s * s + (y[c] ? log(1.0 - (pm - 1e-10)) : log(pm + 1e-10));   /* :470-472 */
                    }
                }
                if (with_lb) {
                    for (int k = 0; k < K; ++k) {
                        const double *C = mkSk + ((size_t)m * K + k) * D * D, *hh = HH + p * D * D;
                        double tr = 0.0;
                        for (int a = 0; a < D; ++a) for (int b = 0; b < D; ++b) tr += hh[a * D + b] * C[b * D + a];
                        acc -= 0.5 * r[(size_t)n * K + k] * tr;
                    }
                    lb_pz_qz += acc;
                }
            }
        }
        if (with_lb) {                                             /* :475-492 */
            const double LOG2PI = log(2.0 * M_PI), MD = (double)M * D, dga = digamma_pos(alpha_k);
            double lb_p_w = 0, lb_p_c = 0, lb_p_pi = 0, lb_p_tau = 0, lb_q_c = 0, lb_q_pi = 0, lb_q_w = 0, lb_q_tau = 0;
            double sd0 = 0, sdk = 0, sl0 = 0, slk = 0;
            for (int k = 0; k < K; ++k) {
                const double dg = dga - log(beta_k[k]), Ea = alpha_k / beta_k[k];
                lb_p_w += -0.5 * MD * LOG2PI + 0.5 * MD * dg - 0.5 * Ea * E_ww[k];
                lb_p_c += colsum[k] * e_log_pi[k];
                lb_p_pi += (delta_0[k] - 1) * e_log_pi[k];
                lb_q_pi += (delta_k[k] - 1) * e_log_pi[k];
                sd0 += delta_0[k]; sdk += delta_k[k]; sl0 += lgamma(delta_0[k]); slk += lgamma(delta_k[k]);
                lb_p_tau += alpha_0 * log(beta_0) + (alpha_0 - 1) * dg - beta_0 * Ea - lgamma(alpha_0);
                lb_q_tau += -lgamma(alpha_k) + (alpha_k - 1) * dga + log(beta_k[k]) - alpha_k;
            }
            lb_p_pi += lgamma(sd0) - sl0;
            lb_q_pi += lgamma(sdk) - slk;
            for (int64_t j = 0; j < (int64_t)N * K; ++j) lb_q_c += r[j] * log_r[j];
            for (int m = 0; m < M; ++m) lb_q_w += -0.5 * logdet_m[m];
            lb_q_w += -0.5 * D * (1.0 + LOG2PI) * (double)M * K;
            const double v = lb_pz_qz + lb_p_c + lb_p_pi + lb_p_w + lb_p_tau - lb_q_c - lb_q_pi - lb_q_w - lb_q_tau;
            if (n_lb < lb_cap) lb[n_lb] = v;
            ++n_lb;
            if (v < lb_prev) flags |= 1;                           /* :495-497 */
            const int conv = fabs(v - lb_prev) < epsilon_conv;     /* :499-501 */
            lb_prev = v;
            if (conv) break;
        }
        if (i == vb_max_iter) flags |= 2;                          /* :511-513 */
    }
    clock_gettime(CLOCK_MONOTONIC, &ts1);
    if (loop_seconds) *loop_seconds = (double)(ts1.tv_sec - ts0.tv_sec) + 1e-9 * (double)(ts1.tv_nsec - ts0.tv_nsec);
    if (W) for (int m = 0; m < M; ++m) for (int k = 0; k < K; ++k) for (int d = 0; d < D; ++d)
        W[m + (size_t)M * (d + (size_t)D * k)] = m_k[((size_t)m * K + k) * D + d];
    if (W_Sigma) for (int k = 0; k < K; ++k) for (int m = 0; m < M; ++m)
        memcpy(W_Sigma + ((size_t)k * M + m) * D * D, S_k + ((size_t)m * K + k) * D * D, sizeof(double) * D * D);
    if (r_out) for (int n = 0; n < N; ++n) for (int k = 0; k < K; ++k) r_out[n + (size_t)N * k] = r[(size_t)n * K + k];
    for (int k = 0; k < K; ++k) {
        if (delta_out) delta_out[k] = delta_k[k];
        if (beta_out) beta_out[k] = beta_k[k];
        if (pi_out) pi_out[k] = pi_k[k];
    }
    if (n_iter_out) *n_iter_out = it;
    if (flags_out) *flags_out = flags;
    free(cell_ptr); free(reg_ptr); free(reg_pair); free(H); free(HH); free(mu); free(Ez); free(Ezz); free(HtEz);
    free(m_k); free(S_k); free(mkSk); free(r); free(log_r); free(Eww_m); free(logdet_m);
    return n_lb;
}
