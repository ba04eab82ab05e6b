/*
 * .Call glue between R and libmelissa_b200.so.  Written without an R
 * installation at hand: it has never been compiled against R itself, only type-checked against a
 * stand-in for the few R declarations it uses (tests/r_stub, tests/test_r_glue.py).  It is the only
 * file that includes R headers; the library itself is R-free.
 *
 *   R CMD SHLIB melissa_R.c -I../../include -L../../melissa_b200 -lmelissa_b200
 */
#include <R.h>
#include <Rinternals.h>
#include <R_ext/Rdynload.h>
#include <R_ext/Utils.h>
#include "melissa_b200.h"

/* User interrupts: R_CheckUserInterrupt() longjmps, which must not happen inside the library (GPU threads are
 * running).  The hook therefore tests for an interrupt inside R_ToplevelExec (no longjmp escapes it) and asks
 * the library to stop; the glue frees the handle and only then raises the R condition. */
extern void Rf_onintr(void);      /* Rinterface.h on unix-alikes: raises the interrupt condition (after the clean-up) */
static void check_interrupt(void *unused) { (void) unused; R_CheckUserInterrupt(); }
static int progress_hook(int32_t iter, double lb, double change, void *verbose) {
    if (*(int *) verbose && iter % 50 == 0)                           /* R/melissa_vb.R:503-507 */
        Rprintf("\rIt:\t%d\tLB:\t%f\tDiff:\t%f\n", iter, lb, change);
    return R_ToplevelExec(check_interrupt, NULL) == FALSE;            /* FALSE = an interrupt is pending */
}

/* X: list (cells) of lists (regions) of I x 2 REALSXP or logical NA.  Walk it twice: count, then fill. */
SEXP C_melissa_fit(SEXP X, SEXP K_, SEXP basis_M, SEXP mus, SEXP gamma, SEXP w0, SEXP delta_0,
                   SEXP alpha_0, SEXP beta_0, SEXP max_iter, SEXP eps, SEXP n_gpus_, SEXP verbose_) {
    const int N = LENGTH(X), M = LENGTH(VECTOR_ELT(X, 0)), K = asInteger(K_), nb = asInteger(basis_M), D = nb + 1;
    R_xlen_t P = 0, nnz = 0;
    for (int n = 0; n < N; ++n) for (int m = 0; m < M; ++m) {
        SEXP r = VECTOR_ELT(VECTOR_ELT(X, n), m);
        if (TYPEOF(r) == REALSXP && isMatrix(r)) { ++P; nnz += nrows(r); }
    }
    int64_t *ptr = (int64_t *) R_alloc(P + 1, sizeof(int64_t));
    int32_t *cell = (int32_t *) R_alloc(P, sizeof(int32_t)), *reg = (int32_t *) R_alloc(P, sizeof(int32_t));
    double *x = (double *) R_alloc(nnz, sizeof(double));
    uint8_t *y = (uint8_t *) R_alloc(nnz, 1);
    R_xlen_t p = 0, c = 0; ptr[0] = 0;
    for (int n = 0; n < N; ++n) for (int m = 0; m < M; ++m) {
        SEXP r = VECTOR_ELT(VECTOR_ELT(X, n), m);
        if (!(TYPEOF(r) == REALSXP && isMatrix(r))) continue;
        const int I = nrows(r); const double *v = REAL(r);          /* column-major: x = v[0..I), y = v[I..2I) */
        for (int i = 0; i < I; ++i) { x[c] = v[i]; y[c] = (uint8_t) v[I + i]; ++c; }
        cell[p] = n; reg[p] = m; ptr[++p] = c;
    }
    /* one object for all GPUs of the box the caller asked for (the role of no_cores, R/utils.R:30-43) */
    melissa_multi *mm = NULL;
    if (melissa_pack_multi(N, M, P, ptr, cell, reg, x, y, MELISSA_BASIS_RBF, nb, REAL(mus), asReal(gamma),
                           asInteger(n_gpus_), NULL, &mm))
        error("%s", melissa_last_error());                           /* e.g. "No coverage across all regions. ..." */
    const int cap = asInteger(max_iter) / 10 + 3;
    SEXP W = PROTECT(alloc3DArray(REALSXP, M, D, K)), S = PROTECT(allocVector(REALSXP, (R_xlen_t) K * M * D * D));
    SEXP r_nk = PROTECT(allocMatrix(REALSXP, N, K)), delta = PROTECT(allocVector(REALSXP, K));
    SEXP alpha = PROTECT(allocVector(REALSXP, K)), beta = PROTECT(allocVector(REALSXP, K));
    SEXP pi_k = PROTECT(allocVector(REALSXP, K)), lb = PROTECT(allocVector(REALSXP, cap));
    int n_lb = 0, n_iter = 0, flags = 0, verbose = asLogical(verbose_) == TRUE;
    const double b0 = isNull(beta_0) ? R_NaN : asReal(beta_0);       /* NaN = default sqrt(alpha_k), :271-273 */
    melissa_set_progress(progress_hook, &verbose);
    const int rc = melissa_vb_run_multi(mm, K, REAL(w0), REAL(delta_0), asReal(alpha_0), b0, asInteger(max_iter),
                                        asReal(eps), 10, REAL(W), REAL(S), REAL(r_nk), REAL(delta), REAL(alpha),
                                        REAL(beta), REAL(pi_k), REAL(lb), cap, &n_lb, &n_iter, &flags);
    melissa_set_progress(NULL, NULL);
    melissa_multi_free(mm);                                          /* free BEFORE any R error / longjmp */
    if (rc) error("%s", melissa_last_error());
    if (flags & MELISSA_FLAG_INTERRUPTED) { UNPROTECT(8); Rf_onintr(); }
    SEXP out = PROTECT(allocVector(VECSXP, 10));
    SET_VECTOR_ELT(out, 0, W); SET_VECTOR_ELT(out, 1, S); SET_VECTOR_ELT(out, 2, r_nk);
    SET_VECTOR_ELT(out, 3, delta); SET_VECTOR_ELT(out, 4, alpha); SET_VECTOR_ELT(out, 5, beta);
    SET_VECTOR_ELT(out, 6, pi_k); SET_VECTOR_ELT(out, 7, lengthgets(lb, n_lb));
    SET_VECTOR_ELT(out, 8, ScalarInteger(n_iter)); SET_VECTOR_ELT(out, 9, ScalarInteger(flags));
    UNPROTECT(9);
    return out;
}

static const R_CallMethodDef call_methods[] = {{"C_melissa_fit", (DL_FUNC) &C_melissa_fit, 13}, {NULL, NULL, 0}};

void R_init_melissaB200(DllInfo *dll) {
    R_registerRoutines(dll, NULL, call_methods, NULL, NULL);
    R_useDynamicSymbols(dll, FALSE);
}
