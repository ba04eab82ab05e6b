/* TEST INFRASTRUCTURE.  A minimal stand-in for R's C API -- only the declarations r-package/src/melissa_R.c uses,
 * written from R's documented interface ("Writing R Extensions", section 5 and 6) -- so that the .Call glue can at
 * least be type-checked against include/melissa_b200.h on a machine without R (tests/test_r_glue.py).  It is not R:
 * nothing here is linked or run. */
#ifndef MELISSA_TEST_R_STUB_RINTERNALS_H
#define MELISSA_TEST_R_STUB_RINTERNALS_H
#include <stddef.h>
#include <stdint.h>

typedef struct SEXPREC *SEXP;
typedef ptrdiff_t R_xlen_t;
typedef int R_len_t;
typedef unsigned int SEXPTYPE;
typedef enum { FALSE = 0, TRUE } Rboolean;
#define REALSXP 14
#define VECSXP 19
extern SEXP R_NilValue;
extern double R_NaN;

int LENGTH(SEXP x);
int TYPEOF(SEXP x);
double *REAL(SEXP x);
SEXP VECTOR_ELT(SEXP x, R_xlen_t i);
SEXP SET_VECTOR_ELT(SEXP x, R_xlen_t i, SEXP v);
SEXP Rf_protect(SEXP);
void Rf_unprotect(int);
#define PROTECT(s) Rf_protect(s)
#define UNPROTECT(n) Rf_unprotect(n)
SEXP Rf_allocVector(SEXPTYPE, R_xlen_t);
SEXP Rf_allocMatrix(SEXPTYPE, int, int);
SEXP Rf_alloc3DArray(SEXPTYPE, int, int, int);
SEXP Rf_lengthgets(SEXP, R_len_t);
SEXP Rf_ScalarInteger(int);
int Rf_asInteger(SEXP);
int Rf_asLogical(SEXP);
double Rf_asReal(SEXP);
Rboolean Rf_isMatrix(SEXP);
Rboolean Rf_isNull(SEXP);
int Rf_nrows(SEXP);
void Rf_error(const char *, ...) __attribute__((noreturn));
#define allocVector Rf_allocVector
#define allocMatrix Rf_allocMatrix
#define alloc3DArray Rf_alloc3DArray
#define lengthgets Rf_lengthgets
#define ScalarInteger Rf_ScalarInteger
#define asInteger Rf_asInteger
#define asLogical Rf_asLogical
#define asReal Rf_asReal
#define isMatrix Rf_isMatrix
#define isNull Rf_isNull
#define nrows Rf_nrows
#define error Rf_error
Rboolean R_ToplevelExec(void (*fun)(void *), void *data);
#endif
