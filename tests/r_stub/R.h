/* TEST INFRASTRUCTURE: see Rinternals.h in this directory. */
#ifndef MELISSA_TEST_R_STUB_R_H
#define MELISSA_TEST_R_STUB_R_H
#include <stddef.h>
char *R_alloc(size_t n, int size);
void Rprintf(const char *, ...);
#endif
