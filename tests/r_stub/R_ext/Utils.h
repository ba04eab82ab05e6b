/* TEST INFRASTRUCTURE: see ../Rinternals.h. */
#ifndef MELISSA_TEST_R_STUB_UTILS_H
#define MELISSA_TEST_R_STUB_UTILS_H
void R_CheckUserInterrupt(void);
#endif
