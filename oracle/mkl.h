/* mkl.h shim -- TEST INFRASTRUCTURE ONLY. Intel MKL is not available on this box; the
 * reference's utils.c:8 and test.c:5 include "mkl.h" for cblas_dgemm (kernel 0 / the self-check oracle).
 * This declares just enough for the reference harness to compile; mkl_shim.c provides cblas_dgemm. */
#ifndef ORACLE_MKL_SHIM_H
#define ORACLE_MKL_SHIM_H
enum { CblasColMajor = 102, CblasNoTrans = 111 };
void cblas_dgemm(int layout, int transa, int transb, int m, int n, int k, double alpha, const double *A, int lda,
                 const double *B, int ldb, double beta, double *C, int ldc);
#endif
