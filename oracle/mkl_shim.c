/* mkl_shim.c -- TEST INFRASTRUCTURE ONLY. Stand-in for MKL's cblas_dgemm (column-major, NoTrans/NoTrans) so the
 * reference's own harness (test.c:38) links on a box without MKL: forwards to the plain-C oracle. */
#include "mkl.h"
void oracle_dgemm_k18(int, int, int, double, const double *, int, const double *, int, double, double *, int);
void cblas_dgemm(int layout, int transa, int transb, int m, int n, int k, double alpha, const double *A, int lda,
                 const double *B, int ldb, double beta, double *C, int ldc) {
  (void)layout; (void)transa; (void)transb;
  oracle_dgemm_k18(m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
}
