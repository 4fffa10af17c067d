/*
 * mkl.h stand-in that serves cblas_dgemm from the B200 library.
 *
 * The reference includes "mkl.h" (test.c:5, utils.c:8) for exactly one function: cblas_dgemm, used as kernel 0
 * (utils.c:87-89) and as the verification oracle of every other kernel (test.c:38).  Put this directory on the include
 * path instead of MKL's and link libdgemm_b200.so: the UNMODIFIED reference harness then runs kernel 0 — and checks
 * its own kernels 1..19 — against the FP64 tensor-core path (`make -C oracle _ref/dgemm_x86_b200`).
 * Only the CBLAS names the reference uses, with CBLAS's numeric values.
 */
#ifndef DGEMM_B200_MKL_STANDIN_H
#define DGEMM_B200_MKL_STANDIN_H
#include "dgemm_b200.h"

typedef enum { CblasRowMajor = 101, CblasColMajor = 102 } CBLAS_LAYOUT;
typedef enum { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 } CBLAS_TRANSPOSE;

static inline void cblas_dgemm(CBLAS_LAYOUT layout, CBLAS_TRANSPOSE transa, CBLAS_TRANSPOSE transb, int m, int n, int k,
                               double alpha, const double *A, int lda, const double *B, int ldb, double beta, double *C,
                               int ldc) {
  dgemm_b200_cblas_dgemm((int)layout, (int)transa, (int)transb, m, n, k, alpha, A, lda, B, ldb, beta, C, ldc);
}
#endif
