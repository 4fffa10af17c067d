/*
 * kernel20.h — "one more kernel_number" for a kernels.h-style aggregator (reference kernels.h:1-19 includes
 * include/kernel1.h ... kernel19.h; each defines mydgemm_cpu_vNN with the 11-argument signature,
 * reference include/kernel18.h:629-640).  This header is pure C and only declares the two GPU entries that
 * libdgemm_b200.so exports; the reference-side patch that adds `case 20` / `case 21` to utils.c's switch
 * (utils.c:167-191) is shown in INTEGRATION.md.
 *
 *   kernel 20  mydgemm_gpu_v20   one B200
 *   kernel 21  mydgemm_gpu_v21   all B200s of the box (column panels of B/C, A replicated over NVLink)
 *   kernel 22..27                the optimisation ladder's six rungs (mydgemm_gpu_ladder below)
 */
#ifndef KERNEL20_H
#define KERNEL20_H
#include "dgemm_b200.h"

/* reference utils.c:159-165 shape: the harness wrappers fix LDA = m, LDB = k, LDC = m */
static inline void test_mydgemm_v20(int m, int n, int k, double alpha, double *A, double *B, double beta, double *C) {
  mydgemm_gpu_v20(m, n, k, alpha, A, m, B, k, beta, C, m);
}
static inline void test_mydgemm_v21(int m, int n, int k, double alpha, double *A, double *B, double beta, double *C) {
  mydgemm_gpu_v21(m, n, k, alpha, A, m, B, k, beta, C, m);
}

/* kernels 22 .. 27: the six rungs of the GPU optimisation ladder, selectable by number the way the reference's tutorial
 * steps are (`./dgemm_x86 1` .. `19`, reference test.c:8-19, utils.c:167-191, README.md:27-162):
 *   22 naive, one thread per entry          23 shared-memory tiles + 8x8 register tile on the DFMA pipe
 *   24 DMMA tensor cores + cp.async stages  25 + TMA / mbarrier, warp-specialised, persistent
 *   26 + stream-K tail (the product kernel) 27 rung 25 on 2-CTA clusters, A panel by TMA multicast
 * Same 11 arguments, host pointers.  A rung is a teaching step, not a pipeline: the operands are copied whole to the device
 * (leading dimensions rounded up to even so that every rung's loader is eligible), the rung runs once through
 * dgemm_b200_device_ex, C is copied back.  On failure the status is left for dgemm_b200_last_status() and C is untouched. */
static inline void mydgemm_gpu_ladder(int rung, int M, int N, int K, double alpha, double *A, int LDA, double *B, int LDB,
                                      double beta, double *C, int LDC) {
  static const int variant_of[6] = {DGEMM_B200_VARIANT_NAIVE, DGEMM_B200_VARIANT_DFMA, DGEMM_B200_VARIANT_CPASYNC16,
                                    DGEMM_B200_VARIANT_TMA, DGEMM_B200_VARIANT_TMA, DGEMM_B200_VARIANT_CLUSTER};
  static const int splits_of[6] = {1, 1, 1, 1, 0, 1}; /* 1 = whole tiles only; 0 = the heuristic (stream-K tail) */
  if (rung < 1 || rung > 6 || M <= 0 || N <= 0) return;
  const long long lda = (M + 1) & ~1, ldb = ((K > 0 ? K : 1) + 1) & ~1, ldc = lda;
  const int need_ab = K > 0 && alpha != 0.0;
  double *dA = (double *)dgemm_b200_malloc(sizeof(double) * (size_t)lda * (size_t)(need_ab ? K : 1));
  double *dB = (double *)dgemm_b200_malloc(sizeof(double) * (size_t)ldb * (size_t)N);
  double *dC = (double *)dgemm_b200_malloc(sizeof(double) * (size_t)ldc * (size_t)N);
  if (dA && dB && dC &&
      (!need_ab || (dgemm_b200_h2d(dA, lda, A, LDA, M, K, NULL) == 0 && dgemm_b200_h2d(dB, ldb, B, LDB, K, N, NULL) == 0)) &&
      (beta == 0.0 || dgemm_b200_h2d(dC, ldc, C, LDC, M, N, NULL) == 0) &&
      dgemm_b200_device_ex(M, N, K, alpha, dA, lda, dB, ldb, beta, dC, ldc, NULL, variant_of[rung - 1], splits_of[rung - 1]) == 0 &&
      dgemm_b200_d2h(C, LDC, dC, ldc, M, N, NULL) == 0)
    dgemm_b200_sync(NULL);
  dgemm_b200_free(dA); dgemm_b200_free(dB); dgemm_b200_free(dC); /* (never touches the status) */
}
static inline void test_mydgemm_ladder(int kernel_num, int m, int n, int k, double alpha, double *A, double *B, double beta,
                                       double *C) {
  mydgemm_gpu_ladder(kernel_num - 21, m, n, k, alpha, A, m, B, k, beta, C, m);
}
#endif
