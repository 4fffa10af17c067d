/*
 * kernel20.h — "one more kernel_number" for a kernels.h-style aggregator (reference kernels.h:1-19 includes
 * include/kernel1.h ... kernel19.h; each defines mydgemm_cpu_vNN with the 11-argument signature,
 * reference include/kernel18.h:629-640).  This header is pure C and only declares the two GPU entries that
 * libdgemm_b200.so exports; the reference-side patch that adds `case 20` / `case 21` to utils.c's switch
 * (utils.c:167-191) is shown in INTEGRATION.md.
 *
 *   kernel 20  mydgemm_gpu_v20   one B200
 *   kernel 21  mydgemm_gpu_v21   all B200s of the box (column panels of B/C, A replicated over NVLink)
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
#endif
