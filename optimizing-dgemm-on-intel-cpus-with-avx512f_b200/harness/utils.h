/* utils.h — harness utilities, same roles as the reference's utils.h:1-12 with size_t-safe signatures
 * (the reference's int element counts overflow past 46340^2). */
#ifndef DGEMM_B200_HARNESS_UTILS_H
#define DGEMM_B200_HARNESS_UTILS_H
#include <stdbool.h>
#include <stddef.h>

double get_sec(void);                                              /* reference utils.c:47-51 */
void randomize_matrix(double *A, size_t m, size_t n, unsigned seed); /* utils.c:53-62, but seeded (fixes F7) */
void randomize_matrix_signed(double *A, size_t m, size_t n, unsigned long long seed); /* U(-1,1), 53-bit */
void copy_matrix(const double *src, double *dest, size_t n);       /* utils.c:64-72 */
/* Sampled verification with the north-star bound |C - C_ref| <= 2*K*eps*(|alpha||A||B| + |beta*C0|):
 * `samples` entries of C are recomputed as long-double dot products. Replaces the reference's
 * verify_matrix (utils.c:74-85: absolute 1e-2 against MKL). Returns the worst |err|/bound ratio. */
double verify_sampled(int m, int n, int k, double alpha, const double *A, int lda, const double *B, int ldb,
                      double beta, const double *C0, const double *C, int ldc, int samples, unsigned seed);
void test_kernel(int kernel_num, int m, int n, int k, double alpha, double *A, double *B, double beta, double *C);
#endif
