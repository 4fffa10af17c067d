/* utils.c — harness utilities and the kernel_number dispatch for the GPU kernels.
 * Same roles as the reference's utils.c (get_sec :47-51, randomize_matrix :53-62, copy_matrix :64-72,
 * verify_matrix :74-85, test_mydgemm_vNN :91-165, test_kernel :167-191); written from scratch, size_t-safe,
 * deterministic, and with a relative error bound instead of the absolute 1e-2. */
#include "utils.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <sys/time.h>

#include "kernel20.h"

double get_sec(void) {
  struct timeval t;
  gettimeofday(&t, NULL);
  return t.tv_sec + 1e-6 * t.tv_usec;
}

/* the reference's value distribution: two-decimal non-negative numbers, third draw discarded */
void randomize_matrix(double *A, size_t m, size_t n, unsigned seed) {
  srand(seed);
  for (size_t i = 0; i < m * n; i++) {
    A[i] = (double)(rand() % 100) + 0.01 * (rand() % 100);
    (void)rand();
  }
}

static unsigned long long mix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

void randomize_matrix_signed(double *A, size_t m, size_t n, unsigned long long seed) {
  for (size_t j = 0; j < n; j++)
    for (size_t i = 0; i < m; i++) {
      unsigned long long h = mix64(seed * 0xD1342543DE82EF95ull + (i + ((unsigned long long)j << 32)));
      A[i + j * m] = (double)(long long)(h >> 11) * (1.0 / 4503599627370496.0) - 1.0;
    }
}

void copy_matrix(const double *src, double *dest, size_t n) {
  for (size_t i = 0; i < n; i++) dest[i] = src[i];
}

double verify_sampled(int m, int n, int k, double alpha, const double *A, int lda, const double *B, int ldb,
                      double beta, const double *C0, const double *C, int ldc, int samples, unsigned seed) {
  const double eps = 2.220446049250313e-16;
  double worst = 0.0;
  unsigned long long s = seed;
  for (int t = 0; t < samples; t++) {
    s = mix64(s + t);
    int i = (int)(s % (unsigned long long)m), j = (int)((s >> 32) % (unsigned long long)n);
    if (t == 0) { i = m - 1; j = n - 1; }   /* always include the last row / column (the reference's F1 bug) */
    if (t == 1) { i = 0; j = 0; }
    long double dot = 0.0L, mag = 0.0L;
    for (int p = 0; p < k; p++) {
      long double a = A[(size_t)i + (size_t)p * lda], b = B[(size_t)p + (size_t)j * ldb];
      dot += a * b;
      mag += fabsl(a * b);
    }
    long double c0 = beta == 0.0 ? 0.0L : (long double)beta * C0[(size_t)i + (size_t)j * ldc];
    long double ref = (long double)alpha * dot + c0;
    double bound = 2.0 * (k > 0 ? k : 1) * eps * (double)(fabsl((long double)alpha) * mag + fabsl(c0));
    double err = fabs((double)(ref - (long double)C[(size_t)i + (size_t)j * ldc]));
    double ratio = err == 0.0 ? 0.0 : (bound > 0.0 ? err / bound : INFINITY);
    if (!(ratio <= worst)) worst = ratio; /* also propagates NaN */
  }
  return worst;
}

void test_kernel(int kernel_num, int m, int n, int k, double alpha, double *A, double *B, double beta, double *C) {
  switch (kernel_num) {
    case 20: test_mydgemm_v20(m, n, k, alpha, A, B, beta, C); break;
    case 21: test_mydgemm_v21(m, n, k, alpha, A, B, beta, C); break;
    case 22: case 23: case 24: case 25: case 26: case 27: /* the ladder's rungs, include/kernel20.h */
      test_mydgemm_ladder(kernel_num, m, n, k, alpha, A, B, beta, C); break;
    default: break;
  }
}
