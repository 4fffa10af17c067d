/*
 * dgemm_oracle.c -- TEST INFRASTRUCTURE ONLY.  Not product code.
 *
 * A plain-C restatement of the arithmetic of the reference's DGEMM hot path
 * (yzhaiustc/Optimizing-DGEMM-on-Intel-CPUs-with-AVX512F, include/kernel18.h and
 * include/kernel19.h; citations below are file:line inside /root/reference).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this file's shared object; the product (csrc/, libdgemm_b200.so)
 * never links or calls it.
 *
 * PINNED: the reference keeps no golden vectors for this path, so the
 * oracle is pinned against the reference ITSELF: oracle/_ref/libref.so is compiled from
 * /root/reference/kernels.h by oracle/Makefile and tests/test_oracle_vs_ref.py requires
 * oracle_dgemm_k18() to be BIT-IDENTICAL to mydgemm_cpu_v18 / mydgemm_cpu_v19 on every
 * shape inside their valid domain (even M, even N, ...), and oracle_dgemm_k1() to be
 * bit-identical to mydgemm_cpu_v1.  Small input/output fixtures produced by the
 * reference kernels are committed under tests/golden/ (tests/golden/make_golden.py).
 *
 * Per-entry arithmetic of kernel18 that is restated here:
 *   1. C(i,j) <- beta*C(i,j) for the whole matrix, unless beta == 1.0     (kernel18.h:644, 12-46)
 *   2. return if alpha == 0 or K == 0                                      (kernel18.h:645)
 *   3. K is cut into blocks of K_BLOCKING = 384                            (kernel18.h:9, 654-655)
 *   4. inside a block A is first rounded to a~(i,k) = fl(alpha*A(i,k))     (packing_a_k18, kernel18.h:69-110)
 *   5. acc <- 0; for k ascending: acc <- fma(a~(i,k), B(k,j), acc)          (vfmadd231pd, kernel18.h:125-170;
 *                                                                           K loop 362-379, sequential in k)
 *   6. C(i,j) <- fl(C(i,j) + acc)                                          (vaddpd in SAVE_*, kernel18.h:262-327)
 * The N- and M-blocking (8640 / 192) and the 24x8 register tile only change which entries
 * are processed together, never the order of operations on one entry, so they are not
 * restated.  kernel19 performs exactly the same per-entry sequence (K-outer loop,
 * kernel19.h:713-745) but returns BEFORE the beta scaling when alpha==0 or K==0
 * (kernel19.h:693); oracle_dgemm_k19 keeps that quirk so it can be checked.
 *
 * Unlike the reference the restatement handles odd M / N and uses
 * 64-bit indexing; on the reference's valid domain the results are bit-identical.
 *
 * Build: see oracle/Makefile (gcc -O2 -mfma -ffp-contract=off -fopenmp).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_K_BLOCKING 384 /* kernel18.h:9, kernel19.h:9 */

#define IDX(i, j, ld) ((size_t)(i) + (size_t)(j) * (size_t)(ld))

/* C *= beta over the M x N window; rows M..LDC-1 are never touched (scale_c_k18, kernel18.h:12-46). */
static void oracle_scale_c(double *C, int M, int N, int LDC, double beta) {
#pragma omp parallel for schedule(static)
  for (int j = 0; j < N; j++)
    for (int i = 0; i < M; i++) C[IDX(i, j, LDC)] *= beta;
}

/* steps 3-6 above for rows [m0, m0+mb) of C */
static void oracle_accumulate(int m0, int mb, int N, int K, double alpha, const double *A, int LDA, const double *B,
                              int LDB, double *C, int LDC) {
  if (mb <= 0 || N <= 0) return;
  double *apack = (double *)malloc(sizeof(double) * (size_t)mb * ORACLE_K_BLOCKING);
  for (int k0 = 0; k0 < K; k0 += ORACLE_K_BLOCKING) {
    int kb = K - k0 > ORACLE_K_BLOCKING ? ORACLE_K_BLOCKING : K - k0;
    /* step 4: alpha is rounded into A once per block (packing_a_k18) */
#pragma omp parallel for schedule(static)
    for (int k = 0; k < kb; k++)
      for (int i = 0; i < mb; i++) apack[(size_t)k * mb + i] = A[IDX(m0 + i, k0 + k, LDA)] * alpha;
    /* 8x4 register tile purely for speed: every entry still sees acc<-0, fma over ascending k, C+=acc */
#pragma omp parallel for schedule(dynamic, 4)
    for (int j0 = 0; j0 < N; j0 += 4) {
      const int jb = N - j0 < 4 ? N - j0 : 4;
      for (int i0 = 0; i0 < mb; i0 += 8) {
        const int ib = mb - i0 < 8 ? mb - i0 : 8;
        double acc[4][8];
        for (int jj = 0; jj < 4; jj++)
          for (int ii = 0; ii < 8; ii++) acc[jj][ii] = 0.0;
        if (ib == 8 && jb == 4) {
          const double *b0 = B + IDX(k0, j0, LDB), *b1 = b0 + LDB, *b2 = b1 + LDB, *b3 = b2 + LDB;
          for (int k = 0; k < kb; k++) {
            const double *ap = apack + (size_t)k * mb + i0;
            for (int ii = 0; ii < 8; ii++) acc[0][ii] = fma(ap[ii], b0[k], acc[0][ii]); /* step 5 */
            for (int ii = 0; ii < 8; ii++) acc[1][ii] = fma(ap[ii], b1[k], acc[1][ii]);
            for (int ii = 0; ii < 8; ii++) acc[2][ii] = fma(ap[ii], b2[k], acc[2][ii]);
            for (int ii = 0; ii < 8; ii++) acc[3][ii] = fma(ap[ii], b3[k], acc[3][ii]);
          }
        } else {
          for (int k = 0; k < kb; k++) {
            const double *ap = apack + (size_t)k * mb + i0;
            for (int jj = 0; jj < jb; jj++) {
              const double b = B[IDX(k0 + k, j0 + jj, LDB)];
              for (int ii = 0; ii < ib; ii++) acc[jj][ii] = fma(ap[ii], b, acc[jj][ii]);
            }
          }
        }
        for (int jj = 0; jj < jb; jj++) {
          double *c = C + IDX(m0 + i0, j0 + jj, LDC);
          for (int ii = 0; ii < ib; ii++) c[ii] += acc[jj][ii]; /* step 6 */
        }
      }
    }
  }
  free(apack);
}

/* Restates mydgemm_cpu_v18 (kernel18.h:629-671). */
void oracle_dgemm_k18(int M, int N, int K, double alpha, const double *A, int LDA, const double *B, int LDB, double beta,
                      double *C, int LDC) {
  if (M <= 0 || N <= 0) return; /* macro kernel guard, kernel18.h:605 */
  if (beta != 1.0) oracle_scale_c(C, M, N, LDC, beta);
  if (alpha == 0.0 || K == 0) return;
  oracle_accumulate(0, M, N, K, alpha, A, LDA, B, LDB, C, LDC);
}

/* Restates partition_m_dim (kernel19.h:636-656): row band of thread ithr out of nthrs. */
void oracle_partition_m_dim(int ithr, int nthrs, int m, int *offset, int *block) {
  int band = m / nthrs;
  int tail = m - (nthrs - 1) * band;
  if (tail > band + 1) band++;
  tail = m - (nthrs - 1) * band;
  *block = ithr < nthrs - 1 ? band : tail;
  *offset = ithr * band;
  if (*offset >= m) {
    *block = 0;
    *offset = 0;
  } else if (*offset + *block > m) {
    *block = m - *offset;
  }
}

/* Restates mydgemm_cpu_v19 (kernel19.h:678-755) for a team of nthr threads: the early return precedes
 * the beta scaling (F4), every thread scales and accumulates only its own row band. Bands are walked
 * one after another here (the arithmetic on an entry does not depend on which thread owns it). */
void oracle_dgemm_k19(int M, int N, int K, double alpha, const double *A, int LDA, const double *B, int LDB, double beta,
                      double *C, int LDC, int nthr) {
  if (alpha == 0.0 || K == 0) return; /* kernel19.h:693 */
  if (nthr < 1) nthr = 1;
  for (int t = 0; t < nthr; t++) {
    int off, blk;
    oracle_partition_m_dim(t, nthr, M, &off, &blk);
    if (blk <= 0) continue;
    if (beta != 1.0) oracle_scale_c(C + off, blk, N, LDC, beta); /* kernel19.h:702 */
    oracle_accumulate(off, blk, N, K, alpha, A, LDA, B, LDB, C, LDC);
  }
}

/* Restates mydgemm_cpu_v1 (kernel1.h:14-24): the reference's any-shape kernel. No FMA: the reference is
 * built -O0, so alpha*A(i,k)*B(k,j) is two rounded multiplies and a rounded add. */
void oracle_dgemm_k1(int M, int N, int K, double alpha, const double *A, int LDA, const double *B, int LDB, double beta,
                     double *C, int LDC) {
  if (beta != 1.0) oracle_scale_c(C, M, N, LDC, beta);
#pragma omp parallel for schedule(static)
  for (int j = 0; j < N; j++)
    for (int i = 0; i < M; i++) {
      double c = C[IDX(i, j, LDC)];
      for (int k = 0; k < K; k++) {
        volatile double t = alpha * A[IDX(i, k, LDA)];
        volatile double u = t * B[IDX(k, j, LDB)];
        c = c + u;
      }
      C[IDX(i, j, LDC)] = c;
    }
}

/* S(i,j) = |alpha| * sum_k |A(i,k)||B(k,j)| + |beta*C0(i,j)| : the magnitude term of the parity bound
 * |C - C_ref| <= c*K*eps*S. S is M x N with leading dimension M. */
void oracle_bound_term(int M, int N, int K, double alpha, const double *A, int LDA, const double *B, int LDB,
                       double beta, const double *C0, int LDC, double *S) {
  double aa = fabs(alpha);
#pragma omp parallel
  {
    double *acc = (double *)malloc(sizeof(double) * (size_t)(M > 0 ? M : 1));
#pragma omp for schedule(static)
    for (int j = 0; j < N; j++) {
      for (int i = 0; i < M; i++) acc[i] = 0.0;
      for (int k = 0; k < K; k++) {
        double b = fabs(B[IDX(k, j, LDB)]);
        const double *a = A + IDX(0, k, LDA);
        for (int i = 0; i < M; i++) acc[i] += fabs(a[i]) * b;
      }
      for (int i = 0; i < M; i++) {
        double c0 = (beta == 0.0) ? 0.0 : fabs(beta * C0[IDX(i, j, LDC)]);
        S[IDX(i, j, M)] = aa * acc[i] + c0;
      }
    }
    free(acc);
  }
}

/* max over the M x N window of |X - Y| / (K*eps*S) (eps = 2^-52); also returns max |X - Y|.
 * Entries with S == 0 must match exactly (ratio = inf otherwise). */
double oracle_max_bound_ratio(int M, int N, int K, const double *X, int LDX, const double *Y, int LDY, const double *S,
                              double *max_abs_err) {
  const double eps = 2.220446049250313e-16;
  double worst = 0.0, worst_abs = 0.0;
  double kk = K > 0 ? (double)K : 1.0;
  for (int j = 0; j < N; j++)
    for (int i = 0; i < M; i++) {
      double d = fabs(X[IDX(i, j, LDX)] - Y[IDX(i, j, LDY)]);
      if (d != d) { worst = INFINITY; worst_abs = INFINITY; continue; }
      if (d > worst_abs) worst_abs = d;
      double s = kk * eps * S[IDX(i, j, M)];
      double r = d == 0.0 ? 0.0 : (s > 0.0 ? d / s : INFINITY);
      if (r > worst) worst = r;
    }
  if (max_abs_err) *max_abs_err = worst_abs;
  return worst;
}

/* ---- deterministic input distributions ------------------------------------------------- */

/* D1, reference style: x = (rand()%100) + 0.01*(rand()%100), third draw discarded -- utils.c:58-59 -- but
 * seeded ONCE per matrix with a caller-chosen seed (fixes F7: the reference reseeds with time(NULL)).
 * Fills n doubles in storage order exactly as randomize_matrix does. */
void oracle_fill_reference_style(double *X, size_t n, unsigned seed) {
  srand(seed);
  for (size_t i = 0; i < n; i++) {
    X[i] = (double)(rand() % 100) + 0.01 * (rand() % 100);
    (void)rand();
  }
}

static inline uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

/* D2, signed: counter-based U(-1,1) with 53-bit mantissas, x(i,j) = f(seed, i + j*2^32). The value of an
 * entry depends only on (seed, i, j), so any window / panel can be generated independently; the device
 * fill kernel in csrc/ uses the same formula. */
double oracle_signed_value(uint64_t seed, uint64_t i, uint64_t j) {
  uint64_t h = splitmix64(seed * 0xD1342543DE82EF95ull + (i + (j << 32)));
  return (double)(int64_t)(h >> 11) * (1.0 / 4503599627370496.0) - 1.0; /* (h>>11) in [0,2^53) -> [-1,1) */
}

void oracle_fill_signed(double *X, int rows, int cols, int ld, uint64_t seed, uint64_t row0, uint64_t col0) {
#pragma omp parallel for schedule(static)
  for (int j = 0; j < cols; j++)
    for (int i = 0; i < rows; i++) X[IDX(i, j, ld)] = oracle_signed_value(seed, row0 + i, col0 + j);
}

int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
