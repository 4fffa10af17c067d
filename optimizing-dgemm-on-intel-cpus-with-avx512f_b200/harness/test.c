/* test.c — the reference's CLI harness (reference test.c:1-54) for the GPU kernel numbers.
 *
 *   dgemm_b200 <kernel_num> [max_size] [alpha beta]      kernel_num: 20 = one B200, 21 = all B200s of the box,
 *                                                        22-27 = the six rungs of the optimisation ladder
 *
 * Same flow and the same two output lines per size as the reference ("M=N=K=%d:" and "Average elasped time: %f
 * second, performance: %f GFLOPS." — test.c:35,49 — so data/data.sh-style scraping keeps working): sizes 100..3000
 * step 100, one verified call, then 3 timed calls through the 11-argument host-pointer entry (so the time INCLUDES
 * H2D/D2H, like any other kernel number), average reported with 2*m*n*k flops.  Additions: deterministic seeds,
 * sampled verification with a relative bound instead of MKL + abs 1e-2, a kernel-only line measured with CUDA
 * events on device-resident operands, a non-zero exit status when the GPU call reports an error, and — replacing the
 * reference's fixed-column scrape (data/data.sh:6, `cut -c 52-60`) — one JSON record per size appended to the file named
 * by the environment variable DGEMM_B200_JSON.
 */
#include <stdio.h>
#include <stdlib.h>

#include "kernel20.h"
#include "utils.h"

int main(int argc, char *argv[]) {
  if (argc < 2) {
    printf("Please select a kernel (20 = one B200, 21 = all B200s; the CPU kernels 0-19 live in the reference).\n");
    exit(-1);
  }
  int kernel_num = atoi(argv[1]);
  if (kernel_num < 20 || kernel_num > 27) { /* 22-27: the six rungs of the GPU optimisation ladder (include/kernel20.h) */
    printf("Please enter a valid kernel number (20-27).\n");
    exit(-2);
  }
  int max_size = argc > 2 ? atoi(argv[2]) : 3000;
  double alpha = argc > 4 ? atof(argv[3]) : 2.0, beta = argc > 4 ? atof(argv[4]) : 0.; /* test.c:25 */
  if (max_size < 100) max_size = 100;
  int N = 3;
  size_t cap = (size_t)max_size * max_size;
  /* pageable host memory, exactly like the reference harness (test.c:27-30) */
  double *A = (double *)malloc(sizeof(double) * cap), *B = (double *)malloc(sizeof(double) * cap);
  double *C = (double *)malloc(sizeof(double) * cap), *C_ref = (double *)malloc(sizeof(double) * cap);
  if (!A || !B || !C || !C_ref) { printf("host allocation failed\n"); exit(-4); }
  randomize_matrix(A, max_size, max_size, 1);
  randomize_matrix(B, max_size, max_size, 2);
  randomize_matrix(C, max_size, max_size, 3);
  copy_matrix(C, C_ref, cap);
  const char *json_path = getenv("DGEMM_B200_JSON");
  FILE *json = json_path ? fopen(json_path, "a") : NULL;
  if (dgemm_b200_device_count() < 1) {
    printf("No usable B200 (sm_100) device: kernel %d has no CPU fallback. Exited.\n", kernel_num);
    exit(-5);
  }
  for (int size = 100; size <= max_size; size += 100) {
    int m = size, n = size, k = size;
    printf("\nM=N=K=%d:\n", m);
    test_kernel(kernel_num, m, n, k, alpha, A, B, beta, C);
    if (dgemm_b200_last_status() != DGEMM_B200_OK) {
      printf("GPU kernel failed: %s. Exited.\n", dgemm_b200_last_error());
      exit(-6);
    }
    double ratio = verify_sampled(m, n, k, alpha, A, m, B, k, beta, C_ref, C, m, 512, (unsigned)size);
    if (!(ratio <= 1.0)) {
      printf("Failed to pass the correctness verification (|err|/bound = %g). Exited.\n", ratio);
      exit(-3);
    }
    double t0 = get_sec();
    for (int r = 0; r < N; r++) test_kernel(kernel_num, m, n, k, alpha, A, B, beta, C);
    double t1 = get_sec();
    printf("Average elasped time: %f second, performance: %f GFLOPS.\n", (t1 - t0) / N,
           2. * 1e-9 * N * m * n * k / (t1 - t0));
    double kernel_ms = -1.0;
    if (kernel_num == 20) { /* kernel-only: operands resident in HBM, CUDA events */
      size_t bytes = sizeof(double) * (size_t)m * m;
      double *dA = (double *)dgemm_b200_malloc(bytes), *dB = (double *)dgemm_b200_malloc(bytes),
             *dC = (double *)dgemm_b200_malloc(bytes);
      if (dA && dB && dC) {
        dgemm_b200_h2d(dA, m, A, m, m, k, NULL);
        dgemm_b200_h2d(dB, k, B, k, k, n, NULL);
        dgemm_b200_h2d(dC, m, C_ref, m, m, n, NULL);
        dgemm_b200_sync(NULL);
        double ms = dgemm_b200_time_device(m, n, k, alpha, dA, m, dB, k, beta, dC, m, DGEMM_B200_VARIANT_AUTO, 0, 2, 5,
                                           (size_t)256 << 20);
        kernel_ms = ms;
        if (ms > 0) printf("Kernel-only time: %f second, performance: %f GFLOPS (max |err|/bound %.3f).\n", ms * 1e-3,
                           2. * 1e-9 * m * n * k / (ms * 1e-3), ratio);
      }
      dgemm_b200_free(dA); dgemm_b200_free(dB); dgemm_b200_free(dC);
    }
    if (json) {
      fprintf(json, "{\"kernel\": %d, \"m\": %d, \"n\": %d, \"k\": %d, \"alpha\": %g, \"beta\": %g, "
                    "\"e2e_seconds\": %.9f, \"e2e_gflops\": %.3f, \"kernel_seconds\": %.9f, \"kernel_gflops\": %.3f, "
                    "\"max_err_over_bound\": %.4g}\n",
              kernel_num, m, n, k, alpha, beta, (t1 - t0) / N, 2. * 1e-9 * N * m * n * k / (t1 - t0),
              kernel_ms > 0 ? kernel_ms * 1e-3 : -1.0, kernel_ms > 0 ? 2. * 1e-9 * m * n * k / (kernel_ms * 1e-3) : -1.0,
              ratio);
      fflush(json);
    }
    copy_matrix(C_ref, C, (size_t)m * n); /* restore C for the next size, as test.c:50 does */
  }
  if (json) fclose(json);
  free(A); free(B); free(C); free(C_ref);
  dgemm_b200_shutdown();
  return 0;
}
