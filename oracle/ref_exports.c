/*
 * ref_exports.c -- TEST INFRASTRUCTURE ONLY: the translation unit that turns the UNMODIFIED reference
 * kernels into oracle/_ref/libref.so.  It contains no reference source: it #includes the reference's
 * own aggregator header (kernels.h:1-19 pulls in include/kernel1.h ... kernel19.h) from where it lies
 * under REF_ROOT (default /root/reference) and adds three one-line exports for symbols the reference
 * declares `static inline` or only reaches through its harness.
 * Compiled by oracle/Makefile with the reference's own flags (Makefile:3): -O0 -march=skylake-avx512 -w -fopenmp.
 */
#include <omp.h>
#include <stdlib.h>
#include REF_KERNELS_H

/* partition_m_dim is static inline in include/kernel19.h:636 */
void ref_partition_m_dim(int ithr, int nthrs, int m, int *offset, int *block) {
  partition_m_dim(ithr, nthrs, m, offset, block);
}

/* mydgemm_cpu_v19 takes its team size from the OpenMP runtime (kernel19.h:695-697) */
void ref_set_num_threads(int n) { omp_set_num_threads(n); }
int ref_get_max_threads(void) { return omp_get_max_threads(); }
