// dgemm_host.h — internal interface between the kernel launchers (dgemm_kernels.cu) and the C-ABI (dgemm_api.cu).
#pragma once
#include <cuda_runtime.h>

#include <vector>

namespace dg {

struct Params;

enum Variant {
  VARIANT_AUTO = 0,       // the TMA kernel; its parity-split variant for odd lda / ldb or 8-byte aligned bases
  VARIANT_TMA = 1,        // dgemm_tma_ws_kernel / dgemm_tma_ws_split_kernel (unaligned operands need K >= 2, N >= 2)
  VARIANT_CPASYNC8 = 2,   // dgemm_cpasync_kernel<1>: any alignment, any leading dimension
  VARIANT_CPASYNC16 = 3,  // dgemm_cpasync_kernel<2>: same eligibility as TMA
  VARIANT_NAIVE = 4,      // ladder rung 1 (dgemm_ladder.cu): never chosen by VARIANT_AUTO
  VARIANT_DFMA = 5,       // ladder rung 2: shared-memory tiles + 8x8 register tile on the CUDA-core FP64 pipe
};

bool tma_eligible(const Params& P);
int tma_split_bits(const Params& P);
int tma_mis_halves(const Params& P);
bool tma_split_ok(const Params& P);   // needs P.A, P.B, P.lda, P.ldb, P.K, P.N
cudaError_t launch_gemm(const Params& P, int variant, int num_sms, cudaStream_t stream);
// the TMA kernel on 64 x 64 tiles (dgemm_kernels_small.cu); maps = {tmA, tmA1, tmB, tmB1} (CUtensorMap*, the second of a
// pair only read when the operand's SPLIT bit is set), boxes 16 x 16 (or 16 x 8 per k parity) / 16 x 64 (16 x 32)
constexpr int SMALL_BM = 64, SMALL_BN = 64;
cudaError_t launch_tma_small(const void* const maps[4], int split, const Params& P, int grid, int dp_units,
                             cudaStream_t stream);
cudaError_t launch_ladder(const Params& P, int variant, cudaStream_t stream);
cudaError_t launch_scale_c(double* C, int M, int N, long long ldc, double beta, int num_sms, cudaStream_t stream);
// C += beta * X, both M x N column-major
cudaError_t launch_add_scaled(double* C, long long ldc, const double* X, long long ldx, int M, int N, double beta,
                              int num_sms, cudaStream_t stream);
// dst = src (rows x cols, column-major) with a new leading dimension
cudaError_t launch_realign(const double* src, long long lds, double* dst, long long ldd, int rows, int cols, int num_sms,
                           cudaStream_t stream);
// dst (cols x rows, ldd) = src^T, src rows x cols with leading dimension lds, both column-major
cudaError_t launch_transpose(const double* src, long long lds, double* dst, long long ldd, int rows, int cols,
                             cudaStream_t stream);
cudaError_t launch_fill_signed(double* X, int rows, int cols, long long ld, unsigned long long seed,
                               unsigned long long row0, unsigned long long col0, int num_sms, cudaStream_t stream);
// schedule of the host-pointer entry (dgemm_hostplan.cpp): column-panel / K-slab boundaries and the order of the
// (slab, panel) device calls; the H2D copies are queued in the order the calls first need them
struct HostPlanItem { int s, p, s_end; };  // device call: slabs [s, s_end) of K against panel p
struct HostPlan {
  std::vector<int> nb, kb;
  std::vector<HostPlanItem> order;
  int deferred_panels = 0;    // panels [0, deferred_panels): beta*C is added AFTER their last call (see dgemm_hostplan.cpp)
  double predicted_ms = 0.0;  // the planner's own estimate of the whole call (model in dgemm_hostplan.cpp)
};
HostPlan plan_host_pipeline(int M, int N, int K, bool need_ab, bool need_c_in, int num_sms, bool pageable = false);

// pageable host memory <-> HBM through per-thread pinned slots (dgemm_staging.cu)
bool host_pointer_is_pageable(const void* p);
bool staging_usable(int dev);
cudaError_t staged_h2d(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                       cudaStream_t stream, int dev);
cudaError_t staged_d2h_async(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                             cudaEvent_t ready, cudaStream_t stream, int dev);
cudaError_t staged_d2h_wait();
void staging_shutdown();

int launch_count();
void reset_launch_count();

}  // namespace dg
