// dgemm_ladder.cu — the lower rungs of the GPU optimisation ladder, selectable as kernel variants.
//
// The reference's product is a teaching sequence: kernel1 (naive) ... kernel18 (README.md:27-162, include/kernel1.h
// ... kernel18.h).  Its Blackwell counterpart, each rung a variant of dgemm_b200_device_ex:
//   rung 1  VARIANT_NAIVE      one thread per C entry, k loop from global memory            (cf. include/kernel1.h:14-24)
//   rung 2  VARIANT_DFMA       128x128x8 shared-memory tiles, 8x8 register tile per thread,  (cf. kernel3..kernel9: register
//                              DFMA on the CUDA cores, register-staged double buffering       blocking, cache blocking, packing)
//   rung 3  VARIANT_CPASYNC16  FP64 tensor cores (DMMA) + cp.async multistage pipeline       (dgemm_kernels.cu)
//   rung 4  VARIANT_TMA, splits=1   TMA + mbarrier, warp-specialised, persistent, data-parallel tiles
//   rung 5  VARIANT_TMA, splits=0   + stream-K tail, fused NVLink fetch                          (the product path)
// Rungs 1-2 are NOT used by the product dispatch (VARIANT_AUTO never picks them); they exist for the ladder table in
// DESIGN.md and as the DFMA-vs-DMMA ablation at GEMM level.  Same semantics and edge handling as the product kernels.
#include "dgemm_device.cuh"
#include "dgemm_host.h"

namespace dg {

// ---------------------------------------------------------------------------------------------------------
// rung 1: naive
// ---------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dgemm_naive_kernel(const Params P) {
  const int i = blockIdx.x * 32 + (threadIdx.x & 31);   // row: consecutive lanes -> consecutive rows (coalesced A, C)
  const int j = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (i >= P.M || j >= P.N) return;
  const double* a = P.A + i;
  const double* b = P.B + (long long)j * P.ldb;
  double acc = 0.0;
  for (int k = 0; k < P.K; k++) acc = fma(a[(long long)k * P.lda], b[k], acc);
  double* c = P.C + (long long)j * P.ldc + i;
  *c = P.beta == 0.0 ? P.alpha * acc : fma(P.beta, *c, P.alpha * acc);
}

// ---------------------------------------------------------------------------------------------------------
// rung 2: shared-memory tiles + 8x8 register tile, DFMA
// ---------------------------------------------------------------------------------------------------------
constexpr int LBM = 128, LBN = 128, LBK = 8;

__global__ void __launch_bounds__(256) dgemm_dfma_kernel(const Params P) {
  __shared__ double As[2][LBK][LBM];       // [k][m]: column-major A tile as it lies in memory
  __shared__ double Bs[2][LBK][LBN + 4];   // [k][n]: B tile transposed on the way in (+4: fewer store conflicts)
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;  // thread grid 16 (m) x 16 (n); rows tx*4+{0..3} and 64+tx*4+{0..3}
  const int m0 = blockIdx.x * LBM, n0 = blockIdx.y * LBN;

  // global -> register staging: A tile 128x8 = 1024 doubles -> 4 per thread; B tile 8x128 -> 4 per thread
  const int a_m = tid & 127, a_k = tid >> 7;          // + 2*e for e = 0..3
  const int b_k = tid & 7, b_n = tid >> 3;            // + 32*e
  double ra[4], rb[4];
  auto load_global = [&](int k0) {
#pragma unroll
    for (int e = 0; e < 4; e++) {
      const int m = m0 + a_m, k = k0 + a_k + 2 * e;
      ra[e] = (m < P.M && k < P.K) ? P.A[(long long)k * P.lda + m] : 0.0;
      const int kk = k0 + b_k, n = n0 + b_n + 32 * e;
      rb[e] = (kk < P.K && n < P.N) ? P.B[(long long)n * P.ldb + kk] : 0.0;
    }
  };
  auto store_shared = [&](int buf) {
#pragma unroll
    for (int e = 0; e < 4; e++) {
      As[buf][a_k + 2 * e][a_m] = ra[e];
      Bs[buf][b_k][b_n + 32 * e] = rb[e];
    }
  };

  double acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.0;

  const int nk = (P.K + LBK - 1) / LBK;
  load_global(0);
  store_shared(0);
  __syncthreads();
  for (int t = 0; t < nk; t++) {
    const int buf = t & 1;
    if (t + 1 < nk) load_global((t + 1) * LBK);  // next tile's loads fly while this tile is multiplied
#pragma unroll
    for (int k = 0; k < LBK; k++) {
      double a[8], b[8];
      const double2* ap = reinterpret_cast<const double2*>(&As[buf][k][0]);
      const double2* bp = reinterpret_cast<const double2*>(&Bs[buf][k][0]);
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const double2 a0 = ap[(h * 64 + tx * 4) / 2], a1 = ap[(h * 64 + tx * 4) / 2 + 1];
        a[h * 4 + 0] = a0.x; a[h * 4 + 1] = a0.y; a[h * 4 + 2] = a1.x; a[h * 4 + 3] = a1.y;
        const double2 b0 = bp[(h * 64 + ty * 4) / 2], b1 = bp[(h * 64 + ty * 4) / 2 + 1];
        b[h * 4 + 0] = b0.x; b[h * 4 + 1] = b0.y; b[h * 4 + 2] = b1.x; b[h * 4 + 3] = b1.y;
      }
#pragma unroll
      for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 8; j++) acc[i][j] = fma(a[i], b[j], acc[i][j]);
    }
    if (t + 1 < nk) {
      store_shared(buf ^ 1);
      __syncthreads();
    }
  }
  // epilogue: thread owns rows {tx*4..+3, 64+tx*4..+3} x cols {ty*4..+3, 64+ty*4..+3}
#pragma unroll
  for (int j = 0; j < 8; j++) {
    const int n = n0 + (j >> 2) * 64 + ty * 4 + (j & 3);
    if (n >= P.N) continue;
    double* col = P.C + (long long)n * P.ldc;
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const int m = m0 + (i >> 2) * 64 + tx * 4 + (i & 3);
      if (m >= P.M) continue;
      const double v = P.alpha * acc[i][j];
      col[m] = P.beta == 0.0 ? v : fma(P.beta, col[m], v);
    }
  }
}

int g_ladder_launches = 0;

cudaError_t launch_ladder(const Params& P, int variant, cudaStream_t stream) {
  if (variant == VARIANT_NAIVE) {
    dim3 grid((P.M + 31) / 32, (P.N + 7) / 8);
    dgemm_naive_kernel<<<grid, 256, 0, stream>>>(P);
  } else {
    dim3 grid((P.M + LBM - 1) / LBM, (P.N + LBN - 1) / LBN);
    dgemm_dfma_kernel<<<grid, 256, 0, stream>>>(P);
  }
  return cudaGetLastError();
}

}  // namespace dg
