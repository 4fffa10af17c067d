// lds_dmma_probe.cu — how much of the FP64 tensor pipe survives the fragment loads of a warp tile?
//
// The three tile builds of the DGEMM kernel run their main loops at different fractions of the DMMA rate (timelines,
// DESIGN.md section 3b): 128 x 128 (64 x 32 warp tiles: 12 LDS.64 per 32 DMMAs) ~98 %, 128 x 64 (32 x 32: 8 per 16) ~93 %,
// 64 x 64 (32 x 16: 6 per 8) ~86 %.  This probe isolates the loop: 8 warps per SM (2 per sub-partition), FM + FN
// conflict-free LDS.64 per k-step feeding FM x FN DMMA.8x8x4, fragments double-buffered exactly as in
// dgemm_tma_kernel.cuh, no TMA, no barriers.  Variants: B fragments of two k-steps by ONE LDS.128 (B128), and 4 warps per
// sub-partition.  Prints flop/clk/SM from clock64().
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o lds_dmma_probe tools/lds_dmma_probe.cu && ./lds_dmma_probe
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}
__device__ __forceinline__ double lds64(uint32_t a) {
  double v; asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a) : "memory"); return v;
}
__device__ __forceinline__ void lds128(uint32_t a, double& x, double& y) {
  asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y) : "r"(a) : "memory");
}

template <int FM, int FN> struct Frag { double a[FM], b[FN]; };

// MODE 0: FM + FN LDS.64 per k-step.  MODE 1: B of two k-steps by one LDS.128 (FM LDS.64 per step + FN LDS.128 per two).
// MODE 2: no loads at all (register operands): the pipe's own rate with this instruction mix.
template <int FM, int FN, int MODE, int W>
__global__ void __launch_bounds__(W * 32) probe_kernel(double* out, unsigned long long* cyc, int iters) {
  extern __shared__ __align__(16) unsigned char smem[];
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem);
  for (int i = threadIdx.x; i < 16384 / 8; i += blockDim.x) reinterpret_cast<double*>(smem)[i] = 1.0 + i * 1e-9;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // conflict-free: lane l reads 8 bytes at l*8 (+256 per fragment): 2 wavefronts per LDS.64, 4 per LDS.128
  const uint32_t a0 = base + (warp & 7) * 1024 + lane * 8, b0 = base + 8192 + lane * 8, b128 = base + 8192 + lane * 16;
  double acc[FN][FM][2];
#pragma unroll
  for (int i = 0; i < FN; i++)
#pragma unroll
    for (int j = 0; j < FM; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
  Frag<FM, FN> f0, f1;
  double bx[FN], by[FN];
  auto load = [&](Frag<FM, FN>& f, int step) {
    if (MODE == 2) {
#pragma unroll
      for (int j = 0; j < FM; j++) f.a[j] = 1.0 + j;
#pragma unroll
      for (int i = 0; i < FN; i++) f.b[i] = 1.0 - i;
      return;
    }
#pragma unroll
    for (int j = 0; j < FM; j++) f.a[j] = lds64(a0 + step * 2048 + (j & 3) * 256);
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < FN; i++) f.b[i] = lds64(b0 + step * 1024 + (i & 3) * 256);
    } else if ((step & 1) == 0) {
#pragma unroll
      for (int i = 0; i < FN; i++) { lds128(b128 + (step >> 1) * 2048 + (i & 3) * 512, bx[i], by[i]); f.b[i] = bx[i]; }
    } else {
#pragma unroll
      for (int i = 0; i < FN; i++) f.b[i] = by[i];
    }
  };
  auto mma = [&](const Frag<FM, FN>& f) {
#pragma unroll
    for (int i = 0; i < FN; i++)
#pragma unroll
      for (int j = 0; j < FM; j++) dmma884(acc[i][j][0], acc[i][j][1], f.b[i], f.a[j]);
  };
  __syncthreads();
  const unsigned long long c0 = clock64();
  load(f0, 0);
#pragma unroll 1
  for (int it = 0; it < iters; it++) {   // one "stage" = 4 k-steps, as in the kernel
    load(f1, 1); mma(f0);
    load(f0, 2); mma(f1);
    load(f1, 3); mma(f0);
    load(f0, 0); mma(f1);
  }
  const unsigned long long c1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < FN; i++)
#pragma unroll
    for (int j = 0; j < FM; j++) s += acc[i][j][0] + acc[i][j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = c1 - c0;
}

template <typename K>
static void run(const char* name, K kern, int fm, int fn, int warps, int iters, int nsm, double* d_out, unsigned long long* d_cyc) {
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
  kern<<<nsm, warps * 32, 16384>>>(d_out, d_cyc, iters / 8 + 1);
  CK(cudaDeviceSynchronize());
  std::vector<unsigned long long> c(nsm);
  double best = 1e300;
  for (int rep = 0; rep < 3; rep++) {
    kern<<<nsm, warps * 32, 16384>>>(d_out, d_cyc, iters);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(c.data(), d_cyc, 8 * nsm, cudaMemcpyDeviceToHost));
    std::sort(c.begin(), c.end());
    best = std::min(best, (double)c[nsm / 2]);
  }
  const double dmma_per_warp = 4.0 * iters * fm * fn;
  const double flop_clk_sm = dmma_per_warp * 512.0 * warps / best;
  printf("{\"probe\": \"%s\", \"frag_m\": %d, \"frag_n\": %d, \"warps_per_sm\": %d, \"lds_per_dmma\": %.3f, "
         "\"flop_per_clk_per_sm\": %.2f, \"pct_of_128\": %.1f}\n",
         name, fm, fn, warps, (double)(fm + fn) / (fm * fn), flop_clk_sm, flop_clk_sm / 1.28);
  fflush(stdout);
}

int main(int argc, char** argv) {
  const int iters = argc > 1 ? atoi(argv[1]) : 4000;
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int nsm = p.multiProcessorCount;
  double* d_out; unsigned long long* d_cyc;
  CK(cudaMalloc(&d_out, 8 * nsm * 512)); CK(cudaMalloc(&d_cyc, 8 * nsm));
#define RUN(FM, FN, MODE, W, NAME) run(NAME, probe_kernel<FM, FN, MODE, W>, FM, FN, W, iters, nsm, d_out, d_cyc)
  RUN(8, 4, 2, 8, "regs only, 64x32 warp tile");
  RUN(8, 4, 0, 8, "LDS.64, 64x32 warp tile (128x128 CTA tile)");
  RUN(8, 4, 1, 8, "B by LDS.128, 64x32 warp tile");
  RUN(4, 4, 2, 8, "regs only, 32x32 warp tile");
  RUN(4, 4, 0, 8, "LDS.64, 32x32 warp tile (128x64 CTA tile)");
  RUN(4, 4, 1, 8, "B by LDS.128, 32x32 warp tile");
  RUN(4, 2, 2, 8, "regs only, 32x16 warp tile");
  RUN(4, 2, 0, 8, "LDS.64, 32x16 warp tile (64x64 CTA tile)");
  RUN(4, 2, 1, 8, "B by LDS.128, 32x16 warp tile");
  RUN(4, 2, 0, 16, "LDS.64, 32x16 warp tile, 4 warps per sub-partition");
  RUN(4, 4, 0, 16, "LDS.64, 32x32 warp tile, 4 warps per sub-partition");
  RUN(2, 2, 0, 16, "LDS.64, 16x16 warp tile, 4 warps per sub-partition");
  RUN(4, 2, 0, 4, "LDS.64, 32x16 warp tile, 1 warp per sub-partition");
  RUN(8, 4, 0, 4, "LDS.64, 64x32 warp tile, 1 warp per sub-partition");
  return 0;
}
