// loop_bubble_probe.cu — which ingredient of the consumers' k-tile loop costs the ~75 clk per k-tile?
//
// tools/lds_dmma_probe.cu runs the fragment loads + DMMAs of a warp tile alone and reaches 100 % of the FP64 pipe for every
// warp tile; the kernel's loop with a producer that loads nothing (make libdbg) loses a constant ~75-80 clk per k-tile and
// sub-partition whatever the tile (profiles/r02_loop_with_and_without_tma.txt, r02_tiny_tile_loop_with_and_without_tma.txt).
// This probe adds the kernel's ingredients to the bare loop ONE AT A TIME, with the same PTX the kernel uses:
//   mode 0  bare loop, fixed shared-memory addresses                      (= lds_dmma_probe)
//   mode 1  + ring addressing: the stage base moves every k-tile, (s, phase) bookkeeping, has_next predicate
//   mode 2  + release: __syncwarp(); lane 0 arrives on the stage's "empty" mbarrier
//   mode 3  + acquire: a producer warp hands every stage over (waits empty, arrives full); consumers probe the next
//             stage's "full" barrier with test_wait one DMMA group ahead and fall back to try_wait  (= the kernel's loop)
//   mode 4  as 3, but no test_wait: try_wait right where the next stage's first fragments are loaded
//   mode 5  as 3, but the release comes BEFORE the last DMMA group (the stage's last loads have been issued by then)
// Prints clk per k-tile and sub-partition against the pipe's minimum (FM x FN x 4 DMMAs x 16 clk x warps per sub-partition).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o loop_bubble_probe tools/loop_bubble_probe.cu && ./loop_bubble_probe
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
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ uint32_t mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done;
}

constexpr int STAGES = 6, SLOT = 4096, CW = 8;   // 8 consumer warps (2 per sub-partition) + 1 producer warp

template <int FM, int FN> struct Frag { double a[FM], b[FN]; };

template <int FM, int FN, int MODE>
__global__ void __launch_bounds__((CW + 1) * 32) probe_kernel(double* out, unsigned long long* cyc, int iters) {
  extern __shared__ __align__(1024) unsigned char smem[];
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t bars = base + STAGES * SLOT;   // full[STAGES], empty[STAGES]
  for (int i = threadIdx.x; i < STAGES * SLOT / 8; i += blockDim.x) reinterpret_cast<double*>(smem)[i] = 1.0 + i * 1e-9;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 8 * (STAGES + s), CW); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp == CW) {   // producer: hands every stage over without loading anything (modes >= 3)
    if (MODE >= 3 && lane == 0) {
      uint32_t s = 0, ph = 0;
      for (int it = 0; it < iters; it++) {
        mbar_wait(bars + 8 * (STAGES + s), ph ^ 1);
        mbar_arrive(bars + 8 * s);
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
    }
    return;
  }
  // conflict-free: lane l reads 8 bytes at l*8 (+256 per fragment): 2 wavefronts per LDS.64
  const uint32_t la = lane * 8, lb = 2048 + lane * 8;
  double acc[FN][FM][2];
#pragma unroll
  for (int i = 0; i < FN; i++)
#pragma unroll
    for (int j = 0; j < FM; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
  Frag<FM, FN> f0, f1;
  auto load = [&](Frag<FM, FN>& f, uint32_t sb, int step) {
#pragma unroll
    for (int j = 0; j < FM; j++) f.a[j] = lds64(sb + la + step * 512 + (j & 1) * 256);
#pragma unroll
    for (int i = 0; i < FN; i++) f.b[i] = lds64(sb + lb + step * 512 + (i & 1) * 256);
  };
  auto mma = [&](const Frag<FM, FN>& f) {
#pragma unroll
    for (int i = 0; i < FN; i++)
#pragma unroll
      for (int j = 0; j < FM; j++) dmma884(acc[i][j][0], acc[i][j][1], f.b[i], f.a[j]);
  };
  uint32_t s = 0, ph = 0;
  asm volatile("bar.sync 1, %0;" ::"n"(CW * 32) : "memory");
  const unsigned long long c0 = clock64();
  if (MODE >= 3) mbar_wait(bars + 8 * s, ph);
  load(f0, base, 0);
#pragma unroll 1
  for (int it = 0; it < iters; it++) {   // one k-tile = 4 k-steps, as in the kernel
    const uint32_t sb = MODE >= 1 ? base + s * SLOT : base;
    const uint32_t cur_empty = bars + 8 * (STAGES + s);
    if (MODE >= 1) { if (++s == STAGES) { s = 0; ph ^= 1; } }
    const bool has_next = MODE >= 1 ? it + 1 < iters : true;
    const uint32_t nb = MODE >= 1 ? base + s * SLOT : base;
    load(f1, sb, 1); mma(f0);
    load(f0, sb, 2); mma(f1);
    load(f1, sb, 3);
    uint32_t next_ready = 1u;
    if (MODE == 3 || MODE == 5) next_ready = has_next ? mbar_test_wait(bars + 8 * s, ph) : 0u;
    mma(f0);
    if (MODE == 5) { __syncwarp(); if (lane == 0) mbar_arrive(cur_empty); }
    if (has_next) {
      if (MODE == 3 || MODE == 5) { if (!next_ready) mbar_wait(bars + 8 * s, ph); }
      if (MODE == 4) mbar_wait(bars + 8 * s, ph);
      load(f0, nb, 0);
    }
    mma(f1);
    if (MODE >= 2 && MODE != 5) { __syncwarp(); if (lane == 0) mbar_arrive(cur_empty); }
  }
  const unsigned long long c1 = clock64();
  double sum = 0;
#pragma unroll
  for (int i = 0; i < FN; i++)
#pragma unroll
    for (int j = 0; j < FM; j++) sum += acc[i][j][0] + acc[i][j][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
  if (threadIdx.x == 0) cyc[blockIdx.x] = c1 - c0;
}

template <typename K>
static void run(const char* tile, int mode, K kern, int fm, int fn, int iters, int nsm, double* d_out, unsigned long long* d_cyc) {
  const int smem = STAGES * SLOT + 2 * STAGES * 8;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<nsm, (CW + 1) * 32, smem>>>(d_out, d_cyc, iters / 8 + 1);
  CK(cudaDeviceSynchronize());
  std::vector<unsigned long long> c(nsm);
  double best = 1e300;
  for (int rep = 0; rep < 3; rep++) {
    kern<<<nsm, (CW + 1) * 32, smem>>>(d_out, d_cyc, iters);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(c.data(), d_cyc, 8 * nsm, cudaMemcpyDeviceToHost));
    std::sort(c.begin(), c.end());
    best = std::min(best, (double)c[nsm / 2]);
  }
  const double ideal = 4.0 * fm * fn * 16.0 * (CW / 4);
  printf("{\"warp_tile\": \"%s\", \"mode\": %d, \"clk_per_ktile\": %.1f, \"pipe_minimum\": %.0f, \"extra_clk\": %.1f, \"pct\": %.1f}\n",
         tile, mode, best / iters, ideal, best / iters - ideal, 100.0 * ideal / (best / iters));
  fflush(stdout);
}

int main(int argc, char** argv) {
  const int iters = argc > 1 ? atoi(argv[1]) : 4000;
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int nsm = p.multiProcessorCount;
  double* d_out; unsigned long long* d_cyc;
  CK(cudaMalloc(&d_out, 8 * nsm * 512)); CK(cudaMalloc(&d_cyc, 8 * nsm));
#define RUN(FM, FN, MODE, NAME) run(NAME, MODE, probe_kernel<FM, FN, MODE>, FM, FN, iters, nsm, d_out, d_cyc)
#define ALLMODES(FM, FN, NAME) RUN(FM, FN, 0, NAME); RUN(FM, FN, 1, NAME); RUN(FM, FN, 2, NAME); RUN(FM, FN, 3, NAME); \
  RUN(FM, FN, 4, NAME); RUN(FM, FN, 5, NAME)
  ALLMODES(2, 2, "16x16");
  ALLMODES(4, 2, "32x16");
  ALLMODES(4, 4, "32x32");
  ALLMODES(8, 4, "64x32");
  return 0;
}
