// loop_bubble_probe2.cu — tools/loop_bubble_probe.cu with the KERNEL'S OWN fragment code: dgemm_device.cuh is included with
// the tile configuration given on the command line, so LaneOffsets / load_frags / mma_frags, the swizzled addresses, the
// stage size and the warp grid are exactly the shipped ones.  The producer hands every stage over without loading it
// (as `make libdbg` does), so what is timed is the consumers' loop alone.
//   mode 0  bare loop on stage 0            mode 1  + ring addressing        mode 2  + release (syncwarp, lane 0 arrives)
//   mode 3  + acquire = the kernel's loop   mode 6  mode 3 in a 12-warp block with the kernel's setmaxnreg split
//   mode 7  mode 6 + warp 9 polling a shared word with nanosleep (the C prefetch warp), warps 10-11 exit at once
//   mode 8  mode 7 with the probe's answer made warp-uniform by a vote (no divergent branch in front of the next DMMA group)
//   mode 9  mode 7 without __syncwarp() in front of the elected arrive (the last mma.sync is the convergence point)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -DDG_BM=128 -DDG_BN=128 [-DDG_WARPS_M=.. -DDG_WARPS_N=..]
//        -DPROBE_STAGES=6 -o loop_bubble_probe2_big tools/loop_bubble_probe2.cu
#include "../optimizing-dgemm-on-intel-cpus-with-avx512f_b200/csrc/dgemm_device.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

#ifndef PROBE_STAGES
#define PROBE_STAGES 6
#endif
using namespace dg;
constexpr int STAGES = PROBE_STAGES;
constexpr int CW = CONSUMER_WARPS;
#if defined(PROBE_UNROLL) && PROBE_UNROLL > 0
constexpr int PROBE_UNROLL_N = PROBE_UNROLL;
#endif

template <int MODE>
__global__ void __launch_bounds__(MODE >= 6 ? CONSUMER_THREADS + 128 : CONSUMER_THREADS + 32, 1)
probe_kernel(double* out, unsigned long long* cyc, int iters) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem_base + STAGES * STAGE_BYTES;
  const uint32_t word = bars + 16 * STAGES;
  for (int i = threadIdx.x; i < STAGES * STAGE_BYTES / 8; i += blockDim.x)
    asm volatile("st.shared.f64 [%0], %1;" ::"r"(smem_base + 8 * i), "d"(1.0 + i * 1e-9) : "memory");
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; s++) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 8 * (STAGES + s), CW); }
    asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(word), "r"(0u) : "memory");
    fence_mbar_init();
  }
  __syncthreads();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (warp >= CW) {
    if (MODE >= 6) asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(PRODUCER_REGS));
    if (warp == CW && lane == 0 && MODE >= 3) {
      uint32_t s = 0, ph = 0;
      for (int it = 0; it < iters; it++) {
        mbar_wait(bars + 8 * (STAGES + s), ph ^ 1);
        mbar_arrive(bars + 8 * s);
        if (++s == STAGES) { s = 0; ph ^= 1; }
      }
      asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(word), "r"(1u) : "memory");
    } else if (warp == CW + 1 && MODE >= 7) {
      uint32_t seen;
      do {
        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(seen) : "r"(word) : "memory");
        if (!seen) __nanosleep(256);
      } while (!seen);
    }
    return;
  }
  if (MODE >= 6) asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CONSUMER_REGS));
  const int warp_m = warp % WARPS_M, warp_n = warp / WARPS_M;
  LaneOffsets lo = make_lane_offsets(lane, warp_m, warp_n);
  double acc[FRAG_N][FRAG_M][2];
  zero_acc(acc);
  Frags f0, f1;
  uint32_t s = 0, ph = 0;
  consumer_bar_sync();
  const unsigned long long c0 = clock64();
  if (MODE >= 3) mbar_wait(bars + 8 * s, ph);
  load_frags(f0, smem_base, smem_base + A_STAGE_BYTES, lo, 0);
#ifdef PROBE_UNROLL   // -DPROBE_UNROLL=n: force n; -DPROBE_UNROLL=0: leave it to the compiler, as the kernel does
#if PROBE_UNROLL > 0
#pragma unroll PROBE_UNROLL_N
#endif
#else
#pragma unroll 1
#endif
  for (int kt = 0; kt < iters; kt++) {
    const uint32_t sA = MODE >= 1 ? smem_base + s * STAGE_BYTES : smem_base, sB = sA + A_STAGE_BYTES;
    const uint32_t cur_empty = bars + 8 * (STAGES + s);
    if (MODE >= 1) { if (++s == STAGES) { s = 0; ph ^= 1; } }
    const bool has_next = MODE >= 1 ? kt + 1 < iters : true;
    const uint32_t nA = MODE >= 1 ? smem_base + s * STAGE_BYTES : smem_base;
    load_frags(f1, sA, sB, lo, 1);
    mma_frags(acc, f0);
    load_frags(f0, sA, sB, lo, 2);
    mma_frags(acc, f1);
    load_frags(f1, sA, sB, lo, 3);
    uint32_t next_ready = 1u;
    if (MODE >= 3) next_ready = has_next ? mbar_test_wait(bars + 8 * s, ph) : 0u;
    if (MODE == 8) next_ready = __all_sync(0xffffffffu, next_ready);
    mma_frags(acc, f0);
    if (has_next) {
      if (MODE >= 3) { if (!next_ready) mbar_wait(bars + 8 * s, ph); }
      load_frags(f0, nA, nA + A_STAGE_BYTES, lo, 0);
    }
    mma_frags(acc, f1);
    if (MODE >= 2) { if (MODE != 9) __syncwarp(); if (lane == 0) mbar_arrive(cur_empty); }
  }
  const unsigned long long c1 = clock64();
  double sum = 0;
#pragma unroll
  for (int i = 0; i < FRAG_N; i++)
#pragma unroll
    for (int j = 0; j < FRAG_M; j++) sum += acc[i][j][0] + acc[i][j][1];
  out[blockIdx.x * CONSUMER_THREADS + tid] = sum;
  if (tid == 0) cyc[blockIdx.x] = c1 - c0;
}

template <typename K>
static void run(int mode, K kern, int threads, int iters, int nsm, double* d_out, unsigned long long* d_cyc) {
  const int smem = STAGES * STAGE_BYTES + 2 * STAGES * 8 + 16 + 1024;
  CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  kern<<<nsm, threads, smem>>>(d_out, d_cyc, iters / 8 + 1);
  CK(cudaDeviceSynchronize());
  std::vector<unsigned long long> c(nsm);
  double best = 1e300;
  for (int rep = 0; rep < 3; rep++) {
    kern<<<nsm, threads, smem>>>(d_out, d_cyc, iters);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(c.data(), d_cyc, 8 * nsm, cudaMemcpyDeviceToHost));
    std::sort(c.begin(), c.end());
    best = std::min(best, (double)c[nsm / 2]);
  }
  const double ideal = 4.0 * FRAG_M * FRAG_N * 16.0 * (CW / 4);
  printf("{\"cta_tile\": \"%dx%d\", \"warp_tile\": \"%dx%d\", \"mode\": %d, \"clk_per_ktile\": %.1f, \"pipe_minimum\": %.0f, "
         "\"extra_clk\": %.1f, \"pct\": %.1f}\n", BM, BN, WM, WN, mode, best / iters, ideal, best / iters - ideal,
         100.0 * ideal / (best / iters));
  fflush(stdout);
}

int main(int argc, char** argv) {
  const int iters = argc > 1 ? atoi(argv[1]) : 4000;
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int nsm = p.multiProcessorCount;
  double* d_out; unsigned long long* d_cyc;
  CK(cudaMalloc(&d_out, 8 * nsm * 1024)); CK(cudaMalloc(&d_cyc, 8 * nsm));
#define RUN(MODE) run(MODE, probe_kernel<MODE>, MODE >= 6 ? CONSUMER_THREADS + 128 : CONSUMER_THREADS + 32, iters, nsm, d_out, d_cyc)
  RUN(0); RUN(1); RUN(2); RUN(3); RUN(6); RUN(7); RUN(8); RUN(9);
  return 0;
}
