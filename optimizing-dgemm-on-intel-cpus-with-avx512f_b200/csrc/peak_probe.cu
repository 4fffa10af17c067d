// peak_probe.cu — FP64 pipe microbenchmarks for B200 (sm_100a).
//
// Measures the denominators the DGEMM roofline is quoted against (the
// 128 flop/clk/SM figure is derived, not measured):
//   * DFMA-only issue rate          (fma.rn.f64, register operands)
//   * DMMA-only issue rate          (mma.sync.m8n8k4.f64 -> SASS DMMA.8x8x4)
//   * the same with 1/2/4 warps per SM sub-partition, to see how many warps the
//     pipe needs to saturate
//   * the SM clock DURING each run, derived in-kernel from clock64() / globaltimer
//
// Stand-alone binary (no torch, no cuBLAS):  peak_probe [iters]
// Prints one JSON object per line.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint64_t gtimer() {
  uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t;
}

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

struct Stamp { unsigned long long cyc; unsigned long long ns; };

// NACC independent accumulator pairs per warp; each loop trip issues NACC DMMAs.
template <int NACC>
__global__ void __launch_bounds__(NACC >= 16 ? 256 : 1024) dmma_kernel(double* out, Stamp* st, int iters) {
  double acc[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; i++) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  __syncthreads();
  unsigned long long c0 = clock64(), t0 = gtimer();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < NACC; i++) dmma884(acc[i][0], acc[i][1], a, b);
  }
  unsigned long long c1 = clock64(), t1 = gtimer();
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += acc[i][0] + acc[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) { st[blockIdx.x].cyc = c1 - c0; st[blockIdx.x].ns = t1 - t0; }
}

template <int NACC>
__global__ void __launch_bounds__(1024) dfma_kernel(double* out, Stamp* st, int iters) {
  double acc[NACC];
#pragma unroll
  for (int i = 0; i < NACC; i++) acc[i] = threadIdx.x * 1e-3 + i;
  double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
  __syncthreads();
  unsigned long long c0 = clock64(), t0 = gtimer();
#pragma unroll 1
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < NACC; i++) asm volatile("fma.rn.f64 %0, %1, %0, %2;" : "+d"(acc[i]) : "d"(a), "d"(b));
  }
  unsigned long long c1 = clock64(), t1 = gtimer();
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; i++) s += acc[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) { st[blockIdx.x].cyc = c1 - c0; st[blockIdx.x].ns = t1 - t0; }
}

template <typename K>
static void run(const char* name, K kern, int warps_per_sm, int nacc, double flop_per_instr_warp, int iters,
                int nsm, double* d_out, Stamp* d_st) {
  int threads = warps_per_sm * 32;
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  kern<<<nsm, threads>>>(d_out, d_st, iters / 8 + 1);  // warm-up
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  std::vector<Stamp> st(nsm);
  double mhz = 0, cyc_med = 0;
  for (int rep = 0; rep < 3; rep++) {
    CK(cudaEventRecord(e0));
    kern<<<nsm, threads>>>(d_out, d_st, iters);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) {
      best = ms;
      CK(cudaMemcpy(st.data(), d_st, sizeof(Stamp) * nsm, cudaMemcpyDeviceToHost));
      std::vector<double> f, c;
      for (auto& s : st) { f.push_back(1e3 * (double)s.cyc / (double)s.ns); c.push_back((double)s.cyc); }
      std::sort(f.begin(), f.end()); std::sort(c.begin(), c.end());
      mhz = f[f.size() / 2]; cyc_med = c[c.size() / 2];
    }
  }
  double instr_per_warp = (double)iters * nacc;
  double flops = instr_per_warp * flop_per_instr_warp * warps_per_sm * nsm;
  double tflops = flops / (best * 1e-3) / 1e12;
  // flop per clock per SM from the in-kernel cycle counter (launch overhead excluded)
  double flop_clk_sm = instr_per_warp * flop_per_instr_warp * warps_per_sm / cyc_med;
  // cycles between successive instructions on one SM sub-partition
  double cyc_per_instr_smsp = cyc_med / (instr_per_warp * warps_per_sm / 4.0);
  printf("{\"probe\": \"%s\", \"warps_per_sm\": %d, \"nacc\": %d, \"ms\": %.4f, \"tflops\": %.3f, "
         "\"sm_mhz_inkernel\": %.1f, \"flop_per_clk_per_sm\": %.2f, \"clk_per_instr_per_smsp\": %.3f}\n",
         name, warps_per_sm, nacc, best, tflops, mhz, flop_clk_sm, cyc_per_instr_smsp);
  fflush(stdout);
}

int main(int argc, char** argv) {
  int iters = argc > 1 ? atoi(argv[1]) : 20000;
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int nsm = p.multiProcessorCount;
  int clk_khz = 0; CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  printf("{\"device\": \"%s\", \"sms\": %d, \"cc\": \"%d.%d\", \"clock_rate_mhz\": %.0f}\n", p.name, nsm, p.major,
         p.minor, clk_khz / 1e3);
  double* d_out; Stamp* d_st;
  CK(cudaMalloc(&d_out, sizeof(double) * nsm * 1024));
  CK(cudaMalloc(&d_st, sizeof(Stamp) * nsm));
  // DMMA m8n8k4: 8*8*4*2 = 512 flop per warp instruction
  run("dmma_m8n8k4", dmma_kernel<1>, 4, 1, 512.0, iters * 4, nsm, d_out, d_st);   // dependent chain = latency
  run("dmma_m8n8k4", dmma_kernel<2>, 4, 2, 512.0, iters * 4, nsm, d_out, d_st);
  run("dmma_m8n8k4", dmma_kernel<4>, 4, 4, 512.0, iters * 2, nsm, d_out, d_st);
  run("dmma_m8n8k4", dmma_kernel<32>, 4, 32, 512.0, iters / 2, nsm, d_out, d_st);  // a 32-accumulator warp tile, 1 warp/SMSP
  run("dmma_m8n8k4", dmma_kernel<32>, 8, 32, 512.0, iters / 2, nsm, d_out, d_st);  // ... 2 warps/SMSP (the GEMM's shape)
  run("dmma_m8n8k4", dmma_kernel<8>, 16, 8, 512.0, iters, nsm, d_out, d_st);       // 4 warps/SMSP
  // DFMA: 32 lanes * 2 flop = 64 flop per warp instruction
  run("dfma", dfma_kernel<1>, 4, 1, 64.0, iters * 8, nsm, d_out, d_st);            // dependent chain = latency
  run("dfma", dfma_kernel<16>, 4, 16, 64.0, iters * 4, nsm, d_out, d_st);
  run("dfma", dfma_kernel<16>, 8, 16, 64.0, iters * 4, nsm, d_out, d_st);
  run("dfma", dfma_kernel<16>, 32, 16, 64.0, iters, nsm, d_out, d_st);
  return 0;
}
