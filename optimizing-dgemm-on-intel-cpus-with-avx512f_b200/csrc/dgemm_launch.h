// dgemm_launch.h — one launch helper for the TMA kernels of both tile configurations (dgemm_kernels.cu,
// dgemm_kernels_small.cu): cudaLaunchKernelEx with the two launch attributes the kernels are written for.
//
//   cooperative   configurations whose CTAs WAIT ON OTHER CTAs OF THE SAME GRID (k-lockstep: every producer waits for every
//                 other one; parallel stream-K fix-up: the ranges of a tile wait for each other) are only correct when the
//                 whole grid is co-resident.  A cooperative launch makes the runtime guarantee that or fail the launch
//                 (cudaErrorCooperativeLaunchTooLarge); run_device then falls back to the schedule without such waits.
//                 (The serial stream-K fix-up needs no guarantee: an owner only waits for CTAs with a LOWER blockIdx,
//                 which were dispatched before it and never wait for anybody.)
//   programmatic  every other launch allows programmatic dependent launch: the kernel's prologue (barrier init, tensor-map
//                 prefetch) runs while the previous kernel of the stream drains, and griddepcontrol.wait holds back its first
//                 global access until that kernel has completed.
//                 INVARIANT: a kernel launched with this attribute MUST execute griddepcontrol.wait before its first global
//                 memory access (the TMA kernels do, dgemm_tma_kernel.cuh).  Every other kernel of the library (cp.async,
//                 split-K reduce, scale / add / realign / transpose, ladder rungs, flag kernels) is launched with <<<>>> and
//                 no attribute, so it starts only after the previous kernel of its stream has completed and needs no wait.
//                 Do not pass pdl = true for a kernel that lacks the instruction.
#pragma once
#include <cuda_runtime.h>

#include <utility>

namespace dg {

struct LaunchCfg {
  int grid = 1, threads = 1, smem = 0;
  cudaStream_t stream = nullptr;
  bool cooperative = false, pdl = false;
};

template <typename... KArgs, typename... Args>
inline cudaError_t launch_ex(void (*kern)(KArgs...), const LaunchCfg& c, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)c.grid);
  cfg.blockDim = dim3((unsigned)c.threads);
  cfg.dynamicSmemBytes = (size_t)c.smem;
  cfg.stream = c.stream;
  cudaLaunchAttribute attr[1];
  int n = 0;
  if (c.cooperative) {
    attr[n].id = cudaLaunchAttributeCooperative;
    attr[n].val.cooperative = 1;
    n++;
  } else if (c.pdl) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    n++;
  }
  cfg.attrs = attr;
  cfg.numAttrs = (unsigned)n;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

}  // namespace dg
