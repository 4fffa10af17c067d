// dgemm_kernels_small.cu — the persistent TMA kernel of dgemm_tma_kernel.cuh compiled for a 64 x 64 CTA tile
// (namespace dg64): 8 DMMA consumer warps with 32 x 16 warp tiles, 16 KB pipeline stages, 32 KB stream-K partials.
//
// Why: a problem with few 128 x 128 tiles (256^3: 4, 512^3: 16) can only fill 148 SMs by cutting K into many ranges per
// tile, and then the partial-tile exchange dominates: R ranges of a tile write and read R x 128 KB through L2 (512^3:
// 16 MB each way, ~6.5 us of an 18 us kernel).  A quarter-size tile needs a quarter of the ranges per tile and moves a
// quarter of the bytes per partial.  Its main loop is less efficient (6 fragment loads per 8 DMMAs instead of 12 per
// 32), which does not matter where the launch is a few microseconds long; run_device picks it only for small,
// TMA-eligible problems (see dgemm_api.cu) and the 128 x 128 kernel stays the performance path.
// Same role in the reference: the 8x8 / 2x8 ... edge micro-kernels beside the 24x8 one (include/kernel18.h:533-600).
#define DG_NAMESPACE dg64
#define DG_SMALL_TILE 1
#define DG_BM 64
#define DG_BN 64
#include "dgemm_tma_kernel.cuh"
#include "dgemm_host.h"
#include "dgemm_launch.h"

#include <atomic>

namespace dg {

static_assert(dg64::BM == SMALL_BM && dg64::BN == SMALL_BN, "dgemm_host.h must state the small tile's size");

constexpr int SMALL_STAGES = 8;   // 8 x 16 KB: the same 128 KB in flight as four stages of the big tile
constexpr int SMALL_SMEM = SMALL_STAGES * dg64::STAGE_BYTES + 2 * SMALL_STAGES * 8 + 16 + 1024;

static cudaError_t small_attrs() {
  constexpr int STAGES = SMALL_STAGES;
  const int smem = SMALL_SMEM;
  static std::atomic<bool> attr_done[64];  // function attributes are per device
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return err;
  if (!attr_done[dev & 63]) {
    const cudaFuncAttribute a = cudaFuncAttributeMaxDynamicSharedMemorySize;
    if ((err = cudaFuncSetAttribute(dg64::dgemm_tma_ws_kernel<STAGES>, a, smem)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(dg64::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A>, a, smem)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(dg64::dgemm_tma_ws_split_kernel<STAGES, SPLIT_B>, a, smem)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(dg64::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A | SPLIT_B>, a, smem)) != cudaSuccess)
      return err;
    attr_done[dev & 63] = true;
  }
  return cudaSuccess;
}

cudaError_t small_tile_occupancy(int* ctas_per_sm) {
  cudaError_t err = small_attrs();
  if (err != cudaSuccess) return err;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, dg64::dgemm_tma_ws_kernel<SMALL_STAGES>, dg64::TMA_THREADS,
                                                       SMALL_SMEM);
}

cudaError_t launch_tma_small(const void* const maps[4], int split, const Params& P, int grid, int dp_units,
                             cudaStream_t stream, bool cooperative, bool pdl) {
  constexpr int STAGES = SMALL_STAGES;
  cudaError_t err = small_attrs();
  if (err != cudaSuccess) return err;
  const CUtensorMap &tmA = *reinterpret_cast<const CUtensorMap*>(maps[0]), &tmA1 = *reinterpret_cast<const CUtensorMap*>(maps[1]),
                    &tmB = *reinterpret_cast<const CUtensorMap*>(maps[2]), &tmB1 = *reinterpret_cast<const CUtensorMap*>(maps[3]);
  LaunchCfg cfg;
  cfg.grid = grid; cfg.threads = dg64::TMA_THREADS; cfg.smem = SMALL_SMEM; cfg.stream = stream;
  cfg.cooperative = cooperative; cfg.pdl = pdl;
  switch (split) {
    case SPLIT_NONE:
      err = launch_ex(dg64::dgemm_tma_ws_kernel<STAGES>, cfg, tmA, tmB, P, dp_units);
      break;
    case SPLIT_A:
      err = launch_ex(dg64::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A>, cfg, tmA, tmA1, tmB, tmB1, P, dp_units);
      break;
    case SPLIT_B:
      err = launch_ex(dg64::dgemm_tma_ws_split_kernel<STAGES, SPLIT_B>, cfg, tmA, tmA1, tmB, tmB1, P, dp_units);
      break;
    default:
      err = launch_ex(dg64::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A | SPLIT_B>, cfg, tmA, tmA1, tmB, tmB1, P, dp_units);
  }
  if (err != cudaSuccess) cudaGetLastError();
  return err;
}

}  // namespace dg
