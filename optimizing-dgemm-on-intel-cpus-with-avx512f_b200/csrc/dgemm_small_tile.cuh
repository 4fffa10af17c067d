// dgemm_small_tile.cuh — the smaller-tile builds of the persistent TMA kernel plus their launchers.  Included by
// dgemm_kernels_small.cu (namespace dg64: 64 x 64, one CTA per SM, 8 stages), dgemm_kernels_small2.cu (namespace dg64x2:
// 64 x 64, two CTAs per SM, 4 stages each) and dgemm_kernels_mid.cu (namespace dg128x64: 128 x 64, 32 x 32 warp tiles),
// which set DG_NAMESPACE / DG_SMALL_STAGES / DG_MIN_CTAS / the register split, optionally DG_BM / DG_BN / the warp grid,
// and DG_SMALL_FN (the suffix of the two host entry points) before including it.
#define DG_SMALL_TILE 1
#ifndef DG_BM
#define DG_BM 64
#define DG_BN 64
#endif
#include "dgemm_tma_kernel.cuh"
#include "dgemm_host.h"
#include "dgemm_launch.h"

#include <algorithm>
#include <atomic>

namespace dg {

namespace ks = DG_NAMESPACE;
static_assert((ks::BM == SMALL_BM && ks::BN == SMALL_BN) || (ks::BM == MID_BM && ks::BN == MID_BN) ||
                  (ks::BM == TINY_BM && ks::BN == TINY_BN),
              "dgemm_host.h must state the tile's size");

namespace {
constexpr int SMALL_STAGES = DG_SMALL_STAGES;
constexpr int SMALL_SMEM = SMALL_STAGES * ks::SLOT_BYTES + 2 * SMALL_STAGES * 8 + 16 + 1024;

cudaError_t small_attrs() {
  constexpr int STAGES = SMALL_STAGES;
  const int smem = SMALL_SMEM;
  static std::atomic<bool> attr_done[64];  // function attributes are per device
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return err;
  if (!attr_done[dev & 63]) {
    const cudaFuncAttribute a = cudaFuncAttributeMaxDynamicSharedMemorySize;
    if ((err = cudaFuncSetAttribute(ks::dgemm_tma_ws_kernel<STAGES>, a, smem)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(ks::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A>, a, smem)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(ks::dgemm_tma_ws_split_kernel<STAGES, SPLIT_B>, a, smem)) != cudaSuccess) return err;
    if ((err = cudaFuncSetAttribute(ks::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A | SPLIT_B>, a, smem)) != cudaSuccess)
      return err;
    attr_done[dev & 63] = true;
  }
  return cudaSuccess;
}
}  // namespace

cudaError_t DG_SMALL_FN(small_tile_occupancy)(int* ctas_per_sm) {
  cudaError_t err = small_attrs();
  if (err != cudaSuccess) return err;
  // the least over the plain and the parity-split kernels (they differ in registers, not in shared memory)
  int n[4] = {0, 0, 0, 0};
  if ((err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n[0], ks::dgemm_tma_ws_kernel<SMALL_STAGES>, ks::TMA_THREADS, SMALL_SMEM))) return err;
  if ((err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n[1], ks::dgemm_tma_ws_split_kernel<SMALL_STAGES, SPLIT_A>, ks::TMA_THREADS, SMALL_SMEM))) return err;
  if ((err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n[2], ks::dgemm_tma_ws_split_kernel<SMALL_STAGES, SPLIT_B>, ks::TMA_THREADS, SMALL_SMEM))) return err;
  if ((err = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n[3], ks::dgemm_tma_ws_split_kernel<SMALL_STAGES, SPLIT_A | SPLIT_B>, ks::TMA_THREADS, SMALL_SMEM))) return err;
  *ctas_per_sm = std::min(std::min(n[0], n[1]), std::min(n[2], n[3]));
  return cudaSuccess;
}

cudaError_t DG_SMALL_FN(launch_tma_small)(const void* const maps[4], int split, const Params& P, int grid, int dp_units,
                                          cudaStream_t stream, bool cooperative, bool pdl) {
  constexpr int STAGES = SMALL_STAGES;
  cudaError_t err = small_attrs();
  if (err != cudaSuccess) return err;
  const CUtensorMap &tmA = *reinterpret_cast<const CUtensorMap*>(maps[0]), &tmA1 = *reinterpret_cast<const CUtensorMap*>(maps[1]),
                    &tmB = *reinterpret_cast<const CUtensorMap*>(maps[2]), &tmB1 = *reinterpret_cast<const CUtensorMap*>(maps[3]);
  LaunchCfg cfg;
  cfg.grid = grid; cfg.threads = ks::TMA_THREADS; cfg.smem = SMALL_SMEM; cfg.stream = stream;
  cfg.cooperative = cooperative; cfg.pdl = pdl;
  switch (split) {
    case SPLIT_NONE:
      err = launch_ex(ks::dgemm_tma_ws_kernel<STAGES>, cfg, tmA, tmB, P, dp_units);
      break;
    case SPLIT_A:
      err = launch_ex(ks::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A>, cfg, tmA, tmA1, tmB, tmB1, P, dp_units);
      break;
    case SPLIT_B:
      err = launch_ex(ks::dgemm_tma_ws_split_kernel<STAGES, SPLIT_B>, cfg, tmA, tmA1, tmB, tmB1, P, dp_units);
      break;
    default:
      err = launch_ex(ks::dgemm_tma_ws_split_kernel<STAGES, SPLIT_A | SPLIT_B>, cfg, tmA, tmA1, tmB, tmB1, P, dp_units);
  }
  if (err != cudaSuccess) cudaGetLastError();
  return err;
}

}  // namespace dg
