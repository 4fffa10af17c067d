// dgemm_params.h — the launch parameter block shared by every tile configuration of the kernels and by the host code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dg {

constexpr int BK = 16;   // k-depth of one pipeline stage (one 128-byte row of B, sixteen 128-byte rows of an A sub-tile)

struct Params {
  int M, N, K;
  double alpha, beta;
  const double* A; long long lda;
  const double* B; long long ldb;
  double* C; long long ldc;
  int tiles_m, tiles_n;
  int k_tiles;            // ceil(K / BK)
  int splits;             // split-K factor, 1 = none
  int k_tiles_per_split;  // ceil(k_tiles / splits)
  int tile_small;         // 1 / 2: tiles_m / tiles_n count 64 x 64 tiles and the dg64 (one CTA per SM) / dg64x2 (two) build runs;
                          // 3: 128 x 64 tiles, the dg128x64 build (dgemm_host.h: TILE_*)
  int group_m;            // rasterisation: tiles are walked in bands of group_m tile-rows
  int c_vec2;             // C base 16-byte aligned and ldc even -> 16-byte epilogue accesses
  double* workspace;      // splits > 1: [unit][ACC_PER_THREAD][CONSUMER_THREADS] partial accumulators;
                          // stream-K: one such slot per CTA range
  // ---- stream-K tail (TMA kernel only): tiles [sk_tile0, sk_tile0 + sk_tiles) are cut into gridDim.x equal
  // ranges of (tile, k-tile) iterations instead of whole tiles
  int sk_tile0, sk_tiles;
  int sk_grid;                // CTAs launched when sk_tiles > 0 (every one of them must be co-resident)
  int sk_parallel;            // 1: parallel fix-up (all-stream-K mode, many ranges per tile): every range stores its
                              // partial (2 slots per range: head / tail) and each participant of a tile reduces a
                              // 1/R slice of it; 0: the tile's owner adds the other ranges' partials serially
  unsigned int sk_epoch;      // flag value that means "partial of this launch is in the workspace"
  unsigned int* sk_flags;     // one flag per CTA range
  // ---- k-lockstep of the data-parallel waves (TMA kernel only).  Persistent CTAs have no reason to stay together: each
  // walks its own tiles and nothing pulls a CTA that has fallen behind back.  But the 148 tiles of a wave share 16 row
  // panels of A and ~9 column panels of B through L2 only while their CTAs read the same k at about the same time; once
  // they drift further apart than the L2 can bridge, every panel is fetched from HBM several times (measured at
  // 16384^3: 175 GB of DRAM reads against a 46 GB floor for this tile size).  Every sync_kt k-tiles the producer
  // thread ticks a per-point counter and will not run more than SYNC_SLACK points ahead of the slowest CTA.
  unsigned int* sync_ctr;     // one counter per sync point, zeroed before the launch; nullptr = off
  int sync_waves;             // full waves (every CTA has a tile) covered by the lockstep
  int sync_kt;                // k-tiles between sync points
  // ---- fused NVLink fetch (TMA kernel only): while the DMMA warps work on this call's K-slab, the three spare
  // warps of the producer warpgroup copy fetch_bytes from fetch_src (normally a PEER GPU's copy of the next K-slab
  // of A, mapped through CUDA IPC / peer access) to fetch_dst in local HBM.  16-byte aligned, multiple of 16.
  const void* fetch_src;
  void* fetch_dst;
  unsigned long long fetch_bytes;
  // ---- optional per-CTA timeline (debug / profiling aid; nullptr in normal use): 16 globaltimer stamps per CTA
  unsigned long long* timeline;
  // ---- parity-split operands (TMA kernel, SPLIT != 0): an operand whose leading dimension is odd or whose base is only
  // 8-byte aligned cannot be described by ONE tensor map (16-byte base and strides).  Its even and its odd columns can
  // (stride 2*ld*8 is a multiple of 16) as long as the column start of that parity is 16-byte aligned; TMA rejects a
  // box that starts 8 bytes off a 16-byte boundary (measured: "illegal instruction" with an odd inner coordinate), so a
  // parity whose columns start at 8 mod 16 (a_mis / b_mis = 1) is staged with 8-byte cp.async by the two spare warps
  // of the producer warpgroup instead, into the same swizzled layout and onto the same mbarrier.
  int a_mis[2], b_mis[2];
};
// SPLIT template bits of dgemm_tma_ws_kernel
enum { SPLIT_NONE = 0, SPLIT_A = 1, SPLIT_B = 2 };

}  // namespace dg
