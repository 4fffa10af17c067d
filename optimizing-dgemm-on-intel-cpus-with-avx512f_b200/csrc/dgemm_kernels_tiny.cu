// dgemm_kernels_tiny.cu — the persistent TMA kernel of dgemm_tma_kernel.cuh compiled for a 64 x 32 CTA tile (namespace
// dg64x32): 8 DMMA consumer warps in a 4 x 2 grid with 16 x 16 warp tiles (8 accumulator registers), 12 KB k-tiles (A 64 x 16
// as four 16 x 16 boxes, B one 16 x 32 box), ONE k-tile per ring slot (DG_TINY_KPS = 1, see below), 16 KB stream-K partials.
//
// Why: 512^3 is 16 big, 32 mid or 64 small tiles for 148 SMs; each of those builds fills the GPU by cutting K into ranges,
// and when every range of a tile ends at the same moment the owner pays the whole exchange after its loop (partial
// store, flag, bulk load, add, C read-modify-write: ~5 of 22.5 us, profiles/r02_timelines_small_tile_one_vs_two_ctas.txt),
// which is why cuBLAS (64 x 32 tiles, 128 data-parallel CTAs, no exchange) was ahead at this one size: 15.2 against
// 11.8 TFLOP/s.  This tile gives 512^3 128 whole tiles: one wave, no workspace, no flags.  A 64 x 32 k-tile is only 512 clk
// of DMMAs, so the ~75 clk a ring-slot boundary costs a sub-partition is 15 %.  Two k-tiles per slot (DG_TINY_KPS=2, `make
// libtiny`) halve the crossings but were measured no faster end to end and are NOT shipped: KPS = 1.
// run_device picks it where one data-parallel wave of these tiles covers most of the GPU (dgemm_api.cu: choose_tile).
// Same role in the reference: the narrowest edge micro-kernels (m8n4 ... m2n4, include/kernel18.h:571-600).
#define DG_NAMESPACE dg64x32
#define DG_BM 64
#define DG_BN 32
#if DG_TINY_WARPS4     // measurement build (make libtiny): 4 consumer warps, one per SM sub-partition, 32 x 16 warp tiles
#define DG_WARPS_M 2
#define DG_WARPS_N 2
#define DG_MIN_CTAS 2          // (a register cap, 128 per thread at compile time; shared memory keeps it at one CTA per SM)
#define DG_PRODUCER_REGS 40
#define DG_CONSUMER_REGS 200
#elif DG_TINY_WARPS16  // measurement build: 16 consumer warps, four per SM sub-partition, 16 x 8 warp tiles
#define DG_WARPS_M 4
#define DG_WARPS_N 4
#else
#define DG_WARPS_M 4
#define DG_WARPS_N 2
#endif
#ifndef DG_TINY_KPS
#define DG_TINY_KPS 1
#endif
#ifndef DG_TINY_DUAL_ACC
#define DG_TINY_DUAL_ACC 1
#endif
#ifndef DG_TINY_EARLY_DESC
#define DG_TINY_EARLY_DESC 1
#endif
#define DG_DUAL_ACC DG_TINY_DUAL_ACC
#define DG_EARLY_DESC_PREFETCH DG_TINY_EARLY_DESC
#undef DG_KT_PER_STAGE
#define DG_KT_PER_STAGE DG_TINY_KPS
#ifndef DG_TINY_STAGES
#define DG_TINY_STAGES 12
#endif
#define DG_SMALL_STAGES (DG_TINY_STAGES / DG_KT_PER_STAGE)   // 144 KB in flight: 12 of the 32 k-tiles of a 512^3 tile
#define DG_SMALL_FN(name) name##_tiny
#include "dgemm_small_tile.cuh"
