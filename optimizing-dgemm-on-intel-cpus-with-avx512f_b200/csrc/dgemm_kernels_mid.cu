// dgemm_kernels_mid.cu — the persistent TMA kernel of dgemm_tma_kernel.cuh compiled for a 128 x 64 CTA tile (namespace
// dg128x64): 8 DMMA consumer warps in a 4 x 2 grid with 32 x 32 warp tiles (64 accumulator registers), 24 KB pipeline
// stages (A 128 x 16 as eight 16 x 16 boxes, B one 16 x 64 box), 64 KB stream-K partials.
//
// Why: every stage boundary costs a sub-partition ~75 clk whatever the tile (profiles/r02_loop_with_and_without_tma.txt),
// so the 64 x 64 build's loop (1024-clk stages) runs at ~92 % of the FP64 pipe and the 128 x 128 build's (4096-clk stages)
// at 98 %, but the latter with four times the bytes per stream-K partial and a quarter of the tiles to balance.  This
// tile sits between them: 2048-clk stages (96 %), half the partial of the big tile.  run_device picks it for the sizes
// between the two (dgemm_api.cu: choose_tile; measured band in profiles/r02_tile_probe_big_small_mid.jsonl).
// Same role in the reference: the 16x8 / 8x8 edge micro-kernels beside the 24x8 one (include/kernel18.h:533-600).
#define DG_NAMESPACE dg128x64
#define DG_BM 128
#define DG_BN 64
#define DG_WARPS_M 4
#define DG_WARPS_N 2
#define DG_SMALL_STAGES (8 / DG_KT_PER_STAGE)   // 8 x 24 KB = 192 KB in flight, as the big tile's 6 x 32 KB
#define DG_SMALL_FN(name) name##_mid
#include "dgemm_small_tile.cuh"
