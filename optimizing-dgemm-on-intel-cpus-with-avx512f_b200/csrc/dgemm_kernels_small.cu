// dgemm_kernels_small.cu — the persistent TMA kernel of dgemm_tma_kernel.cuh compiled for a 64 x 64 CTA tile
// (namespace dg64): 8 DMMA consumer warps with 32 x 16 warp tiles, 16 KB pipeline stages, 32 KB stream-K partials.
//
// Why: a problem with few 128 x 128 tiles (256^3: 4, 512^3: 16) can only fill 148 SMs by cutting K into many ranges per
// tile, and then the partial-tile exchange dominates: R ranges of a tile write and read R x 128 KB through L2 (512^3:
// 16 MB each way, ~6.5 us of an 18 us kernel).  A quarter-size tile needs a quarter of the ranges per tile and moves a
// quarter of the bytes per partial.  Its main loop is less efficient (92 % of the pipe against 98 %: the ~75 clk every
// stage boundary costs are 7 % of a 1024-clk stage and 2 % of a 4096-clk one -- not the fragment loads, see
// tools/lds_dmma_probe.cu), which does not matter where the launch is a few microseconds long; run_device picks it only for small,
// TMA-eligible problems (see dgemm_api.cu) and the 128 x 128 kernel stays the performance path.
// Same role in the reference: the 8x8 / 2x8 ... edge micro-kernels beside the 24x8 one (include/kernel18.h:533-600).
//
// Two builds of it (dgemm_small_tile.cuh): this one keeps ONE CTA per SM with 8 pipeline stages; dgemm_kernels_small2.cu
// runs TWO CTAs per SM with 4 stages each, so that one CTA's epilogue / partial exchange overlaps the other's main loop.
#define DG_NAMESPACE dg64
#ifndef DG_KT_UNROLL
#define DG_KT_UNROLL 1   // the k-tile loop NOT unrolled: left alone, nvcc unrolls this build's loop by two and a whole tile of 128 k-tiles
#endif                   // takes 141.1 kclk instead of 136.2 (profiles/r02_kt_unroll_clk.txt); the other builds keep the compiler's choice
#define DG_SMALL_STAGES (8 / DG_KT_PER_STAGE)   // 8 x 16 KB: the same 128 KB in flight as four stages of the big tile
#define DG_SMALL_FN(name) name
#include "dgemm_small_tile.cuh"
