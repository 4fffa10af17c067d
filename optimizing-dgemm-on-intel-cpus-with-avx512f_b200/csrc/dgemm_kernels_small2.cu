// dgemm_kernels_small2.cu — the 64 x 64 build of the persistent TMA kernel at TWO CTAs per SM (namespace dg64x2).
//
// Why: at small and mid sizes a CTA spends a visible share of its life outside the DMMA loop: start-up (barrier init,
// descriptor fetch, first operands from cold HBM: ~2 us), one epilogue or partial store per segment (1-3 us each: C is
// read, combined, written), the stream-K fix-up.  With one CTA per SM the FP64 pipe idles through all of it (ncu at
// 1024^3, profiles/r01_ncu_full_dg64_1024x1024x1024.json: DMMA busy 77 % of elapsed, the busy cycles exactly the
// algorithmic minimum).  Two resident CTAs with four consumer warps per SM sub-partition between them take turns: while
// one drains, stores or waits for a partial, the other one's main loop has the pipe to itself.
//
// Budget per CTA: 384 threads compiled at 65536 / 768 = 80 registers (launch bounds (384, 2)), re-balanced to
// 40 (producer warpgroup) / 96 (consumers: 32 accumulator registers + two fragment sets; the DMMA loop compiles without
// local-memory traffic, with 24 / 104 it has two LDL per k-tile); 4 stages x 16 KB = 65 KB of
// shared memory, so two CTAs hold the same 128 KB of operands in flight per SM as the one-CTA build.
// run_device sizes the grid and the stream-K ranges for 2 x #SMs CTA slots when DGEMM_B200_SMALL_CTAS=2 selects this build.
//
// RESULT (profiles/r02_small_tile_probe_one_vs_two_ctas.jsonl, r02_timelines_small_tile_one_vs_two_ctas.txt,
// r02_ncu_full_dg64x2_1024.json): never faster than the one-CTA build -- 512^3 10.6 vs 11.9 TFLOP/s, 1024^3 27.4 vs 27.8,
// 1000x1003x997 24.7 vs 25.1, 2048^3 32.8 vs 33.2.  The two CTAs of an SM are launched together, reach their tile switches
// and their fix-up together and finish together, so there is little for one to overlap with in the other; the DMMA pipe
// is busy 76 % of the elapsed cycles (83 % of the active ones), as before.  Kept selectable, not the default.  The reference has no counterpart (one thread per core, kernel19.h:678-755).
#define DG_NAMESPACE dg64x2
#ifndef DG_KT_UNROLL
#define DG_KT_UNROLL 1   // the k-tile loop NOT unrolled: left alone, nvcc unrolls this build's loop by two and a whole tile of 128 k-tiles
#endif                   // takes 141.1 kclk instead of 136.2 (profiles/r02_kt_unroll_clk.txt); the other builds keep the compiler's choice
#define DG_SMALL_STAGES (4 / DG_KT_PER_STAGE)
#define DG_MIN_CTAS 2
#define DG_PRODUCER_REGS 40
#define DG_CONSUMER_REGS 96
#define DG_SMALL_FN(name) name##_x2
#include "dgemm_small_tile.cuh"
