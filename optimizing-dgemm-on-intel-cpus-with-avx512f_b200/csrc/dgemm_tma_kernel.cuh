// dgemm_tma_kernel.cuh — the persistent, warp-specialised TMA kernel (device code only), written against the tile
// configuration of the including translation unit (dgemm_device.cuh: DG_NAMESPACE, DG_BM, DG_BN).  Included by
// dgemm_kernels.cu (128 x 128 tiles, namespace dg) and dgemm_kernels_small.cu (64 x 64 tiles, namespace dg64).
#pragma once
#include "dgemm_device.cuh"

#include <cuda.h>

#ifndef DG_MIN_CTAS
#define DG_MIN_CTAS 1   // resident CTAs per SM the kernels are compiled for (register budget = 65536 / (threads x this))
#endif

namespace DG_NAMESPACE {

// =========================================================================================================
// TMA + mbarrier, warp-specialised, persistent
// =========================================================================================================
// 12 warps: two consumer warpgroups (8 warps) + one producer warpgroup (warp 8 issues TMA, warps 9-11 only
// donate their registers).  Registers are allocated to a CTA in units of 4 warps, so a 9-warp block costs as
// much as a 12-warp one; the block is compiled at 65536/384 -> 168 regs/thread and re-balanced with setmaxnreg:
// producer warpgroup 168 -> 40, consumer warpgroups 168 -> 232 (4*32*40 + 8*32*232 = 64512 <= 65536), which
// keeps the 128 accumulator registers + double-buffered fragments out of local memory.
constexpr int TMA_THREADS = CONSUMER_THREADS + 128;
// A segment = a contiguous run of k-tiles of one output tile.  Phase 1 (data-parallel): whole tiles (or split-K
// units) strided over the grid.  Phase 2 (stream-K): the iteration space sk_tiles x k_tiles is cut into gridDim.x
// equal ranges; range ids run AGAINST blockIdx so that the CTA that finishes a tile (its owner, the range holding
// k-tile 0) only ever waits for CTAs with a lower blockIdx, which the hardware dispatched earlier.
struct Segment {
  int tile, kb, ke;   // rasterised tile index, k-tile range [kb, ke)
  int kind;           // see below
  int slot;           // workspace slot (split-K: unit index; stream-K contributor: own range id)
  int wave;           // phase 1: how many units this CTA has done before this one; -1 in phase 2
};
// Skew between the two consumer warps of an SM sub-partition (w and w + 4), in 64ths of the pair's time per stage; 0 = off.
// Measured (tools/history/r02_call20.sh / r02_call21.sh, profiles/r02_consumer_warp_skew.txt): 2/64 .. 8/64 of a stage take the
// 128 x 128 kernel from 36.52 to 36.70-36.76 TFLOP/s at 16384^3 (8192^3 36.39 -> 36.67, 4096^3 36.19 -> 36.36) and a whole
// 128 x 64 tile from 139.3 to 137.9 us; 12/64 and more are WORSE than none (36.35), and the 64 x 64 tile gains nothing.
#ifndef DG_SKEW_64THS
#define DG_SKEW_64THS (DG_BM * DG_BN >= 128 * 64 ? 4 : 0)
#endif
constexpr long long SKEW_CLK = CONSUMER_WARPS == 8 ? (long long)FRAG_M * FRAG_N * 4 * 16 * 2 * DG_SKEW_64THS / 64 : 0;
// Unroll factor of the consumers' k-tile loop, 0 = ptxas' choice (one stage per trip on the 128-wide tiles, two on the
// 64 x 64 one).  Forcing 2 or 4 is SLOWER (profiles/r02_kt_unroll.txt: 16384^3 36.74 -> 36.18 / 35.91 TFLOP/s, a whole
// 128 x 128 tile 270.8 -> 275.6 / 277.8 us): a stage of the big tile is already ~5 KB of SASS.
#ifndef DG_KT_UNROLL
#define DG_KT_UNROLL 0
#endif
#ifndef DG_UNIFORM_PROBE
#define DG_UNIFORM_PROBE 0
#endif
#ifndef DG_NO_SYNCWARP_ARRIVE
#define DG_NO_SYNCWARP_ARRIVE 0
#endif
#ifndef DG_EARLY_DESC_PREFETCH
#define DG_EARLY_DESC_PREFETCH 0
#endif
constexpr int KT_UNROLL = DG_KT_UNROLL > 0 ? DG_KT_UNROLL : 1;
// Two accumulator sets per warp (even / odd k-steps, added once per segment): with a 16 x 16 warp tile a warp owns only four
// accumulators, so every DMMA of a k-step waits for the same accumulator's DMMA of the previous one (26 clk) unless the
// other warp of the sub-partition fills the gap; the second set doubles the distance.  Only for tiles whose accumulators
// are few (the 64 x 32 build); changes the order of the sums inside a tile (same bound, still deterministic).
#ifndef DG_DUAL_ACC
#define DG_DUAL_ACC 0
#endif
constexpr int SYNC_SLACK = 2;  // k-lockstep: a producer may run this many sync points ahead of the slowest CTA
enum { SEG_DIRECT = 0, SEG_SPLITK = 1, SEG_SK_OWNER = 2, SEG_SK_CONTRIB = 3 };

struct SegmentWalker {
  int u, num_units, stride;   // phase 1
  long long it, end;          // phase 2
  int rid;
  __device__ SegmentWalker(const Params& P, int num_units_)
      : u(blockIdx.x), num_units(num_units_), stride(gridDim.x) {
    rid = gridDim.x - 1 - blockIdx.x;
    const long long total = (long long)P.sk_tiles * P.k_tiles;
    it = total * rid / gridDim.x;
    end = total * (rid + 1) / gridDim.x;
  }
  __device__ bool next(const Params& P, Segment& sg) {
    if (u < num_units) {
      const int tile = u / P.splits, split = u - tile * P.splits;
      sg.tile = tile;
      sg.kb = split * P.k_tiles_per_split;
      sg.ke = min(P.k_tiles, sg.kb + P.k_tiles_per_split);
      sg.kind = P.splits == 1 ? SEG_DIRECT : SEG_SPLITK;
      sg.slot = u;
      sg.wave = (u - (int)blockIdx.x) / stride;
      u += stride;
      return true;
    }
    if (it < end) {
      const int t = (int)(it / P.k_tiles);
      sg.tile = P.sk_tile0 + t;
      sg.kb = (int)(it - (long long)t * P.k_tiles);
      sg.ke = (int)min((long long)P.k_tiles, sg.kb + (end - it));
      sg.kind = sg.kb > 0 ? SEG_SK_CONTRIB : (sg.ke == P.k_tiles ? SEG_DIRECT : SEG_SK_OWNER);
      sg.slot = rid;
      sg.wave = -1;
      it += sg.ke - sg.kb;
      return true;
    }
    return false;
  }
};

// SPLIT (SPLIT_A | SPLIT_B bits): the operand comes in by column parity: a parity whose columns start on a 16-byte
// boundary through its own tensor map (tmA / tmA1, tmB / tmB1), the other one (Params::a_mis / b_mis) through 8-byte
// cp.async issued by warps 10-11.  tmA1 / tmB1 are only read when the bit is set.  SPLIT = 0 is the performance path and compiles to what it was before the
// split variants existed (every difference is an `if constexpr`).
template <int STAGES, int SPLIT>
__device__ __forceinline__ void tma_ws_body(const CUtensorMap& tmA, const CUtensorMap& tmA1, const CUtensorMap& tmB,
                                            const CUtensorMap& tmB1, const Params& P, int num_units) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by SWIZZLE_128B; dynamic smem base is only guaranteed 16 B -> round up
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bars = smem_base + STAGES * SLOT_BYTES;  // full[STAGES] then empty[STAGES], 8 B each (STAGES ring slots of KPS k-tiles)
  const uint32_t progress = bars + 16 * STAGES;            // segments the producer has started (paces the C prefetch)
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  if (tid == 0) stamp(P, 0);   // (debug timeline only; the one global write ahead of griddep_wait, into its own buffer)
#if DG_EARLY_DESC_PREFETCH
  // the descriptors are kernel parameters, not data of an earlier kernel: their fetch may start before the barriers exist
  if constexpr (SPLIT == SPLIT_NONE) {
    if (tid == CONSUMER_THREADS) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
  }
#endif

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < STAGES; s++) {
      // full: one arrive (the producer's expect_tx); split kernels: + one cp.async arrive-on per helper thread
      mbar_init(bars + 8 * s, 1 + (SPLIT != SPLIT_NONE ? HELPER_THREADS : 0));
      mbar_init(bars + 8 * (STAGES + s), CONSUMER_WARPS);  // empty: one arrive per consumer warp
    }
    asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(progress), "r"(0u) : "memory");
    fence_mbar_init();
  }
  __syncthreads();
  // Programmatic dependent launch: everything above overlapped the tail of the previous kernel of the stream; nothing
  // below (first global access) runs before that kernel has completed and its writes are visible.  No-ops when the launch
  // did not carry the attribute.  The dependents of THIS grid may start their own prologue as soon as every CTA is here.
  griddep_wait();
  griddep_launch_dependents();

  if (warp >= CONSUMER_WARPS) {
    // ------------------------------- producer warpgroup: one elected lane issues TMA --------------------------
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(PRODUCER_REGS));
    if (warp == CONSUMER_WARPS && lane == 0) {
#if !DG_EARLY_DESC_PREFETCH
      if constexpr (SPLIT == SPLIT_NONE) {   // (a split kernel may have been handed a dummy map for a misaligned parity)
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
      }
#endif
      // bytes per stage that arrive by TMA (the misaligned parity halves come from the helper warps)
      uint32_t tx_bytes = STAGE_BYTES;
      if constexpr ((SPLIT & SPLIT_A) != 0) tx_bytes -= (uint32_t)(P.a_mis[0] + P.a_mis[1]) * (A_STAGE_BYTES / 2);
      if constexpr ((SPLIT & SPLIT_B) != 0) tx_bytes -= (uint32_t)(P.b_mis[0] + P.b_mis[1]) * (B_STAGE_BYTES / 2);
      uint32_t s = 0, ph = 0;
      SegmentWalker walk(P, num_units);
      Segment sg;
      uint32_t seg_idx = 0;
      while (walk.next(P, sg)) {
        int tm, tn;
        decode_tile(P, sg.tile, tm, tn);
        const int m0 = tm * BM, n0 = tn * BN;
        asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(progress), "r"(++seg_idx) : "memory");
        const bool lockstep = P.sync_ctr != nullptr && sg.wave >= 0 && sg.wave < P.sync_waves;
        const int sync_per_tile = (P.k_tiles + P.sync_kt - 1) / max(P.sync_kt, 1);
        int next_sync = sg.kb, point = sg.wave * sync_per_tile;
        for (int kt = sg.kb; kt < sg.ke; kt++) {
          if (lockstep && kt == next_sync) {  // tick this point; wait for everyone at the point SYNC_SLACK back
            red_add_relaxed_gpu(P.sync_ctr + point, 1u);
            if (point >= SYNC_SLACK)
              while (ld_relaxed_gpu(P.sync_ctr + point - SYNC_SLACK) < gridDim.x) __nanosleep(64);
            point++;
            next_sync += P.sync_kt;
          }
          // k-tile number `sub` of its ring slot; the slot's barrier is armed once, for all the k-tiles it will hold
          const int sub = KPS == 1 ? 0 : (kt - sg.kb) % KPS;
          const bool slot_done = KPS == 1 || sub == KPS - 1 || kt + 1 == sg.ke;
          const uint32_t full = bars + 8 * s;
          const uint32_t dstA = smem_base + s * SLOT_BYTES + sub * STAGE_BYTES, dstB = dstA + A_STAGE_BYTES;
          if (sub == 0) mbar_wait(bars + 8 * (STAGES + s), ph ^ 1);  // slot free?
#ifdef DG_DEBUG_NO_TMA
          // measurement build only (make libdbg): the stage is handed over without loading anything, so the consumers'
          // loop runs against barriers that are always ready -- separates "waiting for operands" from the loop's own cost
          if constexpr (SPLIT == SPLIT_NONE) {
            if (sub == 0) mbar_arrive(full);
            if (slot_done && ++s == STAGES) { s = 0; ph ^= 1; }
            continue;
          }
#endif
          if (sub == 0) {
            const uint32_t n_kt = KPS == 1 ? 1u : (uint32_t)min(KPS, sg.ke - kt);
            if constexpr (SPLIT == SPLIT_NONE) mbar_arrive_expect_tx(full, n_kt * STAGE_BYTES);
            else mbar_arrive_expect_tx(full, n_kt * tx_bytes);
          }
          const int k0 = kt * BK;
          if constexpr ((SPLIT & SPLIT_A) != 0) {
            // A by k parity: per 16-row sub-tile one box of 16(m) x 8(even k) into rows 0..7, one of the odd k into 8..15
#pragma unroll
            for (int sub = 0; sub < BM / 16; sub++) {
              if (!P.a_mis[0]) tma_load_2d(dstA + sub * 2048, &tmA, full, m0 + sub * 16, k0 >> 1);
              if (!P.a_mis[1]) tma_load_2d(dstA + sub * 2048 + 1024, &tmA1, full, m0 + sub * 16, k0 >> 1);
            }
          } else {
#pragma unroll
            for (int sub = 0; sub < BM / 16; sub++)  // A: 8 boxes of 16(m) x 16(k); OOB rows/cols arrive as zeros
              tma_load_2d(dstA + sub * 2048, &tmA, full, m0 + sub * 16, k0);
          }
          if constexpr ((SPLIT & SPLIT_B) != 0) {
            // B by column parity: 16(k) x 64(even n) into rows 0..63, the odd columns into rows 64..127
            if (!P.b_mis[0]) tma_load_2d(dstB, &tmB, full, k0, n0 >> 1);
            if (!P.b_mis[1]) tma_load_2d(dstB + B_STAGE_BYTES / 2, &tmB1, full, k0, n0 >> 1);
          } else {
            tma_load_2d(dstB, &tmB, full, k0, n0);   // B: one box of 16(k) x 128(n)
          }
          if (slot_done && ++s == STAGES) { s = 0; ph ^= 1; }
        }
        if (sg.kind == SEG_SK_OWNER && !P.sk_parallel) {
          // serial stream-K fix-up: the partials of the ranges that hold the rest of this tile come in through the
          // SAME ring, PARTIAL_CHUNKS stages per partial, as soon as their flag is up, so the owner's consumers find them in
          // shared memory when their main loop ends instead of paying two global round trips per partial.  An owner
          // segment is the last of its range, and a contributor never waits for anybody: no cycle.
          const long long total = (long long)P.sk_tiles * P.k_tiles;
          const long long tile_end = (long long)(sg.tile - P.sk_tile0 + 1) * P.k_tiles;
          for (int r = sg.slot + 1; r < (int)gridDim.x && total * r / gridDim.x < tile_end; r++) {
            while (ld_acquire_gpu(P.sk_flags + r) != P.sk_epoch) __nanosleep(64);
            fence_proxy_async_global();
            const char* src = reinterpret_cast<const char*>(P.workspace + (size_t)r * (BM * BN));
#pragma unroll 1
            for (int c = 0; c < PARTIAL_CHUNKS; c++) {
              mbar_wait(bars + 8 * (STAGES + s), ph ^ 1);
              const uint32_t full = bars + 8 * s;
              mbar_arrive_expect_tx(full, PARTIAL_CHUNK_BYTES);
              bulk_load_1d(smem_base + s * SLOT_BYTES, src + (size_t)c * PARTIAL_CHUNK_BYTES, PARTIAL_CHUNK_BYTES, full);
              if (++s == STAGES) { s = 0; ph ^= 1; }
            }
          }
        }
      }
    } else if (warp == CONSUMER_WARPS + 1) {
      // ---------------- warp 9: pull the C tile of every segment that ends in an epilogue into L2 -------------------
      // Without it all CTAs hit their epilogue together and read 148 x 128 KB of C from DRAM in four dependent
      // rounds (~9 us per tile, measured with the timeline stamps); prefetched, the rounds are L2 hits.  Paced by
      // the producer's segment counter so the prefetch runs during the main loop of the same tile, never far ahead.
      if (P.beta != 0.0) {
        SegmentWalker walk(P, num_units);
        Segment sg;
        uint32_t seg_idx = 0;
        while (walk.next(P, sg)) {
          ++seg_idx;
          if (sg.kind != SEG_DIRECT && sg.kind != SEG_SK_OWNER) continue;
          uint32_t seen;
          do {
            asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(seen) : "r"(progress) : "memory");
            if (seen < seg_idx) __nanosleep(256);
          } while (seen < seg_idx);
          int tm, tn;
          decode_tile(P, sg.tile, tm, tn);
          const int m0 = tm * BM, n0 = tn * BN;
          for (int i = lane; i < BN * (BM / 16); i += 32) {   // one 128-byte line (16 doubles) per prefetch
            const int n = n0 + i / (BM / 16), m = m0 + (i % (BM / 16)) * 16;
            if (n < P.N && m < P.M)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(P.C + (long long)n * P.ldc + m));
          }
        }
      }
    } else if (SPLIT != SPLIT_NONE && warp > CONSUMER_WARPS + 1) {
      // ---------------- helper warps (10, 11) of the split kernels: cp.async staging of the misaligned parity halves --------
      // They walk the same segments and stages as the producer, wait for the same empty barriers, and every thread
      // arrives once per stage on the full barrier (arrive-on of its own copies), also for the stages that carry a
      // stream-K partial, so the barrier's arrival count is the same for every phase.
      if constexpr (SPLIT != SPLIT_NONE) {
        const int t = tid - (CONSUMER_THREADS + 64);
        uint32_t s = 0, ph = 0;
        SegmentWalker walk(P, num_units);
        Segment sg;
        while (walk.next(P, sg)) {
          int tm, tn;
          decode_tile(P, sg.tile, tm, tn);
          const int m0 = tm * BM, n0 = tn * BN;
#pragma unroll 1
          for (int kt = sg.kb; kt < sg.ke; kt++) {
            const int sub = KPS == 1 ? 0 : (kt - sg.kb) % KPS;
            const bool slot_done = KPS == 1 || sub == KPS - 1 || kt + 1 == sg.ke;
            if (sub == 0) mbar_wait(bars + 8 * (STAGES + s), ph ^ 1);
            const uint32_t dstA = smem_base + s * SLOT_BYTES + sub * STAGE_BYTES, dstB = dstA + A_STAGE_BYTES;
            const int k0 = kt * BK;
            if constexpr ((SPLIT & SPLIT_A) != 0) {
#pragma unroll 1
              for (int p = 0; p < 2; p++)
                if (P.a_mis[p]) helper_copy_a_half(P, dstA, m0, k0, p, t);
            }
            if constexpr ((SPLIT & SPLIT_B) != 0) {
#pragma unroll 1
              for (int p = 0; p < 2; p++)
                if (P.b_mis[p]) helper_copy_b_half(P, dstB, n0, k0, p, t);
            }
            if (slot_done) {   // one arrival per helper thread and slot, behind the copies of all its k-tiles
              cp_async_mbar_arrive_noinc(bars + 8 * s);
              if (++s == STAGES) { s = 0; ph ^= 1; }
            }
          }
          if (sg.kind == SEG_SK_OWNER && !P.sk_parallel) {
            const long long total = (long long)P.sk_tiles * P.k_tiles;
            const long long tile_end = (long long)(sg.tile - P.sk_tile0 + 1) * P.k_tiles;
            for (int r = sg.slot + 1; r < (int)gridDim.x && total * r / gridDim.x < tile_end; r++) {
#pragma unroll 1
              for (int c = 0; c < PARTIAL_CHUNKS; c++) {
                mbar_wait(bars + 8 * (STAGES + s), ph ^ 1);
                mbar_arrive(bars + 8 * s);
                if (++s == STAGES) { s = 0; ph ^= 1; }
              }
            }
          }
        }
      }
    } else if (warp > CONSUMER_WARPS + 1 && P.fetch_bytes) {
      // ---------------- fetch warps (10, 11): NVLink pull of the NEXT K-slab of A, overlapped with the math ---------
      // every CTA copies one contiguous 1/gridDim.x share; 4 x 16 B per lane in flight
      constexpr int FETCH_THREADS = 64, UNROLL = 4;
      const unsigned long long n16 = P.fetch_bytes >> 4;
      const unsigned long long per = (n16 + gridDim.x - 1) / gridDim.x;
      const unsigned long long beg = per * blockIdx.x, end = beg + per < n16 ? beg + per : n16;
      const int4* src = reinterpret_cast<const int4*>(P.fetch_src);
      int4* dst = reinterpret_cast<int4*>(P.fetch_dst);
      const int t = tid - (CONSUMER_THREADS + 64);
      unsigned long long i = beg + t;
      for (; i + (UNROLL - 1) * FETCH_THREADS < end; i += UNROLL * FETCH_THREADS) {
        int4 v[UNROLL];
#pragma unroll
        for (int j = 0; j < UNROLL; j++) v[j] = ld_stream_16(src + i + j * FETCH_THREADS);
#pragma unroll
        for (int j = 0; j < UNROLL; j++) st_stream_16(dst + i + j * FETCH_THREADS, v[j]);
      }
      for (; i < end; i += FETCH_THREADS) st_stream_16(dst + i, ld_stream_16(src + i));
    }
  } else {
    // ------------------------------- consumer warps: DMMA main loop + epilogue -------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CONSUMER_REGS));
    const int warp_m = warp % WARPS_M, warp_n = warp / WARPS_M;
    typename LaneOffsetsFor<SPLIT>::type lo;
    make_lane_offsets_t(lo, lane, warp_m, warp_n);
    double acc[FRAG_N][FRAG_M][2];
#if DG_DUAL_ACC
    double acc_odd[FRAG_N][FRAG_M][2];
#define DG_ACC_ODD acc_odd
#else
#define DG_ACC_ODD acc
#endif
    Frags f0, f1;
    uint32_t s = 0, ph = 0;
    SegmentWalker walk(P, num_units);
    Segment sg;
    int tslot = 1;
    if (tid == 0) stamp(P, tslot++);
    int fix_tile[2] = {-1, -1};  // parallel fix-up: tiles this range holds a partial of (head, tail)
    while (walk.next(P, sg)) {
      zero_acc(acc);
#if DG_DUAL_ACC
      zero_acc(acc_odd);
#endif
      if (sg.kb < sg.ke) {
        // software pipeline over k-steps: the fragments of step i+1 are loaded while the 32 DMMAs of step i issue,
        // INCLUDING across the stage boundary (wait on the next stage's barrier before the last DMMA group)
        mbar_wait(bars + 8 * s, ph);
        if (tid == 0) stamp(P, tslot++);   // first operands of the segment have landed
        if constexpr (SKEW_CLK > 0) {
          // the two consumer warps of an SM sub-partition (w, w + 4) would otherwise walk the stages in lockstep and
          // reach every stage boundary (barrier probe, branches, arrive: ~75 clk without a DMMA to issue) together;
          // held back by a fraction of a stage, each one's boundary falls into the other's DMMA stream
          consumer_bar_sync();
          if (warp >= CONSUMER_WARPS / 2) {
            const long long t0 = clock64();
            while (clock64() - t0 < SKEW_CLK) {}
          }
        }
        load_frags(f0, smem_base + s * SLOT_BYTES, smem_base + s * SLOT_BYTES + A_STAGE_BYTES, lo, 0);
        int sub = 0;   // k-tile of the current ring slot (KPS > 1 only)
#if DG_KT_UNROLL > 0   // measurement builds only (make libunroll); left to ptxas otherwise (it unrolls the 64 x 64 loop by 2)
#pragma unroll KT_UNROLL
#endif
        for (int kt = sg.kb; kt < sg.ke; kt++) {
          const uint32_t sA = smem_base + s * SLOT_BYTES + (KPS == 1 ? 0 : sub * STAGE_BYTES), sB = sA + A_STAGE_BYTES;
          const uint32_t cur_empty = bars + 8 * (STAGES + s);
          const bool has_next = kt + 1 < sg.ke;
          // boundary: this k-tile is the last one of its slot -- only then is there a barrier to probe and a slot to give back
          const bool boundary = KPS == 1 || sub == KPS - 1 || !has_next;
          if (boundary) { sub = 0; if (++s == STAGES) { s = 0; ph ^= 1; } } else { sub++; }
          const uint32_t nA = smem_base + s * SLOT_BYTES + (KPS == 1 ? 0 : sub * STAGE_BYTES);   // where the next k-tile lies
          load_frags(f1, sA, sB, lo, 1);
          mma_frags(acc, f0);
          load_frags(f0, sA, sB, lo, 2);
          mma_frags(DG_ACC_ODD, f1);
          load_frags(f1, sA, sB, lo, 3);
          // probe the NEXT stage's barrier AFTER this stage's last loads (its acquire would hold them back) and
          // consume the answer one DMMA group (32 x 16 clk) later
#if DG_UNIFORM_PROBE
          // the probe's answer made warp-uniform by a vote: the branch on it is then a uniform one and no reconvergence
          // point sits in front of the next DMMA group (tools/loop_bubble_probe2.cu mode 8: 4138 -> 4118 clk per k-tile)
          const uint32_t next_ready = __all_sync(0xffffffffu, (has_next && boundary) ? mbar_test_wait(bars + 8 * s, ph) : (boundary ? 0u : 1u));
#else
          const uint32_t next_ready = (has_next && boundary) ? mbar_test_wait(bars + 8 * s, ph) : (boundary ? 0u : 1u);
#endif
          mma_frags(acc, f0);
          if (has_next) {
            if (!next_ready) mbar_wait(bars + 8 * s, ph);  // rare: the producer normally runs stages ahead
            load_frags(f0, nA, nA + A_STAGE_BYTES, lo, 0);
          }
          mma_frags(DG_ACC_ODD, f1);  // consumes the last fragments read from the current stage ...
          if (boundary) {
#if !DG_NO_SYNCWARP_ARRIVE
            __syncwarp();
#endif
            if (lane == 0) mbar_arrive(cur_empty);  // ... so the slot can go back to the producer
          }
        }
      }
#if DG_DUAL_ACC
#pragma unroll
      for (int i = 0; i < FRAG_N; i++)
#pragma unroll
        for (int j = 0; j < FRAG_M; j++) { acc[i][j][0] += acc_odd[i][j][0]; acc[i][j][1] += acc_odd[i][j][1]; }
#endif
      int tm, tn;
      decode_tile(P, sg.tile, tm, tn);
      if (tid == 0) stamp(P, tslot++);     // main loop of the segment done
      if (sg.kind == SEG_DIRECT) {
        epilogue_store<SPLIT>(acc, P, tm * BM, tn * BN, warp_m, warp_n, lane);
      } else if (sg.kind == SEG_SPLITK) {
        store_partial(acc, P.workspace + (size_t)sg.slot * (BM * BN), tid);
      } else if (P.sk_parallel) {
        // parallel fix-up, phase A: every partial goes to the workspace (slot 2r = head, 2r+1 = tail of range r)
        const int slot = 2 * sg.slot + (sg.kind == SEG_SK_OWNER ? 1 : 0);
        store_partial(acc, P.workspace + (size_t)slot * (BM * BN), tid);
        __threadfence();
        consumer_bar_sync();
        if (tid == 0) st_release_gpu(P.sk_flags + slot, P.sk_epoch);
        fix_tile[sg.kind == SEG_SK_OWNER ? 1 : 0] = sg.tile;
      } else if (sg.kind == SEG_SK_CONTRIB) {
        store_partial(acc, P.workspace + (size_t)sg.slot * (BM * BN), tid);
        __threadfence();
        fence_proxy_async_global();   // the owner reads this slot with a bulk copy (async proxy)
        consumer_bar_sync();
        if (tid == 0) st_release_gpu(P.sk_flags + sg.slot, P.sk_epoch);
      } else {  // SEG_SK_OWNER: add the partials of the ranges that hold the rest of this tile, in range order
        const long long total = (long long)P.sk_tiles * P.k_tiles;
        const long long tile_end = (long long)(sg.tile - P.sk_tile0 + 1) * P.k_tiles;
        for (int r = sg.slot + 1; r < (int)gridDim.x && total * r / gridDim.x < tile_end; r++) {
#pragma unroll
          for (int c = 0; c < PARTIAL_CHUNKS; c++) {   // the producer fetched partial r into four stages of the ring
            mbar_wait(bars + 8 * s, ph);
            add_partial_smem(acc, smem_base + s * SLOT_BYTES, c, tid);
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + 8 * (STAGES + s));
            if (++s == STAGES) { s = 0; ph ^= 1; }
          }
        }
        epilogue_store<SPLIT>(acc, P, tm * BM, tn * BN, warp_m, warp_n, lane);
      }
      if (tid == 0) stamp(P, tslot++);     // epilogue / partial store / fix-up of the segment done
    }
    if (P.sk_parallel) {
      // parallel fix-up, phase B: the R ranges that hold a piece of a tile each reduce fragments q = s, s+R, ... of it
      // over all R partials (ascending range order -> deterministic) and finish those fragments.
      const int G = gridDim.x, rid = G - 1 - (int)blockIdx.x;
      const long long total = (long long)P.sk_tiles * P.k_tiles;
      auto range_of = [&](long long x) {
        int r = (int)(x * G / total);
        while (r + 1 < G && total * (r + 1) / G <= x) r++;
        while (r > 0 && total * r / G > x) r--;
        return r;
      };
#pragma unroll 1
      for (int h = 0; h < 2; h++) {
        if (fix_tile[h] < 0) continue;
        const long long x0 = (long long)(fix_tile[h] - P.sk_tile0) * P.k_tiles;
        const int r_first = range_of(x0), r_last = range_of(x0 + P.k_tiles - 1);
        const int R = r_last - r_first + 1, me = rid - r_first;
        if (tid < R) {
          const int r = r_first + tid;
          const unsigned int* f = P.sk_flags + 2 * r + (r == r_first ? 1 : 0);
          while (ld_acquire_gpu(f) != P.sk_epoch) __nanosleep(64);
        }
        consumer_bar_sync();
        int tm, tn;
        decode_tile(P, fix_tile[h], tm, tn);
        // four fragments per pass so that 4 partial loads per range and 4 C loads are in flight together
        const bool interior = P.c_vec2 && (tm * BM + BM <= P.M) && (tn * BN + BN <= P.N);
        const int lr = lane >> 2, lq = lane & 3;
#pragma unroll 1
        for (int q0 = me; q0 < FRAG_N * FRAG_M; q0 += 4 * R) {
          double s0[4] = {0.0, 0.0, 0.0, 0.0}, s1[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 4
          for (int r = r_first; r <= r_last; r++) {
            const int slot = 2 * r + (r == r_first ? 1 : 0);
            const double2* base = reinterpret_cast<const double2*>(P.workspace + (size_t)slot * (BM * BN)) + tid;
            double2 v[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
              const int q = min(q0 + u * R, FRAG_N * FRAG_M - 1);  // clamped: the extra value is discarded below
              v[u] = __ldcg(base + q * CONSUMER_THREADS);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) { s0[u] += v[u].x; s1[u] += v[u].y; }
          }
          if (interior) {
            double* p[4];
            double2 c[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
              const int q = min(q0 + u * R, FRAG_N * FRAG_M - 1), i = q / FRAG_M, j = q % FRAG_M;
              p[u] = P.C + (long long)(tn * BN + tile_col<SPLIT>(warp_n * WN + i * 8 + perm_n_t<SPLIT>(lr))) * P.ldc +
                     tm * BM + warp_m * WM +
                     map_m(2 * lq) + (j >> 1) * 16 + (j & 1) * 4;
              c[u] = make_double2(0.0, 0.0);
            }
            if (P.beta != 0.0) {
#pragma unroll
              for (int u = 0; u < 4; u++) c[u] = *reinterpret_cast<const double2*>(p[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
              if (q0 + u * R < FRAG_N * FRAG_M)
                *reinterpret_cast<double2*>(p[u]) =
                    make_double2(fma(P.beta, c[u].x, P.alpha * s0[u]), fma(P.beta, c[u].y, P.alpha * s1[u]));
            }
          } else {
#pragma unroll 1
            for (int u = 0; u < 4; u++) {
              const int q = q0 + u * R;
              if (q < FRAG_N * FRAG_M)
                epilogue_frag<SPLIT>(s0[u], s1[u], P, tm * BM, tn * BN, warp_m, warp_n, lane, q / FRAG_M, q % FRAG_M);
            }
          }
        }
      }
      if (tid == 0) stamp(P, 15);
    }
  }
}

template <int STAGES>
__global__ void __launch_bounds__(TMA_THREADS, DG_MIN_CTAS)
dgemm_tma_ws_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params P,
                    int num_units) {
  tma_ws_body<STAGES, SPLIT_NONE>(tmA, tmA, tmB, tmB, P, num_units);
}
// operands TMA cannot describe with one map (odd LDA / LDB, 8-byte aligned bases): parity-split maps, no copy
template <int STAGES, int SPLIT>
__global__ void __launch_bounds__(TMA_THREADS, DG_MIN_CTAS)
dgemm_tma_ws_split_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmA1,
                          const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmB1,
                          const Params P, int num_units) {
  tma_ws_body<STAGES, SPLIT>(tmA, tmA1, tmB, tmB1, P, num_units);
}

#undef DG_ACC_ODD
}  // namespace DG_NAMESPACE
