// dgemm_device.cuh — device-side building blocks of the B200 DGEMM (column-major C = alpha*A*B + beta*C).
//
// Replaces the reference's micro/macro kernel (include/kernel18.h:113-626: COMPUTE_m24n8 ... macro_kernel_k18),
// its packing routines (packing_a_k18 / packing_b_k18, kernel18.h:48-110) and its separate beta pass
// (scale_c_k18, kernel18.h:12-46).  Nothing here is translated from that code: the register tile is a
// warp-level FP64 tensor-core tile (mma.sync.m8n8k4.f64 -> SASS DMMA.8x8x4, the only native FP64 MMA on
// sm_100a), "packing" is a swizzled shared-memory tile written by TMA (or cp.async), and beta is fused
// into the epilogue.
//
// CTA tile 128(m) x 128(n) x 16(k) per pipeline stage, 8 consumer warps, warp tile 64(m) x 32(n):
//   * A stage  = 8 sub-tiles of [16 k][16 m] doubles (2 KB each, m contiguous = column-major A as it lies
//                in HBM), 128-byte rows, SWIZZLE_128B with the k index as the swizzle row;
//   * B stage  = [128 n][16 k] doubles (k contiguous = column-major B), 128-byte rows, SWIZZLE_128B with
//                the n index as the swizzle row.
// The MMA is fed "transposed": the m8n8k4 row operand comes from B (rows = n), the column operand from A
// (cols = m), so every lane's accumulator pair is two CONSECUTIVE ROWS of one column of C -> 16-byte
// stores into column-major C.  Which n / m a lane feeds is permuted (kPermN / kMapM) so that every
// LDS.64 of a fragment is bank-conflict free under the 128B swizzle (tools/check_swizzle.py enumerates it).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "dgemm_params.h"

// Tile configuration of this translation unit.  The library is built from two: the 128 x 128 CTA tile (namespace dg, the
// performance path and everything else) and a 64 x 64 CTA tile (namespace dg64, dgemm_kernels_small.cu: the same TMA
// kernel body for problems with too few 128 x 128 tiles to fill the GPU; 32 x 16 warp tiles, 32 KB stream-K partials).
#ifndef DG_NAMESPACE
#define DG_NAMESPACE dg
#endif
#ifndef DG_BM
#define DG_BM 128
#endif
#ifndef DG_BN
#define DG_BN 128
#endif

namespace DG_NAMESPACE {
#ifdef DG_SMALL_TILE
using dg::Params;
using dg::SPLIT_NONE;
using dg::SPLIT_A;
using dg::SPLIT_B;
using dg::BK;
#endif

constexpr int BM = DG_BM;
constexpr int BN = DG_BN;
constexpr int A_STAGE_BYTES = BM * BK * 8;  // 16 KiB
constexpr int B_STAGE_BYTES = BN * BK * 8;  // 16 KiB
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
// k-tiles per ring slot of the TMA kernel (one full / empty mbarrier pair per slot).  Every slot boundary costs a
// sub-partition ~75 clk whatever the tile (profiles/r02_loop_with_and_without_tma.txt), which is 2 % of a 128 x 128 stage
// but 15 % of a 64 x 32 one; with two k-tiles per slot the consumers cross half as many.  1 = the layout EVERY shipped
// kernel uses (the 64 x 32 build included: 2 was measured and rejected); `make libkps` builds all of
// them with 2 for A/B timing.
#ifndef DG_KT_PER_STAGE
#define DG_KT_PER_STAGE 1
#endif
constexpr int KPS = DG_KT_PER_STAGE;
constexpr int SLOT_BYTES = KPS * STAGE_BYTES;
// Warp grid inside the CTA tile, a build-time switch (-DDG_WARPS_M=.. -DDG_WARPS_N=..):
//   2 x 4 (default)  8 consumer warps, 64x32 warp tiles, 128 accumulator registers, 2 warps per SM sub-partition
//   4 x 4            16 consumer warps, 32x32 warp tiles, 64 accumulator registers, 4 warps per SM sub-partition
#ifndef DG_WARPS_M
#define DG_WARPS_M 2
#endif
#ifndef DG_WARPS_N
#define DG_WARPS_N 4
#endif
constexpr int WARPS_M = DG_WARPS_M;
constexpr int WARPS_N = DG_WARPS_N;
constexpr int CONSUMER_WARPS = WARPS_M * WARPS_N;
constexpr int CONSUMER_THREADS = CONSUMER_WARPS * 32;
// register split between the producer warpgroup and the consumer warps (setmaxnreg); the CTA is compiled at
// 65536 / (CONSUMER_THREADS + 128) registers per thread, rounded down to 8
#ifndef DG_PRODUCER_REGS
#define DG_PRODUCER_REGS (DG_WARPS_M * DG_WARPS_N == 8 ? 40 : 24)
#endif
#ifndef DG_CONSUMER_REGS
#define DG_CONSUMER_REGS (DG_WARPS_M * DG_WARPS_N == 8 ? 232 : 112)
#endif
constexpr int PRODUCER_REGS = DG_PRODUCER_REGS;
constexpr int CONSUMER_REGS = DG_CONSUMER_REGS;
constexpr int WM = BM / WARPS_M;  // 64
constexpr int WN = BN / WARPS_N;  // 32
constexpr int FRAG_M = WM / 8;    // 8 column fragments (from A)
constexpr int FRAG_N = WN / 8;    // 4 row fragments (from B)
constexpr int ACC_PER_THREAD = FRAG_M * FRAG_N * 2;  // 64 doubles = 128 registers



__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void stamp(const Params& P, int slot) {
#ifdef DG_STAMP_CLOCK64   // measurement build (make libclk): SM cycles instead of nanoseconds -- differences inside ONE CTA only
  if (P.timeline && slot < 16) P.timeline[blockIdx.x * 16 + slot] = (unsigned long long)clock64();
#else
  if (P.timeline && slot < 16) P.timeline[blockIdx.x * 16 + slot] = globaltimer_ns();
#endif
}

// ---- lane -> matrix index maps (see header comment) -------------------------------------------------------
// row fragment (from B): mma row r = lane/4 feeds n = 8*t + kPermN(r)
__device__ __forceinline__ int perm_n(int r) { return ((r & 3) << 1) | (r >> 2); }  // 0,2,4,6,1,3,5,7
// With A split by k parity the k values a lane quad feeds change (see LaneOffsetsKS) and the B fragment loads need a
// different row permutation to stay conflict free: 0,1,4,5,2,3,6,7 (tools/check_swizzle.py)
__device__ __forceinline__ int perm_n_ks(int r) { return (r & 1) | ((r & 2) << 1) | ((r & 4) >> 1); }
template <int SPLIT>
__device__ __forceinline__ int perm_n_t(int r) { return (SPLIT & SPLIT_A) ? perm_n_ks(r) : perm_n(r); }
// shared-memory B row of the CTA tile -> column of the tile.  B split by n parity: rows 0..63 hold the even columns,
// rows 64..127 the odd ones (two TMA boxes of 64 columns each).
template <int SPLIT>
__device__ __forceinline__ int tile_col(int row) {
  return (SPLIT & SPLIT_B) ? (((row % (BN / 2)) << 1) | (row / (BN / 2))) : row;
}
// column fragment (from A): mma col c feeds m = 16*s + 4*f + kMapM(c)
__device__ __forceinline__ int map_m(int c) { return (c & 1) | ((c & 2) << 2) | ((c & 4) >> 1); }  // 0,1,8,9,2,3,10,11

// ---- raw PTX wrappers ----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// volatile + "memory": the front end must not move a fragment load above the mbarrier wait that guards its stage
// (ptxas still schedules the resulting LDS freely between the DMMAs; try_wait has acquire semantics for it).
__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr) : "memory");
  return v;
}

// D(8x8) += Arow(8x4) * Bcol(4x8), FP64 tensor core.  SASS: DMMA.8x8x4
__device__ __forceinline__ void dmma_8x8x4(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(d0), "+d"(d1)
               : "d"(a), "d"(b));
}

// programmatic dependent launch (see dgemm_launch.h)
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(bar), "r"(parity)
      : "memory");
}
// Non-blocking probe: 1 when the phase with this parity has completed.  Issued a whole stage ahead of its use, so
// its latency (~100+ clk) hides behind DMMAs instead of stalling both warps of a sub-partition at every stage edge.
__device__ __forceinline__ uint32_t mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(done)
      : "r"(bar), "r"(parity)
      : "memory");
  return done;
}
// TMA: 2-D tiled load global -> shared, completion counted on an mbarrier.  SASS: UTMALDG.2D
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
// TMA: 1-D bulk copy global -> shared (multiple of 16 bytes, 16-byte aligned), completion on an mbarrier.  SASS: UBLKCP
__device__ __forceinline__ void bulk_load_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
// Orders generic-proxy accesses to global memory (st.global / ld.acquire) against async-proxy ones (bulk copies):
// the stream-K partials are written with st.global by one CTA and read back by another CTA's bulk copy.
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
// cp.async with zero fill (src_bytes = 0 -> the destination is filled with zeros).  SASS: LDGSTS
template <int BYTES>
__device__ __forceinline__ void cp_async_zfill(uint32_t dst, const void* src, int src_bytes) {
  if constexpr (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
// makes the mbarrier track this thread's outstanding cp.async copies: one arrival (already counted in the barrier's
// init count: .noinc) fires when all of them have landed
__device__ __forceinline__ void cp_async_mbar_arrive_noinc(uint32_t bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// ---- swizzled shared-memory offsets (bytes inside one stage's A or B region) ----------------------------
// A element (m_local, k_local): sub-tile m_local/16, row k_local, 16-byte chunk (mm/2) ^ (k_local & 7)
__device__ __forceinline__ uint32_t a_smem_off(int m_local, int k_local) {
  int mm = m_local & 15;
  return (uint32_t)((m_local >> 4) * 2048 + k_local * 128 + ((((mm >> 1) ^ (k_local & 7))) << 4) + ((mm & 1) << 3));
}
// B element (k_local, n_local): row n_local, 16-byte chunk (k_local/2) ^ (n_local & 7)
__device__ __forceinline__ uint32_t b_smem_off(int k_local, int n_local) {
  return (uint32_t)(n_local * 128 + ((((k_local >> 1) ^ (n_local & 7))) << 4) + ((k_local & 1) << 3));
}

// ---- staging of a misaligned parity half by the helper warps (64 threads, t = 0..63), zero-filled at the edges -------
constexpr int HELPER_THREADS = 64;
__device__ __forceinline__ void cp_async_8(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
// A, k parity p: the 8 k-columns k0 + 2j + p of the stage x BM rows -> smem rows p*8 + j of every 16-row sub-tile.
// Thread t copies rows t, t + 64, ... of each column (a warp instruction reads 256 contiguous bytes).  Interior stages
// take a straight-line path with precomputed offsets: the two helper warps share their SM sub-partitions with DMMA warps
// and have about 2 us per stage; the generic index arithmetic alone made them the bottleneck when both operands are split.
__device__ __forceinline__ void helper_copy_a_half(const Params& P, uint32_t dstA, int m0, int k0, int p, int t) {
  constexpr int ROWS_PER_THREAD = BM / HELPER_THREADS;   // 2 (128 x 128 tile) or 1 (64 x 64 tile)
  static_assert(ROWS_PER_THREAD * HELPER_THREADS == BM, "helper threads must tile the rows of the CTA tile");
  const double* src = P.A + (long long)(k0 + p) * P.lda + m0 + t;
  const long long step = 2 * P.lda;
  const uint32_t base = dstA + (uint32_t)(t >> 4) * 2048u + (uint32_t)p * 1024u + (uint32_t)((t & 1) << 3);
  const uint32_t c0 = (uint32_t)(t & 15) >> 1;
  if ((m0 + BM <= P.M) && (k0 + BK <= P.K)) {
#pragma unroll
    for (int j = 0; j < BK / 2; j++) {
      const uint32_t off = (uint32_t)j * 128u + ((c0 ^ (uint32_t)j) << 4);
#pragma unroll
      for (int h = 0; h < ROWS_PER_THREAD; h++)   // row t + 64h: four sub-tiles further, same chunk
        cp_async_8(base + (uint32_t)h * 4u * 2048u + off, src + j * step + 64 * h);
    }
    return;
  }
  // edge stage (last tile row / K tail): same addresses, zero fill for rows >= M and columns >= K
#pragma unroll
  for (int j = 0; j < BK / 2; j++) {
    const uint32_t off = (uint32_t)j * 128u + ((c0 ^ (uint32_t)j) << 4);
    const bool kc = k0 + 2 * j + p < P.K;
#pragma unroll
    for (int h = 0; h < ROWS_PER_THREAD; h++) {
      const bool ok = kc && (m0 + t + 64 * h < P.M);
      cp_async_zfill<8>(base + (uint32_t)h * 4u * 2048u + off, ok ? src + j * step + 64 * h : P.A, ok ? 8 : 0);
    }
  }
}
// B, column parity p: the BN/2 columns n0 + 2c + p of the tile x 16 k -> smem rows p*BN/2 + c.  Thread t copies
// k = t & 15 of columns c = (t >> 4) + 4e, e = 0..BN/8-1.
__device__ __forceinline__ void helper_copy_b_half(const Params& P, uint32_t dstB, int n0, int k0, int p, int t) {
  constexpr int COLS_PER_THREAD = (BN / 2) / (HELPER_THREADS / BK);   // 16 or 8
  const int c0 = t >> 4, kl = t & 15;
  const double* src = P.B + (long long)(n0 + 2 * c0 + p) * P.ldb + k0 + kl;
  const long long step = 8 * P.ldb;   // c += 4 -> n += 8
  const uint32_t base = dstB + (uint32_t)(p * (BN / 2) + c0) * 128u + (uint32_t)((kl & 1) << 3);
  const uint32_t x0 = (uint32_t)((kl >> 1) ^ c0) << 4, x1 = (uint32_t)((kl >> 1) ^ (c0 + 4)) << 4;   // row & 7 = c0 (+4)
  if ((n0 + BN <= P.N) && (k0 + BK <= P.K)) {
#pragma unroll
    for (int e = 0; e < COLS_PER_THREAD; e++) cp_async_8(base + (uint32_t)e * 512u + ((e & 1) ? x1 : x0), src + e * step);
    return;
  }
  const bool kv = k0 + kl < P.K;   // edge stage: zero fill for k >= K and columns >= N
#pragma unroll
  for (int e = 0; e < COLS_PER_THREAD; e++) {
    const bool ok = kv && (n0 + 2 * (c0 + 4 * e) + p < P.N);
    cp_async_zfill<8>(base + (uint32_t)e * 512u + ((e & 1) ? x1 : x0), ok ? src + e * step : P.B, ok ? 8 : 0);
  }
}

// Per-lane fragment offsets, computed once per thread.
struct LaneOffsets {
  uint32_t a[2][2];  // [k-parity p = kk&1][f]: row (p*4+q), col 4f + map_m(r), sub-tile 0
  uint32_t b[4];     // [kk]: row perm_n(r), k = kk*4+q, n-block 0
};
__device__ __forceinline__ LaneOffsets make_lane_offsets(int lane, int warp_m, int warp_n) {
  const int r = lane >> 2, q = lane & 3;
  LaneOffsets lo;
#pragma unroll
  for (int p = 0; p < 2; p++)
#pragma unroll
    for (int f = 0; f < 2; f++) lo.a[p][f] = a_smem_off(warp_m * WM + 4 * f + map_m(r), p * 4 + q);
#pragma unroll
  for (int kk = 0; kk < 4; kk++) lo.b[kk] = b_smem_off(kk * 4 + q, warp_n * WN + perm_n(r));
  return lo;
}

// A split by k parity: a 16-row A sub-tile holds the even k in rows 0..7 and the odd k in rows 8..15 (two TMA boxes of
// 8 k-columns each), i.e. k sits in row (k&1)*8 + (k>>1).  The lanes q = 0..3 of k-step kk then feed
// k = 8*(kk>>1) + 2q + ((q^kk)&1)  ->  {0,3,4,7} {1,2,5,6} {8,11,12,15} {9,10,13,14}: four distinct rows mod 8 for A
// (conflict free exactly like the unsplit layout) and both 8-byte halves of two 16-byte chunks for B.
struct LaneOffsetsKS {
  uint32_t a[4][2];  // [kk][f]
  uint32_t b[4];     // [kk]
};
__device__ __forceinline__ int ks_k(int kk, int q) { return 8 * (kk >> 1) + 2 * q + ((q ^ kk) & 1); }
__device__ __forceinline__ LaneOffsetsKS make_lane_offsets_ks(int lane, int warp_m, int warp_n) {
  const int r = lane >> 2, q = lane & 3;
  LaneOffsetsKS lo;
#pragma unroll
  for (int kk = 0; kk < 4; kk++) {
    const int k = ks_k(kk, q);
#pragma unroll
    for (int f = 0; f < 2; f++) lo.a[kk][f] = a_smem_off(warp_m * WM + 4 * f + map_m(r), (k & 1) * 8 + (k >> 1));
    lo.b[kk] = b_smem_off(k, warp_n * WN + perm_n_ks(r));
  }
  return lo;
}
template <int SPLIT> struct LaneOffsetsFor { typedef LaneOffsets type; };
template <> struct LaneOffsetsFor<SPLIT_A> { typedef LaneOffsetsKS type; };
template <> struct LaneOffsetsFor<SPLIT_A | SPLIT_B> { typedef LaneOffsetsKS type; };
__device__ __forceinline__ void make_lane_offsets_t(LaneOffsets& lo, int lane, int warp_m, int warp_n) {
  lo = make_lane_offsets(lane, warp_m, warp_n);
}
__device__ __forceinline__ void make_lane_offsets_t(LaneOffsetsKS& lo, int lane, int warp_m, int warp_n) {
  lo = make_lane_offsets_ks(lane, warp_m, warp_n);
}

// Register fragments of one k-step (4 k): 8 column fragments from A, 4 row fragments from B.
struct Frags {
  double a[FRAG_M];
  double b[FRAG_N];
};
__device__ __forceinline__ void load_frags(Frags& f, uint32_t sA, uint32_t sB, const LaneOffsets& lo, int kk) {
#pragma unroll
  for (int j = 0; j < FRAG_M; j++)  // sub-tile j/2, half f = j&1; rows kk*4.. -> (kk>>1)*1024 + (kk&1)*512
    f.a[j] = lds_f64(sA + lo.a[kk & 1][j & 1] + (kk >> 1) * 1024 + (j >> 1) * 2048);
#pragma unroll
  for (int i = 0; i < FRAG_N; i++) f.b[i] = lds_f64(sB + lo.b[kk] + i * 1024);
}
__device__ __forceinline__ void load_frags(Frags& f, uint32_t sA, uint32_t sB, const LaneOffsetsKS& lo, int kk) {
#pragma unroll
  for (int j = 0; j < FRAG_M; j++) f.a[j] = lds_f64(sA + lo.a[kk][j & 1] + (j >> 1) * 2048);
#pragma unroll
  for (int i = 0; i < FRAG_N; i++) f.b[i] = lds_f64(sB + lo.b[kk] + i * 1024);
}
__device__ __forceinline__ void mma_frags(double (&acc)[FRAG_N][FRAG_M][2], const Frags& f) {
#pragma unroll
  for (int i = 0; i < FRAG_N; i++)
#pragma unroll
    for (int j = 0; j < FRAG_M; j++) dmma_8x8x4(acc[i][j][0], acc[i][j][1], f.b[i], f.a[j]);
}

// One pipeline stage (16 k) of the warp tile: 4 k-steps x (8 + 4 fragment loads, 32 DMMAs).
__device__ __forceinline__ void mma_stage(double (&acc)[FRAG_N][FRAG_M][2], uint32_t sA, uint32_t sB,
                                          const LaneOffsets& lo) {
  Frags f0, f1;
  load_frags(f0, sA, sB, lo, 0);
  load_frags(f1, sA, sB, lo, 1);
  mma_frags(acc, f0);
  load_frags(f0, sA, sB, lo, 2);
  mma_frags(acc, f1);
  load_frags(f1, sA, sB, lo, 3);
  mma_frags(acc, f0);
  mma_frags(acc, f1);
}

__device__ __forceinline__ void zero_acc(double (&acc)[FRAG_N][FRAG_M][2]) {
#pragma unroll
  for (int i = 0; i < FRAG_N; i++)
#pragma unroll
    for (int j = 0; j < FRAG_M; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
}

// Fused epilogue: C = alpha*acc + beta*C on the lanes' own entries, straight from registers.
// beta == 0 never reads C (BLAS convention; NaN/Inf in C are overwritten).  Rows >= M / cols >= N are
// predicated off, so padding rows M..ldc-1 are never touched.
template <int SPLIT = SPLIT_NONE>
__device__ __forceinline__ void epilogue_store(const double (&acc)[FRAG_N][FRAG_M][2], const Params& P, int m0, int n0,
                                               int warp_m, int warp_n, int lane) {
  const int r = lane >> 2, q = lane & 3;
  const int mq = map_m(2 * q);  // 0, 8, 2, 10
  const double alpha = P.alpha, beta = P.beta;
  const bool use_c = beta != 0.0;
  const int mbase = m0 + warp_m * WM + mq;
  constexpr int HP = FRAG_N >= 2 ? 2 : 1;   // row fragments per pass of the fast paths (one when the warp tile is 8 columns wide)
  // Fast path (warp tile fully inside C, 16-byte aligned columns): straight-line code, 16 independent 16-byte loads
  // in flight per lane before the first use.  The predicated edge path below puts every load in its own basic
  // block, which serialises the round trips (measured: ~9 us per tile instead of ~2).
  // (with B split by column parity a warp's columns are spread over the whole tile width)
  const bool inside = (m0 + warp_m * WM + WM <= P.M) && (n0 + ((SPLIT & SPLIT_B) ? BN : warp_n * WN + WN) <= P.N);
  if (inside && !P.c_vec2) {
    // same, for a C whose columns are only 8-byte aligned (odd LDC or base): scalar accesses, still branch-free
#pragma unroll
    for (int ip = 0; ip < FRAG_N; ip += 2) {
      double* col0 = P.C + (long long)(n0 + tile_col<SPLIT>(warp_n * WN + ip * 8 + perm_n_t<SPLIT>(r))) * P.ldc + mbase;
      double* col1 = col0 + ((SPLIT & SPLIT_B) ? 16 : 8) * P.ldc;   // row + 8 stays in the same parity half
      double c[2][FRAG_M][2];
      if (use_c) {
#pragma unroll
        for (int j = 0; j < FRAG_M; j++) {
          const int o = (j >> 1) * 16 + (j & 1) * 4;
          c[0][j][0] = col0[o]; c[0][j][1] = col0[o + 1];
          if (HP == 2) { c[1][j][0] = col1[o]; c[1][j][1] = col1[o + 1]; }
        }
      }
#pragma unroll
      for (int h = 0; h < HP; h++) {
        double* col = h ? col1 : col0;
#pragma unroll
        for (int j = 0; j < FRAG_M; j++) {
          const int o = (j >> 1) * 16 + (j & 1) * 4;
          double v0 = alpha * acc[ip + h][j][0], v1 = alpha * acc[ip + h][j][1];
          if (use_c) { v0 = fma(beta, c[h][j][0], v0); v1 = fma(beta, c[h][j][1], v1); }
          col[o] = v0; col[o + 1] = v1;
        }
      }
    }
    return;
  }
  const bool interior = P.c_vec2 && inside;
  if (interior) {
#pragma unroll
    for (int ip = 0; ip < FRAG_N; ip += 2) {
      double* col0 = P.C + (long long)(n0 + tile_col<SPLIT>(warp_n * WN + ip * 8 + perm_n_t<SPLIT>(r))) * P.ldc + mbase;
      double* col1 = col0 + ((SPLIT & SPLIT_B) ? 16 : 8) * P.ldc;   // row + 8 stays in the same parity half
      double2 c[2][FRAG_M];
      if (use_c) {
#pragma unroll
        for (int j = 0; j < FRAG_M; j++) {
          c[0][j] = *reinterpret_cast<const double2*>(col0 + (j >> 1) * 16 + (j & 1) * 4);
          if (HP == 2) c[1][j] = *reinterpret_cast<const double2*>(col1 + (j >> 1) * 16 + (j & 1) * 4);
        }
      }
#pragma unroll
      for (int h = 0; h < HP; h++) {
        double* col = h ? col1 : col0;
#pragma unroll
        for (int j = 0; j < FRAG_M; j++) {
          double v0 = alpha * acc[ip + h][j][0], v1 = alpha * acc[ip + h][j][1];
          if (use_c) { v0 = fma(beta, c[h][j].x, v0); v1 = fma(beta, c[h][j].y, v1); }
          *reinterpret_cast<double2*>(col + (j >> 1) * 16 + (j & 1) * 4) = make_double2(v0, v1);
        }
      }
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < FRAG_N; i++) {
    const int n = n0 + tile_col<SPLIT>(warp_n * WN + i * 8 + perm_n_t<SPLIT>(r));
    if (n >= P.N) continue;
    double* col = P.C + (long long)n * P.ldc;
    double c0[FRAG_M], c1[FRAG_M];
    if (use_c) {
#pragma unroll
      for (int j = 0; j < FRAG_M; j++) {
        const int m = mbase + (j >> 1) * 16 + (j & 1) * 4;
        c0[j] = c1[j] = 0.0;
        if (P.c_vec2 && m + 1 < P.M) {
          double2 v = *reinterpret_cast<const double2*>(col + m);
          c0[j] = v.x; c1[j] = v.y;
        } else {
          if (m < P.M) c0[j] = col[m];
          if (m + 1 < P.M) c1[j] = col[m + 1];
        }
      }
    }
#pragma unroll
    for (int j = 0; j < FRAG_M; j++) {
      const int m = mbase + (j >> 1) * 16 + (j & 1) * 4;
      double v0 = alpha * acc[i][j][0], v1 = alpha * acc[i][j][1];
      if (use_c) { v0 = fma(beta, c0[j], v0); v1 = fma(beta, c1[j], v1); }
      if (P.c_vec2 && m + 1 < P.M) {
        *reinterpret_cast<double2*>(col + m) = make_double2(v0, v1);
      } else {
        if (m < P.M) col[m] = v0;
        if (m + 1 < P.M) col[m + 1] = v1;
      }
    }
  }
}

// Epilogue for ONE fragment (i, j) of the lane: used by the parallel stream-K fix-up, where a CTA finishes only a
// slice of the tile.  Fully predicated (edges, unaligned C).
template <int SPLIT = SPLIT_NONE>
__device__ __forceinline__ void epilogue_frag(double a0, double a1, const Params& P, int m0, int n0, int warp_m,
                                              int warp_n, int lane, int i, int j) {
  const int r = lane >> 2, q = lane & 3;
  const int n = n0 + tile_col<SPLIT>(warp_n * WN + i * 8 + perm_n_t<SPLIT>(r));
  const int m = m0 + warp_m * WM + map_m(2 * q) + (j >> 1) * 16 + (j & 1) * 4;
  if (n >= P.N || m >= P.M) return;
  double* p = P.C + (long long)n * P.ldc + m;
  double v0 = P.alpha * a0, v1 = P.alpha * a1;
  const bool two = m + 1 < P.M;
  if (P.c_vec2 && two) {
    if (P.beta != 0.0) {
      const double2 c = *reinterpret_cast<const double2*>(p);
      v0 = fma(P.beta, c.x, v0); v1 = fma(P.beta, c.y, v1);
    }
    *reinterpret_cast<double2*>(p) = make_double2(v0, v1);
  } else {
    if (P.beta != 0.0) {
      v0 = fma(P.beta, p[0], v0);
      if (two) v1 = fma(P.beta, p[1], v1);
    }
    p[0] = v0;
    if (two) p[1] = v1;
  }
}

// split-K: dump / reload raw accumulators in [reg][thread] order (fully coalesced 16-byte accesses)
__device__ __forceinline__ void store_partial(const double (&acc)[FRAG_N][FRAG_M][2], double* ws, int tid) {
  double2* p = reinterpret_cast<double2*>(ws) + tid;
#pragma unroll
  for (int i = 0; i < FRAG_N; i++)
#pragma unroll
    for (int j = 0; j < FRAG_M; j++) p[(i * FRAG_M + j) * CONSUMER_THREADS] = make_double2(acc[i][j][0], acc[i][j][1]);
}
__device__ __forceinline__ void add_partial(double (&acc)[FRAG_N][FRAG_M][2], const double* ws, int tid) {
  const double2* p = reinterpret_cast<const double2*>(ws) + tid;
#pragma unroll
  for (int i = 0; i < FRAG_N; i++)
#pragma unroll
    for (int j = 0; j < FRAG_M; j++) {
      double2 v = p[(i * FRAG_M + j) * CONSUMER_THREADS];
      acc[i][j][0] += v.x; acc[i][j][1] += v.y;
    }
}

__device__ __forceinline__ void st_release_gpu(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_relaxed_gpu(unsigned int* p, unsigned int v) {
  asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_relaxed_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned int ld_acquire_gpu(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// streaming 16-byte copy primitives for the fused fetch: peer loads bypass L1 allocation, stores go straight to L2
__device__ __forceinline__ int4 ld_stream_16(const int4* p) {
  int4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_stream_16(int4* p, const int4& v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// barrier among the 8 consumer warps only (the producer warpgroup never joins): named barrier 1
__device__ __forceinline__ void consumer_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(CONSUMER_THREADS) : "memory"); }

// stream-K partial reload: L2-coherent loads (the slot was written by another SM during this launch)
__device__ __forceinline__ void add_partial_cg(double (&acc)[FRAG_N][FRAG_M][2], const double* ws, int tid) {
  const double2* p = reinterpret_cast<const double2*>(ws) + tid;
#pragma unroll
  for (int i = 0; i < FRAG_N; i++)
#pragma unroll
    for (int j = 0; j < FRAG_M; j++) {
      double2 v = __ldcg(p + (i * FRAG_M + j) * CONSUMER_THREADS);
      acc[i][j][0] += v.x; acc[i][j][1] += v.y;
    }
}

// stream-K partial that the producer brought into a pipeline stage with bulk_load_1d: chunk c of the [reg][thread]
// layout holds fragments q = c*FPC .. c*FPC+FPC-1 for all consumer threads (16-byte LDS, conflict free)
// A partial is PARTIAL_FRAGS fragments of CONSUMER_THREADS x 16 bytes; a chunk is the largest whole number of fragments
// that divides the partial and fits one pipeline stage (128 x 128: 4 chunks of 32 KB = a whole stage; 64 x 64: 2 of 16 KB;
// 128 x 64 with 32 x 32 warp tiles: 4 chunks of 16 KB in 24 KB stages).
constexpr int PARTIAL_FRAGS = FRAG_N * FRAG_M;
constexpr int PARTIAL_FRAG_BYTES = CONSUMER_THREADS * 16;
constexpr int partial_fpc() {
  int f = PARTIAL_FRAGS;
  while (f > 1 && (f * PARTIAL_FRAG_BYTES > STAGE_BYTES || PARTIAL_FRAGS % f != 0)) f--;
  return f;
}
constexpr int PARTIAL_FPC = partial_fpc();                               // fragments per chunk
constexpr int PARTIAL_CHUNKS = PARTIAL_FRAGS / PARTIAL_FPC;              // stages per partial
constexpr int PARTIAL_CHUNK_BYTES = PARTIAL_FPC * PARTIAL_FRAG_BYTES;
static_assert(PARTIAL_CHUNKS * PARTIAL_CHUNK_BYTES == BM * BN * 8 && PARTIAL_FPC * PARTIAL_CHUNKS == PARTIAL_FRAGS &&
              PARTIAL_CHUNK_BYTES <= STAGE_BYTES, "partial tile must split into chunks of at most one stage");
__device__ __forceinline__ void add_partial_smem(double (&acc)[FRAG_N][FRAG_M][2], uint32_t stage, int chunk, int tid) {
#pragma unroll
  for (int c = 0; c < PARTIAL_CHUNKS; c++) {
    if (c != chunk) continue;   // chunk is a compile-time constant at every call site (unrolled loop)
#pragma unroll
    for (int jj = 0; jj < PARTIAL_FPC; jj++) {
      const int q = c * PARTIAL_FPC + jj, i = q / FRAG_M, j = q % FRAG_M;
      double x, y;
      asm volatile("ld.shared.v2.f64 {%0,%1}, [%2];" : "=d"(x), "=d"(y)
                   : "r"(stage + (uint32_t)(jj * CONSUMER_THREADS + tid) * 16u) : "memory");
      acc[i][j][0] += x; acc[i][j][1] += y;
    }
  }
}

// tile index (rasterised order) -> tile coordinates
__device__ __forceinline__ void decode_tile(const Params& P, int tile, int& tm, int& tn) {
  const int band_tiles = P.group_m * P.tiles_n;
  const int band = tile / band_tiles;
  const int in_band = tile - band * band_tiles;
  const int gh = min(P.group_m, P.tiles_m - band * P.group_m);
  tn = in_band / gh;
  tm = band * P.group_m + (in_band - tn * gh);
}

// work unit -> (tile_m, tile_n, split).  Units are ordered tile-major inside rasterisation bands of group_m
// tile-rows so that the CTAs resident at any moment share A row-panels and B column-panels in L2.
struct Unit { int tm, tn, split, k_begin, k_end; };
__device__ __forceinline__ Unit decode_unit(const Params& P, int u) {
  Unit w;
  const int tile = u / P.splits;
  w.split = u - tile * P.splits;
  const int band_tiles = P.group_m * P.tiles_n;
  const int band = tile / band_tiles;
  const int in_band = tile - band * band_tiles;
  const int gh = min(P.group_m, P.tiles_m - band * P.group_m);
  w.tn = in_band / gh;
  w.tm = band * P.group_m + (in_band - w.tn * gh);
  w.k_begin = w.split * P.k_tiles_per_split;
  w.k_end = min(P.k_tiles, w.k_begin + P.k_tiles_per_split);
  return w;
}

}  // namespace DG_NAMESPACE
