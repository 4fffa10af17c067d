// dgemm_kernels_cluster.cu — ladder rung 6: the TMA kernel on 2-CTA thread-block clusters with TMA multicast.
//
// Two CTAs of a cluster (the two SMs of a TPC) work on two horizontally adjacent 128 x 128 output tiles, which need the SAME
// 128 x 16 panel of A at every k-tile.  Each CTA fetches HALF of that panel (4 of the 8 boxes of 16 x 16) and TMA writes it
// into both CTAs' shared memory (cp.async.bulk.tensor ... .multicast::cluster, mask 0b11); its B tile it fetches alone.  A CTA
// then pulls 24 KB per k-tile through L2 instead of 32 KB, and the k-lockstep that the persistent kernel imposes in software
// across the whole grid (Params::sync_ctr) exists in hardware inside every pair: neither CTA can run ahead of the other,
// because a pipeline slot is only refilled once BOTH CTAs' consumers have released it.
//
//   full[s]   (per CTA) one arrive.expect_tx(32 KB) by the CTA's own producer; the bytes come from three sources: own half of
//             A, the peer's half of A (multicast signals the barrier at the same offset in every destination CTA), own B
//   empty[s]  (per CTA) 2 x 8 arrivals: every consumer warp of BOTH CTAs (the peer's arrive remotely, mapa + shared::cluster),
//             because the CTA's producer writes into both CTAs' slot s
//
// Deliberately simple around that: data-parallel tile pairs only (no stream-K tail, no split-K, no fused fetch, one tensor
// map per operand), so that the A/B against the shipped kernel isolates the effect of clusters + multicast.  Selected only
// through dgemm_b200_device_ex(variant = DGEMM_B200_VARIANT_CLUSTER); DESIGN.md section 3a records the measurements and the
// keep / reject decision.  Mirrors nothing in the reference (its "multicast" is one packed B shared by all OpenMP threads,
// include/kernel19.h:708-731).
#include "dgemm_device.cuh"
#include "dgemm_host.h"

#include <cuda.h>

#include <algorithm>
#include <atomic>

namespace dg {

constexpr int CL_STAGES = 6;
constexpr int CL_THREADS = CONSUMER_THREADS + 128;
constexpr int CL_SMEM = CL_STAGES * STAGE_BYTES + 2 * CL_STAGES * 8 + 16 + 1024;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(cta));
  return r;
}
// arrive on the barrier at shared::cluster address `remote_bar` (a peer CTA's).  Default semantics (release at CTA scope), as
// the local arrive: the consumer's reads of the slot are complete once their values have been used (register dependence),
// so no fence is needed.  The first version used .release.cluster here and .acquire.cluster in the waits; ptxas turns those
// into MEMBAR.ALL.GPU + ERRBAR + CGAERRBAR per arrive and CCTL.IVALL (L1 invalidate) per wait, and ncu showed 1.9 warps per
// issue stalled on "membar" and the DMMA pipe 83 % busy instead of 98 % (profiles/r02_ncu_cluster_vs_tma_16384.json, v1).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t remote_bar) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote_bar) : "memory");
}
// TMA 2-D tiled load, written to the same shared-memory offset of every CTA in `mask`, completion bytes counted on the
// mbarrier at the same offset in each of them.  SASS: UTMALDG.2D.MULTICAST
__device__ __forceinline__ void tma_load_2d_multicast(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(CL_THREADS, 1)
dgemm_tma_cluster_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params P,
                         int num_pairs, int pairs_n) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // same offset in both CTAs (same kernel, same layout)
  const uint32_t bars = smem_base + CL_STAGES * STAGE_BYTES;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_ctarank(), peer = rank ^ 1u;
  const int cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (tid == 0) {
#pragma unroll
    for (int s = 0; s < CL_STAGES; s++) {
      mbar_init(bars + 8 * s, 1);
      mbar_init(bars + 8 * (CL_STAGES + s), 2 * CONSUMER_WARPS);
    }
    fence_mbar_init();
  }
  cluster_sync_all();   // both CTAs' barriers exist before anybody multicasts into, or arrives on, the other one's
  griddep_wait();
  griddep_launch_dependents();

  // pair -> (tile row, pair column), rasterised in bands of group_m tile rows like the shipped kernel
  auto decode_pair = [&](int pair, int& tm, int& tnp) {
    const int band_tiles = P.group_m * pairs_n;
    const int band = pair / band_tiles, in_band = pair - band * band_tiles;
    const int gh = min(P.group_m, P.tiles_m - band * P.group_m);
    tnp = in_band / gh;
    tm = band * P.group_m + (in_band - tnp * gh);
  };

  if (warp >= CONSUMER_WARPS) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(PRODUCER_REGS));
    if (warp == CONSUMER_WARPS && lane == 0) {
      tma_prefetch_desc(&tmA);
      tma_prefetch_desc(&tmB);
      uint32_t s = 0, ph = 0;
      for (int pair = cluster_id; pair < num_pairs; pair += num_clusters) {
        int tm, tnp;
        decode_pair(pair, tm, tnp);
        const int m0 = tm * BM, n0 = (2 * tnp + (int)rank) * BN;   // (n0 >= N for the odd last pair's second tile: TMA zero-fills)
        for (int kt = 0; kt < P.k_tiles; kt++) {
          mbar_wait(bars + 8 * (CL_STAGES + s), ph ^ 1);   // slot s is free in BOTH CTAs
          const uint32_t full = bars + 8 * s;
          const uint32_t dstA = smem_base + s * STAGE_BYTES, dstB = dstA + A_STAGE_BYTES;
          mbar_arrive_expect_tx(full, STAGE_BYTES);
          const int k0 = kt * BK;
#pragma unroll
          for (int h = 0; h < BM / 32; h++) {   // my half of the A panel (sub-tiles 4*rank .. 4*rank+3), into both CTAs
            const int sub = (int)rank * (BM / 32) + h;
            tma_load_2d_multicast(dstA + sub * 2048, &tmA, full, m0 + sub * 16, k0, (uint16_t)0x3);
          }
          tma_load_2d(dstB, &tmB, full, k0, n0);
          if (++s == CL_STAGES) { s = 0; ph ^= 1; }
        }
      }
    } else if (warp == CONSUMER_WARPS + 1 && P.beta != 0.0) {
      // C tiles of this CTA into L2 ahead of their epilogues (as in the shipped kernel, unpaced: pairs are walked in order)
      for (int pair = cluster_id; pair < num_pairs; pair += num_clusters) {
        int tm, tnp;
        decode_pair(pair, tm, tnp);
        const int m0 = tm * BM, n0 = (2 * tnp + (int)rank) * BN;
        for (int i = lane; i < BN * (BM / 16); i += 32) {
          const int n = n0 + (i >> 3), m = m0 + ((i & 7) << 4);
          if (n < P.N && m < P.M) asm volatile("prefetch.global.L2 [%0];" ::"l"(P.C + (long long)n * P.ldc + m));
        }
        __nanosleep(20000);
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(CONSUMER_REGS));
    const int warp_m = warp % WARPS_M, warp_n = warp / WARPS_M;
    const LaneOffsets lo = make_lane_offsets(lane, warp_m, warp_n);
    double acc[FRAG_N][FRAG_M][2];
    Frags f0, f1;
    uint32_t s = 0, ph = 0;
    for (int pair = cluster_id; pair < num_pairs; pair += num_clusters) {
      int tm, tnp;
      decode_pair(pair, tm, tnp);
      zero_acc(acc);
      mbar_wait(bars + 8 * s, ph);
      load_frags(f0, smem_base + s * STAGE_BYTES, smem_base + s * STAGE_BYTES + A_STAGE_BYTES, lo, 0);
      for (int kt = 0; kt < P.k_tiles; kt++) {
        const uint32_t sA = smem_base + s * STAGE_BYTES, sB = sA + A_STAGE_BYTES;
        const uint32_t cur_empty = bars + 8 * (CL_STAGES + s);
        if (++s == CL_STAGES) { s = 0; ph ^= 1; }
        const bool has_next = kt + 1 < P.k_tiles;
        load_frags(f1, sA, sB, lo, 1);
        mma_frags(acc, f0);
        load_frags(f0, sA, sB, lo, 2);
        mma_frags(acc, f1);
        load_frags(f1, sA, sB, lo, 3);
        const uint32_t next_ready = has_next ? mbar_test_wait(bars + 8 * s, ph) : 0u;   // as in the shipped kernel
        mma_frags(acc, f0);
        if (has_next) {
          if (!next_ready) mbar_wait(bars + 8 * s, ph);
          load_frags(f0, smem_base + s * STAGE_BYTES, smem_base + s * STAGE_BYTES + A_STAGE_BYTES, lo, 0);
        }
        mma_frags(acc, f1);
        __syncwarp();
        if (lane == 0) {   // the slot goes back to BOTH producers: each of them writes half of A into this CTA
          mbar_arrive(cur_empty);
          mbar_arrive_cluster(mapa_shared(cur_empty, peer));
        }
      }
      epilogue_store<SPLIT_NONE>(acc, P, tm * BM, (2 * tnp + (int)rank) * BN, warp_m, warp_n, lane);
    }
  }
  // nobody leaves while its peer may still multicast into its shared memory or arrive on its barriers
  cluster_sync_all();
}

cudaError_t launch_tma_cluster(const void* const maps[2], const Params& P, int num_sms, cudaStream_t stream) {
  static std::atomic<bool> attr_done[64];
  int dev = 0;
  cudaError_t err = cudaGetDevice(&dev);
  if (err != cudaSuccess) return err;
  if (!attr_done[dev & 63]) {
    if ((err = cudaFuncSetAttribute(dgemm_tma_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CL_SMEM)) != cudaSuccess)
      return err;
    attr_done[dev & 63] = true;
  }
  const int pairs_n = (P.tiles_n + 1) / 2;
  const int num_pairs = P.tiles_m * pairs_n;
  // how many 2-CTA clusters the device can hold at once (GPC boundaries may leave an SM without a partner)
  static std::atomic<int> max_clusters[64];
  int mc = max_clusters[dev & 63].load();
  if (mc == 0) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(num_sms & ~1)); cfg.blockDim = dim3(CL_THREADS); cfg.dynamicSmemBytes = CL_SMEM;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    if (cudaOccupancyMaxActiveClusters(&mc, dgemm_tma_cluster_kernel, &cfg) != cudaSuccess || mc <= 0) {
      cudaGetLastError();
      mc = num_sms / 2;
    }
    max_clusters[dev & 63].store(mc);
  }
  const int clusters = std::max(1, std::min({num_pairs, num_sms / 2, mc}));
  const CUtensorMap &tmA = *reinterpret_cast<const CUtensorMap*>(maps[0]), &tmB = *reinterpret_cast<const CUtensorMap*>(maps[1]);
  dgemm_tma_cluster_kernel<<<2 * clusters, CL_THREADS, CL_SMEM, stream>>>(tmA, tmB, P, num_pairs, pairs_n);
  return cudaGetLastError();
}

}  // namespace dg
