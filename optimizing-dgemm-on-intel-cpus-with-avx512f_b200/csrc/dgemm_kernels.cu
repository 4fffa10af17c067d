// dgemm_kernels.cu — the sm_100a kernels of the DGEMM hot path and their launchers.
//
//   dgemm_tma_ws_kernel     persistent, warp-specialised: 1 TMA producer warp + 8 DMMA consumer warps, N-stage
//                           mbarrier ring (full/empty), fused alpha/beta epilogue.  The performance path.
//                           Plays the role of mydgemm_cpu_v18's N->K->M blocking loops + packing + macro kernel
//                           (include/kernel18.h:629-671) and of kernel19's thread-band driver (kernel19.h:678-755).
//   dgemm_cpasync_kernel    same tile math, operands staged by cp.async (8- or 16-byte, zero-filled at the
//                           edges).  Needed because TMA requires 16-byte aligned bases and strides: an odd LDA/LDB
//                           (e.g. the 1000x1003x997 config) or an odd element offset cannot be described by a
//                           tensor map.  Device-side — there is no CPU fallback.
//   splitk_reduce_kernel    deterministic ordered sum of split-K partial tiles + the fused epilogue.
//   scale_c_kernel          C = beta*C for the degenerate alpha == 0 / K == 0 calls (kernel18.h:644-645 order).
//   fill_signed_kernel      counter-based U(-1,1) generator (same formula as oracle_signed_value) for benches.
#include "dgemm_device.cuh"
#include "dgemm_tma_kernel.cuh"
#include "dgemm_host.h"
#include "dgemm_launch.h"

#include <cuda.h>
#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

namespace dg {

// =========================================================================================================
// generic loader: cp.async multistage, one CTA per work unit
// =========================================================================================================
template <int VEC>  // doubles per cp.async: 1 (any alignment) or 2 (16-byte aligned operands)
__device__ __forceinline__ void cpasync_load_stage(const Params& P, uint32_t sA, uint32_t sB, int m0, int n0, int k0,
                                                   int tid) {
  constexpr int BYTES = VEC * 8;
  // A tile: 16 k-columns of 128 consecutive rows (m fastest across threads -> coalesced)
  constexpr int A_ROW_T = BM / VEC;                   // threads covering one k-column
  constexpr int A_K_STEP = CONSUMER_THREADS / A_ROW_T;  // k-columns covered per pass
#pragma unroll
  for (int e = 0; e < BK / A_K_STEP; e++) {
    const int ml = (tid % A_ROW_T) * VEC, kl = tid / A_ROW_T + e * A_K_STEP;
    const int m = m0 + ml, k = k0 + kl;
    const bool ok = (m + VEC <= P.M) && (k < P.K);  // VEC==2: M is even-reachable only when aligned; tail handled below
    const double* src = ok ? P.A + (long long)k * P.lda + m : P.A;
    int bytes = ok ? BYTES : 0;
    if (VEC == 2 && !ok && m < P.M && k < P.K) { src = P.A + (long long)k * P.lda + m; bytes = 8; }  // odd M tail
    cp_async_zfill<BYTES>(sA + a_smem_off(ml, kl), src, bytes);
  }
  // B tile: 128 n-columns of 16 consecutive k (k fastest across threads)
  constexpr int B_COL_T = BK / VEC;
  constexpr int B_N_STEP = CONSUMER_THREADS / B_COL_T;
#pragma unroll
  for (int e = 0; e < BN / B_N_STEP; e++) {
    const int kl = (tid % B_COL_T) * VEC, nl = tid / B_COL_T + e * B_N_STEP;
    const int k = k0 + kl, n = n0 + nl;
    const bool ok = (k + VEC <= P.K) && (n < P.N);
    const double* src = ok ? P.B + (long long)n * P.ldb + k : P.B;
    int bytes = ok ? BYTES : 0;
    if (VEC == 2 && !ok && k < P.K && n < P.N) { src = P.B + (long long)n * P.ldb + k; bytes = 8; }  // odd K tail
    cp_async_zfill<BYTES>(sB + b_smem_off(kl, nl), src, bytes);
  }
}

template <int VEC, int STAGES>
__global__ void __launch_bounds__(CONSUMER_THREADS, 1) dgemm_cpasync_kernel(const Params P) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int warp_m = warp % WARPS_M, warp_n = warp / WARPS_M;
  const LaneOffsets lo = make_lane_offsets(lane, warp_m, warp_n);
  const Unit w = decode_unit(P, blockIdx.x);
  const int m0 = w.tm * BM, n0 = w.tn * BN;
  const int nk = w.k_end - w.k_begin;

  double acc[FRAG_N][FRAG_M][2];
  zero_acc(acc);
#pragma unroll
  for (int s = 0; s < STAGES - 1; s++) {  // prologue: STAGES-1 tiles in flight
    if (s < nk) {
      const uint32_t sA = smem_base + s * STAGE_BYTES;
      cpasync_load_stage<VEC>(P, sA, sA + A_STAGE_BYTES, m0, n0, (w.k_begin + s) * BK, tid);
    }
    cp_async_commit();
  }
  for (int i = 0; i < nk; i++) {
    cp_async_wait<STAGES - 2>();  // tile i has landed (for this thread's copies) ...
    __syncthreads();              // ... and for everyone's; also: everyone finished reading tile i-1's slot
    const int nxt = i + STAGES - 1;
    if (nxt < nk) {
      const uint32_t sA = smem_base + (nxt % STAGES) * STAGE_BYTES;
      cpasync_load_stage<VEC>(P, sA, sA + A_STAGE_BYTES, m0, n0, (w.k_begin + nxt) * BK, tid);
    }
    cp_async_commit();
    const uint32_t sA = smem_base + (i % STAGES) * STAGE_BYTES;
    mma_stage(acc, sA, sA + A_STAGE_BYTES, lo);
  }
  cp_async_wait<0>();
  if (P.splits == 1)
    epilogue_store(acc, P, m0, n0, warp_m, warp_n, lane);
  else
    store_partial(acc, P.workspace + (size_t)blockIdx.x * (BM * BN), tid);
}

// =========================================================================================================
// split-K fix-up, degenerate scaling, generators
// =========================================================================================================
__global__ void __launch_bounds__(CONSUMER_THREADS) splitk_reduce_kernel(const Params P) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int warp_m = warp % WARPS_M, warp_n = warp / WARPS_M;
  Params Q = P;
  Q.splits = 1;
  const Unit w = decode_unit(Q, blockIdx.x);  // blockIdx.x = tile index in rasterised order
  double acc[FRAG_N][FRAG_M][2];
  zero_acc(acc);
  for (int s = 0; s < P.splits; s++)  // fixed ascending order -> run-to-run deterministic
    add_partial(acc, P.workspace + ((size_t)blockIdx.x * P.splits + s) * (BM * BN), tid);
  epilogue_store(acc, P, w.tm * BM, w.tn * BN, warp_m, warp_n, lane);
}

__global__ void scale_c_kernel(double* C, int M, int N, long long ldc, double beta) {
  const long long total = (long long)M * N;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long j = idx / M, i = idx - j * M;
    double* p = C + j * ldc + i;
    *p = (beta == 0.0) ? 0.0 : *p * beta;
  }
}

// C (M x N, ldc) += beta * X (M x N, ldx): the deferred beta*C term of the host pipeline's start-up panels
// (dgemm_hostplan.cpp).  HBM-bound, 24 bytes per element.
__global__ void __launch_bounds__(256)
add_scaled_kernel(double* __restrict__ C, long long ldc, const double* __restrict__ X, long long ldx, int M, int N,
                  double beta) {
  for (int c = blockIdx.x; c < N; c += gridDim.x) {
    double* cc = C + (long long)c * ldc;
    const double* xc = X + (long long)c * ldx;
    for (int r = threadIdx.x; r < M; r += 256) cc[r] = fma(beta, xc[r], cc[r]);
  }
}

// dst (rows x cols, ldd) = src (rows x cols, lds), column-major, 8-byte elements: brings an operand with an odd leading
// dimension or an 8-byte-aligned base into a TMA-describable layout (run_device's realign pre-pass).  HBM-bound.
__global__ void __launch_bounds__(256)
realign_kernel(const double* __restrict__ src, long long lds, double* __restrict__ dst, long long ldd, int rows, int cols) {
  for (int c = blockIdx.x; c < cols; c += gridDim.x) {
    const double* s = src + (long long)c * lds;
    double* d = dst + (long long)c * ldd;
    for (int r = threadIdx.x; r < rows; r += 256) d[r] = s[r];
  }
}

// dst (cols x rows, column-major, ldd) = transpose of src (rows x cols, column-major, lds).  Pre-pass for op(A) = A^T /
// op(B) = B^T (dgemm_b200_device_op): 32 x 32 tiles through padded shared memory, both sides coalesced.  HBM-bound:
// 16 bytes per element against the 2*N (or 2*M) flops the GEMM spends on it.
constexpr int TR_TILE = 32, TR_ROWS = 8;
__global__ void __launch_bounds__(TR_TILE * TR_ROWS)
transpose_kernel(const double* __restrict__ src, long long lds, double* __restrict__ dst, long long ldd, int rows,
                 int cols, int tiles_r) {
  __shared__ double tile[TR_TILE][TR_TILE + 1];
  const int tr = blockIdx.x % tiles_r, tc = blockIdx.x / tiles_r;
  const int r0 = tr * TR_TILE, c0 = tc * TR_TILE;
  const int tx = threadIdx.x % TR_TILE, ty = threadIdx.x / TR_TILE;
#pragma unroll
  for (int j = ty; j < TR_TILE; j += TR_ROWS) {
    const int r = r0 + tx, c = c0 + j;
    if (r < rows && c < cols) tile[j][tx] = src[(long long)c * lds + r];
  }
  __syncthreads();
#pragma unroll
  for (int j = ty; j < TR_TILE; j += TR_ROWS) {
    const int c = c0 + tx, r = r0 + j;   // dst element (c, r) = src (r, c)
    if (r < rows && c < cols) dst[(long long)r * ldd + c] = tile[tx][j];
  }
}

__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__global__ void fill_signed_kernel(double* X, int rows, int cols, long long ld, unsigned long long seed,
                                   unsigned long long row0, unsigned long long col0) {
  const long long total = (long long)rows * cols;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total;
       idx += (long long)gridDim.x * blockDim.x) {
    const long long j = idx / rows, i = idx - j * rows;
    const unsigned long long h = splitmix64(seed * 0xD1342543DE82EF95ull + ((row0 + i) + ((col0 + j) << 32)));
    X[j * ld + i] = (double)(long long)(h >> 11) * (1.0 / 4503599627370496.0) - 1.0;
  }
}

// =========================================================================================================
// host-side launchers
// =========================================================================================================
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static const PFN_encodeTiled fn = [] {   // thread-safe: mydgemm_gpu_v21 launches from one host thread per GPU
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return (PFN_encodeTiled) nullptr;
    return (PFN_encodeTiled)p;
  }();
  return fn;
}

// Tensor-map cache.  A CUtensorMap is a pure function of (base, dims, stride, box) -- it holds no reference to the
// memory's contents or to a device -- so the 128-byte descriptors of recently used operands are kept: the host pipeline
// and the multi-GPU layer call the kernel dozens of times per pass on the same few (pointer, ld) pairs, and a repeated
// small call (the reference harness times three identical calls per size, test.c:44-48) re-encodes nothing.
// DGEMM_B200_NO_TMAP_CACHE=1 (read once) switches it off for A/B timing.
struct TmapKey {
  const void* base;
  long long rows, cols, ld;
  int box_rows, box_cols;
  bool operator==(const TmapKey& o) const {
    return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld && box_rows == o.box_rows && box_cols == o.box_cols;
  }
};
struct TmapCache {
  static constexpr int N = 64;
  TmapKey key[N];
  CUtensorMap map[N];
  unsigned long long stamp[N];
  unsigned long long clock = 0;
  int used = 0;
  std::mutex mu;
  bool get(const TmapKey& k, CUtensorMap* out) {
    std::lock_guard<std::mutex> lk(mu);
    for (int i = 0; i < used; i++)
      if (key[i] == k) { *out = map[i]; stamp[i] = ++clock; return true; }
    return false;
  }
  void put(const TmapKey& k, const CUtensorMap& m) {
    std::lock_guard<std::mutex> lk(mu);
    int slot = used;
    if (used < N) used++;
    else {
      slot = 0;
      for (int i = 1; i < N; i++) if (stamp[i] < stamp[slot]) slot = i;   // least recently used
    }
    key[slot] = k; map[slot] = m; stamp[slot] = ++clock;
  }
};
static TmapCache g_tmaps;
static std::atomic<unsigned long long> g_tmap_hits{0}, g_tmap_misses{0};
static const bool kNoTmapCache = getenv("DGEMM_B200_NO_TMAP_CACHE") != nullptr;
void tmap_cache_stats(unsigned long long* hits, unsigned long long* misses) {
  *hits = g_tmap_hits.load(); *misses = g_tmap_misses.load();
}

// 2-D FP64 tensor map over a column-major matrix: dim0 = rows (contiguous), dim1 = cols, stride = ld*8 bytes.
static bool make_tmap(CUtensorMap* tm, const double* base, long long rows, long long cols, long long ld, int box_rows,
                      int box_cols) {
  PFN_encodeTiled enc = get_encode_fn();
  if (!enc) return false;
  const TmapKey key{base, rows, cols, ld, box_rows, box_cols};
  if (!kNoTmapCache && g_tmaps.get(key, tm)) { g_tmap_hits++; return true; }
  g_tmap_misses++;
  cuuint64_t dims[2] = {(cuuint64_t)rows, (cuuint64_t)cols};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
  cuuint32_t box[2] = {(cuuint32_t)box_rows, (cuuint32_t)box_cols};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, (void*)base, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_SUCCESS && !kNoTmapCache) g_tmaps.put(key, *tm);
  return r == CUDA_SUCCESS;
}

bool tma_eligible(const Params& P) {
  return ((uintptr_t)P.A % 16 == 0) && ((uintptr_t)P.B % 16 == 0) && (P.lda % 2 == 0) && (P.ldb % 2 == 0) &&
         P.lda * 8 < (1ll << 40) && P.ldb * 8 < (1ll << 40);
}
// which operands need parity-split tensor maps (SPLIT_A | SPLIT_B), and whether the split kernel can take the call:
// every parity needs at least one column, and the doubled stride must fit a tensor map
int tma_split_bits(const Params& P) {
  static const int kForce = getenv("DGEMM_B200_FORCE_SPLIT") ? (atoi(getenv("DGEMM_B200_FORCE_SPLIT")) & 3) | 4 : 0;
  if (kForce) return kForce & 3;   // debug (read once): parity maps on aligned operands
  return (((uintptr_t)P.A % 16 != 0 || P.lda % 2 != 0) ? SPLIT_A : 0) |
         (((uintptr_t)P.B % 16 != 0 || P.ldb % 2 != 0) ? SPLIT_B : 0);
}
// how many parity halves of the split operands start at 8 mod 16 and have to be staged with cp.async
int tma_mis_halves(const Params& P) {
  const int bits = tma_split_bits(P);
  int n = 0;
  for (int p = 0; p < 2; p++) {
    if ((bits & SPLIT_A) && (uintptr_t)(P.A + (long long)p * P.lda) % 16 != 0) n++;
    if ((bits & SPLIT_B) && (uintptr_t)(P.B + (long long)p * P.ldb) % 16 != 0) n++;
  }
  return n;
}
bool tma_split_ok(const Params& P) {
  const int bits = tma_split_bits(P);
  if ((uintptr_t)P.A % 8 != 0 || (uintptr_t)P.B % 8 != 0) return false;
  if (P.lda * 16 >= (1ll << 40) || P.ldb * 16 >= (1ll << 40)) return false;
  if ((bits & SPLIT_A) && P.K < 2) return false;
  if ((bits & SPLIT_B) && P.N < 2) return false;
  return true;
}
// Tensor map over the columns of parity p of a column-major operand (rows x cols, ld): stride 2*ld elements.  A parity
// whose columns start at 8 mod 16 cannot have one (*mis = 1, the kernel stages it with cp.async); *tm is then left alone.
static bool make_parity_tmap(CUtensorMap* tm, const double* base, long long rows, long long cols, long long ld, int p,
                             int box_rows, int box_cols, int* mis) {
  const double* addr = base + (long long)p * ld;
  *mis = ((uintptr_t)addr % 16 != 0) ? 1 : 0;
  if (*mis) return true;
  return make_tmap(tm, addr, rows, (cols - p + 1) / 2, 2 * ld, box_rows, box_cols);
}

// counted from every host thread that launches (mydgemm_gpu_v21 drives one thread per GPU)
static std::atomic<int> g_launches{0};
int launch_count() { return g_launches.load(std::memory_order_relaxed); }
void reset_launch_count() { g_launches.store(0, std::memory_order_relaxed); }

#ifndef DG_TMA_STAGES
#define DG_TMA_STAGES (6 / DG_KT_PER_STAGE)   // ring slots of KPS k-tiles each: 192 KB of operands in flight
#endif
constexpr int TMA_STAGES = DG_TMA_STAGES;
constexpr int CPA_STAGES = 4;

template <typename Kern>
static cudaError_t set_smem(Kern k, int bytes) {
  return cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

// how many CTAs of the TMA kernel (either tile size, any split variant: same block size and shared memory) can be
// resident on the current device at once; what a grid whose CTAs wait on each other must not exceed
int tma_coresident_capacity(int tile_small, int num_sms) {
  static std::atomic<int> per_sm[TILE_COUNT][64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || tile_small < 0 || tile_small >= TILE_COUNT) return 0;
  std::atomic<int>& slot = per_sm[tile_small][dev & 63];
  int n = slot.load(std::memory_order_relaxed);
  if (n == 0) {
    cudaError_t e;
    if (tile_small == 1) e = small_tile_occupancy(&n);
    else if (tile_small == 2) e = small_tile_occupancy_x2(&n);
    else if (tile_small == 3) e = small_tile_occupancy_mid(&n);
    else if (tile_small == TILE_TINY) e = small_tile_occupancy_tiny(&n);
    else {
      const int smem = TMA_STAGES * SLOT_BYTES + 2 * TMA_STAGES * 8 + 16 + 1024;
      e = set_smem(dgemm_tma_ws_kernel<TMA_STAGES>, smem);
      if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, dgemm_tma_ws_kernel<TMA_STAGES>, TMA_THREADS, smem);
    }
    if (e != cudaSuccess) { cudaGetLastError(); return 0; }
    slot.store(n > 0 ? n : -1, std::memory_order_relaxed);
  }
  return n > 0 ? n * num_sms : 0;
}

cudaError_t launch_gemm(const Params& P, int variant, int num_sms, cudaStream_t stream, bool cooperative, bool pdl) {
  const int tiles = P.tiles_m * P.tiles_n;
  const int units = tiles * P.splits;
  cudaError_t err = cudaSuccess;
  if (variant == VARIANT_NAIVE || variant == VARIANT_DFMA) {
    g_launches++;
    return launch_ladder(P, variant, stream);
  }
  if (variant == VARIANT_CLUSTER) {
    if (!tma_eligible(P) || tma_split_bits(P) || P.fetch_bytes || P.splits != 1 || P.sk_tiles > 0 || P.tile_small)
      return cudaErrorInvalidValue;
    CUtensorMap tmA, tmB;
    if (!make_tmap(&tmA, P.A, P.M, P.K, P.lda, 16, BK) || !make_tmap(&tmB, P.B, P.K, P.N, P.ldb, BK, BN)) return cudaErrorInvalidValue;
    const void* maps[2] = {&tmA, &tmB};
    g_launches++;
    return launch_tma_cluster(maps, P, num_sms, stream);
  }
  if (variant == VARIANT_TMA) {
    const int split = tma_split_bits(P);
    if (split && !tma_split_ok(P)) return cudaErrorInvalidValue;
    if (P.tile_small && (P.splits != 1 || P.fetch_bytes)) return cudaErrorInvalidValue;
    const int bn = tile_bn(P.tile_small);   // (the A boxes are 16 rows high in every tile configuration)
    Params Q = P;
    Q.a_mis[0] = Q.a_mis[1] = Q.b_mis[0] = Q.b_mis[1] = 0;
    CUtensorMap tmA, tmA1, tmB, tmB1;
    memset(&tmA, 0, sizeof tmA); memset(&tmA1, 0, sizeof tmA1); memset(&tmB, 0, sizeof tmB); memset(&tmB1, 0, sizeof tmB1);
    bool ok = true;
    if (split & SPLIT_A) {
      if (P.fetch_bytes) return cudaErrorInvalidValue;   // warps 10-11 cannot do both jobs (run_device never asks)
      ok = ok && make_parity_tmap(&tmA, P.A, P.M, P.K, P.lda, 0, 16, BK / 2, &Q.a_mis[0]) &&
           make_parity_tmap(&tmA1, P.A, P.M, P.K, P.lda, 1, 16, BK / 2, &Q.a_mis[1]);
    } else {
      ok = ok && make_tmap(&tmA, P.A, P.M, P.K, P.lda, 16, BK);
    }
    if (split & SPLIT_B) {
      if (P.fetch_bytes) return cudaErrorInvalidValue;
      ok = ok && make_parity_tmap(&tmB, P.B, P.K, P.N, P.ldb, 0, BK, bn / 2, &Q.b_mis[0]) &&
           make_parity_tmap(&tmB1, P.B, P.K, P.N, P.ldb, 1, BK, bn / 2, &Q.b_mis[1]);
    } else {
      ok = ok && make_tmap(&tmB, P.B, P.K, P.N, P.ldb, BK, bn);
    }
    if (!ok) return cudaErrorInvalidValue;
    if (P.tile_small) {   // 64 x 64 tiles: dgemm_kernels_small.cu (one CTA per SM), dgemm_kernels_small2.cu (two); 128 x 64: dgemm_kernels_mid.cu
      const int slots = P.tile_small == 2 ? SMALL_X2_CTAS_PER_SM * num_sms : num_sms;
      const int dp_units = P.sk_tiles > 0 ? P.sk_tile0 : tiles;
      const int grid = P.sk_tiles > 0 ? P.sk_grid : (dp_units < slots ? dp_units : slots);
      const void* maps[4] = {&tmA, &tmA1, &tmB, &tmB1};
      g_launches++;
      if (P.tile_small == TILE_MID) return launch_tma_small_mid(maps, split, Q, grid, dp_units, stream, cooperative, pdl);
      if (P.tile_small == TILE_TINY) return launch_tma_small_tiny(maps, split, Q, grid, dp_units, stream, cooperative, pdl);
      return P.tile_small == 2 ? launch_tma_small_x2(maps, split, Q, grid, dp_units, stream, cooperative, pdl)
                               : launch_tma_small(maps, split, Q, grid, dp_units, stream, cooperative, pdl);
    }
    const int smem = TMA_STAGES * SLOT_BYTES + 2 * TMA_STAGES * 8 + 16 + 1024;
    static std::atomic<bool> attr_done[64];  // function attributes are per device
    int dev = 0;
    if ((err = cudaGetDevice(&dev)) != cudaSuccess) return err;
    if (!attr_done[dev & 63]) {
      if ((err = set_smem(dgemm_tma_ws_kernel<TMA_STAGES>, smem)) != cudaSuccess) return err;
      if ((err = set_smem(dgemm_tma_ws_split_kernel<TMA_STAGES, SPLIT_A>, smem)) != cudaSuccess) return err;
      if ((err = set_smem(dgemm_tma_ws_split_kernel<TMA_STAGES, SPLIT_B>, smem)) != cudaSuccess) return err;
      if ((err = set_smem(dgemm_tma_ws_split_kernel<TMA_STAGES, SPLIT_A | SPLIT_B>, smem)) != cudaSuccess) return err;
      attr_done[dev & 63] = true;
    }
    // data-parallel units first (tiles before sk_tile0, times splits), stream-K ranges after
    const int dp_units = (P.sk_tiles > 0 ? P.sk_tile0 : tiles) * P.splits;
    const int grid = P.sk_tiles > 0 ? P.sk_grid : (dp_units < num_sms ? dp_units : num_sms);
    LaunchCfg cfg;
    cfg.grid = grid; cfg.threads = TMA_THREADS; cfg.smem = smem; cfg.stream = stream;
    cfg.cooperative = cooperative; cfg.pdl = pdl;
    switch (split) {
      case SPLIT_NONE:
        err = launch_ex(dgemm_tma_ws_kernel<TMA_STAGES>, cfg, tmA, tmB, Q, dp_units);
        break;
      case SPLIT_A:
        err = launch_ex(dgemm_tma_ws_split_kernel<TMA_STAGES, SPLIT_A>, cfg, tmA, tmA1, tmB, tmB1, Q, dp_units);
        break;
      case SPLIT_B:
        err = launch_ex(dgemm_tma_ws_split_kernel<TMA_STAGES, SPLIT_B>, cfg, tmA, tmA1, tmB, tmB1, Q, dp_units);
        break;
      default:
        err = launch_ex(dgemm_tma_ws_split_kernel<TMA_STAGES, SPLIT_A | SPLIT_B>, cfg, tmA, tmA1, tmB, tmB1, Q, dp_units);
    }
    if (err != cudaSuccess) { cudaGetLastError(); return err; }
  } else {
    const int smem = CPA_STAGES * STAGE_BYTES + 1024;
    static std::atomic<bool> attr_done[64];
    int dev = 0;
    if ((err = cudaGetDevice(&dev)) != cudaSuccess) return err;
    if (!attr_done[dev & 63]) {
      if ((err = set_smem(dgemm_cpasync_kernel<1, CPA_STAGES>, smem)) != cudaSuccess) return err;
      if ((err = set_smem(dgemm_cpasync_kernel<2, CPA_STAGES>, smem)) != cudaSuccess) return err;
      attr_done[dev & 63] = true;
    }
    if (P.fetch_bytes) {  // no spare warps in this kernel: the copy engine does the fetch, stream-ordered before it
      if ((err = cudaMemcpyAsync(P.fetch_dst, P.fetch_src, P.fetch_bytes, cudaMemcpyDefault, stream)) != cudaSuccess)
        return err;
    }
    if (variant == VARIANT_CPASYNC16)
      dgemm_cpasync_kernel<2, CPA_STAGES><<<units, CONSUMER_THREADS, smem, stream>>>(P);
    else
      dgemm_cpasync_kernel<1, CPA_STAGES><<<units, CONSUMER_THREADS, smem, stream>>>(P);
  }
  g_launches++;
  if ((err = cudaGetLastError()) != cudaSuccess) return err;
  if (P.splits > 1) {
    splitk_reduce_kernel<<<tiles, CONSUMER_THREADS, 0, stream>>>(P);
    g_launches++;
    err = cudaGetLastError();
  }
  return err;
}

cudaError_t launch_scale_c(double* C, int M, int N, long long ldc, double beta, int num_sms, cudaStream_t stream) {
  if (M <= 0 || N <= 0 || beta == 1.0) return cudaSuccess;
  scale_c_kernel<<<num_sms * 4, 256, 0, stream>>>(C, M, N, ldc, beta);
  g_launches++;
  return cudaGetLastError();
}

cudaError_t launch_add_scaled(double* C, long long ldc, const double* X, long long ldx, int M, int N, double beta,
                              int num_sms, cudaStream_t stream) {
  if (M <= 0 || N <= 0) return cudaSuccess;
  add_scaled_kernel<<<std::min(N, num_sms * 8), 256, 0, stream>>>(C, ldc, X, ldx, M, N, beta);
  g_launches++;
  return cudaGetLastError();
}

cudaError_t launch_realign(const double* src, long long lds, double* dst, long long ldd, int rows, int cols, int num_sms,
                           cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  realign_kernel<<<std::min(cols, num_sms * 8), 256, 0, stream>>>(src, lds, dst, ldd, rows, cols);
  g_launches++;
  return cudaGetLastError();
}

cudaError_t launch_transpose(const double* src, long long lds, double* dst, long long ldd, int rows, int cols,
                             cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  const long long tiles_r = (rows + TR_TILE - 1) / TR_TILE, tiles_c = (cols + TR_TILE - 1) / TR_TILE;
  if (tiles_r * tiles_c > 0x7fffffffll) return cudaErrorInvalidValue;
  transpose_kernel<<<(unsigned)(tiles_r * tiles_c), TR_TILE * TR_ROWS, 0, stream>>>(src, lds, dst, ldd, rows, cols,
                                                                                     (int)tiles_r);
  g_launches++;
  return cudaGetLastError();
}

// ---- cross-rank sequence words of the multi-GPU layer (dgemm_rank.cu) ---------------------------------------------------
__global__ void set_flag_kernel(unsigned int* flag, unsigned int value) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}
__global__ void wait_flag_kernel(const unsigned int* flag, unsigned int need, unsigned int* fail, unsigned long long timeout_ns) {
  if (threadIdx.x != 0) return;
  if (*reinterpret_cast<volatile unsigned int*>(fail)) return;   // an earlier wait of this pass already gave up: do not wait again
  const unsigned long long t0 = globaltimer_ns();
  for (;;) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
    if ((int)(v - need) >= 0) return;
    __nanosleep(1000);
    if (globaltimer_ns() - t0 > timeout_ns) { atomicExch(fail, 1u); return; }
  }
}
typedef CUresult (*PFN_streamWriteValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
cudaError_t launch_set_flag(unsigned int* flag, unsigned int value, cudaStream_t stream) {
  static const PFN_streamWriteValue32 fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (getenv("DGEMM_B200_NO_STREAM_MEMOPS") ||
        cudaGetDriverEntryPoint("cuStreamWriteValue32", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
      return (PFN_streamWriteValue32) nullptr;
    return (PFN_streamWriteValue32)p;
  }();
  if (fn && fn((CUstream)stream, (CUdeviceptr)(uintptr_t)flag, value, 0) == CUDA_SUCCESS) return cudaSuccess;
  set_flag_kernel<<<1, 1, 0, stream>>>(flag, value);
  return cudaGetLastError();
}
cudaError_t launch_wait_flag(const unsigned int* flag, unsigned int need, unsigned int* fail, unsigned long long timeout_ns,
                             cudaStream_t stream) {
  wait_flag_kernel<<<1, 32, 0, stream>>>(flag, need, fail, timeout_ns);
  return cudaGetLastError();
}

cudaError_t launch_fill_signed(double* X, int rows, int cols, long long ld, unsigned long long seed,
                               unsigned long long row0, unsigned long long col0, int num_sms, cudaStream_t stream) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  fill_signed_kernel<<<num_sms * 8, 256, 0, stream>>>(X, rows, cols, ld, seed, row0, col0);
  return cudaGetLastError();
}

}  // namespace dg
