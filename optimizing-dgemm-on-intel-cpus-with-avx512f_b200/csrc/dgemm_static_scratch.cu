// dgemm_static_scratch.cu — dgemm_b200_device_static_scratch: the split-K DGEMM for callers that must not allocate.
//
// Who this is for.  Every other device entry of the library goes through a per-device context that cudaMalloc's its
// stream-K flags on first use and grows a workspace buffer whenever a split-K / stream-K plan needs a bigger one
// (dgemm_api.cu: ensure_ctx, grow).  Two kinds of caller cannot live with that:
//   * a process whose own memory pool already owns (nearly) all of the HBM — a solver or a framework allocator that
//     reserved the card at start-up.  A cudaMalloc from inside this library then fails with out-of-memory although the
//     caller has plenty of memory, just not memory the driver can still hand out;
//   * a caller in a latency-critical or multi-stream loop.  cudaMalloc and cudaFree take a device-wide lock and cudaFree
//     waits for the whole device to go idle, so one workspace growth in here stalls every other stream of the process.
// For them the split-K partial tiles live in ONE array that is part of the library image: it is placed in device memory
// when the module is loaded on a GPU and stays there until the process exits (DGEMM_B200_STATIC_SCRATCH_BYTES = 16 MiB on
// every GPU that uses the library — the price, and the reason the maximum is kept modest).  The entry calls no cudaMalloc,
// cudaFree, cudaMallocAsync, stream or event creation, and does not touch the library's per-device context.
//
// What runs.  The same kernels as dgemm_b200_device_ex(..., variant, splits >= 2): every (tile, split) unit multiplies its
// K range and stores a 128 x 128 partial tile, then splitk_reduce_kernel adds a tile's partials in ascending split order
// and applies alpha / beta (dgemm_kernels.cu).  TMA-eligible operands (16-byte aligned, even LDA / LDB) take the TMA
// kernel, anything else the 8-byte cp.async kernel, exactly as a forced variant would.  Same bits as that call with the
// same split count.  No counterpart in the reference: its kernels malloc and free their pack buffers on every call
// (include/kernel18.h:647-648, 670).
//
// Sharing rules.  There is one array per GPU and every call on that GPU uses it from offset 0:
//   * calls on one stream are ordered by the stream and are safe back to back;
//   * calls whose kernels can overlap on the GPU (different streams, or different host threads launching onto different
//     streams) RACE on the partial tiles and produce garbage without any error — order them with events;
//   * the array keeps whatever the previous call left in it.  Nothing in it has to start at zero: every unit overwrites
//     its whole 128 KiB slot before the reduce kernel reads it, and only the slots of this call are read.  (There are no
//     flags or counters in it, on purpose.)
// Set-up.  The FIRST call on each device looks up the array's device address (cudaGetSymbolAddress, once per device, with
// that device current) and reads the SM count; it may also trigger the lazy load of the kernels.  Make that first call
// before any stream capture begins on that device.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "../../include/dgemm_b200.h"
#include "dgemm_host.h"
#include "dgemm_params.h"

// shared with dgemm_api.cu: status / message of the last call, and the mutex that serialises the host side of every entry
int dgemm_b200_internal_fail(int status, const char* msg);
int dgemm_b200_internal_ok();
std::recursive_mutex& dgemm_b200_internal_mutex();

namespace {

constexpr int kTileM = 128, kTileN = 128;                                     // tile of the kernels used here (dg::BM x dg::BN)
constexpr size_t kSlotDoubles = (size_t)kTileM * kTileN;                      // one partial tile
constexpr size_t kScratchDoubles = DGEMM_B200_STATIC_SCRATCH_BYTES / sizeof(double);
constexpr int kMaxSlots = (int)(kScratchDoubles / kSlotDoubles);              // 128 partial tiles
static_assert(kScratchDoubles % kSlotDoubles == 0, "the scratch holds whole partial tiles");

// The scratch itself.  In host code this name is only a handle for the runtime's symbol table — NOT a device address:
// handing `g_static_scratch` to a kernel would pass a host pointer.  Host code must go through cudaGetSymbolAddress.
static __device__ __align__(256) double g_static_scratch[kScratchDoubles];

constexpr int kMaxDev = 64;
struct PerDevice {
  double* scratch = nullptr;   // device address of g_static_scratch ON THIS DEVICE (each device has its own copy)
  int num_sms = 0;
  bool ready = false;
};
PerDevice g_dev[kMaxDev];
std::mutex g_setup_mu;

int failf(int status, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
int failf(int status, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  return dgemm_b200_internal_fail(status, buf);
}

// once per device, with that device current
int setup_current_device(PerDevice** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return failf(DGEMM_B200_ERR_NO_DEVICE, "no usable CUDA device: %s (this library has no CPU fallback)", cudaGetErrorString(e));
  }
  if (dev < 0 || dev >= kMaxDev) return failf(DGEMM_B200_ERR_BAD_ARG, "device %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_setup_mu);
  PerDevice& d = g_dev[dev];
  if (!d.ready) {
    int major = 0;
    e = cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    if (e != cudaSuccess || major != 10) {
      cudaGetLastError();
      return failf(DGEMM_B200_ERR_NO_DEVICE, "device %d is not an sm_100 GPU (the library holds sm_100a code only)", dev);
    }
    e = cudaDeviceGetAttribute(&d.num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) {
      cudaGetLastError();
      return failf(DGEMM_B200_ERR_CUDA, "cudaDeviceGetAttribute(SM count): %s", cudaGetErrorString(e));
    }
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return failf(DGEMM_B200_ERR_CUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", dev, cudaGetErrorString(e));
    }
    d.scratch = static_cast<double*>(p);
    d.ready = true;
  }
  *out = &d;
  return DGEMM_B200_OK;
}

}  // namespace

extern "C" int dgemm_b200_device_static_scratch(int M, int N, int K, double alpha, const double* dA, long long lda,
                                                const double* dB, long long ldb, double beta, double* dC, long long ldc,
                                                void* stream, int splits) {
  std::lock_guard<std::recursive_mutex> host_lk(dgemm_b200_internal_mutex());
  if (M < 0 || N < 0 || K < 0) return failf(DGEMM_B200_ERR_BAD_ARG, "negative dimension M=%d N=%d K=%d", M, N, K);
  if (lda < std::max(1, M)) return failf(DGEMM_B200_ERR_BAD_ARG, "LDA=%lld < max(1,M)=%d", lda, std::max(1, M));
  if (ldb < std::max(1, K)) return failf(DGEMM_B200_ERR_BAD_ARG, "LDB=%lld < max(1,K)=%d", ldb, std::max(1, K));
  if (ldc < std::max(1, M)) return failf(DGEMM_B200_ERR_BAD_ARG, "LDC=%lld < max(1,M)=%d", ldc, std::max(1, M));
  if (M > 0 && N > 0 && !dC) return failf(DGEMM_B200_ERR_BAD_ARG, "C is NULL");
  if (M > 0 && N > 0 && K > 0 && alpha != 0.0 && (!dA || !dB)) return failf(DGEMM_B200_ERR_BAD_ARG, "A or B is NULL");
  if (splits < 0 || splits == 1)
    return failf(DGEMM_B200_ERR_BAD_ARG, "splits=%d: pass 0 (choose) or >= 2 (this entry is the split-K path)", splits);
  if (M == 0 || N == 0) return dgemm_b200_internal_ok();

  PerDevice* d = nullptr;
  int st = setup_current_device(&d);
  if (st) return st;
  cudaStream_t s = static_cast<cudaStream_t>(stream);

  if (alpha == 0.0 || K == 0) {   // C = beta*C only (reference kernel18.h:644-645 order); no scratch involved
    cudaError_t e = dg::launch_scale_c(dC, M, N, ldc, beta, d->num_sms, s);
    if (e != cudaSuccess) return failf(DGEMM_B200_ERR_CUDA, "scale_c launch: %s", cudaGetErrorString(e));
    return dgemm_b200_internal_ok();
  }

  const long long tiles_m = (M + kTileM - 1) / kTileM, tiles_n = (N + kTileN - 1) / kTileN;
  const long long tiles = tiles_m * tiles_n;
  const int k_tiles = (K + dg::BK - 1) / dg::BK;
  // the size limit: two partials per output tile is the least a split-K run needs
  if (tiles * 2 > kMaxSlots)
    return failf(DGEMM_B200_ERR_BAD_ARG,
                 "M=%d N=%d is %lld output tiles of %d x %d; the static scratch (%zu bytes) holds %d partial tiles, i.e. at "
                 "most %d output tiles at 2 splits. Use dgemm_b200_device for larger outputs",
                 M, N, tiles, kTileM, kTileN, (size_t)DGEMM_B200_STATIC_SCRATCH_BYTES, kMaxSlots, kMaxSlots / 2);
  if (splits == 0) {
    // enough units for one wave of the SMs, at least 4 k-tiles (64 k) per unit, never more than the scratch holds
    const int fill = (int)((d->num_sms + tiles - 1) / tiles);
    splits = std::max(2, std::min({fill, std::max(1, k_tiles / 4), (int)(kMaxSlots / tiles)}));
  }
  if ((long long)splits * tiles > kMaxSlots)
    return failf(DGEMM_B200_ERR_BAD_ARG,
                 "splits=%d x %lld output tiles = %lld partial tiles; the static scratch (%zu bytes) holds %d", splits, tiles,
                 (long long)splits * tiles, (size_t)DGEMM_B200_STATIC_SCRATCH_BYTES, kMaxSlots);

  dg::Params P;
  memset(&P, 0, sizeof P);
  P.M = M; P.N = N; P.K = K;
  P.alpha = alpha; P.beta = beta;
  P.A = dA; P.lda = lda; P.B = dB; P.ldb = ldb; P.C = dC; P.ldc = ldc;
  P.tiles_m = (int)tiles_m; P.tiles_n = (int)tiles_n;
  P.k_tiles = k_tiles;
  splits = std::max(1, std::min(splits, k_tiles));   // K < 32: one k-tile, nothing to split (no scratch used then)
  P.k_tiles_per_split = (k_tiles + splits - 1) / splits;
  P.splits = (k_tiles + P.k_tiles_per_split - 1) / P.k_tiles_per_split;   // no empty split: every slot read is written
  P.tile_small = dg::TILE_BIG;
  P.group_m = std::min(P.tiles_m, 16);
  P.c_vec2 = ((uintptr_t)dC % 16 == 0) && (ldc % 2 == 0);
  P.sk_tile0 = (int)tiles;           // no stream-K tail, no lockstep, no fused fetch, no timeline (all zero / null)
  P.workspace = d->scratch;          // the DEVICE address from cudaGetSymbolAddress, never the symbol's name
  const int variant = dg::tma_eligible(P) ? dg::VARIANT_TMA : dg::VARIANT_CPASYNC8;
  cudaError_t e = dg::launch_gemm(P, variant, d->num_sms, s, /*cooperative=*/false, /*pdl=*/false);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return failf(DGEMM_B200_ERR_CUDA, "split-K launch on the static scratch: %s", cudaGetErrorString(e));
  }
  return dgemm_b200_internal_ok();
}
