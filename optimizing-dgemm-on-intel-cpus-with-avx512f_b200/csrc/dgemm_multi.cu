// dgemm_multi.cu — single-process multi-GPU layer behind mydgemm_gpu_v21 (the reference signature,
// include/kernel19.h:678-689; kernel19 shards ROW bands over OpenMP threads with B shared, kernel19.h:700 —
// here COLUMN panels of B and C are sharded over GPUs with A shared, because column panels of column-major
// matrices are contiguous).
//
//   GPU g owns columns [n_g, n_g+1) of B and C (multiples of the 128-wide CTA tile).
//   A (M x K) is cut into G K-slabs (contiguous column ranges).  GPU g copies slab g from the host over its OWN
//   PCIe link; the slabs then travel GPU<->GPU over NVLink.  Each GPU starts on its own slab at once:
//   C_g = beta*C_g + alpha*A[:,slab]*B_g[slab,:] for the first slab, C_g += ... for the rest.  No reduction:
//   column panels are independent.
//
// Two exchange modes:
//   fused (default when every GPU pair has peer access): the DGEMM kernel working on slab i pulls slab i+1 from
//       its owner's HBM with its spare warps (dgemm_device.cuh, "fetch warps"), so the replication of A costs no
//       SMs and no extra launches and is hidden behind the math.  A persistent kernel that owns every SM starves
//       NCCL's own kernels (measured: 2 GPUs 57 TF with ncclBroadcast vs ~72 ideal), which is why this exists.
//   nccl  (DGEMM_B200_EXCHANGE=nccl, or no peer access): one ncclBroadcast per slab on a side stream.
// NCCL is resolved with dlopen at first use so the library loads on boxes without NCCL.
#include "../../include/dgemm_b200.h"
#include "dgemm_device.cuh"
#include "dgemm_host.h"

#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

int dgemm_b200_internal_fail(int status, const char* msg);
int dgemm_b200_internal_ok();
const char* dgemm_b200_internal_errmsg();
int dgemm_b200_internal_run_on(int dev, int M, int N, int K, double alpha, const double* dA, long long lda,
                               const double* dB, long long ldb, double beta, double* dC, long long ldc,
                               cudaStream_t stream, const void* fetch_src, void* fetch_dst, size_t fetch_bytes);

void dgemm_b200_multi_shutdown();

namespace {

struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool load() {
    if (handle) return true;
    handle = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!handle) handle = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!handle) return false;
#define SYM(field, name) field = (decltype(field))dlsym(handle, name)
    SYM(CommInitAll, "ncclCommInitAll");
    SYM(CommDestroy, "ncclCommDestroy");
    SYM(GroupStart, "ncclGroupStart");
    SYM(GroupEnd, "ncclGroupEnd");
    SYM(Broadcast, "ncclBroadcast");
    SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
    return CommInitAll && CommDestroy && GroupStart && GroupEnd && Broadcast && GetErrorString;
  }
};

struct DevState {
  cudaStream_t s_h2d = nullptr, s_comm = nullptr, s_compute = nullptr, s_d2h = nullptr;
  double *a = nullptr, *b = nullptr, *c = nullptr;   // simple path: whole A, the B and C panels
  size_t a_cap = 0, b_cap = 0, c_cap = 0;
  void* xbuf = nullptr;                              // rank pipeline: exchange buffer (A + sequence words)
  size_t x_cap_bytes = 0;
  int x_M = 0, x_K = 0;
  cudaEvent_t ev_slab = nullptr, ev_bc = nullptr, ev_done = nullptr;
  std::vector<cudaEvent_t> ev_bcast;  // per slab
};

struct MultiCtx {
  int g = 0;
  bool fused = false;  // peer access everywhere -> in-kernel NVLink fetch instead of NCCL broadcasts
  bool want_fused = true;  // what DGEMM_B200_EXCHANGE asked for when the context was built
  NcclApi nccl;
  std::vector<ncclComm_t> comms;
  std::vector<DevState> dev;
} g_multi;

#define MCU(expr)                                                                                       \
  do {                                                                                                  \
    cudaError_t e_ = (expr);                                                                            \
    if (e_ != cudaSuccess) {                                                                            \
      char buf[384];                                                                                    \
      snprintf(buf, sizeof(buf), "%s: %s", #expr, cudaGetErrorString(e_));                              \
      return dgemm_b200_internal_fail(e_ == cudaErrorMemoryAllocation ? DGEMM_B200_ERR_OOM : DGEMM_B200_ERR_CUDA, buf); \
    }                                                                                                   \
  } while (0)
#define MNCCL(expr)                                                                \
  do {                                                                             \
    ncclResult_t r_ = (expr);                                                      \
    if (r_ != ncclSuccess) {                                                       \
      char buf[384];                                                               \
      snprintf(buf, sizeof(buf), "%s: %s", #expr, g_multi.nccl.GetErrorString(r_)); \
      return dgemm_b200_internal_fail(DGEMM_B200_ERR_NCCL, buf);                   \
    }                                                                              \
  } while (0)

int ensure_multi(int g) {
  MultiCtx& m = g_multi;
  // peer access for every ordered pair -> fused exchange
  const char* mode = getenv("DGEMM_B200_EXCHANGE");
  bool want_fused = !(mode && strcmp(mode, "nccl") == 0);
  if (m.g == g && m.want_fused == want_fused) return DGEMM_B200_OK;   // (a later change of DGEMM_B200_EXCHANGE rebuilds it)
  if (m.g != 0) {
    dgemm_b200_multi_shutdown();
  }
  m.want_fused = want_fused;
  if (want_fused) {
    for (int a = 0; a < g && want_fused; a++)
      for (int b = 0; b < g && want_fused; b++) {
        if (a == b) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, a, b) != cudaSuccess || !can) want_fused = false;
      }
    for (int a = 0; a < g && want_fused; a++) {
      MCU(cudaSetDevice(a));
      for (int b = 0; b < g; b++) {
        if (a == b) continue;
        cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
        else if (e != cudaSuccess) { cudaGetLastError(); want_fused = false; break; }
      }
    }
  }
  m.fused = want_fused;
  m.comms.assign(g, nullptr);
  if (!m.fused) {
    if (!m.nccl.load()) return dgemm_b200_internal_fail(DGEMM_B200_ERR_NCCL, "libnccl.so.2 not found (dlopen)");
    std::vector<int> devs(g);
    for (int i = 0; i < g; i++) devs[i] = i;
    MNCCL(m.nccl.CommInitAll(m.comms.data(), g, devs.data()));
  }
  m.dev.assign(g, DevState());
  for (int d = 0; d < g; d++) {
    MCU(cudaSetDevice(d));
    DevState& s = m.dev[d];
    MCU(cudaStreamCreateWithFlags(&s.s_h2d, cudaStreamNonBlocking));
    MCU(cudaStreamCreateWithFlags(&s.s_comm, cudaStreamNonBlocking));
    MCU(cudaStreamCreateWithFlags(&s.s_compute, cudaStreamNonBlocking));
    MCU(cudaStreamCreateWithFlags(&s.s_d2h, cudaStreamNonBlocking));
    MCU(cudaEventCreateWithFlags(&s.ev_slab, cudaEventDisableTiming));
    MCU(cudaEventCreateWithFlags(&s.ev_bc, cudaEventDisableTiming));
    MCU(cudaEventCreateWithFlags(&s.ev_done, cudaEventDisableTiming));
    s.ev_bcast.resize(g);
    for (int k = 0; k < g; k++) MCU(cudaEventCreateWithFlags(&s.ev_bcast[k], cudaEventDisableTiming));
  }
  m.g = g;
  return DGEMM_B200_OK;
}

int grow_dev(double** p, size_t* cap, size_t bytes) {
  if (*cap >= bytes) return DGEMM_B200_OK;
  if (*p) {
    MCU(cudaDeviceSynchronize());
    cudaFree(*p);
    *p = nullptr;
    *cap = 0;
  }
  MCU(cudaMalloc(p, bytes));
  *cap = bytes;
  return DGEMM_B200_OK;
}

}  // namespace

void dgemm_b200_multi_shutdown() {
  MultiCtx& m = g_multi;
  if (m.g == 0) return;
  for (int d = 0; d < m.g; d++) {
    cudaSetDevice(d);
    cudaDeviceSynchronize();
    DevState& s = m.dev[d];
    if (s.a) cudaFree(s.a);
    if (s.b) cudaFree(s.b);
    if (s.c) cudaFree(s.c);
    if (s.xbuf) cudaFree(s.xbuf);
    cudaStreamDestroy(s.s_h2d); cudaStreamDestroy(s.s_comm); cudaStreamDestroy(s.s_compute); cudaStreamDestroy(s.s_d2h);
    cudaEventDestroy(s.ev_slab); cudaEventDestroy(s.ev_bc); cudaEventDestroy(s.ev_done);
    for (auto e : s.ev_bcast) cudaEventDestroy(e);
    if (d < (int)m.comms.size() && m.comms[d]) m.nccl.CommDestroy(m.comms[d]);
  }
  m.dev.clear();
  m.comms.clear();
  m.g = 0;
  cudaSetDevice(0);
}

// column / K partition shared by every rank: boundaries are multiples of `align` except the last
static void partition(int total, int parts, int align, std::vector<int>& begin) {
  begin.assign(parts + 1, total);
  int per = ((total + parts - 1) / parts + align - 1) / align * align;
  for (int i = 0; i <= parts; i++) begin[i] = std::min(total, i * per);
}

// ---- fused mode: every GPU runs the rank pipeline (dgemm_rank.cu) from its own host thread ---------------------------------
// The same executor and schedule as the one-process-per-GPU entry dgemm_b200_rank_pipeline: own K-slab uploaded in growing
// pieces and published piece by piece, foreign slabs pulled over NVLink by the DGEMM kernel's spare warps, slab-major start,
// panel-major finish with deferred C and tapered last sub-panels.  The threads never synchronise with each other on the
// host; joining them is the barrier between two passes that the exchange buffers need.
static int run_host_multi_ranked(int M, int N, int K, double alpha, const double* A, int LDA, const double* B, int LDB,
                                 double beta, double* C, int LDC, int G, const std::vector<int>& nb,
                                 const std::vector<int>& kb) {
  MultiCtx& m = g_multi;
  const size_t xbytes = dg::exchange_bytes(M, K);
  for (int d = 0; d < G; d++) {
    MCU(cudaSetDevice(d));
    DevState& s = m.dev[d];
    if (s.x_cap_bytes < xbytes || s.x_M != M || s.x_K != K) {   // (the flag block sits behind A: its offset depends on M, K)
      if (s.xbuf) {
        MCU(cudaDeviceSynchronize());
        cudaFree(s.xbuf);
        s.xbuf = nullptr;
        s.x_cap_bytes = 0;
      }
      MCU(cudaMalloc(&s.xbuf, xbytes));
      s.x_cap_bytes = xbytes; s.x_M = M; s.x_K = K;
    }
    MCU(cudaMemset(static_cast<char*>(s.xbuf) + dg::exchange_flag_offset(M, K), 0, 256));
    // The memset runs on the legacy default stream; the rank threads publish and wait on cudaStreamNonBlocking streams (and
    // the peers' waits run on other devices), which nothing orders behind it.  The buffer is kept across calls, so until the
    // zeroes land the words still hold the LAST call's final values and would satisfy every wait of this one.  Finish it here,
    // before any rank thread exists.
    MCU(cudaStreamSynchronize(0));
  }
  std::vector<void*> xbuf(G);
  for (int d = 0; d < G; d++) xbuf[d] = m.dev[d].xbuf;
  const unsigned int pass = 0;   // the sequence words were just zeroed: every call is pass 0 on fresh flags
  std::vector<int> status(G, 0);
  std::vector<std::string> message(G);
  std::vector<std::thread> threads;
  for (int d = 0; d < G; d++) {
    threads.emplace_back([&, d]() {
      dg::RankArgs a;
      a.M = M; a.ncols = nb[d + 1] - nb[d]; a.K = K; a.alpha = alpha; a.beta = beta;
      a.A_slab = A + (size_t)kb[d] * LDA; a.LDA = LDA;
      a.B = B + (size_t)nb[d] * LDB; a.LDB = LDB;
      a.C = C + (size_t)nb[d] * LDC; a.LDC = LDC;
      a.rank = d; a.world = G; a.kb = kb.data(); a.xbuf = xbuf.data(); a.pass = pass;
      status[d] = dg::rank_pipeline_exec(d, a);
      if (status[d]) message[d] = dgemm_b200_internal_errmsg();   // status and message are per thread: hand them to the parent
    });
  }
  for (auto& t : threads) t.join();
  cudaSetDevice(0);
  for (int d = 0; d < G; d++)
    if (status[d]) return dgemm_b200_internal_fail(status[d], message[d].c_str());
  return dgemm_b200_internal_ok();
}

int dgemm_b200_run_host_multi(int M, int N, int K, double alpha, const double* A, int LDA, const double* B, int LDB,
                              double beta, double* C, int LDC, int G) {
  if (M == 0 || N == 0) return dgemm_b200_internal_ok();
  int st = ensure_multi(G);
  if (st) return st;
  MultiCtx& m = g_multi;
  const bool need_ab = !(alpha == 0.0 || K == 0);
  const bool need_c_in = beta != 0.0;
  const long long ldaD = (M + 1) & ~1ll, ldbD = (K + 1) & ~1ll, ldcD = (M + 1) & ~1ll;
  std::vector<int> nb, kb;
  partition(N, G, dg::BN, nb);
  partition(K, G, dg::BK, kb);
  // fused exchange and every GPU has a K-slab of its own: the rank pipeline (dgemm_rank.cu); a GPU without columns still
  // uploads and publishes its slab
  if (m.fused && need_ab && kb[G] - kb[G - 1] > 0 && !getenv("DGEMM_B200_MULTI_SIMPLE"))
    return run_host_multi_ranked(M, N, K, alpha, A, LDA, B, LDB, beta, C, LDC, G, nb, kb);

  for (int d = 0; d < G; d++) {
    MCU(cudaSetDevice(d));
    DevState& s = m.dev[d];
    const int ncols = nb[d + 1] - nb[d];
    if (need_ab) {
      if ((st = grow_dev(&s.a, &s.a_cap, (size_t)ldaD * K * 8))) return st;
      if ((st = grow_dev(&s.b, &s.b_cap, (size_t)ldbD * std::max(ncols, 1) * 8))) return st;
    }
    if ((st = grow_dev(&s.c, &s.c_cap, (size_t)ldcD * std::max(ncols, 1) * 8))) return st;
  }

  // phase 1: every GPU pulls its A slab, B panel and C panel over its own PCIe link (one host thread per GPU,
  // so pageable sources do not serialise the devices); slab-ready events feed the NVLink exchange
  std::vector<std::promise<int>> slab_recorded(G);
  std::vector<std::future<int>> slab_fut;
  for (auto& p : slab_recorded) slab_fut.push_back(p.get_future());
  std::vector<int> thread_status(G, 0);
  std::vector<std::thread> threads;
  for (int d = 0; d < G; d++) {
    threads.emplace_back([&, d]() {
      DevState& s = m.dev[d];
      cudaError_t e = cudaSetDevice(d);
      const int k0 = kb[d], kc = kb[d + 1] - kb[d];
      if (e == cudaSuccess && need_ab && kc > 0)
        e = cudaMemcpy2DAsync(s.a + (size_t)k0 * ldaD, ldaD * 8, A + (size_t)k0 * LDA, (size_t)LDA * 8, (size_t)M * 8, kc,
                              cudaMemcpyHostToDevice, s.s_h2d);
      if (e == cudaSuccess) e = cudaEventRecord(s.ev_slab, s.s_h2d);
      slab_recorded[d].set_value(e == cudaSuccess ? 0 : 1);
      const int n0 = nb[d], nc = nb[d + 1] - nb[d];
      if (e == cudaSuccess && need_ab && nc > 0)
        e = cudaMemcpy2DAsync(s.b, ldbD * 8, B + (size_t)n0 * LDB, (size_t)LDB * 8, (size_t)K * 8, nc,
                              cudaMemcpyHostToDevice, s.s_h2d);
      if (e == cudaSuccess && need_c_in && nc > 0)
        e = cudaMemcpy2DAsync(s.c, ldcD * 8, C + (size_t)n0 * LDC, (size_t)LDC * 8, (size_t)M * 8, nc,
                              cudaMemcpyHostToDevice, s.s_h2d);
      if (e == cudaSuccess) e = cudaEventRecord(s.ev_bc, s.s_h2d);
      thread_status[d] = (e == cudaSuccess) ? 0 : (int)e;
    });
  }
  // NVLink exchange: slab g is broadcast from GPU g to everyone (grouped -> one collective per slab)
  int nccl_rc = 0;
  if (need_ab && !m.fused) {
    for (int g = 0; g < G && !nccl_rc; g++) {
      if (slab_fut[g].get() != 0) { nccl_rc = -1; break; }
      const size_t count = (size_t)(kb[g + 1] - kb[g]) * ldaD;
      cudaSetDevice(g);
      cudaStreamWaitEvent(m.dev[g].s_comm, m.dev[g].ev_slab, 0);  // root's slab must have landed
      if (count > 0) {
        ncclResult_t r = m.nccl.GroupStart();
        for (int d = 0; d < G && r == ncclSuccess; d++) {
          double* p = m.dev[d].a + (size_t)kb[g] * ldaD;
          r = m.nccl.Broadcast(p, p, count, ncclDouble, g, m.comms[d], m.dev[d].s_comm);
        }
        ncclResult_t r2 = m.nccl.GroupEnd();
        if (r != ncclSuccess || r2 != ncclSuccess) nccl_rc = (int)(r != ncclSuccess ? r : r2);
      }
      for (int d = 0; d < G; d++) {
        cudaSetDevice(d);
        cudaEventRecord(m.dev[d].ev_bcast[g], m.dev[d].s_comm);
      }
    }
  } else {
    for (int g = 0; g < G; g++) slab_fut[g].get();
  }
  for (auto& t : threads) t.join();
  for (int d = 0; d < G; d++)
    if (thread_status[d])
      return dgemm_b200_internal_fail(DGEMM_B200_ERR_CUDA, cudaGetErrorString((cudaError_t)thread_status[d]));
  if (nccl_rc)
    return dgemm_b200_internal_fail(DGEMM_B200_ERR_NCCL,
                                    nccl_rc > 0 ? m.nccl.GetErrorString((ncclResult_t)nccl_rc) : "slab copy failed");

  // phase 2: per GPU, own slab first, then the others in ring order as their broadcasts complete
  for (int d = 0; d < G; d++) {
    MCU(cudaSetDevice(d));
    DevState& s = m.dev[d];
    const int nc = nb[d + 1] - nb[d];
    if (nc <= 0) continue;
    MCU(cudaStreamWaitEvent(s.s_compute, s.ev_bc, 0));
    if (!need_ab) {
      if ((st = dgemm_b200_internal_run_on(d, M, nc, K, alpha, s.a, ldaD, s.b, ldbD, beta, s.c, ldcD, s.s_compute,
                                           nullptr, nullptr, 0)))
        return st;
    } else {
      // non-empty slabs in ring order starting at the own one
      std::vector<int> order;
      for (int i = 0; i < G; i++)
        if (kb[(d + i) % G + 1] - kb[(d + i) % G] > 0) order.push_back((d + i) % G);
      bool first = true;
      if (m.fused && !order.empty() && order[0] != d) {
        // own slab empty: a K = 0 call applies beta and fetches the first foreign slab (copy engine)
        const int nx = order[0];
        MCU(cudaStreamWaitEvent(s.s_compute, m.dev[nx].ev_slab, 0));
        if ((st = dgemm_b200_internal_run_on(d, M, nc, 0, alpha, s.a, ldaD, s.b, ldbD, beta, s.c, ldcD, s.s_compute,
                                             m.dev[nx].a + (size_t)kb[nx] * ldaD, s.a + (size_t)kb[nx] * ldaD,
                                             (size_t)(kb[nx + 1] - kb[nx]) * ldaD * 8)))
          return st;
        first = false;
      }
      for (size_t i = 0; i < order.size(); i++) {
        const int g = order[i];
        const int kc = kb[g + 1] - kb[g];
        const void* fsrc = nullptr;
        void* fdst = nullptr;
        size_t fbytes = 0;
        if (m.fused) {
          if (i + 1 < order.size()) {  // pull the NEXT slab from its owner while this one is multiplied
            const int nx = order[i + 1];
            MCU(cudaStreamWaitEvent(s.s_compute, m.dev[nx].ev_slab, 0));  // owner's H2D of that slab is complete
            fsrc = m.dev[nx].a + (size_t)kb[nx] * ldaD;
            fdst = s.a + (size_t)kb[nx] * ldaD;
            fbytes = (size_t)(kb[nx + 1] - kb[nx]) * ldaD * 8;
          }
          if (g == d) MCU(cudaStreamWaitEvent(s.s_compute, s.ev_slab, 0));
        } else if (g != d) {
          MCU(cudaStreamWaitEvent(s.s_compute, s.ev_bcast[g], 0));
        }
        if ((st = dgemm_b200_internal_run_on(d, M, nc, kc, alpha, s.a + (size_t)kb[g] * ldaD, ldaD, s.b + kb[g], ldbD,
                                             first ? beta : 1.0, s.c, ldcD, s.s_compute, fsrc, fdst, fbytes)))
          return st;
        first = false;
      }
    }
    MCU(cudaEventRecord(s.ev_done, s.s_compute));
  }
  // phase 3: D2H of every C panel, again one host thread per GPU (a pageable destination blocks its caller)
  threads.clear();
  for (int d = 0; d < G; d++) {
    threads.emplace_back([&, d]() {
      DevState& s = m.dev[d];
      const int nc = nb[d + 1] - nb[d];
      cudaError_t e = cudaSetDevice(d);
      if (e == cudaSuccess && nc > 0) {
        e = cudaStreamWaitEvent(s.s_d2h, s.ev_done, 0);
        if (e == cudaSuccess)
          e = cudaMemcpy2DAsync(C + (size_t)nb[d] * LDC, (size_t)LDC * 8, s.c, ldcD * 8, (size_t)M * 8, nc,
                                cudaMemcpyDeviceToHost, s.s_d2h);
      }
      for (cudaStream_t q : {s.s_d2h, s.s_compute, s.s_comm, s.s_h2d})
        if (e == cudaSuccess) e = cudaStreamSynchronize(q);
      thread_status[d] = (e == cudaSuccess) ? 0 : (int)e;
    });
  }
  for (auto& t : threads) t.join();
  for (int d = 0; d < G; d++)
    if (thread_status[d])
      return dgemm_b200_internal_fail(DGEMM_B200_ERR_CUDA, cudaGetErrorString((cudaError_t)thread_status[d]));
  cudaSetDevice(0);
  return dgemm_b200_internal_ok();
}
