// dgemm_rank.cu — executor of one rank's pass of the column-panel sharded DGEMM (schedule: dgemm_rankplan.cpp).
//
// ONE implementation serves both multi-GPU front ends:
//   * dgemm_b200_rank_pipeline   one process per GPU (torchrun): the caller passes its host pieces and the exchange buffers
//                                of all ranks (its own, and its peers' opened through CUDA IPC);
//   * mydgemm_gpu_v21            one process, one host thread per GPU (dgemm_multi.cu) calling the same function with the
//                                peers' buffers reached through peer access.
// It is the multi-GPU counterpart of mydgemm_cpu_v19's parallel region (reference include/kernel19.h:695-745): there the
// threads meet at three OpenMP barriers per (K, N) block around the shared packed B; here ranks never meet at all.  The only
// cross-rank dependency is "piece q of rank g's K-slab is in g's HBM", published by g as a sequence word behind its H2D
// copies and awaited by a peer right before the call that pulls that piece over NVLink.
//
// Exchange buffer (dgemm_b200_exchange_alloc): [ A, M x K column-major, ld = M rounded up to even | 256-byte aligned flag
// block ].  flag word 0 = "landed": pass * 1024 + (upload pieces of the own slab that have landed in this pass).
#include "../../include/dgemm_b200.h"
#include "dgemm_device.cuh"
#include "dgemm_host.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <vector>

int dgemm_b200_internal_fail(int status, const char* msg);
int dgemm_b200_internal_ok();
int dgemm_b200_internal_run_on(int dev, int M, int N, int K, double alpha, const double* dA, long long lda,
                               const double* dB, long long ldb, double beta, double* dC, long long ldc,
                               cudaStream_t stream, const void* fetch_src, void* fetch_dst, size_t fetch_bytes);
int dgemm_b200_internal_rank_ctx(int dev, size_t b_bytes, size_t c_bytes, size_t x_bytes, double** b, double** c, double** x,
                                 cudaStream_t* s_compute, cudaStream_t* s_h2d, cudaStream_t* s_d2h, int* num_sms,
                                 unsigned int** fetch_fail);
double dgemm_b200_internal_flag_timeout_ms();

namespace dg {

size_t exchange_flag_offset(int M, int K) {
  const size_t lda = ((size_t)std::max(M, 1) + 1) & ~(size_t)1;
  return (lda * (size_t)std::max(K, 1) * 8 + 255) & ~(size_t)255;
}
size_t exchange_bytes(int M, int K) { return exchange_flag_offset(M, K) + 256; }

namespace {

#define RCU(expr)                                                                                                      \
  do {                                                                                                                 \
    cudaError_t e_ = (expr);                                                                                           \
    if (e_ != cudaSuccess) {                                                                                           \
      char buf[384];                                                                                                   \
      snprintf(buf, sizeof(buf), "rank pipeline: %s: %s", #expr, cudaGetErrorString(e_));                              \
      return dgemm_b200_internal_fail(e_ == cudaErrorMemoryAllocation ? DGEMM_B200_ERR_OOM : DGEMM_B200_ERR_CUDA, buf); \
    }                                                                                                                  \
  } while (0)

struct EventPool {   // every event of a pass, destroyed together
  std::vector<cudaEvent_t> all;
  bool timing = false;
  cudaError_t make(cudaEvent_t* e) {
    cudaError_t r = cudaEventCreateWithFlags(e, timing ? cudaEventDefault : cudaEventDisableTiming);
    if (r == cudaSuccess) all.push_back(*e);
    return r;
  }
  ~EventPool() { for (cudaEvent_t e : all) cudaEventDestroy(e); }
};

}  // namespace

int rank_pipeline_exec(int dev, const RankArgs& a) {
  const int M = a.M, ncols = a.ncols, K = a.K, G = a.world, r = a.rank;
  if (M < 0 || ncols < 0 || K < 0 || G < 1 || r < 0 || r >= G || !a.kb || !a.xbuf)
    return dgemm_b200_internal_fail(DGEMM_B200_ERR_BAD_ARG, "rank pipeline: bad dimensions / rank / world / kb / xbuf");
  if (a.LDA < std::max(1, M) || a.LDB < std::max(1, K) || a.LDC < std::max(1, M))
    return dgemm_b200_internal_fail(DGEMM_B200_ERR_BAD_ARG, "rank pipeline: leading dimension smaller than the row count");
  for (int g = 0; g < G; g++)
    if (!a.xbuf[g]) return dgemm_b200_internal_fail(DGEMM_B200_ERR_BAD_ARG, "rank pipeline: missing exchange buffer");
  if (M == 0) return dgemm_b200_internal_ok();
  const bool need_ab = !(a.alpha == 0.0 || K == 0);
  const bool need_c_in = a.beta != 0.0;
  const long long ldaD = (M + 1) & ~1ll, ldbD = (std::max(K, 1) + 1) & ~1ll, ldcD = (M + 1) & ~1ll;
  int prev = 0;
  cudaGetDevice(&prev);
  RCU(cudaSetDevice(dev));
  struct Restore { int d; ~Restore() { cudaSetDevice(d); } } restore{prev};

  double *dB = nullptr, *dC = nullptr, *dX = nullptr;
  cudaStream_t s_compute = nullptr, s_h2d = nullptr, s_d2h = nullptr;
  int num_sms = 148;
  unsigned int* fetch_fail = nullptr;
  const size_t nc1 = (size_t)std::max(ncols, 1);
  int st = dgemm_b200_internal_rank_ctx(dev, need_ab ? (size_t)ldbD * nc1 * 8 : 0, (size_t)ldcD * nc1 * 8,
                                        need_c_in && need_ab ? (size_t)ldcD * nc1 * 8 : 0, &dB, &dC, &dX, &s_compute, &s_h2d,
                                        &s_d2h, &num_sms, &fetch_fail);
  if (st) return st;
  double* dA = static_cast<double*>(a.xbuf[r]);
  unsigned int* my_flag = reinterpret_cast<unsigned int*>(static_cast<char*>(a.xbuf[r]) + exchange_flag_offset(M, K));
  const unsigned int seq_base = a.pass * 1024u;
  const unsigned long long wait_ns = (unsigned long long)(dgemm_b200_internal_flag_timeout_ms() * 1e6);

  EventPool ev;
  ev.timing = a.timeline != nullptr && a.timeline_cap >= 3;
  int tl_n = 0;
  struct Mark { int kind, idx; cudaEvent_t e; };
  std::vector<Mark> marks;
  cudaEvent_t ev0 = nullptr;
  auto mark = [&](int kind, int idx, cudaStream_t s) -> cudaError_t {
    if (!ev.timing) return cudaSuccess;
    cudaEvent_t e;
    cudaError_t rc = ev.make(&e);
    if (rc != cudaSuccess) return rc;
    marks.push_back({kind, idx, e});
    return cudaEventRecord(e, s);
  };
  if (ev.timing) {
    RCU(ev.make(&ev0));
    RCU(cudaEventRecord(ev0, s_h2d));
  }
  auto h2d = [&](double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols) {
    return cudaMemcpy2DAsync(dst, (size_t)ldd * 8, src, (size_t)lds * 8, (size_t)rows * 8, (size_t)cols, cudaMemcpyHostToDevice,
                             s_h2d);
  };
  auto d2h_panel = [&](int c0, int nc, int idx) -> int {
    cudaEvent_t done;
    RCU(ev.make(&done));
    RCU(cudaEventRecord(done, s_compute));
    RCU(cudaStreamWaitEvent(s_d2h, done, 0));
    RCU(cudaMemcpy2DAsync(a.C + (size_t)c0 * a.LDC, (size_t)a.LDC * 8, dC + (size_t)c0 * ldcD, (size_t)ldcD * 8, (size_t)M * 8,
                          (size_t)nc, cudaMemcpyDeviceToHost, s_d2h));
    RCU(mark(RANK_TL_D2H_DONE, idx, s_d2h));
    return DGEMM_B200_OK;
  };
  auto finish = [&]() -> int {
    for (cudaStream_t q : {s_d2h, s_compute, s_h2d}) RCU(cudaStreamSynchronize(q));
    unsigned int failed = 0;
    RCU(cudaMemcpy(&failed, fetch_fail, sizeof(failed), cudaMemcpyDeviceToHost));
    if (failed) {
      cudaMemset(fetch_fail, 0, sizeof(failed));
      return dgemm_b200_internal_fail(DGEMM_B200_ERR_CUDA, "rank pipeline: a peer's K-slab of A never arrived (its owner "
                                                           "did not upload it within the time limit)");
    }
    if (ev.timing) {
      for (const Mark& mk : marks) {
        if (tl_n + 3 > a.timeline_cap) break;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ev0, mk.e) != cudaSuccess) { cudaGetLastError(); continue; }
        a.timeline[tl_n++] = mk.kind; a.timeline[tl_n++] = mk.idx; a.timeline[tl_n++] = ms;
      }
    }
    if (a.timeline_n) *a.timeline_n = tl_n;
    return dgemm_b200_internal_ok();
  };

  auto body = [&]() -> int {
  if (!need_ab) {
    // C = beta * C only (kernel18 order, reference kernel18.h:644-645); this rank still publishes "slab landed" so that a
    // peer with a different view of alpha (there is none: alpha is the same on every rank) could not hang -- nothing is pulled
    if (ncols > 0) {
      if (need_c_in) RCU(h2d(dC, ldcD, a.C, a.LDC, M, ncols));
      cudaEvent_t in;
      RCU(ev.make(&in));
      RCU(cudaEventRecord(in, s_h2d));
      RCU(cudaStreamWaitEvent(s_compute, in, 0));
      if ((st = dgemm_b200_internal_run_on(dev, M, ncols, 0, a.alpha, dA, ldaD, dB, ldbD, a.beta, dC, ldcD, s_compute, nullptr,
                                           nullptr, 0)))
        return st;
      if ((st = d2h_panel(0, ncols, 0))) return st;
    }
    return finish();
  }

  const RankPlan plan = plan_rank_pipeline(M, ncols, K, G, r, a.kb, need_c_in, a.p1);
  if (!plan.ok)
    return dgemm_b200_internal_fail(DGEMM_B200_ERR_BAD_ARG,
                                    "rank pipeline: every rank needs a non-empty K-slab of its own (K too small for this many "
                                    "ranks) and kb must be ascending from 0 to K");
  if (plan.pieces.size() > 1000)
    return dgemm_b200_internal_fail(DGEMM_B200_ERR_BAD_ARG, "rank pipeline: K-slab too wide for the sequence word");

  // ---- queue copies and calls in the order the calls need them ----------------------------------------------------------
  size_t next_copy = 0;
  std::vector<cudaEvent_t> copy_ev(plan.copies.size(), nullptr);
  auto issue_copies_through = [&](int last, bool want_event) -> int {
    for (; (int)next_copy <= last; next_copy++) {
      const RankCopy& cp = plan.copies[next_copy];
      if (cp.kind == RANK_COPY_A) {
        RCU(h2d(dA + (size_t)cp.k0 * ldaD, ldaD, a.A_slab + (size_t)(cp.k0 - a.kb[r]) * a.LDA, a.LDA, M, cp.kc));
        RCU(launch_set_flag(my_flag, seq_base + (unsigned)cp.piece + 1u, s_h2d));   // peers may pull this piece now
      } else if (cp.kind == RANK_COPY_B) {
        RCU(h2d(dB + (size_t)cp.c0 * ldbD + cp.k0, ldbD, a.B + (size_t)cp.c0 * a.LDB + cp.k0, a.LDB, cp.kc, cp.nc));
      } else {
        RCU(h2d(dX + (size_t)cp.c0 * ldcD, ldcD, a.C + (size_t)cp.c0 * a.LDC, a.LDC, M, cp.nc));
      }
    }
    if (want_event && last >= 0 && !copy_ev[last]) {
      RCU(ev.make(&copy_ev[last]));
      RCU(cudaEventRecord(copy_ev[last], s_h2d));
    }
    return DGEMM_B200_OK;
  };

  // EVERY copy of the pass is queued before the first kernel.  The peers' pulls depend on this rank's uploads, and this
  // rank's calls depend on theirs: if the host queued uploads only as its own calls came up, anything that blocks the host
  // for a moment (a pageable source, a buffer that has to grow) while a call waits for a peer would close a cycle across
  // ranks.  With the uploads already in the copy queue no call of any rank can ever wait for a host thread.
  for (size_t i = 0; i < plan.calls.size(); i++) {
    const RankCall& c = plan.calls[i];
    // (a sub-panel's C_in rides behind its last B block)
    if ((st = issue_copies_through(c.wait_copy, true))) return st;
    if (c.wait_x_copy >= 0 && (st = issue_copies_through(c.wait_x_copy, true))) return st;
    RCU(mark(RANK_TL_H2D_DONE, (int)i, s_h2d));
  }
  if ((st = issue_copies_through((int)plan.copies.size() - 1, false))) return st;
  for (size_t i = 0; i < plan.calls.size(); i++) {
    const RankCall& c = plan.calls[i];
    RCU(cudaStreamWaitEvent(s_compute, copy_ev[c.wait_copy], 0));
    const void* fsrc = nullptr;
    void* fdst = nullptr;
    size_t fbytes = 0;
    if (c.f_owner >= 0) {
      const char* peer = static_cast<const char*>(a.xbuf[c.f_owner]);
      const unsigned int* flag = reinterpret_cast<const unsigned int*>(peer + exchange_flag_offset(M, K));
      // the owner may still be uploading: hold the compute stream until its sequence word covers the piece (bounded wait)
      RCU(launch_wait_flag(flag, seq_base + (unsigned)c.f_need_piece, fetch_fail, wait_ns, s_compute));
      fsrc = reinterpret_cast<const double*>(peer) + (size_t)c.f_k0 * ldaD;
      fdst = dA + (size_t)c.f_k0 * ldaD;
      fbytes = (size_t)c.f_kc * ldaD * 8;
    }
    RCU(mark(RANK_TL_CALL_START, (int)i, s_compute));
    const double b_eff = c.first ? (need_c_in ? 0.0 : a.beta) : 1.0;
    if ((st = dgemm_b200_internal_run_on(dev, M, c.nc, c.kc, a.alpha, dA + (size_t)c.k0 * ldaD, ldaD,
                                         dB + (size_t)c.c0 * ldbD + c.k0, ldbD, b_eff, dC + (size_t)c.c0 * ldcD, ldcD, s_compute,
                                         fsrc, fdst, fbytes)))
      return st;
    RCU(mark(RANK_TL_CALL_END, (int)i, s_compute));
    if (c.last) {
      if (c.wait_x_copy >= 0) {
        RCU(cudaStreamWaitEvent(s_compute, copy_ev[c.wait_x_copy], 0));
        RCU(launch_add_scaled(dC + (size_t)c.c0 * ldcD, ldcD, dX + (size_t)c.c0 * ldcD, ldcD, M, c.nc, a.beta, num_sms,
                              s_compute));
      }
      if ((st = d2h_panel(c.c0, c.nc, c.sub))) return st;
    }
  }
  if ((st = issue_copies_through((int)plan.copies.size() - 1, false))) return st;   // (ncols == 0: only the own slab)
  return finish();
  };
  const int result = body();
  if (result) {   // nothing of this pass may still be reading or writing the caller's memory when the error is returned
    for (cudaStream_t q : {s_d2h, s_compute, s_h2d}) cudaStreamSynchronize(q);
    cudaGetLastError();
  }
  return result;
}

}  // namespace dg
