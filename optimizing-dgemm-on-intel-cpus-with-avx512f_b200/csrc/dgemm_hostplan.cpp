// dgemm_hostplan.cpp — schedule of the host-pointer entry (mydgemm_gpu_v20): how N is cut into column panels and K
// into slabs, and in which order the (slab, panel) device calls and their H2D copies are queued.
//
// The reference's driver (include/kernel18.h:652-669) walks N -> K -> M blocks over operands that are already in
// host memory.  On the GPU the same three loops have to hide a PCIe link: A, B and (beta != 0) C come in over one
// H2D stream at ~55 GB/s while the kernel runs at ~36 TFLOP/s, and C goes back over a D2H stream.  A device call
// (s, p) = A[:, slab s] * B[slab s, panel p] needs A slab s, C panel p and B block (s, p) to have arrived.  With
// a_s slabs and a_p panels resident, one more slab costs M*kw bytes and enables a_p calls, one more panel costs
// M*nw bytes and enables a_s calls, so the schedule is
//   phase A  grow slabs and panels in balance (whichever enables more flops per byte) — the math can only run
//            as fast as the rectangle of resident data grows; small lead-in pieces shorten the first, exposed copy.
//            With beta != 0 a panel would cost its M*nw bytes of C before its first call; instead the panels opened
//            here start with beta = 0 (C is not read) and get beta*C added by an elementwise pass after their LAST
//            call, when the link has slack ("deferred C"): opening a panel then costs only a B block and the math is
//            compute-bound after the first small slab of A;
//   phase B  once the resident panels are wide enough that one more slab is multiplied for longer than it takes to
//            copy (N_resident >= 4 * TFLOPs / H2D_BW columns), bring in the rest of A;
//   phase C  panel by panel, ONE call per panel over the whole K (all of A is resident by now: one epilogue per tile
//            and one launch instead of one per slab), so panels finish one after another and their D2H hides behind
//            the next panel's math; the last panels are tapered so that the copy left exposed after the last kernel
//            is small.
// The piece widths are chosen per call by simulating the three streams for a few hundred candidates with a
// calibrated model (below) and taking the fastest — microseconds of host time per candidate, cached per shape.
#include "dgemm_host.h"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <mutex>

namespace dg {
namespace {

struct Model {
  double h2d = 55e9, d2h = 55e9;  // bytes/s, pinned host memory, PCIe gen5 x16 (measured; tools/pcie_probe.py)
  // pageable memory staged by dgemm_staging.cu: the fill -> DMA pipeline sustains 50 GB/s alone (tools/stage_probe.cu); with
  // the way back running beside it, planning for 45 gave 27.7 TFLOP/s at 8192^3 against 25.5 when planning for 30
  // (profiles/r02_pageable_matrix.txt); DGEMM_B200_PAGEABLE_GBPS overrides
  double h2d_pageable = 45e9, d2h_pageable = 45e9;
  double hbm = 5.0e12;            // bytes/s an elementwise kernel sustains (beta*C fix-up of deferred panels)
  double t_ktile = 2.13e-6;       // s per 128x128x16 k-tile per SM: 4096 clk at the DMMA issue limit, 1.965 GHz
  double t_epi = 2.2e-6;          // s per output tile per call: the fused epilogue reads and writes the C tile once per
                                  // K-slab, so thin slabs pay it more often (timeline stamps)
  double t_kernel = 10e-6;        // s per device call: launch gap, pipeline fill, stream-K fix-up
  double t_copy = 5e-6;           // s per cudaMemcpy2DAsync
  double t_issue = 14e-6;         // s of host time to queue one device call (3 copies, 2 tensor maps, events, launch)
  int sms = 148;
  Model() {
    if (const char* e = getenv("DGEMM_B200_H2D_GBPS")) h2d = atof(e) * 1e9;
    if (const char* e = getenv("DGEMM_B200_D2H_GBPS")) d2h = atof(e) * 1e9;
    if (const char* e = getenv("DGEMM_B200_PAGEABLE_GBPS")) h2d_pageable = d2h_pageable = atof(e) * 1e9;
  }
  double flops_per_s() const { return sms * 2.0 * BM_ * BN_ * BK_ / t_ktile; }
  static constexpr int BM_ = 128, BN_ = 128, BK_ = 16;
  double kernel(int M, int nc, int kc) const {
    const long long tiles = (long long)((M + BM_ - 1) / BM_) * ((nc + BN_ - 1) / BN_);
    const long long kt = (kc + BK_ - 1) / BK_;
    // stream-K: every SM gets an equal share of the (tile, k-tile) iterations and of the epilogues
    return (double)((tiles * kt + sms - 1) / sms) * t_ktile + (double)tiles / sms * t_epi + t_kernel;
  }
};

std::vector<int> make_bounds(int total, const std::vector<int>& lead, int main_w, const std::vector<int>& tail, int align) {
  int tail_sum = 0;
  for (int t : tail) tail_sum += t;
  std::vector<int> b{0};
  const bool use_tail = tail_sum > 0 && tail_sum * 2 <= total;
  const int head_total = use_tail ? (total - tail_sum) / align * align : total;  // the last tail piece takes the remainder
  for (int w : lead) {
    if (b.back() + w > head_total / 2) break;  // lead-in pieces never take more than half
    b.push_back(b.back() + w);
  }
  main_w = std::max(align, main_w / align * align);
  while (b.back() < head_total) {
    int nxt = b.back() + main_w;
    if (head_total - nxt < main_w / 2) nxt = head_total;  // fold a short remainder into the last main piece
    b.push_back(std::min(nxt, head_total));
  }
  if (use_tail) {
    for (int t : tail) b.push_back(b.back() + t);
    b.back() = total;
  }
  return b;
}

// the three-phase order described in the header comment
// defer_c: the panels opened in phase A start with beta = 0 and get beta*C added after their last call, so opening
// them costs only a B block; *deferred receives how many (they are panels [0, *deferred)).
std::vector<HostPlanItem> make_order(int M, const std::vector<int>& nb, const std::vector<int>& kb, bool need_c_in,
                                     double need_cols, bool merge_k, bool defer_c, int* deferred) {
  const int P = (int)nb.size() - 1, S = (int)kb.size() - 1;
  std::vector<HostPlanItem> order;
  order.reserve((size_t)P * S);
  int aS = 0, aP = 0;
  bool phase_b = false;
  *deferred = 0;
  const bool c_costs = need_c_in && !(defer_c && S > 1);
  while (aS < S || aP < P) {
    const double Kw = kb[aS], Nw = nb[aP];
    if (!phase_b && Nw >= need_cols && Kw > 0) phase_b = true;
    bool pick_slab;
    if (aS == S) pick_slab = false;
    else if (aP == P) pick_slab = true;
    else if (phase_b) pick_slab = true;
    else if (Kw == 0) pick_slab = true;
    else if (Nw == 0) pick_slab = false;
    else {
      const double kw = kb[aS + 1] - kb[aS], nw = nb[aP + 1] - nb[aP];
      const double rs = (2.0 * M * kw * Nw) / (8.0 * (M * kw + kw * Nw));                       // flops per byte moved
      const double rp = (2.0 * M * Kw * nw) / (8.0 * ((c_costs ? (double)M * nw : 0.0) + Kw * nw));
      pick_slab = rs > rp;
    }
    if (pick_slab) {
      const int s = aS++;
      for (int p = 0; p < aP; p++) order.push_back({s, p, s + 1});
    } else {
      const int p = aP++;
      if (need_c_in && defer_c && S > 1 && !phase_b && aS < S) *deferred = aP;  // opened during the start-up phase
      if (merge_k && aS == S && S > 1) {
        order.push_back({0, p, S});  // all of A is resident: the whole K in one call (one epilogue per tile, one launch)
      } else {
        for (int s = 0; s < aS; s++) order.push_back({s, p, s + 1});
      }
    }
  }
  return order;
}

// the three streams, exactly as run_host_single queues them: H2D in need-order, kernels in `order`, D2H of a panel
// after its last call
double simulate(const Model& m, int M, const std::vector<int>& nb, const std::vector<int>& kb,
                const std::vector<HostPlanItem>& order, bool need_c_in, int deferred) {
  const int P = (int)nb.size() - 1, S = (int)kb.size() - 1;
  std::vector<char> a_sent(S, 0), c_sent(P, 0);
  std::vector<int> last(P, -1);
  for (int i = 0; i < (int)order.size(); i++) last[order[i].p] = i;
  double th = 0, tc = 0, td = 0;
  for (int i = 0; i < (int)order.size(); i++) {
    const int p = order[i].p;
    const double nc = nb[p + 1] - nb[p], kc = kb[order[i].s_end] - kb[order[i].s];
    th = std::max(th, i * m.t_issue);  // nothing starts before the host has queued it
    for (int s = order[i].s; s < order[i].s_end; s++)
      if (!a_sent[s]) { th += 8.0 * M * (kb[s + 1] - kb[s]) / m.h2d + m.t_copy; a_sent[s] = 1; }
    if (need_c_in && p >= deferred && !c_sent[p]) { th += 8.0 * M * nc / m.h2d + m.t_copy; c_sent[p] = 1; }
    th += 8.0 * kc * nc / m.h2d + m.t_copy;
    tc = std::max(tc, th) + m.kernel(M, (int)nc, (int)kc);
    if (need_c_in && p < deferred && i == last[p]) {  // C arrives behind the last B block; then C += beta * C_in
      th += 8.0 * M * nc / m.h2d + m.t_copy;
      tc = std::max(tc, th) + 24.0 * M * nc / m.hbm + m.t_kernel * 0.5;
    }
    if (i == last[p]) td = std::max(td, tc) + 8.0 * M * nc / m.d2h + m.t_copy;
  }
  return std::max(td, tc);
}

struct CacheEntry {
  int M, N, K;
  bool need_c_in, pageable;
  HostPlan plan;
};
std::vector<CacheEntry> g_cache;
std::mutex g_cache_mu;

}  // namespace

HostPlan plan_host_pipeline(int M, int N, int K, bool need_ab, bool need_c_in, int num_sms, bool pageable) {
  HostPlan best;
  if (!need_ab || M <= 0 || N <= 0 || K <= 0) {  // C = beta*C only: one piece
    best.nb = {0, std::max(N, 0)};
    best.kb = {0, std::max(K, 0)};
    best.order = {{0, 0, 1}};
    best.predicted_ms = 0.0;
    return best;
  }
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    for (const CacheEntry& e : g_cache)
      if (e.M == M && e.N == N && e.K == K && e.need_c_in == need_c_in && e.pageable == pageable) return e.plan;
  }
  Model m;
  if (num_sms > 0) m.sms = num_sms;
  if (pageable) {  // staged through pinned slots by host threads (dgemm_staging.cu): bounded by the host's memcpy rate
    m.h2d = std::min(m.h2d, m.h2d_pageable);
    m.d2h = std::min(m.d2h, m.d2h_pageable);
  }
  const double need_cols = 1.3 * 4.0 * m.flops_per_s() / m.h2d;
  // whole-K calls in phase C; env switches for A/B measurements
  const bool merge_k = getenv("DGEMM_B200_PLAN_MERGE_K") ? atoi(getenv("DGEMM_B200_PLAN_MERGE_K")) != 0 : true;
  const bool defer_ok = need_c_in && (getenv("DGEMM_B200_PLAN_DEFER_C") ? atoi(getenv("DGEMM_B200_PLAN_DEFER_C")) != 0 : true);
  best.predicted_ms = 1e300;
  // lead-in pieces: none, 2 / 4 / 6 equal ones, or geometric growth towards the main width
  static const std::vector<int> kLeads[] = {{}, {256, 256}, {256, 256, 256, 256}, {256, 256, 256, 256, 256, 256},
                                            {512, 512}, {512, 512, 512, 512}, {512, 512, 512, 512, 512, 512},
                                            {1024, 1024}, {1024, 1024, 1024, 1024}, {1024, 1024, 1024, 1024, 1024, 1024},
                                            {256, 256, 512, 1024}, {256, 256, 512, 512, 1024, 1024},
                                            {512, 512, 1024, 1024}, {512, 512, 1024, 1024, 2048},
                                            {512, 512, 512, 1024, 1024, 1024}, {512, 1024, 1024, 1024, 1024}};
  static const int kMainW[] = {8192, 4096, 2048, 1024, 512, 256};  // coarse first: ties keep fewer pieces
  static const std::vector<int> kTails[] = {{}, {256, 128}, {512, 256, 128}};
  const int max_items = 448;
  for (const std::vector<int>& lead : kLeads) {
      const int lw = lead.empty() ? 0 : lead.back();
      for (int wn : kMainW) {
        if (wn < lw || (wn >= 2 * N && wn > 256)) continue;
        for (const std::vector<int>& tail : kTails) {
          const std::vector<int> nb = make_bounds(N, lead, wn, tail, Model::BN_);
          for (int wk : kMainW) {
            if (wk < lw || (wk >= 2 * K && wk > 256)) continue;
            const std::vector<int> kb = make_bounds(K, lead, wk, {}, Model::BK_);
            if ((nb.size() - 1) * (kb.size() - 1) > (size_t)max_items) continue;
            for (int dc = 0; dc <= (defer_ok ? 1 : 0); dc++) {
              int deferred = 0;
              std::vector<HostPlanItem> order = make_order(M, nb, kb, need_c_in, need_cols, merge_k, dc != 0, &deferred);
              if (dc && deferred == 0) continue;
              const double t = simulate(m, M, nb, kb, order, need_c_in, deferred);
              if (t < best.predicted_ms * 0.999) {  // ties: keep the earlier (coarser lead-in, fewer pieces) candidate
                best.predicted_ms = t;
                best.nb = nb;
                best.kb = kb;
                best.order.swap(order);
                best.deferred_panels = deferred;
              }
            }
          }
        }
      }
    }
  if (best.order.empty()) {
    // No candidate fits the item budget (N or K in the millions with the other dimensions thin: even 8192-wide pieces
    // give more than max_items panels x slabs).  A schedule that always fits: pieces widened until it does, uniform, no
    // lead-in, panels x slabs in the same three-phase order.
    int wn = 8192, wk = 8192;
    auto pieces = [](int total, int w) { return (total + w - 1) / w; };
    while ((long long)pieces(N, wn) * pieces(K, wk) > max_items) {
      if (pieces(N, wn) >= pieces(K, wk)) wn *= 2; else wk *= 2;
    }
    const std::vector<int> nb = make_bounds(N, {}, wn, {}, Model::BN_), kb = make_bounds(K, {}, wk, {}, Model::BK_);
    int deferred = 0;
    best.order = make_order(M, nb, kb, need_c_in, need_cols, merge_k, defer_ok, &deferred);
    best.predicted_ms = simulate(m, M, nb, kb, best.order, need_c_in, deferred);
    best.nb = nb;
    best.kb = kb;
    best.deferred_panels = deferred;
  }
  best.predicted_ms *= 1e3;
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    if (g_cache.size() >= 32) g_cache.erase(g_cache.begin());
    g_cache.push_back({M, N, K, need_c_in, pageable, best});
  }
  return best;
}

}  // namespace dg
