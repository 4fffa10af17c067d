// dgemm_rankplan.cpp — schedule of ONE rank's pass of the column-panel sharded DGEMM (mydgemm_gpu_v21 and the
// one-process-per-GPU entry dgemm_b200_rank_pipeline).  Pure host logic, no CUDA: tests drive it without a GPU.
//
// The reference's parallel kernel gives every OpenMP thread a row band of C and lets all threads share one packed copy of
// B (include/kernel19.h:636-656 partition_m_dim, :700-745).  Across GPUs the roles are swapped (column panels, A shared):
// rank r owns the column panel J_r of B and C and the K-slab K_r of A; its pass computes
//     C[:, J_r] = beta * C[:, J_r] + alpha * sum_g A[:, K_g] * B[K_g, J_r]
// from operands that start in HOST memory.  What has to be hidden behind ~36 TFLOP/s of math:
//     PCIe in   the own K-slab of A, the whole B panel, the C panel (beta != 0)          ~55 GB/s
//     NVLink    the K-slabs of the other ranks, pulled from their owners' HBM            (fused into the DGEMM kernel)
//     PCIe out  the C panel
// The schedule (same three ideas as the single-GPU planner, dgemm_hostplan.cpp):
//   * lead-in: the own slab goes up in pieces that start small (first call waits for ~128 MB of A and one B block) and
//     grow geometrically; the owner publishes "pieces landed" in a sequence word its peers can read, so a peer pulls piece q
//     as soon as it is in the owner's HBM -- there is no "every slab is uploaded" barrier anywhere in the pass;
//   * phase 1 (slab-major): the first p1 slabs in ring order are multiplied slab by slab against all column sub-panels.
//     The calls of this phase carry the NVLink pulls of ALL foreign slabs, spread in proportion to their flops (a slab is
//     always complete before the first call that multiplies it: stream order);
//   * phase 2 (panel-major): all of A is resident; every column sub-panel takes ONE call per remaining contiguous K range
//     (ring order wraps at most once), gets beta*C_in added (C is deferred: sub-panels start with beta = 0, C_in comes up
//     behind the B blocks) and goes back to the host while the next sub-panel is multiplied.  The last sub-panels are
//     narrow, so the copy left exposed after the last kernel is small; D2H is spread over the whole phase instead of
//     piling up behind the last slab.
#include "dgemm_host.h"

#include <algorithm>
#include <cstdlib>

namespace dg {

namespace {
int round_up(int x, int a) { return (x + a - 1) / a * a; }
}  // namespace

// Upload pieces of the K-slab [k_begin, k_end) of A (absolute k boundaries).  Every rank evaluates this for every owner:
// piece q of owner g is "landed" when g's sequence word reaches base + q + 1.
std::vector<int> rank_slab_pieces(int M, int k_begin, int k_end) {
  std::vector<int> b{k_begin};
  const int kc = k_end - k_begin;
  if (kc <= 0) return b;
  // first piece ~128 MB of A (2-3 ms of an idle PCIe link), at least 128 columns; widths w0, w0, 2 w0, 2 w0, then 4 w0.
  // The pieces stay this small on purpose.  With 8 GPUs uploading at once the links of this pool's hosts drop to 23-35 GB/s
  // (measured, profiles/r02_e2e_multi_n8.jsonl), which is barely faster than the own slab is multiplied (a 65536 x 8192 panel
  // consumes a column of A in 29 us and receives one in 17-26 us): a piece that is twice as wide as the one before it then
  // arrives AFTER the previous one has been multiplied, and a 2048-wide last piece would leave 60 ms of math behind the last
  // upload.  A 1024-wide call costs ~1 % more than a 2048-wide one (one more epilogue per tile).
  long long w0 = (128ll << 20) / (8ll * std::max(M, 1));
  w0 = std::max(128ll, std::min(2048ll, (w0 + 127) / 128 * 128));
  const int main_w = (int)std::min(4 * w0, std::max(2048ll, w0));
  int w = (int)w0, reps = 2;
  while (b.back() < k_end) {
    int nxt = b.back() + w;
    if (k_end - nxt < w / 2) nxt = k_end;   // fold a short remainder into the last piece
    b.push_back(std::min(nxt, k_end));
    if (--reps == 0) { w = std::min(main_w, 2 * w); reps = 2; }
  }
  return b;
}

// Column sub-panels of a panel of `ncols` columns: equal main pieces (multiples of 128) and a tapered tail.
std::vector<int> rank_sub_panels(int ncols) {
  std::vector<int> sp{0};
  if (ncols <= 0) return sp;
  std::vector<int> tail;
  if (ncols >= 6144) tail = {512, 256, 128};
  else if (ncols >= 2048) tail = {256, 128};
  int tails = 0;
  for (int t : tail) tails += t;
  const int head = ncols - tails;
  const int main_w = getenv("DGEMM_B200_RANK_SUBPANEL") ? std::max(128, atoi(getenv("DGEMM_B200_RANK_SUBPANEL"))) : 2048;
  const int parts = std::max(1, std::min(8, (head + main_w / 2) / main_w));
  const int per = round_up((head + parts - 1) / parts, 128);
  for (int i = 1; i <= parts; i++) {
    const int e = std::min(head, i * per);
    if (e > sp.back()) sp.push_back(e);
  }
  for (size_t i = 0; i < tail.size(); i++) sp.push_back(i + 1 == tail.size() ? ncols : sp.back() + tail[i]);
  return sp;
}

static double call_flops(int M, const RankCall& c) { return 2.0 * M * (double)c.nc * c.kc; }

// Attach the pulls of the foreign slabs `targets` (ring order) to the calls [c_begin, c_end): every slab gets a contiguous
// group of carrier calls (group sizes in proportion to flops), its columns are spread over the group in proportion to
// flops.  One fetch per call, never across a slab boundary.  Returns false when there are fewer carriers than slabs.
static bool spread_pulls(RankPlan& R, int M, const int* kb, const std::vector<int>& targets, int c_begin, int c_end) {
  const int nt = (int)targets.size(), nc = c_end - c_begin;
  if (nt == 0) return true;
  if (nc < nt) return false;
  std::vector<double> cum(nc + 1, 0.0);
  for (int i = 0; i < nc; i++) cum[i + 1] = cum[i] + call_flops(M, R.calls[c_begin + i]);
  double total_cols = 0.0;
  for (int g : targets) total_cols += kb[g + 1] - kb[g];
  // group boundaries in carrier indices: slab t gets carriers [gb[t], gb[t+1])
  std::vector<int> gb(nt + 1, 0);
  double cols_done = 0.0;
  for (int t = 0; t < nt; t++) {
    cols_done += kb[targets[t] + 1] - kb[targets[t]];
    const double want = cum[nc] * cols_done / total_cols;
    int e = gb[t] + 1;
    while (e < nc && cum[e] < want * (1.0 - 1e-9)) e++;       // first index whose cumulative flops reach the share ...
    if (e > gb[t] + 1 && cum[e] - want > want - cum[e - 1]) e--;  // ... or the one before, whichever is closer
    e = std::max(e, gb[t] + 1);
    e = std::min(e, nc - (nt - 1 - t));                        // leave one carrier for every later slab
    gb[t + 1] = (t == nt - 1) ? nc : e;
  }
  for (int t = 0; t < nt; t++) {
    const int g = targets[t], cols = kb[g + 1] - kb[g];
    const double f0 = cum[gb[t]], f1 = cum[gb[t + 1]];
    int pos = 0;
    for (int i = gb[t]; i < gb[t + 1]; i++) {
      int end = (i + 1 == gb[t + 1]) ? cols : (int)((double)cols * (cum[i + 1] - f0) / std::max(f1 - f0, 1.0));
      end = std::min(cols, end / 16 * 16);   // whole multiples of 16 columns (16-byte aligned for any even ld)
      if (i + 1 == gb[t + 1]) end = cols;
      if (end > pos) {
        RankCall& c = R.calls[c_begin + i];
        c.f_owner = g; c.f_k0 = kb[g] + pos; c.f_kc = end - pos;
        pos = end;
      }
    }
  }
  return true;
}

// Shift the pulls of the calls [c_begin, c_end) (all of ONE slab, in column order) one call later: call i carries what call
// i-1 carried, the last call carries the last two shares, the first none.  The owner uploads its slab piece by piece at the
// pace of ITS link; a puller on a faster link would otherwise ask for piece q the moment its own piece q has landed.
static void lag_pulls(RankPlan& R, int c_begin, int c_end) {
  if (c_end - c_begin < 2) return;
  RankCall& last = R.calls[c_end - 1];
  const RankCall& before_last = R.calls[c_end - 2];
  if (before_last.f_owner >= 0) {
    if (last.f_owner < 0) {
      last.f_owner = before_last.f_owner; last.f_k0 = before_last.f_k0; last.f_kc = before_last.f_kc;
    } else {
      last.f_kc += last.f_k0 - before_last.f_k0;   // contiguous: [before_last.f_k0, last.f_k0 + last.f_kc)
      last.f_k0 = before_last.f_k0;
    }
  }
  for (int i = c_end - 2; i > c_begin; i--) {
    R.calls[i].f_owner = R.calls[i - 1].f_owner; R.calls[i].f_k0 = R.calls[i - 1].f_k0; R.calls[i].f_kc = R.calls[i - 1].f_kc;
  }
  R.calls[c_begin].f_owner = -1; R.calls[c_begin].f_k0 = 0; R.calls[c_begin].f_kc = 0;
}

RankPlan plan_rank_pipeline(int M, int ncols, int K, int G, int rank, const int* kb, bool need_c_in, int p1_override) {
  RankPlan R;
  R.ok = false;
  if (M <= 0 || K <= 0 || G < 1 || rank < 0 || rank >= G || ncols < 0) return R;
  if (kb[0] != 0 || kb[G] != K) return R;
  for (int g = 0; g < G; g++)
    if (kb[g + 1] < kb[g]) return R;
  std::vector<int> order;   // non-empty slabs in ring order; the own one must be first (something local to start on)
  for (int i = 0; i < G; i++) {
    const int g = (rank + i) % G;
    if (kb[g + 1] > kb[g]) order.push_back(g);
  }
  if (order.empty() || order[0] != rank) return R;
  const int S = (int)order.size();
  R.order = order;
  R.pieces = rank_slab_pieces(M, kb[rank], kb[rank + 1]);
  R.sp = rank_sub_panels(ncols);
  const int SP = (int)R.sp.size() - 1, Q = (int)R.pieces.size() - 1;

  // how many slabs run slab-major (phase 1).  The rest is panel-major, which needs all of A resident: the pulls are
  // spread over the phase-1 slabs, so a longer phase 1 means a gentler NVLink rate and a shorter window for the D2H.
  //   S = 1: 1;  S = 2: 1;  S >= 3: about half of the pass.  DGEMM_B200_RANK_P1 / p1_override force a value; p1 = S is the
  //   plain ring (slab i+1 pulled while slab i is multiplied, every sub-panel finishes behind the last slab).
  int p1 = S <= 2 ? 1 : std::max(2, S / 2);
  if (const char* e = getenv("DGEMM_B200_RANK_P1")) p1 = atoi(e);
  if (p1_override > 0) p1 = p1_override;
  p1 = std::max(1, std::min(S, p1));

  for (int attempt = 0; attempt < 2; attempt++) {
    R.calls.clear();
    R.copies.clear();
    R.p1 = p1;
    if (ncols == 0) {   // nothing to compute: the rank only uploads its slab (and publishes it) for its peers
      for (int q = 0; q < Q; q++) R.copies.push_back({RANK_COPY_A, R.pieces[q], R.pieces[q + 1] - R.pieces[q], 0, 0, q});
      R.ok = true;
      return R;
    }
    std::vector<char> started(SP, 0);
    // j >= 0: one column sub-panel; j < 0: ALL columns in one call (slab-major calls that finish no sub-panel: the B block
    // of a K piece is small, and one launch over 64 tile columns has less ramp and tail than seven over 15 or fewer)
    auto add_call = [&](int k0, int kc, int j, bool last, int a_piece) {
      RankCall c = {};
      c.k0 = k0; c.kc = kc; c.sub = j;
      c.c0 = j >= 0 ? R.sp[j] : 0;
      c.nc = j >= 0 ? R.sp[j + 1] - R.sp[j] : ncols;
      const int j0 = j >= 0 ? j : 0, j1 = j >= 0 ? j + 1 : SP;
      c.first = started[j0] ? 0 : 1;   // (sub-panels are always started together: phase 1 precedes phase 2)
      c.last = last ? 1 : 0;
      c.f_owner = -1; c.f_k0 = 0; c.f_kc = 0;
      c.wait_x_copy = -1;
      for (int jj = j0; jj < j1; jj++) started[jj] = 1;
      if (a_piece >= 0) {   // own slab: the piece goes up once, ahead of its first B block
        bool sent = false;
        for (const RankCopy& cp : R.copies) sent |= cp.kind == RANK_COPY_A && cp.piece == a_piece;
        if (!sent) R.copies.push_back({RANK_COPY_A, R.pieces[a_piece], R.pieces[a_piece + 1] - R.pieces[a_piece], 0, 0, a_piece});
      }
      R.copies.push_back({RANK_COPY_B, k0, kc, c.c0, c.nc, -1});
      c.wait_copy = (int)R.copies.size() - 1;
      if (last && need_c_in) {
        R.copies.push_back({RANK_COPY_X, 0, 0, c.c0, c.nc, -1});
        c.wait_x_copy = (int)R.copies.size() - 1;
      }
      R.calls.push_back(c);
    };
    // phase 1
    std::vector<int> phase_begin;
    for (int i = 0; i < p1; i++) {
      phase_begin.push_back((int)R.calls.size());
      const int g = order[i];
      const bool final_slab = (i == S - 1);
      if (i == 0) {
        for (int q = 0; q < Q; q++) {
          if (final_slab && q == Q - 1)   // the calls that finish the sub-panels go one by one: staggered D2H
            for (int j = 0; j < SP; j++) add_call(R.pieces[q], R.pieces[q + 1] - R.pieces[q], j, true, q);
          else add_call(R.pieces[q], R.pieces[q + 1] - R.pieces[q], -1, false, q);
        }
      } else if (final_slab) {
        for (int j = 0; j < SP; j++) add_call(kb[g], kb[g + 1] - kb[g], j, true, -1);
      } else {
        // in K pieces of about 2048: every call carries one NVLink pull, and more, equal carriers pace the pulls evenly
        const int kc = kb[g + 1] - kb[g], parts = std::max(1, (kc + 1024) / 2048);
        const int per = round_up((kc + parts - 1) / parts, 16);
        for (int k0 = 0; k0 < kc; k0 += per) add_call(kb[g] + k0, std::min(per, kc - k0), -1, false, -1);
      }
    }
    phase_begin.push_back((int)R.calls.size());
    // phase 2: the remaining slabs as contiguous K ranges (ring order wraps at most once)
    std::vector<std::pair<int, int>> ranges;
    for (int i = p1; i < S; i++) {
      const int g = order[i];
      if (!ranges.empty() && ranges.back().second == kb[g]) ranges.back().second = kb[g + 1];
      else ranges.push_back({kb[g], kb[g + 1]});
    }
    for (int j = 0; j < SP; j++)
      for (size_t r = 0; r < ranges.size(); r++)
        add_call(ranges[r].first, ranges[r].second - ranges[r].first, j, r + 1 == ranges.size(), -1);
    // pulls: slab order[i] (i >= 1) must be complete before phase i (i < p1) or before phase 2 (i >= p1)
    bool ok = true;
    if (S > 1) {
      if (p1 == S) {   // plain ring: one slab per phase
        for (int i = 0; i + 1 < S && ok; i++) ok = spread_pulls(R, M, kb, {order[i + 1]}, phase_begin[i], phase_begin[i + 1]);
        if (ok && Q >= 2 && phase_begin[1] - phase_begin[0] == Q) lag_pulls(R, phase_begin[0], phase_begin[1]);
      } else {
        // The calls of the OWN slab pull the next slab only, column for column at the pace the own slab is multiplied --
        // which is the pace at which its owner uploads it (every rank's link carries the same load): asking for more, sooner,
        // would make the DGEMM wait for another rank's PCIe link.  By the end of the own slab every owner's upload is
        // complete, and the remaining slabs are spread over the other slab-major calls.
        if (p1 >= 2) {
          ok = spread_pulls(R, M, kb, {order[1]}, phase_begin[0], phase_begin[1]);
          std::vector<int> targets(order.begin() + 2, order.end());
          ok = ok && spread_pulls(R, M, kb, targets, phase_begin[1], phase_begin[p1]);
        } else {
          std::vector<int> targets(order.begin() + 1, order.end());
          ok = spread_pulls(R, M, kb, targets, phase_begin[0], phase_begin[1]);
        }
        if (ok && Q >= 2 && R.calls[phase_begin[1] - 1].f_owner == order[1] && R.calls[phase_begin[0]].f_owner == order[1] &&
            (p1 >= 2 || S == 2))
          lag_pulls(R, phase_begin[0], phase_begin[1]);
        // deadline check: every pull of slab order[i] sits in a call before the first call that multiplies it
        for (int i = 1; i < S && ok; i++) {
          const int g = order[i];
          int last_pull = -1, first_use = (int)R.calls.size();
          for (int c = 0; c < (int)R.calls.size(); c++) {
            if (R.calls[c].f_owner == g) last_pull = c;
            if (R.calls[c].k0 < kb[g + 1] && R.calls[c].k0 + R.calls[c].kc > kb[g] && c < first_use) first_use = c;
          }
          ok = last_pull >= 0 && last_pull < first_use;
        }
      }
    }
    if (ok) {
      // which piece of the owner's upload has to have landed before a pull may start
      for (RankCall& c : R.calls) {
        c.f_need_piece = 0;
        if (c.f_owner < 0) continue;
        const std::vector<int> pc = rank_slab_pieces(M, kb[c.f_owner], kb[c.f_owner + 1]);
        int q = 0;
        while (q + 1 < (int)pc.size() - 1 && pc[q + 1] < c.f_k0 + c.f_kc) q++;
        c.f_need_piece = q + 1;
      }
      R.ok = true;
      return R;
    }
    if (p1 == S) break;
    p1 = S;   // fall back to the plain ring, which always has a carrier per slab
  }
  R.calls.clear();
  R.copies.clear();
  return R;
}

}  // namespace dg
