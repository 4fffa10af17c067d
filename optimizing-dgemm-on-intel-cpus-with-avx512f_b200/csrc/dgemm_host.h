// dgemm_host.h — internal interface between the kernel launchers (dgemm_kernels.cu) and the C-ABI (dgemm_api.cu).
#pragma once
#include <cuda_runtime.h>

#include <vector>

namespace dg {

struct Params;

enum Variant {
  VARIANT_AUTO = 0,       // the TMA kernel; its parity-split variant for odd lda / ldb or 8-byte aligned bases
  VARIANT_TMA = 1,        // dgemm_tma_ws_kernel / dgemm_tma_ws_split_kernel (unaligned operands need K >= 2, N >= 2)
  VARIANT_CPASYNC8 = 2,   // dgemm_cpasync_kernel<1>: any alignment, any leading dimension
  VARIANT_CPASYNC16 = 3,  // dgemm_cpasync_kernel<2>: same eligibility as TMA
  VARIANT_NAIVE = 4,      // ladder rung 1 (dgemm_ladder.cu): never chosen by VARIANT_AUTO
  VARIANT_DFMA = 5,       // ladder rung 2: shared-memory tiles + 8x8 register tile on the CUDA-core FP64 pipe
  VARIANT_CLUSTER = 6,    // ladder rung 6 (dgemm_kernels_cluster.cu): 2-CTA clusters, A panel by TMA multicast; data-parallel
                          // tile pairs only, TMA-eligible operands only; never chosen by VARIANT_AUTO
};

bool tma_eligible(const Params& P);
int tma_split_bits(const Params& P);
int tma_mis_halves(const Params& P);
bool tma_split_ok(const Params& P);   // needs P.A, P.B, P.lda, P.ldb, P.K, P.N
// cooperative: the launch must fail (cudaErrorCooperativeLaunchTooLarge) rather than start a grid that is not wholly
// co-resident -- for configurations whose CTAs wait on each other (k-lockstep, parallel stream-K fix-up); pdl: allow
// programmatic dependent launch behind the previous kernel of the stream (ignored when cooperative).  TMA variants only.
cudaError_t launch_gemm(const Params& P, int variant, int num_sms, cudaStream_t stream, bool cooperative = false,
                        bool pdl = false);
// CTAs of the TMA kernel the current device can hold at once; tile_small as in Params (0: 128 x 128, 1: dg64, 2: dg64x2, 3: dg128x64)
int tma_coresident_capacity(int tile_small, int num_sms);   // (4: dg64x32)
cudaError_t small_tile_occupancy(int* ctas_per_sm);
cudaError_t small_tile_occupancy_x2(int* ctas_per_sm);
void tmap_cache_stats(unsigned long long* hits, unsigned long long* misses);
// the TMA kernel on 64 x 64 tiles (dgemm_kernels_small.cu); maps = {tmA, tmA1, tmB, tmB1} (CUtensorMap*, the second of a
// pair only read when the operand's SPLIT bit is set), boxes 16 x 16 (or 16 x 8 per k parity) / 16 x 64 (16 x 32)
constexpr int SMALL_BM = 64, SMALL_BN = 64;
cudaError_t launch_tma_small(const void* const maps[4], int split, const Params& P, int grid, int dp_units,
                             cudaStream_t stream, bool cooperative, bool pdl);
// the same at two CTAs per SM (dgemm_kernels_small2.cu); Params::tile_small == 2
cudaError_t launch_tma_small_x2(const void* const maps[4], int split, const Params& P, int grid, int dp_units,
                                cudaStream_t stream, bool cooperative, bool pdl);
constexpr int SMALL_X2_CTAS_PER_SM = 2;
// the TMA kernel on 128 x 64 tiles (dgemm_kernels_mid.cu); Params::tile_small == 3; B boxes 16 x 64 as the 64 x 64 build's
constexpr int MID_BM = 128, MID_BN = 64;
cudaError_t launch_tma_small_mid(const void* const maps[4], int split, const Params& P, int grid, int dp_units,
                                 cudaStream_t stream, bool cooperative, bool pdl);
cudaError_t small_tile_occupancy_mid(int* ctas_per_sm);
// the TMA kernel on 64 x 32 tiles (dgemm_kernels_tiny.cu); Params::tile_small == 4; B boxes 16 x 32, one k-tile per ring slot
constexpr int TINY_BM = 64, TINY_BN = 32;
cudaError_t launch_tma_small_tiny(const void* const maps[4], int split, const Params& P, int grid, int dp_units,
                                  cudaStream_t stream, bool cooperative, bool pdl);
cudaError_t small_tile_occupancy_tiny(int* ctas_per_sm);
enum { TILE_BIG = 0, TILE_SMALL = 1, TILE_SMALL_X2 = 2, TILE_MID = 3, TILE_TINY = 4, TILE_COUNT = 5 };   // Params::tile_small
inline int tile_bm(int t) { return t == TILE_BIG ? 128 : (t == TILE_MID ? MID_BM : (t == TILE_TINY ? TINY_BM : SMALL_BM)); }
inline int tile_bn(int t) { return t == TILE_BIG ? 128 : (t == TILE_MID ? MID_BN : (t == TILE_TINY ? TINY_BN : SMALL_BN)); }
cudaError_t launch_ladder(const Params& P, int variant, cudaStream_t stream);
cudaError_t launch_tma_cluster(const void* const maps[2], const Params& P, int num_sms, cudaStream_t stream);
cudaError_t launch_scale_c(double* C, int M, int N, long long ldc, double beta, int num_sms, cudaStream_t stream);
// C += beta * X, both M x N column-major
cudaError_t launch_add_scaled(double* C, long long ldc, const double* X, long long ldx, int M, int N, double beta,
                              int num_sms, cudaStream_t stream);
// dst = src (rows x cols, column-major) with a new leading dimension
cudaError_t launch_realign(const double* src, long long lds, double* dst, long long ldd, int rows, int cols, int num_sms,
                           cudaStream_t stream);
// dst (cols x rows, ldd) = src^T, src rows x cols with leading dimension lds, both column-major
cudaError_t launch_transpose(const double* src, long long lds, double* dst, long long ldd, int rows, int cols,
                             cudaStream_t stream);
cudaError_t launch_fill_signed(double* X, int rows, int cols, long long ld, unsigned long long seed,
                               unsigned long long row0, unsigned long long col0, int num_sms, cudaStream_t stream);
// schedule of the host-pointer entry (dgemm_hostplan.cpp): column-panel / K-slab boundaries and the order of the
// (slab, panel) device calls; the H2D copies are queued in the order the calls first need them
struct HostPlanItem { int s, p, s_end; };  // device call: slabs [s, s_end) of K against panel p
struct HostPlan {
  std::vector<int> nb, kb;
  std::vector<HostPlanItem> order;
  int deferred_panels = 0;    // panels [0, deferred_panels): beta*C is added AFTER their last call (see dgemm_hostplan.cpp)
  double predicted_ms = 0.0;  // the planner's own estimate of the whole call (model in dgemm_hostplan.cpp)
};
HostPlan plan_host_pipeline(int M, int N, int K, bool need_ab, bool need_c_in, int num_sms, bool pageable = false);

// schedule of one rank's pass of the column-panel sharded DGEMM (dgemm_rankplan.cpp)
enum { RANK_COPY_A = 0, RANK_COPY_B = 1, RANK_COPY_X = 2 };
struct RankCopy {      // one H2D copy, queued in this order on the rank's copy stream
  int kind;            // RANK_COPY_A: piece `piece` of the own K-slab of A (columns [k0, k0+kc)); the owner's sequence word
                       //              is advanced to piece+1 behind it;  RANK_COPY_B: rows [k0, k0+kc) of columns
                       //              [c0, c0+nc) of the B panel;  RANK_COPY_X: C_in of columns [c0, c0+nc) (deferred C)
  int k0, kc, c0, nc, piece;
};
struct RankCall {      // one device call: C[:, c0:c0+nc] (+)= alpha * A[:, k0:k0+kc] * B[k0:k0+kc, c0:c0+nc]
  int k0, kc, c0, nc, sub;
  int first, last;     // first call on this sub-panel (beta_eff = 0, or beta when C is not deferred); last: add beta*C_in, D2H
  int f_owner, f_k0, f_kc;   // NVLink pull carried by this call: columns [f_k0, f_k0+f_kc) of A from rank f_owner (-1: none)
  int f_need_piece;    // ... which may start once this many upload pieces of the owner have landed
  int wait_copy;       // index in copies of the last H2D copy the call needs
  int wait_x_copy;     // last call of a sub-panel with beta != 0: index of its C_in copy, else -1
};
struct RankPlan {
  bool ok = false;
  int p1 = 0;                    // slabs multiplied slab-major (phase 1); the rest is panel-major
  std::vector<int> order;        // non-empty slabs in ring order, own first
  std::vector<int> pieces;       // absolute k boundaries of the own slab's upload pieces
  std::vector<int> sp;           // column sub-panel boundaries inside the rank's panel
  std::vector<RankCopy> copies;
  std::vector<RankCall> calls;
};
std::vector<int> rank_slab_pieces(int M, int k_begin, int k_end);
std::vector<int> rank_sub_panels(int ncols);
RankPlan plan_rank_pipeline(int M, int ncols, int K, int G, int rank, const int* kb, bool need_c_in, int p1_override = 0);

// executor of that schedule (dgemm_rank.cu)
struct RankArgs {
  int M = 0, ncols = 0, K = 0;
  double alpha = 1.0, beta = 0.0;
  const double* A_slab = nullptr; int LDA = 1;   // host: columns [kb[rank], kb[rank+1]) of A
  const double* B = nullptr; int LDB = 1;        // host: the rank's column panel of B (K x ncols)
  double* C = nullptr; int LDC = 1;              // host: the rank's column panel of C (M x ncols), in place
  int rank = 0, world = 1;
  const int* kb = nullptr;                       // world + 1 K-slab boundaries
  void* const* xbuf = nullptr;                   // world exchange buffers (own: local; peers': IPC / peer-access pointers)
  unsigned int pass = 0;                         // same value on every rank, +1 per pass on these exchange buffers
  int p1 = 0;                                    // > 0: force the number of slab-major slabs (0: planner's choice)
  double* timeline = nullptr; int timeline_cap = 0; int* timeline_n = nullptr;   // optional (kind, index, ms) triples
};
enum { RANK_TL_H2D_DONE = 0, RANK_TL_CALL_START = 1, RANK_TL_CALL_END = 2, RANK_TL_D2H_DONE = 3 };
int rank_pipeline_exec(int dev, const RankArgs& a);
size_t exchange_flag_offset(int M, int K);
size_t exchange_bytes(int M, int K);
// *flag = value, stream-ordered (cuStreamWriteValue32; a one-thread kernel where the driver has no stream memory operations)
cudaError_t launch_set_flag(unsigned int* flag, unsigned int value, cudaStream_t stream);
// holds `stream` until (int)(*flag - need) >= 0; flag may live in a peer GPU's memory.  Bounded: after timeout_ns *fail is set
cudaError_t launch_wait_flag(const unsigned int* flag, unsigned int need, unsigned int* fail, unsigned long long timeout_ns,
                             cudaStream_t stream);

// pageable host memory <-> HBM through per-thread pinned slots (dgemm_staging.cu)
bool host_pointer_is_pageable(const void* p);
bool staging_usable(int dev);
cudaError_t staged_h2d(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                       cudaStream_t stream, int dev);
cudaError_t staged_d2h_async(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                             cudaEvent_t ready, cudaStream_t stream, int dev);
cudaError_t staged_d2h_wait();
// the streaming form of staged_h2d: submit every block ahead (ticket per block, < 0 on error), wait per ticket
long long staging_submit_h2d(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                             cudaStream_t stream, int dev, bool pageable, cudaError_t* err);
cudaError_t staging_wait_submitted(long long ticket);
void staging_trace_dump(double t0_ms);   // DGEMM_B200_HOST_TRACE
void staging_drain();   // returns when neither team holds a block any more (error paths)
void staging_shutdown();

int launch_count();
void reset_launch_count();

}  // namespace dg
