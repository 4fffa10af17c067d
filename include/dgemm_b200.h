/*
 * dgemm_b200.h — C ABI of the B200 (sm_100a) DGEMM hot path.  Plain C: pointers and sizes only.
 *
 * Drop-in boundary: the reference dispatches twenty kernels with one 11-argument signature,
 *     void mydgemm_cpu_vNN(int M, int N, int K, double alpha, double *A, int LDA, double *B, int LDB,
 *                          double beta, double *C, int LDC)
 * (reference include/kernel18.h:629-640, include/kernel19.h:678-689) through test_mydgemm_vNN / test_kernel
 * (reference utils.c:159-165, 167-191).  mydgemm_gpu_v20 / mydgemm_gpu_v21 below are "one more kernel_number":
 * same parameters, same meaning, HOST pointers, column-major, NoTrans/NoTrans, C updated in place.
 * include/kernel20.h wraps them for a kernels.h-style aggregator; INTEGRATION.md shows the reference-side patch.
 *
 * Semantics (BLAS / kernel18 order, reference kernel18.h:644-645):
 *   M == 0 or N == 0      -> no-op
 *   alpha == 0 or K == 0  -> C = beta*C only (kernel19's skip of the beta scaling, kernel19.h:693, is NOT kept)
 *   beta == 0             -> C is not read (NaN/Inf in C are overwritten)
 *   rows M..LDC-1 of every column of C (caller padding) are never written; A and B are read-only
 *   odd M / N are computed correctly (the reference drops them, kernel18.h:612-625: that bug is NOT kept)
 * The 11-argument entries return void like the reference's; failures print one line to stderr and set a
 * status readable with dgemm_b200_last_status().  There is NO CPU fallback: without a usable CUDA device the
 * call fails with DGEMM_B200_ERR_NO_DEVICE and leaves C untouched.
 * Threading: entries may be called from any host thread, one call at a time (a process-wide mutex serialises).
 */
#ifndef DGEMM_B200_H
#define DGEMM_B200_H
#include <stddef.h>
#ifdef __cplusplus
extern "C" {
#endif

enum {
  DGEMM_B200_OK = 0,
  DGEMM_B200_ERR_NO_DEVICE = 1,    /* no CUDA device / driver, or not compute capability 10.x */
  DGEMM_B200_ERR_BAD_ARG = 2,      /* negative dimension, LDA < M, LDB < K, LDC < M, null pointer */
  DGEMM_B200_ERR_OOM = 3,          /* device or pinned-host allocation failed */
  DGEMM_B200_ERR_CUDA = 4,         /* any other CUDA runtime / driver error (message in last_error) */
  DGEMM_B200_ERR_NCCL = 5          /* multi-GPU entry: NCCL error */
};

/* kernel variants, for tests / ablations (dgemm_b200_device_ex) */
enum {
  DGEMM_B200_VARIANT_AUTO = 0,       /* the TMA kernel for every operand layout (see DGEMM_B200_ODD_LD in INTEGRATION.md) */
  DGEMM_B200_VARIANT_TMA = 1,        /* TMA + mbarrier warp-specialised persistent kernel; operands with an odd leading
                                        dimension or an 8-byte-aligned base are read in place by its parity-split
                                        variant (one parity by TMA, the other by cp.async from spare warps) */
  DGEMM_B200_VARIANT_CPASYNC8 = 2,   /* cp.async 8-byte loader: any alignment / odd leading dimensions */
  DGEMM_B200_VARIANT_CPASYNC16 = 3,  /* cp.async 16-byte loader */
  DGEMM_B200_VARIANT_NAIVE = 4,      /* optimisation-ladder rung 1: one thread per entry (never auto-selected) */
  DGEMM_B200_VARIANT_DFMA = 5,       /* ladder rung 2: smem tiles + 8x8 register tile on the CUDA-core FP64 pipe */
  DGEMM_B200_VARIANT_CLUSTER = 6     /* ladder rung 6: 2-CTA thread-block clusters, the shared A panel by TMA multicast
                                        (data-parallel tile pairs, TMA-eligible operands only; never auto-selected) */
};

/* ---- the reference's kernel signature (replaces mydgemm_cpu_v18 / v19 at utils.c:159-165) ---------------- */
/* single GPU; host pointers (pageable or pinned); includes H2D of A, B (and C when beta != 0) and D2H of C */
void mydgemm_gpu_v20(int M, int N, int K, double alpha, double *A, int LDA, double *B, int LDB, double beta,
                     double *C, int LDC);
/* all visible GPUs of the box: column panels of B and C per GPU, A replicated over NVLink with NCCL */
void mydgemm_gpu_v21(int M, int N, int K, double alpha, double *A, int LDA, double *B, int LDB, double beta,
                     double *C, int LDC);

/* ---- the CBLAS signature the reference harness calls for kernel 0 and for its own verification (reference
 * test.c:38, utils.c:87-89: cblas_dgemm(CblasColMajor, CblasNoTrans, CblasNoTrans, m,n,k, alpha, A,m, B,k, beta, C,m)).
 * Host pointers, single GPU.  layout: 101 row-major / 102 column-major; transa / transb: 111 NoTrans, 112 Trans,
 * 113 ConjTrans (= Trans for real data) — the numeric values of CBLAS_LAYOUT / CBLAS_TRANSPOSE, so an `mkl.h`
 * stand-in can forward cblas_dgemm to it unchanged (harness/mkl_b200/mkl.h).  NoTrans x NoTrans in either layout is
 * the pipelined mydgemm_gpu_v20 path; a transposed operand is copied whole and transposed on the device first. */
enum {
  DGEMM_B200_ROW_MAJOR = 101, DGEMM_B200_COL_MAJOR = 102,
  DGEMM_B200_NO_TRANS = 111, DGEMM_B200_TRANS = 112, DGEMM_B200_CONJ_TRANS = 113
};
void dgemm_b200_cblas_dgemm(int layout, int transa, int transb, int M, int N, int K, double alpha, const double *A,
                            int lda, const double *B, int ldb, double beta, double *C, int ldc);

int dgemm_b200_last_status(void);        /* status of the most recent call on this thread's process */
const char *dgemm_b200_last_error(void); /* human-readable message for the last failure ("" if none) */

/* ---- lifetime ------------------------------------------------------------------------------------------- */
int dgemm_b200_device_count(void);        /* usable sm_100 devices; 0 when there is none */
int dgemm_b200_init(int device);          /* optional: create the context for `device` now; returns status */
int dgemm_b200_shutdown(void);            /* frees every cached device / pinned buffer, stream and communicator */
int dgemm_b200_set_max_gpus(int n);       /* cap the GPUs v21 uses (0 = all visible) */

/* ---- device-pointer API (kernel-only path; what the 11-argument entries call after staging) ----------------
 * dA, dB, dC are device pointers on the CURRENT device; `stream` is a cudaStream_t (NULL = default stream).
 * Asynchronous with respect to the host. Returns a status code. */
int dgemm_b200_device(int M, int N, int K, double alpha, const double *dA, long long lda, const double *dB,
                      long long ldb, double beta, double *dC, long long ldc, void *stream);
/* same, forcing a kernel variant and a K-decomposition: splits = 0 heuristic (TMA kernel: data-parallel waves +
 * stream-K tail; cp.async kernel: split-K when there are too few tiles), 1 = pure data-parallel, > 1 = split-K
 * with that many splits and a deterministic reduce pass */
int dgemm_b200_device_ex(int M, int N, int K, double alpha, const double *dA, long long lda, const double *dB,
                         long long ldb, double beta, double *dC, long long ldc, void *stream, int variant,
                         int splits);
/* C = alpha*op(A)*op(B) + beta*C on device pointers, same layout / transpose codes as dgemm_b200_cblas_dgemm.
 * lda / ldb are the leading dimensions of A and B AS STORED. */
int dgemm_b200_device_op(int layout, int transa, int transb, int M, int N, int K, double alpha, const double *dA,
                         long long lda, const double *dB, long long ldb, double beta, double *dC, long long ldc,
                         void *stream);
/* The split-K DGEMM with NO allocation of any kind.  (No counterpart in the reference, whose kernels malloc and free their
 * pack buffers inside every call, include/kernel18.h:647-648,670; this entry exists for GPU callers only.)
 * Meant for programs that cannot let a library call cudaMalloc: a process whose own memory pool has already taken nearly all
 * of the HBM (the library's request would fail with out-of-memory, or eat into what the pool believes it owns), and code in a
 * latency-critical or many-stream loop, where the device-wide lock of cudaMalloc and the device-wide wait of cudaFree stall
 * every other stream.  All other device entries may allocate (context on first use, workspace growth later).
 * Here the partial tiles of the split-K kernels live in an array that is part of the library image: every GPU the library
 * runs on carries DGEMM_B200_STATIC_SCRATCH_BYTES of it from module load to process exit, used or not.
 *   operands   device pointers on the CURRENT device, as for dgemm_b200_device; `stream` a cudaStream_t (NULL = default)
 *   splits     0 = choose (2 .. what the scratch holds), or >= 2; 1 is rejected (nothing to keep in scratch: a caller that
 *              wants no split has no use for this entry)
 *   limit      ceil(M/128) * ceil(N/128) * splits <= DGEMM_B200_STATIC_SCRATCH_BYTES / (128*128*8) = 128 partial tiles, so
 *              at most 64 output tiles (e.g. 1024 x 1024, any K).  Larger calls fail with DGEMM_B200_ERR_BAD_ARG and a
 *              message naming the limit; nothing is launched and C is untouched
 *   result     bit-identical to dgemm_b200_device_ex(variant = TMA for 16-byte aligned operands with even LDA / LDB, else
 *              CPASYNC8, same splits); same degenerate-case semantics as every entry (alpha == 0 or K == 0: C = beta*C)
 *   sharing    every call on a GPU uses the same array.  Calls on one stream may follow each other freely (nothing in the
 *              array has to be zero at the start: each unit overwrites its whole slot before it is read).  Calls whose
 *              kernels can run at the same time on one GPU — different streams — race and return garbage SILENTLY:
 *              order them with events
 *   first call on each device: looks up the array's device address and the SM count (cudaGetSymbolAddress, no
 *              allocation).  Make it before any stream capture starts on that device */
#define DGEMM_B200_STATIC_SCRATCH_BYTES (16u << 20)
int dgemm_b200_device_static_scratch(int M, int N, int K, double alpha, const double *dA, long long lda,
                                     const double *dB, long long ldb, double beta, double *dC, long long ldc,
                                     void *stream, int splits);
/* C (M x N, ldc) += beta * X (M x N, ldx), device pointers, asynchronous on `stream`.  The "deferred C" step of a
 * pipelined caller: start the panel's DGEMM calls with beta = 0 while C is still on its way over PCIe, add beta*C_in
 * after the last one (mydgemm_gpu_v20 does this internally; bench.py's multi-GPU end-to-end pipeline uses this entry).
 * The reference applies beta first (scale_c_k18, kernel18.h:12-46, 644); the sum is the same up to rounding. */
int dgemm_b200_add_scaled(int M, int N, double beta, const double *dX, long long ldx, double *dC, long long ldc,
                          void *stream);
/* DGEMM fused with an NVLink fetch: while the FP64 tensor-core warps compute this call, spare warps of the same
 * kernel copy fetch_bytes from fetch_src to fetch_dst (16-byte aligned, multiple of 16).  fetch_src is normally a
 * PEER GPU's buffer (dgemm_b200_ipc_open / peer access): the multi-GPU layer uses it to pull the NEXT K-slab of A
 * while the current one is being multiplied, so the replication of A never waits for free SMs.  fetch_bytes = 0
 * makes it identical to dgemm_b200_device.  The copy is complete when the call's kernel is (stream order). */
int dgemm_b200_device_fetch(int M, int N, int K, double alpha, const double *dA, long long lda, const double *dB,
                            long long ldb, double beta, double *dC, long long ldc, void *stream,
                            const void *fetch_src, void *fetch_dst, size_t fetch_bytes);
/* CUDA IPC plumbing for one-process-per-GPU launches: export a dgemm_b200_malloc'ed buffer as a 64-byte handle,
 * open a peer's handle as a device pointer usable as fetch_src, close it */
int dgemm_b200_ipc_export(const void *dptr, void *handle64);
int dgemm_b200_ipc_open(const void *handle64, void **dptr);
int dgemm_b200_ipc_close(void *dptr);
/* ---- one rank's share of a column-panel sharded DGEMM, host operands in and out (one process per GPU) ------------------
 * The multi-GPU counterpart of the reference's parallel kernel (include/kernel19.h:678-755: one OpenMP thread per row
 * band, partition_m_dim :636-656).  Rank r of `world` owns the column panel J_r of B and C and the K-slab
 * [kb[r], kb[r+1]) of A, all in HOST memory (pinned for full speed), and computes
 *     C[:, J_r] = beta * C[:, J_r] + alpha * A * B[:, J_r]
 * on the CURRENT device.  A is completed over NVLink: every rank uploads only its own slab into its exchange buffer
 * (dgemm_b200_exchange_alloc; peers' buffers opened with dgemm_b200_ipc_open) and the DGEMM kernels of the other ranks pull
 * it from there with their spare warps while they compute.  Ranks never wait for each other on the host and there is no
 * collective: the only dependency is "piece q of rank g's slab has landed", a sequence word behind g's H2D copies.
 *   A_slab    M x (kb[r+1]-kb[r]), leading dimension LDA          B_panel  K x ncols, LDB          C_panel  M x ncols, LDC
 *   kb        world + 1 ascending K boundaries, kb[0] = 0, kb[world] = K, the same on every rank; every rank needs a
 *             non-empty slab (multiples of 16 keep the pulls 16-byte aligned)
 *   xbuf      world exchange buffers for this (M, K): xbuf[rank] the rank's own, xbuf[g] rank g's as mapped here
 *   pass      the same number on every rank, increased by one for every pass on these exchange buffers; the caller keeps
 *             pass p+1 from starting on any rank before every rank has returned from pass p (one host barrier)
 *   timeline  optional: receives (kind, index, milliseconds since the pass began) triples, kind 0 = H2D for call `index`
 *             complete, 1 / 2 = call `index` started / finished, 3 = D2H of column sub-panel `index` complete
 * Same semantics as mydgemm_gpu_v20 on the panel (alpha == 0 or K == 0 -> C = beta*C; beta == 0 -> C not read).
 * Errors: when a peer's slab does not arrive within DGEMM_B200_FLAG_TIMEOUT_MS the call returns DGEMM_B200_ERR_CUDA AFTER the
 * pass has drained, and the caller's C_panel is then UNDEFINED (sub-panels computed on the incomplete A have already been
 * written back): keep a copy of C if beta != 0 and the pass must be repeatable.
 * Host memory: the overlap of H2D, compute and D2H (and the quoted throughput) needs PINNED A_slab / B_panel / C_panel;
 * pageable operands are correct but every copy then blocks the launching thread (also true for mydgemm_gpu_v21, which
 * runs this pipeline on one thread per GPU: the reference harness's malloc'ed matrices take that slower road). */
int dgemm_b200_rank_pipeline(int M, int ncols, int K, double alpha, const double *A_slab, int LDA,
                             const double *B_panel, int LDB, double beta, double *C_panel, int LDC, int rank,
                             int world, const int *kb, void *const *xbuf, unsigned int pass, double *timeline,
                             int timeline_cap, int *timeline_n);
size_t dgemm_b200_exchange_bytes(int M, int K);   /* size of an exchange buffer: A (ld = M rounded up to even) + sequence words */
void *dgemm_b200_exchange_alloc(int M, int K);    /* on the current device, sequence words zeroed; free with dgemm_b200_free */
/* the schedule dgemm_b200_rank_pipeline uses (no device needed): n_calls x 12 ints {k0, kc, c0, nc, first, last, pull owner
 * (-1: none), pull k0, pull kc, owner pieces needed, H2D copy awaited, C_in copy awaited (-1: none)} and n_copies x 6 ints
 * {kind 0 = own A piece / 1 = B block / 2 = C_in, k0, kc, c0, nc, piece}; p1 > 0 forces the number of slab-major slabs */
int dgemm_b200_rank_plan(int M, int ncols, int K, int world, int rank, const int *kb, int need_c_in, int p1, int *calls,
                         int cap_calls, int *n_calls, int *copies, int cap_copies, int *n_copies, int *p1_out);
/* re-read the DGEMM_B200_* environment switches (they are read once, at the first call) */
int dgemm_b200_reload_env(void);

/* plan the heuristic would use: writes variant, split-K factor, CTA count and the number of output tiles handled
 * by the stream-K tail (any out pointer may be NULL); returns status */
int dgemm_b200_plan(int M, int N, int K, const double *dA, long long lda, const double *dB, long long ldb,
                    int *variant, int *splits, int *ctas, int *sk_tiles);

/* schedule mydgemm_gpu_v20 would use for this shape (no device needed, nothing is launched): the n_panels + 1
 * column-panel boundaries, the n_slabs + 1 K-slab boundaries and the order of the device calls as n_items triples
 * {first slab, panel, end slab}: the call multiplies K-slabs [first, end) of A with the matching rows of panel `panel`.  Every array holds `cap` ints (any may be NULL); predicted_ms is the planner's model
 * estimate of the whole call from pinned host memory; deferred_panels (may be NULL) is the number of leading panels
 * whose beta*C term is added after their last call instead of before their first (so that the start of the pipeline
 * does not wait for C to arrive).  Plays the role of the reference's N -> K -> M blocking
 * loops (include/kernel18.h:652-669) when the operands start in host memory. */
int dgemm_b200_host_plan(int M, int N, int K, double alpha, double beta, int *panels, int *n_panels, int *slabs,
                         int *n_slabs, int *order, int *n_items, int cap, double *predicted_ms, int *deferred_panels);

/* Out of core (one GPU, host-pointer entries): when A, B, C and the pipeline's side buffers exceed the HBM budget
 * (DGEMM_B200_HBM_BUDGET bytes, default 90 % of the free HBM), mydgemm_gpu_v20 / dgemm_b200_host cut C into Mb x Nb blocks
 * (and K into chunks of Kb only if one block's operands still do not fit) and run the in-core pipeline block by block.
 * This reports the blocking a budget leads to and the PCIe bytes it moves (no device needed; whole dimensions = in core).
 * Counterpart of the reference's outer N / K blocking loops (include/kernel18.h:652-669), one memory level further out. */
int dgemm_b200_host_blocking(int M, int N, int K, double alpha, double beta, double budget_bytes, int *Mb, int *Nb, int *Kb,
                             double *pcie_bytes);
/* Which build of the TMA kernel a device call of this shape runs (16-byte aligned operands, even leading dimensions, no fused
 * fetch) on a GPU with num_sms SMs, how many CTAs it launches and how many of its tiles are cut into stream-K ranges.  Pure
 * host logic, no device needed.  tile: 0 = 128x128, 1 = 64x64, 2 = 64x64 at two CTAs per SM, 3 = 128x64, 4 = 64x32 (the
 * builds play the role of the reference's edge micro-kernels beside the 24x8 one, include/kernel18.h:533-600). */
int dgemm_b200_tile_plan(int M, int N, int K, int num_sms, int *tile, int *ctas, int *stream_k_tiles);
/* bytes of device memory the library's cached buffers hold on the current device (operand staging, workspaces) */
size_t dgemm_b200_device_bytes(void);

/* ---- helpers so a plain-C harness needs no CUDA headers -------------------------------------------------------- */
void *dgemm_b200_malloc(size_t bytes);                 /* device memory on the current device (NULL on failure) */
void dgemm_b200_free(void *dptr);
void *dgemm_b200_host_alloc(size_t bytes);             /* pinned host memory */
void dgemm_b200_host_free(void *hptr);
/* 2-D column-major copies: `cols` columns of `rows` doubles, leading dimensions in elements */
int dgemm_b200_h2d(double *dst_dev, long long ldd, const double *src_host, long long lds, long long rows,
                   long long cols, void *stream);
int dgemm_b200_d2h(double *dst_host, long long ldd, const double *src_dev, long long lds, long long rows,
                   long long cols, void *stream);
int dgemm_b200_sync(void *stream);                      /* cudaStreamSynchronize (NULL: whole device) */
/* counter-based U(-1,1) fill on the device; entry (i,j) = f(seed, row0+i, col0+j), identical to the oracle's
 * oracle_signed_value, so huge matrices never have to exist on the host */
int dgemm_b200_fill_signed(double *dX, int rows, int cols, long long ld, unsigned long long seed,
                           unsigned long long row0, unsigned long long col0, void *stream);
/* runs the device call `reps` times between two CUDA events on `stream`; returns average milliseconds
 * per call (< 0 on failure). flush_l2_bytes > 0: a buffer of that size is overwritten before every rep. */
double dgemm_b200_time_device(int M, int N, int K, double alpha, const double *dA, long long lda, const double *dB,
                              long long ldb, double beta, double *dC, long long ldc, int variant, int splits,
                              int warmup, int reps, size_t flush_l2_bytes);
/* profiling aid: when dbuf != NULL (device buffer of 16 unsigned long long per CTA of the grid; 1024 * 16 is always enough),
 * every CTA of the TMA kernel writes
 * up to 16 %globaltimer stamps (entry, consumer start, then per segment: operands landed / main loop done / epilogue
 * done) into dbuf[cta*16 ...].  NULL switches it off (default). */
int dgemm_b200_debug_timeline(unsigned long long *dbuf);
/* test aid for the co-residency guarantee of the kernels that wait for each other (k-lockstep, parallel stream-K fix-up):
 * launches, on a stream of its own, `ctas` CTAs that each hold `smem_bytes` of shared memory for `ms` milliseconds
 * (<= 5000) and returns at once.  A DGEMM issued while they run must still finish with the right result. */
int dgemm_b200_debug_occupy_sms(int ctas, unsigned smem_bytes, double ms);
/* measurement aid: microseconds an EMPTY kernel of the given launch shape takes when timed the way dgemm_b200_time_device
 * times a DGEMM (event, launch, event, synchronise).  flags: 1 cooperative launch, 2 programmatic-dependent-launch
 * attribute, 4 with 512 bytes of kernel parameters.  < 0 on failure. */
double dgemm_b200_debug_launch_floor(int ctas, int threads, unsigned smem_bytes, int flags, int reps);
/* number of kernels this library has launched since the last reset (bench.py's gpu_launches) */
int dgemm_b200_launch_count(void);
void dgemm_b200_reset_launch_count(void);
/* SM clock (MHz) as seen by a short in-kernel clock64()/globaltimer measurement; < 0 on failure */
double dgemm_b200_sm_clock_mhz(void);

#ifdef __cplusplus
}
#endif
#endif /* DGEMM_B200_H */
