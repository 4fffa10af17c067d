// dgemm_staging.cu — pageable host memory <-> HBM at PCIe speed.
//
// The reference harness hands its kernels plain malloc'ed matrices (test.c:27-30), and mydgemm_gpu_v20 keeps that
// contract.  cudaMemcpy from pageable memory is staged by the driver through one internal bounce buffer on the
// calling thread (11 GB/s measured on this pool's hosts, against 55 GB/s for pinned memory), which would make the
// drop-in call several times slower than the same call on pinned buffers.  This file does the staging itself.
//
// Up (host -> device): a STREAM of copy jobs.  The caller submits every block it will need, in the order it will need
// them, and gets a ticket per block; a team of host threads walks the blocks chunk by chunk (4 MB) without stopping at
// block boundaries: copy the chunk of the caller's matrix into one of the thread's two pinned slots, queue the slot's DMA on
// the copy stream, go on with the other slot.  staging_wait_submitted(ticket) returns once every chunk of every block up
// to that ticket has been queued, i.e. an event recorded on the copy stream from then on covers them.  The first
// version ran one blocking parallel-for per block and reached 30 GB/s (every block ramped the team up and down and
// the caller could not look ahead); tools/stage_probe.cu measures what the host can do: one thread fills slots at
// 11.5 GB/s, 6-8 threads with two 4 MB slots each sustain 50 GB/s through the whole fill -> DMA pipeline, any copy
// routine (glibc memcpy, rep movsb, AVX-512 non-temporal: 50.6 / 50.4 / 51.5 GB/s; profiles/r02_stage_probe.jsonl).
//
// Down (device -> host): the mirror image, two DMAs in flight per thread (the next chunk's DMA is requested before
// the previous one is copied out); 4 threads reach 40 GB/s, 8 reach 50 (same probe).  A block becomes visible to the
// team when a host function queued behind its last kernel has run; until then the threads sleep.
//
// No arithmetic happens here — the bytes are moved, never touched.
#include "dgemm_host.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <memory>
#include <mutex>
#include <thread>
#include <vector>

namespace dg {
namespace {

constexpr size_t SLOT_BYTES = 4u << 20;   // size of a pinned slot = largest chunk
// chunk size in use (DGEMM_B200_STAGE_CHUNK_KB, 64 ... 4096; read when the teams start)
size_t g_chunk_bytes = SLOT_BYTES;
#define CHUNK_BYTES g_chunk_bytes

struct Slot {
  char* host = nullptr;
  cudaEvent_t done = nullptr;  // the DMA that last used this slot
  bool used = false;
};

struct Chunk {
  long long row0, rows, col0, cols;
};

std::vector<Chunk> make_chunks(long long rows, long long cols, size_t chunk_bytes) {
  std::vector<Chunk> out;
  const long long max_elems = chunk_bytes / 8;
  if (rows <= max_elems) {
    const long long cpc = std::max(1ll, max_elems / rows);
    for (long long c = 0; c < cols; c += cpc) out.push_back({0, rows, c, std::min(cpc, cols - c)});
  } else {
    for (long long c = 0; c < cols; c++)
      for (long long r = 0; r < rows; r += max_elems) out.push_back({r, std::min(max_elems, rows - r), c, 1});
  }
  return out;
}

struct Job {
  double* dev_ptr = nullptr;     // device side
  long long ldd = 0;
  double* host_ptr = nullptr;    // caller's side (read for up, written for down)
  long long ldh = 0;
  long long rows = 0, cols = 0;
  bool staged = true;            // false: pinned caller memory, one direct 2-D copy (up only)
  cudaEvent_t ready = nullptr;   // down only: the device block is complete once this has fired
  cudaStream_t stream = nullptr;
  int dev = 0;
  std::vector<Chunk> chunks;
  size_t next = 0;               // first chunk nobody has taken yet
  size_t remaining = 0;          // chunks not yet finished
  std::atomic<bool> is_ready{false};  // down only
  double t_submit = 0.0, t_ready = 0.0, t_first = 0.0;   // trace
};

// DGEMM_B200_HOST_TRACE: when each block of the way back became ready / was taken first / was retired, and when each
// block of the way up was retired (steady-clock milliseconds; staging_trace_dump prints and clears them)
bool g_trace = false;
std::mutex g_trace_mu;
struct TraceRow { char dir; long long bytes; double t_submit, t_ready, t_first, t_done; };
std::vector<TraceRow> g_trace_rows;
double trace_now() {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

std::atomic<int> g_err{0};
void note(cudaError_t e) {
  if (e != cudaSuccess) {
    int expected = 0;
    g_err.compare_exchange_strong(expected, (int)e);
  }
}

// FIFO of jobs handed out chunk by chunk to a fixed team of threads; jobs retire in submission order
class Engine {
 public:
  Engine(int n, int dev, bool up) : n_(n), dev_(dev), up_(up), slots_(2 * n) {}
  ~Engine() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    work_cv_.notify_all();
    for (auto& t : threads_) t.join();
    for (Slot& s : slots_) {
      if (s.host) cudaFreeHost(s.host);
      if (s.done) cudaEventDestroy(s.done);
    }
    if (sig_) cudaStreamDestroy(sig_);
  }
  int dev() const { return dev_; }
  cudaError_t start() {
    for (Slot& s : slots_) {
      if (s.host) continue;
      cudaError_t e = cudaHostAlloc((void**)&s.host, SLOT_BYTES, cudaHostAllocPortable);
      if (e != cudaSuccess) return e;
      if ((e = cudaEventCreateWithFlags(&s.done, cudaEventDisableTiming)) != cudaSuccess) return e;
    }
    if (threads_.empty())
      for (int w = 0; w < n_; w++) threads_.emplace_back([this, w] { up_ ? loop_up(w) : loop_down(w); });
    return cudaSuccess;
  }
  long long submit(std::unique_ptr<Job> j) {
    j->remaining = j->chunks.size();
    if (g_trace) j->t_submit = trace_now();
    std::lock_guard<std::mutex> lk(mu_);
    live_.push_back(std::move(j));
    const long long ticket = base_ticket_ + (long long)live_.size() - 1;
    work_cv_.notify_all();
    return ticket;
  }
  void wait_retired(long long ticket) {
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [&] { return base_ticket_ > ticket; });
  }
  void mark_ready(Job* j) {   // (called from a CUDA host function: no CUDA calls in here)
    std::lock_guard<std::mutex> lk(mu_);
    j->is_ready.store(true, std::memory_order_release);
    if (g_trace) j->t_ready = trace_now();
    work_cv_.notify_all();
  }
  cudaStream_t signal_stream() {
    if (!sig_) cudaStreamCreateWithFlags(&sig_, cudaStreamNonBlocking);
    return sig_;
  }
  void wait_idle() {
    std::unique_lock<std::mutex> lk(mu_);
    done_cv_.wait(lk, [&] { return live_.empty(); });
  }

 private:
  // next chunk in FIFO order; block == false returns at once when there is none.  Way back: a block can only be taken
  // once it is complete on the device (Job::is_ready, set by a host function queued behind the block's last kernel), so no
  // thread ever waits for the GPU while it still holds a chunk it could copy out, and nobody spins.
  bool take(bool block, Job** job, size_t* chunk) {
    std::unique_lock<std::mutex> lk(mu_);
    auto avail = [&] { return issue_ < live_.size() && (up_ || live_[issue_]->is_ready.load(std::memory_order_acquire)); };
    if (block) work_cv_.wait(lk, [&] { return stop_ || avail(); });
    if (!avail()) return false;
    Job* j = live_[issue_].get();
    *job = j;
    if (g_trace && j->next == 0) j->t_first = trace_now();
    *chunk = j->next++;
    if (j->next == j->chunks.size()) issue_++;
    return true;
  }
  void finish(Job* j) {
    std::lock_guard<std::mutex> lk(mu_);
    if (--j->remaining != 0) return;
    while (!live_.empty() && live_.front()->remaining == 0) {   // (a retiring job is fully issued: it sits before issue_)
      if (g_trace) {
        const Job& f = *live_.front();
        std::lock_guard<std::mutex> tl(g_trace_mu);
        g_trace_rows.push_back({up_ ? 'u' : 'd', f.rows * f.cols * 8, f.t_submit, f.t_ready, f.t_first, trace_now()});
      }
      live_.pop_front();
      base_ticket_++;
      issue_--;
    }
    done_cv_.notify_all();
  }
  bool stopping() {
    std::lock_guard<std::mutex> lk(mu_);
    return stop_;
  }

  void loop_up(int w) {
    int turn = 0;
    for (;;) {
      Job* j;
      size_t ci;
      if (!take(true, &j, &ci)) {
        if (stopping()) return;
        continue;
      }
      cudaSetDevice(j->dev);
      const Chunk& c = j->chunks[ci];
      if (j->rows <= 0 || j->cols <= 0) {
        // (an empty block: only there to keep the tickets dense)
      } else if (!j->staged) {
        note(cudaMemcpy2DAsync(j->dev_ptr, (size_t)j->ldd * 8, j->host_ptr, (size_t)j->ldh * 8, (size_t)j->rows * 8, (size_t)j->cols,
                               cudaMemcpyHostToDevice, j->stream));
      } else {
        Slot& s = slots_[2 * w + (turn++ & 1)];
        if (s.used) note(cudaEventSynchronize(s.done));  // the slot's previous DMA has read it
        const double* src = j->host_ptr;
        if (j->ldh == c.rows && c.row0 == 0)
          memcpy(s.host, src + c.col0 * j->ldh, (size_t)c.rows * c.cols * 8);
        else
          for (long long cc = 0; cc < c.cols; cc++)
            memcpy(s.host + (size_t)cc * c.rows * 8, src + (c.col0 + cc) * j->ldh + c.row0, (size_t)c.rows * 8);
        if (j->ldd == c.rows && c.row0 == 0)
          note(cudaMemcpyAsync(j->dev_ptr + c.col0 * j->ldd, s.host, (size_t)c.rows * c.cols * 8, cudaMemcpyHostToDevice, j->stream));
        else
          note(cudaMemcpy2DAsync(j->dev_ptr + c.col0 * j->ldd + c.row0, (size_t)j->ldd * 8, s.host, (size_t)c.rows * 8,
                                 (size_t)c.rows * 8, (size_t)c.cols, cudaMemcpyHostToDevice, j->stream));
        note(cudaEventRecord(s.done, j->stream));
        s.used = true;
      }
      finish(j);
    }
  }

  void loop_down(int w) {
    struct Pending { Job* j; size_t ci; int slot; };
    std::deque<Pending> pending;   // DMAs requested, not yet copied out (at most two: one per slot)
    int turn = 0;
    for (;;) {
      while (pending.size() < 2) {
        Job* j;
        size_t ci;
        if (!take(pending.empty(), &j, &ci)) break;
        cudaSetDevice(j->dev);
        const Chunk& c = j->chunks[ci];
        const int k = 2 * w + (turn++ & 1);   // (the other slot is the pending one, or free)
        Slot& s = slots_[k];
        note(cudaMemcpy2DAsync(s.host, (size_t)c.rows * 8, j->dev_ptr + c.col0 * j->ldd + c.row0, (size_t)j->ldd * 8,
                               (size_t)c.rows * 8, (size_t)c.cols, cudaMemcpyDeviceToHost, j->stream));
        note(cudaEventRecord(s.done, j->stream));
        pending.push_back({j, ci, k});
      }
      if (pending.empty()) {
        if (stopping()) return;
        continue;
      }
      const Pending p = pending.front();
      pending.pop_front();
      Slot& s = slots_[p.slot];
      note(cudaEventSynchronize(s.done));
      const Chunk& c = p.j->chunks[p.ci];
      double* dst = p.j->host_ptr;
      if (p.j->ldh == c.rows && c.row0 == 0)
        memcpy(dst + c.col0 * p.j->ldh, s.host, (size_t)c.rows * c.cols * 8);
      else
        for (long long cc = 0; cc < c.cols; cc++)
          memcpy(dst + (c.col0 + cc) * p.j->ldh + c.row0, s.host + (size_t)cc * c.rows * 8, (size_t)c.rows * 8);
      finish(p.j);
    }
  }

  const int n_, dev_;
  const bool up_;
  std::vector<Slot> slots_;   // two per thread
  std::vector<std::thread> threads_;
  std::mutex mu_;
  std::condition_variable work_cv_, done_cv_;
  std::deque<std::unique_ptr<Job>> live_;   // submitted, not yet retired, in ticket order
  long long base_ticket_ = 0;               // ticket of live_.front()
  size_t issue_ = 0;                        // index in live_ of the first job with chunks nobody has taken
  bool stop_ = false;
  cudaStream_t sig_ = nullptr;              // way back: carries the "block is complete" host functions
};

struct State {
  std::mutex mu;
  Engine* up = nullptr;
  Engine* down = nullptr;
};
State& state() {
  static State* s = new State;  // leaked on purpose: worker threads may outlive static destruction order
  return *s;
}

int stage_threads() {
  if (const char* e = getenv("DGEMM_B200_STAGE_THREADS")) return std::max(1, std::min(32, atoi(e)));
  // tools/stage_probe.cu on this pool's 16-vCPU hosts: the fill -> DMA pipeline sustains 24 / 47 / 50.6 / 50.1 / 49 GB/s with
  // 2 / 4 / 6 / 8 / 12 threads (4 MB slots, two per thread); the main thread and the four threads of the way back
  // need cores too
  const unsigned hc = std::thread::hardware_concurrency();
  return (int)std::max(2u, std::min(8u, hc / 2));
}

cudaError_t ensure_started(int dev) {
  State& st = state();
  std::lock_guard<std::mutex> lk(st.mu);
  if (!st.up) {
    g_trace = getenv("DGEMM_B200_HOST_TRACE") != nullptr;
    if (const char* e = getenv("DGEMM_B200_STAGE_CHUNK_KB"))
      g_chunk_bytes = std::max<size_t>(64u << 10, std::min<size_t>(SLOT_BYTES, (size_t)atoll(e) << 10));
    const int n = stage_threads();
    st.up = new Engine(n, dev, true);
    st.down = new Engine(n, dev, false);   // (asleep unless a finished block is waiting; see Engine::take)
  }
  cudaError_t e = st.up->start();
  if (e == cudaSuccess) e = st.down->start();
  return e;
}

cudaError_t take_error() { return (cudaError_t)g_err.exchange(0); }

}  // namespace

// the slots' events belong to the device of the first staged call; calls on another device use the driver's path
bool staging_usable(int dev) {
  State& st = state();
  std::lock_guard<std::mutex> lk(st.mu);
  return !st.up || st.up->dev() == dev;
}

bool host_pointer_is_pageable(const void* p) {
  if (!p) return false;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return true;
  }
  return a.type == cudaMemoryTypeUnregistered;
}

// Queues one block for the team (host -> device) and returns its ticket at once (< 0: error, see *err).  pageable: the
// block goes through the pinned slots; otherwise (pinned caller memory) it is one direct copy, kept in the same queue so
// that the copy stream sees all blocks in submission order.
long long staging_submit_h2d(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                             cudaStream_t stream, int dev, bool pageable, cudaError_t* err) {
  *err = ensure_started(dev);
  if (*err != cudaSuccess) return -1;
  std::unique_ptr<Job> j(new Job);
  j->dev_ptr = dst; j->ldd = ldd; j->host_ptr = const_cast<double*>(src); j->ldh = lds; j->rows = rows; j->cols = cols;
  j->staged = pageable; j->stream = stream; j->dev = dev;
  if (rows > 0 && cols > 0) {
    if (pageable) j->chunks = make_chunks(rows, cols, CHUNK_BYTES);
    else j->chunks.push_back({0, rows, 0, cols});
  }
  if (j->chunks.empty()) j->chunks.push_back({0, 0, 0, 0}), j->staged = false, j->rows = j->cols = 0;   // keeps tickets dense
  return state().up->submit(std::move(j));
}

// returns once every chunk of every block up to `ticket` has been queued on its stream
cudaError_t staging_wait_submitted(long long ticket) {
  State& st = state();
  if (!st.up || ticket < 0) return cudaSuccess;
  st.up->wait_retired(ticket);
  return take_error();
}

// Blocking form: returns once the DMA of every chunk has been queued on `stream` (so an event recorded on `stream` after
// this call covers the whole block).  `src` is pageable host memory, `dst` device memory.
cudaError_t staged_h2d(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                       cudaStream_t stream, int dev) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  cudaError_t e;
  const long long t = staging_submit_h2d(dst, ldd, src, lds, rows, cols, stream, dev, true, &e);
  if (t < 0) return e;
  return staging_wait_submitted(t);
}

// Asynchronous: once `ready` has fired, the block is copied device -> pinned slots -> `dst` (pageable) by the down team.
// staged_d2h_wait() returns when every queued block has reached the caller's memory.
cudaError_t staged_d2h_async(double* dst, long long ldd, const double* src, long long lds, long long rows, long long cols,
                             cudaEvent_t ready, cudaStream_t stream, int dev) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  cudaError_t e = ensure_started(dev);
  if (e != cudaSuccess) return e;
  std::unique_ptr<Job> j(new Job);
  j->dev_ptr = const_cast<double*>(src); j->ldd = lds; j->host_ptr = dst; j->ldh = ldd; j->rows = rows; j->cols = cols;
  j->ready = ready; j->stream = stream; j->dev = dev;
  // (1 MB chunks on the way back: a C panel of a mid-size call is a few MB, and all threads of the team should get a part)
  j->chunks = make_chunks(rows, cols, std::min<size_t>(CHUNK_BYTES, 1u << 20));
  Engine* down = state().down;
  Job* raw = j.get();
  down->submit(std::move(j));   // (alive until its last chunk has been copied out, which needs is_ready first)
  struct Signal { Engine* e; Job* j; };
  Signal* sg = new Signal{down, raw};
  cudaStream_t sig = down->signal_stream();
  e = cudaStreamWaitEvent(sig, ready, 0);
  if (e == cudaSuccess)
    e = cudaLaunchHostFunc(sig, [](void* p) {
      Signal* s = static_cast<Signal*>(p);
      s->e->mark_ready(s->j);
      delete s;
    }, sg);
  if (e != cudaSuccess) {   // cannot be signalled: release the block now, the DMA itself is ordered behind `ready` below
    cudaStreamWaitEvent(stream, ready, 0);
    down->mark_ready(raw);
    delete sg;
  }
  return cudaSuccess;
}

cudaError_t staged_d2h_wait() {
  State& st = state();
  if (!st.down) return cudaSuccess;
  st.down->wait_idle();
  return take_error();
}

// DGEMM_B200_HOST_TRACE: one line per block since the last dump, times in ms relative to t0_ms (a steady-clock reading)
void staging_trace_dump(double t0_ms) {
  if (!g_trace) return;
  std::lock_guard<std::mutex> tl(g_trace_mu);
  for (const TraceRow& r : g_trace_rows)
    fprintf(stderr, "{\"staging_block\": \"%s\", \"mb\": %.2f, \"submitted\": %.3f, \"ready\": %.3f, \"first_chunk_taken\": %.3f, \"retired\": %.3f}\n",
            r.dir == 'u' ? "up" : "down", r.bytes / 1048576.0, r.t_submit - t0_ms, r.t_ready > 0 ? r.t_ready - t0_ms : -1.0,
            r.t_first - t0_ms, r.t_done - t0_ms);
  g_trace_rows.clear();
}

// (error paths: nothing of the caller's memory may still be read by the up team when the call returns)
void staging_drain() {
  State& st = state();
  if (st.up) st.up->wait_idle();
  if (st.down) st.down->wait_idle();
}

void staging_shutdown() {
  State& st = state();
  std::lock_guard<std::mutex> lk(st.mu);
  delete st.up;
  delete st.down;
  st.up = st.down = nullptr;
}

}  // namespace dg
