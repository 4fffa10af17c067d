// stage_probe.cu — what limits pageable -> HBM staging on this host?  (tools/, not part of the library)
//   1. host-only: T threads copy a pageable matrix into their pinned slots (no DMA), per copy method and slot size
//   2. the whole staged pipeline (fill slot -> DMA -> event) per method / threads / slot size / slots per thread
//   3. references: DMA from pinned memory alone, cudaMemcpy from pageable memory (the driver's own staging)
// One JSON line per measurement.   build: make stage_probe    run: .../bin/stage_probe [GB]
#include <cuda_runtime.h>
#include <immintrin.h>

#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

static double now() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

__attribute__((target("avx512f"))) static void copy_nt512(char* dst, const char* src, size_t n) {
  size_t i = 0;
  for (; i + 256 <= n; i += 256) {
    __m512i a = _mm512_loadu_si512((const void*)(src + i)), b = _mm512_loadu_si512((const void*)(src + i + 64));
    __m512i c = _mm512_loadu_si512((const void*)(src + i + 128)), d = _mm512_loadu_si512((const void*)(src + i + 192));
    _mm512_stream_si512((__m512i*)(dst + i), a);
    _mm512_stream_si512((__m512i*)(dst + i + 64), b);
    _mm512_stream_si512((__m512i*)(dst + i + 128), c);
    _mm512_stream_si512((__m512i*)(dst + i + 192), d);
  }
  if (i < n) memcpy(dst + i, src + i, n - i);
  _mm_sfence();
}
__attribute__((target("avx512f"))) static void copy_nt512_prefetch(char* dst, const char* src, size_t n) {
  size_t i = 0;
  for (; i + 256 <= n; i += 256) {
    _mm_prefetch(src + i + 2048, _MM_HINT_NTA);
    _mm_prefetch(src + i + 2048 + 64, _MM_HINT_NTA);
    _mm_prefetch(src + i + 2048 + 128, _MM_HINT_NTA);
    _mm_prefetch(src + i + 2048 + 192, _MM_HINT_NTA);
    __m512i a = _mm512_loadu_si512((const void*)(src + i)), b = _mm512_loadu_si512((const void*)(src + i + 64));
    __m512i c = _mm512_loadu_si512((const void*)(src + i + 128)), d = _mm512_loadu_si512((const void*)(src + i + 192));
    _mm512_stream_si512((__m512i*)(dst + i), a);
    _mm512_stream_si512((__m512i*)(dst + i + 64), b);
    _mm512_stream_si512((__m512i*)(dst + i + 128), c);
    _mm512_stream_si512((__m512i*)(dst + i + 192), d);
  }
  if (i < n) memcpy(dst + i, src + i, n - i);
  _mm_sfence();
}
static void copy_movsb(char* dst, const char* src, size_t n) {
  asm volatile("rep movsb" : "+D"(dst), "+S"(src), "+c"(n) : : "memory");
}
static void copy_glibc(char* dst, const char* src, size_t n) { memcpy(dst, src, n); }

typedef void (*copy_fn)(char*, const char*, size_t);
struct Method { const char* name; copy_fn fn; };

int main(int argc, char** argv) {
  const double gb = argc > 1 ? atof(argv[1]) : 2.0;
  const size_t total = (size_t)(gb * (1u << 30)) & ~((size_t)(64u << 20) - 1);
  const bool avx512 = __builtin_cpu_supports("avx512f");
  std::vector<Method> methods = {{"glibc memcpy", copy_glibc}, {"rep movsb", copy_movsb}};
  if (avx512) { methods.push_back({"avx512 non-temporal", copy_nt512}); methods.push_back({"avx512 non-temporal + prefetchnta", copy_nt512_prefetch}); }
  char* src = (char*)malloc(total);
  if (!src) return 1;
  {  // first touch by several threads
    std::vector<std::thread> th;
    for (int t = 0; t < 8; t++) th.emplace_back([&, t] { memset(src + total / 8 * t, t + 1, total / 8); });
    for (auto& x : th) x.join();
  }
  char* dev = nullptr;
  if (cudaMalloc(&dev, total) != cudaSuccess) { printf("{\"error\": \"cudaMalloc\"}\n"); return 1; }
  const unsigned hc = std::thread::hardware_concurrency();
  printf("{\"host_threads\": %u, \"avx512f\": %s, \"bytes\": %zu}\n", hc, avx512 ? "true" : "false", total);
  // ---- references
  {
    char* pin = nullptr;
    cudaHostAlloc((void**)&pin, total, cudaHostAllocPortable);
    memcpy(pin, src, total);
    for (int rep = 0; rep < 2; rep++) {
      double t0 = now();
      cudaMemcpy(dev, pin, total, cudaMemcpyHostToDevice);
      double dt = now() - t0;
      if (rep) printf("{\"what\": \"DMA from pinned memory\", \"gb_per_s\": %.1f}\n", total / dt / 1e9);
    }
    for (int rep = 0; rep < 2; rep++) {
      double t0 = now();
      cudaMemcpy(dev, src, total, cudaMemcpyHostToDevice);
      double dt = now() - t0;
      if (rep) printf("{\"what\": \"cudaMemcpy from pageable memory (driver staging)\", \"gb_per_s\": %.1f}\n", total / dt / 1e9);
    }
    cudaFreeHost(pin);
  }
  const int max_threads = 16, max_slots = 4;
  const size_t max_slot = 16u << 20;
  std::vector<char*> slots(max_threads * max_slots);
  std::vector<cudaEvent_t> ev(max_threads * max_slots);
  for (size_t i = 0; i < slots.size(); i++) {
    if (cudaHostAlloc((void**)&slots[i], max_slot, cudaHostAllocPortable) != cudaSuccess) { printf("{\"error\": \"cudaHostAlloc\"}\n"); return 1; }
    memset(slots[i], 0, max_slot);
    cudaEventCreateWithFlags(&ev[i], cudaEventDisableTiming);
  }
  cudaStream_t stream;
  cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking);
  // ---- 1. host-only fill rate
  for (const Method& m : methods)
    for (int T : {1, 2, 4, 8, 16}) {
      if ((unsigned)T > hc) continue;
      for (size_t S : {(size_t)1 << 20, (size_t)4 << 20, (size_t)16 << 20}) {
        const size_t chunks = total / S;
        std::atomic<size_t> next{0};
        double t0 = now();
        std::vector<std::thread> th;
        for (int t = 0; t < T; t++)
          th.emplace_back([&, t] {
            int turn = 0;
            for (;;) {
              const size_t i = next.fetch_add(1);
              if (i >= chunks) break;
              m.fn(slots[t * max_slots + (turn++ & 1)], src + i * S, S);
            }
          });
        for (auto& x : th) x.join();
        double dt = now() - t0;
        printf("{\"what\": \"host copy only\", \"method\": \"%s\", \"threads\": %d, \"slot_mb\": %zu, \"gb_per_s\": %.1f}\n", m.name, T,
               S >> 20, total / dt / 1e9);
        fflush(stdout);
      }
    }
  // ---- 2. the staged pipeline
  for (const Method& m : methods)
    for (int T : {2, 4, 6, 8, 12, 16}) {
      if ((unsigned)T > hc) continue;
      for (size_t S : {(size_t)1 << 20, (size_t)4 << 20, (size_t)16 << 20})
        for (int NS : {2, 4}) {
          const size_t chunks = total / S;
          std::atomic<size_t> next{0};
          cudaDeviceSynchronize();
          double t0 = now();
          std::vector<std::thread> th;
          for (int t = 0; t < T; t++)
            th.emplace_back([&, t] {
              int turn = 0;
              std::vector<char> used(NS, 0);
              for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= chunks) break;
                const int k = turn++ % NS, id = t * max_slots + k;
                if (used[k]) cudaEventSynchronize(ev[id]);
                m.fn(slots[id], src + i * S, S);
                cudaMemcpyAsync(dev + i * S, slots[id], S, cudaMemcpyHostToDevice, stream);
                cudaEventRecord(ev[id], stream);
                used[k] = 1;
              }
            });
          for (auto& x : th) x.join();
          cudaStreamSynchronize(stream);
          double dt = now() - t0;
          printf("{\"what\": \"staged pipeline\", \"method\": \"%s\", \"threads\": %d, \"slot_mb\": %zu, \"slots_per_thread\": %d, \"gb_per_s\": %.1f}\n",
                 m.name, T, S >> 20, NS, total / dt / 1e9);
          fflush(stdout);
        }
    }
  // ---- 3. the way back: DMA into slots, threads copy out into pageable memory
  for (const Method& m : methods)
    for (int T : {2, 4, 8}) {
      if ((unsigned)T > hc) continue;
      if (m.fn == copy_nt512 || m.fn == copy_nt512_prefetch) continue;   // (their 64-byte stores need an aligned destination)
      const size_t S = 4u << 20, chunks = total / S;
      std::atomic<size_t> next{0};
      cudaDeviceSynchronize();
      double t0 = now();
      std::vector<std::thread> th;
      for (int t = 0; t < T; t++)
        th.emplace_back([&, t] {
          // two slots in flight per thread: request chunk i+1 before draining chunk i
          size_t mine[2];
          int have = 0, turn = 0;
          auto request = [&](int k) -> bool {
            const size_t i = next.fetch_add(1);
            if (i >= chunks) return false;
            mine[k] = i;
            cudaMemcpyAsync(slots[t * max_slots + k], dev + i * S, S, cudaMemcpyDeviceToHost, stream);
            cudaEventRecord(ev[t * max_slots + k], stream);
            return true;
          };
          bool live[2] = {request(0), request(1)};
          (void)have;
          while (live[0] || live[1]) {
            const int k = turn++ & 1;
            if (!live[k]) continue;
            cudaEventSynchronize(ev[t * max_slots + k]);
            m.fn(src + mine[k] * S, slots[t * max_slots + k], S);
            live[k] = request(k);
          }
        });
      for (auto& x : th) x.join();
      double dt = now() - t0;
      printf("{\"what\": \"staged way back (d2h)\", \"method\": \"%s\", \"threads\": %d, \"slot_mb\": 4, \"gb_per_s\": %.1f}\n", m.name, T,
             total / dt / 1e9);
      fflush(stdout);
    }
  return 0;
}
