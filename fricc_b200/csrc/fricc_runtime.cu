// fricc_runtime.cu -- error state, stream, stream-ordered device allocations, launch accounting.
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <iterator>
#include <map>
#include <mutex>
#include <unordered_map>

#include "fricc_common.cuh"

namespace fricc {

static thread_local std::string g_last_error;
static cudaStream_t g_stream = nullptr;
static std::atomic<int64_t> g_launches{0};
static int g_num_sms = 0;
static std::once_flag g_pool_once;

void set_last_error(const std::string& msg) { g_last_error = msg; }
const std::string& last_error() { return g_last_error; }

static double g_slow_ms = -2.0;  // -2: environment not read yet
double slow_trace_ms() {
  if (g_slow_ms == -2.0) {
    const char* e = getenv("FRICC_TRACE_SLOW_MS");
    g_slow_ms = e ? atof(e) : -1.0;
  }
  return g_slow_ms;
}
static inline long long now_ns() {
  return std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
SlowScope::SlowScope(const char* w, const char* f, int l) : what(w), file(f), line(l), t0(0) {
  if (slow_trace_ms() >= 0.0) t0 = now_ns();
}
SlowScope::~SlowScope() {
  if (t0) {
    const double ms = (double)(now_ns() - t0) * 1e-6;
    if (ms > g_slow_ms) fprintf(stderr, "[fricc slow] %.1f ms in %s (%s:%d)\n", ms, what, file, line);
  }
}

void fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  throw Error(code, buf);
}

static int g_uniform_carveout = -1;
int uniform_carveout() { return g_uniform_carveout; }
void set_uniform_carveout(int pct) { g_uniform_carveout = pct < 0 ? -1 : (pct > 100 ? 100 : pct); }
cudaStream_t stream() { return g_stream; }
void set_stream(cudaStream_t s) { g_stream = s; }
void count_launch(int64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
int64_t launches() { return g_launches.load(std::memory_order_relaxed); }

void check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) fail(FRICC_ECUDA, "launch of %s failed: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  count_launch(1);
}

int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    FRICC_CUDA(cudaGetDevice(&dev));
    FRICC_CUDA(cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev));
  }
  return g_num_sms;
}
void reset_device_cache() { g_num_sms = 0; }

static void init_pool() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return;
  cudaMemPool_t pool;
  if (cudaDeviceGetDefaultMemPool(&pool, dev) != cudaSuccess) return;
  uint64_t thresh = UINT64_MAX;  // keep freed blocks cached in the pool
  cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thresh);
}

// ---------------------------------------------------------------------------------------------------------------------
// Block cache in front of cudaMallocAsync.
//
// Why: a plugin call allocates its scratch (sign matrix, CSR remainder, sort buffers ...) per call.  The driver's pool
// recycles freed memory, but whenever it has to map more (its free space is fragmented differently from step to step)
// the call goes into the kernel driver, and there it queues behind whoever else holds the driver's lock -- e.g. an
// `nvidia-smi -lms` clock sampler.  Measured on B200 with the tracer above: cudaMallocAsync calls of 30 ms ... 2.7 s
// inside fricc_contract, a few per hundred steps, each one long enough to drain the device (the host runs only ~300 ms
// ahead).  With the cache a steady-state step makes NO allocation call at all: freed blocks are kept per
// (device, stream) and handed out again by size.
//
// Ordering: a block freed while the library issues on stream S is only reused by allocations made while it issues on S,
// so every use of the block is ordered behind the previous owner's work exactly as with cudaFreeAsync(p, S).
// ---------------------------------------------------------------------------------------------------------------------
struct CacheKey {
  int dev;
  cudaStream_t stream;
  bool operator<(const CacheKey& o) const { return dev != o.dev ? dev < o.dev : stream < o.stream; }
};
// (heap objects that are never destroyed: buffers owned by static objects of other translation units are released
// during process exit, after this unit's own statics would be gone)
static std::mutex& g_cache_mu = *new std::mutex;
static std::unordered_map<void*, size_t>& g_block_bytes = *new std::unordered_map<void*, size_t>;   // live or cached block -> size
static std::map<CacheKey, std::multimap<size_t, void*>>& g_cached = *new std::map<CacheKey, std::multimap<size_t, void*>>;  // free blocks by size
static size_t g_cached_bytes = 0;
static size_t g_cache_cap = (size_t)24 << 30;       // "alloc_cache_mb": total bytes kept (0 = no caching)
static size_t g_cache_max_block = (size_t)2 << 30;  // larger blocks (digit planes of a ladder slice) go back to the driver
static std::atomic<int64_t> g_alloc_calls{0}, g_alloc_hits{0};

int64_t alloc_driver_calls() { return g_alloc_calls.load(std::memory_order_relaxed); }
int64_t alloc_cache_hits() { return g_alloc_hits.load(std::memory_order_relaxed); }

// sizes are rounded to 4 significant bits (<= 12.5 % slack) so that data-dependent sizes fall into few classes
static size_t round_block(size_t bytes) {
  if (bytes <= 4096) return (bytes + 511) & ~(size_t)511;
  int k = 63 - __builtin_clzll((unsigned long long)bytes);
  const size_t step = (size_t)1 << (k - 3);
  return (bytes + step - 1) & ~(step - 1);
}

static void release_cached_locked(size_t keep_bytes) {
  // give the largest blocks back to the driver first, until at most keep_bytes stay cached
  while (g_cached_bytes > keep_bytes) {
    std::multimap<size_t, void*>* best = nullptr;
    cudaStream_t st = nullptr;
    for (auto& kv : g_cached)
      if (!kv.second.empty() && (!best || kv.second.rbegin()->first > best->rbegin()->first)) { best = &kv.second; st = kv.first.stream; }
    if (!best) break;
    auto it = std::prev(best->end());
    g_cached_bytes -= it->first;
    g_block_bytes.erase(it->second);
    // a block of the stream the library is issuing on goes back in stream order; one cached under another stream (which
    // may not even exist any more) is not ordered against that stream's work by a free on this one: cudaFree waits
    if (st == g_stream) cudaFreeAsync(it->second, st);
    else cudaFree(it->second);
    best->erase(it);
  }
}

void set_alloc_cache_mb(int mb) {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  g_cache_cap = mb <= 0 ? 0 : (size_t)mb << 20;
  release_cached_locked(g_cache_cap);
}

void release_cached_blocks() {
  std::lock_guard<std::mutex> lk(g_cache_mu);
  release_cached_locked(0);
}

void* dev_alloc(size_t bytes) {
  std::call_once(g_pool_once, init_pool);
  const size_t want = round_block(bytes ? bytes : 1);
  int dev = 0;
  cudaGetDevice(&dev);
  {
    std::lock_guard<std::mutex> lk(g_cache_mu);
    auto f = g_cached.find(CacheKey{dev, g_stream});
    if (f != g_cached.end()) {
      auto it = f->second.lower_bound(want);
      // a larger block will do if it wastes little (blocks of one size class differ by < 12.5 %; small ones may take 2x)
      if (it != f->second.end() && it->first <= want + (want >= ((size_t)64 << 20) ? want / 8 : want)) {
        void* p = it->second;
        g_cached_bytes -= it->first;
        f->second.erase(it);
        g_alloc_hits.fetch_add(1, std::memory_order_relaxed);
        return p;
      }
    }
  }
  void* p = nullptr;
  cudaError_t e;
  {
    SlowScope slow("cudaMallocAsync", __FILE__, __LINE__);
    e = cudaMallocAsync(&p, want, g_stream);
  }
  if (e != cudaSuccess) {  // out of memory: everything the cache holds goes back first
    cudaGetLastError();
    release_cached_blocks();
    cudaStreamSynchronize(g_stream);
    e = cudaMallocAsync(&p, want, g_stream);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    fail(FRICC_ECUDA, "cudaMallocAsync(%zu bytes) failed: %s", want, cudaGetErrorString(e));
  }
  g_alloc_calls.fetch_add(1, std::memory_order_relaxed);
  std::lock_guard<std::mutex> lk(g_cache_mu);
  g_block_bytes[p] = want;
  return p;
}

void dev_free(void* p) {
  if (!p) return;
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lk(g_cache_mu);
  auto it = g_block_bytes.find(p);
  const size_t bytes = it == g_block_bytes.end() ? 0 : it->second;
  if (bytes == 0 || bytes > g_cache_max_block || g_cache_cap == 0) {
    if (it != g_block_bytes.end()) g_block_bytes.erase(it);
    SlowScope slow("cudaFreeAsync", __FILE__, __LINE__);
    cudaFreeAsync(p, g_stream);
    return;
  }
  g_cached[CacheKey{dev, g_stream}].emplace(bytes, p);
  g_cached_bytes += bytes;
  if (g_cached_bytes > g_cache_cap) release_cached_locked(g_cache_cap - g_cache_cap / 4);
}

}  // namespace fricc
