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
// device and handed out again by size.
//
// Ordering and ownership (round 3).  Every block remembers the device it was allocated on.  A block is cached under
// THAT device -- never under whatever device happens to be current -- and a free that arrives while another device is
// current does not cache at all (cudaFree, which needs no stream).  When a block is freed an event is recorded on the
// stream the library issues on at that moment (the point cudaFreeAsync(p, S) would have ordered the free at) and kept
// with the block; whoever takes the block out again makes its own stream wait for that event.  So reuse is ordered
// behind the previous owner's work whichever stream asks, and a stream handle that is destroyed and whose value the
// driver hands out again cannot inherit blocks without an ordering (the cache used to be keyed by the raw handle).
// ---------------------------------------------------------------------------------------------------------------------
struct BlockInfo {
  size_t bytes;
  int dev;  // device the block was allocated on
};
struct CachedBlock {
  void* p;
  cudaEvent_t freed;  // recorded behind the last owner's work
};
// (heap objects that are never destroyed: buffers owned by static objects of other translation units are released
// during process exit, after this unit's own statics would be gone)
static std::mutex& g_cache_mu = *new std::mutex;
static std::unordered_map<void*, BlockInfo>& g_block_bytes = *new std::unordered_map<void*, BlockInfo>;   // live or cached block -> size, device
static std::map<int, std::multimap<size_t, CachedBlock>>& g_cached = *new std::map<int, std::multimap<size_t, CachedBlock>>;  // device -> free blocks by size
static std::map<int, std::vector<cudaEvent_t>>& g_event_pool = *new std::map<int, std::vector<cudaEvent_t>>;  // device -> idle events
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

// an event of device `dev` (which must be current); nullptr if the runtime refuses (e.g. during process exit)
static cudaEvent_t take_event_locked(int dev) {
  auto& pool = g_event_pool[dev];
  if (!pool.empty()) {
    cudaEvent_t e = pool.back();
    pool.pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return e;
}

static void release_cached_locked(size_t keep_bytes) {
  // give the largest blocks back to the driver first, until at most keep_bytes stay cached
  int cur = -1;
  if (cudaGetDevice(&cur) != cudaSuccess) { cudaGetLastError(); cur = -1; }
  while (g_cached_bytes > keep_bytes) {
    std::multimap<size_t, CachedBlock>* best = nullptr;
    int bdev = -1;
    for (auto& kv : g_cached)
      if (!kv.second.empty() && (!best || kv.second.rbegin()->first > best->rbegin()->first)) { best = &kv.second; bdev = kv.first; }
    if (!best) break;
    auto it = std::prev(best->end());
    const CachedBlock b = it->second;
    g_cached_bytes -= it->first;
    g_block_bytes.erase(b.p);
    best->erase(it);
    if (bdev == cur && cudaStreamWaitEvent(g_stream, b.freed, 0) == cudaSuccess) {
      // a block of the current device goes back in stream order, behind the work its last owner had queued
      cudaFreeAsync(b.p, g_stream);
      g_event_pool[bdev].push_back(b.freed);
    } else {
      // another device is current (or the stream refused the wait): cudaFree needs no stream and waits by itself
      cudaGetLastError();
      cudaFree(b.p);
      cudaEventDestroy(b.freed);
    }
    cudaGetLastError();
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
    std::unique_lock<std::mutex> lk(g_cache_mu);
    auto f = g_cached.find(dev);
    if (f != g_cached.end()) {
      auto it = f->second.lower_bound(want);
      // a larger block will do if it wastes little (blocks of one size class differ by < 12.5 %; small ones may take 2x)
      if (it != f->second.end() && it->first <= want + (want >= ((size_t)64 << 20) ? want / 8 : want)) {
        const CachedBlock b = it->second;
        g_cached_bytes -= it->first;
        f->second.erase(it);
        // the new owner's stream waits for the previous owner's work (a no-op on the device when it is the same stream);
        // the wait is taken at this moment, so the event can be recorded again by the next free straight away
        const cudaError_t we = cudaStreamWaitEvent(g_stream, b.freed, 0);
        g_event_pool[dev].push_back(b.freed);
        lk.unlock();
        if (we != cudaSuccess) {
          cudaGetLastError();
          FRICC_CUDA(cudaDeviceSynchronize());  // no stream-side ordering to be had: order by waiting
        }
        g_alloc_hits.fetch_add(1, std::memory_order_relaxed);
        return b.p;
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
  g_block_bytes[p] = BlockInfo{want, dev};
  return p;
}

void dev_free(void* p) {
  if (!p) return;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); dev = -1; }
  std::lock_guard<std::mutex> lk(g_cache_mu);
  auto it = g_block_bytes.find(p);
  const bool known = it != g_block_bytes.end();
  const size_t bytes = known ? it->second.bytes : 0;
  const int owner = known ? it->second.dev : dev;
  if (owner != dev) {
    // freed while another device is current: g_stream belongs to that other device, so neither a stream-ordered free nor
    // an event on it can order anything for this block
    if (known) g_block_bytes.erase(it);
    SlowScope slow("cudaFree (block of another device)", __FILE__, __LINE__);
    cudaFree(p);
    cudaGetLastError();
    return;
  }
  cudaEvent_t ev = nullptr;
  const bool cacheable = bytes != 0 && bytes <= g_cache_max_block && g_cache_cap != 0;
  if (cacheable) {
    ev = take_event_locked(dev);
    if (ev && cudaEventRecord(ev, g_stream) != cudaSuccess) {
      cudaGetLastError();
      g_event_pool[dev].push_back(ev);
      ev = nullptr;
    }
  }
  if (!ev) {
    if (known) g_block_bytes.erase(it);
    SlowScope slow("cudaFreeAsync", __FILE__, __LINE__);
    cudaFreeAsync(p, g_stream);
    return;
  }
  g_cached[dev].emplace(bytes, CachedBlock{p, ev});
  g_cached_bytes += bytes;
  if (g_cached_bytes > g_cache_cap) release_cached_locked(g_cache_cap - g_cache_cap / 4);
}

}  // namespace fricc
