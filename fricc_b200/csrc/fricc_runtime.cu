// fricc_runtime.cu -- error state, stream, stream-ordered device allocations, launch accounting.
#include <atomic>
#include <mutex>

#include "fricc_common.cuh"

namespace fricc {

static thread_local std::string g_last_error;
static cudaStream_t g_stream = nullptr;
static std::atomic<int64_t> g_launches{0};
static int g_num_sms = 0;
static std::once_flag g_pool_once;

void set_last_error(const std::string& msg) { g_last_error = msg; }
const std::string& last_error() { return g_last_error; }

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

void* dev_alloc(size_t bytes) {
  std::call_once(g_pool_once, init_pool);
  void* p = nullptr;
  cudaError_t e = cudaMallocAsync(&p, bytes, g_stream);
  if (e != cudaSuccess) {
    cudaGetLastError();
    fail(FRICC_ECUDA, "cudaMallocAsync(%zu bytes) failed: %s", bytes, cudaGetErrorString(e));
  }
  return p;
}

void dev_free(void* p) {
  if (p) cudaFreeAsync(p, g_stream);
}

}  // namespace fricc
