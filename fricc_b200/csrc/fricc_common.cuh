// fricc_common.cuh -- shared plumbing for libfricc_b200 (error handling, stream, device buffers, key helpers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/fricc_b200.h"

namespace fricc {

// ---------------------------------------------------------------------------------------------------------------
// errors: C++ exceptions inside the library, converted to status codes at the C-ABI (see FRICC_API_BEGIN/END)
// ---------------------------------------------------------------------------------------------------------------
struct Error : public std::runtime_error {
  int code;
  Error(int c, const std::string& msg) : std::runtime_error(msg), code(c) {}
};

void set_last_error(const std::string& msg);
[[noreturn]] void fail(int code, const char* fmt, ...);

// Diagnostic: with FRICC_TRACE_SLOW_MS=<ms> in the environment every runtime call / launch / allocation that keeps the
// HOST longer than that is reported on stderr (finds what a stalled plugin call was waiting in).  Off: one load + branch.
double slow_trace_ms();  // < 0: off
struct SlowScope {
  const char* what;
  const char* file;
  int line;
  long long t0;
  SlowScope(const char* w, const char* f, int l);
  ~SlowScope();
};

#define FRICC_CUDA(expr)                                                                              \
  do {                                                                                                \
    ::fricc::SlowScope _slow(#expr, __FILE__, __LINE__);                                              \
    cudaError_t _e = (expr);                                                                          \
    if (_e != cudaSuccess)                                                                            \
      ::fricc::fail(FRICC_ECUDA, "CUDA error %s at %s:%d: %s", cudaGetErrorName(_e), __FILE__, __LINE__, \
                    cudaGetErrorString(_e));                                                          \
  } while (0)

#define FRICC_API_BEGIN try { ::fricc::SlowScope _slow_api(__func__, "C-ABI call, whole", 0);
#define FRICC_API_END                                  \
  }                                                    \
  catch (const ::fricc::Error& e) {                    \
    ::fricc::set_last_error(e.what());                 \
    return e.code;                                     \
  }                                                    \
  catch (const std::exception& e) {                    \
    ::fricc::set_last_error(e.what());                 \
    return FRICC_EINVAL;                               \
  }                                                    \
  return FRICC_OK;

// ---------------------------------------------------------------------------------------------------------------
// stream + launch accounting
// ---------------------------------------------------------------------------------------------------------------
cudaStream_t stream();
void count_launch(int64_t n = 1);
void check_launch(const char* what);  // cudaGetLastError after a launch + count it

// "uniform_carveout" option (>= 0): every kernel launched through FRICC_LAUNCH asks for the same L1/shared split (percent
// shared).  CTAs of kernels that want different splits cannot share an SM; with the peer-store kernel resident on every
// SM (it asks for the tensor-core kernel's split so that the two can run together) the short kernels of the next term
// otherwise wait for it.  -1 (default): each kernel keeps the driver's choice.
int uniform_carveout();
#define FRICC_LAUNCH(name, grid, block, smem, ...)                  \
  do {                                                              \
    if (::fricc::uniform_carveout() >= 0) {                         \
      static int fricc_carveout_ = -2;                              \
      if (fricc_carveout_ != ::fricc::uniform_carveout()) {         \
        fricc_carveout_ = ::fricc::uniform_carveout();              \
        cudaFuncSetAttribute(name, cudaFuncAttributePreferredSharedMemoryCarveout, fricc_carveout_); \
      }                                                             \
    }                                                               \
    {                                                               \
      ::fricc::SlowScope _slow("launch " #name, __FILE__, __LINE__); \
      name<<<(grid), (block), (smem), ::fricc::stream()>>>(__VA_ARGS__); \
    }                                                               \
    ::fricc::check_launch(#name);                                   \
  } while (0)

int num_sms();

// ---------------------------------------------------------------------------------------------------------------
// device memory: stream-ordered allocations (cudaMallocAsync pool, release threshold raised at init so that the
// per-call scratch buffers are recycled instead of going back to the driver)
// ---------------------------------------------------------------------------------------------------------------
void* dev_alloc(size_t bytes);
void dev_free(void* p);

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t n = 0;
  DevBuf() = default;
  explicit DevBuf(size_t count) { alloc(count); }
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  DevBuf(DevBuf&& o) noexcept : p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
  DevBuf& operator=(DevBuf&& o) noexcept {
    if (this != &o) {
      release();
      p = o.p; n = o.n; o.p = nullptr; o.n = 0;
    }
    return *this;
  }
  ~DevBuf() { release(); }
  void alloc(size_t count) {
    release();
    n = count;
    p = static_cast<T*>(dev_alloc((count ? count : 1) * sizeof(T)));
  }
  void release() {
    if (p) dev_free(p);
    p = nullptr; n = 0;
  }
  void zero() { FRICC_CUDA(cudaMemsetAsync(p, 0, (n ? n : 1) * sizeof(T), stream())); }
  void to_host(T* dst, size_t count) const {
    FRICC_CUDA(cudaMemcpyAsync(dst, p, count * sizeof(T), cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
  void to_host_at(T* dst, size_t index) const {  // one element
    FRICC_CUDA(cudaMemcpyAsync(dst, p + index, sizeof(T), cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
  void from_host(const T* src, size_t count) {
    FRICC_CUDA(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, stream()));
  }
};

// A read-only view of a caller array that is guaranteed to be on the device: either the caller's own device
// pointer (zero-copy) or a staged copy of host memory that lives as long as the view.
template <typename T>
struct DevIn {
  const T* p = nullptr;
  DevBuf<T> staged;
  DevIn(const T* src, int mem, size_t count) {
    if (mem == FRICC_DEVICE) {
      p = src;
    } else {
      staged.alloc(count);
      if (count) staged.from_host(src, count);
      p = staged.p;
    }
  }
};

// ---------------------------------------------------------------------------------------------------------------
// keys: key(x) = bit pattern of |x| -- monotone in |x| for every non-NaN double
// ---------------------------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ uint64_t abs_key(double x) {
#ifdef __CUDA_ARCH__
  return (uint64_t)__double_as_longlong(x) & 0x7fffffffffffffffULL;
#else
  uint64_t u;
  memcpy(&u, &x, 8);
  return u & 0x7fffffffffffffffULL;
#endif
}
__host__ __device__ __forceinline__ double key_to_double(uint64_t k) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)k);
#else
  double d;
  memcpy(&d, &k, 8);
  return d;
#endif
}

static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------------------------------------------
// the SparseTensor4d handle (reference: include/sparse_tensor.hpp:13-100)
// ---------------------------------------------------------------------------------------------------------------
// One bucketed layout of the entries for a given owner pair R / contracted pair K:
// entries sorted by (R, K); rowptr has nR+1 offsets; kidx is the packed K index; vals the entry values.
struct Layout {
  int r0 = -1, r1 = -1;  // which T axes form R = (r0, r1)   (K is the remaining two axes in ascending order)
  int64_t nR = 0, nK = 0;
  bool csr_built = false;
  DevBuf<int64_t> rowptr;  // [nR+1]
  DevBuf<uint32_t> kidx;   // [nnz]
  DevBuf<double> vals;     // [nnz]
  int64_t max_row = 0;     // longest bucket (entries)
  // ---- "exact order" form (fricc_contract.cu: spmm_exact): buckets by R only, entries of a bucket in STORAGE order --
  // the order in which the reference's loop nest adds them (sparse_tensor.cpp:66-73: `for s in 0..m`)
  bool xo_built = false;
  DevBuf<int64_t> xo_rowptr;  // [nR+1]
  DevBuf<uint32_t> xo_kidx;   // [nnz]
  DevBuf<double> xo_vals;     // [nnz]
  // ---- tensor-core form (fricc_dense.cu), built lazily: the entries whose value is +-c (every sampled entry of an FRI
  // compression carries the same magnitude c = r/n') as a dense int8 sign matrix A[Mpad][Kpad]; everything else (the
  // exactly preserved set D, duplicates) stays a CSR remainder that goes through the fp64 SpMM.
  bool dense_built = false;
  bool rest_built = false;     // cmag + rest_* are valid (possibly without the sign matrix: fused pairs)
  int64_t Mpad = 0, Kpad = 0;
  DevBuf<int8_t> dense;        // [Mpad][Kpad], K-major
  DevBuf<double> cmag;         // [1]  c
  DevBuf<int32_t> rowsum;      // [Mpad] sum_K dense[R][K]
  int64_t n_rest = -1;         // entries in the CSR remainder (-1: only known on the device, rest_rowptr[nR])
  DevBuf<int64_t> rest_rowptr; // [nR+1]
  DevBuf<uint32_t> rest_kidx;
  DevBuf<double> rest_vals;
};

}  // namespace fricc

struct fricc_st {
  int64_t dims[4] = {0, 0, 0, 0};
  int64_t nnz = 0;
  int64_t n_exact = -1;  // leading entries that are NOT sampled ones (|D| of an FRI compression; nnz for "largest"); -1 = unknown
  fricc::DevBuf<int64_t> flat;   // [nnz] flat row-major index, storage order = reference order [D..., S...]
  fricc::DevBuf<double> vals;    // [nnz]
  std::vector<fricc::Layout> layouts;  // built lazily per term family, invalidated by set_element
  // host mirror for get_element / print (downloaded on first use)
  std::vector<int64_t> h_flat;
  std::vector<double> h_vals;
  bool host_valid = false;
};
