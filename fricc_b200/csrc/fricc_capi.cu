// fricc_capi.cu -- the extern "C" surface declared in include/fricc_b200.h.
#include <algorithm>
#include <cmath>
#include <numeric>

#include "fricc_common.cuh"
#include "fricc_dense.cuh"
#include "fricc_fri.cuh"
#include "fricc_psrc.cuh"  // host-side construction only (no kernels in this TU)
#include "fricc_select.cuh"

namespace fricc {
const std::string& last_error();
void set_stream(cudaStream_t s);
int64_t launches();
void reset_device_cache();
void set_push_ctas_per_sm(int n);  // fricc_peer.cu: grid cap of the peer-store kernels (CTAs per SM)
void set_push_carveout(int pct);   // fricc_peer.cu: L1/shared split the peer-store kernels ask for (-1: default)
void set_push_copy_engine(int on);  // fricc_peer.cu: long-row views go through cudaMemcpy(2D)Async instead of the store kernel
void set_push_min_row_bytes(int n);
void set_peer_barrier_timeout_s(int s);       // fricc_peer.cu: how long a barrier waits for a missing rank (0 = for ever)
unsigned long long peer_barrier_timed_out();  // epoch of the first barrier that gave up since the last check (clears it)
void set_uniform_carveout(int pct);  // fricc_runtime.cu
void set_alloc_cache_mb(int mb);     // fricc_runtime.cu: bytes of freed scratch blocks kept for the next call (0: none)
void release_cached_blocks();
int64_t alloc_driver_calls();
int64_t alloc_cache_hits();
int64_t push_copies();
}  // namespace fricc

using namespace fricc;

static int64_t g_tc_calls = 0, g_spmm_calls = 0, g_fused_pairs = 0, g_const_plane_hits = 0, g_prepared_hits = 0;

namespace {

Sampling parse_sampling(const char* s) {
  if (s && strcmp(s, "pivotal") == 0) return Sampling::Pivotal;
  if (s && strcmp(s, "systematic") == 0) return Sampling::Systematic;
  // the reference's check (fri_utils.cpp:443-444) can never fire; an unknown method leaves S empty
  return Sampling::None;
}

int64_t prod4(const int64_t d[4]) { return d[0] * d[1] * d[2] * d[3]; }

// number of elements spanned by a strided 4-D view with non-negative strides (the pointer addresses element 0,0,0,0)
int64_t span_of(const int64_t shape[4], const int64_t strides[4]) {
  int64_t s = 1;
  for (int i = 0; i < 4; i++) {
    if (shape[i] <= 0) return 0;
    if (strides[i] < 0) fail(FRICC_EINVAL, "negative strides are not supported at the C-ABI (make the array contiguous)");
    s += (shape[i] - 1) * strides[i];
  }
  return s;
}

void build_from_compressed(fricc_st* t, Compressed& c) {
  t->nnz = c.m;
  t->n_exact = c.nD;
  t->flat = std::move(c.flat);
  t->vals = std::move(c.vals);
  t->layouts.clear();
  t->host_valid = false;
}

void ensure_host(fricc_st* t) {
  if (t->host_valid) return;
  t->h_flat.resize((size_t)t->nnz);
  t->h_vals.resize((size_t)t->nnz);
  if (t->nnz > 0) {
    FRICC_CUDA(cudaMemcpyAsync(t->h_flat.data(), t->flat.p, (size_t)t->nnz * 8, cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaMemcpyAsync(t->h_vals.data(), t->vals.p, (size_t)t->nnz * 8, cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
  t->host_valid = true;
}

void check_handle(fricc_st_t t) {
  if (!t) fail(FRICC_EINVAL, "null SparseTensor4d handle");
}

int create_impl(const double* tensor, int mem, const int64_t dims[4], int64_t m, const char* compression,
                const char* sampling_method, int verbose, const DrawSource& draws, int64_t* n_consumed,
                fricc_st_t* out) {
  FRICC_API_BEGIN
  if (!tensor || !dims || !out || !compression) fail(FRICC_EINVAL, "null argument");
  for (int i = 0; i < 4; i++)
    if (dims[i] <= 0) fail(FRICC_EINVAL, "dims must be positive");
  const int64_t n = prod4(dims);
  if (m < 0 || m > n)
    fail(FRICC_EINVAL, "Bad m value: need 0 <= m <= tensor size (m=%lld, size=%lld)", (long long)m, (long long)n);
  const bool largest = strcmp(compression, "largest") == 0;
  const bool fri = strcmp(compression, "fri") == 0;
  if (!largest && !fri)
    fail(FRICC_EINVAL, "The compression method you chose (%s) isn't supported: compression must be 'largest' or 'fri'",
         compression);
  DevIn<double> x(tensor, mem, (size_t)n);
  Compressed c;
  if (largest) compress_largest(x.p, n, m, c);
  else compress_fri(x.p, n, m, parse_sampling(sampling_method), draws, verbose != 0, c);
  if (n_consumed) *n_consumed = c.consumed;
  fricc_st* t = new fricc_st();
  for (int i = 0; i < 4; i++) t->dims[i] = dims[i];
  build_from_compressed(t, c);
  if (largest) t->n_exact = t->nnz;  // no common magnitude to expect
  // an unknown sampling method leaves S empty and the tail unfilled (fri_utils.cpp:443-444 can never fire): no promise
  // about the entries behind D -- the tensor-core form then classifies every entry instead of trusting the hint
  else if (parse_sampling(sampling_method) == Sampling::None) t->n_exact = -1;
  FRICC_CUDA(cudaStreamSynchronize(stream()));  // staged input must outlive the kernels that read it
  *out = t;
  FRICC_API_END
}

}  // namespace

extern "C" {

const char* fricc_last_error(void) { return fricc::last_error().c_str(); }
const char* fricc_version(void) { return "fricc_b200 0.1 (sm_100a)"; }

int fricc_device_count(int* count) {
  FRICC_API_BEGIN
  int c = 0;
  cudaError_t e = cudaGetDeviceCount(&c);
  if (e != cudaSuccess) {
    cudaGetLastError();
    c = 0;
  }
  *count = c;
  FRICC_API_END
}
int fricc_set_device(int device) {
  FRICC_API_BEGIN
  FRICC_CUDA(cudaSetDevice(device));
  fricc::reset_device_cache();
  FRICC_API_END
}
int fricc_set_stream(void* s) {
  fricc::set_stream(reinterpret_cast<cudaStream_t>(s));
  return FRICC_OK;
}
int fricc_stream_synchronize(void) {
  FRICC_API_BEGIN
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  if (const unsigned long long e = peer_barrier_timed_out())
    fail(FRICC_ECUDA, "fricc_peer_barrier number %llu gave up waiting for a peer (option \"peer_barrier_timeout_s\"): "
                      "results queued behind it are void", e);
  FRICC_API_END
}
int64_t fricc_kernel_launches(void) { return fricc::launches(); }

int fricc_set_option(const char* key, int value) {
  FRICC_API_BEGIN
  if (key && strcmp(key, "escan") == 0) set_parallel_scan(value);
  else if (key && strcmp(key, "d_candidates") == 0) set_first_candidates(value);
  else if (key && strcmp(key, "d_estimate") == 0) set_d_estimate(value);
  else if (key && strcmp(key, "pivotal_parallel") == 0) set_pivotal_parallel(value);
  else if (key && strcmp(key, "d_parallel") == 0) set_d_parallel(value);
  else if (key && strcmp(key, "push_ctas_per_sm") == 0) set_push_ctas_per_sm(value);
  else if (key && strcmp(key, "push_carveout") == 0) set_push_carveout(value);
  else if (key && strcmp(key, "push_copy_engine") == 0) set_push_copy_engine(value);
  else if (key && strcmp(key, "push_min_row_bytes") == 0) set_push_min_row_bytes(value);
  else if (key && strcmp(key, "peer_barrier_timeout_s") == 0) set_peer_barrier_timeout_s(value);
  else if (key && strcmp(key, "uniform_carveout") == 0) set_uniform_carveout(value);
  else if (key && strcmp(key, "alloc_cache_mb") == 0) set_alloc_cache_mb(value);
  else if (key && strcmp(key, "release_cached") == 0) release_cached_blocks();
  else if (key && strncmp(key, "spmm_", 5) == 0) set_spmm_option(key, value);
  else if (key && strncmp(key, "tc_", 3) == 0) set_tc_option(key, value);
  else fail(FRICC_EINVAL, "unknown option '%s'", key ? key : "(null)");
  FRICC_API_END
}

int fricc_get_stat(const char* key, int64_t* value) {
  FRICC_API_BEGIN
  if (!key || !value) fail(FRICC_EINVAL, "null argument");
  if (strcmp(key, "pivotal_restarts") == 0) *value = pivotal_restarts_last();
  else if (strcmp(key, "tc_contract_calls") == 0) *value = g_tc_calls;      // fricc_contract calls that took the tensor-core form
  else if (strcmp(key, "spmm_contract_calls") == 0) *value = g_spmm_calls;
  else if (strcmp(key, "prepared_plane_hits") == 0) *value = g_prepared_hits;   // tensor-core calls that found their planes prepared
  else if (strcmp(key, "constant_plane_hits") == 0) *value = g_const_plane_hits;   // plugin calls that reused the planes of a declared-constant operand
  else if (strcmp(key, "fused_pairs") == 0) *value = g_fused_pairs;   // fricc_contract_pair calls that ran as ONE dense product
  else if (strcmp(key, "alloc_driver_calls") == 0) *value = alloc_driver_calls();   // cudaMallocAsync calls so far
  else if (strcmp(key, "alloc_cache_hits") == 0) *value = alloc_cache_hits();       // allocations served from cached blocks
  else if (strcmp(key, "peer_barrier_timeouts") == 0) *value = (int64_t)peer_barrier_timed_out();   // epoch of the first barrier that gave up (0 = none); clears it
  else if (strcmp(key, "push_copies") == 0) *value = push_copies();   // cudaMemcpy(2D)Async calls issued by fricc_push_view
  else if (strcmp(key, "planes_predicted_builds") == 0) *value = planes_predicted_builds();  // builds that reused exponents  // ... and the fp64 SpMM
  else fail(FRICC_EINVAL, "unknown stat '%s'", key);
  FRICC_API_END
}

int fricc_srand(uint32_t seed) {
  rand_seed(seed);
  return FRICC_OK;
}
int fricc_rand_fill(int32_t* out, int64_t n) {
  FRICC_API_BEGIN
  if (n < 0 || (!out && n > 0)) fail(FRICC_EINVAL, "bad arguments");
  rand_fill(out, n);
  FRICC_API_END
}
int64_t fricc_rand_consumed(void) { return rand_consumed(); }
int fricc_rand_skip(int64_t n) {
  FRICC_API_BEGIN
  if (n < 0) fail(FRICC_EINVAL, "bad arguments");
  rand_skip(n);
  FRICC_API_END
}
int fricc_rand_state_size(void) { return (int)sizeof(RandSnapshot); }
int fricc_rand_get_state(void* buf) {
  FRICC_API_BEGIN
  if (!buf) fail(FRICC_EINVAL, "null argument");
  RandSnapshot s = rand_snapshot();
  memcpy(buf, &s, sizeof(s));
  FRICC_API_END
}
int fricc_rand_set_state(const void* buf) {
  FRICC_API_BEGIN
  if (!buf) fail(FRICC_EINVAL, "null argument");
  RandSnapshot s;
  memcpy(&s, buf, sizeof(s));
  rand_restore(s);
  FRICC_API_END
}
int fricc_rand_fill_device(int32_t* out_dev, int64_t n) {
  FRICC_API_BEGIN
  if (n < 0 || (!out_dev && n > 0)) fail(FRICC_EINVAL, "bad arguments");
  rand_fill_device(out_dev, n);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  FRICC_API_END
}

// ---- SparseTensor4d ------------------------------------------------------------------------------------------
int fricc_st_create_empty(const int64_t dims[4], int64_t nnz, fricc_st_t* out) {
  FRICC_API_BEGIN
  if (!dims || !out || nnz < 0) fail(FRICC_EINVAL, "bad arguments");
  fricc_st* t = new fricc_st();
  for (int i = 0; i < 4; i++) t->dims[i] = dims[i];
  t->nnz = nnz;
  t->flat.alloc((size_t)std::max<int64_t>(nnz, 1));
  t->vals.alloc((size_t)std::max<int64_t>(nnz, 1));
  t->flat.zero();
  t->vals.zero();
  *out = t;
  FRICC_API_END
}

int fricc_st_create(const double* tensor, int mem, const int64_t dims[4], int64_t m, const char* compression,
                    const char* sampling_method, int verbose, fricc_st_t* out) {
  DrawSource ds;
  ds.from_stream = true;
  return create_impl(tensor, mem, dims, m, compression, sampling_method, verbose, ds, nullptr, out);
}

int fricc_st_create_with_draws(const double* tensor, int mem, const int64_t dims[4], int64_t m,
                               const char* compression, const char* sampling_method, int verbose, const int32_t* draws,
                               int64_t n_draws, int64_t* n_consumed, fricc_st_t* out) {
  DrawSource ds;
  ds.from_stream = false;
  ds.host = draws;
  ds.n = draws ? n_draws : 0;
  return create_impl(tensor, mem, dims, m, compression, sampling_method, verbose, ds, n_consumed, out);
}

int fricc_st_create_from_coo(const int64_t* flat_idx, const double* vals, int mem, const int64_t dims[4], int64_t m,
                             fricc_st_t* out) {
  FRICC_API_BEGIN
  if (!dims || !out || m < 0 || (m > 0 && (!flat_idx || !vals))) fail(FRICC_EINVAL, "bad arguments");
  fricc_st* t = new fricc_st();
  for (int i = 0; i < 4; i++) t->dims[i] = dims[i];
  t->nnz = m;
  t->flat.alloc((size_t)std::max<int64_t>(m, 1));
  t->vals.alloc((size_t)std::max<int64_t>(m, 1));
  if (m > 0) {
    cudaMemcpyKind kind = mem == FRICC_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    FRICC_CUDA(cudaMemcpyAsync(t->flat.p, flat_idx, (size_t)m * 8, kind, stream()));
    FRICC_CUDA(cudaMemcpyAsync(t->vals.p, vals, (size_t)m * 8, kind, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
  *out = t;
  FRICC_API_END
}

int fricc_st_get_n_exact(fricc_st_t t, int64_t* n_exact) {
  FRICC_API_BEGIN
  check_handle(t);
  if (!n_exact) fail(FRICC_EINVAL, "null argument");
  *n_exact = t->n_exact;
  FRICC_API_END
}

int fricc_st_set_n_exact(fricc_st_t t, int64_t n_exact) {
  FRICC_API_BEGIN
  check_handle(t);
  if (n_exact < -1 || n_exact > t->nnz) fail(FRICC_EINVAL, "n_exact out of range");
  t->n_exact = n_exact;
  FRICC_API_END
}

int fricc_st_destroy(fricc_st_t t) {
  FRICC_API_BEGIN
  delete t;
  FRICC_API_END
}

int fricc_st_size(fricc_st_t t, int64_t* nnz) {
  FRICC_API_BEGIN
  check_handle(t);
  *nnz = t->nnz;
  FRICC_API_END
}
int fricc_st_dims(fricc_st_t t, int64_t dims[4]) {
  FRICC_API_BEGIN
  check_handle(t);
  for (int i = 0; i < 4; i++) dims[i] = t->dims[i];
  FRICC_API_END
}

int fricc_st_get_element(fricc_st_t t, int64_t mi, int64_t idx[4], double* value) {
  FRICC_API_BEGIN
  check_handle(t);
  if (mi < 0 || mi >= t->nnz) fail(FRICC_ERANGE, "Index you requested is larger than the specified size");
  ensure_host(t);
  int64_t f = t->h_flat[(size_t)mi];
  const int64_t* d = t->dims;
  // unpack_idx (sparse_tensor.hpp:108-125)
  idx[0] = f / (d[1] * d[2] * d[3]);
  idx[1] = f % (d[1] * d[2] * d[3]) / (d[2] * d[3]);
  idx[2] = f % (d[2] * d[3]) / d[3];
  idx[3] = f % d[3];
  *value = t->h_vals[(size_t)mi];
  FRICC_API_END
}

int fricc_st_set_element(fricc_st_t t, int64_t mi, int64_t flat_idx, double value) {
  FRICC_API_BEGIN
  check_handle(t);
  if (mi < 0 || mi >= t->nnz) fail(FRICC_ERANGE, "Index you requested is larger than the specified size");
  if (flat_idx < 0 || flat_idx >= prod4(t->dims)) fail(FRICC_ERANGE, "Indices are out of range for given tensor");
  FRICC_CUDA(cudaMemcpyAsync(t->flat.p + mi, &flat_idx, 8, cudaMemcpyHostToDevice, stream()));
  FRICC_CUDA(cudaMemcpyAsync(t->vals.p + mi, &value, 8, cudaMemcpyHostToDevice, stream()));
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  t->layouts.clear();
  t->n_exact = -1;
  t->host_valid = false;
  FRICC_API_END
}

int fricc_st_export(fricc_st_t t, int64_t* flat_idx, double* vals, int mem) {
  FRICC_API_BEGIN
  check_handle(t);
  if (t->nnz > 0) {
    cudaMemcpyKind kind = mem == FRICC_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost;
    if (flat_idx) FRICC_CUDA(cudaMemcpyAsync(flat_idx, t->flat.p, (size_t)t->nnz * 8, kind, stream()));
    if (vals) FRICC_CUDA(cudaMemcpyAsync(vals, t->vals.p, (size_t)t->nnz * 8, kind, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
  FRICC_API_END
}

}  // extern "C"

// ---- contractions --------------------------------------------------------------------------------------------
namespace {

OutMap out_map(const TermSpec& ts, const int64_t* t_dims, const int64_t out_shape[4], const int64_t out_strides[4]) {
  OutMap om;
  om.nR2 = t_dims[ts.tR[1]];
  om.nL2 = out_shape[ts.oL[1]];
  om.sR1 = out_strides[ts.oR[0]];
  om.sR2 = out_strides[ts.oR[1]];
  om.sL1 = out_strides[ts.oL[0]];
  om.sL2 = out_strides[ts.oL[1]];
  return om;
}

// The owner pair R and the free pair L of a term are unordered as far as the mathematics goes; the kernels walk the
// MINOR index of R along the lanes of a warp and the minor index of L along a thread's consecutive columns.  Put
// the index with the smaller output stride in the minor position, so that e.g. out(i,j,a,b) written by R = (b,j)
// (terms 1302/1202/1303) is stored along b, its unit-stride axis, instead of one 32-byte sector per element.
TermSpec oriented(const TermSpec& ts, const int64_t out_strides[4]) {
  TermSpec e = ts;
  if (out_strides[e.oR[0]] < out_strides[e.oR[1]]) {
    std::swap(e.tR[0], e.tR[1]);
    std::swap(e.oR[0], e.oR[1]);
  }
  if (out_strides[e.oL[0]] < out_strides[e.oL[1]]) {
    std::swap(e.wL[0], e.wL[1]);
    std::swap(e.oL[0], e.oL[1]);
  }
  return e;
}

// ---- digit planes kept across calls (option "tc_plane_cache", off by default) ---------------------------------
// FRILCCSD hands the SAME operands to contract_DTSpT in every iteration (fricc/frilccsd.py:171-179); with the cache on,
// their planes are built once.  An entry is keyed by (address, shape, strides, K/L axes, planes) and guarded by a
// fingerprint of 4096 evenly spaced elements, re-read on every call -- a changed operand at an old address is rebuilt
// unless every sampled element is unchanged, which is why the cache is opt-in: the caller states that operands at one
// address are either rewritten wholesale or not at all.
struct PlaneEntry {
  const double* p = nullptr;
  int64_t shape[4] = {0, 0, 0, 0}, strides[4] = {0, 0, 0, 0};
  int axes[4] = {0, 0, 0, 0};
  int S = 0;
  uint64_t print = 0;
  uint64_t stamp = 0;
  const double* prepared = nullptr;   // history entries: the operand fricc_prepare_operand encoded these planes from
  fricc::WSlices ws;
  // History entries are shared by every operand of one geometry and may be rebuilt on one stream (fricc_prepare_operand
  // on the caller's second stream) while a contraction on another stream still reads them, or read before the build on
  // the other stream has finished.  Two events close both holes inside the library, whatever the caller's own stream
  // discipline is: `ready` is recorded behind every build and waited for by the stream of every reader; `last_use` is
  // recorded behind every reader and waited for by the stream of the next build.
  cudaEvent_t ready = nullptr, last_use = nullptr;
  static void mark(cudaEvent_t& ev) {   // record on the library's stream (event created on first use)
    if (!ev) FRICC_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    FRICC_CUDA(cudaEventRecord(ev, fricc::stream()));
  }
  static void wait(cudaEvent_t ev) {    // the library's stream waits (nothing to wait for before the first record)
    if (ev) FRICC_CUDA(cudaStreamWaitEvent(fricc::stream(), ev, 0));
  }
  ~PlaneEntry() {
    if (ready) cudaEventDestroy(ready);
    if (last_use) cudaEventDestroy(last_use);
  }
};
PlaneEntry* g_planes_in_use = nullptr;   // the history entry planes_for handed out last (its reader records last_use)
void planes_used() {
  if (g_planes_in_use) PlaneEntry::mark(g_planes_in_use->last_use);
  g_planes_in_use = nullptr;
}
std::vector<PlaneEntry*> g_plane_cache;
// Operands the caller has declared constant (fricc_constant_operand): address ranges, and their planes keyed by
// (address, shape, strides, K/L axes, planes) -- no fingerprint, no read-back: the declaration is the guarantee.
struct ConstRange { const double* lo; const double* hi; };
std::vector<ConstRange> g_const_ranges;
std::vector<PlaneEntry*> g_plane_const;
WSlices* g_pair_planes = nullptr;        // planes of the combined operand of the last fused pair (fricc_contract_pair)
std::vector<PlaneEntry*> g_plane_hist;   // geometry-keyed buffers + exponents ("tc_history_mb"), see planes_for
uint64_t g_plane_stamp = 0;
constexpr size_t kPlaneCacheEntries = 12;

}  // namespace

namespace fricc {
uint64_t view_fingerprint(const WView& w, int64_t l_begin, int64_t l_count);  // fricc_dense.cu
}

namespace {

bool is_constant_operand(const double* p) {
  if (!p) return false;
  for (const ConstRange& r : g_const_ranges)
    if (p >= r.lo && p < r.hi) return true;
  return false;
}

void const_planes_drop(const double* lo, const double* hi) {   // entries whose operand starts inside [lo, hi)
  for (size_t i = 0; i < g_plane_const.size();) {
    if (g_plane_const[i]->p >= lo && g_plane_const[i]->p < hi) {
      delete g_plane_const[i];
      g_plane_const.erase(g_plane_const.begin() + (long)i);
    } else {
      i++;
    }
  }
}

void plane_cache_clear() {
  for (PlaneEntry* e : g_plane_const) delete e;
  g_plane_const.clear();
  for (PlaneEntry* e : g_plane_cache) delete e;
  g_plane_cache.clear();
  for (PlaneEntry* e : g_plane_hist) delete e;
  g_plane_hist.clear();
  g_planes_in_use = nullptr;
  delete g_pair_planes;
  g_pair_planes = nullptr;
}

const WSlices& planes_for(const double* Wd, const int64_t shape[4], const int64_t strides[4], const TermSpec& ts,
                          const WView& view, int64_t nL, WSlices& local, bool prepare_only = false) {
  const int S = tc_planes();
  g_planes_in_use = nullptr;
  if (is_constant_operand(Wd)) {
    // declared constant by the caller (the ERI blocks of a CC driver, fricc/friccsd.py:347-378 pass eris.ovov to five
    // of the eleven terms in every iteration): planes are built on first use and kept
    const int axes[4] = {ts.wK[0], ts.wK[1], ts.wL[0], ts.wL[1]};
    for (PlaneEntry* e : g_plane_const)
      if (e->p == Wd && e->S == S && memcmp(e->shape, shape, sizeof(e->shape)) == 0 &&
          memcmp(e->strides, strides, sizeof(e->strides)) == 0 && memcmp(e->axes, axes, sizeof(axes)) == 0) {
        g_const_plane_hits++;
        return e->ws;
      }
    PlaneEntry* e = new PlaneEntry();
    e->p = Wd; e->S = S;
    memcpy(e->shape, shape, sizeof(e->shape));
    memcpy(e->strides, strides, sizeof(e->strides));
    memcpy(e->axes, axes, sizeof(axes));
    g_plane_const.push_back(e);
    wslices_build_view(e->ws, view, 0, nL, S);
    return e->ws;
  }
  if (!tc_plane_cache()) {
    // no caching of CONTENT, but the buffers and column exponents of the last operand of this geometry are kept: an
    // iterative driver comes back with the same shapes every iteration, and the exponents predict the next build
    // (planes_predict_kernel: one streaming pass instead of the two-look build) while the 19 GB of planes of a ladder
    // slice are not re-allocated.  Bounded by "tc_history_mb" (least recently used geometries go first; 0 = keep nothing).
    if (tc_history_mb() <= 0) {
      wslices_build_view(local, view, 0, nL, S);
      return local;
    }
    const int axes[4] = {ts.wK[0], ts.wK[1], ts.wL[0], ts.wL[1]};
    PlaneEntry* hit = nullptr;
    for (PlaneEntry* e : g_plane_hist)
      if (e->S == S && memcmp(e->shape, shape, sizeof(e->shape)) == 0 && memcmp(e->strides, strides, sizeof(e->strides)) == 0 &&
          memcmp(e->axes, axes, sizeof(axes)) == 0) {
        hit = e;
        break;
      }
    if (!hit) {
      hit = new PlaneEntry();
      hit->S = S;
      memcpy(hit->shape, shape, sizeof(hit->shape));
      memcpy(hit->strides, strides, sizeof(hit->strides));
      memcpy(hit->axes, axes, sizeof(axes));
      g_plane_hist.push_back(hit);
    }
    hit->stamp = ++g_plane_stamp;
    if (Wd && hit->prepared == Wd && !prepare_only) {
      // encoded ahead of this call by fricc_prepare_operand (on whatever stream the caller chose): used once
      hit->prepared = nullptr;
      g_prepared_hits++;
      PlaneEntry::wait(hit->ready);       // the reader's stream is ordered behind the encoding
      g_planes_in_use = hit;
      return hit->ws;
    }
    PlaneEntry::wait(hit->last_use);      // the build overwrites planes a reader on another stream may still be using
    wslices_build_view(hit->ws, view, 0, nL, S);
    PlaneEntry::mark(hit->ready);
    hit->prepared = prepare_only ? Wd : nullptr;
    g_planes_in_use = prepare_only ? nullptr : hit;
    // stay under the byte budget (never evicting the entry just built)
    const size_t budget = (size_t)tc_history_mb() << 20;
    for (;;) {
      size_t total = 0;
      PlaneEntry* lru = nullptr;
      for (PlaneEntry* e : g_plane_hist) {
        total += e->ws.q.n;
        if (e != hit && (!lru || e->stamp < lru->stamp)) lru = e;
      }
      if (total <= budget || !lru) break;
      g_plane_hist.erase(std::find(g_plane_hist.begin(), g_plane_hist.end(), lru));
      if (g_planes_in_use == lru) g_planes_in_use = nullptr;
      delete lru;
    }
    return hit->ws;
  }
  const uint64_t fp = view_fingerprint(view, 0, nL);
  const int axes[4] = {ts.wK[0], ts.wK[1], ts.wL[0], ts.wL[1]};
  for (PlaneEntry* e : g_plane_cache) {
    if (e->p == Wd && e->S == S && memcmp(e->shape, shape, sizeof(e->shape)) == 0 &&
        memcmp(e->strides, strides, sizeof(e->strides)) == 0 && memcmp(e->axes, axes, sizeof(axes)) == 0) {
      if (e->print != fp) {
        wslices_build_view(e->ws, view, 0, nL, S);
        e->print = fp;
      }
      e->stamp = ++g_plane_stamp;
      return e->ws;
    }
  }
  if (g_plane_cache.size() >= kPlaneCacheEntries) {  // evict the least recently used
    size_t lru = 0;
    for (size_t i = 1; i < g_plane_cache.size(); i++)
      if (g_plane_cache[i]->stamp < g_plane_cache[lru]->stamp) lru = i;
    delete g_plane_cache[lru];
    g_plane_cache.erase(g_plane_cache.begin() + (long)lru);
  }
  PlaneEntry* e = new PlaneEntry();
  e->p = Wd; e->S = S; e->print = fp; e->stamp = ++g_plane_stamp;
  memcpy(e->shape, shape, sizeof(e->shape));
  memcpy(e->strides, strides, sizeof(e->strides));
  memcpy(e->axes, axes, sizeof(axes));
  g_plane_cache.push_back(e);
  wslices_build_view(e->ws, view, 0, nL, S);
  return e->ws;
}

// fricc_contract's choice between the two forms of a contraction ("tc_mode": 0 never, 1 this cost model, 2 always).
// Rates measured on B200 at the headline shape (profiles/): the bucketed fp64 SpMM sustains ~2.7e12 FMA/s (L1 data
// pipe), the int8 form ~1.5e15 MAC/s over the PADDED dense product, plane building runs at HBM speed.
bool use_tensor_cores(fricc_st* T, int64_t nR, int64_t nK, int64_t nL, bool panel_in_place) {
  const int mode = tc_mode();
  if (mode == 0 || spmm_exact_order() || T->nnz == 0) return false;
  if (mode == 2) return true;
  if (T->n_exact < 0) return false;                        // unknown provenance: would have to classify every entry
  const double m = (double)T->nnz, nS = (double)(T->nnz - T->n_exact);
  if (2.0 * nS < m || nK < 256 || nL < 64 || nR < 32) return false;
  const double S = tc_planes();
  const double Mpad = (double)(ceil_div(nR, 128) * 128), Kpad = (double)(ceil_div(nK, 128) * 128);
  // fp64 SpMM: ~2.7e12 FMA/s on wide column ranges, about half of that on a few hundred columns, plus the (R,K) bucket
  // sort of the entries that only this form needs (~0.8 ms at m = 1e7)
  const double spmm_rate = nL >= 2048 ? 2.7e12 : 1.5e12;
  const double t_spmm = m * (double)nL / spmm_rate + m * 0.8e-10 +
                        (panel_in_place ? 0.0 : 16.0 * (double)nK * (double)nL / 3.0e12);
  // the pair kernel works in whole rounds of (256 rows x 64 columns) tiles over SMs/2 CTA pairs, 128 S cycles of tensor
  // work per 128-deep k-block at ~1.6 GHz under the power cap (a 200-column slice of a 1600-row term is ONE round that
  // leaves most pairs idle: the SpMM wins there); then the sign matrix (sort-free scatter + row sums), the planes and
  // the fp64 remainder of the exact entries
  const double tiles = (double)ceil_div(nR, 256) * (double)ceil_div(nL, 64);
  const double rounds = std::ceil(tiles / std::max(1, num_sms() / 2));
  const double t_tc = rounds * (Kpad / 128.0) * (128.0 * S) / 1.6e9 +
                      (double)nK * (double)nL * (16.0 + S) / 3.0e12 + Mpad * Kpad / 1.0e12 + 0.3e-3 +
                      (m - nS) * (double)nL / 1.0e12;
  return t_tc < t_spmm;
}

}  // namespace

extern "C" {

int fricc_contract(const double* W, int w_mem, const int64_t w_shape[4], const int64_t w_strides[4], fricc_st_t T,
                   double* out, int out_mem, const int64_t out_shape[4], const int64_t out_strides[4],
                   const char* term) {
  FRICC_API_BEGIN
  check_handle(T);
  if (!W || !out || !w_shape || !w_strides || !out_shape || !out_strides) fail(FRICC_EINVAL, "null argument");
  const TermSpec* ts0 = find_term(term);
  if (!ts0) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  const int64_t* d = T->dims;
  // shape agreement (the reference indexes unchecked; here a mismatch is an error instead of a wild read)
  for (int q = 0; q < 2; q++) {
    if (w_shape[ts0->wK[q]] != d[ts0->tK[q]])
      fail(FRICC_EINVAL, "term %s (%s): W axis %d has extent %lld, T axis %d has %lld", ts0->name, ts0->einsum,
           ts0->wK[q], (long long)w_shape[ts0->wK[q]], ts0->tK[q], (long long)d[ts0->tK[q]]);
    if (out_shape[ts0->oR[q]] != d[ts0->tR[q]])
      fail(FRICC_EINVAL, "term %s (%s): out axis %d has extent %lld, T axis %d has %lld", ts0->name, ts0->einsum,
           ts0->oR[q], (long long)out_shape[ts0->oR[q]], ts0->tR[q], (long long)d[ts0->tR[q]]);
    if (out_shape[ts0->oL[q]] != w_shape[ts0->wL[q]])
      fail(FRICC_EINVAL, "term %s (%s): out axis %d has extent %lld, W axis %d has %lld", ts0->name, ts0->einsum,
           ts0->oL[q], (long long)out_shape[ts0->oL[q]], ts0->wL[q], (long long)w_shape[ts0->wL[q]]);
  }
  const TermSpec tsv = oriented(*ts0, out_strides);
  const TermSpec* ts = &tsv;
  const int64_t nK1 = w_shape[ts->wK[0]], nK2 = w_shape[ts->wK[1]];
  const int64_t nL1 = w_shape[ts->wL[0]], nL2 = w_shape[ts->wL[1]];
  const int64_t nK = nK1 * nK2, nL = nL1 * nL2;
  const int64_t nR = d[ts->tR[0]] * d[ts->tR[1]];
  if (nK == 0 || nL == 0 || T->nnz == 0) return FRICC_OK;

  const int64_t w_span = span_of(w_shape, w_strides);
  const int64_t o_span = span_of(out_shape, out_strides);
  DevIn<double> Wd(W, w_mem, (size_t)w_span);

  DevBuf<double> out_stage;
  double* od = out;
  if (out_mem != FRICC_DEVICE) {
    out_stage.alloc((size_t)o_span);
    out_stage.from_host(out, (size_t)o_span);
    od = out_stage.p;
  }
  const OutMap om = out_map(*ts, d, out_shape, out_strides);

  // is W already the [K1][K2][L1][L2] row-major panel?  then the SpMM streams it in place
  const bool in_place = w_strides[ts->wL[1]] == 1 && w_strides[ts->wL[0]] == nL2 && w_strides[ts->wK[1]] == nL &&
                        w_strides[ts->wK[0]] == nK2 * nL;
  if (use_tensor_cores(T, nR, nK, nL, in_place)) {
    // tensor-core form, W read where it lies: digit planes straight from the 4-D operand, sign matrix x planes on
    // tcgen05.mma.kind::i8, the exact part of T through the fp64 remainder kernel on the same view
    g_tc_calls++;
    const WView view = term_view(Wd.p, w_shape, w_strides, *ts);
    WSlices local;
    const WSlices& ws = planes_for(w_mem == FRICC_DEVICE ? W : nullptr, w_shape, w_strides, *ts, view, nL, local);
    Layout& lay = get_layout(T, *ts, false);
    contract_dense(T, *ts, lay, ws, view, 0, od, om, ts->scale, false);
    planes_used();
  } else {
    g_spmm_calls++;
    if (w_mem == FRICC_DEVICE)
      for (PlaneEntry* e : g_plane_hist)
        if (e->prepared == W) e->prepared = nullptr;   // prepared for a call that did not take the tensor-core form
    DevBuf<double> panel;
    const double* Wp = Wd.p;
    if (!in_place) {
      panel.alloc((size_t)(nK * nL));
      permute_w(Wd.p, w_shape, w_strides, *ts, 0, nL, panel.p);
      Wp = panel.p;
    }
    spmm_term(T, *ts, Wp, nL, 0, nL, od, om);
  }
  if (out_mem != FRICC_DEVICE) out_stage.to_host(out, (size_t)o_span);
  // (all-device calls return without synchronising: scratch buffers are stream-ordered allocations)
  FRICC_API_END
}

int fricc_contract_pair(const double* W_a, const int64_t a_shape[4], const int64_t a_strides[4], const double* W_b,
                        const int64_t b_shape[4], const int64_t b_strides[4], fricc_st_t T, double* out,
                        const int64_t out_shape[4], const int64_t out_strides[4], const char* term_a, const char* term_b) {
  FRICC_API_BEGIN
  check_handle(T);
  if (!W_a || !W_b || !out || !a_shape || !a_strides || !b_shape || !b_strides || !out_shape || !out_strides)
    fail(FRICC_EINVAL, "null argument");
  const TermSpec* ta0 = find_term(term_a);
  const TermSpec* tb0 = find_term(term_b);
  if (!ta0 || !tb0) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term");
  // the two plain calls: what this function must equal, and what it falls back to
  auto two_calls = [&]() {
    int rc = fricc_contract(W_a, FRICC_DEVICE, a_shape, a_strides, T, out, FRICC_DEVICE, out_shape, out_strides, term_a);
    if (rc != FRICC_OK) return rc;
    return fricc_contract(W_b, FRICC_DEVICE, b_shape, b_strides, T, out, FRICC_DEVICE, out_shape, out_strides, term_b);
  };
  const TermSpec tsa = oriented(*ta0, out_strides), tsb = oriented(*tb0, out_strides);
  const bool same_out = memcmp(tsa.oR, tsb.oR, sizeof(tsa.oR)) == 0 && memcmp(tsa.oL, tsb.oL, sizeof(tsa.oL)) == 0;
  const bool same_t = memcmp(tsa.tR, tsb.tR, sizeof(tsa.tR)) == 0 && memcmp(tsa.tK, tsb.tK, sizeof(tsa.tK)) == 0;
  const bool same_w = W_a == W_b && memcmp(a_shape, b_shape, 4 * sizeof(int64_t)) == 0 &&
                      memcmp(a_strides, b_strides, 4 * sizeof(int64_t)) == 0 &&
                      memcmp(tsa.wK, tsb.wK, sizeof(tsa.wK)) == 0 && memcmp(tsa.wL, tsb.wL, sizeof(tsa.wL)) == 0;
  // small-integer ratio of the prefactors: s_a = sigma q_a, s_b = sigma q_b
  const double sigma = std::min(std::fabs(tsa.scale), std::fabs(tsb.scale));
  const double qa = tsa.scale / sigma, qb = tsb.scale / sigma;
  const bool int_ratio = sigma > 0.0 && qa == std::nearbyint(qa) && qb == std::nearbyint(qb) && std::fabs(qa) <= 8 && std::fabs(qb) <= 8;
  const int64_t* d = T->dims;
  const int64_t nK = a_shape[tsa.wK[0]] * a_shape[tsa.wK[1]], nL = a_shape[tsa.wL[0]] * a_shape[tsa.wL[1]];
  const int64_t nR = d[tsa.tR[0]] * d[tsa.tR[1]];
  const bool shapes_ok = b_shape[tsb.wK[0]] * b_shape[tsb.wK[1]] == nK && b_shape[tsb.wL[0]] == a_shape[tsa.wL[0]] &&
                         b_shape[tsb.wL[1]] == a_shape[tsa.wL[1]] && d[tsb.tR[0]] * d[tsb.tR[1]] == nR;
  if (!same_out || !int_ratio || !shapes_ok || !(same_t || same_w) || T->n_exact < 0 || nK == 0 || nL == 0 || T->nnz == 0 ||
      !use_tensor_cores(T, nR, nK, nL, false))
    return two_calls();
  // shape agreement of each term with T / out: let the plain path report it
  for (int q = 0; q < 2; q++)
    if (a_shape[tsa.wK[q]] != d[tsa.tK[q]] || b_shape[tsb.wK[q]] != d[tsb.tK[q]] || out_shape[tsa.oR[q]] != d[tsa.tR[q]] ||
        out_shape[tsa.oL[q]] != a_shape[tsa.wL[q]] ||
        // term b axis by axis as well: on the same-operand path its owner pair comes from OTHER axes of T than term a's,
        // and the product test above lets a T whose two extents are swapped against out (3,4 vs 4,3) through
        out_shape[tsb.oR[q]] != d[tsb.tR[q]] || out_shape[tsb.oL[q]] != b_shape[tsb.wL[q]])
      return two_calls();
  (void)span_of(a_shape, a_strides);
  (void)span_of(b_shape, b_strides);
  (void)span_of(out_shape, out_strides);
  const OutMap om = out_map(tsa, d, out_shape, out_strides);
  g_tc_calls++;
  g_fused_pairs++;
  if (same_w) {
    // one operand view, (R, K) from different axes of T: one weighted sign matrix
    const WView view = term_view(W_a, a_shape, a_strides, tsa);
    WSlices local;
    const WSlices& ws = planes_for(W_a, a_shape, a_strides, tsa, view, nL, local);
    get_layout(T, tsa, false);
    get_layout(T, tsb, false);                       // (the second call may move the first entry: fetch both again)
    Layout& la = get_layout(T, tsa, false);
    Layout& lb = get_layout(T, tsb, false);
    contract_dense_pair(T, tsa, la, (int)qa, tsb, lb, (int)qb, sigma, ws, view, 0, out, om);
    planes_used();
  } else {
    // one sign matrix, two operands: planes of the combination q_a W_a + q_b W_b, cut straight from the two views (no fp64
    // copy of it); the exact entries take the fp64 remainder on each view
    const WView va = term_view(W_a, a_shape, a_strides, tsa), vb = term_view(W_b, b_shape, b_strides, tsb);
    if (!g_pair_planes) g_pair_planes = new WSlices();   // buffers kept between calls (one geometry at a time)
    WSlices& pair_planes = *g_pair_planes;
    wslices_build_view2(pair_planes, va, qa, vb, qb, 0, nL, tc_planes());
    Layout& lay = get_layout(T, tsa, false);
    dense_form(T, tsa, lay);
    contract_dense_signs(T, lay, pair_planes, 0, out, om, sigma);
    if (lay.n_rest != 0) {
      contract_rest(lay, va, 0, nL, out, om, sigma * qa);
      contract_rest(lay, vb, 0, nL, out, om, sigma * qb);
    }
  }
  FRICC_API_END
}

int fricc_prepare_operand(const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4],
                          const int64_t out_strides[4], const char* term) {
  FRICC_API_BEGIN
  if (!W_dev || !w_shape || !w_strides || !out_strides) fail(FRICC_EINVAL, "null argument");
  const TermSpec* ts0 = find_term(term);
  if (!ts0) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  const TermSpec ts = oriented(*ts0, out_strides);
  const int64_t nK = w_shape[ts.wK[0]] * w_shape[ts.wK[1]], nL = w_shape[ts.wL[0]] * w_shape[ts.wL[1]];
  if (nK == 0 || nL == 0) return FRICC_OK;
  (void)span_of(w_shape, w_strides);
  if (tc_plane_cache() || tc_history_mb() <= 0) return FRICC_OK;   // content cache: nothing to prepare; no history: nowhere to keep it
  const WView view = term_view(W_dev, w_shape, w_strides, ts);
  WSlices local;
  (void)planes_for(W_dev, w_shape, w_strides, ts, view, nL, local, /*prepare_only=*/!is_constant_operand(W_dev));
  FRICC_API_END
}

int fricc_constant_operand(const double* p, int64_t n_elements, int on) {
  FRICC_API_BEGIN
  if (!p || n_elements <= 0) fail(FRICC_EINVAL, "bad arguments");
  const double* lo = p;
  const double* hi = p + n_elements;
  // planes built from anything inside the range are dropped either way: a new declaration starts from the current content
  const_planes_drop(lo, hi);
  for (size_t i = 0; i < g_const_ranges.size();) {
    if (g_const_ranges[i].lo < hi && lo < g_const_ranges[i].hi) g_const_ranges.erase(g_const_ranges.begin() + (long)i);
    else i++;
  }
  if (on) g_const_ranges.push_back(ConstRange{lo, hi});
  FRICC_API_END
}

int fricc_plane_cache_clear(void) {
  FRICC_API_BEGIN
  plane_cache_clear();
  FRICC_API_END
}

int fricc_permute_w(const double* W, int w_mem, const int64_t w_shape[4], const int64_t w_strides[4], const char* term,
                    int64_t l_begin, int64_t l_count, double* Wp_dev) {
  FRICC_API_BEGIN
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  const int64_t nL = w_shape[ts->wL[0]] * w_shape[ts->wL[1]];
  if (l_begin < 0 || l_count < 0 || l_begin + l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  DevIn<double> Wd(W, w_mem, (size_t)span_of(w_shape, w_strides));
  permute_w(Wd.p, w_shape, w_strides, *ts, l_begin, l_count, Wp_dev);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  FRICC_API_END
}

int fricc_contract_panel(const double* Wp_dev, int64_t ldw, int64_t l_begin, int64_t l_count, fricc_st_t T,
                         double* out_panel_dev, int64_t ld_out, const char* term) {
  FRICC_API_BEGIN
  check_handle(T);
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!Wp_dev || !out_panel_dev || l_count < 0 || ldw < l_count || ld_out < l_count) fail(FRICC_EINVAL, "bad arguments");
  if (T->nnz == 0 || l_count == 0) return FRICC_OK;
  // out panel is [R][ld_out] row-major: address = R*ld_out + (l - l_begin)
  OutMap om;
  om.nR2 = 1;
  om.nL2 = 1;
  om.sR1 = ld_out;
  om.sR2 = 0;
  om.sL1 = 1;
  om.sL2 = 0;
  // the kernel maps column l = l_begin + col through sL1; shift the base so that column l_begin lands at offset 0
  spmm_term(T, *ts, Wp_dev, ldw, l_begin, l_count, out_panel_dev - l_begin, om);
  FRICC_API_END
}

int fricc_contract_panel_into(const double* Wp_dev, int64_t ldw, int64_t l_begin, int64_t l_count, fricc_st_t T,
                              double* out_dev, const int64_t out_shape[4], const int64_t out_strides[4],
                              const char* term) {
  FRICC_API_BEGIN
  check_handle(T);
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!Wp_dev || !out_dev || !out_shape || !out_strides || l_begin < 0 || l_count < 0 || ldw < l_count)
    fail(FRICC_EINVAL, "bad arguments");
  const int64_t* d = T->dims;
  for (int q = 0; q < 2; q++)
    if (out_shape[ts->oR[q]] != d[ts->tR[q]]) fail(FRICC_EINVAL, "term %s: out does not match T on the owner pair", ts->name);
  const int64_t nL = out_shape[ts->oL[0]] * out_shape[ts->oL[1]];
  if (l_begin + l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  if (T->nnz == 0 || l_count == 0) return FRICC_OK;
  const OutMap om = out_map(*ts, d, out_shape, out_strides);
  spmm_term(T, *ts, Wp_dev, ldw, l_begin, l_count, out_dev, om);
  FRICC_API_END
}

int fricc_accumulate_panels(const double* panels_dev, int64_t n_panels, int64_t ld, fricc_st_t T, const char* term,
                            double* out_dev, const int64_t out_shape[4], const int64_t out_strides[4]) {
  FRICC_API_BEGIN
  check_handle(T);
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!panels_dev || !out_dev || n_panels <= 0 || ld <= 0) fail(FRICC_EINVAL, "bad arguments");
  const int64_t* d = T->dims;
  for (int q = 0; q < 2; q++)
    if (out_shape[ts->oR[q]] != d[ts->tR[q]]) fail(FRICC_EINVAL, "term %s: out does not match T on the owner pair", ts->name);
  OutMap om;
  om.nR2 = d[ts->tR[1]];
  om.nL2 = out_shape[ts->oL[1]];
  om.sR1 = out_strides[ts->oR[0]];
  om.sR2 = out_strides[ts->oR[1]];
  om.sL1 = out_strides[ts->oL[0]];
  om.sL2 = out_strides[ts->oL[1]];
  const int64_t nR = d[ts->tR[0]] * d[ts->tR[1]];
  const int64_t nL = out_shape[ts->oL[0]] * out_shape[ts->oL[1]];
  accumulate_panels(panels_dev, n_panels, ld, nR, nL, out_dev, om);
  FRICC_API_END
}

// ---- tensor-core form of the contractions (fricc_dense.cu) ---------------------------------------------------
struct fricc_wslices {
  fricc::WSlices ws;
};

int fricc_wslices_create(const double* Wp_dev, int64_t ldw, int64_t nK, int64_t l_count, int n_planes,
                         fricc_wslices_t* out) {
  FRICC_API_BEGIN
  if (!Wp_dev || !out) fail(FRICC_EINVAL, "null argument");

  fricc_wslices* h = new fricc_wslices();
  try {
    wslices_build(h->ws, Wp_dev, ldw, nK, l_count, n_planes);
  } catch (...) {
    delete h;
    throw;
  }
  *out = h;
  FRICC_API_END
}

int fricc_wslices_create_view(const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4], const char* term,
                              int64_t l_begin, int64_t l_count, int n_planes, fricc_wslices_t* out) {
  FRICC_API_BEGIN
  if (!W_dev || !w_shape || !w_strides || !out) fail(FRICC_EINVAL, "null argument");
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  const int64_t nL = w_shape[ts->wL[0]] * w_shape[ts->wL[1]];
  if (l_begin < 0 || l_count <= 0 || l_begin + l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  (void)span_of(w_shape, w_strides);  // rejects negative strides
  fricc_wslices* h = new fricc_wslices();
  try {
    wslices_build_view(h->ws, term_view(W_dev, w_shape, w_strides, *ts), l_begin, l_count, n_planes);
  } catch (...) {
    delete h;
    throw;
  }
  *out = h;
  FRICC_API_END
}

int fricc_wslices_update_view(fricc_wslices_t h, const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4],
                              const char* term, int64_t l_begin) {
  FRICC_API_BEGIN
  if (!h || !W_dev || !w_shape || !w_strides) fail(FRICC_EINVAL, "null argument");
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  const int64_t nL = w_shape[ts->wL[0]] * w_shape[ts->wL[1]];
  if (l_begin < 0 || l_begin + h->ws.l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  if (w_shape[ts->wK[0]] * w_shape[ts->wK[1]] != h->ws.nK) fail(FRICC_EINVAL, "contracted pair does not match the planes");
  wslices_build_view(h->ws, term_view(W_dev, w_shape, w_strides, *ts), l_begin, h->ws.l_count, h->ws.S);
  FRICC_API_END
}

int fricc_wslices_redone(fricc_wslices_t h, int64_t* n_columns) {
  FRICC_API_BEGIN
  if (!h || !n_columns) fail(FRICC_EINVAL, "null argument");
  *n_columns = -1;                       // the last build was not a predicted one
  if (h->ws.redone.n == 1) {
    int v = 0;
    h->ws.redone.to_host(&v, 1);
    *n_columns = v;
  }
  FRICC_API_END
}

int fricc_int8_peak(double seconds, double* tops) {
  FRICC_API_BEGIN
  if (!tops || !(seconds > 0.0)) fail(FRICC_EINVAL, "bad arguments");
  *tops = int8_peak_tops(seconds);
  FRICC_API_END
}

int fricc_wslices_update(fricc_wslices_t h, const double* Wp_dev, int64_t ldw) {
  FRICC_API_BEGIN
  if (!h || !Wp_dev) fail(FRICC_EINVAL, "null argument");
  wslices_build(h->ws, Wp_dev, ldw, h->ws.nK, h->ws.l_count, h->ws.S);
  FRICC_API_END
}

int fricc_wslices_destroy(fricc_wslices_t h) {
  FRICC_API_BEGIN
  delete h;
  FRICC_API_END
}

int fricc_wslices_bytes(fricc_wslices_t h, int64_t* bytes) {
  FRICC_API_BEGIN
  if (!h || !bytes) fail(FRICC_EINVAL, "null argument");
  *bytes = (int64_t)h->ws.q.n;
  FRICC_API_END
}

int fricc_wslices_reconstruct(fricc_wslices_t h, double* Wrec_dev) {
  FRICC_API_BEGIN
  if (!h || !Wrec_dev) fail(FRICC_EINVAL, "null argument");
  wslices_reconstruct(h->ws, Wrec_dev);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  FRICC_API_END
}

int fricc_st_sign_fraction(fricc_st_t T, const char* term, double* fraction) {
  FRICC_API_BEGIN
  check_handle(T);
  const TermSpec* ts = find_term(term);
  if (!ts || !fraction) fail(FRICC_EINVAL, "bad arguments");
  *fraction = T->nnz > 0 ? dense_form(T, *ts, get_layout(T, *ts, false)) : 0.0;
  FRICC_API_END
}

static void contract_tc_common(fricc_wslices_t h, const double* Wp_dev, int64_t ldw, int64_t l_begin, fricc_st_t T,
                               double* out, const OutMap& om, const TermSpec* ts) {
  if (T->nnz == 0 || h->ws.l_count == 0) return;
  if (spmm_exact_order()) {  // parity mode: the reference's own order and roundings on the fp64 panel
    spmm_term(T, *ts, Wp_dev, ldw, l_begin, h->ws.l_count, out, om);
    return;
  }
  Layout& lay = get_layout(T, *ts, false);
  contract_dense(T, *ts, lay, h->ws, Wp_dev, ldw, l_begin, out, om, ts->scale);
}

int fricc_contract_panel_tc(fricc_wslices_t h, const double* Wp_dev, int64_t ldw, int64_t l_begin, fricc_st_t T,
                            double* out_panel_dev, int64_t ld_out, const char* term, int overwrite) {
  FRICC_API_BEGIN
  check_handle(T);
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!h || !Wp_dev || !out_panel_dev || ldw < h->ws.l_count || ld_out < h->ws.l_count) fail(FRICC_EINVAL, "bad arguments");
  OutMap om;
  om.nR2 = 1; om.nL2 = 1; om.sR1 = ld_out; om.sR2 = 0; om.sL1 = 1; om.sL2 = 0;
  const int64_t nR = T->dims[ts->tR[0]] * T->dims[ts->tR[1]];
  bool ow = overwrite != 0;
  if (spmm_exact_order()) {
    if (ow && h->ws.l_count > 0 && nR > 0)
      FRICC_CUDA(cudaMemset2DAsync(out_panel_dev, (size_t)ld_out * 8, 0, (size_t)h->ws.l_count * 8, (size_t)nR, stream()));
    if (T->nnz > 0 && h->ws.l_count > 0)
      spmm_term(T, *ts, Wp_dev, ldw, l_begin, h->ws.l_count, out_panel_dev - l_begin, om);
    return FRICC_OK;
  }
  if (ow) {
    // "out = result": the tensor-core kernel writes every element of the panel columns without reading them.  When it
    // will not run (no entries, or nothing in the sign matrix) the columns are cleared here instead.
    bool gemm_runs = false;
    if (T->nnz > 0 && h->ws.l_count > 0) {
      Layout& lay = get_layout(T, *ts, false);
      dense_form(T, *ts, lay);
      gemm_runs = lay.n_rest != T->nnz;
    }
    if (!gemm_runs) {
      if (h->ws.l_count > 0 && nR > 0)
        FRICC_CUDA(cudaMemset2DAsync(out_panel_dev, (size_t)ld_out * 8, 0, (size_t)h->ws.l_count * 8, (size_t)nR, stream()));
      ow = false;
    }
  }
  if (T->nnz == 0 || h->ws.l_count == 0) return FRICC_OK;
  Layout& lay = get_layout(T, *ts, false);
  contract_dense(T, *ts, lay, h->ws, Wp_dev, ldw, l_begin, out_panel_dev - l_begin, om, ts->scale, ow);
  FRICC_API_END
}

int fricc_contract_panel_into_tc(fricc_wslices_t h, const double* Wp_dev, int64_t ldw, int64_t l_begin, fricc_st_t T,
                                 double* out_dev, const int64_t out_shape[4], const int64_t out_strides[4],
                                 const char* term) {
  FRICC_API_BEGIN
  check_handle(T);
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!h || !Wp_dev || !out_dev || !out_shape || !out_strides || l_begin < 0 || ldw < h->ws.l_count)
    fail(FRICC_EINVAL, "bad arguments");
  const int64_t* d = T->dims;
  for (int q = 0; q < 2; q++)
    if (out_shape[ts->oR[q]] != d[ts->tR[q]]) fail(FRICC_EINVAL, "term %s: out does not match T on the owner pair", ts->name);
  const int64_t nL = out_shape[ts->oL[0]] * out_shape[ts->oL[1]];
  if (l_begin + h->ws.l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  OutMap om;
  om.nR2 = d[ts->tR[1]];
  om.nL2 = out_shape[ts->oL[1]];
  om.sR1 = out_strides[ts->oR[0]];
  om.sR2 = out_strides[ts->oR[1]];
  om.sL1 = out_strides[ts->oL[0]];
  om.sL2 = out_strides[ts->oL[1]];
  contract_tc_common(h, Wp_dev, ldw, l_begin, T, out_dev, om, ts);
  FRICC_API_END
}

int fricc_contract_planes(fricc_wslices_t h, const double* W_dev, const int64_t w_shape[4], const int64_t w_strides[4],
                          int64_t l_begin, fricc_st_t T, double* out_dev, const int64_t out_shape[4],
                          const int64_t out_strides[4], const char* term) {
  FRICC_API_BEGIN
  check_handle(T);
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!h || !W_dev || !out_dev || !w_shape || !w_strides || !out_shape || !out_strides || l_begin < 0)
    fail(FRICC_EINVAL, "bad arguments");
  const int64_t* d = T->dims;
  for (int q = 0; q < 2; q++) {
    if (w_shape[ts->wK[q]] != d[ts->tK[q]]) fail(FRICC_EINVAL, "term %s: W does not match T on the contracted pair", ts->name);
    if (out_shape[ts->oR[q]] != d[ts->tR[q]]) fail(FRICC_EINVAL, "term %s: out does not match T on the owner pair", ts->name);
    if (out_shape[ts->oL[q]] != w_shape[ts->wL[q]]) fail(FRICC_EINVAL, "term %s: out does not match W on the free pair", ts->name);
  }
  const int64_t nL = w_shape[ts->wL[0]] * w_shape[ts->wL[1]];
  if (l_begin + h->ws.l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  if (T->nnz == 0 || h->ws.l_count == 0) return FRICC_OK;
  const OutMap om = out_map(*ts, d, out_shape, out_strides);
  const WView view = term_view(W_dev, w_shape, w_strides, *ts);
  if (spmm_exact_order()) fail(FRICC_EINVAL, "parity mode (spmm_exact_order) runs through fricc_contract");
  Layout& lay = get_layout(T, *ts, false);
  contract_dense(T, *ts, lay, h->ws, view, l_begin, out_dev, om, ts->scale, false);
  FRICC_API_END
}

// ---- stage-wise utilities ------------------------------------------------------------------------------------
int fricc_one_norm(const double* x, int mem, int64_t n, double* norm) {
  FRICC_API_BEGIN
  if (n < 0 || !norm || (n > 0 && !x)) fail(FRICC_EINVAL, "bad arguments");
  DevIn<double> xd(x, mem, (size_t)n);
  *norm = one_norm_sequential(xd.p, n);
  FRICC_API_END
}

int fricc_argsort(const double* x, int mem, int64_t n, int64_t m, int64_t* idx_out) {
  FRICC_API_BEGIN
  if (n < 0 || m < 0 || m > n || (m > 0 && !idx_out)) fail(FRICC_EINVAL, "argsort: need 0 <= m <= N");
  DevIn<double> xd(x, mem, (size_t)n);
  TopM top;
  select_top_m(xd.p, n, m, top);
  if (m > 0) top.idx.to_host(reinterpret_cast<uint64_t*>(idx_out), (size_t)m);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  FRICC_API_END
}

int fricc_get_d_largest(const double* x, int mem, int64_t n, int64_t n_sample, int64_t* D_out, int64_t* n_D,
                        double* remaining_norm) {
  FRICC_API_BEGIN
  if (n < 0 || n_sample < 0 || n_sample > n || !n_D || !remaining_norm) fail(FRICC_EINVAL, "get_d_largest: bad arguments");
  DevIn<double> xd(x, mem, (size_t)n);
  double norm = one_norm_sequential(xd.p, n);
  TopM top;
  select_top_m(xd.p, n, n_sample, top);
  int64_t nD = 0;
  double r = norm;
  if (n_sample > 0) d_greedy(top, n_sample, norm, &nD, &r);
  if (nD > 0 && D_out) top.idx.to_host(reinterpret_cast<uint64_t*>(D_out), (size_t)nD);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  *n_D = nD;
  *remaining_norm = r;
  FRICC_API_END
}

int fricc_make_probability_vector(const double* x, int mem, int64_t n, int64_t n_sample, const int64_t* D_host,
                                  int64_t n_D, double remaining_norm, double* p_out, int p_mem) {
  FRICC_API_BEGIN
  if (n < 0 || n_D < 0 || (n > 0 && (!x || !p_out))) fail(FRICC_EINVAL, "bad arguments");
  DevIn<double> xd(x, mem, (size_t)n);
  for (int64_t i = 0; i < n_D; i++)
    if (D_host[i] < 0 || D_host[i] >= n) fail(FRICC_ERANGE, "D index out of range");
  DevIn<int64_t> Dd(D_host, FRICC_HOST, (size_t)n_D);
  DevBuf<double> pstage;
  double* pd = p_out;
  if (p_mem != FRICC_DEVICE) {
    pstage.alloc((size_t)n);
    pd = pstage.p;
  }
  // copy out even on failure?  no: the reference exits before returning anything
  make_probability_vector(xd.p, n, n_sample, Dd.p, n_D, remaining_norm, pd);
  if (p_mem != FRICC_DEVICE && n > 0) pstage.to_host(p_out, (size_t)n);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  FRICC_API_END
}

static int sample_impl(bool pivotal, int64_t n_sample, const double* probs, int mem, int64_t n, const int32_t* draws,
                       int64_t n_draws, int64_t* n_consumed, int64_t* S_out) {
  FRICC_API_BEGIN
  if (n < 0 || n_sample < 0 || (n > 0 && !probs) || (n_sample > 0 && !S_out)) fail(FRICC_EINVAL, "bad arguments");
  DevIn<double> pd(probs, mem, (size_t)n);
  DrawSource ds;
  ds.from_stream = (draws == nullptr);
  ds.host = draws;
  ds.n = draws ? n_draws : 0;
  DevBuf<int64_t> S((size_t)std::max<int64_t>(n_sample, 1));
  int64_t consumed = 0;
  if (pivotal) sample_pivotal_probs(n_sample, pd.p, n, ds, &consumed, S.p);
  else sample_systematic_probs(n_sample, pd.p, n, ds, &consumed, S.p);
  if (n_sample > 0) S.to_host(S_out, (size_t)n_sample);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  if (n_consumed) *n_consumed = consumed;
  FRICC_API_END
}

int fricc_sample_pivotal(int64_t n_sample, const double* probs, int mem, int64_t n, const int32_t* draws,
                         int64_t n_draws, int64_t* n_consumed, int64_t* S_out) {
  return sample_impl(true, n_sample, probs, mem, n, draws, n_draws, n_consumed, S_out);
}
int fricc_sample_systematic(int64_t n_sample, const double* probs, int mem, int64_t n, const int32_t* draws,
                            int64_t n_draws, int64_t* n_consumed, int64_t* S_out) {
  return sample_impl(false, n_sample, probs, mem, n, draws, n_draws, n_consumed, S_out);
}

int fricc_parallel_sample(int64_t n_sample, const double* probs, int mem, int64_t n, int n_chunks, const int32_t* draws,
                          int64_t n_draws, int64_t* n_consumed, int64_t* S_out) {
  FRICC_API_BEGIN
  if (n < 0 || n_sample < 0 || (n > 0 && !probs) || (n_sample > 0 && !S_out)) fail(FRICC_EINVAL, "bad arguments");
  DevIn<double> pd(probs, mem, (size_t)n);
  DrawSource ds;
  ds.from_stream = (draws == nullptr);
  ds.host = draws;
  ds.n = draws ? n_draws : 0;
  DevBuf<int64_t> S((size_t)std::max<int64_t>(n_sample, 1));
  int64_t consumed = 0;
  parallel_sample(n_sample, pd.p, n, n_chunks, ds, &consumed, S.p);
  if (n_sample > 0) S.to_host(S_out, (size_t)n_sample);
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  if (n_consumed) *n_consumed = consumed;
  FRICC_API_END
}

int fricc_fri_compression(const double* x, int mem, int64_t n, int64_t n_sample, const char* sampling_method,
                          int verbose, const int32_t* draws, int64_t n_draws, int64_t* n_consumed, int64_t* idx_out,
                          double* vals_out) {
  FRICC_API_BEGIN
  if (n < 0 || n_sample < 0 || n_sample > n) fail(FRICC_EINVAL, "fri_compression: need 0 <= n_sample <= N");
  DevIn<double> xd(x, mem, (size_t)n);
  DrawSource ds;
  ds.from_stream = (draws == nullptr);
  ds.host = draws;
  ds.n = draws ? n_draws : 0;
  Compressed c;
  compress_fri(xd.p, n, n_sample, parse_sampling(sampling_method), ds, verbose != 0, c);
  if (n_sample > 0) {
    c.flat.to_host(idx_out, (size_t)n_sample);
    c.vals.to_host(vals_out, (size_t)n_sample);
  }
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  if (n_consumed) *n_consumed = c.consumed;
  FRICC_API_END
}

int fricc_key_histogram(const double* x_dev, int64_t n, int shift, uint64_t prefix, int prefix_bits,
                        unsigned long long* hist_dev) {
  FRICC_API_BEGIN
  if (shift != 0 && shift != 3 && shift != 15 && shift != 27 && shift != 39 && shift != 51)
    fail(FRICC_EINVAL, "shift must be one of 51,39,27,15,3,0");
  key_histogram(x_dev, n, shift, shift == 0 ? 7u : 4095u, prefix, prefix_bits, hist_dev);
  FRICC_API_END
}

// ---- range-sharded compression -------------------------------------------------------------------------------------
struct fricc_shard {
  const double* x = nullptr;
  int64_t n = 0, j0 = 0, n_total = 0;
  PSrc ps{};
  ShardScan* scan = nullptr;
};

int fricc_shard_create(const double* x_dev, int64_t n, int64_t j0, int64_t n_total, fricc_shard_t* out) {
  FRICC_API_BEGIN
  if (!out) fail(FRICC_EINVAL, "null output handle");
  if (n < 0 || j0 < 0 || j0 + n > n_total) fail(FRICC_EINVAL, "shard [j0, j0+n) must lie inside [0, n_total)");
  if (n > 0 && !x_dev) fail(FRICC_EINVAL, "null shard pointer");
  (void)num_sms();  // no device -> FRICC_ECUDA (there is no CPU path)
  fricc_shard* s = new fricc_shard();
  s->x = x_dev; s->n = n; s->j0 = j0; s->n_total = n_total;
  s->ps = PSrc{nullptr, x_dev, 1.0, 0.0, ~0ULL, 0, 1, j0};
  s->scan = shard_scan_create();
  *out = s;
  FRICC_API_END
}

int fricc_shard_destroy(fricc_shard_t s) {
  FRICC_API_BEGIN
  if (s) {
    shard_scan_destroy(s->scan);
    delete s;
  }
  FRICC_API_END
}

static void check_shard(fricc_shard_t s) {
  if (!s) fail(FRICC_EINVAL, "null shard handle");
}

int fricc_shard_use_abs(fricc_shard_t s) {
  FRICC_API_BEGIN
  check_shard(s);
  s->ps = PSrc{nullptr, s->x, 1.0, 0.0, ~0ULL, 0, 1, s->j0};
  FRICC_API_END
}

int fricc_shard_use_probabilities(fricc_shard_t s, double remaining_norm, int64_t n_sample, uint64_t key_cut,
                                  uint64_t idx_cut) {
  FRICC_API_BEGIN
  check_shard(s);
  s->ps = PSrc{nullptr, s->x, remaining_norm, (double)n_sample, key_cut, idx_cut, 0, s->j0};
  FRICC_API_END
}

int fricc_shard_scan_local(fricc_shard_t s, double* approx_total_dev) {
  FRICC_API_BEGIN
  check_shard(s);
  shard_scan_local(s->scan, s->ps, s->n, approx_total_dev);
  FRICC_API_END
}

int fricc_shard_scan_maps(fricc_shard_t s, const double* approx_carry_dev) {
  FRICC_API_BEGIN
  check_shard(s);
  shard_scan_maps(s->scan, approx_carry_dev);
  FRICC_API_END
}

int fricc_shard_scan_resolve(fricc_shard_t s, const double* carry_in_dev, double* carry_out_dev) {
  FRICC_API_BEGIN
  check_shard(s);
  if (!carry_in_dev || !carry_out_dev) fail(FRICC_EINVAL, "null carry pointer");
  shard_scan_resolve(s->scan, carry_in_dev, carry_out_dev);
  FRICC_API_END
}

int fricc_shard_emit_systematic(fricc_shard_t s, int64_t n_sample, int32_t draw, int64_t* S_dev) {
  FRICC_API_BEGIN
  check_shard(s);
  if (s->ps.absval) fail(FRICC_EINVAL, "fricc_shard_use_probabilities first");
  const double rn = (double)draw / 2147483647.0;  // (double)rand() / RAND_MAX   (fri_utils.cpp:47)
  shard_scan_emit(s->scan, n_sample, rn, S_dev);
  FRICC_API_END
}

int fricc_shard_emit_systematic_multi(fricc_shard_t s, int64_t n_sample, const int32_t* draws, int n_draws, int64_t* S_dev,
                                      int64_t stride) {
  FRICC_API_BEGIN
  check_shard(s);
  if (s->ps.absval) fail(FRICC_EINVAL, "fricc_shard_use_probabilities first");
  if (n_draws < 0 || (n_draws > 0 && (!draws || !S_dev))) fail(FRICC_EINVAL, "bad arguments");
  std::vector<double> rn((size_t)n_draws);
  for (int d = 0; d < n_draws; d++) rn[d] = (double)draws[d] / 2147483647.0;  // (double)rand() / RAND_MAX   (fri_utils.cpp:47)
  shard_scan_emit_multi(s->scan, n_sample, rn.data(), n_draws, S_dev, stride);
  FRICC_API_END
}

int fricc_shard_sample_chunk(fricc_shard_t s, int64_t budget, double t, double a, double chunk_norm, const int32_t* draws,
                             int64_t n_draws, int64_t* n_consumed, int64_t* S_dev) {
  FRICC_API_BEGIN
  check_shard(s);
  if (s->ps.absval) fail(FRICC_EINVAL, "fricc_shard_use_probabilities first");
  if (budget > 0 && !S_dev) fail(FRICC_EINVAL, "null sample buffer");
  DrawSource ds;
  ds.from_stream = (draws == nullptr);
  ds.host = draws;
  ds.n = draws ? n_draws : 0;
  int64_t used = 0;
  sample_chunk(s->ps, s->n, s->j0, budget, t, a, chunk_norm, ds, &used, S_dev);
  if (n_consumed) *n_consumed = used;
  FRICC_API_END
}

int fricc_shard_check(fricc_shard_t s) {
  FRICC_API_BEGIN
  check_shard(s);
  raise_error_flags(shard_scan_errors(s->scan));
  FRICC_API_END
}

int fricc_shard_candidates(fricc_shard_t s, uint64_t key_min, uint64_t* keys_out_dev, uint64_t* idx_out_dev, int64_t cap,
                           unsigned long long* count_dev) {
  FRICC_API_BEGIN
  check_shard(s);
  if (cap < 0 || !count_dev) fail(FRICC_EINVAL, "bad candidate buffer");
  shard_candidates(s->x, s->n, s->j0, key_min, keys_out_dev, idx_out_dev, cap, count_dev);
  FRICC_API_END
}

int fricc_sort_candidates(uint64_t* keys_dev, uint64_t* idx_dev, int64_t count, int64_t n_total) {
  FRICC_API_BEGIN
  (void)num_sms();  // no device -> FRICC_ECUDA (there is no CPU path)
  sort_candidates(keys_dev, idx_dev, count, n_total);
  FRICC_API_END
}

int fricc_d_greedy(const uint64_t* keys_sorted_dev, int64_t n_candidates, int64_t n_sample, double norm, int64_t* n_D,
                   double* remaining_norm) {
  FRICC_API_BEGIN
  (void)num_sms();  // no device -> FRICC_ECUDA (there is no CPU path)
  if (!n_D || !remaining_norm) fail(FRICC_EINVAL, "null output");
  *n_D = 0;
  *remaining_norm = norm;
  if (n_candidates > 0) {
    if (n_candidates >= (1 << 15)) d_greedy_parallel(keys_sorted_dev, n_candidates, n_sample, norm, n_D, remaining_norm);
    else d_greedy_serial(keys_sorted_dev, 0, n_candidates, n_sample, norm, n_D, remaining_norm);
  }
  FRICC_API_END
}

int fricc_shard_values(fricc_shard_t s, const int64_t* flat_dev, int64_t m, int64_t n_D, double* vals_dev) {
  FRICC_API_BEGIN
  check_shard(s);
  if (s->ps.absval) fail(FRICC_EINVAL, "fricc_shard_use_probabilities first");
  shard_values(s->ps, s->n, flat_dev, m, n_D, vals_dev);
  FRICC_API_END
}

}  // extern "C"
