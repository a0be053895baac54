// oracle/ref_adapter.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product library).
//
// A thin C-ABI shim of OUR OWN around the reference's C++ functions so that tests can reach the
// stage-wise FRI utilities that the reference's pybind11 surface declares but cannot call from
// Python (std::span arguments, /root/reference/src/py_rccsd.cpp:17-30).  This translation unit is
// compiled TOGETHER WITH the untouched reference sources, in place, by oracle/Makefile; no
// reference source is copied into this repository.
//
// It also provides __wrap_rand (link flag -Wl,--wrap=rand): the reference draws its uniforms with
// libc rand() (/root/reference/src/fri_utils.cpp:47,74); with the wrap in place the same integer
// stream can be fed to the reference and to the CUDA implementation.
#include <omp.h>

#include <cstdlib>
#include <cstring>
#include <span>
#include <string>
#include <vector>

#include <fri_utils.hpp>  // the reference's header (include path set by the Makefile)

// defined with external linkage at /root/reference/src/fri_utils.cpp:164 but not declared in the header
double one_norm(std::span<double const> q);

extern "C" int __real_rand(void);

namespace {
std::vector<int> g_draws;
size_t g_pos = 0;
bool g_inject = false;
long g_calls = 0;
}  // namespace

extern "C" {

// rand() as seen by every object linked into this .so
int __wrap_rand(void) {
  g_calls++;
  if (g_inject) {
    if (g_pos < g_draws.size()) return g_draws[g_pos++];
    return 0;  // stream exhausted: deterministic filler
  }
  return __real_rand();
}

void ref_set_draws(const int* draws, long n) {
  g_draws.assign(draws, draws + n);
  g_pos = 0;
  g_inject = true;
  g_calls = 0;
}
void ref_clear_draws(void) {
  g_inject = false;
  g_draws.clear();
  g_pos = 0;
  g_calls = 0;
}
long ref_rand_calls(void) { return g_calls; }
void ref_reset_rand_calls(void) { g_calls = 0; }

// OpenMP threads of the reference for the calls that follow (the contraction kernels give the same bits for any
// thread count: one owner thread per output pair, sparse_tensor.cpp:62-64; one_norm does not)
void ref_set_num_threads(int n) { omp_set_num_threads(n > 0 ? n : 1); }
int ref_get_max_threads(void) { return omp_get_max_threads(); }

double ref_one_norm(const double* x, long n) { return one_norm(std::span<const double>(x, (size_t)n)); }

void ref_argsort(const double* x, long n, long m, long* out) {
  auto v = argsort(std::span<const double>(x, (size_t)n), (size_t)m);
  for (long i = 0; i < m; i++) out[i] = (long)v[i];
}

long ref_get_d_largest(const double* x, long n, long n_sample, long* D, double* remaining_norm) {
  auto [d, r] = get_d_largest(std::span<const double>(x, (size_t)n), (size_t)n_sample);
  for (size_t i = 0; i < d.size(); i++) D[i] = (long)d[i];
  *remaining_norm = r;
  return (long)d.size();
}

void ref_make_probability_vector(const double* x, long n, long n_sample, const long* D, long nD,
                                 double remaining_norm, double* p) {
  std::vector<size_t> d(D, D + nD);
  auto v = make_probability_vector(std::span<const double>(x, (size_t)n), (size_t)n_sample, d, remaining_norm);
  std::memcpy(p, v.data(), sizeof(double) * (size_t)n);
}

void ref_sample_pivotal(long n_sample, const double* p, long n, long* S) {
  std::vector<double> probs(p, p + n);
  auto s = sample_pivotal((size_t)n_sample, probs);
  for (size_t i = 0; i < s.size(); i++) S[i] = (long)s[i];
}

void ref_sample_systematic(long n_sample, const double* p, long n, long* S) {
  std::vector<double> probs(p, p + n);
  auto s = sample_systematic((size_t)n_sample, probs);
  for (size_t i = 0; i < s.size(); i++) S[i] = (long)s[i];
}

// parallel_sample (fri_utils.cpp:176-362) cuts q into omp_get_max_threads() chunks and lets every thread call rand()
// concurrently: with more than one thread the ORDER of the draws is a race, so only two cases are reproducible --
// one thread (any draws), or any thread count with a constant draw stream.  n_threads is set for this call only.
// Returns the number of samples written.
long ref_parallel_sample(long n_sample, const double* q, long n, int n_threads, long* S) {
  const int before = omp_get_max_threads();
  omp_set_num_threads(n_threads);
  auto s = parallel_sample((size_t)n_sample, std::span<const double>(q, (size_t)n));
  omp_set_num_threads(before);
  for (size_t i = 0; i < s.size(); i++) S[i] = (long)s[i];
  return (long)s.size();
}

void ref_fri_compression(const double* x, long n, long n_sample, const char* method, int verbose, long* idx,
                         double* vals) {
  auto [ci, cv] = fri_compression(std::span<const double>(x, (size_t)n), (size_t)n_sample, std::string(method),
                                  verbose != 0);
  for (long i = 0; i < n_sample; i++) {
    idx[i] = (long)ci[i];
    vals[i] = cv[i];
  }
}

}  // extern "C"
