// tests/host/exact_scan_host.cpp -- CPU harness for fricc_b200/csrc/fricc_exact.cuh (the same header the kernels use).
// Mimics the GPU chunk algorithm: groups of E elements are composed into one map each ("threads"), the group maps are
// prefix-composed ("block scan"), every group replays its elements from its exact start state, the first state >= 2^53
// is a binade crossing that is redone with one real addition.  Compared by tests/test_exact_scan_host.py against the
// plain left-to-right sum.
#include <cstdint>
#include <vector>

#include "../../fricc_b200/csrc/fricc_exact.cuh"

using namespace fricc;

extern "C" int exact_running_sum(const double* p, int64_t n, int chunk, int group, double* out, int64_t* n_cross) {
  double c = 0.0;
  int64_t pos = 0, crossings = 0;
  std::vector<EMap> gm;
  std::vector<int64_t> st((size_t)chunk);
  while (pos < n) {
    int64_t end = (pos / chunk + 1) * chunk;
    if (end > n) end = n;
    const bool zero = (c == 0.0);
    if (!zero && !fx_is_normal_positive(c)) {  // subnormal running sum: real additions
      c = c + p[pos];
      out[pos++] = c;
      continue;
    }
    const int k = zero ? 0 : fx_binade(c);
    const int64_t C0 = zero ? 0 : fx_to_state(c);
    const int64_t cnt = end - pos;
    const int64_t ng = (cnt + group - 1) / group;
    gm.assign((size_t)ng, emap_identity());
    for (int64_t g = 0; g < ng; g++) {  // "thread" maps
      EMap m = emap_identity();
      for (int64_t j = g * group; j < cnt && j < (g + 1) * group; j++)
        m = emap_compose(m, zero ? emap_of_zero_state(p[pos + j]) : emap_of(p[pos + j], k));
      gm[(size_t)g] = m;
    }
    EMap run = emap_identity();  // exclusive prefix composition
    int64_t first_cross = -1;
    for (int64_t g = 0; g < ng; g++) {
      int64_t S = emap_apply(run, C0);
      for (int64_t j = g * group; j < cnt && j < (g + 1) * group; j++) {
        S = emap_apply(zero ? emap_of_zero_state(p[pos + j]) : emap_of(p[pos + j], k), S);
        st[(size_t)j] = S;
        if (S >= FX_TWO53 && first_cross < 0) first_cross = j;
      }
      run = emap_compose(run, gm[(size_t)g]);
    }
    int64_t valid = first_cross < 0 ? cnt : first_cross;
    for (int64_t j = 0; j < valid; j++) out[pos + j] = zero ? 0.0 : fx_from_state(st[(size_t)j], k);
    if (first_cross >= 0) {
      double prev = first_cross == 0 ? c : out[pos + first_cross - 1];
      c = prev + p[pos + first_cross];
      out[pos + first_cross] = c;
      pos += first_cross + 1;
      crossings++;
    } else {
      c = out[end - 1];
      pos = end;
    }
  }
  if (n_cross) *n_cross = crossings;
  return 0;
}

// composed map of a whole block under an assumed binade, then applied to an exact start (the E2/E3 path of the GPU scan)
extern "C" int block_map_apply(const double* p, int64_t n, double c_in, double* c_out, int* valid) {
  int k = fx_binade(c_in);
  EMap m = emap_identity();
  for (int64_t j = 0; j < n; j++) m = emap_compose(m, emap_of(p[j], k));
  int64_t S = emap_apply(m, fx_to_state(c_in));
  *valid = S < FX_TWO53;
  *c_out = *valid ? fx_from_state(S, k) : 0.0;
  return 0;
}
