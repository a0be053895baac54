// fricc_select.cuh -- top-m selection by magnitude (radix-histogram select) -- see fricc_select.cu
#pragma once
#include "fricc_common.cuh"
#include "fricc_sort.cuh"

namespace fricc {

// the m largest-|x| elements, ordered by (|x| descending, flat index ascending)
struct TopM {
  int64_t m = 0;
  DevBuf<uint64_t> keys;  // bit pattern of |x|
  DevBuf<uint64_t> idx;   // flat index
};

void select_top_m(const double* x_dev, int64_t n, int64_t m, TopM& out);

// one histogram pass (also the building block of the distributed select: the caller all-reduces hist_dev)
void key_histogram(const double* x_dev, int64_t n, int shift, unsigned mask, uint64_t prefix, int prefix_bits,
                   unsigned long long* hist_dev);

std::vector<Digit> candidate_digits(int64_t n);

// upper estimate of how many of the largest magnitudes the greedy D test can accept (sizes compress_fri's candidate list)
int64_t estimate_d_candidates(const double* x_dev, int64_t n, int64_t n_sample, double norm);

}  // namespace fricc
