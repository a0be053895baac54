// fricc_sort.cuh -- hand-written stable LSD radix sort of (uint64 key, uint64 payload) pairs and a device scan.
#pragma once
#include "fricc_common.cuh"

namespace fricc {

// One 8-bit digit of the sort: taken from the key (src=0) or from the payload (src=1) at bit `shift`;
// invert=1 sorts that digit in descending order.
struct Digit {
  int src;
  int shift;
  int invert;
};

// Sorts n pairs held in (k[0], v[0]) using (k[1], v[1]) as the ping-pong buffer, digit by digit in the order given
// (least significant first).  Digits on which all items agree are skipped.  Returns which buffer (0/1) holds the result.
int radix_sort_pairs(uint64_t* k[2], uint64_t* v[2], int64_t n, const std::vector<Digit>& digits);

// exclusive prefix sum of n uint32 values, in place (n up to a few million; single pass over three kernels)
void exclusive_scan_u32(uint32_t* d, int64_t n);
// exclusive prefix sum of n int64 values, in place
void exclusive_scan_i64(int64_t* d, int64_t n);
// exclusive prefix sum of n doubles, in place (tree order: an APPROXIMATION of the left-to-right sum)
void exclusive_scan_f64(double* d, int64_t n);

}  // namespace fricc
