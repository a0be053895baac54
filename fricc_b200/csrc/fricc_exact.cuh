// fricc_exact.cuh -- the algebra that lets a LEFT-TO-RIGHT fp64 running sum be computed by a PARALLEL scan, bit for bit.
//
// Why: both samplers of the reference, and its one-norm, are sequential fp64 recurrences
//      c_j = fl(c_{j-1} + p_j),  p_j >= 0                       (/root/reference/src/fri_utils.cpp:73,143,169)
// and a tree-ordered parallel sum rounds differently (tens of differing samples out of 1e7 at N=1e8).
//
// How: while the running sum stays inside one binade [2^k, 2^(k+1)) it is an integer multiple C of ulp_k = 2^(k-52), and
//      fl(c + p) = (C + RN(p / ulp_k)) * ulp_k
// where RN is round-to-nearest and the tie case (frac(p/ulp_k) == 1/2 exactly) rounds to the EVEN neighbour of C + p/ulp_k,
// i.e. depends only on the parity of the running integer.  So each element acts on the integer state as one of
//      plain(A):    S -> S + A
//      even(A, B):  S -> 2*ceil((S + A) / 2) + B
// and these maps are closed under composition (below) -- an associative monoid, hence a normal parallel scan.  The model is
// exact as long as the result stays below 2^53 (i.e. inside the binade); the first element whose result reaches 2^53 is the
// binade crossing: it is redone with one real fp64 addition and the scan restarts in the new binade.
//
// Everything here is __host__ __device__ so that tests/host/exact_scan_host.cpp can exercise the same code on the CPU.
#pragma once
#include <stdint.h>
#include <string.h>

#ifdef __CUDACC__
#define FX_HD __host__ __device__ __forceinline__
#else
#define FX_HD inline
#endif

namespace fricc {

constexpr int64_t FX_TWO52 = (int64_t)1 << 52;
constexpr int64_t FX_TWO53 = (int64_t)1 << 53;

struct EMap {
  int64_t A;
  int64_t B;
  int32_t even;  // 0: S -> S + A          1: S -> 2*ceil((S+A)/2) + B
};

FX_HD EMap emap_identity() { return EMap{0, 0, 0}; }

FX_HD int64_t sat_add(int64_t a, int64_t b) {  // operands are >= 0; saturate far above 2^53 so overflow cannot wrap
  int64_t s = a + b;
  const int64_t cap = (int64_t)1 << 60;
  return s > cap ? cap : s;
}

// h = g o f   (f is applied first)
FX_HD EMap emap_compose(const EMap& f, const EMap& g) {
  EMap h;
  if (!f.even) {
    h.even = g.even;
    h.A = sat_add(f.A, g.A);
    h.B = g.B;
  } else if (!g.even) {
    h.even = 1;
    h.A = f.A;
    h.B = sat_add(f.B, g.A);
  } else {
    h.even = 1;
    h.A = f.A;
    int64_t t = sat_add(f.B, g.A);
    h.B = sat_add(((t + 1) >> 1) << 1, g.B);
  }
  return h;
}

FX_HD int64_t emap_apply(const EMap& f, int64_t S) {
  int64_t t = sat_add(S, f.A);
  if (!f.even) return t;
  return sat_add(((t + 1) >> 1) << 1, f.B);
}

FX_HD uint64_t fx_bits(double x) {
#ifdef __CUDA_ARCH__
  return (uint64_t)__double_as_longlong(x);
#else
  uint64_t u;
  memcpy(&u, &x, 8);
  return u;
#endif
}
FX_HD double fx_from_bits(uint64_t u) {
#ifdef __CUDA_ARCH__
  return __longlong_as_double((long long)u);
#else
  double d;
  memcpy(&d, &u, 8);
  return d;
#endif
}

// unbiased exponent k of a positive NORMAL double c, i.e. c in [2^k, 2^(k+1)).  (Subnormal running sums are treated as
// "no binade" by the callers: they fall back to real additions until the sum is normal.)
FX_HD int fx_binade(double c) { return (int)((fx_bits(c) >> 52) & 0x7ff) - 1023; }
FX_HD bool fx_is_normal_positive(double c) {
  uint64_t e = (fx_bits(c) >> 52) & 0x7ff;
  return c > 0.0 && e != 0 && e != 0x7ff;
}

// running sum c (in binade k) <-> integer state C = c / ulp_k in [2^52, 2^53)
FX_HD int64_t fx_to_state(double c) { return (int64_t)((fx_bits(c) & 0x000fffffffffffffULL) | 0x0010000000000000ULL); }
FX_HD double fx_from_state(int64_t C, int k) {  // requires 2^52 <= C < 2^53
  return fx_from_bits(((uint64_t)(k + 1023) << 52) | ((uint64_t)C & 0x000fffffffffffffULL));
}

// the map of one addend p >= 0 when the running sum is in binade k
FX_HD EMap emap_of(double p, int k) {
  uint64_t b = fx_bits(p) & 0x7fffffffffffffffULL;
  uint64_t ebits = b >> 52;
  uint64_t mant = b & 0x000fffffffffffffULL;
  int e;
  if (ebits == 0) {
    e = -1022;  // subnormal (or zero): no hidden bit
  } else {
    mant |= 0x0010000000000000ULL;
    e = (int)ebits - 1023;
  }
  EMap m{0, 0, 0};
  if (mant == 0) return m;
  if (ebits == 0x7ff) {  // inf / nan: force a crossing, the real addition reproduces whatever IEEE does
    m.A = FX_TWO53;
    return m;
  }
  // p = mant * 2^(e-52);  p / ulp_k = mant * 2^(e-k)
  int sh = k - e;
  if (sh <= 0) {  // p >= 2^k: the sum certainly leaves the binade
    m.A = FX_TWO53;
    return m;
  }
  if (sh >= 64) return m;  // far below half an ulp
  uint64_t f = mant >> sh;
  uint64_t rem = mant & ((1ULL << sh) - 1ULL);
  uint64_t half = 1ULL << (sh - 1);
  if (rem < half) {
    m.A = (int64_t)f;
  } else if (rem > half) {
    m.A = (int64_t)f + 1;
  } else {
    m.even = 1;
    m.A = (int64_t)f;
  }
  return m;
}

// the map used while the running sum is exactly 0: 0 + p = p is exact, so the first non-zero addend "crosses"
FX_HD EMap emap_of_zero_state(double p) {
  EMap m{0, 0, 0};
  if (p != 0.0) m.A = FX_TWO53;
  return m;
}

}  // namespace fricc
