// fricc_dense.cuh -- tensor-core form of the contractions (see fricc_dense.cu).
#pragma once
#include "fricc_contract.cuh"

namespace fricc {

// Error-free int8 digit expansion of a W panel (columns [0, l_count) of Wp[K][ldw]):
//   W[K][l] + 2^e_l = 2^(e_l - 7) * sum_s digit_s[K][l] * 256^(-s)  (+ a tail below 2^(e_l - 8*S)),   0 <= digit <= 255,
// e_l = exponent of the largest magnitude in column l.  Stored K-major per 64-column tile:
//   q[((tile*S + s)*64 + col)][Kpad]  (int8)
struct WSlices {
  int64_t nK = 0, Kpad = 0, l_count = 0, n_tiles = 0;
  int S = 0;
  DevBuf<int8_t> q;
  DevBuf<int32_t> expo;      // [n_tiles*64]
  DevBuf<double> colscale;   // [n_tiles*64]  2^(e_l - 7)
};

constexpr int TC_MAX_SLICES = 7;

void wslices_build(WSlices& ws, const double* Wp_dev, int64_t ldw, int64_t nK, int64_t l_count, int S);
// Wrec[K][l_count] = the value the digits represent (fp64), for tests of the expansion
void wslices_reconstruct(const WSlices& ws, double* Wrec_dev);

// Builds (once per layout) the int8 sign matrix + CSR remainder; returns the fraction of entries in the sign matrix.
double dense_form(fricc_st* T, const TermSpec& ts, Layout& lay);

// out[R, l] += scale * sum_K T[R,K] * W[K,l]   for the panel columns of `ws` (panel-local column j <-> l = l_begin + j):
// sign matrix x digit slices on the int8 tensor cores (tcgen05.mma.kind::i8, exact int32 accumulation), the remainder
// through the fp64 SpMM on Wp_dev.
void contract_dense(fricc_st* T, const TermSpec& ts, Layout& lay, const WSlices& ws, const double* Wp_dev, int64_t ldw,
                    int64_t l_begin, double* out_dev, const OutMap& om, double scale, bool overwrite = false);

void set_tc_option(const char* key, int value);

// CSR SpMM on raw arrays (fricc_contract.cu)
void spmm_csr(const int64_t* rowptr, const uint32_t* kidx, const double* vals, int64_t nR, const double* Wp_dev,
              int64_t ldw, int64_t l_begin, int64_t l_count, double* out_dev, const OutMap& om, double scale);

}  // namespace fricc
