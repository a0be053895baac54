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
  bool expo_valid = false;   // expo / colscale describe planes of this geometry (they can predict the next build's)
  DevBuf<int> redone;        // [1] columns whose exponent moved in the last predicted build (device counter)
};
int64_t planes_predicted_builds();
int tc_history_mb();

constexpr int TC_MAX_SLICES = 7;

// The [K][L] operand of a term as a VIEW of the caller's 4-D array: element (K, l) -- K = k1*nK2 + k2 the contracted
// pair, l = l1*nL2 + l2 the free pair, both GLOBAL indices -- is p[k1*sK1 + k2*sK2 + l1*sL1 + l2*sL2] (element strides).
struct WView {
  const double* p;
  int64_t sK1, sK2, sL1, sL2;
  int64_t nK1, nK2, nL1, nL2;
};
// a [K][ldw] row-major panel holding the columns [l_begin, ...) of an nL-column operand
WView panel_view(const double* Wp, int64_t ldw, int64_t nK, int64_t l_begin, int64_t nL);
// the operand of `ts` inside a 4-D array with the given shape / element strides
WView term_view(const double* W, const int64_t shape[4], const int64_t strides[4], const TermSpec& ts);
bool view_is_panel(const WView& w);
// planes of the columns [l_begin, l_begin + l_count) of a view, read where the operand lies (no fp64 panel).  Operands
// whose contracted run is contiguous (the ladder's native W[a,b,c,d]) go through a single-read cluster kernel.
void wslices_build_view(WSlices& ws, const WView& w, int64_t l_begin, int64_t l_count, int S);
void wslices_build_view2(WSlices& ws, const WView& w, double q1, const WView& w2, double q2, int64_t l_begin, int64_t l_count,
                         int S);
double dense_form(fricc_st* T, const TermSpec& ts, Layout& lay);
void contract_dense_signs(fricc_st* T, Layout& lay, const WSlices& ws, int64_t l_begin, double* out, const OutMap& om, double scale);
void contract_rest(Layout& lay, const WView& w, int64_t l_begin, int64_t l_count, double* out, const OutMap& om, double scale);
void panel_axpby(double* dst, const double* a, double qa, const double* b, double qb, int64_t n);
// two terms on the same operand view and output map, (R, K) from different axes of T: one dense product (fricc_dense.cu)
void contract_dense_pair(fricc_st* T, const TermSpec& ts1, Layout& lay1, int q1, const TermSpec& ts2, Layout& lay2, int q2,
                         double sigma, const WSlices& ws, const WView& w, int64_t l_begin, double* out, const OutMap& om);
void contract_dense(fricc_st* T, const TermSpec& ts, Layout& lay, const WSlices& ws, const WView& w, int64_t l_begin,
                    double* out_dev, const OutMap& om, double scale, bool overwrite = false);
void spmm_csr_view(const int64_t* rowptr, const uint32_t* kidx, const double* vals, int64_t nR, const WView& w,
                   int64_t l_begin, int64_t l_count, double* out_dev, const OutMap& om, double scale);
// int8 TOP/s of back-to-back tcgen05.mma.kind::i8 on resident operands, over a run of about `seconds`
double int8_peak_tops(double seconds);
int tc_mode();         // option "tc_mode": 0 never / 1 cost model / 2 always -- what fricc_contract does
int tc_planes();       // option "tc_planes"
int tc_plane_cache();  // option "tc_plane_cache"

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
