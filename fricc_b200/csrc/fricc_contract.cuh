// fricc_contract.cuh -- bucketed sparse-row x dense-panel contractions -- see fricc_contract.cu
#pragma once
#include "fricc_common.cuh"

namespace fricc {

// One contraction term of the reference's dispatcher (/root/reference/src/sparse_tensor.cpp:349-382), expressed as
//      out[R, L] += scale * sum_K T[R, K] * W[K, L]
// by naming which axes of T (dims o,o,v,v), of W and of out play the roles R (owner pair), K (contracted pair) and
// L (free pair).  Each pair is ordered (major, minor).
struct TermSpec {
  const char* name;
  int tR[2];   // T axes forming R
  int tK[2];   // T axes forming K
  int wK[2];   // W axes carrying K (same order as tK)
  int wL[2];   // W axes carrying L
  int oR[2];   // out axes carrying R (same order as tR)
  int oL[2];   // out axes carrying L (same order as wL)
  double scale;
  const char* einsum;  // the caller's formula, for documentation and error messages
};

const TermSpec* find_term(const char* name);

// bucketed layout of T for owner pair (tR) / contracted pair (tK); cached in the handle.  need_csr = false returns the
// entry without sorting the entries (the tensor-core form of an FRI compression does not need the full CSR)
Layout& get_layout(fricc_st* T, const TermSpec& ts, bool need_csr = true);
void build_csr(fricc_st* T, const TermSpec& ts, int64_t first, int64_t count, int64_t nR, int64_t nK,
               DevBuf<int64_t>& rowptr, DevBuf<uint32_t>& kidx, DevBuf<double>& vals);

// W (4-D, element strides) -> panel Wp[K][l_count] for the L columns [l_begin, l_begin + l_count)
void permute_w(const double* W_dev, const int64_t shape[4], const int64_t strides[4], const TermSpec& ts,
               int64_t l_begin, int64_t l_count, double* Wp_dev);

// out[r1*sr1 + r2*sr2 + l1*sl1 + l2*sl2] += scale * sum_e vals[e] * Wp[kidx[e]][l - l_begin]
struct OutMap {
  int64_t nR2, nL2;      // minor extents of the R and L pairs
  int64_t sR1, sR2, sL1, sL2;  // element strides of out along (R major, R minor, L major, L minor)
};
void spmm(const Layout& lay, const double* Wp_dev, int64_t ldw, int64_t l_begin, int64_t l_count, double* out_dev,
          const OutMap& om, double scale);

// The contraction of one term on a [K][ldw] panel, in whichever form the "spmm_exact_order" option selects:
//   0 (default): (R,K)-sorted buckets, FMA accumulation from zero, one read-modify-write per output element;
//   1: the reference's own order and roundings -- entries of a bucket in storage order, out += (scale*W)*value with a
//      rounded product and a rounded sum per entry (sparse_tensor.cpp:66-73) -- bit-identical to the compiled reference.
void spmm_term(fricc_st* T, const TermSpec& ts, const double* Wp_dev, int64_t ldw, int64_t l_begin, int64_t l_count,
               double* out_dev, const OutMap& om);
bool spmm_exact_order();

// tuning switches of the SpMM launch ("spmm_nv2": double2 per thread 0/1/2/4, "spmm_threads")
void set_spmm_option(const char* key, int value);

void accumulate_panels(const double* panels_dev, int64_t n_panels, int64_t ld, int64_t nR, int64_t nL, double* out_dev,
                       const OutMap& om);

}  // namespace fricc
