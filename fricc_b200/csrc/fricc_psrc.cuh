// fricc_psrc.cuh -- the probability source shared by the sampler kernels, and the kernel error flags.
#pragma once
#include "fricc_common.cuh"

namespace fricc {

// error flag bits raised by kernels
constexpr int ERRF_PROB = 1;     // some p > 1                       (fri_utils.cpp:383-395)
constexpr int ERRF_ACCEPT = 2;   // acceptance prob. not in [0,1]    (fri_utils.cpp:84-94)
constexpr int ERRF_DRAWS = 4;    // draw buffer too short

// probability source: an explicit vector, or computed on the fly from x exactly as make_probability_vector does
// (/root/reference/src/fri_utils.cpp:367-398) -- the compression path never materialises p.
// Requires -fmad=false in the including translation unit (no contraction of the divide/multiply).
struct PSrc {
  const double* probs;  // explicit probabilities (stage-wise API) or nullptr
  const double* x;      // dense tensor
  double r;             // remaining norm
  double nprime;        // (double)(m - |D|)
  uint64_t key_cut;     // D = { key > key_cut } U { key == key_cut and idx <= idx_cut };  no D: key_cut = ~0
  uint64_t idx_cut;
  int absval;           // 1: return |x| (one_norm)
  int64_t j0 = 0;       // flat index of x[0] (range-sharded compression: x is the shard [j0, j0+n) of the tensor)
  // the addend of element j given its value (x[j], or probs[j] for an explicit vector)
  __device__ __forceinline__ double of_value(double xv, int64_t j) const {
    if (probs) return xv;
    if (absval) return fabs(xv);
    uint64_t key = abs_key(xv);
    bool inD = key > key_cut || (key == key_cut && (uint64_t)(j + j0) <= idx_cut);
    if (inD) return 0.0;
    return __dmul_rn(__ddiv_rn(fabs(xv), r), nprime);
  }
  __device__ __forceinline__ const double* values() const { return probs ? probs : x; }
  __device__ __forceinline__ double operator()(int64_t j) const {
    if (probs) return probs[j];
    double xv = x[j];
    if (absval) return fabs(xv);
    uint64_t key = abs_key(xv);
    bool inD = key > key_cut || (key == key_cut && (uint64_t)(j + j0) <= idx_cut);
    if (inD) return 0.0;
    // abs(x[i]) / remaining_norm * n_sample   (fri_utils.cpp:375), left to right, no contraction
    return __dmul_rn(__ddiv_rn(fabs(xv), r), nprime);
  }
};

// smallest i in [0, nsamp] with fl((double)i + rn) >= X   (uk_i = i*1.0 + rn, fri_utils.cpp:45-58)
__device__ __forceinline__ int64_t first_sample_at_or_above(double X, double rn, int64_t nsamp) {
  double g = __dsub_rn(X, rn);
  int64_t i;
  if (!(g > 0.0)) i = 0;
  else if (g >= (double)nsamp) i = nsamp;
  else i = (int64_t)ceil(g);
  while (i > 0 && __dadd_rn((double)(i - 1), rn) >= X) i--;
  while (i < nsamp && __dadd_rn((double)i, rn) < X) i++;
  return i;
}

// exact (left-to-right order) sums computed by the parallel binade-segmented scan  (fricc_escan.cu)
double exact_sum_parallel(const PSrc& ps, int64_t n, int* err_flags_out);
void systematic_parallel(const PSrc& ps, int64_t n, int64_t nsamp, double rn, int64_t* S_dev, int* err_flags_out);

// The same scan cut into its stages for a tensor sharded by contiguous flat-index ranges over several ranks
// (fricc_shard.cu): the stages between which the caller exchanges one double per rank.  All scalars stay on the device,
// so a rank-to-rank chain of carries is ordered by the stream alone.
struct ShardScan;
ShardScan* shard_scan_create();
void shard_scan_destroy(ShardScan* s);
// E1: approximate chunk sums of the shard; *approx_total_dev = approximate sum of the whole shard
void shard_scan_local(ShardScan* s, const PSrc& ps, int64_t n, double* approx_total_dev);
// E2: chunk maps in the binades predicted by (*approx_carry_dev + approximate running sum inside the shard)
void shard_scan_maps(ShardScan* s, const double* approx_carry_dev);
// E3: exact running sums at chunk starts from the exact sum *carry_in_dev of everything before the shard;
// *carry_out_dev = exact sum behind the shard (may alias carry_in_dev)
void shard_scan_resolve(ShardScan* s, const double* carry_in_dev, double* carry_out_dev);
// E4: systematic samples owned by the shard's elements, written as GLOBAL flat indices into S_dev[nsamp]
void shard_scan_emit(ShardScan* s, int64_t nsamp, double rn, int64_t* S_dev);
// the same for up to EMIT_MAX_DRAWS draws from ONE pass over the shard (the running sums do not depend on the draw):
// draw d writes S_dev + d * stride
constexpr int EMIT_MAX_DRAWS = 4;
void shard_scan_emit_multi(ShardScan* s, int64_t nsamp, const double* rn, int n_draws, int64_t* S_dev, int64_t stride);
int shard_scan_errors(ShardScan* s);  // ERRF_* raised so far (synchronises)

}  // namespace fricc
