// fricc_fri.cuh -- FRI compression stages (reference: /root/reference/src/fri_utils.cpp) -- see fricc_fri.cu
#pragma once
#include "fricc_common.cuh"
#include "fricc_select.cuh"

namespace fricc {

enum class Sampling { Pivotal = 0, Systematic = 1, None = 2 };

// where the uniform draws come from: a caller buffer (host) or the library's glibc-compatible stream
struct DrawSource {
  const int32_t* host = nullptr;  // draws[t] = value of the t-th rand() call
  int64_t n = 0;
  bool from_stream = true;
};

struct Compressed {
  int64_t m = 0;
  int64_t nD = 0;            // |D|
  double remaining_norm = 0; // r after the greedy subtraction
  int64_t consumed = 0;      // draws the reference would have consumed
  DevBuf<int64_t> flat;      // [m]  [D..., S...]
  DevBuf<double> vals;       // [m]
};

// select the implementation of the exact sums: 1 = parallel binade-segmented scan (default), 0 = one-thread chain
void set_parallel_scan(int on);

// exact sequential-order sums (bit-identical to a left-to-right fp64 accumulation)
double one_norm_sequential(const double* x_dev, int64_t n);

// get_d_largest on a sorted candidate list
// (top holds the top.m largest magnitudes; n_sample is the reference's n = m_keep, which may exceed top.m)
void d_greedy(const TopM& top, int64_t n_sample, double norm, int64_t* nD, double* remaining_norm);
// the plain one-warp walk from candidate `lo` with remaining norm r_start, and the parallel form (fricc_escan.cu)
void d_greedy_serial(const uint64_t* keys_dev, int64_t lo, int64_t cand, int64_t n_sample, double r_start, int64_t* nD,
                     double* remaining_norm);
void d_greedy_parallel(const uint64_t* keys_dev, int64_t cand, int64_t n_sample, double norm, int64_t* nD,
                       double* remaining_norm);
// first size of the candidate list compress_fri orders when it looks for D (grows x16 until the greedy test stops)
void set_first_candidates(int64_t c);
void set_d_estimate(int on);
void set_d_parallel(int mode);

// make_probability_vector (materialised; stage-wise API only -- the compression itself never stores p)
void make_probability_vector(const double* x_dev, int64_t n, int64_t n_sample, const int64_t* D_dev, int64_t nD,
                             double remaining_norm, double* p_dev);

// samplers on an explicit probability vector (stage-wise API); S_dev has n_sample entries
void sample_systematic_probs(int64_t n_sample, const double* p_dev, int64_t n, const DrawSource& draws,
                             int64_t* consumed, int64_t* S_dev);
void sample_pivotal_probs(int64_t n_sample, const double* p_dev, int64_t n, const DrawSource& draws, int64_t* consumed,
                          int64_t* S_dev);

// parallel_sample (fri_utils.cpp:176-362) with n_chunks chunks, draws consumed in chunk order   (fricc_psample.cu)
void parallel_sample(int64_t n_sample, const double* p_dev, int64_t n, int n_chunks, const DrawSource& draws,
                     int64_t* consumed, int64_t* S_dev);

// the per-chunk part of parallel_sample for ONE chunk held by this rank (fri_utils.cpp:296-356); budgets by the caller
struct PSrc;
void sample_chunk(const PSrc& ps, int64_t n, int64_t j0, int64_t g, double t, double a, double norm, const DrawSource& draws,
                  int64_t* consumed, int64_t* S_dev);

// final state of a pivotal walk (device struct shared by the chain kernel and the parallel sampler)
struct PivotalState {
  double a;           // running sum after the last visited element
  long long isel;     // current candidate i
  long long count;    // samples written
  long long visited;  // number of loop iterations the reference would have executed (= draws consumed)
};
struct PSrc;
// parallel exact pivotal sampling (fricc_pivotal.cu); false = could not finish (caller falls back to the chain kernel)
bool pivotal_parallel(const PSrc& ps, int64_t n, int64_t nsamp, const int32_t* draws_dev, int64_t ndraws,
                      int64_t* S_dev, PivotalState* st_dev, int* err_dev, int* restarts_out);
// "pivotal_parallel" option: -1 auto (parallel for n >= 65536), 0 one-thread chain, 1 always parallel
void set_pivotal_parallel(int mode);
int pivotal_restarts_last();

// the two compressions behind SparseTensor4d's constructor
void compress_largest(const double* x_dev, int64_t n, int64_t m, Compressed& out);
void compress_fri(const double* x_dev, int64_t n, int64_t m, Sampling method, const DrawSource& draws, bool verbose,
                  Compressed& out);

// per-shard stages of the range-sharded compression (fricc_shard.cu)
void shard_candidates(const double* x_dev, int64_t n, int64_t j0, uint64_t key_min, uint64_t* keys_out_dev,
                      uint64_t* idx_out_dev, int64_t cap, unsigned long long* count_dev);
void sort_candidates(uint64_t* keys_dev, uint64_t* idx_dev, int64_t count, int64_t n_total);
void shard_values(const PSrc& ps, int64_t n, const int64_t* flat_dev, int64_t m, int64_t nD, double* vals_dev);
void raise_error_flags(int flags);  // ERRF_* -> the reference's error (fail)

// glibc-compatible rand() stream (host side)
void rand_seed(uint32_t seed);
void rand_fill(int32_t* out, int64_t n);  // advances
void rand_fill_device(int32_t* out_dev, int64_t n);  // next n draws into device memory; does NOT advance the stream
void rand_skip(int64_t count);                        // advance the stream by count draws (jump-ahead)
int64_t rand_consumed();
struct RandSnapshot {
  uint32_t ring[31];
  int64_t i;
  int64_t consumed;
};
RandSnapshot rand_snapshot();
void rand_restore(const RandSnapshot& s);

}  // namespace fricc
