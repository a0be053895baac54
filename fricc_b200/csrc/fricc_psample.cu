// fricc_psample.cu -- parallel_sample ("Alg. A.2", /root/reference/src/fri_utils.cpp:176-362): chunked pivotal sampling.
//
// The reference cuts q into omp_get_max_threads() contiguous chunks (the last one takes the remainder), gives chunk j
// the budget a_j = n_sample * |q_j|_1 / |q|_1, rounds the budgets to integers g_j with a pivotal draw over the fractional
// parts, raises / lowers the LEADING probabilities of a chunk until its mass is g_j, and samples every chunk
// independently with sample_pivotal; indices are offset by the chunk start and concatenated in chunk order.
// Its threads call rand() concurrently, so for more than one chunk the order of its draws is a race.  Here the draws
// are consumed in chunk order (chunk j after chunk j-1): the reference's behaviour whenever its threads run one after
// another, and exactly its behaviour with one thread.  The CPU restatement with the same convention
// (oracle/fricc_oracle.c: orc_parallel_sample) is pinned against the compiled reference (tests/test_oracle.py).
//
// Everything that touches n elements runs on the device: the norms are the exact left-to-right sums of fricc_escan.cu,
// the chunks are sampled by the parallel pivotal sampler (fricc_pivotal.cu), the rescaling walk below is one thread
// per chunk (it stops after the first few elements: the mass to move is < 1 and the leading probabilities add up to
// that within ~N / n_sample elements).  The O(n_chunks) budget arithmetic is host code, as in the reference.
//
// Compiled with -fmad=false (same IEEE operations, in the same order, as the reference's x86-64 build).
#include "fricc_fri.cuh"
#include "fricc_psrc.cuh"

#include <cmath>
#include <vector>

namespace fricc {

// fri_utils.cpp:303-322.  q: this call's private copy of the probabilities (the reference copies every chunk into
// qj[j], :196-198,219), modified in place.
__device__ void rescale_chunk(double* __restrict__ qj, int64_t len, double tj, double aj, double gj, double sj) {
  if (gj > aj) {
    for (int64_t i = 0; i < len; i++) {
      const double qi = qj[i];
      double y = __ddiv_rn(qi, tj);
      if (!(y < 1.)) y = 1.;  // std::min(1., y)
      sj = __dadd_rn(sj, __dsub_rn(y, qi));
      if (sj >= gj) {
        qj[i] = __dsub_rn(__dadd_rn(y, gj), sj);
        break;
      }
      qj[i] = y;
    }
  } else if (gj < aj) {
    const double omt = __dsub_rn(1., tj);
    for (int64_t i = 0; i < len; i++) {
      const double qi = qj[i];
      double y = __ddiv_rn(__dsub_rn(qi, tj), omt);
      if (!(0. < y)) y = 0.;  // std::max(0., y)
      sj = __dadd_rn(sj, __dsub_rn(y, qi));
      if (sj <= gj) {
        qj[i] = __dsub_rn(__dadd_rn(y, gj), sj);
        break;
      }
      qj[i] = y;
    }
  }
}

__global__ void rescale_chunks_kernel(double* __restrict__ q, int64_t step, int64_t rem, int n_chunks,
                                      const double* __restrict__ t, const double* __restrict__ a,
                                      const int64_t* __restrict__ g, const double* __restrict__ norms) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n_chunks) return;
  rescale_chunk(q + (int64_t)j * step, step + (j + 1 == n_chunks ? rem : 0), t[j], a[j], (double)g[j], norms[j]);
}

// one chunk, its budget by value (the rank-distributed form below)
__global__ void rescale_one_chunk_kernel(double* __restrict__ q, int64_t len, double t, double a, double g, double norm) {
  if (blockIdx.x == 0 && threadIdx.x == 0) rescale_chunk(q, len, t, a, g, norm);
}

__global__ void materialize_kernel(PSrc ps, int64_t n, double* __restrict__ q) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) q[i] = ps(i);
}

__global__ void add_offset_kernel(int64_t* __restrict__ S, int64_t count, int64_t offset) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) S[i] += offset;
}

// the next `count` draws of `src` starting `skip` draws in (explicit buffers only)
static DrawSource advance(const DrawSource& src, int64_t skip) {
  DrawSource d = src;
  if (!src.from_stream) {
    const int64_t s = std::min<int64_t>(skip, src.n);
    d.host = src.host + s;
    d.n = src.n - s;
  }
  return d;
}

// sample_pivotal(n_sample, probs[0..n)) for the cases the reference's loop leaves after at most one iteration, decided
// on the host from the first two probabilities (fri_utils.cpp:70-105): n == 1 (the loop does not run) and n_sample == 0
// (one draw is taken, then `S.size() == n_sample` ends the loop).  Returns the draws consumed.
static int64_t trivial_pivotal(int64_t n_sample, const double* p_dev, int64_t n) {
  if (n == 1) {
    if (n_sample != 0) fail(FRICC_ECOUNT, "ERROR IN PIVOTAL SAMPLING: wrong number of samples (reference: fri_utils.cpp:121-124)");
    return 0;
  }
  double p01[2];
  FRICC_CUDA(cudaMemcpyAsync(p01, p_dev, sizeof(p01), cudaMemcpyDeviceToHost, stream()));
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  double a = p01[0] + p01[1];
  if (!(a < 1.)) {
    a -= 1.;
    const double pa = (1. - p01[1]) / (1. - a);
    if (pa > 1 || pa < 0)
      fail(FRICC_EPIVOTAL, "ERROR IN PIVOTAL SAMPLING: acceptance probability outside [0,1] (reference: fri_utils.cpp:84-94)");
    // a crossing pushes one sample: S.size() == 1 != n_sample == 0, the loop runs on to the end and the final count
    // check fails
    fail(FRICC_ECOUNT, "ERROR IN PIVOTAL SAMPLING: wrong number of samples (reference: fri_utils.cpp:121-124)");
  }
  return 1;
}

void parallel_sample(int64_t n_sample, const double* p_dev, int64_t n, int n_chunks, const DrawSource& draws,
                     int64_t* consumed, int64_t* S_dev) {
  if (n_chunks < 1) fail(FRICC_EINVAL, "parallel_sample: need at least one chunk");
  if (n_sample < 0 || n < n_chunks)
    fail(FRICC_EINVAL, "parallel_sample: need n_sample >= 0 and at least one element per chunk (n=%lld, chunks=%d)",
         (long long)n, n_chunks);
  const int64_t step = n / n_chunks, rem = n % n_chunks;
  const int T = n_chunks;
  auto len_of = [&](int j) { return step + (j + 1 == T ? rem : 0); };
  // ---- budgets (fri_utils.cpp:191, 214-237): exact left-to-right norms on the device, O(T) arithmetic on the host
  const double q_norm = one_norm_sequential(p_dev, n);
  std::vector<double> t(T), a(T), norms(T);
  std::vector<int64_t> g(T);
  int64_t c = n_sample;
  for (int j = 0; j < T; j++) {
    norms[j] = one_norm_sequential(p_dev + (int64_t)j * step, len_of(j));
    a[j] = (double)n_sample * norms[j] / q_norm;
    t[j] = a[j] - std::floor(a[j]);
    g[j] = (int64_t)(int)std::floor(a[j]);
    c -= g[j];
  }
  if (c < 0) fail(FRICC_ECOUNT, "parallel_sample: the integer budgets exceed n_sample");
  // ---- which chunks take one more sample: s_prime = sample_pivotal(c, t)   (:268-275)
  int64_t used = 0;
  {
    DevBuf<double> t_dev((size_t)T);
    t_dev.from_host(t.data(), (size_t)T);
    if (T == 1 || c == 0) {
      const int64_t u = trivial_pivotal(c, t_dev.p, T);
      if (draws.from_stream && u > 0) rand_skip(u);  // leave the stream where the reference's rand() would be
      used += u;
    } else {
      DevBuf<int64_t> sp_dev((size_t)c);
      int64_t u = 0;
      sample_pivotal_probs(c, t_dev.p, T, advance(draws, used), &u, sp_dev.p);
      used += u;
      std::vector<int64_t> sp((size_t)c);
      sp_dev.to_host(sp.data(), (size_t)c);
      for (int64_t k = 0; k < c; k++) g[(size_t)sp[(size_t)k]] += 1;
    }
  }
  // ---- rescale the leading probabilities of every chunk to the integer mass g_j   (:303-322)
  DevBuf<double> q((size_t)n);
  FRICC_CUDA(cudaMemcpyAsync(q.p, p_dev, (size_t)n * sizeof(double), cudaMemcpyDeviceToDevice, stream()));
  {
    DevBuf<double> tad((size_t)T * 3);
    DevBuf<int64_t> gd((size_t)T);
    std::vector<double> pack((size_t)T * 3);
    for (int j = 0; j < T; j++) {
      pack[(size_t)j] = t[j];
      pack[(size_t)(T + j)] = a[j];
      pack[(size_t)(2 * T + j)] = norms[j];
    }
    tad.from_host(pack.data(), (size_t)T * 3);
    gd.from_host(g.data(), (size_t)T);
    FRICC_LAUNCH(rescale_chunks_kernel, (unsigned)ceil_div(T, 64), 64, 0, q.p, step, rem, T, tad.p, tad.p + T, gd.p,
                 tad.p + 2 * T);
    FRICC_CUDA(cudaStreamSynchronize(stream()));  // pack / g are host temporaries of this scope
  }
  // ---- sample every chunk, in chunk order (draws consumed in that order), offset, concatenate   (:345-356)
  int64_t lb = 0;
  for (int j = 0; j < T; j++) {
    const int64_t lo = (int64_t)j * step, len = len_of(j);
    if (len == 1 || g[j] == 0) {
      const int64_t u = trivial_pivotal(g[j], q.p + lo, len);
      if (draws.from_stream && u > 0) rand_skip(u);
      used += u;
    } else {
      int64_t u = 0;
      sample_pivotal_probs(g[j], q.p + lo, len, advance(draws, used), &u, S_dev + lb);
      used += u;
      if (lo > 0) FRICC_LAUNCH(add_offset_kernel, (unsigned)ceil_div(g[j], 256), 256, 0, S_dev + lb, g[j], lo);
    }
    lb += g[j];
  }
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  *consumed = used;
}

// ---------------------------------------------------------------------------------------------------------------------
// Alg. A.2 with the chunks on DIFFERENT RANKS (parallel_sample as the distributed sampler).  A rank
// holds one chunk: the probabilities `ps` of its n elements (implicit: |x| / r * n' with D zeroed, or explicit).  The
// budget arithmetic -- |q|_1, |q_j|_1, a_j, t_j, g_j, the pivotal draw over the fractional parts (fri_utils.cpp:191-275)
// -- needs one double per rank and is done by the caller (fricc_b200/ranges.py) with the exact sums of fricc_escan.cu;
// this is the per-chunk part (:296-356): copy, rescale the leading probabilities to the integer mass g, sample_pivotal,
// offset by the chunk's first flat index.  No state crosses the chunk boundary, which is what lets the chunks run on
// different GPUs at the same time; each takes its own draws (the reference's threads share rand() in a race).
// ---------------------------------------------------------------------------------------------------------------------
void sample_chunk(const PSrc& ps, int64_t n, int64_t j0, int64_t g, double t, double a, double norm, const DrawSource& draws,
                  int64_t* consumed, int64_t* S_dev) {
  *consumed = 0;
  if (n < 1) {
    if (g != 0) fail(FRICC_ECOUNT, "sample_chunk: an empty chunk with a non-zero budget");
    return;
  }
  if (g < 0 || g > n) fail(FRICC_ECOUNT, "sample_chunk: budget %lld outside [0, %lld]", (long long)g, (long long)n);
  DevBuf<double> q((size_t)n);
  FRICC_LAUNCH(materialize_kernel, (unsigned)std::min<int64_t>(ceil_div(n, 256), (int64_t)num_sms() * 8), 256, 0, ps, n, q.p);
  FRICC_LAUNCH(rescale_one_chunk_kernel, 1, 32, 0, q.p, n, t, a, (double)g, norm);
  int64_t u = 0;
  if (n == 1 || g == 0) {
    u = trivial_pivotal(g, q.p, n);
    if (draws.from_stream && u > 0) rand_skip(u);
  } else {
    sample_pivotal_probs(g, q.p, n, draws, &u, S_dev);
    if (j0 > 0) FRICC_LAUNCH(add_offset_kernel, (unsigned)ceil_div(g, 256), 256, 0, S_dev, g, j0);
  }
  FRICC_CUDA(cudaStreamSynchronize(stream()));  // q is a temporary of this scope
  *consumed = u;
}

}  // namespace fricc
