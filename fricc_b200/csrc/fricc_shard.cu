// fricc_shard.cu -- FRI compression of a tensor that is SHARDED by contiguous flat-index ranges over several GPUs.
//
// Rank g holds x[j0_g, j0_g + n_g) (rank order = index order).  The reference compresses one vector in one process
// (/root/reference/src/fri_utils.cpp:439-522); this file provides the per-shard stages between which the host
// (fricc_b200/dist.py: compress_ranges) exchanges a few bytes, such that the result is bit-identical to the
// single-process compression:
//
//   norm       one_norm (fri_utils.cpp:164-171) is a left-to-right sum: every shard builds its chunk maps in parallel
//              (binades predicted from all-gathered approximate shard sums), then the exact sum is carried from rank to
//              rank through ONE small kernel per rank (shard_scan_resolve) -- device scalars only, stream-ordered.
//   top-|D|    argsort (fri_utils.cpp:11-36): key histograms of every shard are all-reduced, the host walks the digits
//              to a key threshold that at least `cand` elements reach, every shard compacts its elements at or above
//              it (shard_candidates_kernel), the lists are all-gathered and ordered identically on every rank
//              (sort_candidates), and get_d_largest's greedy test (fri_utils.cpp:400-437) runs on the ordered list.
//   sampling   sample_systematic (fri_utils.cpp:129-159): the same staged scan over p = |x|/r*n' (never materialised);
//              every shard emits the samples its own elements own into a zero-initialised list that is summed over ranks
//              (unmatched samples stay 0 exactly as in the reference).
//   values     fri_utils.cpp:498-504: the owner of a flat index computes x or x/p, everyone else contributes 0.0.
//
// Compiled with -fmad=false (PSrc).
#include "fricc_fri.cuh"
#include "fricc_psrc.cuh"
#include "fricc_sort.cuh"

#include <algorithm>

namespace fricc {

constexpr int SC_THREADS = 512;
constexpr int SC_ITEMS = 8;

// Appends (key, GLOBAL index) of every shard element with key >= key_min.  Same tile scheme as compact_kernel
// (fricc_select.cu): classify in registers, CTA-wide scan, one atomic per tile.  Order of the appended candidates is
// arbitrary; sort_candidates makes the result deterministic.
__global__ void __launch_bounds__(SC_THREADS)
shard_candidates_kernel(const double* __restrict__ x, int64_t n, int64_t j0, uint64_t key_min,
                        uint64_t* __restrict__ out_k, uint64_t* __restrict__ out_i, unsigned long long cap,
                        unsigned long long* __restrict__ count) {
  __shared__ unsigned s_w[SC_THREADS / 32];
  __shared__ unsigned long long s_base;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile_elems = (int64_t)SC_THREADS * SC_ITEMS;
  const int64_t ntiles = (n + tile_elems - 1) / tile_elems;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * tile_elems + threadIdx.x;
    uint64_t key[SC_ITEMS];
    unsigned bits = 0, c = 0;
#pragma unroll
    for (int q = 0; q < SC_ITEMS; q++) {
      const int64_t i = base + (int64_t)q * SC_THREADS;
      key[q] = 0;
      if (i < n) {
        key[q] = abs_key(x[i]);
        if (key[q] >= key_min) {
          bits |= 1u << q;
          c++;
        }
      }
    }
    unsigned inc = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    unsigned off = 0, tot = 0;
#pragma unroll
    for (unsigned w = 0; w < SC_THREADS / 32; w++) {
      const unsigned v = s_w[w];
      if (w < warp) off += v;
      tot += v;
    }
    if (threadIdx.x == 0) s_base = tot ? atomicAdd(count, (unsigned long long)tot) : 0ULL;
    __syncthreads();
    unsigned long long p = s_base + off + (inc - c);
#pragma unroll
    for (int q = 0; q < SC_ITEMS; q++) {
      if (bits & (1u << q)) {
        if (p < cap) {
          out_k[p] = key[q];
          out_i[p] = (uint64_t)(j0 + base + (int64_t)q * SC_THREADS);
        }
        p++;
      }
    }
    __syncthreads();
  }
}

void shard_candidates(const double* x_dev, int64_t n, int64_t j0, uint64_t key_min, uint64_t* keys_out_dev,
                      uint64_t* idx_out_dev, int64_t cap, unsigned long long* count_dev) {
  FRICC_CUDA(cudaMemsetAsync(count_dev, 0, sizeof(unsigned long long), stream()));
  if (n <= 0) return;
  const int grid = (int)std::max<int64_t>(
      1, std::min<int64_t>(ceil_div(n, (int64_t)SC_THREADS * SC_ITEMS), (int64_t)num_sms() * 8));
  FRICC_LAUNCH(shard_candidates_kernel, grid, SC_THREADS, 0, x_dev, n, j0, key_min, keys_out_dev, idx_out_dev,
               (unsigned long long)cap, count_dev);
}

// order (key, index) pairs by key descending, index ascending -- the canonical order of select_top_m -- in place
void sort_candidates(uint64_t* keys_dev, uint64_t* idx_dev, int64_t count, int64_t n_total) {
  if (count <= 1) return;
  DevBuf<uint64_t> k2((size_t)count), i2((size_t)count);
  uint64_t* kb[2] = {keys_dev, k2.p};
  uint64_t* ib[2] = {idx_dev, i2.p};
  const int r = radix_sort_pairs(kb, ib, count, candidate_digits(n_total));
  if (r == 1) {
    FRICC_CUDA(cudaMemcpyAsync(keys_dev, k2.p, (size_t)count * 8, cudaMemcpyDeviceToDevice, stream()));
    FRICC_CUDA(cudaMemcpyAsync(idx_dev, i2.p, (size_t)count * 8, cudaMemcpyDeviceToDevice, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));  // k2 / i2 are released on return
  }
}

// values of the entries this shard owns (fri_utils.cpp:498-504): entry e < nD is exactly preserved (x), every other
// entry is a sample (x / p); entries owned by another shard get 0.0 so that a sum over ranks assembles the list.
__global__ void shard_values_kernel(PSrc ps, int64_t n, const int64_t* __restrict__ flat, int64_t m, int64_t nD,
                                    double* __restrict__ vals) {
  const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  const int64_t j = flat[e] - ps.j0;
  double v = 0.0;
  if (j >= 0 && j < n) v = e < nD ? ps.x[j] : __ddiv_rn(ps.x[j], ps(j));
  vals[e] = v;
}

void shard_values(const PSrc& ps, int64_t n, const int64_t* flat_dev, int64_t m, int64_t nD, double* vals_dev) {
  if (m <= 0) return;
  FRICC_LAUNCH(shard_values_kernel, (unsigned)ceil_div(m, 256), 256, 0, ps, n, flat_dev, m, nD, vals_dev);
}

}  // namespace fricc
