// fricc_mscan.cuh -- ordered device-wide EXCLUSIVE scan over an arbitrary monoid (non-commutative allowed).
//
// A monoid is a struct M with   typedef ... T;   static T identity();   static T op(const T& first, const T& second);
// static T shfl_up(const T&, int)   (all __device__).  Three kernels: per-tile totals, one-CTA scan of the totals,
// per-tile apply.  Used for the shift maps and the prefix maximum of the parallel pivotal sampler (fricc_pivotal.cu)
// and for the rounding maps of the parallel D-selection (fricc_escan.cu).
#pragma once
#include "fricc_common.cuh"

namespace fricc {

constexpr int MS_THREADS = 256;
constexpr int MS_IPT = 4;
constexpr int MS_TILE = MS_THREADS * MS_IPT;

// ordered inclusive scan across the block of one item per thread; *excl = ordered product of all earlier threads' items
template <class M>
__device__ __forceinline__ typename M::T ms_block_scan(typename M::T v, typename M::T* smem /*[32]*/,
                                                       typename M::T* excl, typename M::T* total) {
  typedef typename M::T T;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned nwarps = (blockDim.x + 31) >> 5;
  T inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T t = M::shfl_up(inc, o);
    if (lane >= (unsigned)o) inc = M::op(t, inc);
  }
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    T s = (lane < nwarps) ? smem[lane] : M::identity();
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      T t = M::shfl_up(s, o);
      if (lane >= (unsigned)o) s = M::op(t, s);
    }
    smem[lane] = s;
  }
  __syncthreads();
  T woff = warp ? smem[warp - 1] : M::identity();
  *total = smem[nwarps - 1];
  T prev = M::shfl_up(inc, 1);
  if (lane == 0) prev = M::identity();
  *excl = M::op(woff, prev);
  __syncthreads();
  return M::op(woff, inc);
}

template <class M>
__global__ void __launch_bounds__(MS_THREADS)
ms_tile_totals(const typename M::T* __restrict__ in, int64_t n, typename M::T* __restrict__ totals) {
  typedef typename M::T T;
  __shared__ T sm[32];
  const int64_t base = (int64_t)blockIdx.x * MS_TILE + (int64_t)threadIdx.x * MS_IPT;
  T s = M::identity();
#pragma unroll
  for (int i = 0; i < MS_IPT; i++)
    if (base + i < n) s = M::op(s, in[base + i]);
  T ex, tot;
  ms_block_scan<M>(s, sm, &ex, &tot);
  if (threadIdx.x == 0) totals[blockIdx.x] = tot;
}

// in-place ordered EXCLUSIVE scan of a short array by one CTA
template <class M>
__global__ void __launch_bounds__(MS_THREADS)
ms_single_block(typename M::T* __restrict__ d, int64_t n) {
  typedef typename M::T T;
  __shared__ T sm[32];
  T carry = M::identity();
  for (int64_t base = 0; base < n; base += MS_THREADS) {
    const int64_t idx = base + threadIdx.x;
    T v = (idx < n) ? d[idx] : M::identity();
    T ex, tot;
    ms_block_scan<M>(v, sm, &ex, &tot);
    if (idx < n) d[idx] = M::op(carry, ex);
    carry = M::op(carry, tot);
  }
}

// out[i] = ordered product of in[0..i)   (exclusive), given the exclusive scan of the tile totals
template <class M>
__global__ void __launch_bounds__(MS_THREADS)
ms_apply(const typename M::T* __restrict__ in, int64_t n, const typename M::T* __restrict__ tile_excl,
         typename M::T* __restrict__ out) {
  typedef typename M::T T;
  __shared__ T sm[32];
  const int64_t base = (int64_t)blockIdx.x * MS_TILE + (int64_t)threadIdx.x * MS_IPT;
  T v[MS_IPT];
  T s = M::identity();
#pragma unroll
  for (int i = 0; i < MS_IPT; i++) {
    v[i] = (base + i < n) ? in[base + i] : M::identity();
    s = M::op(s, v[i]);
  }
  T ex, tot;
  ms_block_scan<M>(s, sm, &ex, &tot);
  T run = M::op(tile_excl[blockIdx.x], ex);
#pragma unroll
  for (int i = 0; i < MS_IPT; i++) {
    if (base + i < n) out[base + i] = run;
    run = M::op(run, v[i]);
  }
}

template <class M>
static inline void ms_exclusive_scan(const typename M::T* in, int64_t n, typename M::T* out) {
  typedef typename M::T T;
  if (n <= 0) return;
  const int64_t ntiles = ceil_div(n, MS_TILE);
  DevBuf<T> totals((size_t)ntiles);
  FRICC_LAUNCH(ms_tile_totals<M>, (unsigned)ntiles, MS_THREADS, 0, in, n, totals.p);
  FRICC_LAUNCH(ms_single_block<M>, 1, MS_THREADS, 0, totals.p, ntiles);
  FRICC_LAUNCH(ms_apply<M>, (unsigned)ntiles, MS_THREADS, 0, in, n, totals.p, out);
}

}  // namespace fricc
