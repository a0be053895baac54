// fricc_escan.cu -- exact LEFT-TO-RIGHT fp64 running sums computed in parallel (binade-segmented integer scan).
//
// The reference accumulates sequentially:  one_norm (fri_utils.cpp:164-171, 1 thread), std::partial_sum in
// sample_systematic (fri_utils.cpp:142-143).  These kernels reproduce those sums bit for bit while reading x at HBM
// speed; the algebra (maps plain(A) / even(A,B), crossings) is in fricc_exact.cuh and is unit-tested on the CPU by
// tests/test_exact_scan_host.py.
//
//   E1  escan_approx_kernel   per-chunk tree sum (approximate)            -> approximate running sum at chunk starts
//   E2  escan_map_kernel      per-chunk composed integer map, in the binade the approximate sum predicts
//   E3  escan_resolve_kernel  one CTA walks the chunk maps 32 at a time (warp scan of maps), carrying the EXACT sum;
//                             a chunk whose map is not applicable (binade crossing inside it, or the prediction was
//                             off) is redone cooperatively with real additions at the crossings
//   E4  escan_emit_kernel     per chunk, from its exact start: exact running sums of all elements -> systematic samples
//
// Compiled with -fmad=false (the probabilities are computed exactly as the reference does).
#include "fricc_exact.cuh"
#include "fricc_psrc.cuh"
#include "fricc_sort.cuh"

namespace fricc {

constexpr int ES_THREADS = 256;
constexpr int ES_E = 8;
constexpr int ES_CHUNK = ES_THREADS * ES_E;  // 2048 elements
constexpr int ES_WARPS = ES_THREADS / 32;

struct ChunkMap {
  int64_t A, B;
  int32_t even;
  int32_t k;  // assumed binade; ES_K_NONE = no usable prediction (zero / subnormal / non-finite running sum)
};
constexpr int ES_K_NONE = -100000;

__device__ __forceinline__ EMap shfl_up_emap(const EMap& m, int o) {
  EMap r;
  r.A = __shfl_up_sync(0xffffffffu, m.A, o);
  r.B = __shfl_up_sync(0xffffffffu, m.B, o);
  r.even = __shfl_up_sync(0xffffffffu, m.even, o);
  return r;
}

// ordered inclusive scan inside a warp: lane i gets map_0 then map_1 ... then map_i
__device__ __forceinline__ EMap warp_incl_scan_emap(EMap m) {
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    EMap other = shfl_up_emap(m, o);
    if (lane >= (unsigned)o) m = emap_compose(other, m);
  }
  return m;
}

// ordered exclusive scan over the CTA (one map per thread); smem: one entry per warp; all threads must call
__device__ __forceinline__ EMap block_excl_scan_emap(EMap mine, EMap* smem) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  EMap inc = warp_incl_scan_emap(mine);
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  EMap pre = emap_identity();
  for (unsigned w = 0; w < warp; w++) pre = emap_compose(pre, smem[w]);
  EMap ex = shfl_up_emap(inc, 1);
  if (lane == 0) ex = emap_identity();
  __syncthreads();
  return emap_compose(pre, ex);
}

// ---- E1 ----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ES_THREADS)
escan_approx_kernel(PSrc ps, int64_t n, double* __restrict__ approx, int* __restrict__ err) {
  __shared__ double sm[ES_WARPS];
  const int64_t base = (int64_t)blockIdx.x * ES_CHUNK;
  double s = 0.0;
  int bad = 0;
#pragma unroll
  for (int q = 0; q < ES_E; q++) {
    int64_t j = base + (int64_t)q * ES_THREADS + threadIdx.x;  // coalesced; the order of this sum is irrelevant
    if (j < n) {
      double p = ps(j);
      if (p > 1.0) bad = 1;
      s += p;
    }
  }
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int w = 0; w < ES_WARPS; w++) t += sm[w];
    approx[blockIdx.x] = t;
  }
  if (bad && !ps.absval) atomicOr(err, ERRF_PROB);
}

// ---- E2 ----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ES_THREADS)
escan_map_kernel(PSrc ps, int64_t n, const double* __restrict__ approx_excl, ChunkMap* __restrict__ maps) {
  __shared__ EMap sm[ES_WARPS];
  const int64_t base = (int64_t)blockIdx.x * ES_CHUNK + (int64_t)threadIdx.x * ES_E;
  const double c0 = approx_excl[blockIdx.x];
  const bool usable = fx_is_normal_positive(c0);
  const int k = usable ? fx_binade(c0) : 0;
  EMap m = emap_identity();
  if (usable) {
#pragma unroll
    for (int q = 0; q < ES_E; q++) {
      int64_t j = base + q;
      if (j < n) m = emap_compose(m, emap_of(ps(j), k));
    }
  }
  EMap inc = warp_incl_scan_emap(m);
  if ((threadIdx.x & 31) == 31) sm[threadIdx.x >> 5] = inc;
  __syncthreads();
  if (threadIdx.x == 0) {
    EMap t = emap_identity();
    for (int w = 0; w < ES_WARPS; w++) t = emap_compose(t, sm[w]);
    ChunkMap cm;
    cm.A = t.A; cm.B = t.B; cm.even = t.even;
    cm.k = usable ? k : ES_K_NONE;
    maps[blockIdx.x] = cm;
  }
}

// ---- exact scan of one chunk by the whole CTA ----------------------------------------------------------------------
// p[0..cnt) are the chunk's addends (shared memory), c_in the exact running sum before the chunk.  On return
// cs[j] = exact running sum after element j, and the function returns the exact running sum after the chunk.
constexpr int ES_MAX_WARPS = 32;  // the resolve kernel runs the same chunk scan with a 1024-thread CTA
struct ChunkScratch {
  EMap warp_maps[ES_MAX_WARPS];
  int first_cross[ES_MAX_WARPS];
  double c;
  int start;
};

__device__ double chunk_exact_scan(const double* __restrict__ p, int cnt, double c_in, double* __restrict__ cs,
                                   ChunkScratch* sc) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int j0 = threadIdx.x * ES_E;
  if (threadIdx.x == 0) {
    sc->c = c_in;
    sc->start = 0;
  }
  __syncthreads();
  while (true) {
    const double c = sc->c;
    const int start = sc->start;
    if (start >= cnt) break;
    const bool zero = (c == 0.0);
    if (!zero && !fx_is_normal_positive(c)) {
      // subnormal (or non-finite) running sum: one real addition, then look again
      __syncthreads();
      if (threadIdx.x == 0) {
        double cn = __dadd_rn(c, p[start]);
        cs[start] = cn;
        sc->c = cn;
        sc->start = start + 1;
      }
      __syncthreads();
      continue;
    }
    const int k = zero ? 0 : fx_binade(c);
    const int64_t C0 = zero ? 0 : fx_to_state(c);
    EMap mine = emap_identity();
#pragma unroll
    for (int q = 0; q < ES_E; q++) {
      int j = j0 + q;
      if (j >= start && j < cnt) mine = emap_compose(mine, zero ? emap_of_zero_state(p[j]) : emap_of(p[j], k));
    }
    EMap ex = block_excl_scan_emap(mine, sc->warp_maps);
    int64_t S = emap_apply(ex, C0);
    int64_t st[ES_E];
    int myfirst = ES_CHUNK;
#pragma unroll
    for (int q = 0; q < ES_E; q++) {
      int j = j0 + q;
      if (j >= start && j < cnt) {
        S = emap_apply(zero ? emap_of_zero_state(p[j]) : emap_of(p[j], k), S);
        if (S >= FX_TWO53 && myfirst == ES_CHUNK) myfirst = j;
      }
      st[q] = S;
    }
    int wf = myfirst;
    for (int o = 16; o; o >>= 1) wf = min(wf, __shfl_xor_sync(0xffffffffu, wf, o));
    if (lane == 0) sc->first_cross[warp] = wf;
    __syncthreads();
    int jx = ES_CHUNK;
    for (int w = 0; w < (int)(blockDim.x >> 5); w++) jx = min(jx, sc->first_cross[w]);
    const int valid_end = min(jx, cnt);
#pragma unroll
    for (int q = 0; q < ES_E; q++) {
      int j = j0 + q;
      if (j >= start && j < valid_end) cs[j] = zero ? 0.0 : fx_from_state(st[q], k);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      if (jx < cnt) {
        // binade crossing at element jx: one real addition, restart behind it
        double prev = (jx == start) ? c : cs[jx - 1];
        double cn = __dadd_rn(prev, p[jx]);
        cs[jx] = cn;
        sc->c = cn;
        sc->start = jx + 1;
      } else {
        sc->c = cs[cnt - 1];
        sc->start = cnt;
      }
    }
    __syncthreads();
  }
  double out = sc->c;
  __syncthreads();
  return out;
}

// ---- E3 ----------------------------------------------------------------------------------------------------------
// One CTA of RS_THREADS threads walks the chunk maps RS_THREADS at a time (CTA-wide ordered scan of maps), carrying the
// EXACT sum; the first chunk of a batch whose map is not applicable (binade crossing inside it, or the prediction was
// off) is redone cooperatively with real additions at the crossings, and the walk resumes behind it.
// exact_in[i] = exact running sum before chunk i; *total = exact sum of everything.
constexpr int RS_THREADS = 1024;
constexpr int RS_WARPS = RS_THREADS / 32;

__global__ void __launch_bounds__(RS_THREADS)
escan_resolve_kernel(PSrc ps, int64_t n, const ChunkMap* __restrict__ maps, int64_t nchunks,
                     double* __restrict__ exact_in, double* __restrict__ total, int* __restrict__ stats) {
  __shared__ double pbuf[ES_CHUNK];
  __shared__ double cs[ES_CHUNK];
  __shared__ ChunkScratch sc;
  __shared__ EMap s_wmap[RS_WARPS];
  __shared__ int s_wgood[RS_WARPS];
  __shared__ double s_c;
  __shared__ long long s_next;
  __shared__ int s_dirty;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    s_c = 0.0;
    s_next = 0;
  }
  __syncthreads();
  int ndirty = 0;
  while (true) {
    const int64_t i0 = s_next;
    if (i0 >= nchunks) break;
    __syncthreads();
    {
      const double c = s_c;
      const int64_t i = i0 + threadIdx.x;
      ChunkMap cm;
      cm.A = 0; cm.B = 0; cm.even = 0; cm.k = ES_K_NONE;
      if (i < nchunks) cm = maps[i];
      const bool cnormal = fx_is_normal_positive(c);
      const int kc = cnormal ? fx_binade(c) : ES_K_NONE;
      const int64_t C0 = cnormal ? fx_to_state(c) : 0;
      EMap m{cm.A, cm.B, cm.even};
      EMap inc = warp_incl_scan_emap(m);
      if (lane == 31) s_wmap[warp] = inc;
      __syncthreads();
      // composed map of all preceding warps: lanes load the warp totals below `warp`, ordered warp scan, take the last
      EMap wm = (lane < warp) ? s_wmap[lane] : emap_identity();
      wm = warp_incl_scan_emap(wm);
      EMap pre;
      pre.A = __shfl_sync(0xffffffffu, wm.A, 31);
      pre.B = __shfl_sync(0xffffffffu, wm.B, 31);
      pre.even = __shfl_sync(0xffffffffu, wm.even, 31);
      const int64_t S_after = emap_apply(emap_compose(pre, inc), C0);
      // a chunk is fine if its prediction matches the current binade and the sum stays inside the binade behind it
      const bool ok = (i < nchunks) && cnormal && cm.k == kc && S_after < FX_TWO53;
      const unsigned okmask = __ballot_sync(0xffffffffu, ok);
      if (lane == 0) s_wgood[warp] = (okmask == 0xffffffffu) ? 32 : (__ffs(~okmask) - 1);
      // exact state BEFORE chunk i = state after chunk i-1
      int64_t S_before = __shfl_up_sync(0xffffffffu, S_after, 1);
      if (lane == 0) S_before = emap_apply(pre, C0);
      __syncthreads();
      int ngood = 0;  // number of leading fine chunks of the batch (0..RS_THREADS)
      for (int w = 0; w < RS_WARPS; w++) {
        const int g = s_wgood[w];
        ngood += g;
        if (g < 32) break;
      }
      const int t = (int)threadIdx.x;
      if (i < nchunks) {
        if (t < ngood || (t == ngood && ngood > 0)) exact_in[i] = fx_from_state(S_before, kc);
        else if (t == ngood) exact_in[i] = c;  // ngood == 0: the batch's first chunk starts from the carried sum
      }
      if (ngood > 0 && t == ngood - 1) s_c = fx_from_state(S_after, kc);
      if (t == 0) {
        s_next = i0 + ngood;
        s_dirty = (ngood < RS_THREADS && i0 + ngood < nchunks) ? 1 : 0;  // chunk i0+ngood is redone cooperatively
      }
    }
    __syncthreads();
    if (s_dirty) {
      const int64_t ic = s_next;
      const int64_t base = ic * ES_CHUNK;
      const int cnt = (int)min((int64_t)ES_CHUNK, n - base);
      for (int k = threadIdx.x; k < cnt; k += RS_THREADS) pbuf[k] = ps(base + k);
      __syncthreads();
      double c_out = chunk_exact_scan(pbuf, cnt, s_c, cs, &sc);
      if (threadIdx.x == 0) {
        s_c = c_out;
        s_next = ic + 1;
      }
      ndirty++;
      __syncthreads();
    }
  }
  if (threadIdx.x == 0) {
    *total = s_c;
    if (stats) stats[0] = ndirty;
  }
}

// ---- E4 ----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ES_THREADS)
escan_emit_kernel(PSrc ps, int64_t n, const double* __restrict__ exact_in, int64_t nsamp, double rn,
                  int64_t* __restrict__ S) {
  __shared__ double pbuf[ES_CHUNK];
  __shared__ double cs[ES_CHUNK];
  __shared__ ChunkScratch sc;
  const int64_t base = (int64_t)blockIdx.x * ES_CHUNK;
  const int cnt = (int)min((int64_t)ES_CHUNK, n - base);
  for (int k = threadIdx.x; k < cnt; k += ES_THREADS) pbuf[k] = ps(base + k);
  __syncthreads();
  const double c_in = exact_in[blockIdx.x];
  chunk_exact_scan(pbuf, cnt, c_in, cs, &sc);
  // interval of element j is [excl, incl): samples i with excl <= uk_i < incl land on j   (fri_utils.cpp:147-157)
  for (int k = threadIdx.x; k < cnt; k += ES_THREADS) {
    double incl = cs[k];
    double excl = k ? cs[k - 1] : c_in;
    if (incl > excl) {
      int64_t lo = first_sample_at_or_above(excl, rn, nsamp);
      int64_t hi = first_sample_at_or_above(incl, rn, nsamp);
      for (int64_t i = lo; i < hi; i++) S[i] = base + k;
    }
  }
}

// ---- host ----------------------------------------------------------------------------------------------------------
struct EscanPlan {
  int64_t nchunks = 0;
  DevBuf<double> approx;     // approximate exclusive running sum at chunk starts
  DevBuf<ChunkMap> maps;
  DevBuf<double> exact_in;   // exact running sum at chunk starts
  DevBuf<double> total;
  DevBuf<int> err;
  DevBuf<int> stats;
};

static void escan_prepare(const PSrc& ps, int64_t n, EscanPlan& pl) {
  pl.nchunks = ceil_div(n, ES_CHUNK);
  pl.approx.alloc((size_t)pl.nchunks);
  pl.maps.alloc((size_t)pl.nchunks);
  pl.exact_in.alloc((size_t)pl.nchunks);
  pl.total.alloc(1);
  pl.err.alloc(1);
  pl.stats.alloc(1);
  pl.err.zero();
  FRICC_LAUNCH(escan_approx_kernel, (unsigned)pl.nchunks, ES_THREADS, 0, ps, n, pl.approx.p, pl.err.p);
  exclusive_scan_f64(pl.approx.p, pl.nchunks);
  FRICC_LAUNCH(escan_map_kernel, (unsigned)pl.nchunks, ES_THREADS, 0, ps, n, pl.approx.p, pl.maps.p);
  FRICC_LAUNCH(escan_resolve_kernel, 1, RS_THREADS, 0, ps, n, pl.maps.p, pl.nchunks, pl.exact_in.p, pl.total.p,
               pl.stats.p);
}

double exact_sum_parallel(const PSrc& ps, int64_t n, int* err_flags_out) {
  if (err_flags_out) *err_flags_out = 0;
  if (n <= 0) return 0.0;
  EscanPlan pl;
  escan_prepare(ps, n, pl);
  double t = 0.0;
  pl.total.to_host(&t, 1);
  if (err_flags_out) pl.err.to_host(err_flags_out, 1);
  return t;
}

void systematic_parallel(const PSrc& ps, int64_t n, int64_t nsamp, double rn, int64_t* S_dev, int* err_flags_out) {
  if (err_flags_out) *err_flags_out = 0;
  if (n <= 0 || nsamp <= 0) return;
  EscanPlan pl;
  escan_prepare(ps, n, pl);
  FRICC_LAUNCH(escan_emit_kernel, (unsigned)pl.nchunks, ES_THREADS, 0, ps, n, pl.exact_in.p, nsamp, rn, S_dev);
  if (err_flags_out) pl.err.to_host(err_flags_out, 1);
  else FRICC_CUDA(cudaStreamSynchronize(stream()));
}

}  // namespace fricc
