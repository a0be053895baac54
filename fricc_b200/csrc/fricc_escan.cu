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
#include <algorithm>

#include "fricc_exact.cuh"
#include "fricc_fri.cuh"
#include "fricc_mscan.cuh"
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
escan_map_kernel(PSrc ps, int64_t n, const double* __restrict__ approx_excl, const double* __restrict__ carry,
                 ChunkMap* __restrict__ maps) {
  __shared__ EMap sm[ES_WARPS];
  const int64_t base = (int64_t)blockIdx.x * ES_CHUNK + (int64_t)threadIdx.x * ES_E;
  const double c0 = approx_excl[blockIdx.x] + (carry ? *carry : 0.0);  // carry: approximate sum before this shard
  const bool usable = fx_is_normal_positive(c0);
  const int k = usable ? fx_binade(c0) : 0;
  EMap m = emap_identity();
  if (usable) {
    const double* src = ps.values() + base;
    if (base + ES_E <= n && (reinterpret_cast<uintptr_t>(src) & 15) == 0) {   // 16-byte loads of the thread's 8 elements
      double xv[ES_E];
#pragma unroll
      for (int q = 0; q < ES_E; q += 2) {
        const double2 t = *reinterpret_cast<const double2*>(src + q);
        xv[q] = t.x; xv[q + 1] = t.y;
      }
#pragma unroll
      for (int q = 0; q < ES_E; q++) m = emap_compose(m, emap_of(ps.of_value(xv[q], base + q), k));
    } else {
#pragma unroll
      for (int q = 0; q < ES_E; q++) {
        int64_t j = base + q;
        if (j < n) m = emap_compose(m, emap_of(ps(j), k));
      }
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
                     double* __restrict__ exact_in, const double* carry_in, double* total, int* __restrict__ stats) {
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
    s_c = carry_in ? *carry_in : 0.0;  // exact sum of everything before this shard (may alias `total`)
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
// Up to EMIT_MAX_DRAWS systematic draws from one pass: the exact running sums do not depend on the draw.
struct EmitDraws {
  int nd;
  double rn[EMIT_MAX_DRAWS];
  int64_t* S[EMIT_MAX_DRAWS];
};

// Fast path (a full chunk whose running sum starts normal and stays inside its binade -- all but a few dozen chunks of
// a tensor): everything in registers.  A thread owns ES_E consecutive elements, keeps their one-element maps from the
// first pass (A and the tie flag; B of a one-element map is 0), gets its exact start from the CTA scan and applies the
// kept maps.  Samples are then located per SAMPLE, not per element: the thread's elements cover [excl, incl_last), the
// samples inside are i in [first_sample(excl), first_sample(incl_last)), and sample i lands on the first element whose
// inclusive sum exceeds uk_i = fl(i + rn) -- the same comparisons as the per-element interval test of the general path
// (fri_utils.cpp:147-157), since i -> uk_i is non-decreasing.  ncu (profiles/ncu_full_r02_escan_emit.csv) had the
// general path issue-bound at ~390 instructions per element with 16-way shared-memory bank conflicts on the
// thread-contiguous reads; this path needs neither the staging buffers nor the second evaluation of the maps.
__global__ void __launch_bounds__(ES_THREADS)
escan_emit_kernel(PSrc ps, int64_t n, const double* __restrict__ exact_in, int64_t nsamp, EmitDraws dr) {
  __shared__ double pbuf[ES_CHUNK];
  __shared__ double cs[ES_CHUNK];
  __shared__ ChunkScratch sc;
  const int64_t base = (int64_t)blockIdx.x * ES_CHUNK;
  const int cnt = (int)min((int64_t)ES_CHUNK, n - base);
  const double c_in = exact_in[blockIdx.x];
  if (cnt == ES_CHUNK && fx_is_normal_positive(c_in)) {   // uniform over the CTA
    const int k = fx_binade(c_in);
    const int64_t C0 = fx_to_state(c_in);
    const int64_t j0 = base + (int64_t)threadIdx.x * ES_E;
    double xv[ES_E];
    const double* src = ps.values() + j0;
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
#pragma unroll
      for (int q = 0; q < ES_E; q += 2) {
        const double2 t = *reinterpret_cast<const double2*>(src + q);
        xv[q] = t.x; xv[q + 1] = t.y;
      }
    } else {
#pragma unroll
      for (int q = 0; q < ES_E; q++) xv[q] = src[q];
    }
    int64_t A[ES_E];
    unsigned even = 0;
    EMap mine = emap_identity();
#pragma unroll
    for (int q = 0; q < ES_E; q++) {
      const EMap m = emap_of(ps.of_value(xv[q], j0 + q), k);
      A[q] = m.A;
      even |= (unsigned)m.even << q;
      mine = emap_compose(mine, m);
    }
    const EMap ex = block_excl_scan_emap(mine, sc.warp_maps);
    const int64_t S0 = emap_apply(ex, C0);
    int64_t st[ES_E];
    int64_t S = S0;
#pragma unroll
    for (int q = 0; q < ES_E; q++) {
      S = emap_apply(EMap{A[q], 0, (int32_t)((even >> q) & 1u)}, S);
      st[q] = S;
    }
    // the states only grow, so the CTA's last one tells whether the sum left the binade anywhere in the chunk
    if (!__syncthreads_or(st[ES_E - 1] >= FX_TWO53)) {
      const double excl = fx_from_state(S0, k);
      double incl[ES_E];
#pragma unroll
      for (int q = 0; q < ES_E; q++) incl[q] = fx_from_state(st[q], k);
      if (incl[ES_E - 1] > excl) {
#pragma unroll
        for (int d = 0; d < EMIT_MAX_DRAWS; d++) {
          if (d >= dr.nd) break;
          const double rn = dr.rn[d];
          const int64_t lo = first_sample_at_or_above(excl, rn, nsamp);
          const int64_t hi = first_sample_at_or_above(incl[ES_E - 1], rn, nsamp);
          for (int64_t i = lo; i < hi; i++) {
            const double uk = __dadd_rn((double)i, rn);
            int q = 0;
#pragma unroll
            for (int t = 0; t < ES_E - 1; t++) q += (incl[t] <= uk) ? 1 : 0;
            dr.S[d][i] = ps.j0 + j0 + q;
          }
        }
      }
      return;
    }
    __syncthreads();  // sc.warp_maps is reused below
  }
  for (int k = threadIdx.x; k < cnt; k += ES_THREADS) pbuf[k] = ps(base + k);
  __syncthreads();
  chunk_exact_scan(pbuf, cnt, c_in, cs, &sc);
  // interval of element j is [excl, incl): samples i with excl <= uk_i < incl land on j   (fri_utils.cpp:147-157)
  for (int k = threadIdx.x; k < cnt; k += ES_THREADS) {
    double incl = cs[k];
    double excl = k ? cs[k - 1] : c_in;
    if (incl > excl) {
#pragma unroll
      for (int d = 0; d < EMIT_MAX_DRAWS; d++) {
        if (d >= dr.nd) break;
        int64_t lo = first_sample_at_or_above(excl, dr.rn[d], nsamp);
        int64_t hi = first_sample_at_or_above(incl, dr.rn[d], nsamp);
        for (int64_t i = lo; i < hi; i++) dr.S[d][i] = ps.j0 + base + k;
      }
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

static void escan_stage_approx(const PSrc& ps, int64_t n, EscanPlan& pl) {
  pl.nchunks = ceil_div(n, ES_CHUNK);
  pl.approx.alloc((size_t)pl.nchunks + 1);  // one extra zero addend: the exclusive scan leaves the total behind it
  pl.maps.alloc((size_t)pl.nchunks);
  pl.exact_in.alloc((size_t)pl.nchunks);
  pl.total.alloc(1);
  pl.err.alloc(1);
  pl.stats.alloc(1);
  pl.err.zero();
  FRICC_CUDA(cudaMemsetAsync(pl.approx.p + pl.nchunks, 0, sizeof(double), stream()));
  FRICC_LAUNCH(escan_approx_kernel, (unsigned)pl.nchunks, ES_THREADS, 0, ps, n, pl.approx.p, pl.err.p);
  exclusive_scan_f64(pl.approx.p, pl.nchunks + 1);
}

static void escan_prepare(const PSrc& ps, int64_t n, EscanPlan& pl) {
  escan_stage_approx(ps, n, pl);
  FRICC_LAUNCH(escan_map_kernel, (unsigned)pl.nchunks, ES_THREADS, 0, ps, n, pl.approx.p, (const double*)nullptr,
               pl.maps.p);
  FRICC_LAUNCH(escan_resolve_kernel, 1, RS_THREADS, 0, ps, n, pl.maps.p, pl.nchunks, pl.exact_in.p,
               (const double*)nullptr, pl.total.p, pl.stats.p);
}

// ---- the scan in stages (range-sharded tensors; see fricc_psrc.cuh) -------------------------------------------------
struct ShardScan {
  EscanPlan pl;
  PSrc ps{};
  int64_t n = 0;
};
ShardScan* shard_scan_create() { return new ShardScan(); }
void shard_scan_destroy(ShardScan* s) { delete s; }

void shard_scan_local(ShardScan* s, const PSrc& ps, int64_t n, double* approx_total_dev) {
  s->ps = ps;
  s->n = n;
  if (n <= 0) {
    s->pl.nchunks = 0;
    s->pl.err.alloc(1);
    s->pl.err.zero();
    FRICC_CUDA(cudaMemsetAsync(approx_total_dev, 0, sizeof(double), stream()));
    return;
  }
  escan_stage_approx(ps, n, s->pl);
  FRICC_CUDA(cudaMemcpyAsync(approx_total_dev, s->pl.approx.p + s->pl.nchunks, sizeof(double), cudaMemcpyDeviceToDevice,
                             stream()));
}

void shard_scan_maps(ShardScan* s, const double* approx_carry_dev) {
  if (s->n <= 0) return;
  FRICC_LAUNCH(escan_map_kernel, (unsigned)s->pl.nchunks, ES_THREADS, 0, s->ps, s->n, s->pl.approx.p, approx_carry_dev,
               s->pl.maps.p);
}

void shard_scan_resolve(ShardScan* s, const double* carry_in_dev, double* carry_out_dev) {
  if (s->n <= 0) {
    if (carry_out_dev != carry_in_dev)
      FRICC_CUDA(cudaMemcpyAsync(carry_out_dev, carry_in_dev, sizeof(double), cudaMemcpyDeviceToDevice, stream()));
    return;
  }
  FRICC_LAUNCH(escan_resolve_kernel, 1, RS_THREADS, 0, s->ps, s->n, s->pl.maps.p, s->pl.nchunks, s->pl.exact_in.p,
               carry_in_dev, carry_out_dev, s->pl.stats.p);
}

void shard_scan_emit_multi(ShardScan* s, int64_t nsamp, const double* rn, int n_draws, int64_t* S_dev, int64_t stride) {
  if (s->n <= 0 || nsamp <= 0) return;
  for (int d0 = 0; d0 < n_draws; d0 += EMIT_MAX_DRAWS) {
    EmitDraws dr{};
    dr.nd = std::min(EMIT_MAX_DRAWS, n_draws - d0);
    for (int d = 0; d < dr.nd; d++) {
      dr.rn[d] = rn[d0 + d];
      dr.S[d] = S_dev + (int64_t)(d0 + d) * stride;
    }
    FRICC_LAUNCH(escan_emit_kernel, (unsigned)s->pl.nchunks, ES_THREADS, 0, s->ps, s->n, s->pl.exact_in.p, nsamp, dr);
  }
}

void shard_scan_emit(ShardScan* s, int64_t nsamp, double rn, int64_t* S_dev) {
  shard_scan_emit_multi(s, nsamp, &rn, 1, S_dev, 0);
}

int shard_scan_errors(ShardScan* s) {
  int e = 0;
  s->pl.err.to_host(&e, 1);
  return e;
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
  EmitDraws dr{};
  dr.nd = 1;
  dr.rn[0] = rn;
  dr.S[0] = S_dev;
  FRICC_LAUNCH(escan_emit_kernel, (unsigned)pl.nchunks, ES_THREADS, 0, ps, n, pl.exact_in.p, nsamp, dr);
  if (err_flags_out) pl.err.to_host(err_flags_out, 1);
  else FRICC_CUDA(cudaStreamSynchronize(stream()));
}

// ---------------------------------------------------------------------------------------------------------------------
// Parallel D-selection.  get_d_largest (fri_utils.cpp:400-437) walks the descending candidates with
//     if ((n - d) * x_d >= r - x_d && r > 1e-14) { D.push(d); r -= x_d; } else break;
// i.e. a left-to-right fp64 running SUBTRACTION whose value decides every test.  While r stays in one binade it is
// C * ulp with C in [2^52, 2^53); in the mirrored state M = 2^54 - C a subtraction of x is the round-half-even integer
// ADDITION  M' = RN_even(M + x/ulp)  (2^54 is even, so the parities of C' and M' agree) -- the same maps as the running
// sums above.  One pass = chunk maps -> ordered scan -> every candidate's exact r before it, its test, and whether r
// leaves the binade behind it.  The first event ends the pass: a failed test is the answer; a binade crossing is redone
// with one real subtraction on the host and the pass restarts behind it (at most a few times: r only shrinks).
// ---------------------------------------------------------------------------------------------------------------------
struct EMapOp {
  typedef EMap T;
  __device__ __forceinline__ static EMap identity() { return emap_identity(); }
  __device__ __forceinline__ static EMap op(const EMap& f, const EMap& g) { return emap_compose(f, g); }
  __device__ __forceinline__ static EMap shfl_up(const EMap& m, int o) { return shfl_up_emap(m, o); }
};

constexpr int64_t DG_MIRROR = (int64_t)1 << 54;
constexpr int64_t DG_LEFT = 3 * FX_TWO52;  // M >= DG_LEFT  <=>  C <= 2^52: r is at or below the binade's lower edge
constexpr int DG_THREADS = 256;
constexpr int DG_E = 8;
constexpr int DG_CHUNK = DG_THREADS * DG_E;
constexpr unsigned long long DG_NONE = ~0ULL;

__global__ void __launch_bounds__(DG_THREADS)
dg_chunk_maps_kernel(const uint64_t* __restrict__ keys, int64_t lo, int64_t hi, int k, EMap* __restrict__ maps) {
  __shared__ EMap sm[32];
  const int64_t base = lo + (int64_t)blockIdx.x * DG_CHUNK + (int64_t)threadIdx.x * DG_E;
  EMap m = emap_identity();
#pragma unroll
  for (int q = 0; q < DG_E; q++)
    if (base + q < hi) m = emap_compose(m, emap_of(key_to_double(keys[base + q]), k));
  EMap ex, tot;
  ms_block_scan<EMapOp>(m, sm, &ex, &tot);
  if (threadIdx.x == 0) maps[blockIdx.x] = tot;
}

// rbefore[d] = exact r before candidate d (d in [lo, hi)), rbefore[hi] = r after all of them (valid up to the first event)
// event = min over candidates of  2*d (test fails at d)  /  2*d+1 (test holds, r leaves the binade behind d)
__global__ void __launch_bounds__(DG_THREADS)
dg_apply_kernel(const uint64_t* __restrict__ keys, int64_t lo, int64_t hi, int k, int64_t M0,
                const EMap* __restrict__ chunk_excl, int64_t n_sample, double* __restrict__ rbefore,
                unsigned long long* __restrict__ event) {
  __shared__ EMap sm[32];
  const int64_t base = lo + (int64_t)blockIdx.x * DG_CHUNK + (int64_t)threadIdx.x * DG_E;
  EMap mine = emap_identity();
  uint64_t key[DG_E];
#pragma unroll
  for (int q = 0; q < DG_E; q++) {
    key[q] = (base + q < hi) ? keys[base + q] : 0ULL;
    if (base + q < hi) mine = emap_compose(mine, emap_of(key_to_double(key[q]), k));
  }
  EMap ex, tot;
  ms_block_scan<EMapOp>(mine, sm, &ex, &tot);
  int64_t M = emap_apply(ex, emap_apply(chunk_excl[blockIdx.x], M0));
#pragma unroll
  for (int q = 0; q < DG_E; q++) {
    const int64_t d = base + q;
    if (d >= hi) break;
    if (M > DG_LEFT) break;  // r left the binade before this candidate: an earlier candidate has reported it
    const double r = fx_from_state(DG_MIRROR - M, k);
    const double xi = key_to_double(key[q]);
    rbefore[d] = r;
    // if ((n_sample - d) * xi >= remaining_norm - xi && remaining_norm > 1e-14)   (fri_utils.cpp:415)
    const bool pass = __dmul_rn((double)(n_sample - d), xi) >= __dsub_rn(r, xi) && r > 1e-14;
    if (!pass) {
      atomicMin(event, (unsigned long long)d * 2ULL);
      break;
    }
    M = emap_apply(emap_of(xi, k), M);
    if (M >= DG_LEFT) {  // C' <= 2^52: r - x_d is below the binade, or lands on its lower edge (possibly from below)
      atomicMin(event, (unsigned long long)d * 2ULL + 1ULL);
      break;
    }
    if (d == hi - 1) rbefore[hi] = fx_from_state(DG_MIRROR - M, k);
  }
}

void d_greedy_parallel(const uint64_t* keys, int64_t cand, int64_t n_sample, double norm, int64_t* nD, double* r_out) {
  DevBuf<double> rbefore((size_t)cand + 1);
  DevBuf<unsigned long long> event(1);
  int64_t lo = 0;
  double r = norm;
  while (lo < cand) {
    if (!fx_is_normal_positive(r)) {  // zero / subnormal / non-finite remaining norm: the plain walk from here
      d_greedy_serial(keys, lo, cand, n_sample, r, nD, r_out);
      return;
    }
    const int k = fx_binade(r);
    const int64_t M0 = DG_MIRROR - fx_to_state(r);
    const int64_t nchunks = ceil_div(cand - lo, DG_CHUNK);
    DevBuf<EMap> maps((size_t)nchunks), excl((size_t)nchunks);
    FRICC_LAUNCH(dg_chunk_maps_kernel, (unsigned)nchunks, DG_THREADS, 0, keys, lo, cand, k, maps.p);
    ms_exclusive_scan<EMapOp>(maps.p, nchunks, excl.p);
    FRICC_CUDA(cudaMemsetAsync(event.p, 0xff, sizeof(unsigned long long), stream()));
    FRICC_LAUNCH(dg_apply_kernel, (unsigned)nchunks, DG_THREADS, 0, keys, lo, cand, k, M0, excl.p, n_sample, rbefore.p,
                 event.p);
    unsigned long long ev = 0;
    event.to_host(&ev, 1);
    if (ev == DG_NONE) {
      rbefore.to_host_at(r_out, (size_t)cand);
      *nD = cand;
      return;
    }
    const int64_t e = (int64_t)(ev >> 1);
    double re = 0;
    rbefore.to_host_at(&re, (size_t)e);
    if ((ev & 1ULL) == 0) {  // the test fails at e
      *nD = e;
      *r_out = re;
      return;
    }
    // the test holds at e and r - x_e drops into a lower binade: one real subtraction, continue behind it
    uint64_t ke = 0;
    FRICC_CUDA(cudaMemcpyAsync(&ke, keys + e, 8, cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
    volatile double diff = re - key_to_double(ke);  // IEEE binary64 subtraction (x86-64 SSE2), as the reference does
    r = diff;
    lo = e + 1;
  }
  *nD = cand;
  *r_out = r;
}


// ---------------------------------------------------------------------------------------------------------------------
// fricc_one_norm_static: the exact one_norm with its working memory in a fixed device array of this library.
//
// Who needs it.  Every other entry point takes its working buffers from the library's block cache, i.e. in the end from
// cudaMallocAsync.  Some programs that drive this library may not do that while they run:
//   - a coupled-cluster driver that has handed (nearly) all of the 180 GB to its own arena -- Wvvvv alone is 205 GB at
//     v = 400, sharded to fill each GPU -- has nothing left for the library's pool to grow into, and a failed
//     cudaMallocAsync in the middle of an iteration is not something it can recover from;
//   - a host thread inside a latency-critical section (an MPI/NCCL progress loop, a real-time monitor): an allocation
//     that has to map memory enters the kernel driver, where this project measured stalls of 0.03 - 2.7 s behind other
//     processes' queries (block cache);
//   - a process confined to a fixed memory budget by its launcher (MPS memory limits, a container's device quota), for
//     which every byte must be accounted for before the run starts.
// For those callers the convergence check of an iteration -- |t2new - t2|_1, the one_norm of fri_utils.cpp:164-171 --
// can go through this entry point: same kernels (E1-E3 above), same bits, but the chunk sums, chunk maps and carries
// live in g_norm_scratch below.  The library image carries that array, so every GPU that loads the library holds the
// 1.3 MB from load to exit whether or not this entry point is ever called.
//
// Limits and rules:
//   - n <= FRICC_ONE_NORM_STATIC_MAX_N (2^26 elements = 512 MB of doubles = 32768 chunks); larger inputs are refused
//     with FRICC_EINVAL, they are not split (a split sum would still be exact, but the caller should decide that);
//   - x must already be on the device (a host array would need a staging buffer, i.e. an allocation);
//   - there is ONE array per GPU and every call uses it: two calls that may overlap on the GPU (different streams
//     without an ordering between them, or two host threads) race on it and return garbage.  Calls on one stream are
//     safe -- stream order separates them;
//   - the array keeps what the previous call left there.  The pipeline overwrites all it reads (E1 writes the chunk
//     sums, E2 the maps, E3 the carries) except the one extra addend behind the chunk sums and the error word, which
//     are cleared with cudaMemsetAsync at the start of every call;
//   - the first call on each device looks up the array's device address (cudaGetSymbolAddress, with that device
//     current) and keeps it; that look-up may initialise the module on the device, so make the first call on a device
//     during set-up, before any stream capture is begun.  Later calls only launch kernels, one memset and -- for a host
//     result -- one 8-byte copy.
// No cudaMalloc*, cudaFree*, DevBuf or library allocator is reached from here.
// ---------------------------------------------------------------------------------------------------------------------
constexpr int64_t NS_MAX_N = (int64_t)FRICC_ONE_NORM_STATIC_MAX_N;
constexpr int64_t NS_MAX_CHUNKS = (NS_MAX_N + ES_CHUNK - 1) / ES_CHUNK;  // 32768
constexpr size_t ns_round256(size_t b) { return (b + 255) / 256 * 256; }
constexpr size_t NS_OFF_APPROX = 0;                                                                  // double[chunks + 1]
constexpr size_t NS_OFF_MAPS = NS_OFF_APPROX + ns_round256((NS_MAX_CHUNKS + 1) * sizeof(double));    // ChunkMap[chunks]
constexpr size_t NS_OFF_EXACT = NS_OFF_MAPS + ns_round256(NS_MAX_CHUNKS * sizeof(ChunkMap));         // double[chunks]
constexpr size_t NS_OFF_TOTAL = NS_OFF_EXACT + ns_round256(NS_MAX_CHUNKS * sizeof(double));          // double
constexpr size_t NS_OFF_FLAGS = NS_OFF_TOTAL + 256;                                                  // int err, int stats
constexpr size_t NS_BYTES = NS_OFF_FLAGS + 256;                                                      // 1 311 488 bytes

// In host code this name is only a handle for cudaGetSymbolAddress: it is NOT a device address and must never be
// passed to a kernel or a memset.
static __device__ __align__(256) unsigned char g_norm_scratch[NS_BYTES];

constexpr int NS_MAX_DEVICES = 64;
static unsigned char* g_norm_scratch_addr[NS_MAX_DEVICES];  // per device: the array's address on that device, 0 = not looked up

static unsigned char* norm_scratch_on_current_device() {
  int dev = -1;
  FRICC_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NS_MAX_DEVICES) fail(FRICC_EINVAL, "one_norm_static: device ordinal %d not supported", dev);
  unsigned char* p = __atomic_load_n(&g_norm_scratch_addr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    void* q = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&q, g_norm_scratch);  // the current device's copy of the array
    if (e != cudaSuccess || !q) {
      cudaGetLastError();
      fail(FRICC_ECUDA, "one_norm_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    }
    p = static_cast<unsigned char*>(q);
    __atomic_store_n(&g_norm_scratch_addr[dev], p, __ATOMIC_RELEASE);  // racing first calls store the same value
  }
  return p;
}

// exclusive running sum of d[0..n) in place by ONE block (n <= 32769 here: at most 33 rounds).  These sums only predict
// the binade of the exact running sum at each chunk start; E3 checks every prediction, so their rounding is free.
constexpr int NS_SCAN_THREADS = 1024;
__global__ void __launch_bounds__(NS_SCAN_THREADS) ns_excl_scan_kernel(double* __restrict__ d, int64_t n) {
  __shared__ double s_warp[32];
  __shared__ double s_carry;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_carry = 0.0;
  __syncthreads();
  for (int64_t base = 0; base < n; base += NS_SCAN_THREADS) {
    const int64_t i = base + threadIdx.x;
    const double v = (i < n) ? d[i] : 0.0;
    double inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      double w = s_warp[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const double t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= (unsigned)o) w += t;
      }
      s_warp[lane] = w;
    }
    __syncthreads();
    const double carry = s_carry;
    const double before = carry + (warp ? s_warp[warp - 1] : 0.0) + (inc - v);
    const double round_total = s_warp[31];
    if (i < n) d[i] = before;
    __syncthreads();
    if (threadIdx.x == 0) s_carry = carry + round_total;
    __syncthreads();
  }
}

static void one_norm_static(const double* x_dev, int64_t n, double* norm, int norm_mem) {
  if (n < 0 || !norm || (n > 0 && !x_dev)) fail(FRICC_EINVAL, "one_norm_static: bad arguments");
  if (norm_mem != FRICC_HOST && norm_mem != FRICC_DEVICE) fail(FRICC_EINVAL, "one_norm_static: norm_mem must be FRICC_HOST or FRICC_DEVICE");
  if (n > NS_MAX_N)
    fail(FRICC_EINVAL,
         "one_norm_static: n = %lld exceeds FRICC_ONE_NORM_STATIC_MAX_N = %lld (the fixed scratch array holds %lld chunks); "
         "use fricc_one_norm, or sum ranges of at most that many elements",
         (long long)n, (long long)NS_MAX_N, (long long)NS_MAX_CHUNKS);
  unsigned char* base = norm_scratch_on_current_device();  // first call on this device: the address look-up
  double* approx = reinterpret_cast<double*>(base + NS_OFF_APPROX);
  ChunkMap* maps = reinterpret_cast<ChunkMap*>(base + NS_OFF_MAPS);
  double* exact_in = reinterpret_cast<double*>(base + NS_OFF_EXACT);
  double* total = reinterpret_cast<double*>(base + NS_OFF_TOTAL);
  int* err = reinterpret_cast<int*>(base + NS_OFF_FLAGS);
  int* stats = err + 1;
  double* result_dev = (norm_mem == FRICC_DEVICE) ? norm : total;
  if (n == 0) {
    FRICC_CUDA(cudaMemsetAsync(result_dev, 0, sizeof(double), stream()));  // the empty sum is +0.0
  } else {
    const int64_t nchunks = ceil_div(n, ES_CHUNK);
    PSrc ps{nullptr, x_dev, 0, 0, 0, 0, 1};
    // what must start at zero: the addend behind the last chunk sum, and the error/statistics words (left-overs of the
    // previous call otherwise)
    FRICC_CUDA(cudaMemsetAsync(approx + nchunks, 0, sizeof(double), stream()));
    FRICC_CUDA(cudaMemsetAsync(err, 0, 2 * sizeof(int), stream()));
    FRICC_LAUNCH(escan_approx_kernel, (unsigned)nchunks, ES_THREADS, 0, ps, n, approx, err);
    FRICC_LAUNCH(ns_excl_scan_kernel, 1, NS_SCAN_THREADS, 0, approx, nchunks + 1);
    FRICC_LAUNCH(escan_map_kernel, (unsigned)nchunks, ES_THREADS, 0, ps, n, (const double*)approx, (const double*)nullptr, maps);
    FRICC_LAUNCH(escan_resolve_kernel, 1, RS_THREADS, 0, ps, n, (const ChunkMap*)maps, nchunks, exact_in,
                 (const double*)nullptr, result_dev, stats);
  }
  if (norm_mem == FRICC_HOST) {
    FRICC_CUDA(cudaMemcpyAsync(norm, result_dev, sizeof(double), cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
}

}  // namespace fricc

extern "C" int fricc_one_norm_static(const double* x_dev, int64_t n, double* norm, int norm_mem) {
  FRICC_API_BEGIN
  ::fricc::one_norm_static(x_dev, n, norm, norm_mem);
  FRICC_API_END
}
