// fricc_pivotal.cu -- ordered pivotal sampling (/root/reference/src/fri_utils.cpp:63-127) in PARALLEL, bit for bit.
//
// The reference walks the probabilities once:   a += p_j;  if (a >= 1) { a -= 1; emit a sample }   -- a mod-1 running
// sum in fp64 whose every later decision depends on the exact bits of `a`.  One thread doing that costs ~15 cycles per
// element (2.7 s at N = 2.56e8).  Here the chain is cut at PREDICTED wrap positions and the pieces run concurrently:
//
//   P0  fixed-point fractions.  q_j = RN(p_j * 2^64) as uint64; prefix sums mod 2^64 are frac(sum p) to ~1e-11, and a
//       carry out of  F_{j-1} + q_j  predicts a wrap (a >= 1) at j.                       (pp_segsum, scan, pp_predict)
//   A   every 32-element segment that contains a predicted wrap is an ANCHOR: the chain is restarted right behind its
//       first wrap from the predicted state s0 (a multiple of u = 2^-52, as every post-wrap value is) and run, with the
//       real fp64 operations, up to the next anchor.  Shift invariance: if the true state at the anchor is s0 + d*u
//       (d integer), every value of the run is shifted by exactly d*u -- all intermediate ulps are <= u/2 so an even
//       multiple is added, roundings and parities are unchanged -- except at a wrap step that is an exact tie in
//       [1,2) (ulp u), where an odd d flips the ties-to-even choice by one unit.  So a run acts on the shift as
//           d -> d + g        (d even),       d -> d + g + t   (d odd),   t in {-1,0,+1} from the run's first tie
//       and these maps form a monoid.                                                      (pp_run_kernel<false>)
//   B   ordered parallel scan of the maps -> the exact start state of every run.          (mscan_*)
//   C   every run is repeated from its exact start: per-element decisions (cross / accept / set) are recorded, and the
//       end state of each run is compared, bitwise, with the start state of the next one.   (pp_run_kernel<true>)
//       If all comparisons hold, the pieces ARE the sequential chain (induction from the exact first run).  A mismatch
//       (a predicted wrap that does not happen, a binade boundary or a wrap threshold within |d|*u of a value: ~1e-12
//       per element) makes the first offending anchor a hard restart point with the exact state the verified prefix
//       ends in, and A-C are repeated behind it.  Correctness never rests on the prediction, only the speed does.
//   S   samples: crossing ranks (prefix count) and the current candidate i (prefix max of `set` positions) over the
//       recorded decisions -> S[rank] = accept ? i : j, early stop at n_sample, tail rule.    (pp_segflags, pp_emit)
//
// Compiled with -fmad=false like fricc_fri.cu (the chain and the decisions are the reference's IEEE operations).
#include <algorithm>

#include "fricc_fri.cuh"
#include "fricc_mscan.cuh"
#include "fricc_psrc.cuh"
#include "fricc_sort.cuh"

namespace fricc {

constexpr int PP_SEG = 32;          // elements per segment (one lane of a warp-cooperative pass, one thread of a run)
constexpr int PP_THREADS = 128;
constexpr double PP_U = 2.220446049250313080847263336181640625e-16;       // 2^-52
constexpr double PP_HALF_U = 1.1102230246251565404236316680908203125e-16;  // 2^-53
constexpr double PP_TWO52 = 4503599627370496.0;
constexpr double PP_TWO64 = 18446744073709551616.0;

// decision bits recorded per element
constexpr unsigned char PF_CROSS = 1, PF_SET = 2, PF_BAD = 4, PF_NODRAW = 8;

__device__ __forceinline__ uint64_t pp_q(double p) {  // RN(p * 2^64), saturating; p in [0, 1]
  double s = p * PP_TWO64;                            // exact (power-of-two scaling)
  if (!(s > 0.0)) return 0ULL;
  if (s >= PP_TWO64) return ~0ULL;
  return __double2ull_rn(s);
}

__device__ __forceinline__ uint64_t shfl_up_u64(uint64_t v, int o) {
  return (uint64_t)__shfl_up_sync(0xffffffffu, (unsigned long long)v, o);
}

// ---------------------------------------------------------------------------------------------------------------------
// P0: per-segment sums of q (mod 2^64) and, after their scan, the predicted first wrap of every segment.
// Warp-cooperative: a warp owns 32 consecutive segments; for segment i the 32 lanes read its 32 elements (coalesced).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PP_THREADS)
pp_segsum_kernel(PSrc ps, int64_t n, int64_t nseg, int64_t* __restrict__ segsum, int* __restrict__ err) {
  const unsigned lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * PP_THREADS + threadIdx.x) >> 5;
  const int64_t seg0 = warp * 32;
  if (seg0 >= nseg) return;
  uint64_t mine = 0;
  int bad = 0;
  for (int i = 0; i < 32; i++) {
    const int64_t j = (seg0 + i) * PP_SEG + lane;
    uint64_t q = 0;
    if (j < n) {
      double p = ps(j);
      if (p > 1.0) bad = 1;
      q = pp_q(p);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) q += (uint64_t)__shfl_xor_sync(0xffffffffu, (unsigned long long)q, o);
    if ((int)lane == i) mine = q;
  }
  if (seg0 + lane < nseg) segsum[seg0 + lane] = (int64_t)mine;
  if (bad) atomicOr(err, ERRF_PROB);
}

// fw[t] = offset (0..31) of the first predicted wrap inside segment t, or -1; s0[t] = predicted state right behind it.
// Segment 0 is forced to be an anchor at position 0 with the exact initial state a = p_0.
__global__ void __launch_bounds__(PP_THREADS)
pp_predict_kernel(PSrc ps, int64_t n, int64_t nseg, const int64_t* __restrict__ segstart, int* __restrict__ fw,
                  double* __restrict__ s0) {
  const unsigned lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * PP_THREADS + threadIdx.x) >> 5;
  const int64_t seg0 = warp * 32;
  if (seg0 >= nseg) return;
  int my_fw = -1;
  double my_s0 = 0.0;
  const uint64_t my_start = (seg0 + lane < nseg) ? (uint64_t)segstart[seg0 + lane] : 0ULL;
  for (int i = 0; i < 32; i++) {
    if (seg0 + i >= nseg) break;  // warp-uniform
    const int64_t j = (seg0 + i) * PP_SEG + lane;
    double p = 0.0;
    if (j < n) p = ps(j);
    uint64_t inc = (j < n) ? pp_q(p) : 0ULL;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint64_t t = shfl_up_u64(inc, o);
      if (lane >= (unsigned)o) inc += t;
    }
    const uint64_t F0 = (uint64_t)__shfl_sync(0xffffffffu, (unsigned long long)my_start, i);
    const uint64_t Finc = F0 + inc;              // fraction after element j (mod 2^64)
    uint64_t Fprev = shfl_up_u64(Finc, 1);
    if (lane == 0) Fprev = F0;
    const bool wrap = (j < n) && (j > 0) && (Finc < Fprev || p >= 1.0);
    const unsigned mask = __ballot_sync(0xffffffffu, wrap);
    int f = mask ? (__ffs(mask) - 1) : -1;
    const uint64_t Fw = (uint64_t)__shfl_sync(0xffffffffu, (unsigned long long)Finc, f < 0 ? 0 : f);
    double s = (double)(Fw >> 12) * PP_U;        // multiple of 2^-52 in [0, 1)
    if (seg0 + i == 0) {                         // the chain starts at j = 0 with a = p_0   (fri_utils.cpp:70)
      f = 0;
      s = __shfl_sync(0xffffffffu, p, 0);
    }
    if ((int)lane == i) {
      my_fw = f;
      my_s0 = s;
    }
  }
  if (seg0 + lane < nseg) {
    fw[seg0 + lane] = my_fw;
    s0[seg0 + lane] = my_s0;
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// the shift maps and their ordered scans
// ---------------------------------------------------------------------------------------------------------------------
struct PMap {  // d -> d + c0 (d even),  d -> d + c1 (d odd)
  long long c0, c1;
};
struct PMapOp {
  typedef PMap T;
  __device__ __forceinline__ static PMap identity() { return PMap{0, 0}; }
  __device__ __forceinline__ static PMap op(const PMap& f, const PMap& g) {  // f first, then g
    PMap h;
    h.c0 = f.c0 + ((f.c0 & 1LL) ? g.c1 : g.c0);
    h.c1 = f.c1 + (((1LL + f.c1) & 1LL) ? g.c1 : g.c0);
    return h;
  }
  __device__ __forceinline__ static PMap shfl_up(const PMap& v, int o) {
    return PMap{__shfl_up_sync(0xffffffffu, v.c0, o), __shfl_up_sync(0xffffffffu, v.c1, o)};
  }
};
struct MaxOp {  // prefix maximum of positions; -1 = none
  typedef long long T;
  __device__ __forceinline__ static long long identity() { return -1; }
  __device__ __forceinline__ static long long op(long long a, long long b) { return a > b ? a : b; }
  __device__ __forceinline__ static long long shfl_up(long long v, int o) { return __shfl_up_sync(0xffffffffu, v, o); }
};

// ---------------------------------------------------------------------------------------------------------------------
// A / C: the runs.  One thread per segment; a thread whose segment is an anchor walks the chain from the position behind
// its anchor up to (and including) the next anchor's position.
// ---------------------------------------------------------------------------------------------------------------------
struct RunArgs {
  int64_t n, nseg, seg_lo;        // segments below seg_lo are final (the verified prefix); seg_lo is the exact restart
  const int* fw;
  const double* s0;
  // A: out
  PMap* maps;
  // C: in
  const PMap* delta;              // exclusive scan of the maps (c0 = shift of the run's start)
  const int32_t* draws;
  int64_t ndraws;
  // C: out
  unsigned char* flags;           // per element decisions
  double* exact_at;               // exact state behind each anchor position, written by the run that ENDS there
  unsigned long long* first_bad;  // smallest anchor segment whose start state disagrees with the chain arriving there
  double* final_a;                // state after element n-1
  int* err;
};

template <bool EXACT>
__global__ void __launch_bounds__(PP_THREADS)
pp_run_kernel(PSrc ps, RunArgs ra) {
  const int64_t t = ra.seg_lo + (int64_t)blockIdx.x * PP_THREADS + threadIdx.x;
  if (t >= ra.nseg) return;
  const int f = ra.fw[t];
  if (f < 0) {
    if (!EXACT) ra.maps[t] = PMap{0, 0};
    return;
  }
  // next anchor
  int64_t tn = t + 1;
  while (tn < ra.nseg && ra.fw[tn] < 0) tn++;
  const int64_t P = t * PP_SEG + f;
  const int64_t Pend = (tn < ra.nseg) ? tn * PP_SEG + ra.fw[tn] : ra.n - 1;
  double a = ra.s0[t];
  if (EXACT) {
    a = __dadd_rn(a, __dmul_rn((double)ra.delta[t].c0, PP_U));  // exact: both are multiples of 2^-52 (delta = 0 at seg_lo)
    if (t != ra.seg_lo && !(a >= 0.0 && a <= 1.0)) atomicMin(ra.first_bad, (unsigned long long)t);
  }
  int tie = 0;
  bool last_cross = false;
  int errf = 0;
  for (int64_t j = P + 1; j <= Pend; j++) {
    const double p = ps(j);
    const double apre = __dadd_rn(a, p);          // a += probs[j]
    const bool c = !(apre < 1.0);                 // else-branch of `if (a < 1.)`
    const double anew = c ? __dsub_rn(apre, 1.0) : apre;  // a -= 1.
    if (!EXACT) {
      if (c && tie == 0 && apre < 2.0) {
        // exact residual of a + p (TwoSum): an exact tie between two multiples of 2^-52 is |res| == 2^-53
        const double bb = __dsub_rn(apre, a);
        const double res = __dadd_rn(__dsub_rn(a, __dsub_rn(apre, bb)), __dsub_rn(p, bb));
        if (res == PP_HALF_U) tie = 1;            // rounded down to the even neighbour; an odd shift rounds up
        else if (res == -PP_HALF_U) tie = -1;
      }
    } else {
      unsigned char fl = 0;
      double rnd = 0.0;
      if (j - 1 < ra.ndraws) rnd = __ddiv_rn((double)ra.draws[j - 1], 2147483647.0);  // (double)rand() / RAND_MAX
      else fl |= PF_NODRAW;
      if (!c) {
        if (rnd < __ddiv_rn(p, anew)) fl |= PF_SET;                                   // rn < probs[j] / a
      } else {
        const double pa = __ddiv_rn(__dsub_rn(1.0, p), __dsub_rn(1.0, anew));         // (1. - probs[j]) / (1. - a)
        fl |= PF_CROSS;
        if (pa > 1.0 || pa < 0.0) fl |= PF_BAD;
        if (rnd < pa) fl |= PF_SET;
      }
      ra.flags[j] = fl;
    }
    a = anew;
    last_cross = c;
  }
  if (!EXACT) {
    PMap m{0, 0};
    if (tn < ra.nseg && last_cross) {
      const double d = __dsub_rn(a, ra.s0[tn]);   // both multiples of 2^-52, close: exact
      const long long g = (long long)__dmul_rn(d, PP_TWO52);
      m.c0 = g;
      m.c1 = g + tie;
    }
    ra.maps[t] = m;
  } else {
    if (tn < ra.nseg) {
      ra.exact_at[tn] = a;
      const double expect = __dadd_rn(ra.s0[tn], __dmul_rn((double)ra.delta[tn].c0, PP_U));
      if (__double_as_longlong(expect) != __double_as_longlong(a)) atomicMin(ra.first_bad, (unsigned long long)tn);
    } else {
      *ra.final_a = a;
    }
    if (errf) atomicOr(ra.err, errf);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// S: from the recorded decisions to the sample list.  Warp-cooperative like P0 (32 lanes = the 32 elements of a segment).
// ---------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PP_THREADS)
pp_segflags_kernel(const unsigned char* __restrict__ flags, int64_t n, int64_t nseg, uint32_t* __restrict__ ncross,
                   long long* __restrict__ lastset) {
  const unsigned lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * PP_THREADS + threadIdx.x) >> 5;
  const int64_t seg0 = warp * 32;
  if (seg0 >= nseg) return;
  uint32_t my_c = 0;
  long long my_l = -1;
  for (int i = 0; i < 32; i++) {
    if (seg0 + i >= nseg) break;
    const int64_t j = (seg0 + i) * PP_SEG + lane;
    const unsigned char fl = (j < n && j > 0) ? flags[j] : 0;
    const unsigned cm = __ballot_sync(0xffffffffu, fl & PF_CROSS);
    const unsigned sm = __ballot_sync(0xffffffffu, fl & PF_SET);
    if ((int)lane == i) {
      my_c = __popc(cm);
      my_l = sm ? ((seg0 + i) * PP_SEG + (31 - __clz(sm))) : -1;
    }
  }
  if (seg0 + lane < nseg) {
    ncross[seg0 + lane] = my_c;
    lastset[seg0 + lane] = my_l;
  }
}

struct EmitOut {
  long long visited;  // position of the n_sample-th crossing (-1: not reached)
};

__global__ void __launch_bounds__(PP_THREADS)
pp_emit_kernel(const unsigned char* __restrict__ flags, int64_t n, int64_t nseg, const uint32_t* __restrict__ rank0,
               const long long* __restrict__ cand0, int64_t nsamp, int64_t* __restrict__ S,
               EmitOut* __restrict__ eo, int* __restrict__ err) {
  const unsigned lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * PP_THREADS + threadIdx.x) >> 5;
  const int64_t seg0 = warp * 32;
  if (seg0 >= nseg) return;
  int errf = 0;
  for (int i = 0; i < 32; i++) {
    if (seg0 + i >= nseg) break;
    const int64_t seg = seg0 + i;
    const int64_t j = seg * PP_SEG + lane;
    const unsigned char fl = (j < n && j > 0) ? flags[j] : 0;
    const unsigned cm = __ballot_sync(0xffffffffu, fl & PF_CROSS);
    const unsigned sm = __ballot_sync(0xffffffffu, fl & PF_SET);
    const long long base_rank = (long long)rank0[seg];
    if (base_rank >= nsamp) continue;  // warp-uniform: everything behind the n_sample-th crossing is never visited
    const unsigned below = (1u << lane) - 1u;
    const long long rank = base_rank + __popc(cm & below);
    if ((fl & PF_NODRAW) && rank < nsamp) errf |= ERRF_DRAWS;  // (visited elements only; refined on the host)
    if (fl & PF_CROSS) {
      if (rank < nsamp) {
        // current candidate i: the last `set` position strictly before j   (initially i = 0, fri_utils.cpp:66)
        const unsigned sb = sm & below;
        long long icur = sb ? (seg * PP_SEG + (31 - __clz(sb))) : cand0[seg];
        if (icur < 0) icur = 0;
        if (fl & PF_BAD) errf |= ERRF_ACCEPT;
        S[rank] = (fl & PF_SET) ? icur : j;  // accepted: S.push_back(i), i = j;   rejected: S.push_back(j)
        if (rank == nsamp - 1) eo->visited = j;
      }
    }
  }
  if (errf) atomicOr(err, errf);
}

__global__ void pp_finalize_kernel(int64_t n, int64_t nseg, int64_t nsamp, const uint32_t* __restrict__ rank0,
                                   const long long* __restrict__ cand0, const double* __restrict__ final_a,
                                   const EmitOut* __restrict__ eo, int64_t* __restrict__ S,
                                   PivotalState* __restrict__ st) {
  long long count = (long long)rank0[nseg];  // total crossings
  long long isel = cand0[nseg];
  if (isel < 0) isel = 0;
  long long visited;
  if (n < 2) visited = 0;
  else if (nsamp == 0) visited = 1;  // the loop body runs once, then S.size()==0==n_sample breaks
  else visited = (eo->visited >= 0) ? eo->visited : (n - 1);
  const double a = *final_a;
  if (count > nsamp) count = nsamp;
  if (count < nsamp && n >= 2) {
    // reached j == N-1 without filling S: `if (abs(a - 1.) < 1e-5) S.push_back(i)`   (fri_utils.cpp:106-118)
    if (fabs(__dsub_rn(a, 1.0)) < 1e-5) {
      S[count] = isel;
      count++;
    }
  }
  st->a = a;
  st->isel = isel;
  st->count = count;
  st->visited = visited;
}

// ---------------------------------------------------------------------------------------------------------------------
// host
// ---------------------------------------------------------------------------------------------------------------------
static inline unsigned pp_warp_grid(int64_t nseg) {  // warp-cooperative kernels: one warp per 32 segments
  return (unsigned)std::max<int64_t>(1, ceil_div(ceil_div(nseg, 32) * 32, PP_THREADS));
}

bool pivotal_parallel(const PSrc& ps, int64_t n, int64_t nsamp, const int32_t* draws_dev, int64_t ndraws,
                      int64_t* S_dev, PivotalState* st_dev, int* err_dev, int* restarts_out) {
  if (restarts_out) *restarts_out = 0;
  if (n < 1) return false;
  const int64_t nseg = ceil_div(n, PP_SEG);
  DevBuf<int64_t> segsum((size_t)nseg);
  DevBuf<int> fw((size_t)nseg);
  DevBuf<double> s0((size_t)nseg), exact_at((size_t)nseg), final_a(1);
  DevBuf<PMap> maps((size_t)nseg), delta((size_t)nseg);
  DevBuf<unsigned char> flags((size_t)n);
  DevBuf<unsigned long long> first_bad(1);

  // ---- P0: fixed-point fractions -> predicted wraps
  FRICC_LAUNCH(pp_segsum_kernel, pp_warp_grid(nseg), PP_THREADS, 0, ps, n, nseg, segsum.p, err_dev);
  exclusive_scan_i64(segsum.p, nseg);  // two's-complement wrap-around = arithmetic mod 2^64
  FRICC_LAUNCH(pp_predict_kernel, pp_warp_grid(nseg), PP_THREADS, 0, ps, n, nseg, segsum.p, fw.p, s0.p);

  // ---- A / B / C, repeated behind a hard restart point while some run does not join the next one
  RunArgs ra{};
  ra.n = n; ra.nseg = nseg; ra.seg_lo = 0;
  ra.fw = fw.p; ra.s0 = s0.p; ra.maps = maps.p; ra.delta = delta.p;
  ra.draws = draws_dev; ra.ndraws = ndraws;
  ra.flags = flags.p; ra.exact_at = exact_at.p; ra.first_bad = first_bad.p; ra.final_a = final_a.p; ra.err = err_dev;
  int restarts = 0;
  while (true) {
    const unsigned grid = (unsigned)std::max<int64_t>(1, ceil_div(nseg - ra.seg_lo, PP_THREADS));
    if (ra.seg_lo > 0) FRICC_CUDA(cudaMemsetAsync(maps.p, 0, (size_t)ra.seg_lo * sizeof(PMap), stream()));
    FRICC_LAUNCH(pp_run_kernel<false>, grid, PP_THREADS, 0, ps, ra);
    ms_exclusive_scan<PMapOp>(maps.p, nseg, delta.p);
    FRICC_CUDA(cudaMemsetAsync(first_bad.p, 0xff, sizeof(unsigned long long), stream()));
    FRICC_LAUNCH(pp_run_kernel<true>, grid, PP_THREADS, 0, ps, ra);
    unsigned long long fb = 0;
    first_bad.to_host(&fb, 1);
    if (fb >= (unsigned long long)nseg) break;
    // the chain is verified up to the anchor of segment fb; restart there from the exact state it arrives with
    if ((int64_t)fb <= ra.seg_lo || ++restarts > 256) return false;  // no progress: the caller uses the one-thread chain
    ra.seg_lo = (int64_t)fb;
    FRICC_CUDA(cudaMemcpyAsync(s0.p + fb, exact_at.p + fb, sizeof(double), cudaMemcpyDeviceToDevice, stream()));
  }
  if (restarts_out) *restarts_out = restarts;

  // ---- S: decisions -> samples
  DevBuf<uint32_t> ncross((size_t)nseg + 1);
  DevBuf<long long> lastset((size_t)nseg + 1), cand0((size_t)nseg + 1);
  DevBuf<EmitOut> eo(1);
  FRICC_CUDA(cudaMemsetAsync(ncross.p + nseg, 0, sizeof(uint32_t), stream()));
  FRICC_CUDA(cudaMemsetAsync(lastset.p + nseg, 0xff, sizeof(long long), stream()));  // -1
  FRICC_CUDA(cudaMemsetAsync(eo.p, 0xff, sizeof(EmitOut), stream()));                // visited = -1
  FRICC_LAUNCH(pp_segflags_kernel, pp_warp_grid(nseg), PP_THREADS, 0, flags.p, n, nseg, ncross.p, lastset.p);
  exclusive_scan_u32(ncross.p, nseg + 1);
  ms_exclusive_scan<MaxOp>(lastset.p, nseg + 1, cand0.p);
  FRICC_LAUNCH(pp_emit_kernel, pp_warp_grid(nseg), PP_THREADS, 0, flags.p, n, nseg, ncross.p, cand0.p, nsamp, S_dev,
               eo.p, err_dev);
  FRICC_LAUNCH(pp_finalize_kernel, 1, 1, 0, n, nseg, nsamp, ncross.p, cand0.p, final_a.p, eo.p, S_dev, st_dev);
  return true;
}

}  // namespace fricc
