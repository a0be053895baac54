// fricc_fri.cu -- FRI compression on the GPU: norm, D-selection, probabilities, pivotal / systematic samplers.
//
// Compiled with -fmad=false: every product/quotient/sum that decides a sample is the same IEEE binary64 operation,
// in the same order, as the reference's x86-64 build (no FMA contraction; /root/reference/CMakeLists.txt:41).
//
//   reference                                   here
//   one_norm            fri_utils.cpp:164-171   one_norm_sequential   (left-to-right sum, = the 1-thread reference)
//   get_d_largest       fri_utils.cpp:400-437   select_top_m + d_greedy_kernel
//   make_probability_v. fri_utils.cpp:367-398   PSrc::operator() (never materialised on the compression path)
//   sample_systematic   fri_utils.cpp:129-159   systematic_chain_kernel (exact running sum + interval -> sample map)
//   sample_pivotal      fri_utils.cpp:63-127    pivotal_chain_kernel    (exact mod-1 running sum + scan-form decisions)
//   fri_compression     fri_utils.cpp:439-522   compress_fri
#include "fricc_fri.cuh"
#include "fricc_psrc.cuh"

#include <chrono>
#include <cmath>

namespace fricc {

// ---------------------------------------------------------------------------------------------------------------
// exact sequential running sums: one CTA; all threads stage a tile of terms in shared memory, thread 0 performs the
// left-to-right fp64 additions (the only order-dependent part of either sampler), all threads consume the result.
// ---------------------------------------------------------------------------------------------------------------
constexpr int CH_THREADS = 512;
constexpr int CH_TILE = 4096;

__global__ void __launch_bounds__(CH_THREADS)
norm_chain_kernel(PSrc ps, int64_t n, double* __restrict__ out) {
  __shared__ double pbuf[CH_TILE];
  double carry = 0.0;
  for (int64_t base = 0; base < n; base += CH_TILE) {
    const int cnt = (int)min((int64_t)CH_TILE, n - base);
    for (int k = threadIdx.x; k < cnt; k += CH_THREADS) pbuf[k] = ps(base + k);
    __syncthreads();
    if (threadIdx.x == 0) {
      double a = carry;
#pragma unroll 8
      for (int k = 0; k < cnt; k++) a = __dadd_rn(a, pbuf[k]);
      carry = a;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = carry;
}

__global__ void __launch_bounds__(CH_THREADS)
systematic_chain_kernel(PSrc ps, int64_t n, int64_t nsamp, double rn, int64_t* __restrict__ S, int* __restrict__ err) {
  __shared__ double pbuf[CH_TILE];
  __shared__ double s_carry;
  double carry = 0.0;
  int bad = 0;
  for (int64_t base = 0; base < n; base += CH_TILE) {
    const int cnt = (int)min((int64_t)CH_TILE, n - base);
    for (int k = threadIdx.x; k < cnt; k += CH_THREADS) {
      double p = ps(base + k);
      if (p > 1.0) bad = 1;
      pbuf[k] = p;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      // intervals[j+1] = intervals[j] + p[j]   (std::partial_sum, fri_utils.cpp:142-143)
      double a = carry;
#pragma unroll 8
      for (int k = 0; k < cnt; k++) {
        a = __dadd_rn(a, pbuf[k]);
        pbuf[k] = a;
      }
      s_carry = a;
    }
    __syncthreads();
    for (int k = threadIdx.x; k < cnt; k += CH_THREADS) {
      double incl = pbuf[k];
      double excl = k ? pbuf[k - 1] : carry;
      if (incl > excl) {
        // samples i with intervals[j] <= uk_i < intervals[j+1] land on element j (fri_utils.cpp:147-157)
        int64_t lo = first_sample_at_or_above(excl, rn, nsamp);
        int64_t hi = first_sample_at_or_above(incl, rn, nsamp);
        for (int64_t i = lo; i < hi; i++) S[i] = base + k;
      }
    }
    double nc = s_carry;
    __syncthreads();
    carry = nc;
  }
  if (bad) atomicOr(err, ERRF_PROB);
}

// --- pivotal -----------------------------------------------------------------------------------------------------
constexpr int PV_THREADS = 512;
constexpr int PV_IPT = 4;
constexpr int PV_TILE = PV_THREADS * PV_IPT;


__global__ void __launch_bounds__(PV_THREADS)
pivotal_chain_kernel(PSrc ps, int64_t n, int64_t nsamp, const int32_t* __restrict__ draws, int64_t ndraws,
                     int64_t* __restrict__ S, PivotalState* __restrict__ st, int* __restrict__ err) {
  __shared__ double pbuf[PV_TILE];
  __shared__ double abuf[PV_TILE];
  __shared__ unsigned char cbuf[PV_TILE];
  __shared__ int s_cnt[PV_THREADS / 32];
  __shared__ long long s_max[PV_THREADS / 32];
  __shared__ double s_a;
  __shared__ long long s_visited;
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

  int errf = 0;
  if (threadIdx.x == 0) {
    double p0 = ps(0);
    if (p0 > 1.0) errf |= ERRF_PROB;
    s_a = p0;  // a = probs[0]   (fri_utils.cpp:70)
    s_visited = -1;
  }
  __syncthreads();
  long long isel = 0;   // i = 0
  long long count = 0;  // S.size()
  bool done = false;

  for (int64_t base = 1; base < n && !done; base += PV_TILE) {
    const int cnt = (int)min((int64_t)PV_TILE, n - base);
    for (int k = threadIdx.x; k < cnt; k += PV_THREADS) {
      double p = ps(base + k);
      if (p > 1.0) errf |= ERRF_PROB;
      pbuf[k] = p;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      double a = s_a;
#pragma unroll 4
      for (int k = 0; k < cnt; k++) {
        double apre = __dadd_rn(a, pbuf[k]);  // a += probs[j]
        bool c = !(apre < 1.0);               // else-branch of `if (a < 1.)`
        a = c ? __dsub_rn(apre, 1.0) : apre;  // a -= 1.
        abuf[k] = a;
        cbuf[k] = c;
      }
      s_a = a;
    }
    __syncthreads();

    // decisions for this thread's PV_IPT consecutive elements
    const int k0 = threadIdx.x * PV_IPT;
    bool setf[PV_IPT], crossf[PV_IPT], accf[PV_IPT], badf[PV_IPT];
    int ncross = 0;
    long long lastset = -1;
#pragma unroll
    for (int q = 0; q < PV_IPT; q++) {
      int k = k0 + q;
      setf[q] = crossf[q] = accf[q] = badf[q] = false;
      if (k < cnt) {
        int64_t j = base + k;
        double rnd = 0.0;
        if (j - 1 < ndraws) rnd = __ddiv_rn((double)draws[j - 1], 2147483647.0);  // (double)rand() / RAND_MAX
        else errf |= ERRF_DRAWS;
        double p = pbuf[k], a = abuf[k];
        if (!cbuf[k]) {
          setf[q] = rnd < __ddiv_rn(p, a);  // rn < probs[j] / a
        } else {
          double pa = __ddiv_rn(__dsub_rn(1.0, p), __dsub_rn(1.0, a));  // (1. - probs[j]) / (1. - a)
          badf[q] = (pa > 1.0 || pa < 0.0);
          accf[q] = rnd < pa;
          setf[q] = accf[q];
          crossf[q] = true;
          ncross++;
        }
        if (setf[q]) lastset = j;
      }
    }
    // block-wide exclusive prefix: number of crossings, and last `set` index
    int inc = ncross;
    long long mx = lastset;
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, inc, o);
      long long u = __shfl_up_sync(0xffffffffu, mx, o);
      if (lane >= (unsigned)o) {
        inc += t;
        mx = max(mx, u);
      }
    }
    if (lane == 31) {
      s_cnt[warp] = inc;
      s_max[warp] = mx;
    }
    __syncthreads();
    int woff = 0, total = 0;
    long long wmax = -1, tmax = -1;
    for (unsigned w = 0; w < PV_THREADS / 32; w++) {
      if (w < warp) {
        woff += s_cnt[w];
        wmax = max(wmax, s_max[w]);
      }
      total += s_cnt[w];
      tmax = max(tmax, s_max[w]);
    }
    int rank = woff + inc - ncross;  // crossings before this thread's first element (within the tile)
    long long prev = __shfl_up_sync(0xffffffffu, mx, 1);
    long long ib = (lane == 0) ? -1 : prev;  // last set among earlier lanes of this warp
    ib = max(ib, wmax);
    long long icur = (ib >= 0) ? ib : isel;  // candidate i before this thread's first element
#pragma unroll
    for (int q = 0; q < PV_IPT; q++) {
      int k = k0 + q;
      if (k < cnt) {
        int64_t j = base + k;
        if (crossf[q]) {
          long long pos = count + rank;
          if (pos < nsamp) {
            if (badf[q]) errf |= ERRF_ACCEPT;
            S[pos] = accf[q] ? icur : j;  // S.push_back(i) : S.push_back(j)
            if (pos == nsamp - 1) s_visited = j;  // `if (S.size() == n_sample) break;`
          }
          rank++;
        }
        if (setf[q]) icur = j;
      }
    }
    count += total;
    if (tmax >= 0) isel = tmax;
    __syncthreads();
    done = (count >= nsamp);
  }

  if (errf) atomicOr(err, errf);
  if (threadIdx.x == 0) {
    long long visited;
    if (n < 2) visited = 0;
    else if (nsamp == 0) visited = 1;  // the loop body runs once, then S.size()==0==n_sample breaks
    else visited = (s_visited >= 0) ? s_visited : (n - 1);
    double a = s_a;
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
}

// ---------------------------------------------------------------------------------------------------------------
// D-selection (greedy over the descending candidate list) -- one warp, lanes prefetch, the test itself is sequential
// ---------------------------------------------------------------------------------------------------------------
// keys[0..ncand) = the ncand largest magnitudes in descending order; n_sample = the reference's n (m_keep); the walk
// starts at candidate `lo` with remaining norm r_start.
__global__ void d_greedy_kernel(const uint64_t* __restrict__ keys, int64_t lo, int64_t ncand, int64_t n_sample,
                                double r_start, long long* __restrict__ nD, double* __restrict__ rout) {
  const unsigned lane = threadIdx.x & 31;
  double r = r_start;
  long long d = lo;
  bool stop = false;
  for (int64_t base = lo; base < ncand && !stop; base += 32) {
    uint64_t key = (base + lane < ncand) ? keys[base + lane] : 0;
    double mine = key_to_double(key);
    int lim = (int)min((int64_t)32, ncand - base);
    for (int t = 0; t < lim; t++) {
      double xi = __shfl_sync(0xffffffffu, mine, t);
      // if ((n_sample - d) * xi >= remaining_norm - xi && remaining_norm > 1e-14)   (fri_utils.cpp:415)
      if (__dmul_rn((double)(n_sample - d), xi) >= __dsub_rn(r, xi) && r > 1e-14) {
        d++;
        r = __dsub_rn(r, xi);
      } else {
        stop = true;
        break;
      }
    }
  }
  if (lane == 0) {
    *nD = d;
    *rout = r;
  }
}

// ---------------------------------------------------------------------------------------------------------------
// elementwise helpers
// ---------------------------------------------------------------------------------------------------------------
__global__ void prob_vector_kernel(const double* __restrict__ x, int64_t n, double r, double ns, double* __restrict__ p) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = __dmul_rn(__ddiv_rn(fabs(x[i]), r), ns);
}
__global__ void zero_at_kernel(double* __restrict__ p, const int64_t* __restrict__ D, int64_t nD) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nD) p[D[i]] = 0.0;
}
__global__ void check_prob_kernel(const double* __restrict__ p, int64_t n, int* __restrict__ err) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && p[i] > 1.0) atomicOr(err, ERRF_PROB);
}

// D part: value = x[idx]; S part: value = x[idx] / p[idx]   (fri_utils.cpp:498-504)
__global__ void gather_values_kernel(PSrc ps, const uint64_t* __restrict__ cand_idx, int64_t nD,
                                     const int64_t* __restrict__ S, int64_t m, int64_t* __restrict__ flat,
                                     double* __restrict__ vals) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  if (e < nD) {
    int64_t idx = (int64_t)cand_idx[e];
    flat[e] = idx;
    vals[e] = ps.x[idx];
  } else {
    int64_t idx = S[e - nD];
    flat[e] = idx;
    vals[e] = __ddiv_rn(ps.x[idx], ps(idx));
  }
}
__global__ void gather_largest_kernel(const double* __restrict__ x, const uint64_t* __restrict__ cand_idx, int64_t m,
                                      int64_t* __restrict__ flat, double* __restrict__ vals) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  int64_t idx = (int64_t)cand_idx[e];
  flat[e] = idx;
  vals[e] = x[idx];
}

static inline unsigned blocks_for(int64_t n, int threads) { return (unsigned)std::max<int64_t>(1, ceil_div(n, threads)); }

// ---------------------------------------------------------------------------------------------------------------
// host orchestration
// ---------------------------------------------------------------------------------------------------------------
// FRICC_ESCAN=0 selects the single-thread chain kernels (reference implementation of the exact sums, kept for A/B
// testing); the default is the parallel binade-segmented scan of fricc_escan.cu.  Both are bit-identical.
static int64_t g_first_candidates = 1 << 16;  // first size of the D candidate list ("d_candidates" option)
static int g_d_parallel = -1;                 // "d_parallel" option: -1 auto (lists >= 32768), 0 serial, 1 parallel
void set_d_parallel(int mode) { g_d_parallel = mode; }
static int g_d_estimate = 1;                  // size the list from the histogram estimate ("d_estimate" option)
void set_first_candidates(int64_t c) { g_first_candidates = c > 0 ? c : 1; }
void set_d_estimate(int on) { g_d_estimate = on ? 1 : 0; }
static int g_pivotal_parallel = -1;  // "pivotal_parallel" option: -1 auto, 0 chain kernel, 1 always parallel
static int g_pivotal_restarts = 0;
void set_pivotal_parallel(int mode) { g_pivotal_parallel = mode; }
int pivotal_restarts_last() { return g_pivotal_restarts; }
static int g_escan = -1;
void set_parallel_scan(int on) { g_escan = on ? 1 : 0; }
static bool use_parallel_scan() {
  if (g_escan < 0) {
    const char* e = getenv("FRICC_ESCAN");
    g_escan = (e && e[0] == '0') ? 0 : 1;
  }
  return g_escan == 1;
}

double one_norm_sequential(const double* x, int64_t n) {
  if (n <= 0) return 0.0;
  PSrc ps{nullptr, x, 0, 0, 0, 0, 1};
  if (use_parallel_scan()) return exact_sum_parallel(ps, n, nullptr);
  DevBuf<double> out(1);
  FRICC_LAUNCH(norm_chain_kernel, 1, CH_THREADS, 0, ps, n, out.p);
  double h = 0;
  out.to_host(&h, 1);
  return h;
}

void d_greedy_serial(const uint64_t* keys, int64_t lo, int64_t cand, int64_t n_sample, double r_start, int64_t* nD,
                     double* remaining_norm) {
  DevBuf<long long> d(1);
  DevBuf<double> r(1);
  FRICC_LAUNCH(d_greedy_kernel, 1, 32, 0, keys, lo, cand, n_sample, r_start, d.p, r.p);
  long long hd = 0;
  d.to_host(&hd, 1);
  r.to_host(remaining_norm, 1);
  *nD = hd;
}

void d_greedy(const TopM& top, int64_t n_sample, double norm, int64_t* nD, double* remaining_norm) {
  // short lists: the one-warp walk; long ones: the parallel rounding-map scan (bit-identical, "d_parallel" option)
  if (g_d_parallel == 1 || (g_d_parallel < 0 && top.m >= (1 << 15)))
    d_greedy_parallel(top.keys.p, top.m, n_sample, norm, nD, remaining_norm);
  else
    d_greedy_serial(top.keys.p, 0, top.m, n_sample, norm, nD, remaining_norm);
}

static void raise_flags(int flags) {
  if (flags & ERRF_PROB)
    fail(FRICC_EPROB, "ERROR MAKING PROBABILITY VECTOR: some p[i] > 1 (reference: fri_utils.cpp:383-395)");
  if (flags & ERRF_DRAWS) fail(FRICC_EDRAWS, "draw buffer too short for the pivotal sampler");
  if (flags & ERRF_ACCEPT)
    fail(FRICC_EPIVOTAL, "ERROR IN PIVOTAL SAMPLING: acceptance probability outside [0,1] (reference: fri_utils.cpp:84-94)");
}

void raise_error_flags(int flags) { raise_flags(flags); }

void make_probability_vector(const double* x, int64_t n, int64_t n_sample, const int64_t* D, int64_t nD,
                             double remaining_norm, double* p) {
  if (n <= 0) return;
  FRICC_LAUNCH(prob_vector_kernel, blocks_for(n, 256), 256, 0, x, n, remaining_norm, (double)n_sample, p);
  if (nD > 0) FRICC_LAUNCH(zero_at_kernel, blocks_for(nD, 256), 256, 0, p, D, nD);
  DevBuf<int> err(1);
  err.zero();
  FRICC_LAUNCH(check_prob_kernel, blocks_for(n, 256), 256, 0, p, n, err.p);
  int h = 0;
  err.to_host(&h, 1);
  raise_flags(h);
}

// fetch `count` draws starting at the current position of the source; for the library stream the caller later
// rewinds with rand_restore + rand_fill(consumed)
static void fetch_draws(const DrawSource& src, int64_t count, std::vector<int32_t>& out) {
  out.resize((size_t)count);
  if (src.from_stream) {
    rand_fill(out.data(), count);
  } else {
    if (src.n < count) out.resize((size_t)src.n);
    if (!out.empty()) memcpy(out.data(), src.host, out.size() * sizeof(int32_t));
  }
}

static void run_systematic(const PSrc& ps, int64_t n, int64_t nsamp, const DrawSource& draws, int64_t* consumed,
                           int64_t* S_dev) {
  // lin_space_add_const draws exactly one uniform, whatever n_sample is   (fri_utils.cpp:47)
  std::vector<int32_t> d;
  fetch_draws(draws, 1, d);
  if (d.empty()) fail(FRICC_EDRAWS, "systematic sampling needs one draw");
  double rn = (double)d[0] / 2147483647.0;
  *consumed = 1;
  if (nsamp <= 0) return;
  FRICC_CUDA(cudaMemsetAsync(S_dev, 0, (size_t)nsamp * sizeof(int64_t), stream()));  // std::vector<size_t> S(n_sample)
  if (n <= 0) return;
  int h = 0;
  if (use_parallel_scan()) {
    systematic_parallel(ps, n, nsamp, rn, S_dev, &h);
  } else {
    DevBuf<int> err(1);
    err.zero();
    FRICC_LAUNCH(systematic_chain_kernel, 1, CH_THREADS, 0, ps, n, nsamp, rn, S_dev, err.p);
    err.to_host(&h, 1);
  }
  raise_flags(h);
}

static void run_pivotal(const PSrc& ps, int64_t n, int64_t nsamp, const DrawSource& draws, int64_t* consumed,
                        int64_t* S_dev) {
  if (n < 1) fail(FRICC_EINVAL, "pivotal sampling needs a non-empty probability vector");
  // one draw per visited element (fri_utils.cpp:74).  Library stream: the next n-1 values of rand() are generated on
  // the device (jump-ahead, fricc_rand.cu) and the host stream is advanced afterwards by exactly the number consumed.
  DevBuf<int32_t> dd;
  int64_t ndraws = 0;
  if (draws.from_stream) {
    ndraws = n - 1;
    dd.alloc((size_t)std::max<int64_t>(ndraws, 1));
    rand_fill_device(dd.p, ndraws);
  } else {
    ndraws = std::min<int64_t>(draws.n, n - 1);
    dd.alloc((size_t)std::max<int64_t>(ndraws, 1));
    if (ndraws > 0) dd.from_host(draws.host, (size_t)ndraws);
  }
  DevBuf<int> err(1);
  err.zero();
  DevBuf<PivotalState> st(1);
  DevBuf<int64_t> dummy;
  int64_t* Sp = S_dev;
  if (nsamp <= 0) {
    dummy.alloc(1);
    Sp = dummy.p;
  }
  bool done = false;
  g_pivotal_restarts = 0;
  if (g_pivotal_parallel == 1 || (g_pivotal_parallel < 0 && n >= (1 << 16))) {
    done = pivotal_parallel(ps, n, nsamp, dd.p, ndraws, Sp, st.p, err.p, &g_pivotal_restarts);
    if (!done) err.zero();  // (flags raised by the abandoned attempt are raised again by the chain kernel)
  }
  if (!done)
    FRICC_LAUNCH(pivotal_chain_kernel, 1, PV_THREADS, 0, ps, n, nsamp, dd.p, ndraws, Sp, st.p, err.p);
  int h = 0;
  err.to_host(&h, 1);
  PivotalState hs{};
  st.to_host(&hs, 1);
  *consumed = hs.visited;
  if (draws.from_stream && hs.visited > 0) rand_skip(hs.visited);  // leave the stream where the reference's rand() would be
  // a short caller buffer only matters if the sampler actually walked past its end
  if ((h & ERRF_DRAWS) && hs.visited <= (long long)ndraws) h &= ~ERRF_DRAWS;
  raise_flags(h);
  if (hs.count != nsamp)
    fail(FRICC_ECOUNT,
         "ERROR IN PIVOTAL SAMPLING: wrong number of elements (got %lld, want %lld, a-1=%g) (reference: fri_utils.cpp:110-124)",
         hs.count, (long long)nsamp, fabs(hs.a - 1.0));
}

void sample_systematic_probs(int64_t n_sample, const double* p, int64_t n, const DrawSource& draws, int64_t* consumed,
                             int64_t* S) {
  PSrc ps{p, nullptr, 0, 0, ~0ULL, 0, 0};
  run_systematic(ps, n, n_sample, draws, consumed, S);
}
void sample_pivotal_probs(int64_t n_sample, const double* p, int64_t n, const DrawSource& draws, int64_t* consumed,
                          int64_t* S) {
  PSrc ps{p, nullptr, 0, 0, ~0ULL, 0, 0};
  run_pivotal(ps, n, n_sample, draws, consumed, S);
}

void compress_largest(const double* x, int64_t n, int64_t m, Compressed& out) {
  TopM top;
  select_top_m(x, n, m, top);
  out.m = m;
  out.nD = m;
  out.consumed = 0;
  out.flat.alloc((size_t)m);
  out.vals.alloc((size_t)m);
  if (m > 0) FRICC_LAUNCH(gather_largest_kernel, blocks_for(m, 256), 256, 0, x, top.idx.p, m, out.flat.p, out.vals.p);
}

static double now_s() {
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

void compress_fri(const double* x, int64_t n, int64_t m, Sampling method, const DrawSource& draws, bool verbose,
                  Compressed& out) {
  if (m < 0 || m > n) fail(FRICC_EINVAL, "fri_compression: need 0 <= m <= N (m=%lld, N=%lld)", (long long)m, (long long)n);
  double t0 = verbose ? now_s() : 0;
  // ---- get_d_largest: norm, candidates, greedy
  double norm = one_norm_sequential(x, n);
  // D is a PREFIX of the descending-|x| order (the greedy test stops at its first failure, fri_utils.cpp:415-431) and
  // nothing behind that prefix is ever used, so only the top `cand` magnitudes are selected and ordered; cand grows
  // geometrically until the greedy test fails inside the candidate list (or the list is all m entries).
  TopM top;
  int64_t nD = 0;
  double r = norm;
  bool estimated = false;
  for (int64_t cand = std::min<int64_t>(m, g_first_candidates); m > 0;) {
    select_top_m(x, n, cand, top);
    d_greedy(top, m, norm, &nD, &r);
    if (nD < cand || cand == m) break;
    // the list was exhausted: size the next one from a count+sum histogram estimate of where the greedy test stops
    // (plus slack), at least x16
    int64_t next = (cand > m / 16) ? m : cand * 16;
    if (g_d_estimate && !estimated) {
      int64_t est = estimate_d_candidates(x, n, m, norm);
      next = std::max<int64_t>(next, est + est / 8 + 1024);
      estimated = true;
    }
    cand = std::min<int64_t>(m, next);
  }
  double t1 = verbose ? now_s() : 0;

  // ---- probability source
  uint64_t cut[2] = {~0ULL, 0};
  if (nD > 0) {
    FRICC_CUDA(cudaMemcpyAsync(&cut[0], top.keys.p + (nD - 1), 8, cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaMemcpyAsync(&cut[1], top.idx.p + (nD - 1), 8, cudaMemcpyDeviceToHost, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
  const int64_t nsamp = m - nD;
  PSrc ps{nullptr, x, r, (double)nsamp, cut[0], cut[1], 0};
  double t2 = verbose ? now_s() : 0;

  // ---- sample
  DevBuf<int64_t> S((size_t)std::max<int64_t>(nsamp, 1));
  int64_t consumed = 0;
  if (method == Sampling::Systematic) run_systematic(ps, n, nsamp, draws, &consumed, S.p);
  else if (method == Sampling::Pivotal) run_pivotal(ps, n, nsamp, draws, &consumed, S.p);
  else FRICC_CUDA(cudaMemsetAsync(S.p, 0, (size_t)std::max<int64_t>(nsamp, 1) * 8, stream()));
  double t3 = verbose ? now_s() : 0;

  // ---- wrap up: indices [D..., S...] and values
  out.m = m;
  out.nD = nD;
  out.remaining_norm = r;
  out.consumed = consumed;
  out.flat.alloc((size_t)m);
  out.vals.alloc((size_t)m);
  if (m > 0)
    FRICC_LAUNCH(gather_values_kernel, blocks_for(m, 256), 256, 0, ps, top.idx.p, nD, S.p, m, out.flat.p, out.vals.p);
  if (verbose) {
    double t4 = now_s();
    double tot = t4 - t0;
    // same 5-row table as fri_utils.cpp:509-519 so logs stay comparable
    printf("Subsection    Fraction of Total    Total Time\n");
    printf("==========    =================    ==========\n");
    printf("Sorting       %6.4f               %6.4f (s)\n", (t1 - t0) / tot, t1 - t0);
    printf("P vector      %6.4f               %6.4f (s)\n", (t2 - t1) / tot, t2 - t1);
    printf("Sampling      %6.4f               %6.4f (s)\n", (t3 - t2) / tot, t3 - t2);
    printf("Wrap up       %6.4f               %6.4f (s)\n", (t4 - t3) / tot, t4 - t3);
    printf("Total         %6.4f               %6.4f (s)\n", 1.0, tot);
    fflush(stdout);
  }
}

}  // namespace fricc
