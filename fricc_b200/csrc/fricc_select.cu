// fricc_select.cu -- deterministic top-m selection by magnitude: radix-histogram select + candidate sort.
//
// Replaces the reference's argsort (/root/reference/src/fri_utils.cpp:11-36), which builds N (value,index) pairs and
// merge-sorts ALL of them by |value| descending only to keep the first m.  Here: the key of an element is the bit
// pattern of |x| (monotone in |x|), the m-th largest key is located by 12-bit histogram passes over the 63 magnitude
// bits (each pass is one streaming read of x), the <= m survivors are compacted, and only they are sorted.
// Tie order: the reference inherits libstdc++'s introsort/inplace_merge order inside runs of equal |x|; this library
// defines the canonical order (|x| descending, flat index ascending) -- see DESIGN.md "tie order".
#include "fricc_select.cuh"

#include <algorithm>

#include "fricc_sort.cuh"

namespace fricc {

constexpr int SEL_BITS = 12;
constexpr int SEL_BINS = 1 << SEL_BITS;
constexpr int SEL_THREADS = 512;

// histogram of the SEL_BITS-bit digit at `shift` over elements whose key bits above (shift+SEL_BITS) equal `prefix`.
// prefix_bits == 0 disables the filter.  (For the final 3-bit digit, shift=0 and the digit is masked to `mask`.)
__global__ void __launch_bounds__(SEL_THREADS)
key_hist_kernel(const double* __restrict__ x, int64_t n, int shift, unsigned mask, uint64_t prefix, int prefix_shift,
                int use_prefix, unsigned long long* __restrict__ hist) {
  __shared__ unsigned sh[SEL_BINS];
  for (int i = threadIdx.x; i < SEL_BINS; i += SEL_THREADS) sh[i] = 0;
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * SEL_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * SEL_THREADS + threadIdx.x; i < n; i += stride) {
    uint64_t key = abs_key(x[i]);
    bool ok = !use_prefix || ((key >> prefix_shift) == prefix);
    if (ok) atomicAdd(&sh[(unsigned)(key >> shift) & mask], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SEL_BINS; i += SEL_THREADS)
    if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// counts AND magnitude sums per 12-bit digit (same filter as key_hist_kernel); used only to ESTIMATE where the greedy
// D test stops (the estimate sizes the candidate list, it never decides anything -- see estimate_d_candidates).
__global__ void __launch_bounds__(SEL_THREADS)
key_hist_sum_kernel(const double* __restrict__ x, int64_t n, int shift, uint64_t prefix, int prefix_shift, int use_prefix,
                    unsigned long long* __restrict__ hist, double* __restrict__ sums) {
  __shared__ unsigned sh[SEL_BINS];
  __shared__ double ss[SEL_BINS];
  for (int i = threadIdx.x; i < SEL_BINS; i += SEL_THREADS) {
    sh[i] = 0;
    ss[i] = 0.0;
  }
  __syncthreads();
  const int64_t stride = (int64_t)gridDim.x * SEL_THREADS;
  for (int64_t i = (int64_t)blockIdx.x * SEL_THREADS + threadIdx.x; i < n; i += stride) {
    uint64_t key = abs_key(x[i]);
    bool ok = !use_prefix || ((key >> prefix_shift) == prefix);
    if (ok) {
      unsigned b = (unsigned)(key >> shift) & (SEL_BINS - 1);
      atomicAdd(&sh[b], 1u);
      atomicAdd(&ss[b], key_to_double(key));
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < SEL_BINS; i += SEL_THREADS)
    if (sh[i]) {
      atomicAdd(&hist[i], (unsigned long long)sh[i]);
      atomicAdd(&sums[i], ss[i]);
    }
}

// Appends (key, index) of every element with (key >> cshift) > cut to `above`, and of every element with
// (key >> cshift) == cut to `amb` (skipped when amb_k == nullptr).  One CTA handles tiles of SEL_THREADS*CP_ITEMS
// elements: classify in registers, CTA-wide exclusive scan of the two class counts, ONE atomicAdd pair per tile
// (a per-warp atomic on a single counter serialises: 4.5 ms at N=2.56e8, see profiles/), coalesced-ish appends.
// The order of the appended candidates is arbitrary (they are sorted afterwards, which makes the result deterministic).
constexpr int CP_ITEMS = 8;
__global__ void __launch_bounds__(SEL_THREADS)
compact_kernel(const double* __restrict__ x, int64_t n, int cshift, uint64_t cut, uint64_t* __restrict__ above_k,
               uint64_t* __restrict__ above_i, unsigned long long* __restrict__ above_n, uint64_t* __restrict__ amb_k,
               uint64_t* __restrict__ amb_i, unsigned long long* __restrict__ amb_n, unsigned long long above_cap,
               unsigned long long amb_cap) {
  __shared__ unsigned s_wa[SEL_THREADS / 32], s_wb[SEL_THREADS / 32];
  __shared__ unsigned long long s_base[2];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t tile_elems = (int64_t)SEL_THREADS * CP_ITEMS;
  const int64_t ntiles = (n + tile_elems - 1) / tile_elems;
  const bool want_amb = amb_k != nullptr;
  for (int64_t tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int64_t base = tile * tile_elems + threadIdx.x;
    uint64_t key[CP_ITEMS];
    unsigned clsbits = 0;  // 2 bits per item: 1 = above, 2 = ambiguous
    unsigned ca = 0, cb = 0;
#pragma unroll
    for (int q = 0; q < CP_ITEMS; q++) {
      int64_t i = base + (int64_t)q * SEL_THREADS;
      key[q] = 0;
      if (i < n) {
        key[q] = abs_key(x[i]);
        uint64_t kp = key[q] >> cshift;
        unsigned cls = kp > cut ? 1u : ((kp == cut && want_amb) ? 2u : 0u);
        clsbits |= cls << (2 * q);
        ca += cls == 1u;
        cb += cls == 2u;
      }
    }
    // inclusive warp scans of (ca, cb)
    unsigned ia = ca, ib = cb;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      unsigned ta = __shfl_up_sync(0xffffffffu, ia, o), tb = __shfl_up_sync(0xffffffffu, ib, o);
      if (lane >= (unsigned)o) { ia += ta; ib += tb; }
    }
    if (lane == 31) { s_wa[warp] = ia; s_wb[warp] = ib; }
    __syncthreads();
    unsigned offa = 0, offb = 0, tota = 0, totb = 0;
#pragma unroll
    for (unsigned w = 0; w < SEL_THREADS / 32; w++) {
      unsigned va = s_wa[w], vb = s_wb[w];
      if (w < warp) { offa += va; offb += vb; }
      tota += va; totb += vb;
    }
    if (threadIdx.x == 0) {
      s_base[0] = tota ? atomicAdd(above_n, (unsigned long long)tota) : 0ULL;
      s_base[1] = totb ? atomicAdd(amb_n, (unsigned long long)totb) : 0ULL;
    }
    __syncthreads();
    unsigned long long pa = s_base[0] + offa + (ia - ca), pb = s_base[1] + offb + (ib - cb);
#pragma unroll
    for (int q = 0; q < CP_ITEMS; q++) {
      unsigned cls = (clsbits >> (2 * q)) & 3u;
      int64_t i = base + (int64_t)q * SEL_THREADS;
      if (cls == 1u) {
        if (pa < above_cap) { above_k[pa] = key[q]; above_i[pa] = (uint64_t)i; }
        pa++;
      } else if (cls == 2u) {
        if (pb < amb_cap) { amb_k[pb] = key[q]; amb_i[pb] = (uint64_t)i; }
        pb++;
      }
    }
    __syncthreads();  // s_base / s_w* are reused by the next tile
  }
}

// exact ties at the cut: pick the `take` lowest-index elements whose key == cut (canonical tie order).
constexpr int TIE_TILE = SEL_THREADS * 4;
__global__ void __launch_bounds__(SEL_THREADS)
tie_count_kernel(const double* __restrict__ x, int64_t n, uint64_t cutkey, int64_t* __restrict__ counts) {
  __shared__ int sm[SEL_THREADS / 32];
  const int64_t base = (int64_t)blockIdx.x * TIE_TILE + (int64_t)threadIdx.x * 4;
  int c = 0;
#pragma unroll
  for (int i = 0; i < 4; i++)
    if (base + i < n && abs_key(x[base + i]) == cutkey) c++;
  for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < SEL_THREADS / 32; w++) t += sm[w];
    counts[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(SEL_THREADS)
tie_emit_kernel(const double* __restrict__ x, int64_t n, uint64_t cutkey, const int64_t* __restrict__ offsets,
                int64_t take, uint64_t* __restrict__ out_k, uint64_t* __restrict__ out_i) {
  __shared__ int sm[SEL_THREADS / 32];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t base = (int64_t)blockIdx.x * TIE_TILE + (int64_t)threadIdx.x * 4;
  bool f[4];
  int c = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    f[i] = base + i < n && abs_key(x[base + i]) == cutkey;
    c += f[i];
  }
  int inc = c;
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= (unsigned)o) inc += t;
  }
  if (lane == 31) sm[warp] = inc;
  __syncthreads();
  int woff = 0;
  for (unsigned w = 0; w < warp; w++) woff += sm[w];
  int64_t pos = offsets[blockIdx.x] + woff + inc - c;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    if (f[i]) {
      if (pos < take) {
        out_k[pos] = cutkey;
        out_i[pos] = (uint64_t)(base + i);
      }
      pos++;
    }
  }
}

static int compact_grid(int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, SEL_THREADS * CP_ITEMS), (int64_t)num_sms() * 8));
}
static int hist_grid(int64_t n) {
  return (int)std::max<int64_t>(1, std::min<int64_t>(ceil_div(n, SEL_THREADS * 8), (int64_t)num_sms() * 4));
}

void key_histogram(const double* x, int64_t n, int shift, unsigned mask, uint64_t prefix, int prefix_bits,
                   unsigned long long* hist_dev) {
  int prefix_shift = shift + (mask == (SEL_BINS - 1) ? SEL_BITS : 3);
  if (prefix_shift > 63) prefix_shift = 63;
  FRICC_LAUNCH(key_hist_kernel, hist_grid(n), SEL_THREADS, 0, x, n, shift, mask, prefix, prefix_shift,
               prefix_bits > 0 ? 1 : 0, hist_dev);
}

// digit schedule over the 63 magnitude bits: five 12-bit digits (shifts 51,39,27,15,3) and one 3-bit digit (shift 0)
static const int kShifts[6] = {51, 39, 27, 15, 3, 0};
static const unsigned kMasks[6] = {SEL_BINS - 1, SEL_BINS - 1, SEL_BINS - 1, SEL_BINS - 1, SEL_BINS - 1, 7};

void select_top_m(const double* x, int64_t n, int64_t m, TopM& out) {
  if (m < 0 || m > n) fail(FRICC_EINVAL, "select_top_m: need 0 <= m <= N (m=%lld, N=%lld)", (long long)m, (long long)n);
  out.m = m;
  out.keys.alloc((size_t)m);
  out.idx.alloc((size_t)m);
  if (m == 0) return;

  // ---- locate the cut: after the loop, every key with (key >> cshift) > prefix is selected (`above` of them),
  //      and `rem` more must come from the bin (key >> cshift) == prefix, which holds bin_count elements
  const int64_t SMALL_BIN = (int64_t)1 << 21;
  DevBuf<unsigned long long> hist(SEL_BINS);
  std::vector<unsigned long long> h(SEL_BINS);
  uint64_t prefix = 0;
  int prefix_bits = 0, cshift = 63;
  int64_t above = 0, rem = m, bin_count = n;
  for (int pass = 0; pass < 6; pass++) {
    hist.zero();
    key_histogram(x, n, kShifts[pass], kMasks[pass], prefix, prefix_bits, hist.p);
    hist.to_host(h.data(), SEL_BINS);
    int nb = (int)kMasks[pass] + 1;
    int64_t acc = 0;
    int b = nb - 1;
    for (; b >= 0; b--) {
      if (acc + (int64_t)h[b] >= rem) break;
      acc += (int64_t)h[b];
    }
    if (b < 0) fail(FRICC_ECUDA, "select_top_m: histogram inconsistent (pass %d)", pass);
    above += acc;
    rem -= acc;
    bin_count = (int64_t)h[b];
    int digit_bits = (kMasks[pass] == 7) ? 3 : SEL_BITS;
    prefix = (prefix << digit_bits) | (uint64_t)b;
    prefix_bits += digit_bits;
    cshift = kShifts[pass];
    if (bin_count == rem || bin_count <= SMALL_BIN) break;
  }

  // ---- compaction
  uint64_t* kbuf[2];
  uint64_t* ibuf[2];
  DevBuf<uint64_t> k1((size_t)m), i1((size_t)m);
  kbuf[0] = out.keys.p; ibuf[0] = out.idx.p;
  kbuf[1] = k1.p; ibuf[1] = i1.p;
  DevBuf<unsigned long long> cnt(2);
  cnt.zero();

  if (bin_count == rem) {
    // the whole bin is selected: everything with (key >> cshift) >= prefix, i.e. > prefix-1
    // (prefix == 0 with cshift covering all bits means "everything": handle by cut = prefix and taking both classes)
    FRICC_LAUNCH(compact_kernel, compact_grid(n), SEL_THREADS, 0, x, n, cshift, prefix, kbuf[0], ibuf[0], cnt.p,
                 kbuf[0] + above, ibuf[0] + above, cnt.p + 1, (unsigned long long)above, (unsigned long long)rem);
  } else if (cshift == 0 && bin_count > SMALL_BIN) {
    // a large group of exactly equal magnitudes straddles the cut: take the lowest indices (canonical order)
    FRICC_LAUNCH(compact_kernel, compact_grid(n), SEL_THREADS, 0, x, n, cshift, prefix, kbuf[0], ibuf[0], cnt.p,
                 (uint64_t*)nullptr, (uint64_t*)nullptr, cnt.p + 1, (unsigned long long)above, 0ULL);
    int64_t nb = ceil_div(n, TIE_TILE);
    DevBuf<int64_t> counts((size_t)nb);
    FRICC_LAUNCH(tie_count_kernel, (unsigned)nb, SEL_THREADS, 0, x, n, prefix, counts.p);
    exclusive_scan_i64(counts.p, nb);
    FRICC_LAUNCH(tie_emit_kernel, (unsigned)nb, SEL_THREADS, 0, x, n, prefix, counts.p, rem, kbuf[0] + above,
                 ibuf[0] + above);
  } else {
    // small ambiguous bin: collect it, order it by (key desc, index asc), keep the first `rem`
    DevBuf<uint64_t> ak((size_t)bin_count), ai((size_t)bin_count), ak2((size_t)bin_count), ai2((size_t)bin_count);
    FRICC_LAUNCH(compact_kernel, compact_grid(n), SEL_THREADS, 0, x, n, cshift, prefix, kbuf[0], ibuf[0], cnt.p, ak.p,
                 ai.p, cnt.p + 1, (unsigned long long)above, (unsigned long long)bin_count);
    uint64_t* ka[2] = {ak.p, ak2.p};
    uint64_t* ia[2] = {ai.p, ai2.p};
    int r = radix_sort_pairs(ka, ia, bin_count, candidate_digits(n));
    FRICC_CUDA(cudaMemcpyAsync(kbuf[0] + above, ka[r], (size_t)rem * 8, cudaMemcpyDeviceToDevice, stream()));
    FRICC_CUDA(cudaMemcpyAsync(ibuf[0] + above, ia[r], (size_t)rem * 8, cudaMemcpyDeviceToDevice, stream()));
  }

  // ---- order all m candidates: |x| descending, index ascending
  int r = radix_sort_pairs(kbuf, ibuf, m, candidate_digits(n));
  if (r == 1) {
    FRICC_CUDA(cudaMemcpyAsync(out.keys.p, k1.p, (size_t)m * 8, cudaMemcpyDeviceToDevice, stream()));
    FRICC_CUDA(cudaMemcpyAsync(out.idx.p, i1.p, (size_t)m * 8, cudaMemcpyDeviceToDevice, stream()));
  }
}

// Upper estimate of |D|+1 for get_d_largest (fri_utils.cpp:400-437): the greedy test at rank d is
//   (n_sample - d) * x_d >= r_d - x_d   with  r_d = norm - sum_{i<d} x_i .
// It is evaluated at the lower edges of the magnitude bins (counts and sums above the edge from one or two histogram
// passes); the first edge where it fails bounds the prefix from above when the test is monotone, and compress_fri grows
// the list if the bound turns out too small, so this only ever sizes buffers -- the exact greedy test decides D.
int64_t estimate_d_candidates(const double* x, int64_t n, int64_t n_sample, double norm) {
  if (n_sample <= 0 || n <= 0) return 0;
  DevBuf<unsigned long long> hist(SEL_BINS);
  DevBuf<double> sums(SEL_BINS);
  std::vector<unsigned long long> h(SEL_BINS);
  std::vector<double> hs(SEL_BINS);
  uint64_t prefix = 0;
  int prefix_bits = 0;
  int64_t c_above = 0;     // elements strictly above the bin under refinement
  double s_above = 0.0;
  const int64_t budget = std::max<int64_t>(1 << 16, n_sample / 64);
  int64_t est = n_sample;
  for (int pass = 0; pass < 3; pass++) {
    const int shift = kShifts[pass];
    hist.zero();
    sums.zero();
    FRICC_LAUNCH(key_hist_sum_kernel, hist_grid(n), SEL_THREADS, 0, x, n, shift, prefix, shift + SEL_BITS, prefix_bits > 0 ? 1 : 0,
                 hist.p, sums.p);
    hist.to_host(h.data(), SEL_BINS);
    sums.to_host(hs.data(), SEL_BINS);
    int64_t c = c_above;
    double sm = s_above;
    int b = SEL_BINS - 1;
    int64_t c_before = c;
    double s_before = sm;
    bool failed = false;
    for (; b >= 0; b--) {
      c_before = c;
      s_before = sm;
      c += (int64_t)h[b];
      sm += hs[b];
      if (h[b] == 0) continue;
      // lower edge of bin b: every element of the bin is >= t
      const double t = key_to_double(((prefix << SEL_BITS) | (uint64_t)b) << shift);
      const bool holds = c < n_sample && (double)(n_sample - c) * t >= (norm - sm) - t && (norm - sm) > 1e-14;
      if (!holds) {
        failed = true;
        break;
      }
    }
    if (!failed) {  // the test holds down to the smallest magnitude seen: D may be everything
      est = n_sample;
      break;
    }
    est = std::min<int64_t>(n_sample, c + 1);
    if ((int64_t)h[b] <= budget || shift == 0) break;
    // the failing bin is fat: look inside it
    prefix = (prefix << SEL_BITS) | (uint64_t)b;
    prefix_bits += SEL_BITS;
    c_above = c_before;
    s_above = s_before;
  }
  return std::max<int64_t>(1, std::min<int64_t>(est, n_sample));
}

// LSD digit order for (key descending, index ascending): index digits first (least significant), then key digits
std::vector<Digit> candidate_digits(int64_t n) {
  std::vector<Digit> d;
  int idx_bits = 1;
  while (idx_bits < 63 && ((int64_t)1 << idx_bits) < n) idx_bits++;
  for (int s = 0; s < idx_bits; s += 8) d.push_back(Digit{1, s, 0});
  for (int s = 0; s < 63; s += 8) d.push_back(Digit{0, s, 1});
  return d;
}

}  // namespace fricc
