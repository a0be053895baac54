// fricc_sort.cu -- stable LSD radix sort (8-bit digits) of (key, payload) pairs + device-wide exclusive scans.
//
// Used for (a) ordering the top-m candidates by (|x| descending, index ascending) -- the reference gets this order
// from a full merge sort of all N pairs (/root/reference/src/fri_utils.cpp:11-36); only the m selected candidates are
// sorted here -- and (b) bucketing the compressed entries by owner pair for the contraction kernels.
#include "fricc_sort.cuh"

namespace fricc {

// ---------------------------------------------------------------------------------------------------------------
// device-wide exclusive scan: tile sums -> scan of tile sums (one CTA) -> tile-local scan + offset
// ---------------------------------------------------------------------------------------------------------------
constexpr int SCAN_THREADS = 1024;
constexpr int SCAN_IPT = 4;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_IPT;

template <typename T>
__device__ __forceinline__ T warp_incl_scan(T v) {
  const unsigned lane = threadIdx.x & 31;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= (unsigned)o) v += t;
  }
  return v;
}

// inclusive scan across the block of one value per thread; returns the inclusive value, *total gets the block sum
template <typename T>
__device__ __forceinline__ T block_incl_scan(T v, T* smem /*[32]*/, T* total) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned nwarps = (blockDim.x + 31) >> 5;
  T inc = warp_incl_scan(v);
  if (lane == 31) smem[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    T s = (lane < nwarps) ? smem[lane] : T(0);
    s = warp_incl_scan(s);
    smem[lane] = s;
  }
  __syncthreads();
  T offset = warp ? smem[warp - 1] : T(0);
  *total = smem[nwarps - 1];
  __syncthreads();
  return inc + offset;
}

template <typename T>
__global__ void scan_tile_sums(const T* __restrict__ d, int64_t n, T* __restrict__ sums) {
  __shared__ T sm[32];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  T s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_IPT; i++) {
    int64_t idx = base + (int64_t)i * SCAN_THREADS + threadIdx.x;
    if (idx < n) s += d[idx];
  }
  T tot;
  block_incl_scan(s, sm, &tot);
  if (threadIdx.x == 0) sums[blockIdx.x] = tot;
}

template <typename T>
__global__ void scan_single_block(T* __restrict__ d, int64_t n) {
  __shared__ T sm[32];
  T carry = 0;
  for (int64_t base = 0; base < n; base += SCAN_THREADS) {
    int64_t idx = base + threadIdx.x;
    T v = (idx < n) ? d[idx] : T(0);
    T tot;
    T inc = block_incl_scan(v, sm, &tot);
    if (idx < n) d[idx] = carry + inc - v;
    carry += tot;
  }
}

template <typename T>
__global__ void scan_tiles(T* __restrict__ d, int64_t n, const T* __restrict__ tile_offsets) {
  __shared__ T sm[32];
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_IPT;
  T v[SCAN_IPT];
  T s = 0;
#pragma unroll
  for (int i = 0; i < SCAN_IPT; i++) {
    v[i] = (base + i < n) ? d[base + i] : T(0);
    s += v[i];
  }
  T tot;
  T inc = block_incl_scan(s, sm, &tot);
  T run = tile_offsets[blockIdx.x] + inc - s;
#pragma unroll
  for (int i = 0; i < SCAN_IPT; i++) {
    if (base + i < n) d[base + i] = run;
    run += v[i];
  }
}

template <typename T>
static void exclusive_scan_impl(T* d, int64_t n) {
  if (n <= 0) return;
  if (n <= 8 * SCAN_THREADS) {
    FRICC_LAUNCH(scan_single_block<T>, 1, SCAN_THREADS, 0, d, n);
    return;
  }
  int64_t ntiles = ceil_div(n, SCAN_TILE);
  DevBuf<T> sums((size_t)ntiles);
  FRICC_LAUNCH(scan_tile_sums<T>, (unsigned)ntiles, SCAN_THREADS, 0, d, n, sums.p);
  exclusive_scan_impl<T>(sums.p, ntiles);
  FRICC_LAUNCH(scan_tiles<T>, (unsigned)ntiles, SCAN_THREADS, 0, d, n, sums.p);
}

void exclusive_scan_u32(uint32_t* d, int64_t n) { exclusive_scan_impl<uint32_t>(d, n); }
void exclusive_scan_i64(int64_t* d, int64_t n) { exclusive_scan_impl<int64_t>(d, n); }
void exclusive_scan_f64(double* d, int64_t n) { exclusive_scan_impl<double>(d, n); }

// ---------------------------------------------------------------------------------------------------------------
// radix sort
// ---------------------------------------------------------------------------------------------------------------
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_IPT = 8;
constexpr int RS_TILE = RS_THREADS * RS_IPT;  // 2048 items per CTA
constexpr int RS_MAX_DIGITS = 16;

struct DigitList {
  int n;
  Digit d[RS_MAX_DIGITS];
};

__device__ __forceinline__ unsigned extract_digit(uint64_t key, uint64_t val, const Digit& dg) {
  uint64_t s = dg.src ? val : key;
  unsigned d = (unsigned)(s >> dg.shift) & 255u;
  return dg.invert ? (d ^ 255u) : d;
}

// global histograms of every digit position in one read of the data (decides which passes can be skipped)
__global__ void rs_global_hist(const uint64_t* __restrict__ k, const uint64_t* __restrict__ v, int64_t n, DigitList dl,
                               unsigned long long* __restrict__ hist /*[ndigits][256]*/) {
  extern __shared__ unsigned sh[];  // [ndigits][256]
  for (int i = threadIdx.x; i < dl.n * 256; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    uint64_t kk = k[i], vv = v[i];
    for (int g = 0; g < dl.n; g++) atomicAdd(&sh[g * 256 + extract_digit(kk, vv, dl.d[g])], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < dl.n * 256; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

// per-CTA digit counts, digit-major: counts[digit * nblocks + block]
__global__ void rs_block_hist(const uint64_t* __restrict__ k, const uint64_t* __restrict__ v, int64_t n, Digit dg,
                              uint32_t* __restrict__ counts, int nblocks) {
  __shared__ unsigned sh[256];
  sh[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = (int64_t)blockIdx.x * RS_TILE;
#pragma unroll
  for (int r = 0; r < RS_IPT; r++) {
    int64_t idx = base + (int64_t)r * RS_THREADS + threadIdx.x;
    if (idx < n) atomicAdd(&sh[extract_digit(k[idx], v[idx], dg)], 1u);
  }
  __syncthreads();
  counts[(int64_t)threadIdx.x * nblocks + blockIdx.x] = sh[threadIdx.x];
}

// stable scatter: item order inside a CTA is (warp, round, lane); ranks are built warp by warp with match_any
__global__ void __launch_bounds__(RS_THREADS)
rs_scatter(const uint64_t* __restrict__ kin, const uint64_t* __restrict__ vin, uint64_t* __restrict__ kout,
           uint64_t* __restrict__ vout, int64_t n, Digit dg, const uint32_t* __restrict__ offsets, int nblocks) {
  __shared__ unsigned wcount[RS_WARPS][257];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < RS_WARPS * 257; i += RS_THREADS) (&wcount[0][0])[i] = 0;
  __syncthreads();

  const int64_t wbase = (int64_t)blockIdx.x * RS_TILE + (int64_t)warp * (32 * RS_IPT);
  uint64_t key[RS_IPT], val[RS_IPT];
  unsigned dig[RS_IPT], rank[RS_IPT];
  const unsigned lt = (1u << lane) - 1u;
#pragma unroll
  for (int r = 0; r < RS_IPT; r++) {
    int64_t idx = wbase + r * 32 + lane;
    bool valid = idx < n;
    key[r] = valid ? kin[idx] : 0;
    val[r] = valid ? vin[idx] : 0;
    unsigned d = valid ? extract_digit(key[r], val[r], dg) : 256u;
    dig[r] = d;
    unsigned peers = __match_any_sync(0xffffffffu, d);
    unsigned leader = __ffs(peers) - 1;
    unsigned old = 0;
    if (lane == leader) {
      old = wcount[warp][d];
      wcount[warp][d] = old + __popc(peers);
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[r] = old + __popc(peers & lt);
    __syncwarp();
  }
  __syncthreads();
  {  // one thread per digit: exclusive prefix over the warps, seeded with this CTA's global offset for the digit
    unsigned d = threadIdx.x;
    unsigned run = offsets[(int64_t)d * nblocks + blockIdx.x];
#pragma unroll
    for (int w = 0; w < RS_WARPS; w++) {
      unsigned c = wcount[w][d];
      wcount[w][d] = run;
      run += c;
    }
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < RS_IPT; r++) {
    if (dig[r] < 256u) {
      unsigned pos = wcount[warp][dig[r]] + rank[r];
      kout[pos] = key[r];
      vout[pos] = val[r];
    }
  }
}

int radix_sort_pairs(uint64_t* k[2], uint64_t* v[2], int64_t n, const std::vector<Digit>& digits) {
  if (n <= 1 || digits.empty()) return 0;
  if (n >= (int64_t)0xffffffffLL) fail(FRICC_EINVAL, "radix_sort_pairs: n=%lld too large", (long long)n);
  if ((int)digits.size() > RS_MAX_DIGITS) fail(FRICC_EINVAL, "radix_sort_pairs: too many digits");
  DigitList dl;
  dl.n = (int)digits.size();
  for (int i = 0; i < dl.n; i++) dl.d[i] = digits[i];

  // Every digit gets its pass.  (An earlier version read a global histogram back to skip digits on which all keys
  // agree: one blocking device-to-host copy per sort -- seven per iteration on the contraction path, each a host
  // synchronisation, and each queued behind whatever else the device-to-host engine was doing.  The callers' digit
  // lists are already cut to the key range, so a skippable digit is rare; the host now never waits here.)
  const int nblocks = (int)ceil_div(n, RS_TILE);
  DevBuf<uint32_t> counts((size_t)256 * nblocks);
  int cur = 0;
  for (int g = 0; g < dl.n; g++) {
    FRICC_LAUNCH(rs_block_hist, nblocks, RS_THREADS, 0, k[cur], v[cur], n, dl.d[g], counts.p, nblocks);
    exclusive_scan_u32(counts.p, (int64_t)256 * nblocks);
    FRICC_LAUNCH(rs_scatter, nblocks, RS_THREADS, 0, k[cur], v[cur], k[cur ^ 1], v[cur ^ 1], n, dl.d[g], counts.p,
                 nblocks);
    cur ^= 1;
  }
  return cur;
}

}  // namespace fricc
