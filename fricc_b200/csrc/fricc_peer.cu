// fricc_peer.cu -- the exchange step of the sharded contractions over NVLink peer memory (no NCCL on the data path).
//
// Every rank owns the L-columns [l_begin, l_begin + l_count) of each contraction output and every rank needs the full
// output afterwards (the next contraction reads it as W, the next compression reads all of t2new).  Instead of
// all-gathering column panels and re-adding them into the 4-D array (one NCCL launch + a 6 GB read-modify-write pass per
// output), each rank accumulates its columns straight into ITS copy of the 4-D output and then pushes exactly those
// elements into the same addresses of every peer's copy with plain stores over NVLink / NVSwitch: the copies live in
// one symmetric allocation per rank (cudaMalloc + CUDA IPC), so "the same element on peer p" is base_p + offset.
// A column is written by its owner only, so no two ranks ever store to the same address; consumers are ordered behind
// the writers by a stream-ordered barrier (fricc_b200/dist.py).
#include <algorithm>

#include "fricc_contract.cuh"

namespace fricc {

constexpr int MAX_PEERS = 16;

struct PeerPtrs {
  double* p[MAX_PEERS];
};

// thread <-> one element of this rank's column slice; FAST = which index runs along threadIdx.x:
//   0: l (the free pair; contiguous when out's unit-stride axis is the L-minor one)      1: R (owner pair)
// The grid is CAPPED (push_grid below) and every CTA walks "virtual blocks": the stores are NVLink-bound, a couple of
// CTAs per SM saturate the links, and the kernel then leaves thread slots, registers and shared memory free for the
// tensor-core kernel of the NEXT term, which runs concurrently on the compute stream (dist.SideStream).
template <int FAST>
__global__ void __launch_bounds__(256)
push_columns_kernel(const double* __restrict__ src, PeerPtrs peers, int n_peers, int self, OutMap om, int64_t nR,
                    int64_t l_begin, int64_t l_count, int64_t nbx, int64_t nvb) {
  for (int64_t vb = blockIdx.x; vb < nvb; vb += gridDim.x) {
    const int64_t by = vb / nbx, bx = vb - by * nbx;
    int64_t R, c;
    if (FAST == 0) {
      c = bx * 256 + threadIdx.x;
      R = by;
    } else {
      R = bx * 256 + threadIdx.x;
      c = by;
    }
    if (R >= nR || c >= l_count) continue;
    const int64_t l = l_begin + c;
    const int64_t r1 = R / om.nR2, r2 = R - r1 * om.nR2;
    const int64_t l1 = l / om.nL2, l2 = l - l1 * om.nL2;
    const int64_t o = r1 * om.sR1 + r2 * om.sR2 + l1 * om.sL1 + l2 * om.sL2;
    const double v = src[o];
    for (int d = 1; d < n_peers; d++) {
      int q = self + d;
      if (q >= n_peers) q -= n_peers;
      peers.p[q][o] = v;
    }
  }
}

static int g_push_ctas_per_sm = 2;  // 2 x 256 threads x 32 registers fit beside the tensor-core CTA (320 x 128); 3 do not
static int g_push_carveout = 100;  // percent of the L1/shared array configured as shared memory; -1 = leave the default
void set_push_ctas_per_sm(int n) { g_push_ctas_per_sm = n < 1 ? 1 : n; }
void set_push_carveout(int pct) { g_push_carveout = pct; }
static int g_push_copy_engine = 0;          // "push_copy_engine": 1 = views of long contiguous rows go through cudaMemcpy(2D)Async
static int64_t g_push_min_row_bytes = 4096;  // shorter rows: the store kernel
static int64_t g_push_copies = 0;
void set_push_copy_engine(int on) { g_push_copy_engine = on < 0 ? 0 : on; }
void set_push_min_row_bytes(int n) { g_push_min_row_bytes = n < 8 ? 8 : n; }
int64_t push_copies() { return g_push_copies; }
// An SM runs CTAs of two kernels side by side only under ONE L1/shared split.  The tensor-core kernel needs ~198 KB of
// shared memory, i.e. the maximum-shared split; these kernels ask for the same split so that they can share the SMs
// with it instead of waiting for them to drain.
template <typename K>
static void prefer_split(K kernel) {
  if (g_push_carveout >= 0)
    FRICC_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, g_push_carveout));
}
static unsigned push_grid(int64_t nvb) {
  return (unsigned)std::max<int64_t>(1, std::min<int64_t>(nvb, (int64_t)num_sms() * g_push_ctas_per_sm));
}

void push_columns(const double* out_local, double* const* peer_out, int n_peers, int self, const OutMap& om, int64_t nR,
                  int64_t l_begin, int64_t l_count) {
  if (n_peers <= 1 || nR == 0 || l_count == 0) return;
  if (n_peers > MAX_PEERS) fail(FRICC_EINVAL, "at most %d peers", MAX_PEERS);
  PeerPtrs pp;
  for (int q = 0; q < MAX_PEERS; q++) pp.p[q] = q < n_peers ? peer_out[q] : nullptr;
  const bool r_fast = om.sL2 != 1 && (om.sR2 == 1 || (om.nR2 == 1 && om.sR1 == 1));
  if (!r_fast) {
    const int64_t nbx = ceil_div(l_count, 256), nvb = nbx * nR;
    prefer_split(push_columns_kernel<0>);
    FRICC_LAUNCH(push_columns_kernel<0>, push_grid(nvb), 256, 0, out_local, pp, n_peers, self, om, nR, l_begin, l_count,
                 nbx, nvb);
  } else {
    const int64_t nbx = ceil_div(nR, 256), nvb = nbx * l_count;
    prefer_split(push_columns_kernel<1>);
    FRICC_LAUNCH(push_columns_kernel<1>, push_grid(nvb), 256, 0, out_local, pp, n_peers, self, om, nR, l_begin, l_count,
                 nbx, nvb);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// a 4-D strided VIEW of this rank's copy -> the same view in every peer's copy
// ---------------------------------------------------------------------------------------------------------------
// The sharded engine (fricc_b200/engine.py) hands the plugin call an L-slice of the output -- a 4-D view with the
// output's own strides -- and then publishes exactly that view.  Axes are walked in order of ascending stride, the
// smallest-stride axis along the threads, so the loads and the peer stores are as contiguous as the view allows.
struct View4 {
  int64_t n[4];  // extents, n[0] along the smallest stride
  int64_t s[4];  // element strides, ascending
};

__global__ void __launch_bounds__(256)
push_view_kernel(const double* __restrict__ src, PeerPtrs peers, int n_peers, int self, View4 v, int64_t nbx, int64_t nvb) {
  for (int64_t vb = blockIdx.x; vb < nvb; vb += gridDim.x) {
    const int64_t by = vb / nbx, bx = vb - by * nbx;
    const int64_t i0 = bx * 256 + threadIdx.x;
    if (i0 >= v.n[0]) continue;
    int64_t r = by;
    const int64_t i1 = r % v.n[1]; r /= v.n[1];
    const int64_t i2 = r % v.n[2]; r /= v.n[2];
    const int64_t o = i0 * v.s[0] + i1 * v.s[1] + i2 * v.s[2] + r * v.s[3];
    const double x = src[o];
    for (int d = 1; d < n_peers; d++) {
      int q = self + d;
      if (q >= n_peers) q -= n_peers;
      peers.p[q][o] = x;
    }
  }
}

void push_view(const double* local, double* const* peer, int n_peers, int self, const int64_t shape[4],
               const int64_t strides[4]) {
  if (n_peers <= 1) return;
  if (n_peers > MAX_PEERS) fail(FRICC_EINVAL, "at most %d peers", MAX_PEERS);
  int order[4] = {0, 1, 2, 3};
  for (int i = 0; i < 4; i++)  // extent-1 axes last, the rest by ascending stride
    for (int j = i + 1; j < 4; j++) {
      const bool ui = shape[order[i]] == 1, uj = shape[order[j]] == 1;
      if ((ui && !uj) || (ui == uj && strides[order[j]] < strides[order[i]])) std::swap(order[i], order[j]);
    }
  View4 v;
  int64_t total = 1;
  for (int i = 0; i < 4; i++) {
    v.n[i] = shape[order[i]];
    v.s[i] = strides[order[i]];
    if (v.s[i] < 0) fail(FRICC_EINVAL, "negative strides are not supported");
    total *= v.n[i];
  }
  if (total == 0) return;
  // merge neighbouring axes that are contiguous with each other (longer runs per warp / per copy row)
  int nd = 4;
  while (nd > 1 && v.n[nd - 1] == 1) nd--;
  for (int i = 0; i + 1 < nd;) {
    if (v.s[i + 1] == v.n[i] * v.s[i]) {
      v.n[i] *= v.n[i + 1];
      for (int j = i + 1; j + 1 < nd; j++) { v.n[j] = v.n[j + 1]; v.s[j] = v.s[j + 1]; }
      nd--;
      v.n[nd] = 1; v.s[nd] = 0;
    } else {
      i++;
    }
  }
  // Optional ("push_copy_engine" = 1): views made of long contiguous rows through the COPY ENGINES, one (2-D) device-to-
  // device copy per peer on the library's stream.  They need no SM, so the terms they run under are not slowed down
  // (8 GPUs, headline shape: the four terms before the first barrier 10.2 ms instead of 15.8) -- but the copies
  // themselves drain late (barrier waits 4.3 + 5-10 ms instead of 0.35 + 2.1) and, with the host running far ahead,
  // the terms behind the first barrier stalled for tens of ms in the measurement; the step as a whole was slower
  // (0.080 s against 0.071), so the store kernel stays the default.
  if (g_push_copy_engine && v.s[0] == 1 && v.n[0] * 8 >= g_push_min_row_bytes && (nd <= 2 || (nd == 3 && v.n[2] <= 64))) {
    const int64_t outer = nd == 3 ? v.n[2] : 1;
    // "push_copy_engine" = 2: one private stream per peer, forked from and joined back into the library's stream, so
    // that the copies to different peers run on different engines at the same time (one stream serialises them).
    const bool fan = g_push_copy_engine >= 2;
    static cudaStream_t fan_stream[MAX_PEERS] = {};
    static cudaEvent_t fan_fork = nullptr, fan_done[MAX_PEERS] = {};
    if (fan) {
      if (!fan_fork) FRICC_CUDA(cudaEventCreateWithFlags(&fan_fork, cudaEventDisableTiming));
      FRICC_CUDA(cudaEventRecord(fan_fork, stream()));
    }
    for (int d = 1; d < n_peers; d++) {
      const int q = (self + d) % n_peers;
      cudaStream_t st = stream();
      if (fan) {
        if (!fan_stream[q]) {
          FRICC_CUDA(cudaStreamCreateWithFlags(&fan_stream[q], cudaStreamNonBlocking));
          FRICC_CUDA(cudaEventCreateWithFlags(&fan_done[q], cudaEventDisableTiming));
        }
        st = fan_stream[q];
        FRICC_CUDA(cudaStreamWaitEvent(st, fan_fork, 0));
      }
      for (int64_t k = 0; k < outer; k++) {
        const int64_t off = nd == 3 ? k * v.s[2] : 0;
        if (nd == 1)
          FRICC_CUDA(cudaMemcpyAsync(peer[q] + off, local + off, (size_t)v.n[0] * 8, cudaMemcpyDefault, st));
        else
          FRICC_CUDA(cudaMemcpy2DAsync(peer[q] + off, (size_t)v.s[1] * 8, local + off, (size_t)v.s[1] * 8, (size_t)v.n[0] * 8,
                                       (size_t)v.n[1], cudaMemcpyDefault, st));
      }
      if (fan) {
        FRICC_CUDA(cudaEventRecord(fan_done[q], st));
        FRICC_CUDA(cudaStreamWaitEvent(stream(), fan_done[q], 0));
      }
    }
    g_push_copies += (n_peers - 1) * outer;
    return;
  }
  PeerPtrs pp;
  for (int q = 0; q < MAX_PEERS; q++) pp.p[q] = q < n_peers ? peer[q] : nullptr;
  const int64_t nbx = ceil_div(v.n[0], 256), nvb = nbx * v.n[1] * v.n[2] * v.n[3];
  prefer_split(push_view_kernel);
  FRICC_LAUNCH(push_view_kernel, push_grid(nvb), 256, 0, local, pp, n_peers, self, v, nbx, nvb);
}

// ---------------------------------------------------------------------------------------------------------------
// stream-ordered barrier across the ranks of one node, through the symmetric allocation (no NCCL launch)
// ---------------------------------------------------------------------------------------------------------------
// flags_q = MAX_PEERS 64-bit epoch counters in rank q's arena.  Barrier number e: rank r stores e into flags_q[r] of every
// peer q (release, system scope: everything this rank's earlier kernels stored into q's memory is ordered before it),
// then waits until its own flags[q] >= e for every q (acquire).  The kernel sits on the rank's stream, so work queued
// behind it starts only when every rank has reached the barrier; the host never blocks.  Each rank spins on ITS OWN GPU.
// A rank that never arrives: by default the wait gives up after g_barrier_timeout_s (option "peer_barrier_timeout_s",
// default 300; 0 = wait for ever).  It does NOT trap -- a trap destroys the CUDA context of a healthy rank whose peer
// is merely slow (first-iteration plane build of a 25 GB slice, a host stall) -- it records the epoch in
// g_barrier_timed_out and lets the stream run on; fricc_stream_synchronize and fricc_get_stat("peer_barrier_timeouts")
// read the word and the former turns it into an error, so the host learns that results behind that barrier are void.
static __device__ unsigned long long g_barrier_timed_out;  // epoch of the first barrier that gave up since the last check
static int g_barrier_timeout_s = 300;
static bool g_barrier_used = false;
void set_peer_barrier_timeout_s(int s) { g_barrier_timeout_s = s < 0 ? 0 : s; }

__global__ void peer_barrier_kernel(PeerPtrs flags, int n_peers, int self, unsigned long long epoch,
                                    unsigned long long timeout_ns) {
  const int q = threadIdx.x;
  if (q >= n_peers) return;
  unsigned long long* mine = reinterpret_cast<unsigned long long*>(flags.p[self]);
  if (q != self) {
    unsigned long long* theirs = reinterpret_cast<unsigned long long*>(flags.p[q]) + self;
    // peer stores of earlier kernels on this stream -- and of a joined side stream -- are complete at the kernel boundary;
    // the fence makes that explicit for every scope before the flag goes out
    __threadfence_system();
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(theirs), "l"(epoch) : "memory");
    unsigned long long t0, now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    unsigned long long seen = 0;
    unsigned spins = 0;
    do {
      asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(mine + q) : "memory");
      if (seen < epoch && timeout_ns && (++spins & 1023u) == 0) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (now - t0 > timeout_ns) {
          atomicCAS(&g_barrier_timed_out, 0ULL, epoch);
          break;
        }
      }
    } while (seen < epoch);
  }
}

void peer_barrier(double* const* flags_peer, int n_peers, int self, unsigned long long epoch) {
  if (n_peers <= 1) return;
  if (n_peers > MAX_PEERS) fail(FRICC_EINVAL, "at most %d peers", MAX_PEERS);
  PeerPtrs pp;
  for (int q = 0; q < MAX_PEERS; q++) pp.p[q] = q < n_peers ? flags_peer[q] : nullptr;
  g_barrier_used = true;
  FRICC_LAUNCH(peer_barrier_kernel, 1, 32, 0, pp, n_peers, self, epoch,
               (unsigned long long)g_barrier_timeout_s * 1000000000ULL);
}

// epoch of the first barrier that gave up since the last call (0 = none); clears the word.  Synchronises the stream.
unsigned long long peer_barrier_timed_out() {
  if (!g_barrier_used) return 0;
  unsigned long long h = 0;
  FRICC_CUDA(cudaMemcpyFromSymbolAsync(&h, g_barrier_timed_out, sizeof(h), 0, cudaMemcpyDeviceToHost, stream()));
  FRICC_CUDA(cudaStreamSynchronize(stream()));
  if (h) {
    const unsigned long long zero = 0;
    FRICC_CUDA(cudaMemcpyToSymbolAsync(g_barrier_timed_out, &zero, sizeof(zero), 0, cudaMemcpyHostToDevice, stream()));
    FRICC_CUDA(cudaStreamSynchronize(stream()));
  }
  return h;
}

// counter-based uniform fill: dst[i] = U(0,1) of (seed, first_index + i) -- a synthetic operand whose values depend on
// the GLOBAL element index only, so that any sharding or sub-sharding of it holds the same numbers (bench.py)
__global__ void fill_uniform_kernel(double* __restrict__ dst, int64_t n, unsigned long long first, unsigned long long seed) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    unsigned long long z = (first + (unsigned long long)i) * 0x9E3779B97F4A7C15ULL + seed;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    dst[i] = (double)(z >> 11) * 1.1102230246251565e-16;  // 2^-53
  }
}

// ---------------------------------------------------------------------------------------------------------------
// panel -> 4-D output (+ peers), transposing.  For terms whose output's unit-stride axis is neither the L-minor nor
// the R-minor index (out(i,j,a,b) with R = (b,j), L = (a,i): 1302 / 1202 / 1303, and 1203) a direct epilogue write is
// one 32-byte sector per element.  There the kernels write a coalesced [R][ld] panel and this kernel moves it through a
// 32x32 shared-memory tile: reads follow the panel's contiguous index l2, writes follow out's contiguous index X
// (r1 or l1).  new = out_local + panel is stored locally AND into every peer's copy (plain stores over NVLink), so
// the accumulate pass and the exchange are one pass.   XMODE 0: X = r1 (fixed r2, l1)   XMODE 2: X = l1 (fixed r1, r2)
// ---------------------------------------------------------------------------------------------------------------
template <int XMODE>
__global__ void __launch_bounds__(256)
scatter_panel_tiled_kernel(const double* __restrict__ panel, int64_t ld, int64_t l_begin, int64_t l_count, int64_t nR1,
                           int64_t l1_first, int64_t n_l1, double* __restrict__ out, PeerPtrs peers, int n_peers,
                           int self, OutMap om, int64_t n_fixed, int64_t n_xb, int64_t nvb) {
  __shared__ double tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int64_t vb = blockIdx.x; vb < nvb; vb += gridDim.x) {  // capped grid, see push_columns_kernel
  const int64_t fixed = vb % n_fixed;
  const int64_t vb2 = vb / n_fixed;
  const int64_t x0 = (vb2 % n_xb) * 32, q0 = (vb2 / n_xb) * 32;  // q = l2
  int64_t r1 = 0, r2 = 0, l1 = 0, nX;
  if (XMODE == 0) { r2 = fixed % om.nR2; l1 = l1_first + fixed / om.nR2; nX = nR1; }
  else { r2 = fixed % om.nR2; r1 = fixed / om.nR2; nX = n_l1; }
  const int64_t l_end = l_begin + l_count;
  for (int r = ty; r < 32; r += 8) {
    const int64_t x = x0 + r, q = q0 + tx;
    if (x < nX && q < om.nL2) {
      int64_t R, l;
      if (XMODE == 0) { R = x * om.nR2 + r2; l = l1 * om.nL2 + q; }
      else { R = r1 * om.nR2 + r2; l = (l1_first + x) * om.nL2 + q; }
      if (l >= l_begin && l < l_end) tile[r][tx] = panel[R * ld + (l - l_begin)];
    }
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int64_t q = q0 + r, x = x0 + tx;
    if (x < nX && q < om.nL2) {
      int64_t l, o;
      if (XMODE == 0) { l = l1 * om.nL2 + q; o = x * om.sR1 + r2 * om.sR2 + l1 * om.sL1 + q * om.sL2; }
      else { l = (l1_first + x) * om.nL2 + q; o = r1 * om.sR1 + r2 * om.sR2 + (l1_first + x) * om.sL1 + q * om.sL2; }
      if (l >= l_begin && l < l_end) {
        const double v = out[o] + tile[tx][r];
        out[o] = v;
        for (int d = 1; d < n_peers; d++) {
          int p = self + d;
          if (p >= n_peers) p -= n_peers;
          peers.p[p][o] = v;
        }
      }
    }
  }
  __syncthreads();  // the tile is reused by the next virtual block
  }
}

__global__ void scatter_panel_kernel(const double* __restrict__ panel, int64_t ld, int64_t l_begin, int64_t l_count,
                                     int64_t nR, double* __restrict__ out, PeerPtrs peers, int n_peers, int self,
                                     OutMap om) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nR * l_count) return;
  const int64_t R = idx / l_count, c = idx - R * l_count;
  const int64_t l = l_begin + c;
  const int64_t r1 = R / om.nR2, r2 = R - r1 * om.nR2;
  const int64_t l1 = l / om.nL2, l2 = l - l1 * om.nL2;
  const int64_t o = r1 * om.sR1 + r2 * om.sR2 + l1 * om.sL1 + l2 * om.sL2;
  const double v = out[o] + panel[R * ld + c];
  out[o] = v;
  for (int d = 1; d < n_peers; d++) {
    int p = self + d;
    if (p >= n_peers) p -= n_peers;
    peers.p[p][o] = v;
  }
}

void scatter_panel(const double* panel, int64_t ld, int64_t l_begin, int64_t l_count, int64_t nR, double* out,
                   double* const* peer_out, int n_peers, int self, const OutMap& om) {
  if (nR == 0 || l_count == 0) return;
  if (n_peers > MAX_PEERS) fail(FRICC_EINVAL, "at most %d peers", MAX_PEERS);
  PeerPtrs pp;
  for (int q = 0; q < MAX_PEERS; q++) pp.p[q] = (peer_out && q < n_peers) ? peer_out[q] : nullptr;
  if (!peer_out) { n_peers = 1; self = 0; }
  const int64_t nR1 = nR / om.nR2;
  const int64_t l1_first = l_begin / om.nL2, l1_last = (l_begin + l_count - 1) / om.nL2;
  const int64_t n_l1 = l1_last - l1_first + 1;
  int xmode = -1;
  int64_t fixed = 0, nX = 0;
  if (om.sL2 != 1) {
    if (om.sR1 == 1) { xmode = 0; fixed = om.nR2 * n_l1; nX = nR1; }
    else if (om.sL1 == 1) { xmode = 2; fixed = nR1 * om.nR2; nX = n_l1; }
  }
  if (xmode >= 0 && fixed > 0 && fixed < ((int64_t)1 << 31) && ceil_div(nX, 32) <= 65535 && ceil_div(om.nL2, 32) <= 65535) {
    const int64_t n_xb = ceil_div(nX, 32), nvb = fixed * n_xb * ceil_div(om.nL2, 32);
    // without peers this is a local HBM-bound pass: let it fill the GPU
    const unsigned grid = peer_out ? push_grid(nvb) : (unsigned)std::min<int64_t>(nvb, (int64_t)num_sms() * 8);
    if (peer_out) {
      prefer_split(scatter_panel_tiled_kernel<0>);
      prefer_split(scatter_panel_tiled_kernel<2>);
    }
    if (xmode == 0)
      FRICC_LAUNCH(scatter_panel_tiled_kernel<0>, grid, 256, 0, panel, ld, l_begin, l_count, nR1, l1_first, n_l1, out, pp,
                   n_peers, self, om, fixed, n_xb, nvb);
    else
      FRICC_LAUNCH(scatter_panel_tiled_kernel<2>, grid, 256, 0, panel, ld, l_begin, l_count, nR1, l1_first, n_l1, out, pp,
                   n_peers, self, om, fixed, n_xb, nvb);
    return;
  }
  FRICC_LAUNCH(scatter_panel_kernel, (unsigned)ceil_div(nR * l_count, 256), 256, 0, panel, ld, l_begin, l_count, nR, out,
               pp, n_peers, self, om);
}

}  // namespace fricc

using namespace fricc;

extern "C" {

int fricc_symm_alloc(int64_t bytes, void** ptr, void* handle64) {
  FRICC_API_BEGIN
  if (bytes <= 0 || !ptr || !handle64) fail(FRICC_EINVAL, "bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  FRICC_CUDA(cudaMalloc(&p, (size_t)bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    cudaFree(p);
    fail(FRICC_ECUDA, "cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
  }
  memcpy(handle64, &h, 64);
  *ptr = p;
  FRICC_API_END
}

int fricc_symm_open(const void* handle64, void** ptr) {
  FRICC_API_BEGIN
  if (!handle64 || !ptr) fail(FRICC_EINVAL, "bad arguments");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  void* p = nullptr;
  FRICC_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *ptr = p;
  FRICC_API_END
}

int fricc_symm_close(void* ptr) {
  FRICC_API_BEGIN
  if (ptr) FRICC_CUDA(cudaIpcCloseMemHandle(ptr));
  FRICC_API_END
}

int fricc_symm_free(void* ptr) {
  FRICC_API_BEGIN
  if (ptr) FRICC_CUDA(cudaFree(ptr));
  FRICC_API_END
}

int fricc_push_columns(const double* out_local_dev, void* const* peer_out_dev, int n_peers, int self, const char* term,
                       const int64_t t_dims[4], const int64_t out_shape[4], const int64_t out_strides[4],
                       int64_t l_begin, int64_t l_count) {
  FRICC_API_BEGIN
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!out_local_dev || !peer_out_dev || !t_dims || !out_shape || !out_strides || n_peers < 1 || self < 0 ||
      self >= n_peers || l_begin < 0 || l_count < 0)
    fail(FRICC_EINVAL, "bad arguments");
  for (int q = 0; q < 2; q++)
    if (out_shape[ts->oR[q]] != t_dims[ts->tR[q]]) fail(FRICC_EINVAL, "term %s: out does not match T on the owner pair", ts->name);
  const int64_t nL = out_shape[ts->oL[0]] * out_shape[ts->oL[1]];
  if (l_begin + l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  OutMap om;
  om.nR2 = t_dims[ts->tR[1]];
  om.nL2 = out_shape[ts->oL[1]];
  om.sR1 = out_strides[ts->oR[0]];
  om.sR2 = out_strides[ts->oR[1]];
  om.sL1 = out_strides[ts->oL[0]];
  om.sL2 = out_strides[ts->oL[1]];
  const int64_t nR = t_dims[ts->tR[0]] * t_dims[ts->tR[1]];
  double* peers[MAX_PEERS];
  if (n_peers > MAX_PEERS) fail(FRICC_EINVAL, "at most %d peers", MAX_PEERS);
  for (int q = 0; q < n_peers; q++) peers[q] = static_cast<double*>(peer_out_dev[q]);
  push_columns(out_local_dev, peers, n_peers, self, om, nR, l_begin, l_count);
  FRICC_API_END
}

int fricc_push_view(const double* view_local_dev, void* const* view_peer_dev, int n_peers, int self,
                    const int64_t shape[4], const int64_t strides[4]) {
  FRICC_API_BEGIN
  if (!view_local_dev || !view_peer_dev || !shape || !strides || n_peers < 1 || self < 0 || self >= n_peers)
    fail(FRICC_EINVAL, "bad arguments");
  double* peers[MAX_PEERS];
  if (n_peers > MAX_PEERS) fail(FRICC_EINVAL, "at most %d peers", MAX_PEERS);
  for (int q = 0; q < n_peers; q++) peers[q] = static_cast<double*>(view_peer_dev[q]);
  push_view(view_local_dev, peers, n_peers, self, shape, strides);
  FRICC_API_END
}

int fricc_peer_barrier(void* const* flags_peer_dev, int n_peers, int self, uint64_t epoch) {
  FRICC_API_BEGIN
  if (!flags_peer_dev || n_peers < 1 || self < 0 || self >= n_peers || n_peers > MAX_PEERS) fail(FRICC_EINVAL, "bad arguments");
  double* peers[MAX_PEERS];
  for (int q = 0; q < n_peers; q++) peers[q] = static_cast<double*>(flags_peer_dev[q]);
  peer_barrier(peers, n_peers, self, (unsigned long long)epoch);
  FRICC_API_END
}

int fricc_fill_uniform(double* dst_dev, int64_t n, uint64_t first_index, uint64_t seed) {
  FRICC_API_BEGIN
  if (n < 0 || (n > 0 && !dst_dev)) fail(FRICC_EINVAL, "bad arguments");
  if (n > 0)
    FRICC_LAUNCH(fill_uniform_kernel, (unsigned)std::min<int64_t>(ceil_div(n, 256), (int64_t)num_sms() * 16), 256, 0, dst_dev, n,
                 (unsigned long long)first_index, (unsigned long long)seed);
  FRICC_API_END
}

int fricc_scatter_panel(const double* panel_dev, int64_t ld, int64_t l_begin, int64_t l_count, const char* term,
                        const int64_t t_dims[4], double* out_local_dev, void* const* peer_out_dev, int n_peers, int self,
                        const int64_t out_shape[4], const int64_t out_strides[4]) {
  FRICC_API_BEGIN
  const TermSpec* ts = find_term(term);
  if (!ts) fail(FRICC_EINVAL, "CASE NOT IMPLEMENTED: unknown contraction term '%s'", term ? term : "(null)");
  if (!panel_dev || !out_local_dev || !t_dims || !out_shape || !out_strides || ld < l_count || l_begin < 0 || l_count < 0 ||
      (peer_out_dev && (n_peers < 1 || self < 0 || self >= n_peers)))
    fail(FRICC_EINVAL, "bad arguments");
  for (int q = 0; q < 2; q++)
    if (out_shape[ts->oR[q]] != t_dims[ts->tR[q]]) fail(FRICC_EINVAL, "term %s: out does not match T on the owner pair", ts->name);
  const int64_t nL = out_shape[ts->oL[0]] * out_shape[ts->oL[1]];
  if (l_begin + l_count > nL) fail(FRICC_EINVAL, "column range out of bounds");
  OutMap om;
  om.nR2 = t_dims[ts->tR[1]];
  om.nL2 = out_shape[ts->oL[1]];
  om.sR1 = out_strides[ts->oR[0]];
  om.sR2 = out_strides[ts->oR[1]];
  om.sL1 = out_strides[ts->oL[0]];
  om.sL2 = out_strides[ts->oL[1]];
  const int64_t nR = t_dims[ts->tR[0]] * t_dims[ts->tR[1]];
  double* peers[MAX_PEERS];
  if (peer_out_dev) {
    if (n_peers > MAX_PEERS) fail(FRICC_EINVAL, "at most %d peers", MAX_PEERS);
    for (int q = 0; q < n_peers; q++) peers[q] = static_cast<double*>(peer_out_dev[q]);
  }
  scatter_panel(panel_dev, ld, l_begin, l_count, nR, out_local_dev, peer_out_dev ? peers : nullptr, n_peers, self, om);
  FRICC_API_END
}

}  // extern "C"
