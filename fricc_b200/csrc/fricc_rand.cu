// fricc_rand.cu -- a glibc-compatible rand() stream (host side).
//
// The reference draws its uniforms with libc rand() and never seeds it (/root/reference/src/fri_utils.cpp:47,74), so a
// stock build sees glibc's TYPE_3 additive-feedback generator seeded with 1.  To be a drop-in with the SAME sample
// sequence for the same seed, the library carries its own copy of that generator:
//     r[0] = seed (0 -> 1);  r[i] = 16807 * r[i-1] mod (2^31 - 1)            i = 1..30   (Schrage form)
//     r[i] = r[i-31]                                                          i = 31..33
//     r[i] = r[i-31] + r[i-3]  (mod 2^32)                                     i >= 34
//     k-th output of rand() = r[k + 344] >> 1
// (tests/test_capi_symbols.py checks this against the C library's srand()/rand()).
//
// Pivotal sampling consumes one draw per visited element (2.56e8 at the headline shape), far too many to produce on the
// host, so the same stream is also generated ON THE DEVICE: the recurrence is linear over Z/2^32 with characteristic
// polynomial P(x) = x^31 - x^28 - 1, hence  r[m+L] = sum_j c_j r[m+j]  with  c(x) = x^L mod P(x).  Jump polynomials for
// L = CH * 2^b are built on the host by square-and-reduce; every device thread jumps to the start of its own chunk of
// CH draws (binary decomposition of the chunk number) and then runs the plain recurrence.  The host stream is advanced
// by the same jump (rand_skip) instead of stepping.
#include <mutex>

#include "fricc_fri.cuh"

namespace fricc {

namespace {
struct State {
  uint32_t ring[31];  // ring[i % 31] == r[i] for the 31 most recent i
  int64_t i;          // index of the next r[] to produce
  int64_t consumed;
};
State g_state;
bool g_seeded = false;
std::mutex g_mu;

void seed_locked(uint32_t seed) {
  if (seed == 0) seed = 1;
  int32_t r[34];
  r[0] = (int32_t)seed;
  for (int i = 1; i < 31; i++) {
    int64_t hi = r[i - 1] / 127773, lo = r[i - 1] % 127773;
    int64_t word = 16807 * lo - 2836 * hi;
    if (word < 0) word += 2147483647;
    r[i] = (int32_t)word;
  }
  for (int i = 0; i < 31; i++) g_state.ring[i] = (uint32_t)r[i];
  // r[31..33] = r[0..2]: the ring slots already hold them (slot i%31)
  g_state.i = 34;
  g_state.consumed = 0;
  // discard r[34..343]
  for (int64_t i = 34; i < 344; i++) {
    uint32_t v = g_state.ring[(i - 31) % 31] + g_state.ring[(i - 3) % 31];
    g_state.ring[i % 31] = v;
  }
  g_state.i = 344;
  g_seeded = true;
}
}  // namespace

void rand_seed(uint32_t seed) {
  std::lock_guard<std::mutex> lk(g_mu);
  seed_locked(seed);
}

void rand_fill(int32_t* out, int64_t n) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_seeded) seed_locked(1);
  int64_t i = g_state.i;
  for (int64_t k = 0; k < n; k++, i++) {
    uint32_t v = g_state.ring[(i - 31) % 31] + g_state.ring[(i - 3) % 31];
    g_state.ring[i % 31] = v;
    out[k] = (int32_t)(v >> 1);
  }
  g_state.i = i;
  g_state.consumed += n;
}

// ---------------------------------------------------------------------------------------------------------------
// jump-ahead:  c(x) = x^L mod P(x),  P(x) = x^31 - x^28 - 1  over Z/2^32
// ---------------------------------------------------------------------------------------------------------------
namespace {
struct Poly {
  uint32_t c[31];
};

Poly poly_mul(const Poly& a, const Poly& b) {
  uint32_t t[61] = {0};
  for (int i = 0; i < 31; i++)
    for (int j = 0; j < 31; j++) t[i + j] += a.c[i] * b.c[j];
  for (int d = 60; d >= 31; d--) {  // x^d = x^(d-3) + x^(d-31)
    t[d - 3] += t[d];
    t[d - 31] += t[d];
  }
  Poly r;
  for (int i = 0; i < 31; i++) r.c[i] = t[i];
  return r;
}

Poly poly_x_pow(uint64_t L) {
  Poly r{}, b{};
  r.c[0] = 1;
  b.c[1] = 1;
  while (L) {
    if (L & 1) r = poly_mul(r, b);
    b = poly_mul(b, b);
    L >>= 1;
  }
  return r;
}

// w[0..30] = r[i-31..i-1]  ->  r[i-31+L .. i-1+L]
void jump_host(uint32_t w[31], const Poly& c) {
  uint32_t e[61];
  for (int s = 0; s < 31; s++) e[s] = w[s];
  for (int s = 0; s < 30; s++) e[31 + s] = e[s] + e[28 + s];
  for (int s = 0; s < 31; s++) {
    uint32_t acc = 0;
    for (int j = 0; j < 31; j++) acc += c.c[j] * e[s + j];
    w[s] = acc;
  }
}

void window_from_state(const State& st, uint32_t w[31]) {  // oldest first
  for (int s = 0; s < 31; s++) w[s] = st.ring[(st.i - 31 + s) % 31];
}
void state_from_window(State& st, const uint32_t w[31], int64_t new_i) {
  st.i = new_i;
  for (int s = 0; s < 31; s++) st.ring[(new_i - 31 + s) % 31] = w[s];
}

constexpr int RG_CH = 31 * 32;  // draws per device thread
constexpr int RG_BITS = 40;
Poly g_chunk_polys[RG_BITS];    // x^(RG_CH * 2^b) mod P
bool g_polys_ready = false;
void ensure_polys() {
  if (g_polys_ready) return;
  Poly p = poly_x_pow(RG_CH);
  for (int b = 0; b < RG_BITS; b++) {
    g_chunk_polys[b] = p;
    p = poly_mul(p, p);
  }
  g_polys_ready = true;
}

struct Window {
  uint32_t w[31];
};

__global__ void __launch_bounds__(128)
rand_generate_kernel(Window base, const Poly* __restrict__ polys, int64_t nchunks, int64_t n, int32_t* __restrict__ out) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= nchunks) return;
  uint32_t w[31];
#pragma unroll
  for (int s = 0; s < 31; s++) w[s] = base.w[s];
  // jump to the start of chunk c
  for (int b = 0; b < RG_BITS; b++) {
    if (!((c >> b) & 1)) continue;
    uint32_t e[61];
#pragma unroll
    for (int s = 0; s < 31; s++) e[s] = w[s];
#pragma unroll
    for (int s = 0; s < 30; s++) e[31 + s] = e[s] + e[28 + s];
    uint32_t cf[31];
#pragma unroll
    for (int j = 0; j < 31; j++) cf[j] = polys[b].c[j];
#pragma unroll
    for (int s = 0; s < 31; s++) {
      uint32_t acc = 0;
#pragma unroll
      for (int j = 0; j < 31; j++) acc += cf[j] * e[s + j];
      w[s] = acc;
    }
  }
  // the plain recurrence: 31 new values per round, window renamed in place
  int64_t k = c * RG_CH;
  for (int it = 0; it < RG_CH / 31 && k < n; it++) {
#pragma unroll
    for (int s = 0; s < 31; s++) {
      const uint32_t v = w[s] + (s < 3 ? w[s + 28] : w[s - 3]);  // r[i] = r[i-31] + r[i-3]
      w[s] = v;
      if (k + s < n) out[k + s] = (int32_t)(v >> 1);
    }
    k += 31;
  }
}
}  // namespace

// the next n values of rand() into DEVICE memory; the host stream is left where it was (callers advance it with rand_skip)
void rand_fill_device(int32_t* out_dev, int64_t n) {
  if (n <= 0) return;
  Window base;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_seeded) seed_locked(1);
    ensure_polys();
    window_from_state(g_state, base.w);
  }
  static DevBuf<Poly>* d_polys = nullptr;  // uploaded once per process (40 x 124 bytes)
  if (!d_polys) {
    d_polys = new DevBuf<Poly>(RG_BITS);
    d_polys->from_host(g_chunk_polys, RG_BITS);
  }
  const int64_t nchunks = ceil_div(n, RG_CH);
  FRICC_LAUNCH(rand_generate_kernel, (unsigned)ceil_div(nchunks, 128), 128, 0, base, d_polys->p, nchunks, n, out_dev);
}

// advance the host stream by `count` draws without producing them
void rand_skip(int64_t count) {
  if (count <= 0) return;
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_seeded) seed_locked(1);
  uint32_t w[31];
  window_from_state(g_state, w);
  Poly b{};
  b.c[1] = 1;  // x^(2^bit)
  for (uint64_t L = (uint64_t)count; L; L >>= 1) {
    if (L & 1) jump_host(w, b);
    b = poly_mul(b, b);
  }
  state_from_window(g_state, w, g_state.i + count);
  g_state.consumed += count;
}

int64_t rand_consumed() {
  std::lock_guard<std::mutex> lk(g_mu);
  return g_seeded ? g_state.consumed : 0;
}

RandSnapshot rand_snapshot() {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_seeded) seed_locked(1);
  RandSnapshot s;
  memcpy(s.ring, g_state.ring, sizeof(s.ring));
  s.i = g_state.i;
  s.consumed = g_state.consumed;
  return s;
}

void rand_restore(const RandSnapshot& s) {
  std::lock_guard<std::mutex> lk(g_mu);
  memcpy(g_state.ring, s.ring, sizeof(s.ring));
  g_state.i = s.i;
  g_state.consumed = s.consumed;
  g_seeded = true;
}

}  // namespace fricc
