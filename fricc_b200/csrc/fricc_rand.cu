// fricc_rand.cu -- a glibc-compatible rand() stream (host side).
//
// The reference draws its uniforms with libc rand() and never seeds it (/root/reference/src/fri_utils.cpp:47,74), so a
// stock build sees glibc's TYPE_3 additive-feedback generator seeded with 1.  To be a drop-in with the SAME sample
// sequence for the same seed, the library carries its own copy of that generator:
//     r[0] = seed (0 -> 1);  r[i] = 16807 * r[i-1] mod (2^31 - 1)            i = 1..30   (Schrage form)
//     r[i] = r[i-31]                                                          i = 31..33
//     r[i] = r[i-31] + r[i-3]  (mod 2^32)                                     i >= 34
//     k-th output of rand() = r[k + 344] >> 1
// (tests/test_rand.py checks this against the C library's srand()/rand()).
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
