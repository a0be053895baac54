// fricc_contract.cu -- SparseTensor4d x dense contractions as ONE primitive: bucketed sparse-row x dense-panel SpMM.
//
// The reference implements 11 terms (/root/reference/src/sparse_tensor.cpp:12-346) with the same loop nest: an OpenMP
// loop over an owner pair R of output indices; every owner scans ALL m stored entries for the ones that belong to it
// (o^2*m or o*v*m index tests) and then walks W with a large stride.  Here the entries are bucketed by R once
// (sort by (R,K), CSR row pointers), W is brought into a [K][L] row-major panel so that the row a nonzero needs is
// contiguous, and one kernel does   out[R, L] += scale * sum_{e in bucket R} val_e * Wp[K_e, L]
// with the accumulators in registers and exactly one read-modify-write of each output element (no atomics).
#include "fricc_dense.cuh"

#include <algorithm>

#include "fricc_sort.cuh"

namespace fricc {

// ---------------------------------------------------------------------------------------------------------------
// term table (every row cites the reference kernel it replaces).  T entry idx = (idx0..idx3).
// ---------------------------------------------------------------------------------------------------------------
static const TermSpec kTerms[] = {
    // name   tR      tK      wK      wL      oR      oL     scale  einsum
    // sparse_tensor.cpp:40-86    out(i,j,a,b) += W(a,b,c,d) * t(i,j,c,d)
    {"2323", {0, 1}, {2, 3}, {2, 3}, {0, 1}, {0, 1}, {2, 3}, 1.0, "abcd,ijcd->ijab"},
    // sparse_tensor.cpp:12-37    out(i,j,a,b) += W(k,l,i,j) * t(k,l,a,b)
    {"0101", {2, 3}, {0, 1}, {0, 1}, {2, 3}, {2, 3}, {0, 1}, 1.0, "klij,klab->ijab"},
    // sparse_tensor.cpp:205-230  out(k,l,i,j) += ovov(k,c,l,d) * t(i,j,c,d)
    {"1323", {0, 1}, {2, 3}, {1, 3}, {0, 2}, {2, 3}, {0, 1}, 1.0, "kcld,ijcd->klij"},
    // sparse_tensor.cpp:89-115   out(i,j,a,b) += 2 * W(a,k,i,c) * t(k,j,c,b)      R=(b,j) K=(k,c) L=(a,i)
    {"1302", {3, 1}, {0, 2}, {1, 3}, {0, 2}, {3, 1}, {2, 0}, 2.0, "2*akic,kjcb->ijab"},
    // sparse_tensor.cpp:118-144  out(i,j,a,b) -= W(a,k,c,i) * t(k,j,c,b)          R=(b,j) K=(k,c) L=(a,i)
    {"1202", {3, 1}, {0, 2}, {1, 2}, {0, 3}, {3, 1}, {2, 0}, -1.0, "-akci,kjcb->ijab"},
    // sparse_tensor.cpp:147-173  out(i,j,a,b) += W(a,k,i,c) * t(k,j,b,c)          R=(b,j) K=(k,c) L=(a,i)
    {"1303", {2, 1}, {0, 3}, {1, 3}, {0, 2}, {3, 1}, {2, 0}, 1.0, "akic,kjbc->ijab"},
    // sparse_tensor.cpp:176-202  out(i,j,a,b) += W(b,k,c,i) * t(k,j,a,c)          R=(a,j) K=(k,c) L=(b,i)
    {"1203", {2, 1}, {0, 3}, {1, 2}, {0, 3}, {2, 1}, {3, 0}, 1.0, "bkci,kjac->ijab"},
    // sparse_tensor.cpp:233-259  out(a,k,i,c) -= 0.5 * ovov(l,d,k,c) * t(i,l,d,a) R=(a,i) K=(l,d) L=(k,c)
    {"0112", {3, 0}, {1, 2}, {0, 1}, {2, 3}, {0, 2}, {1, 3}, -0.5, "-0.5*ldkc,ilda->akic"},
    // sparse_tensor.cpp:262-289  out(a,k,i,c) -= 0.5 * ovov(l,c,k,d) * t(i,l,a,d) R=(a,i) K=(l,d) L=(k,c)
    {"0313", {2, 0}, {1, 3}, {0, 3}, {2, 1}, {0, 2}, {1, 3}, -0.5, "-0.5*lckd,ilad->akic"},
    // sparse_tensor.cpp:292-317  out(a,k,i,c) += ovov(l,d,k,c) * t(i,l,a,d)       R=(a,i) K=(l,d) L=(k,c)
    {"0113", {2, 0}, {1, 3}, {0, 1}, {2, 3}, {0, 2}, {1, 3}, 1.0, "ldkc,ilad->akic"},
    // sparse_tensor.cpp:320-346  out(a,k,c,i) -= 0.5 * ovov(l,c,k,d) * t(i,l,d,a) R=(a,i) K=(l,d) L=(k,c)
    {"0312", {3, 0}, {1, 2}, {0, 3}, {2, 1}, {0, 3}, {1, 2}, -0.5, "-0.5*lckd,ilda->akci"},
    // experimental 2323 variants (sparse_tensor.cpp:389-459): same sum, different CPU threading
    {"taskloop", {0, 1}, {2, 3}, {2, 3}, {0, 1}, {0, 1}, {2, 3}, 1.0, "abcd,ijcd->ijab"},
    {"atomic", {0, 1}, {2, 3}, {2, 3}, {0, 1}, {0, 1}, {2, 3}, 1.0, "abcd,ijcd->ijab"},
};

const TermSpec* find_term(const char* name) {
  if (!name) return nullptr;
  for (const TermSpec& t : kTerms)
    if (strcmp(t.name, name) == 0) return &t;
  return nullptr;
}

// ---------------------------------------------------------------------------------------------------------------
// layout: entries -> (R*nK + K, value) -> sort -> CSR
// ---------------------------------------------------------------------------------------------------------------
struct Dims4 {
  int64_t d[4];
};

__global__ void make_keys_kernel(const int64_t* __restrict__ flat, const double* __restrict__ vals, int64_t m, Dims4 dm,
                                 int r0, int r1, int k0, int k1, uint64_t* __restrict__ keys,
                                 uint64_t* __restrict__ pay) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  int64_t f = flat[e];
  int64_t ix[4];
  ix[3] = f % dm.d[3]; f /= dm.d[3];
  ix[2] = f % dm.d[2]; f /= dm.d[2];
  ix[1] = f % dm.d[1]; f /= dm.d[1];
  ix[0] = f;
  int64_t R = ix[r0] * dm.d[r1] + ix[r1];
  int64_t K = ix[k0] * dm.d[k1] + ix[k1];
  int64_t nK = dm.d[k0] * dm.d[k1];
  keys[e] = (uint64_t)(R * nK + K);
  pay[e] = (uint64_t)__double_as_longlong(vals[e]);
}

__global__ void split_keys_kernel(const uint64_t* __restrict__ keys, const uint64_t* __restrict__ pay, int64_t m,
                                  int64_t nK, uint32_t* __restrict__ kidx, double* __restrict__ vals) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  kidx[e] = (uint32_t)(keys[e] % (uint64_t)nK);
  vals[e] = __longlong_as_double((long long)pay[e]);
}

// rowptr[R] = first sorted position whose key >= R*nK   (R = 0..nR)
__global__ void rowptr_kernel(const uint64_t* __restrict__ keys, int64_t m, int64_t nK, int64_t nR,
                              int64_t* __restrict__ rowptr) {
  int64_t R = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (R > nR) return;
  uint64_t target = (uint64_t)R * (uint64_t)nK;
  int64_t lo = 0, hi = m;
  while (lo < hi) {
    int64_t mid = (lo + hi) >> 1;
    if (keys[mid] < target) lo = mid + 1;
    else hi = mid;
  }
  rowptr[R] = lo;
}

// entries [first, first + count) of T, bucketed by the owner pair of `ts`: sort by (R, K), CSR row pointers
void build_csr(fricc_st* T, const TermSpec& ts, int64_t first, int64_t count, int64_t nR, int64_t nK,
               DevBuf<int64_t>& rowptr, DevBuf<uint32_t>& kidx, DevBuf<double>& vals) {
  const int64_t* d = T->dims;
  rowptr.alloc((size_t)nR + 1);
  kidx.alloc((size_t)std::max<int64_t>(count, 1));
  vals.alloc((size_t)std::max<int64_t>(count, 1));
  if (count == 0) {
    rowptr.zero();
    return;
  }
  const int64_t m = count;
  DevBuf<uint64_t> ka((size_t)m), kb((size_t)m), pa((size_t)m), pb((size_t)m);
  Dims4 dm{{d[0], d[1], d[2], d[3]}};
  unsigned g = (unsigned)ceil_div(m, 256);
  FRICC_LAUNCH(make_keys_kernel, g, 256, 0, T->flat.p + first, T->vals.p + first, m, dm, ts.tR[0], ts.tR[1], ts.tK[0],
               ts.tK[1], ka.p, pa.p);
  int bits = 1;
  while (bits < 63 && ((int64_t)1 << bits) < nR * nK) bits++;
  std::vector<Digit> digits;
  for (int s = 0; s < bits; s += 8) digits.push_back(Digit{0, s, 0});
  uint64_t* kk[2] = {ka.p, kb.p};
  uint64_t* pp[2] = {pa.p, pb.p};
  int r = radix_sort_pairs(kk, pp, m, digits);  // stable: duplicates keep their storage order
  FRICC_LAUNCH(split_keys_kernel, g, 256, 0, kk[r], pp[r], m, nK, kidx.p, vals.p);
  FRICC_LAUNCH(rowptr_kernel, (unsigned)ceil_div(nR + 1, 256), 256, 0, kk[r], m, nK, nR, rowptr.p);
}

Layout& get_layout(fricc_st* T, const TermSpec& ts, bool need_csr) {
  Layout* found = nullptr;
  for (Layout& l : T->layouts)
    if (l.r0 == ts.tR[0] && l.r1 == ts.tR[1]) found = &l;
  if (!found) {
    T->layouts.emplace_back();
    Layout& lay = T->layouts.back();
    lay.r0 = ts.tR[0];
    lay.r1 = ts.tR[1];
    const int64_t* d = T->dims;
    lay.nR = d[ts.tR[0]] * d[ts.tR[1]];
    lay.nK = d[ts.tK[0]] * d[ts.tK[1]];
    if (lay.nK >= ((int64_t)1 << 32)) fail(FRICC_EINVAL, "contracted pair too large for 32-bit packing");
    found = &lay;
  }
  if (need_csr && !found->csr_built) {
    build_csr(T, ts, 0, T->nnz, found->nR, found->nK, found->rowptr, found->kidx, found->vals);
    found->csr_built = true;
  }
  return *found;
}

// ---------------------------------------------------------------------------------------------------------------
// W -> panel
// ---------------------------------------------------------------------------------------------------------------
struct PermArgs {
  int64_t sK1, sK2, sL1, sL2;  // element strides of W along (K major, K minor, L major, L minor)
  int64_t nK1, nK2, nL2;
  int64_t l_begin, l_count;
};

__global__ void permute_gather_kernel(const double* __restrict__ W, PermArgs a, int64_t total, double* __restrict__ Wp) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  int64_t K = idx / a.l_count, c = idx - K * a.l_count;
  int64_t l = a.l_begin + c;
  int64_t src = (K / a.nK2) * a.sK1 + (K % a.nK2) * a.sK2 + (l / a.nL2) * a.sL1 + (l % a.nL2) * a.sL2;
  Wp[idx] = W[src];
}

// K-minor axis contiguous in W: 32x32 shared-memory transpose, coalesced on both sides
__global__ void __launch_bounds__(256)
permute_tiled_kernel(const double* __restrict__ W, PermArgs a, double* __restrict__ Wp) {
  __shared__ double tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t k1 = blockIdx.z;
  const int64_t k2_0 = (int64_t)blockIdx.y * 32, c0 = (int64_t)blockIdx.x * 32;
  for (int r = ty; r < 32; r += 8) {
    int64_t c = c0 + r, k2 = k2_0 + tx;
    if (c < a.l_count && k2 < a.nK2) {
      int64_t l = a.l_begin + c;
      tile[r][tx] = W[k1 * a.sK1 + k2 * a.sK2 + (l / a.nL2) * a.sL1 + (l % a.nL2) * a.sL2];
    }
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    int64_t k2 = k2_0 + r, c = c0 + tx;
    if (c < a.l_count && k2 < a.nK2) Wp[(k1 * a.nK2 + k2) * a.l_count + c] = tile[tx][r];
  }
}

void permute_w(const double* W, const int64_t shape[4], const int64_t strides[4], const TermSpec& ts, int64_t l_begin,
               int64_t l_count, double* Wp) {
  PermArgs a;
  a.sK1 = strides[ts.wK[0]]; a.sK2 = strides[ts.wK[1]];
  a.sL1 = strides[ts.wL[0]]; a.sL2 = strides[ts.wL[1]];
  a.nK1 = shape[ts.wK[0]]; a.nK2 = shape[ts.wK[1]];
  a.nL2 = shape[ts.wL[1]];
  a.l_begin = l_begin; a.l_count = l_count;
  const int64_t nK = a.nK1 * a.nK2;
  if (nK == 0 || l_count == 0) return;
  if (a.sK2 == 1 && a.sL2 != 1 && a.nK1 <= 65535 && ceil_div(a.nK2, 32) <= 65535) {
    dim3 grid((unsigned)ceil_div(l_count, 32), (unsigned)ceil_div(a.nK2, 32), (unsigned)a.nK1);
    FRICC_LAUNCH(permute_tiled_kernel, grid, 256, 0, W, a, Wp);
  } else {
    int64_t total = nK * l_count;
    FRICC_LAUNCH(permute_gather_kernel, (unsigned)ceil_div(total, 256), 256, 0, W, a, total, Wp);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// SpMM: one CTA = one bucket R x one tile of L columns; register accumulators; rows of Wp streamed with vector loads
// ---------------------------------------------------------------------------------------------------------------
constexpr int SPMM_CHUNK = 256;

template <int VEC>
__global__ void __launch_bounds__(256)
spmm_rows_kernel(const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ kidx, const double* __restrict__ vals,
                 const double* __restrict__ Wp, int64_t ldw, int64_t l_begin, int64_t l_count, double* __restrict__ out,
                 OutMap om, double scale) {
  __shared__ uint32_t sk[SPMM_CHUNK];
  __shared__ double sv[SPMM_CHUNK];
  const int64_t R = blockIdx.x;
  const int64_t e0 = rowptr[R], e1 = rowptr[R + 1];
  if (e0 == e1) return;
  const int64_t col = ((int64_t)blockIdx.y * blockDim.x + threadIdx.x) * VEC;
  const bool active = col < l_count;
  const double* wcol = Wp + col;
  double acc[VEC];
#pragma unroll
  for (int v = 0; v < VEC; v++) acc[v] = 0.0;

  for (int64_t c = e0; c < e1; c += SPMM_CHUNK) {
    const int cnt = (int)min((int64_t)SPMM_CHUNK, e1 - c);
    __syncthreads();
    for (int t = threadIdx.x; t < cnt; t += blockDim.x) {
      sk[t] = kidx[c + t];
      sv[t] = vals[c + t];
    }
    __syncthreads();
    if (active) {
#pragma unroll 8
      for (int t = 0; t < cnt; t++) {
        const double* wp = wcol + (int64_t)sk[t] * ldw;
        const double s = sv[t];
        if (VEC == 2) {
          double2 w = __ldg(reinterpret_cast<const double2*>(wp));
          acc[0] = fma(s, w.x, acc[0]);
          acc[VEC - 1] = fma(s, w.y, acc[VEC - 1]);
        } else {
          acc[0] = fma(s, __ldg(wp), acc[0]);
        }
      }
    }
  }
  if (active) {
    const int64_t r1 = R / om.nR2, r2 = R - r1 * om.nR2;
    const int64_t rbase = r1 * om.sR1 + r2 * om.sR2;
#pragma unroll
    for (int v = 0; v < VEC; v++) {
      int64_t l = l_begin + col + v;
      int64_t l1 = l / om.nL2, l2 = l - l1 * om.nL2;
      int64_t o = rbase + l1 * om.sL1 + l2 * om.sL2;
      out[o] += scale * acc[v];
    }
  }
}

// Wide-tile variant: each thread owns NV2 double2 column pairs, interleaved so that every 128-bit load of a warp covers
// 512 contiguous bytes; the bucket's (K, value) pairs are staged as ONE 16-byte shared-memory entry per nonzero, so a
// nonzero costs one broadcast LDS.128 + NV2 LDG.128 + 2*NV2 DFMA per thread.  (The operand path -- one 8-byte W element
// per FMA through L1 -- is the bound; the wider tile amortises the staged-entry reads and the address arithmetic.)
struct __align__(16) StagedEntry {
  unsigned long long koff;  // K * ldw (element offset of the panel row)
  double val;
};

template <int NV2>
__global__ void __launch_bounds__(256)
spmm_rows_wide_kernel(const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ kidx,
                      const double* __restrict__ vals, const double* __restrict__ Wp, int64_t ldw, int64_t l_begin,
                      int64_t l_count, double* __restrict__ out, OutMap om, double scale) {
  __shared__ StagedEntry se[SPMM_CHUNK];
  const int64_t R = blockIdx.x;
  const int64_t e0 = rowptr[R], e1 = rowptr[R + 1];
  if (e0 == e1) return;
  const int stride = blockDim.x * 2;  // columns between a thread's consecutive double2
  const int64_t col0 = (int64_t)blockIdx.y * (stride * NV2) + threadIdx.x * 2;
  bool act[NV2];
  double acc[2 * NV2];
#pragma unroll
  for (int j = 0; j < NV2; j++) {
    act[j] = col0 + (int64_t)j * stride < l_count;
    acc[2 * j] = 0.0;
    acc[2 * j + 1] = 0.0;
  }
  const double* wcol = Wp + col0;

  for (int64_t c = e0; c < e1; c += SPMM_CHUNK) {
    const int cnt = (int)min((int64_t)SPMM_CHUNK, e1 - c);
    __syncthreads();
    for (int t = threadIdx.x; t < cnt; t += blockDim.x) {
      StagedEntry e;
      e.koff = (unsigned long long)kidx[c + t] * (unsigned long long)ldw;
      e.val = vals[c + t];
      se[t] = e;
    }
    __syncthreads();
#pragma unroll 4
    for (int t = 0; t < cnt; t++) {
      const StagedEntry e = se[t];
      const double* wp = wcol + e.koff;
#pragma unroll
      for (int j = 0; j < NV2; j++) {
        if (act[j]) {
          double2 w = __ldg(reinterpret_cast<const double2*>(wp + j * stride));
          acc[2 * j] = fma(e.val, w.x, acc[2 * j]);
          acc[2 * j + 1] = fma(e.val, w.y, acc[2 * j + 1]);
        }
      }
    }
  }
  const int64_t r1 = R / om.nR2, r2 = R - r1 * om.nR2;
  const int64_t rbase = r1 * om.sR1 + r2 * om.sR2;
#pragma unroll
  for (int j = 0; j < NV2; j++) {
    if (act[j]) {
#pragma unroll
      for (int v = 0; v < 2; v++) {
        int64_t l = l_begin + col0 + (int64_t)j * stride + v;
        int64_t l1 = l / om.nL2, l2 = l - l1 * om.nL2;
        int64_t o = rbase + l1 * om.sL1 + l2 * om.sL2;
        out[o] += scale * acc[2 * j + v];
      }
    }
  }
}

static int g_spmm_exact = 0;    // "spmm_exact_order": 1 = the reference's summation order and roundings (bit-identical)
static int g_spmm_nv2 = 2;      // double2 per thread in the wide kernel (1, 2 or 4); 0 = narrow kernel
static int g_spmm_threads = 64;  // measured best on B200 at the headline shape: 64 threads x 2 double2 (256-column tiles)
void set_spmm_option(const char* key, int value) {
  if (strcmp(key, "spmm_nv2") == 0) g_spmm_nv2 = value;
  else if (strcmp(key, "spmm_threads") == 0) g_spmm_threads = value;
  else if (strcmp(key, "spmm_exact_order") == 0) g_spmm_exact = value;
  else fail(FRICC_EINVAL, "unknown option '%s'", key);
}

void spmm_csr(const int64_t* rowptr, const uint32_t* kidx, const double* vals, int64_t nR, const double* Wp, int64_t ldw,
              int64_t l_begin, int64_t l_count, double* out, const OutMap& om, double scale) {
  if (nR == 0 || l_count == 0) return;
  bool vec2 = (ldw % 2 == 0) && (l_count % 2 == 0) && ((reinterpret_cast<uintptr_t>(Wp) & 15) == 0);
  if (vec2 && g_spmm_nv2 > 0) {
    int64_t pairs = l_count / 2;
    int threads = 32;
    while (threads < g_spmm_threads && threads < pairs) threads <<= 1;
    int nv2 = g_spmm_nv2;
    while (nv2 > 1 && (int64_t)threads * (nv2 / 2) >= pairs) nv2 >>= 1;  // do not over-provision small free pairs
    int64_t tiles = ceil_div(pairs, (int64_t)threads * nv2);
    if (tiles > 65535) fail(FRICC_EINVAL, "spmm: free pair too large (%lld columns)", (long long)l_count);
    dim3 grid((unsigned)nR, (unsigned)tiles);
    if (nv2 >= 4)
      FRICC_LAUNCH(spmm_rows_wide_kernel<4>, grid, threads, 0, rowptr, kidx, vals, Wp, ldw, l_begin,
                   l_count, out, om, scale);
    else if (nv2 == 2)
      FRICC_LAUNCH(spmm_rows_wide_kernel<2>, grid, threads, 0, rowptr, kidx, vals, Wp, ldw, l_begin,
                   l_count, out, om, scale);
    else
      FRICC_LAUNCH(spmm_rows_wide_kernel<1>, grid, threads, 0, rowptr, kidx, vals, Wp, ldw, l_begin,
                   l_count, out, om, scale);
    return;
  }
  int vec = vec2 ? 2 : 1;
  int64_t threads_needed = ceil_div(l_count, vec);
  int threads = 32;
  while (threads < 256 && threads < g_spmm_threads * 4 && threads < threads_needed) threads <<= 1;
  int64_t tiles = ceil_div(threads_needed, threads);
  if (tiles > 65535) fail(FRICC_EINVAL, "spmm: free pair too large (%lld columns)", (long long)l_count);
  dim3 grid((unsigned)nR, (unsigned)tiles);
  if (vec2)
    FRICC_LAUNCH(spmm_rows_kernel<2>, grid, threads, 0, rowptr, kidx, vals, Wp, ldw, l_begin, l_count,
                 out, om, scale);
  else
    FRICC_LAUNCH(spmm_rows_kernel<1>, grid, threads, 0, rowptr, kidx, vals, Wp, ldw, l_begin, l_count,
                 out, om, scale);
}

void spmm(const Layout& lay, const double* Wp, int64_t ldw, int64_t l_begin, int64_t l_count, double* out,
          const OutMap& om, double scale) {
  spmm_csr(lay.rowptr.p, lay.kidx.p, lay.vals.p, lay.nR, Wp, ldw, l_begin, l_count, out, om, scale);
}

// ---------------------------------------------------------------------------------------------------------------
// "exact order" form: bit-identical to the reference's loop nest (option "spmm_exact_order")
// ---------------------------------------------------------------------------------------------------------------
// The reference adds the stored entries of an owner pair in STORAGE order s = 0..m-1, one rounded product and one
// rounded sum per entry, straight onto the output element (sparse_tensor.cpp:66-73 and the ten kernels like it; a plain
// x86-64 build: no FMA).  Every prefactor is a power of two, so (scale * W) * value has the one rounding of W * value.
// The fast kernels above sum in (R, K) order with FMAs into a zero accumulator -- within 1e-10, not the same bits.  An
// FRI-CCSD trajectory feeds every residual back into the next compression, and the pivotal sampler's decisions can sit
// on an exact tie, so the energy parity of SURVEY.md 8c (P7) needs identical bits; this form provides them.
__global__ void make_rkeys_kernel(const int64_t* __restrict__ flat, int64_t m, Dims4 dm, int r0, int r1,
                                  uint64_t* __restrict__ keys, uint64_t* __restrict__ pay) {
  int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m) return;
  int64_t f = flat[e];
  int64_t ix[4];
  ix[3] = f % dm.d[3]; f /= dm.d[3];
  ix[2] = f % dm.d[2]; f /= dm.d[2];
  ix[1] = f % dm.d[1]; f /= dm.d[1];
  ix[0] = f;
  keys[e] = (uint64_t)(ix[r0] * dm.d[r1] + ix[r1]);
  pay[e] = (uint64_t)e;
}

__global__ void gather_storage_kernel(const uint64_t* __restrict__ pay, const int64_t* __restrict__ flat,
                                      const double* __restrict__ vals, int64_t m, Dims4 dm, int k0, int k1,
                                      uint32_t* __restrict__ kidx, double* __restrict__ ovals) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= m) return;
  const int64_t e = (int64_t)pay[t];
  int64_t f = flat[e];
  int64_t ix[4];
  ix[3] = f % dm.d[3]; f /= dm.d[3];
  ix[2] = f % dm.d[2]; f /= dm.d[2];
  ix[1] = f % dm.d[1]; f /= dm.d[1];
  ix[0] = f;
  kidx[t] = (uint32_t)(ix[k0] * dm.d[k1] + ix[k1]);
  ovals[t] = vals[e];
}

static void build_exact_order(fricc_st* T, const TermSpec& ts, Layout& lay) {
  const int64_t m = T->nnz;
  const int64_t* d = T->dims;
  lay.xo_rowptr.alloc((size_t)lay.nR + 1);
  lay.xo_kidx.alloc((size_t)std::max<int64_t>(m, 1));
  lay.xo_vals.alloc((size_t)std::max<int64_t>(m, 1));
  if (m == 0) {
    lay.xo_rowptr.zero();
    lay.xo_built = true;
    return;
  }
  DevBuf<uint64_t> ka((size_t)m), kb((size_t)m), pa((size_t)m), pb((size_t)m);
  Dims4 dm{{d[0], d[1], d[2], d[3]}};
  const unsigned g = (unsigned)ceil_div(m, 256);
  FRICC_LAUNCH(make_rkeys_kernel, g, 256, 0, T->flat.p, m, dm, ts.tR[0], ts.tR[1], ka.p, pa.p);
  int bits = 1;
  while (bits < 63 && ((int64_t)1 << bits) < lay.nR) bits++;
  std::vector<Digit> digits;
  for (int s = 0; s < bits; s += 8) digits.push_back(Digit{0, s, 0});
  uint64_t* kk[2] = {ka.p, kb.p};
  uint64_t* pp[2] = {pa.p, pb.p};
  const int r = radix_sort_pairs(kk, pp, m, digits);  // stable: a bucket keeps its storage order
  FRICC_LAUNCH(gather_storage_kernel, g, 256, 0, pp[r], T->flat.p, T->vals.p, m, dm, ts.tK[0], ts.tK[1], lay.xo_kidx.p,
               lay.xo_vals.p);
  FRICC_LAUNCH(rowptr_kernel, (unsigned)ceil_div(lay.nR + 1, 256), 256, 0, kk[r], m, (int64_t)1, lay.nR, lay.xo_rowptr.p);
  lay.xo_built = true;
}

// one CTA = one bucket R x 128 columns; a thread owns one output element and walks the bucket in storage order
__global__ void __launch_bounds__(128)
spmm_rows_exact_kernel(const int64_t* __restrict__ rowptr, const uint32_t* __restrict__ kidx,
                       const double* __restrict__ vals, const double* __restrict__ Wp, int64_t ldw, int64_t l_begin,
                       int64_t l_count, double* __restrict__ out, OutMap om, double scale) {
  __shared__ StagedEntry se[SPMM_CHUNK];
  const int64_t R = blockIdx.x;
  const int64_t e0 = rowptr[R], e1 = rowptr[R + 1];
  if (e0 == e1) return;
  const int64_t col = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
  const bool active = col < l_count;
  int64_t o = 0;
  double acc = 0.0;
  if (active) {
    const int64_t r1 = R / om.nR2, r2 = R - r1 * om.nR2;
    const int64_t l = l_begin + col;
    const int64_t l1 = l / om.nL2, l2 = l - l1 * om.nL2;
    o = r1 * om.sR1 + r2 * om.sR2 + l1 * om.sL1 + l2 * om.sL2;
    acc = out[o];
  }
  for (int64_t c = e0; c < e1; c += SPMM_CHUNK) {
    const int cnt = (int)min((int64_t)SPMM_CHUNK, e1 - c);
    __syncthreads();
    for (int t = threadIdx.x; t < cnt; t += blockDim.x) {
      StagedEntry e;
      e.koff = (unsigned long long)kidx[c + t] * (unsigned long long)ldw;
      e.val = vals[c + t];
      se[t] = e;
    }
    __syncthreads();
    if (active) {
#pragma unroll 4
      for (int t = 0; t < cnt; t++) {
        const StagedEntry e = se[t];
        const double w = __ldg(Wp + e.koff + col);
        acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(scale, w), e.val));  // output += (scale * W) * value, two roundings
      }
    }
  }
  if (active) out[o] = acc;
}

bool spmm_exact_order() { return g_spmm_exact != 0; }

void spmm_exact(fricc_st* T, const TermSpec& ts, const double* Wp, int64_t ldw, int64_t l_begin, int64_t l_count,
                double* out, const OutMap& om) {
  Layout& lay = get_layout(T, ts, false);
  if (!lay.xo_built) build_exact_order(T, ts, lay);
  if (lay.nR == 0 || l_count == 0 || T->nnz == 0) return;
  const int64_t tiles = ceil_div(l_count, 128);
  if (tiles > 65535) fail(FRICC_EINVAL, "spmm: free pair too large (%lld columns)", (long long)l_count);
  dim3 grid((unsigned)lay.nR, (unsigned)tiles);
  FRICC_LAUNCH(spmm_rows_exact_kernel, grid, 128, 0, lay.xo_rowptr.p, lay.xo_kidx.p, lay.xo_vals.p, Wp, ldw, l_begin,
               l_count, out, om, ts.scale);
}

void spmm_term(fricc_st* T, const TermSpec& ts, const double* Wp, int64_t ldw, int64_t l_begin, int64_t l_count,
               double* out, const OutMap& om) {
  if (g_spmm_exact) {
    spmm_exact(T, ts, Wp, ldw, l_begin, l_count, out, om);
    return;
  }
  spmm(get_layout(T, ts), Wp, ldw, l_begin, l_count, out, om, ts.scale);
}

// gathered column-shard panels [g][R][ld] -> strided 4-D output (one read-modify-write per element, no atomics)
__global__ void accumulate_panels_kernel(const double* __restrict__ panels, int64_t total, int64_t ld, int64_t nR,
                                         int64_t nL, double* __restrict__ out, OutMap om) {
  int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  int64_t c = idx % ld;
  int64_t t = idx / ld;
  int64_t R = t % nR, g = t / nR;
  int64_t l = g * ld + c;
  if (l >= nL) return;
  int64_t r1 = R / om.nR2, r2 = R - r1 * om.nR2;
  int64_t l1 = l / om.nL2, l2 = l - l1 * om.nL2;
  out[r1 * om.sR1 + r2 * om.sR2 + l1 * om.sL1 + l2 * om.sL2] += panels[idx];
}

// Tiled variant for outputs whose unit-stride axis is NOT the L-minor one (the panel's unit-stride axis): a 32x32
// shared-memory transpose between the out-contiguous index X (R major, R minor or L major) and the panel-contiguous
// index l2, with the two remaining indices fixed per CTA.  Reads are 256-byte runs of the panel, writes 256-byte runs
// of out; one read-modify-write per output element, no atomics.
//   XMODE 0: X = r1 (fixed: r2, l1)    XMODE 1: X = r2 (fixed: r1, l1)    XMODE 2: X = l1 (fixed: r1, r2)
template <int XMODE>
__global__ void __launch_bounds__(256)
accumulate_panels_tiled_kernel(const double* __restrict__ panels, int64_t ld, int64_t nR, int64_t nL, int64_t nR1,
                               int64_t nL1, double* __restrict__ out, OutMap om) {
  __shared__ double tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t fixed = blockIdx.x;
  const int64_t x0 = (int64_t)blockIdx.y * 32, q0 = (int64_t)blockIdx.z * 32;  // q = l2
  int64_t r1 = 0, r2 = 0, l1 = 0, nX;
  if (XMODE == 0) { r2 = fixed % om.nR2; l1 = fixed / om.nR2; nX = nR1; }
  else if (XMODE == 1) { r1 = fixed % nR1; l1 = fixed / nR1; nX = om.nR2; }
  else { r2 = fixed % om.nR2; r1 = fixed / om.nR2; nX = nL1; }
  // load: rows of the tile = X, columns = l2 (contiguous in the panel)
  for (int r = ty; r < 32; r += 8) {
    const int64_t x = x0 + r, q = q0 + tx;
    if (x < nX && q < om.nL2) {
      int64_t R, l;
      if (XMODE == 0) { R = x * om.nR2 + r2; l = l1 * om.nL2 + q; }
      else if (XMODE == 1) { R = r1 * om.nR2 + x; l = l1 * om.nL2 + q; }
      else { R = r1 * om.nR2 + r2; l = x * om.nL2 + q; }
      const int64_t g = l / ld, c = l - g * ld;
      tile[r][tx] = panels[(g * nR + R) * ld + c];
    }
  }
  __syncthreads();
  // store: rows = l2, columns = X (contiguous in out)
  for (int r = ty; r < 32; r += 8) {
    const int64_t q = q0 + r, x = x0 + tx;
    if (x < nX && q < om.nL2) {
      int64_t o;
      if (XMODE == 0) o = x * om.sR1 + r2 * om.sR2 + l1 * om.sL1 + q * om.sL2;
      else if (XMODE == 1) o = r1 * om.sR1 + x * om.sR2 + l1 * om.sL1 + q * om.sL2;
      else o = r1 * om.sR1 + r2 * om.sR2 + x * om.sL1 + q * om.sL2;
      out[o] += tile[tx][r];
    }
  }
}

void accumulate_panels(const double* panels, int64_t n_panels, int64_t ld, int64_t nR, int64_t nL, double* out,
                       const OutMap& om) {
  int64_t total = n_panels * nR * ld;
  if (total <= 0) return;
  const int64_t nR1 = nR / om.nR2, nL1 = nL / om.nL2;
  // pick the transposing kernel when out's unit-stride axis is r1, r2 or l1 (and the grid fits)
  int xmode = -1;
  int64_t fixed = 0, nX = 0;
  if (om.sL2 != 1 && n_panels * ld >= nL) {
    if (om.sR1 == 1) { xmode = 0; fixed = om.nR2 * nL1; nX = nR1; }
    else if (om.sR2 == 1) { xmode = 1; fixed = nR1 * nL1; nX = om.nR2; }
    else if (om.sL1 == 1) { xmode = 2; fixed = nR1 * om.nR2; nX = nL1; }
  }
  if (xmode >= 0 && fixed > 0 && fixed < ((int64_t)1 << 31) && ceil_div(nX, 32) <= 65535 && ceil_div(om.nL2, 32) <= 65535) {
    dim3 grid((unsigned)fixed, (unsigned)ceil_div(nX, 32), (unsigned)ceil_div(om.nL2, 32));
    if (xmode == 0)
      FRICC_LAUNCH(accumulate_panels_tiled_kernel<0>, grid, 256, 0, panels, ld, nR, nL, nR1, nL1, out, om);
    else if (xmode == 1)
      FRICC_LAUNCH(accumulate_panels_tiled_kernel<1>, grid, 256, 0, panels, ld, nR, nL, nR1, nL1, out, om);
    else
      FRICC_LAUNCH(accumulate_panels_tiled_kernel<2>, grid, 256, 0, panels, ld, nR, nL, nR1, nL1, out, om);
    return;
  }
  FRICC_LAUNCH(accumulate_panels_kernel, (unsigned)ceil_div(total, 256), 256, 0, panels, total, ld, nR, nL, out, om);
}

}  // namespace fricc
