/*
 * oracle/fricc_oracle.c -- TEST INFRASTRUCTURE ONLY.  A plain-C, single-threaded restatement of the reference's
 * FRI-CCSD hot path (jamesETsmith/fri-cc @ 8b481f0), used as the checker in tests/, __graft_entry__.smoke() and as
 * bench.py's CPU-baseline fallback.  Nothing under fricc_b200/ imports, links or executes it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function below against the compiled, unmodified reference
 * (oracle/_ref, built by oracle/Makefile from /root/reference) and against the golden vectors in tests/golden/ that
 * were generated from that build (tests/golden/make_golden.py).
 *
 * Compile with -ffp-contract=off (oracle/Makefile does): the reference is a baseline x86-64 build without FMA.
 * Each function cites the reference lines it follows.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_OK 0
#define ORC_EPROB (-3)    /* p[i] > 1                          fri_utils.cpp:383-395 (reference exits) */
#define ORC_EPIVOTAL (-4) /* acceptance probability not in [0,1] fri_utils.cpp:84-94  (reference exits) */
#define ORC_ECOUNT (-5)   /* wrong number of samples           fri_utils.cpp:110-124 (reference exits) */
#define ORC_EDRAWS (-6)   /* ran out of injected draws                                                  */
#define ORC_EINVAL (-1)

/* ---- glibc rand(): TYPE_3 additive feedback generator (what the reference's rand() calls resolve to) ------------- */
void orc_glibc_rand_fill(uint32_t seed, int64_t skip, int64_t n, int32_t* out) {
  uint32_t ring[31];
  int32_t r = (int32_t)(seed ? seed : 1);
  int64_t i;
  ring[0] = (uint32_t)r;
  for (i = 1; i < 31; i++) {
    int64_t hi = r / 127773, lo = r % 127773;
    int64_t word = 16807 * lo - 2836 * hi;
    if (word < 0) word += 2147483647;
    r = (int32_t)word;
    ring[i] = (uint32_t)r;
  }
  for (i = 34; i < 344 + skip + n; i++) {
    uint32_t v = ring[(i - 31) % 31] + ring[(i - 3) % 31];
    ring[i % 31] = v;
    if (i >= 344 + skip) out[i - 344 - skip] = (int32_t)(v >> 1);
  }
}

/* ---- one_norm: fri_utils.cpp:164-171 with OMP_NUM_THREADS=1 (left-to-right accumulation) ------------------------- */
double orc_one_norm(const double* q, int64_t n) {
  double s = 0.0;
  int64_t i;
  for (i = 0; i < n; i++) s += fabs(q[i]);
  return s;
}

/* ---- argsort: fri_utils.cpp:11-36.  The reference merge-sorts (value,index) pairs by |value| descending and keeps
 * the first m indices; inside runs of equal |value| its order is whatever libstdc++'s introsort leaves (SURVEY.md
 * 7.2-5).  The restatement uses the canonical order (|value| descending, index ascending): identical wherever the
 * magnitudes are distinct, and the same multiset of magnitudes always.                                            */
typedef struct {
  double a;
  int64_t i;
} orc_pair;
static int orc_pair_cmp(const void* pa, const void* pb) {
  const orc_pair* a = (const orc_pair*)pa;
  const orc_pair* b = (const orc_pair*)pb;
  if (a->a > b->a) return -1;
  if (a->a < b->a) return 1;
  return (a->i > b->i) - (a->i < b->i);
}
int orc_argsort(const double* x, int64_t n, int64_t m, int64_t* out) {
  orc_pair* p;
  int64_t i;
  if (m > n || m < 0) return ORC_EINVAL;
  p = (orc_pair*)malloc(sizeof(orc_pair) * (size_t)(n ? n : 1));
  for (i = 0; i < n; i++) {
    p[i].a = fabs(x[i]);
    p[i].i = i;
  }
  qsort(p, (size_t)n, sizeof(orc_pair), orc_pair_cmp);
  for (i = 0; i < m; i++) out[i] = p[i].i;
  free(p);
  return ORC_OK;
}

/* ---- get_d_largest: fri_utils.cpp:400-437 ---------------------------------------------------------------------- */
int orc_get_d_largest(const double* x, int64_t n, int64_t n_sample, int64_t* D, int64_t* nD, double* remaining_norm) {
  double r = orc_one_norm(x, n);
  int64_t* idx = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_sample ? n_sample : 1));
  int64_t i, d = 0;
  int rc = orc_argsort(x, n, n_sample, idx);
  if (rc) {
    free(idx);
    return rc;
  }
  for (i = 0; i < n_sample; i++) {
    double xi = fabs(x[idx[i]]);
    /* (n_sample - d) * xi >= remaining_norm - xi && remaining_norm > 1e-14      :415 */
    if ((double)(n_sample - d) * xi >= r - xi && r > 1e-14) {
      D[d++] = idx[i];
      r -= xi;
    } else {
      break;
    }
  }
  free(idx);
  *nD = d;
  *remaining_norm = r;
  return ORC_OK;
}

/* ---- make_probability_vector: fri_utils.cpp:367-398 ------------------------------------------------------------ */
int orc_make_probability_vector(const double* x, int64_t n, int64_t n_sample, const int64_t* D, int64_t nD,
                                double remaining_norm, double* p) {
  int64_t i;
  for (i = 0; i < n; i++) p[i] = fabs(x[i]) / remaining_norm * (double)n_sample; /* :375 */
  for (i = 0; i < nD; i++) p[D[i]] = 0.0;                                       /* :379-381 */
  for (i = 0; i < n; i++)
    if (p[i] > 1.) return ORC_EPROB; /* :383-395 */
  return ORC_OK;
}

/* ---- sample_pivotal: fri_utils.cpp:63-127.  draws[t] is the value the t-th rand() call returns ------------------- */
int orc_sample_pivotal(int64_t n_sample, const double* probs, int64_t n, const int32_t* draws, int64_t n_draws,
                       int64_t* S, int64_t* consumed) {
  int64_t cnt = 0, i = 0, j, used = 0;
  double a;
  if (n < 1) return ORC_EINVAL;
  a = probs[0];
  for (j = 1; j < n; j++) {
    double rn;
    a += probs[j];
    if (used >= n_draws) {
      *consumed = used;
      return ORC_EDRAWS;
    }
    rn = (double)draws[used++] / 2147483647.0; /* (double)rand() / RAND_MAX */
    if (a < 1.) {
      if (rn < probs[j] / a) i = j;
    } else {
      double prob_accept;
      a -= 1.;
      prob_accept = (1. - probs[j]) / (1. - a);
      if (prob_accept > 1 || prob_accept < 0) {
        *consumed = used;
        return ORC_EPIVOTAL;
      }
      if (rn < prob_accept) {
        if (cnt < n_sample) S[cnt] = i;
        cnt++;
        i = j;
      } else {
        if (cnt < n_sample) S[cnt] = j;
        cnt++;
      }
    }
    if (cnt == n_sample) break; /* :103-105 */
    if (j == n - 1) {           /* :106-118 */
      if (fabs(a - 1.) < 1e-5) {
        if (cnt < n_sample) S[cnt] = i;
        cnt++;
      } else {
        *consumed = used;
        return ORC_ECOUNT;
      }
    }
  }
  *consumed = used;
  if (cnt != n_sample) return ORC_ECOUNT; /* :121-124 */
  return ORC_OK;
}

/* ---- parallel_sample ("Alg. A.2"): fri_utils.cpp:176-362 -----------------------------------------------------------
 * q is cut into n_chunks contiguous chunks (the reference uses omp_get_max_threads(); the last chunk takes the
 * remainder, :183-184,199,216-219).  Per chunk: budget a_j = n_sample * |q_j|_1 / |q|_1, g_j = floor(a_j),
 * t_j = a_j - g_j (:226-231); the c = n_sample - sum g_j left-over samples are given to chunks by a pivotal draw over t
 * (:268-275); a chunk whose budget was rounded up / down has its leading probabilities raised / lowered until its mass
 * is the integer g_j (:303-322); each chunk is then sampled with sample_pivotal (:345) and the indices are offset by
 * j*step and concatenated in chunk order (:353-356).
 * Draw order: the reference's threads call rand() concurrently, i.e. the order is a race for n_chunks > 1.  This
 * restatement consumes the draws in chunk order (chunk j after chunk j-1) -- what the reference does whenever its
 * threads happen to run one after another, and exactly what it does with one thread.  Pinned against oracle/_ref with
 * one thread (random draws) and with several threads on a constant draw stream (where the race cannot matter).
 * work: n doubles of scratch (the per-chunk copies qj, :196-198).                                                   */
int orc_parallel_sample(int64_t n_sample, const double* q, int64_t n, int n_chunks, const int32_t* draws,
                        int64_t n_draws, int64_t* S, int64_t* consumed, double* work) {
  int64_t step, rem, c, used = 0, k, lb = 0;
  double q_norm;
  double *t, *a, *norms;
  int64_t *g, *sp;
  int j, rc = ORC_OK;
  if (n_chunks < 1 || n < 0 || n_sample < 0) return ORC_EINVAL;
  step = n / n_chunks;
  rem = n % n_chunks;
  t = (double*)malloc(sizeof(double) * (size_t)n_chunks * 3);
  g = (int64_t*)malloc(sizeof(int64_t) * (size_t)n_chunks * 2 + sizeof(int64_t));
  if (!t || !g) {
    free(t);
    free(g);
    return ORC_EINVAL;
  }
  a = t + n_chunks;
  norms = a + n_chunks;
  sp = g + n_chunks;
  *consumed = 0;
  q_norm = orc_one_norm(q, n); /* :191 */
  c = n_sample;
  for (j = 0; j < n_chunks; j++) { /* :214-237 */
    int64_t lo = (int64_t)j * step, len = step + ((j + 1 == n_chunks) ? rem : 0);
    memcpy(work + lo, q + lo, sizeof(double) * (size_t)len);
    norms[j] = orc_one_norm(work + lo, len);
    a[j] = (double)n_sample * norms[j] / q_norm; /* size_t * double / double, left to right */
    t[j] = a[j] - floor(a[j]);
    g[j] = (int64_t)(int)floor(a[j]); /* std::vector<int> g */
    c -= g[j];
  }
  /* :268  s_prime = sample_pivotal(c, t);  c is a size_t in the reference: a negative value cannot be sampled */
  if (c < 0) {
    rc = ORC_ECOUNT;
    goto done;
  }
  {
    int64_t u = 0;
    rc = orc_sample_pivotal(c, t, n_chunks, draws, n_draws, sp, &u);
    used += u;
    if (rc != ORC_OK) goto done;
    for (k = 0; k < c; k++) g[sp[k]] += 1; /* :269-275 */
  }
  for (j = 0; j < n_chunks; j++) { /* :297-357, chunk order */
    int64_t lo = (int64_t)j * step, len = step + ((j + 1 == n_chunks) ? rem : 0), i, u = 0;
    double* qj = work + lo;
    double sj = norms[j];
    const double gj = (double)g[j];
    if (gj > a[j]) { /* :303-312 */
      for (i = 0; i < len; i++) {
        double yji = qj[i] / t[j];
        if (!(yji < 1.)) yji = 1.; /* std::min(1., x): returns 1. unless x < 1. */
        sj += yji - qj[i];
        qj[i] = yji;
        if (sj >= gj) {
          qj[i] = yji + gj - sj;
          break;
        }
      }
    } else if (gj < a[j]) { /* :313-322 */
      for (i = 0; i < len; i++) {
        double yji = (qj[i] - t[j]) / (1. - t[j]);
        if (!(0. < yji)) yji = 0.; /* std::max(0., x): returns 0. unless 0. < x */
        sj += yji - qj[i];
        qj[i] = yji;
        if (sj <= gj) {
          qj[i] = yji + gj - sj;
          break;
        }
      }
    }
    rc = orc_sample_pivotal(g[j], qj, len, draws + used, n_draws - used, S + lb, &u); /* :345 */
    used += u;
    if (rc != ORC_OK) goto done;
    for (i = 0; i < g[j]; i++) S[lb + i] += lo; /* :351-353  incr = j * step */
    lb += g[j];
  }
done:
  *consumed = used;
  free(t);
  free(g);
  return rc;
}

/* ---- sample_systematic + lin_space_add_const: fri_utils.cpp:129-159, 45-58 --------------------------------------- */
int orc_sample_systematic(int64_t n_sample, const double* probs, int64_t n, const int32_t* draws, int64_t n_draws,
                          int64_t* S, int64_t* consumed) {
  double rn, step;
  double* uk;
  double* intervals;
  int64_t i, j, last_interval = 0;
  if (n_draws < 1) return ORC_EDRAWS;
  rn = (double)draws[0] / 2147483647.0; /* :47 -- exactly one draw */
  *consumed = 1;
  step = ((double)n_sample - 0.0) / (double)n_sample; /* (stop - start) / num       :50 */
  uk = (double*)malloc(sizeof(double) * (size_t)(n_sample ? n_sample : 1));
  for (i = 0; i < n_sample; i++) uk[i] = (double)i * step + rn; /* :54 */
  intervals = (double*)malloc(sizeof(double) * (size_t)(n + 1));
  intervals[0] = 0.0;
  if (n > 0) { /* std::partial_sum                                                  :142-143 */
    double acc = probs[0];
    intervals[1] = acc;
    for (j = 1; j < n; j++) {
      acc = acc + probs[j];
      intervals[j + 1] = acc;
    }
  }
  for (i = 0; i < n_sample; i++) S[i] = 0; /* std::vector<size_t> S(n_sample)    :132 */
  for (i = 0; i < n_sample; i++) {         /* :147-157 */
    for (j = last_interval; j < n; j++) {
      if (uk[i] >= intervals[j] && uk[i] < intervals[j + 1]) {
        S[i] = j;
        last_interval = j;
        break;
      }
    }
  }
  free(uk);
  free(intervals);
  return ORC_OK;
}

/* ---- fri_compression: fri_utils.cpp:439-522.  method: 0 pivotal, 1 systematic ----------------------------------- */
int orc_fri_compression(const double* x, int64_t n, int64_t n_sample, int method, const int32_t* draws,
                        int64_t n_draws, int64_t* idx, double* vals, int64_t* nD_out, double* rnorm_out,
                        int64_t* consumed) {
  int64_t* D = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_sample ? n_sample : 1));
  double* p = (double*)malloc(sizeof(double) * (size_t)(n ? n : 1));
  int64_t nD = 0, i, used = 0;
  double r = 0;
  int rc = orc_get_d_largest(x, n, n_sample, D, &nD, &r); /* :466 */
  if (!rc) rc = orc_make_probability_vector(x, n, n_sample - nD, D, nD, r, p); /* :475 */
  if (!rc) {
    for (i = 0; i < nD; i++) idx[i] = D[i]; /* :494 */
    if (method == 0) rc = orc_sample_pivotal(n_sample - nD, p, n, draws, n_draws, idx + nD, &used); /* :485-489 */
    else rc = orc_sample_systematic(n_sample - nD, p, n, draws, n_draws, idx + nD, &used);
  }
  if (!rc) {
    for (i = 0; i < nD; i++) vals[i] = x[idx[i]];                       /* :498-500 */
    for (i = nD; i < n_sample; i++) vals[i] = x[idx[i]] / p[idx[i]];    /* :502-504 */
  }
  if (nD_out) *nD_out = nD;
  if (rnorm_out) *rnorm_out = r;
  if (consumed) *consumed = used;
  free(D);
  free(p);
  return rc;
}

/* ---- "largest" compression: sparse_tensor.hpp:46-55 -------------------------------------------------------------- */
int orc_largest(const double* x, int64_t n, int64_t m, int64_t* idx, double* vals) {
  int64_t i;
  int rc = orc_argsort(x, n, m, idx);
  if (rc) return rc;
  for (i = 0; i < m; i++) vals[i] = x[idx[i]];
  return ORC_OK;
}

/* ---- unpack_idx: sparse_tensor.hpp:108-125 ------------------------------------------------------------------------ */
int orc_unpack_idx(int64_t idx, const int64_t dims[4], int64_t out[4]) {
  int q;
  out[0] = idx / (dims[1] * dims[2] * dims[3]);
  out[1] = idx % (dims[1] * dims[2] * dims[3]) / (dims[2] * dims[3]);
  out[2] = idx % (dims[2] * dims[3]) / dims[3];
  out[3] = idx % dims[3];
  for (q = 0; q < 4; q++)
    if (out[q] >= dims[q] || out[q] < 0) return ORC_EINVAL;
  return ORC_OK;
}

/* ---- contractions: sparse_tensor.cpp:12-346.  Every term is the same loop nest -- owner pair, scan of the stored
 * entries in storage order, two free loops -- so each output element receives its contributions in ascending entry
 * order; the restatement iterates entries outermost, which gives every output element the same sequence of updates.
 * W and out are addressed through element strides (numpy semantics, like pybind11's unchecked accessors).          */
#define W4(a, b, c, d) W[(a)*ws[0] + (b)*ws[1] + (c)*ws[2] + (d)*ws[3]]
#define O4(a, b, c, d) out[(a)*os[0] + (b)*os[1] + (c)*os[2] + (d)*os[3]]

int orc_contract(const char* term, const double* W, const int64_t ws[4], const int64_t* flat, const double* vals,
                 int64_t m, const int64_t dims[4], double* out, const int64_t os[4]) {
  const int64_t no = dims[0], nv = dims[2];
  int64_t s, ix[4];
  int code;
  if (!strcmp(term, "0101")) code = 0;
  else if (!strcmp(term, "2323") || !strcmp(term, "taskloop") || !strcmp(term, "atomic")) code = 1;
  else if (!strcmp(term, "1302")) code = 2;
  else if (!strcmp(term, "1202")) code = 3;
  else if (!strcmp(term, "1303")) code = 4;
  else if (!strcmp(term, "1203")) code = 5;
  else if (!strcmp(term, "1323")) code = 6;
  else if (!strcmp(term, "0112")) code = 7;
  else if (!strcmp(term, "0313")) code = 8;
  else if (!strcmp(term, "0113")) code = 9;
  else if (!strcmp(term, "0312")) code = 10;
  else return ORC_EINVAL; /* "CASE NOT IMPLEMENTED"  :379-380 */
  for (s = 0; s < m; s++) {
    const double value = vals[s];
    int64_t p, q;
    if (orc_unpack_idx(flat[s], dims, ix)) return ORC_EINVAL;
    switch (code) {
      case 0: { /* 0101  :12-37   a=ix2 b=ix3 k=ix0 l=ix1 */
        for (p = 0; p < no; p++)
          for (q = 0; q < no; q++) O4(p, q, ix[2], ix[3]) += W4(ix[0], ix[1], p, q) * value;
      } break;
      case 1: { /* 2323  :40-86   i=ix0 j=ix1 c=ix2 d=ix3 */
        for (p = 0; p < nv; p++)
          for (q = 0; q < nv; q++) O4(ix[0], ix[1], p, q) += W4(p, q, ix[2], ix[3]) * value;
      } break;
      case 2: { /* 1302  :89-115  b=ix3 j=ix1 k=ix0 c=ix2 ; loops i (p), a (q) */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(p, ix[1], q, ix[3]) += 2 * W4(q, ix[0], p, ix[2]) * value;
      } break;
      case 3: { /* 1202  :118-144 b=ix3 j=ix1 k=ix0 c=ix2 */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(p, ix[1], q, ix[3]) -= W4(q, ix[0], ix[2], p) * value;
      } break;
      case 4: { /* 1303  :147-173 b=ix2 j=ix1 k=ix0 c=ix3 */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(p, ix[1], q, ix[2]) += W4(q, ix[0], p, ix[3]) * value;
      } break;
      case 5: { /* 1203  :176-202 a=ix2 j=ix1 k=ix0 c=ix3 ; loops i (p), b (q) */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(p, ix[1], ix[2], q) += W4(q, ix[0], ix[3], p) * value;
      } break;
      case 6: { /* 1323  :205-230 i=ix0 j=ix1 c=ix2 d=ix3 ; loops k (p), l (q) */
        for (p = 0; p < no; p++)
          for (q = 0; q < no; q++) O4(p, q, ix[0], ix[1]) += W4(p, ix[2], q, ix[3]) * value;
      } break;
      case 7: { /* 0112  :233-259 a=ix3 i=ix0 l=ix1 d=ix2 ; loops k (p), c (q) */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(ix[3], p, ix[0], q) -= 0.5 * W4(ix[1], ix[2], p, q) * value;
      } break;
      case 8: { /* 0313  :262-289 a=ix2 i=ix0 l=ix1 d=ix3 */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(ix[2], p, ix[0], q) -= 0.5 * W4(ix[1], q, p, ix[3]) * value;
      } break;
      case 9: { /* 0113  :292-317 a=ix2 i=ix0 l=ix1 d=ix3 */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(ix[2], p, ix[0], q) += W4(ix[1], ix[3], p, q) * value;
      } break;
      case 10: { /* 0312 :320-346 a=ix3 i=ix0 l=ix1 d=ix2 */
        for (p = 0; p < no; p++)
          for (q = 0; q < nv; q++) O4(ix[3], p, q, ix[0]) -= 0.5 * W4(ix[1], q, p, ix[2]) * value;
      } break;
    }
  }
  return ORC_OK;
}
