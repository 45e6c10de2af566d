/*
 * oracle/kmeans_oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT.
 *
 * CPU restatement of the Lloyd iteration the GPU library (libkmeans.so) is checked
 * against.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` legs may call into this file; the product never does.
 *
 * PARITY UNPINNED.  The reference (schusto/k-means-demo) holds no numerical k-means:
 * its "centroids" carry no coordinates (app.mjs:128), its "assignment" is a person
 * dragging a card (app.mjs:146-156, 358-372), its "update" renames a centroid after the
 * dominant trait tokens of its members (app.mjs:450-461, 476-480, 571-573), and its
 * "iteration" is a hand-typed counter (app.mjs:286-288, 498-508).  It has no tests, no
 * golden vectors and cannot be executed here (app.mjs:88-103 does not parse; no JS
 * engine on the image).  The arithmetic below therefore follows SURVEY.md Appendix A
 * ("OUR SPEC"): what a plain JavaScript Lloyd over `Number`s would compute —
 * IEEE binary64, left-to-right evaluation, no FMA contraction, strict-`<` argmin.
 * The few idioms that *do* come from the reference are cited where used.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -shared -fPIC (oracle/build.py).
 * -ffp-contract=off is load-bearing: x86-64 SSE2 doubles then behave like V8's.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define KO_F32 0
#define KO_F64 1
#define KO_SUM_SEQUENTIAL 0 /* A3: sum_j[t] += x_it for i = 0..N-1, binary64 */
#define KO_SUM_EXACT 1      /* order-free two-limb fixed point */

/* ------------------------------------------------------------------ A8: generator */

uint64_t ko_splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

uint64_t ko_hash3(uint64_t seed, uint64_t a, uint64_t b) {
  return ko_splitmix64(seed ^ (a * 0x9E3779B97F4A7C15ULL) ^ (b * 0xBF58476D1CE4E5B9ULL));
}

#define KO_CENTRE_TAG 0xC3A5C85C97CB3127ULL

/* centre of blob b, dim t: uniform in [-10, 10) from the top 53 hash bits */
static double blob_centre(uint64_t seed, int64_t b, int32_t t) {
  uint64_t h = ko_hash3(seed ^ KO_CENTRE_TAG, (uint64_t)b, (uint64_t)t);
  double u = (double)(h >> 11) * 0x1.0p-53;
  double v = u * 20.0;
  return v - 10.0;
}

void ko_blob_centres(uint64_t seed, int32_t k_true, int32_t d, double *mu) {
  for (int32_t b = 0; b < k_true; b++)
    for (int32_t t = 0; t < d; t++) mu[(size_t)b * d + t] = blob_centre(seed, b, t);
}

/* Irwin-Hall(12) - 6 from twelve 16-bit hash fields: integer adds + one exact scale. */
static double blob_noise(uint64_t seed, int64_t i, int32_t t) {
  uint32_t s = 0;
  for (int r = 0; r < 3; r++) {
    uint64_t h = ko_hash3(seed, (uint64_t)i, (uint64_t)(3 * (int64_t)t + r));
    s += (uint32_t)(h & 0xFFFF) + (uint32_t)((h >> 16) & 0xFFFF) +
         (uint32_t)((h >> 32) & 0xFFFF) + (uint32_t)(h >> 48);
  }
  double z = (double)s * 0x1.0p-16;
  return z - 6.0;
}

/* rows [i0, i0+n) of the N x d matrix; point i belongs to blob i mod k_true */
void ko_gen_blobs(uint64_t seed, int64_t i0, int64_t n, int32_t d, int32_t k_true,
                  double sigma, float *x) {
  for (int64_t r = 0; r < n; r++) {
    int64_t i = i0 + r;
    int64_t b = i % k_true;
    for (int32_t t = 0; t < d; t++) {
      double mu = blob_centre(seed, b, t);
      double z = blob_noise(seed, i, t);
      double sz = sigma * z;
      double v = mu + sz;
      x[(size_t)r * d + t] = (float)v;
    }
  }
}

/* ------------------------------------------------------------------ A7: Forgy init
 * Partial Fisher-Yates, the loop shape of app.mjs:163
 *   for(let i=U.length-1;i>0;i--){ const j=Math.floor(Math.random()*(i+1)); swap }
 * cut off after k swaps and driven by a seeded splitmix64 stream instead of
 * Math.random().  idx[] is held sparsely (only displaced slots), so N = 1e9 is fine. */
typedef struct { int64_t key, val; } ko_slot;

static int64_t sparse_get(const ko_slot *tab, int n, int64_t key) {
  for (int i = 0; i < n; i++) if (tab[i].key == key) return tab[i].val;
  return key;
}
static void sparse_set(ko_slot *tab, int *n, int64_t key, int64_t val) {
  for (int i = 0; i < *n; i++) if (tab[i].key == key) { tab[i].val = val; return; }
  tab[*n].key = key; tab[*n].val = val; (*n)++;
}

int32_t ko_forgy_indices(int64_t n, int32_t k, uint64_t seed, int64_t *out) {
  if (k < 0 || k > n) return -1;
  ko_slot *tab = (ko_slot *)malloc(sizeof(ko_slot) * (size_t)(2 * k + 2));
  if (!tab) return -2;
  int used = 0;
  uint64_t state = seed;
  for (int32_t s = 0; s < k; s++) {
    int64_t i = n - 1 - s;
    uint64_t r = ko_splitmix64(state);
    state += 0x9E3779B97F4A7C15ULL;
    double u = (double)(r >> 11) * 0x1.0p-53;
    int64_t j = (int64_t)floor(u * (double)(i + 1));
    int64_t vi = sparse_get(tab, used, i), vj = sparse_get(tab, used, j);
    sparse_set(tab, &used, i, vj);
    sparse_set(tab, &used, j, vi);
    out[s] = vj; /* idx[i] after the swap */
  }
  free(tab);
  return 0;
}

/* ------------------------------------------------------------------ A1/A2/A6: assign */

static inline double load_x(const void *x, int dtype, size_t off) {
  return dtype == KO_F32 ? (double)((const float *)x)[off] : ((const double *)x)[off];
}

/* dist^2 in binary64, dims ascending, separate multiply and add (A1) */
static inline double dist2(const void *x, int dtype, size_t row, int32_t d, const double *cj) {
  double acc = 0.0;
  for (int32_t t = 0; t < d; t++) {
    double v = load_x(x, dtype, row * (size_t)d + t);
    double df = v - cj[t];
    double sq = df * df;
    acc = acc + sq;
  }
  return acc;
}

/* ------------------------------------------------------------------ A7b: k-means++ (f1)
 * OUR SPEC (the reference seeds nothing numerically; its only seeded-choice idiom is the
 * Math.random() draw of /root/reference/app.mjs:163, kept here as u = 53 bits / 2^53):
 *   u_0, u_1, ... from splitmix64 exactly as in A7;  first centre = point floor(u_0 * N)
 *   for s = 1 .. k-1:
 *     D_i  = min(D_i, dist2(x_i, centre_{s-1}))            binary64, A1 operation order
 *     q_i  = trunc(D_i * 2^qd)  (int64),  qd = 62 - NB - (2E + 2 + bitlen(d)):  q_i sums to < 2^62
 *     T    = sum_i q_i                                      exact integers: order-free
 *     if T == 0: next = floor(u_s * N)  (every point sits on a centre already)
 *     else target = ((r_s >> 11) * T) >> 53  (128-bit product),  next = least i with q_0+..+q_i > target
 * Quantising D to integers is what makes the draw independent of how the sum is split across
 * blocks and GPUs (the same device the per-cluster sums use). */
void ko_fixed_params(double maxabs, int64_t n_total, int32_t *E, int32_t *NB, int32_t *QB, int32_t *qa);

int32_t ko_kmeanspp_indices(const void *x, int dtype, int64_t n, int32_t d, int32_t k, uint64_t seed,
                            double maxabs, int64_t n_total, int64_t *out) {
  if (k < 1 || k > n) return -1;
  int32_t E, NB, QB, qa;
  ko_fixed_params(maxabs, n_total, &E, &NB, &QB, &qa);
  int32_t dbits = 0;
  while (((int32_t)1 << dbits) <= d) dbits++;
  const int32_t qd = 62 - NB - (2 * E + 2 + dbits);
  double *D = (double *)malloc(sizeof(double) * (size_t)n);
  double *c = (double *)malloc(sizeof(double) * (size_t)d);
  if (!D || !c) { free(D); free(c); return -2; }
  for (int64_t i = 0; i < n; i++) D[i] = INFINITY;
  uint64_t state = seed;
  for (int32_t s = 0; s < k; s++) {
    uint64_t r = ko_splitmix64(state);
    state += 0x9E3779B97F4A7C15ULL;
    int64_t pick;
    if (s == 0) {
      pick = (int64_t)floor((double)(r >> 11) * 0x1.0p-53 * (double)n);
    } else {
      int64_t total = 0;
      for (int64_t i = 0; i < n; i++) {
        double di = dist2(x, dtype, (size_t)i, d, c);
        if (di < D[i]) D[i] = di;
        total += (int64_t)ldexp(D[i], qd); /* the cast truncates toward zero */
      }
      if (total == 0) {
        pick = (int64_t)floor((double)(r >> 11) * 0x1.0p-53 * (double)n);
      } else {
        const int64_t target = (int64_t)(((unsigned __int128)(r >> 11) * (unsigned __int128)(uint64_t)total) >> 53);
        int64_t run = 0;
        pick = n - 1;
        for (int64_t i = 0; i < n; i++) {
          run += (int64_t)ldexp(D[i], qd);
          if (run > target) { pick = i; break; }
        }
      }
    }
    out[s] = pick;
    for (int32_t t = 0; t < d; t++) c[t] = load_x(x, dtype, (size_t)pick * (size_t)d + t);
  }
  free(D);
  free(c);
  return 0;
}

static inline int32_t argmin_row(const void *x, int dtype, size_t i, int32_t d, const double *c,
                                 int32_t k, double *best_out) {
  double best = INFINITY;
  int32_t lab = 0;
  for (int32_t j = 0; j < k; j++) {
    double dj = dist2(x, dtype, i, d, c + (size_t)j * d);
    if (dj < best) { best = dj; lab = j; } /* strict <: lowest index wins ties, NaN never wins */
  }
  *best_out = best;
  return lab;
}

/* Single thread, index order: the restatement proper.  labels[] is in/out
 * (compare against the previous labels; -1 = never assigned).  Returns `changed`. */
int64_t ko_assign(const void *x, int dtype, int64_t n, int32_t d, const double *c, int32_t k,
                  int32_t *labels, double *inertia_out) {
  int64_t changed = 0;
  double inertia = 0.0;
  for (int64_t i = 0; i < n; i++) {
    double best;
    int32_t lab = argmin_row(x, dtype, (size_t)i, d, c, k, &best);
    if (lab != labels[i]) changed++;
    labels[i] = lab;
    inertia = inertia + best; /* A6: sequential, index order */
  }
  if (inertia_out) *inertia_out = inertia;
  return changed;
}

/* All host threads — NOT the restatement: labels are identical (per-point work is
 * independent) but the inertia is a sum of per-thread partials taken in thread order.
 * Used only as the many-core CPU timing arm. */
int64_t ko_assign_mt(const void *x, int dtype, int64_t n, int32_t d, const double *c, int32_t k,
                     int32_t *labels, double *inertia_out, int32_t nthreads) {
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
  int64_t changed = 0;
  double *part = (double *)calloc((size_t)nthreads, sizeof(double));
#pragma omp parallel num_threads(nthreads) reduction(+ : changed)
  {
    int tid = omp_get_thread_num(), nt = omp_get_num_threads();
    int64_t lo = n * tid / nt, hi = n * (tid + 1) / nt;
    double acc = 0.0;
    for (int64_t i = lo; i < hi; i++) {
      double best;
      int32_t lab = argmin_row(x, dtype, (size_t)i, d, c, k, &best);
      if (lab != labels[i]) changed++;
      labels[i] = lab;
      acc = acc + best;
    }
    part[tid] = acc;
  }
  double inertia = 0.0;
  for (int t = 0; t < nthreads; t++) inertia = inertia + part[t];
  free(part);
  if (inertia_out) *inertia_out = inertia;
  return changed;
#else
  (void)nthreads;
  return ko_assign(x, dtype, n, d, c, k, labels, inertia_out);
#endif
}

/* ------------------------------------------------------------------ A3/A4: update */

/* Sequential binary64 sums in index order; empty cluster keeps its centroid — the
 * analog of a member-less centroid simply persisting in the reference
 * (app.mjs:322, 484-488; cohesionFor's n<=1 -> 1 at app.mjs:463). */
void ko_update_seq(const void *x, int dtype, int64_t n, int32_t d, const int32_t *labels,
                   int32_t k, double *c, int64_t *counts) {
  double *sum = (double *)calloc((size_t)k * d, sizeof(double));
  memset(counts, 0, sizeof(int64_t) * (size_t)k);
  for (int64_t i = 0; i < n; i++) {
    int32_t j = labels[i];
    if (j < 0 || j >= k) continue;
    counts[j] += 1;
    for (int32_t t = 0; t < d; t++)
      sum[(size_t)j * d + t] = sum[(size_t)j * d + t] + load_x(x, dtype, (size_t)i * d + t);
  }
  for (int32_t j = 0; j < k; j++)
    if (counts[j] > 0)
      for (int32_t t = 0; t < d; t++) c[(size_t)j * d + t] = sum[(size_t)j * d + t] / (double)counts[j];
  free(sum);
}

/* ---- order-free sums: two int64 limbs per (cluster, dim) -------------------------
 * With maxabs < 2^E and n_total < 2^NB:  QB = 62 - NB,  qa = QB - E.
 *   xs = x * 2^qa            (exact scaling; |xs| < 2^QB)
 *   a  = trunc(xs)           (int64)
 *   b  = rint((xs - a) * 2^QB)   (int64, ties to even; exact for every binary32 x
 *                                 whose ulp is >= 2^-(qa+QB))
 * Sum_a, Sum_b are plain int64 adds (associative => identical on any partition of
 * the points, in any order, on any number of devices).  value = (Sa*2^QB + Sb) * 2^-(qa+QB),
 * rounded to binary64 once (ties to even), then one IEEE division by the count. */
void ko_fixed_params(double maxabs, int64_t n_total, int32_t *E, int32_t *NB, int32_t *QB,
                     int32_t *qa) {
  int32_t e = 0, nb = 1;
  if (maxabs > 0.0) e = (int32_t)ilogb(maxabs) + 1;
  while (nb < 62 && ((int64_t)1 << nb) <= n_total) nb++;
  *E = e; *NB = nb; *QB = 62 - nb; *qa = 62 - nb - e;
}

void ko_fixed_split(double v, int32_t qa, int32_t QB, int64_t *a, int64_t *b) {
  double xs = ldexp(v, qa);
  double tr = trunc(xs);
  double fr = xs - tr;
  *a = (int64_t)tr;
  *b = (int64_t)llrint(ldexp(fr, QB)); /* default rounding mode: to nearest even */
}

static int clz128(unsigned __int128 m) {
  uint64_t hi = (uint64_t)(m >> 64), lo = (uint64_t)m;
  if (hi) return __builtin_clzll(hi);
  return 64 + __builtin_clzll(lo);
}

/* (Sa*2^QB + Sb) * 2^-shift rounded once to binary64, ties to even */
double ko_fixed_to_double(int64_t Sa, int64_t Sb, int32_t QB, int32_t shift) {
  __int128 T = ((__int128)Sa << QB) + (__int128)Sb;
  if (T == 0) return 0.0;
  int neg = T < 0;
  unsigned __int128 m = neg ? (unsigned __int128)(-T) : (unsigned __int128)T;
  int hb = 127 - clz128(m);
  uint64_t mant;
  int e2 = 0;
  if (hb <= 52) {
    mant = (uint64_t)m;
  } else {
    int sh = hb - 52;
    unsigned __int128 q = m >> sh;
    unsigned __int128 rem = m & ((((unsigned __int128)1) << sh) - 1);
    unsigned __int128 half = ((unsigned __int128)1) << (sh - 1);
    mant = (uint64_t)q;
    if (rem > half || (rem == half && (mant & 1))) mant++;
    e2 = sh;
  }
  double v = ldexp((double)mant, e2 - shift);
  return neg ? -v : v;
}

void ko_update_exact(const void *x, int dtype, int64_t n, int32_t d, const int32_t *labels,
                     int32_t k, double maxabs, int64_t n_total, double *c, int64_t *counts,
                     int64_t *Sa_out, int64_t *Sb_out) {
  int32_t E, NB, QB, qa;
  ko_fixed_params(maxabs, n_total, &E, &NB, &QB, &qa);
  int64_t *Sa = (int64_t *)calloc((size_t)k * d, sizeof(int64_t));
  int64_t *Sb = (int64_t *)calloc((size_t)k * d, sizeof(int64_t));
  memset(counts, 0, sizeof(int64_t) * (size_t)k);
  for (int64_t i = 0; i < n; i++) {
    int32_t j = labels[i];
    if (j < 0 || j >= k) continue;
    counts[j] += 1;
    for (int32_t t = 0; t < d; t++) {
      int64_t a, b;
      ko_fixed_split(load_x(x, dtype, (size_t)i * d + t), qa, QB, &a, &b);
      Sa[(size_t)j * d + t] += a;
      Sb[(size_t)j * d + t] += b;
    }
  }
  if (c)
    for (int32_t j = 0; j < k; j++)
      if (counts[j] > 0)
        for (int32_t t = 0; t < d; t++) {
          double s = ko_fixed_to_double(Sa[(size_t)j * d + t], Sb[(size_t)j * d + t], QB, qa + QB);
          c[(size_t)j * d + t] = s / (double)counts[j];
        }
  if (Sa_out) memcpy(Sa_out, Sa, sizeof(int64_t) * (size_t)k * d);
  if (Sb_out) memcpy(Sb_out, Sb, sizeof(int64_t) * (size_t)k * d);
  free(Sa); free(Sb);
}

/* exact-sum form of the inertia: sum of the per-point best distances, order-free.
 * scale: every best < 2^Ei (caller passes Ei), qi = QB - Ei. */
double ko_inertia_exact(const double *best, int64_t n, int64_t n_total, int32_t Ei) {
  int32_t E, NB, QB, qa;
  ko_fixed_params(1.0, n_total, &E, &NB, &QB, &qa);
  int32_t qi = QB - Ei;
  int64_t Sa = 0, Sb = 0;
  for (int64_t i = 0; i < n; i++) {
    int64_t a, b;
    ko_fixed_split(best[i], qi, QB, &a, &b);
    Sa += a; Sb += b;
  }
  return ko_fixed_to_double(Sa, Sb, QB, qi + QB);
}

/* per-point best distances (for the exact inertia and for near-tie analysis) */
void ko_best_distances(const void *x, int dtype, int64_t n, int32_t d, const double *c, int32_t k,
                       double *best, double *second) {
  for (int64_t i = 0; i < n; i++) {
    double b1 = INFINITY, b2 = INFINITY;
    for (int32_t j = 0; j < k; j++) {
      double dj = dist2(x, dtype, (size_t)i, d, c + (size_t)j * d);
      if (dj < b1) { b2 = b1; b1 = dj; } else if (dj < b2) b2 = dj;
    }
    best[i] = b1;
    if (second) second[i] = b2;
  }
}

/* f2: per-cluster inertia for given labels and centroids — members' A1 distances to their own
 * centroid, summed sequentially in index order (A6 restricted to one cluster) */
void ko_cluster_inertia(const void *x, int dtype, int64_t n, int32_t d, const int32_t *labels,
                        const double *c, int32_t k, double *out) {
  for (int32_t j = 0; j < k; j++) out[j] = 0.0;
  for (int64_t i = 0; i < n; i++) {
    int32_t j = labels[i];
    if (j < 0 || j >= k) continue;
    out[j] = out[j] + dist2(x, dtype, (size_t)i, d, c + (size_t)j * d);
  }
}

double ko_maxabs(const void *x, int dtype, int64_t count) {
  double m = 0.0;
  for (int64_t i = 0; i < count; i++) {
    double v = fabs(load_x(x, dtype, (size_t)i));
    if (v > m) m = v;
  }
  return m;
}

/* ------------------------------------------------------------------ A5: run
 *   it=0; repeat { assign -> changed; it+=1; if (changed==0 || it==max_iter) break; update }
 * labels start at -1, so changed == N on the first pass.  Returns the number of assign
 * steps; c[] ends as the centroids the last assign used.  nthreads: 1 = the
 * restatement; >1 or 0 (= all) = the OpenMP timing arm. */
int32_t ko_run(const void *x, int dtype, int64_t n, int32_t d, int32_t k, double *c,
               int32_t *labels, int32_t max_iter, int32_t sum_mode, double maxabs,
               int64_t n_total, int32_t nthreads, double *inertia_out, int64_t *counts_out,
               int64_t *changed_out) {
  int64_t *counts = (int64_t *)calloc((size_t)k, sizeof(int64_t));
  for (int64_t i = 0; i < n; i++) labels[i] = -1;
  int32_t it = 0;
  int64_t changed = 0;
  double inertia = 0.0;
  for (;;) {
    changed = (nthreads == 1) ? ko_assign(x, dtype, n, d, c, k, labels, &inertia)
                              : ko_assign_mt(x, dtype, n, d, c, k, labels, &inertia, nthreads);
    it += 1;
    if (changed == 0 || it >= max_iter) break;
    if (sum_mode == KO_SUM_EXACT)
      ko_update_exact(x, dtype, n, d, labels, k, maxabs, n_total, c, counts, NULL, NULL);
    else
      ko_update_seq(x, dtype, n, d, labels, k, c, counts);
  }
  /* member counts of the final labelling (a6: the reference's per-centroid counts,
   * app.mjs:485-486) */
  memset(counts, 0, sizeof(int64_t) * (size_t)k);
  for (int64_t i = 0; i < n; i++) if (labels[i] >= 0 && labels[i] < k) counts[labels[i]]++;
  if (counts_out) memcpy(counts_out, counts, sizeof(int64_t) * (size_t)k);
  if (inertia_out) *inertia_out = inertia;
  if (changed_out) *changed_out = changed;
  free(counts);
  return it;
}

/* a6: balance{max,min,gap,ratio} with the reference's own formulas (app.mjs:489-493):
 *   ratio = min ? max/min : (max ? Infinity : 1) */
void ko_balance(const int64_t *counts, int32_t k, int64_t *mx, int64_t *mn, int64_t *gap,
                double *ratio) {
  int64_t a = 0, b = 0;
  if (k > 0) { a = counts[0]; b = counts[0]; }
  for (int32_t j = 1; j < k; j++) { if (counts[j] > a) a = counts[j]; if (counts[j] < b) b = counts[j]; }
  *mx = a; *mn = b; *gap = a - b;
  *ratio = b ? (double)a / (double)b : (a ? INFINITY : 1.0);
}

int32_t ko_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
