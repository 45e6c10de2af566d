// km_common.cuh — shared device/host definitions for libkmeans.so (sm_100a).
//
// Semantics are SURVEY.md Appendix A; the reference (schusto/k-means-demo) has no
// arithmetic to cite for them.  See DESIGN.md for the data layout.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define KM_MAX_K (1 << 20)
#define KM_MAX_D (1 << 16)

// ---------------------------------------------------------------------------------
// Accumulator table `acc` (int64, global memory, one per ctx; a second copy holds the
// all-reduced table when world > 1).  Integer adds are associative, so any order of
// atomics, any blocking and any number of GPUs give the same bits.
//   [0, kd)           Sa   coarse limb of sum_j[t]
//   [kd, 2kd)         Sb   fine limb
//   [2kd, 2kd+k)      member counts
//   [2kd+k + 0]       changed      (per iteration)
//   [2kd+k + 1]       inertia limb a
//   [2kd+k + 2]       inertia limb b
//   [2kd+k + 3]       rescored
// ---------------------------------------------------------------------------------
#define KM_ACC_SCALARS 4
#define KM_ACC_CHANGED 0
#define KM_ACC_INERTIA_A 1
#define KM_ACC_INERTIA_B 2
#define KM_ACC_RESCORED 3

struct KmAccLayout {
  int64_t sa, sb, cnt, scal, len;
};
static inline __host__ __device__ KmAccLayout km_acc_layout(int32_t k, int32_t d) {
  KmAccLayout L;
  int64_t kd = (int64_t)k * d;
  L.sa = 0; L.sb = kd; L.cnt = 2 * kd; L.scal = 2 * kd + k; L.len = 2 * kd + k + KM_ACC_SCALARS;
  return L;
}

// Device-resident per-ctx state that one kernel hands to the next without the host.
struct KmState {
  double scale_a;        // 2^qa : point value -> coarse limb units
  double scale_b;        // 2^QB : fraction    -> fine limb units
  double inertia_scale;  // 2^qi : winner distance -> coarse limb units
  double cmax;           // max |c_jt| over the current centroids
  float tau_hilo;        // absolute error term of the hi/lo binary32 filter (small-d path)
  float tau_hi;          // absolute error term of the hi-only binary32 filter (tiled path)
  int32_t QB, qa, qi, E;
  unsigned int ticket;   // last-block election in the finalize kernel
  int32_t it;            // assign steps since the last label reset
  unsigned long long cmax_bits;  // scratch for the finalize kernel's max reduction
  float cnmax2;          // tensor path: max_j |chi_j|^2 of the current centroids
  float pad0;
  // inertia from the sums: T = sum_i |x_i|^2 of the whole job as order-free limbs
  // on the grid 2^-(qt + QBt), and the accumulator of the per-element terms of one finalize
  long long Ta, Tb;
  unsigned long long g_a, g_b;
  int32_t qt, QBt;       // T's grid: QBt = 61 - NB - bitlen(d), qt = QBt - 2E
};

// Small-d path ("prologue-finalize" protocol): what one pass reads and writes.
// Every launch reads (S_in, C_in) plus the pending pass's deltas D_in, and block 0 writes the
// merged sums and the centroids it derived to (S_out, C_out).  Points are accumulated into
// D_out; D_zero is the table the pass after this one will accumulate into.
struct KmPass {
  const long long *S_in;
  long long *S_out;
  const double *C_in;
  double *C_out;
  const long long *D_in;   // nullptr: no pass pending
  long long *D_out;
  long long *D_zero;       // may be nullptr
  struct KmStatSlot *slot; // where block 0 publishes the pending pass's stats (or nullptr)
  long long it_pending;
  int full_mode;           // D_in holds complete sums, not deltas
  int apply_update;        // derive centroids from the sums ...
  int needs_change;        // ... only if the pending pass changed a label
  int QB, qa;
  double maxabs_x;
  long long Ta, Tb;        // sum_i |x_i|^2 of the whole job (order-free limbs), for the inertia
  int qt, QBt;
  int qg, QBg;             // grid of the inertia terms: QBg = 60 - NB - bitlen(d), qg = QBg - 2 E(max(max|x|, max|c|))
};

// d == 2 candidate pruning.  A point's cell is RN(x*sx + Kx), RN(y*sy + Ky):
// one FMA per axis against (K + 1.5*2^23), the integer then sits in the low mantissa bits —
// and because the FMA rounds the exact value once, "x belongs to cell ix" means exactly
// |x*sx + Kx - ix| <= 1/2: cells are exact boxes, no slack needed for the mapping.
struct KmPrune2 {
  int gp;              // cells per side (grid + margins)
  float sx, sy;        // cells per unit length
  float kxm, kym;      // K + 12582912.0f, both integers
  float kx, ky;        // K
  float inv_sx, inv_sy;
  float r;             // half diagonal of a cell
  float absmax;        // max |coordinate| of the bounding box
};

// One slot of the pinned-host stats ring the finalize kernel writes.
struct KmStatSlot {
  long long it;
  long long changed;
  double inertia;
  long long rescored;
  int updated;   // 1 if this finalize rewrote the centroids
  int pad;
};

// ---------------------------------------------------------------------------------
// A8 generator (libm-free; bit-identical on host and device)
// ---------------------------------------------------------------------------------
#define KM_GAMMA 0x9E3779B97F4A7C15ULL
#define KM_CENTRE_TAG 0xC3A5C85C97CB3127ULL

static inline __host__ __device__ uint64_t km_splitmix64(uint64_t z) {
  z += KM_GAMMA;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
static inline __host__ __device__ uint64_t km_hash3(uint64_t seed, uint64_t a, uint64_t b) {
  return km_splitmix64(seed ^ (a * KM_GAMMA) ^ (b * 0xBF58476D1CE4E5B9ULL));
}

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------
// Order-free fixed point: value -> (coarse, fine) int64 limbs.  All operations are exact
// or a single deterministic rounding of that one value, so the sum of limbs depends on
// nothing but the multiset of inputs.
// ---------------------------------------------------------------------------------
__device__ __forceinline__ void km_fixed_split(double v, double scale_a, double scale_b,
                                               long long &a, long long &b) {
  double xs = v * scale_a;            // power-of-two scaling: exact
  double tr = trunc(xs);
  a = __double2ll_rz(xs);
  double fr = __dsub_rn(xs, tr);      // exact
  b = __double2ll_rn(fr * scale_b);   // ties to even
}

// (Sa * 2^QB + Sb) * 2^-shift, rounded once to binary64 (ties to even).
__device__ __forceinline__ double km_fixed_to_double(long long Sa, long long Sb, int QB, int shift) {
  // 128-bit signed T = Sa << QB  +  Sb   (QB in [1, 61])
  unsigned long long lo = (unsigned long long)Sa << QB;
  long long hi = Sa >> (64 - QB);
  unsigned long long lo2 = lo + (unsigned long long)Sb;
  long long hi2 = hi + (Sb >> 63) + (lo2 < lo ? 1 : 0);
  if (hi2 == 0 && lo2 == 0) return 0.0;
  bool neg = hi2 < 0;
  unsigned long long mh = (unsigned long long)hi2, ml = lo2;
  if (neg) { ml = ~ml + 1ULL; mh = ~mh + (ml == 0 ? 1ULL : 0ULL); }
  int hb = mh ? 127 - __clzll((long long)mh) : 63 - __clzll((long long)ml);
  unsigned long long mant;
  int e2 = 0;
  if (hb <= 52) {
    mant = ml;
  } else {
    int sh = hb - 52;  // 1..75
    // q = m >> sh ; rem = m & (2^sh - 1)
    unsigned long long q, rem_hi, rem_lo, half_hi, half_lo;
    if (sh >= 64) {
      q = mh >> (sh - 64);
      rem_hi = (sh == 64) ? 0ULL : (mh & ((1ULL << (sh - 64)) - 1ULL));
      rem_lo = ml;
      half_hi = (sh == 64) ? 0ULL : (1ULL << (sh - 65));
      half_lo = (sh == 64) ? (1ULL << 63) : 0ULL;
    } else {
      q = (ml >> sh) | (mh << (64 - sh));   // mh << (64-sh) is safe: sh in [1,63]
      rem_hi = 0ULL;
      rem_lo = ml & ((1ULL << sh) - 1ULL);
      half_hi = 0ULL;
      half_lo = 1ULL << (sh - 1);
    }
    mant = q;
    bool gt = (rem_hi > half_hi) || (rem_hi == half_hi && rem_lo > half_lo);
    bool eq = (rem_hi == half_hi) && (rem_lo == half_lo);
    if (gt || (eq && (mant & 1ULL))) mant++;
    e2 = sh;
  }
  double v = scalbn((double)mant, e2 - shift);
  return neg ? -v : v;
}

__device__ __forceinline__ void km_atomic_add_ll(long long *p, long long v) {
  atomicAdd(reinterpret_cast<unsigned long long *>(p), (unsigned long long)v);
}

// binary64 distance in the oracle's exact operation order (A1): no FMA contraction.
template <typename XT>
__device__ __forceinline__ double km_dist2_f64(const XT *__restrict__ xrow, const double *__restrict__ c, int d) {
  double acc = 0.0;
  for (int t = 0; t < d; t++) {
    double df = __dsub_rn((double)xrow[t], c[t]);
    acc = __dadd_rn(acc, __dmul_rn(df, df));
  }
  return acc;
}

// ---------------------------------------------------------------------------------
// Inertia from the sums.  With labels l(i), centroids c and the exact per-cluster sums S_j, n_j:
//     sum_i |x_i - c_l(i)|^2  =  T  -  sum_{j,t} c_jt (2 S_jt - n_j c_jt),     T = sum_i |x_i|^2
// an identity of real numbers, so no per-point distance is needed for the total.  T is a
// constant of the data set (one exact pass at upload), S and n are the order-free sums the update
// keeps anyway.  Every term is evaluated in double-double arithmetic (~104 bits) and put on the
// fixed-point grid of T individually, so the result is a sum of integers: identical for any
// blocking or GPU count, and good to ~2^-100 * T even when T is many orders of magnitude larger
// than the inertia (data far from the origin).
// ---------------------------------------------------------------------------------
struct KmDD { double hi, lo; };
__device__ __forceinline__ KmDD km_two_sum(double a, double b) {
  KmDD r;
  r.hi = __dadd_rn(a, b);
  const double bb = __dsub_rn(r.hi, a);
  r.lo = __dadd_rn(__dsub_rn(a, __dsub_rn(r.hi, bb)), __dsub_rn(b, bb));
  return r;
}
__device__ __forceinline__ KmDD km_two_prod(double a, double b) {
  KmDD r;
  r.hi = __dmul_rn(a, b);
  r.lo = __fma_rn(a, b, -r.hi);
  return r;
}
__device__ __forceinline__ KmDD km_dd_add(KmDD a, KmDD b) {
  KmDD s = km_two_sum(a.hi, b.hi);
  const KmDD t = km_two_sum(a.lo, b.lo);
  s.lo = __dadd_rn(s.lo, t.hi);
  s = km_two_sum(s.hi, s.lo);
  s.lo = __dadd_rn(s.lo, t.lo);
  return km_two_sum(s.hi, s.lo);
}
__device__ __forceinline__ KmDD km_dd_mul_d(KmDD a, double b) {
  KmDD p = km_two_prod(a.hi, b);
  p.lo = __fma_rn(a.lo, b, p.lo);
  return km_two_sum(p.hi, p.lo);
}
// exact value of the limb pair (Sa * 2^QB + Sb) * 2^-shift as a double-double (error < 2^-105 relative)
__device__ __forceinline__ KmDD km_fixed_to_dd(long long Sa, long long Sb, int QB, int shift) {
  // four exactly representable pieces: Sa = ah * 2^32 + al, Sb = bh * 2^32 + bl
  const double ah = (double)(Sa >> 32), al = (double)(unsigned)(Sa & 0xFFFFFFFFLL);
  const double bh = (double)(Sb >> 32), bl = (double)(unsigned)(Sb & 0xFFFFFFFFLL);
  KmDD r = km_two_sum(scalbn(ah, 32 + QB - shift), scalbn(al, QB - shift));
  r = km_dd_add(r, km_two_sum(scalbn(bh, 32 - shift), scalbn(bl, -shift)));
  return r;
}
// one (cluster, dimension) term  c (2 S - n c)  on the grid 2^-(qg+QBg), as two limbs
__device__ __forceinline__ void km_inertia_term(double c, long long Sa, long long Sb, long long cnt, int QB, int shift,
                                                double scale_g, double scale_gb, long long &ga, long long &gb) {
  ga = 0; gb = 0;
  if (cnt <= 0) return;   // no members: S = 0, n = 0
  KmDD S = km_fixed_to_dd(Sa, Sb, QB, shift);
  S.hi = __dmul_rn(S.hi, 2.0); S.lo = __dmul_rn(S.lo, 2.0);
  KmDD nc = km_two_prod((double)cnt, c);           // cnt < 2^53: exact as a double
  nc.hi = -nc.hi; nc.lo = -nc.lo;
  const KmDD g = km_dd_mul_d(km_dd_add(S, nc), c);
  long long a, b;
  km_fixed_split(g.hi, scale_g, scale_gb, a, b);
  ga += a; gb += b;
  km_fixed_split(g.lo, scale_g, scale_gb, a, b);
  ga += a; gb += b;
}
// T (limbs on the grid 2^-(qt+QBt), set by max|x|) minus G (the sum of the terms above, limbs on the
// grid 2^-(qg+QBg), set by max(max|x|, max|c|)) -> binary64.  Both limb pairs are exact as
// double-doubles (<= 92 bits); the identity is >= 0 but its rounding may not be, hence the clamp.
__device__ __forceinline__ double km_inertia_from_limbs(long long Ta, long long Tb, int QBt, int qt, long long Ga,
                                                        long long Gb, int QBg, int qg) {
  const KmDD T = km_fixed_to_dd(Ta, Tb, QBt, qt + QBt);
  KmDD G = km_fixed_to_dd(Ga, Gb, QBg, qg + QBg);
  G.hi = -G.hi; G.lo = -G.lo;
  const KmDD r = km_dd_add(T, G);
  const double v = __dadd_rn(r.hi, r.lo);
  return v > 0.0 ? v : 0.0;
}

__device__ __forceinline__ long long km_warp_sum_ll(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
#endif  // __CUDACC__
